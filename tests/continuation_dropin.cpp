// continuation_dropin.cpp -- the drop-in claim for the continuation driver, compiled and run: the reference's UNMODIFIED
// Continuation<Model> template (src/continuation/Continuation.H, read in place from the reference tree, never copied) is
// instantiated with Model = std::shared_ptr<OceanB200> -- the B200 host mirror of Ocean -- and drives the first leg of the
// reference's own integration test src/tests/intt_2dmoc.C:28-40 (3x6x6 2-D MOC box, continuation in "Combined Forcing"
// from 0 to 1 with test/2dmoc/continuation_params.xml) or, with "chain <outdir>", the whole chain of seven continuations of
// that test (:28-120), writing the two end states whose overturning streamfunction the reference asserts (psi_max = 14.7 Sv,
// mirror symmetry of the branches).  Teuchos / Utils / JDQZ are replaced by the test-only stand-ins of tests/compat (a box
// with Trilinos would use the real ones).
// Linear solves: GMRES(500) on the device, CGS2, device-built point-block Jacobi.  The Jacobian of this configuration is
// singular (no fixed pressure points); the solves of Newton's last steps have right-hand sides at the rounding floor of F and
// end at the accuracy guard of the driver (iemic_b200/csrc/gmres_guard.h, return code 2) -- without it the second leg aborts
// at the minimum step size inside the secant destination process (tests/test_continuation_oracle.py shows the same on the
// CPU walk-through of the driver).  On one B200 the seven legs take 21 s and end at psi_max = 14.7043 Sv, psi_min = 0, mirror
// symmetry 5.6e-5 (profiles/r09_continuation_chain_gpu.txt; asserted by tests/test_host_cpp.py).
//
// Build (the test does this): g++ -std=c++17 -Itests/compat -I<reference>/src/continuation -Iinclude -Iiemic_b200/host
//                             tests/continuation_dropin.cpp -Liemic_b200/lib -liemic_b200
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>
#include "OceanB200.H"
#include "Continuation.H"
#include "continuation_cases.H"

using namespace iemic_b200;

static void write_state(OceanB200& ocean, const std::string& path) {
  VectorPtr x = ocean.getState('V');
  std::vector<double> h((size_t)x->MyLength());
  x->toHost(h.data());
  FILE* f = std::fopen(path.c_str(), "wb");
  if (!f) throw std::runtime_error("cannot write " + path);
  std::fwrite(h.data(), sizeof(double), h.size(), f);
  std::fclose(f);
}

int main(int argc, char** argv) {
  if (argc > 2 && std::string(argv[1]) == "ocean") { // ocean <maskfile>: TEST(Ocean, Continuation), test_ocean.C:178-223
    try {
      const thcm_b200_config c = test_ocean_config<thcm_b200_config>();
      const std::vector<int> mask = read_mask_file(argv[2], (size_t)(c.n + 2) * (c.m + 2) * (c.l + 2));
      auto ocean = std::make_shared<OceanB200>(c, mask.data());
      ocean->setSolverParameters(1e-10, 4000, 500, GMRES_B200_CGS2);
      ocean->usePointBlockJacobi(true);
      ocean->setVerbose(std::getenv("IEMIC_COMPAT_VERBOSE") != nullptr);
      return run_test_ocean_continuation<Continuation<std::shared_ptr<OceanB200>>>(ocean) == 0 ? 0 : 1;
    } catch (const std::exception& e) {
      std::printf("continuation drop-in failed: %s\n", e.what());
      return 1;
    }
  }
  const bool chain = argc > 2 && std::string(argv[1]) == "chain";
  const std::string outdir = chain ? argv[2] : "";
  const double destination = (!chain && argc > 1) ? std::atof(argv[1]) : 1.0;
  const double d2r = 3.14159265358979323846 / 180.0;
  const int n = 3, m = 6, l = 6; // test/2dmoc/ocean_params.xml
  thcm_b200_config c{};
  c.n = n; c.m = m; c.l = l;
  c.xmin = 286 * d2r; c.xmax = 350 * d2r; c.ymin = -60 * d2r; c.ymax = 60 * d2r;
  c.periodic = 1; c.hdim = 4000.0; c.qz = 1.0; c.alphaT = 1.0e-4; c.alphaS = 7.6e-4;
  c.ih = 0; c.vmix = 1; c.tap = 1; c.rho_mixing = 0; c.coriolis_on = 0; c.forcing_type = 1;
  c.tres = 1; c.sres = 0; c.iza = 2; c.ite = 1; c.its = 1; c.coupled_T = 0; c.coupled_S = 0;
  c.int_sign = -1; c.intcond_i = -1; c.intcond_j = -1; c.fix_pressure_points = 0;
  std::vector<int> mask((size_t)(n + 2) * (m + 2) * (l + 2), 1); // LAND ring, OCEAN inside, PERIO in the x border columns
  for (int k = 1; k <= l; k++) for (int j = 1; j <= m; j++) for (int i = 0; i <= n + 1; i++)
    mask[(size_t)i + (size_t)(n + 2) * (j + (size_t)(m + 2) * k)] = (i == 0 || i == n + 1) ? 3 : 0;
  try {
    auto ocean = std::make_shared<OceanB200>(c, mask.data());
    // "Starting Parameters" of test/2dmoc/ocean_params.xml
    ocean->setPar("Combined Forcing", 0.0); ocean->setPar("Solar Forcing", 0.0); ocean->setPar("Salinity Forcing", 0.0);
    ocean->setPar("Wind Forcing", 0.0); ocean->setPar("Temperature Forcing", 10.0); ocean->setPar("SPL1", 2.0e3);
    ocean->setPar("SPL2", 0.01); ocean->setPar("P_VC", 0.0); ocean->setPar("Rossby-Number", 0.0); ocean->setPar("CMPR", 0.0);
    ocean->setPar("Horizontal Ekman-Number", 371.764); ocean->setPar("Rayleigh-Number", 15.6869);
    // the linear solver of this path: restarted GMRES on the device with the device-built point-block Jacobi
    ocean->setSolverParameters(1e-10, 4000, 500, GMRES_B200_CGS2);
    ocean->usePointBlockJacobi(true);
    ocean->setVerbose(std::getenv("IEMIC_COMPAT_VERBOSE") != nullptr);

    Teuchos::ParameterList pars("Continuation parameters"); // test/2dmoc/continuation_params.xml
    pars.set("continuation parameter", "Combined Forcing");
    pars.set("initial step size", 1.0e-1); pars.set("minimum step size", 1.0e-8); pars.set("maximum step size", 1.0);
    pars.set("increase step size", 2.0); pars.set("decrease step size", 2.0); pars.set("destination 0", destination);
    pars.set("maximum number of steps", -1); pars.set("Newton tolerance", 1.0e-3); pars.set("destination tolerance", 1.0e-6);
    pars.set("maximum Newton iterations", 15); pars.set("backtracking steps", 5); pars.set("corrector residual test", 'D');
    pars.set("epsilon increment", 1.0e-6); pars.set("state tangent scaling", 1.0);
    pars.set("enable Newton Chord hybrid solve", false); pars.set("backtracking increase", 1.0); pars.set("tangent type", 'S');
    pars.set("predictor bound", 100.0);

    Continuation<std::shared_ptr<OceanB200>> continuation(ocean, pars);
    const int status = continuation.run();
    const double par = ocean->getPar("Combined Forcing");
    ocean->computeRHS();
    const double normF = ocean->getRHS('V')->norm(), normX = ocean->getState('V')->norm();
    std::printf("continuation status %d, Combined Forcing = %.9f, ||x|| = %.6e, ||F|| = %.3e\n", status, par, normX, normF);
    if (status != 0) return 1;
    if (std::fabs(par - destination) > 1e-5) return 1;
    if (!(normX > 0.0) || !(normF < 1e-3)) return 1;
    if (chain) { // src/tests/intt_2dmoc.C:42-120
      struct Leg { const char* par; double dest, ds; const char* save; };
      const Leg legs[] = {{"CMPR", -0.2, -0.5, nullptr}, {"Salinity Forcing", 0.02, 0.5, nullptr}, {"CMPR", 0.0, 0.5, nullptr},
                          {"Salinity Forcing", 0.04, 0.5, "state1.bin"}, {"Salinity Forcing", 0.03, -0.5, nullptr},
                          {"Salinity Forcing", 0.04, -0.5, "state2.bin"}};
      for (const Leg& g : legs) {
        pars.set("continuation parameter", g.par);
        pars.set("destination 0", g.dest);
        pars.set("initial step size", g.ds);
        Continuation<std::shared_ptr<OceanB200>> leg(ocean, pars);
        const int st = leg.run();
        ocean->computeRHS();
        std::printf("leg %s -> %g: status %d, value %.9f, ||x|| = %.6e, ||F|| = %.3e\n", g.par, g.dest, st, ocean->getPar(g.par),
                    ocean->getState('V')->norm(), ocean->getRHS('V')->norm());
        if (st != 0 || std::fabs(ocean->getPar(g.par) - g.dest) > 1e-5) return 1;
        if (g.save) write_state(*ocean, outdir + "/" + g.save);
      }
    }
    std::printf("continuation drop-in ok\n");
  } catch (const std::exception& e) {
    std::printf("continuation drop-in failed: %s\n", e.what());
    return 1;
  }
  return 0;
}
