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
  if (argc > 4 && std::string(argv[1]) == "xml") { // xml <ocean_params.xml> <continuation_params.xml> <mask dir>: run_ocean.C:44-75
    try {
      auto make = [&](const thcm_b200_config& c, const std::vector<int>& mask) {
        auto ocean = std::make_shared<OceanB200>(c, mask.data());
        ocean->setSolverParameters(1e-10, 4000, 500, GMRES_B200_CGS2);
        ocean->usePointBlockJacobi(true);
        ocean->setVerbose(std::getenv("IEMIC_COMPAT_VERBOSE") != nullptr);
        return ocean;
      };
      return run_ocean_from_files<Continuation<std::shared_ptr<OceanB200>>, thcm_b200_config>(argv[2], argv[3], argv[4], make) == 0 ? 0 : 1;
    } catch (const std::exception& e) {
      std::printf("continuation drop-in failed: %s\n", e.what());
      return 1;
    }
  }
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
  const thcm_b200_config c = moc2d_config<thcm_b200_config>(); // test/2dmoc/ocean_params.xml
  const std::vector<int> mask = moc2d_mask(c.n, c.m, c.l);
  try {
    auto ocean = std::make_shared<OceanB200>(c, mask.data());
    moc2d_starting_parameters(ocean); // "Starting Parameters" of test/2dmoc/ocean_params.xml
    // the linear solver of this path: restarted GMRES on the device with the device-built point-block Jacobi
    ocean->setSolverParameters(1e-10, 4000, 500, GMRES_B200_CGS2);
    ocean->usePointBlockJacobi(true);
    ocean->setVerbose(std::getenv("IEMIC_COMPAT_VERBOSE") != nullptr);

    Teuchos::ParameterList pars = moc2d_continuation_params(); // test/2dmoc/continuation_params.xml
    pars.set("destination 0", destination);

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
