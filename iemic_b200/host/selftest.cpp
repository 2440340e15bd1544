// selftest.cpp -- exercises the C++ host mirror (OceanB200 / DeviceVector / Newton) end to end on a GPU.
// Build: g++ -std=c++17 -Iinclude -Iiemic_b200/host iemic_b200/host/selftest.cpp -Liemic_b200/lib -liemic_b200 -Wl,-rpath,...
// With -DWITH_REFERENCE_GMRES and -I<reference>/src/gmressolver -Itests/compat the reference's own, unmodified
// GMRESSolver<Model,VectorPointer> template is instantiated on OceanB200/DeviceVector (the drop-in claim, compiled).
#include <cmath>
#include <cstdio>
#include <vector>
#include "NewtonB200.H"
#ifdef WITH_REFERENCE_GMRES
#include <iostream>
#include "GMRESSolver.H"
#endif

using namespace iemic_b200;

static thcm_b200_config box_config(int n, int m, int l) {
  const double d2r = 3.14159265358979323846 / 180.0;
  thcm_b200_config c{};
  c.n = n; c.m = m; c.l = l;
  c.xmin = 286 * d2r; c.xmax = 350 * d2r; c.ymin = 10 * d2r; c.ymax = 74 * d2r;
  c.periodic = 0; c.hdim = 4000.0; c.qz = 1.0; c.alphaT = 1.0e-4; c.alphaS = 7.6e-4;
  c.ih = 0; c.vmix = 1; c.tap = 1; c.rho_mixing = 0; c.coriolis_on = 1; c.forcing_type = 0;
  c.tres = 1; c.sres = 1; c.iza = 2; c.ite = 1; c.its = 1; c.coupled_T = 0; c.coupled_S = 0;
  c.int_sign = -1; c.intcond_i = -1; c.intcond_j = -1; c.fix_pressure_points = 0;
  return c;
}

int main() {
  const int n = 16, m = 16, l = 8;
  std::vector<int> mask((size_t)(n + 2) * (m + 2) * (l + 2), 1);
  for (int k = 1; k <= l; k++) for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++)
    mask[(size_t)i + (size_t)(n + 2) * (j + (size_t)(m + 2) * k)] = 0;
  try {
    OceanB200 ocean(box_config(n, m, l), mask.data());
    ocean.setPar("Combined Forcing", 0.05);
    ocean.setPar("Salinity Forcing", 1.0);
    ocean.setPar("Wind Forcing", 1.0);
    ocean.setPar("Temperature Forcing", 10.0);
    if (ocean.getPar("Combined Forcing") != 0.05) { std::printf("getPar mismatch\n"); return 1; }

    // trivial state, no forcing => F = 0   (src/tests/test_ocean.C:33-41)
    ocean.setPar("Combined Forcing", 0.0);
    ocean.computeRHS();
    const double f0 = ocean.getRHS('V')->norm();
    if (!(f0 < 1e-6)) { std::printf("trivial-state residual %g\n", f0); return 1; }
    ocean.setPar("Combined Forcing", 0.05);

    // directional derivative: (F(x + h d) - F(x)) / h ~ J d
    VectorPtr x = ocean.getState('V');
    const int64_t N = x->MyLength();
    std::vector<double> hx((size_t)N), hd((size_t)N);
    unsigned s = 12345u;
    auto rnd = [&]() { s = s * 1664525u + 1013904223u; return (double)(s >> 8) / 16777216.0 - 0.5; };
    for (int64_t q = 0; q < N; q++) { hx[q] = 0.02 * rnd(); hd[q] = rnd(); }
    for (int64_t c = 0; c < N / 6; c++) { hx[6 * c + 2] *= 1e-2; if (c >= (int64_t)n * m * (l - 1)) { hx[6 * c + 2] = 0; hd[6 * c + 2] = 0; } }
    x->fromHost(hx.data());
    ocean.computeRHS();
    DeviceVector F0(*ocean.getRHS('V'));
    ocean.computeJacobian();
    DeviceVector d(ocean.handle()), Jd(ocean.handle());
    d.fromHost(hd.data());
    ocean.applyMatrix(d, Jd);
    const double h = 1e-6;
    x->Update(h, d, 1.0);
    ocean.computeRHS();
    DeviceVector fd(*ocean.getRHS('V'));
    fd.update(-1.0 / h, F0, 1.0 / h);
    fd.update(-1.0, Jd, 1.0);
    const double rel = fd.norm() / Jd.norm();
    std::printf("directional derivative mismatch %.3e\n", rel);
    if (!(rel < 1e-3)) return 1;
    x->Update(-h, d, 1.0);

    // linear solve J y = J d  -> residual must drop monotonically
    ocean.setSolverParameters(1e-6, 200, 100, GMRES_B200_CGS2);
    VectorPtr b = std::make_shared<DeviceVector>(Jd);
    int rc = ocean.solve(b);
    DeviceVector r(ocean.handle());
    ocean.applyMatrix(*ocean.getSolution('V'), r);
    r.update(1.0, *b, -1.0);
    std::printf("GMRES rc=%d iters=%d true relative residual %.3e\n", rc, ocean.getNumIters(), r.norm() / b->norm());
    if (!(r.norm() / b->norm() < 0.9)) return 1;

    // Newton driver (control flow of src/newton/Newton.H) runs and does not increase the residual
    ocean.computeRHS();
    const double before = ocean.getNormRHS();
    Newton<OceanB200*> newton(&ocean);
    newton.SetMaxIterations(2);
    newton.Run();
    std::printf("Newton: |F| %.3e -> %.3e after %d iterations\n", before, newton.NormRHS(), newton.Iterations());
    if (!(newton.NormRHS() <= before * (1 + 1e-12))) return 1;

#ifdef WITH_REFERENCE_GMRES
    { // the reference's own template, unmodified, on the B200 model
      VectorPtr xs = std::make_shared<DeviceVector>(ocean.handle());
      GMRESSolver<OceanB200, VectorPtr> ref(ocean, xs, b);
      int rrc = ref.solve();
      std::printf("reference GMRESSolver on OceanB200: rc=%d iters=%d resid=%g\n", rrc, ref.getNumIters(), ref.residual());
    }
#endif
    { // coupled model, ocean side (row N2): attach an atmosphere and sea ice, blocks in the reference's container and on the device
      thcm_b200_config cc = box_config(n, m, l);
      cc.coupled_T = 1;
      OceanB200 oc(cc, mask.data());
      oc.setPar("Combined Forcing", 0.05); oc.setPar("Temperature Forcing", 10.0); oc.setPar("Solar Forcing", 1.0);
      oc.getState('V')->fromHost(hx.data());
      oc.computeRHS();
      const double f_alone = oc.getRHS('V')->norm();
      const size_t nm = (size_t)n * m;
      std::vector<double> Ta(nm), Qa(nm), Aa(nm), Ms(nm);
      for (size_t q = 0; q < nm; q++) { Ta[q] = 5.0 * rnd(); Qa[q] = 1e-3 * rnd(); Aa[q] = 0.5 + rnd(); Ms[q] = (q / n) >= (size_t)(m - 3) ? 1.0 : 0.0; }
      thcm_b200_atmos_pars ap{}; ap.tdim = 1.0; ap.qdim = 0.01; ap.nuq = 2.58e-2; ap.eta = 4.68e-2; ap.dqso = 5.6e-4; ap.Eo0 = 3.1e-8; ap.a0 = 0.3; ap.da = 0.5;
      thcm_b200_seaice_pars sp{}; sp.zeta = 1.7e3; sp.a0 = -0.0575; sp.Lf = 3.347e5; sp.s0 = 35.0; sp.rhoo = 1.024e3;
      oc.synchronizeAtmosphere(Ta.data(), Qa.data(), Aa.data(), ap);
      oc.synchronizeSeaIce(Ms.data(), sp);
      oc.computeRHS();
      const double f_coupled = oc.getRHS('V')->norm();
      if (!(f_coupled != f_alone) || !std::isfinite(f_coupled)) { std::printf("coupling left the residual unchanged (%g)\n", f_coupled); return 1; }
      auto A = oc.getAtmosBlock();
      auto S = oc.getSeaIceBlock();
      if (A->beg.size() != (size_t)N + 1 || A->beg.back() != (int)(3 * nm) || S->beg.back() != (int)nm) { std::printf("coupling block sizes %zu %d %d\n", A->beg.size(), A->beg.back(), S->beg.back()); return 1; }
      // J d + C_OA v on the device against the same sum on the host
      oc.computeJacobian();
      DeviceVector dd(oc.handle()), out(oc.handle());
      dd.fromHost(hd.data());
      oc.applyMatrix(dd, out);
      std::vector<double> base((size_t)N), got((size_t)N), v(3 * nm);
      out.toHost(base.data());
      for (auto& q : v) q = rnd();
      double* d_v = nullptr;
      check(iemic_b200_malloc((void**)&d_v, (int64_t)(v.size() * sizeof(double))));
      check(iemic_b200_memcpy_h2d(d_v, v.data(), (int64_t)(v.size() * sizeof(double))));
      OceanB200::CouplingBlock C(oc, false, (int64_t)(3 * nm));
      C.addTo(d_v, out);
      out.toHost(got.data());
      iemic_b200_free(d_v);
      double worst = 0.0;
      for (int64_t r = 0; r < N; r++) {
        double sum = 0.0;
        for (int p = A->beg[r]; p < A->beg[r + 1]; p++) sum += A->co[p] * v[A->jco[p]];
        double mag = std::fabs(base[r]);
        for (int p = A->beg[r]; p < A->beg[r + 1]; p++) mag += std::fabs(A->co[p] * v[A->jco[p]]);
        // (the device adds a row's entries in column order, this loop in the reference's push order: rounding only)
        worst = std::fmax(worst, std::fabs(got[r] - (base[r] + sum)) / (mag > 0 ? mag : 1.0));
      }
      std::printf("coupled: |F| %.6e -> %.6e, block product deviation %.3e\n", f_alone, f_coupled, worst);
      if (!(worst <= 1e-14)) return 1;
    }
    std::printf("selftest ok\n");
  } catch (const std::exception& e) {
    std::printf("selftest failed: %s\n", e.what());
    return 1;
  }
  return 0;
}
