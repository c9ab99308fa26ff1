// tests/native/riemann_host.cu -- TEST AID: runs the __host__ __device__ functions of
// minicfd_b200/csrc/riemann.cuh on the CPU, so that the device math (fast powers, fused
// value/derivative evaluation, single-loop FindRoot) can be checked against the oracle in the
// `-m "not gpu"` suite.  Never linked into the product.
#include <cstdint>

#include "riemann.cuh"

using namespace mcfd;

static Frame load_frame(const double *f) {
  Frame fr;
  for (int d = 0; d < 3; ++d) { fr.n[d] = f[d]; fr.t[d] = f[3 + d]; fr.b[d] = f[6 + d]; }
  return fr;
}

extern "C" int host_riemann_flux(int kind, double gamma, int64_t n, const double *frames, const double *ul,
                                 const double *ur, double *flux) {
  for (int64_t i = 0; i < n; ++i) {
    Frame fr = load_frame(frames + 9 * i);
    Cons l, r, f;
    for (int k = 0; k < 5; ++k) { l.c[k] = ul[5 * i + k]; r.c[k] = ur[5 * i + k]; }
    switch (kind) {
      case 0: f = rotated_flux_upwind<0>(fr, l, r, gamma); break;
      case 1: f = rotated_flux_upwind<1>(fr, l, r, gamma); break;
      case 2: f = rotated_flux_upwind<2>(fr, l, r, gamma); break;
      case 3: f = rotated_flux_upwind<3>(fr, l, r, gamma); break;
      default: return 1;
    }
    for (int k = 0; k < 5; ++k) flux[5 * i + k] = f.c[k];
  }
  return 0;
}

extern "C" int host_boundary_flux(int kind, double gamma, int64_t n, const double *frames, const double *ui,
                                  const double *uo, double *flux) {
  for (int64_t i = 0; i < n; ++i) {
    Frame fr = load_frame(frames + 9 * i);
    Cons a, b;
    for (int k = 0; k < 5; ++k) { a.c[k] = ui[5 * i + k]; b.c[k] = uo[5 * i + k]; }
    Cons f = boundary_flux(kind, fr, a, b, gamma);
    for (int k = 0; k < 5; ++k) flux[5 * i + k] = f.c[k];
  }
  return 0;
}

// iteration statistics of the fast exact solver (how many pressure-function evaluations a solve takes)
extern "C" int host_expansion_powers(int gamma14, int64_t n, const double *r, double e_val, double *t, double *s) {
  for (int64_t i = 0; i < n; ++i) {
    if (gamma14) expansion_powers<true>(r[i], e_val, t[i], s[i]);
    else expansion_powers<false>(r[i], e_val, t[i], s[i]);
  }
  return 0;
}

// the Markstein quotient used where the reference's IEEE division must be reproduced bit for bit
extern "C" int host_rn_div(int64_t n, const double *a, const double *b, double *q) {
  for (int64_t i = 0; i < n; ++i) q[i] = rn_div(a[i], b[i], rn_rcp(b[i]));
  return 0;
}
extern "C" int host_fast_ops(int64_t n, const double *x, double *rcp, double *rsq) {
  for (int64_t i = 0; i < n; ++i) { rcp[i] = fast_rcp(x[i]); rsq[i] = fast_rsqrt(x[i]); }
  return 0;
}
