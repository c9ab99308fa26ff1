// riemann.cuh -- device-side Euler gas relations, Riemann solvers in the face-normal frame and
// boundary fluxes.  FP64 throughout; everything lives in registers (no per-object state as in
// the reference's Riemann objects, SURVEY.md section 5 "not re-entrant").
//
// Replaces (paths relative to /root/reference/include/mini):
//   riemann/euler/types.hpp:197-290   IdealGas conversions, PrimitiveToFlux, GetFluxMatrix
//   riemann/euler/exact.hpp:49-343    exact Godunov flux
//   riemann/rotated/euler.hpp:80-193  frame rotation, GetFluxUpwind, the six boundary fluxes
// ROE / HLLC / AUSM: the reference's Hllc/Ausm do not compile through rotated::Euler in
// 3-D and Roe exists only in python/riemann.py:397-443.  AUSM follows Ausm<Gas,3> (ausm.hpp:153-217), HLLC is
// Hllc<Gas,2> (hllc.hpp:100-180) with w carried like v, Roe the Python form with the shear waves added; all
// three are pinned against those reference classes (tests/test_riemann_approx.py).
//
// Parity note: the exact solver's `while (f(p) > 0)` / `while (f(p) < -1e-8)` loops
// (exact.hpp:163-174) are discontinuous in their inputs: an iterate whose f lies within rounding of
// a threshold can take one Newton step more or fewer than on the CPU and then differ by up to the
// solver's own 1e-8 stopping tolerance.  One such test is degenerate in practice: when the data are
// continuous across a face (any smooth initial condition sampled at the nodes), f(p_L) is pure
// rounding noise and its SIGN decides between "halve p" and "keep p_L".  To follow the reference
// there, everything feeding the first evaluation of f (U^/det J, ConservativeToPrimitive, the frame
// rotation, f(p_L) itself) uses the reference's operation order without FMA contraction, IEEE
// division and sqrt, and a correctly rounded power when p_L/p_R is within 2^-20 of 1.  Everything
// after that differs from the CPU evaluation by a few ulp (FMA, rsqrt, the fast powers below).
#pragma once
#include <cuda_runtime.h>

#include <cmath>

// Every function here is __host__ __device__ so that tests/native/riemann_host.cu can run the very
// same source on the CPU against the oracle (a test aid: the product only ever calls them in kernels).
#define MCFD_HD __host__ __device__ __forceinline__
#define MCFD_HD_NOINLINE inline __host__ __device__ __noinline__

namespace mcfd {

// Operations that feed a sign test the reference takes on rounding noise (see the parity note
// above) are kept un-contracted and in the reference's order: __dmul_rn & co are never fused.
#ifdef __CUDA_ARCH__
#define MCFD_MUL(a, b) __dmul_rn((a), (b))
#define MCFD_ADD(a, b) __dadd_rn((a), (b))
#define MCFD_SUB(a, b) __dsub_rn((a), (b))
#else
#define MCFD_MUL(a, b) ((a) * (b))
#define MCFD_ADD(a, b) ((a) + (b))
#define MCFD_SUB(a, b) ((a) - (b))
#endif

// ---- branch-free FP64 reciprocal / division / rsqrt from the 20-bit SFU seeds (MUFU.RCP64H /
// MUFU.RSQ64H) and cubically convergent refinement.  CUDA's own `/`, sqrt() and rsqrt() carry
// special-case branches (denormals, overflow) that cost more than the arithmetic here; callers
// guarantee normal, positive arguments.
MCFD_HD double mcfd_fma(double a, double b, double c) {
#ifdef __CUDA_ARCH__
  return fma(a, b, c);
#else
  return std::fma(a, b, c);
#endif
}
MCFD_HD double rcp_seed(double b) {
#ifdef __CUDA_ARCH__
  double y; asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(b)); return y;
#else
  return double(1.0f / float(b));
#endif
}
MCFD_HD double rsqrt_seed(double x) {
#ifdef __CUDA_ARCH__
  double y; asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x)); return y;
#else
  return double(1.0f / std::sqrt(float(x)));
#endif
}
// 1/b to ~1 ulp: r <- r (1 + e + e^2), e = 1 - b r  (2^-20 -> 2^-60)
MCFD_HD double fast_rcp(double b) {
  double r = rcp_seed(b);
  double e = mcfd_fma(-b, r, 1.0);
  return mcfd_fma(r, mcfd_fma(e, e, e), r);
}
// a/b to ~1 ulp
MCFD_HD double fast_div(double a, double b) {
  double r = fast_rcp(b), q = a * r;
  return mcfd_fma(mcfd_fma(-q, b, a), r, q);
}
// 1/sqrt(x) to ~1 ulp: y <- y (1 + e/2 + 3 e^2/8), e = 1 - x y^2
MCFD_HD double fast_rsqrt(double x) {
  double y = rsqrt_seed(x);
  double e = mcfd_fma(-(x * y), y, 1.0);
  return mcfd_fma(y, e * mcfd_fma(0.375, e, 0.5), y);
}
// RN(1/b) (one more correction step; correctly rounded except for rare hard cases) and the
// Markstein sequence for RN(a/b): the quotients that feed a noise-level sign test (parity note).
MCFD_HD double rn_rcp(double b) {
  double r = fast_rcp(b);
  return mcfd_fma(mcfd_fma(-b, r, 1.0), r, r);
}
MCFD_HD double rn_div(double a, double b, double r) {
  double q = a * r;
  return mcfd_fma(mcfd_fma(-q, b, a), r, q);
}
MCFD_HD bool mid_range(double x) { return x > 1e-100 && x < 1e100; }  // safe for the seeds above

struct Prim { double rho, u, v, w, p; };
struct Cons { double c[5]; };
struct Frame { double n[3], t[3], b[3]; };

MCFD_HD void set_zero_if_negative(Prim &s) {  // types.hpp:197-204
  if (s.rho < 0) { s.rho = s.u = s.v = s.w = s.p = 0; }
  if (s.p < 0) { s.rho = s.u = s.v = s.w = s.p = 0; }
}

// types.hpp:246-258, straight transcription: the path of states with rho <= 0 or E < 0 (vacuum, garbage)
MCFD_HD_NOINLINE Prim cons_to_prim_plain(Cons c, double gamma) {
  Prim s{ c.c[0], c.c[1], c.c[2], c.c[3], c.c[4] };
  set_zero_if_negative(s);
  if (s.rho != 0) {
    s.u /= s.rho; s.v /= s.rho; s.w /= s.rho;
    double e_k = MCFD_ADD(MCFD_MUL(s.u, s.u), MCFD_ADD(MCFD_MUL(s.v, s.v), MCFD_MUL(s.w, s.w)));
    e_k = MCFD_MUL(e_k, 0.5);
    s.p = MCFD_SUB(s.p, MCFD_MUL(s.rho, e_k));
    s.p = MCFD_MUL(s.p, gamma - 1);
    set_zero_if_negative(s);
  }
  return s;
}

// The same conversion for an ordinary state (rho > 0, E >= 0), branch-free: RN quotients through one
// reciprocal, and the second SetZeroIfNegative (p < 0 zeroes the whole tuple) as selects.
MCFD_HD Prim cons_to_prim(const Cons &c, double gamma) {
  if (!(mid_range(c.c[0]) && c.c[4] >= 0)) return cons_to_prim_plain(c, gamma);
  Prim s;
  const double r = rn_rcp(c.c[0]);
  s.rho = c.c[0];
  s.u = rn_div(c.c[1], s.rho, r); s.v = rn_div(c.c[2], s.rho, r); s.w = rn_div(c.c[3], s.rho, r);
  double e_k = MCFD_ADD(MCFD_MUL(s.u, s.u), MCFD_ADD(MCFD_MUL(s.v, s.v), MCFD_MUL(s.w, s.w)));
  e_k = MCFD_MUL(e_k, 0.5);
  s.p = MCFD_MUL(MCFD_SUB(c.c[4], MCFD_MUL(s.rho, e_k)), gamma - 1);
  const bool neg = s.p < 0;
  s.rho = neg ? 0.0 : s.rho; s.u = neg ? 0.0 : s.u; s.v = neg ? 0.0 : s.v; s.w = neg ? 0.0 : s.w; s.p = neg ? 0.0 : s.p;
  return s;
}

MCFD_HD double speed_of_sound(const Prim &s, double gamma) {  // types.hpp:240-244
  return (s.rho <= 0 || s.p <= 0) ? 0.0 : sqrt(gamma * s.p / s.rho);
}

MCFD_HD Cons prim_to_flux(const Prim &s, double gamma) {  // types.hpp:259-278
  double e_k = s.u * s.u + (s.v * s.v + (s.w * s.w));
  e_k *= 0.5;
  double e = s.p * (1 / (gamma - 1));
  e += s.rho * e_k;
  Cons f;
  f.c[0] = s.rho * s.u;
  f.c[1] = (s.u * s.rho) * s.u + s.p;
  f.c[2] = (s.v * s.rho) * s.u;
  f.c[3] = (s.w * s.rho) * s.u;
  f.c[4] = e * s.u + s.p * s.u;
  return f;
}

// rotated/euler.hpp:80-97 : momentum/velocity components into and out of the frame (n, t, b)
MCFD_HD void global_to_normal(const Frame &fr, double &x, double &y, double &z) {
  double a = MCFD_ADD(MCFD_ADD(MCFD_MUL(x, fr.n[0]), MCFD_MUL(y, fr.n[1])), MCFD_MUL(z, fr.n[2]));
  double b = MCFD_ADD(MCFD_ADD(MCFD_MUL(x, fr.t[0]), MCFD_MUL(y, fr.t[1])), MCFD_MUL(z, fr.t[2]));
  double c = MCFD_ADD(MCFD_ADD(MCFD_MUL(x, fr.b[0]), MCFD_MUL(y, fr.b[1])), MCFD_MUL(z, fr.b[2]));
  x = a; y = b; z = c;
}
MCFD_HD void normal_to_global(const Frame &fr, double &a, double &b, double &c) {
  double x = a * fr.n[0] + b * fr.t[0] + c * fr.b[0];
  double y = a * fr.n[1] + b * fr.t[1] + c * fr.b[1];
  double z = a * fr.n[2] + b * fr.t[2] + c * fr.b[2];
  a = x; b = y; c = z;
}

// ---------------------------------------------------------------------------------------------
// exact solver, riemann/euler/exact.hpp
struct SpeedChange {  // exact.hpp:175-222
  double rho_before, p_before, a_before, p_const;
  double g_p1_o2, g_p1_o4, e_val, e_prime, gm1_u2;
  MCFD_HD void init(const Prim &before, double gamma) {
    rho_before = before.rho; p_before = before.p;
    a_before = before.rho == 0 ? 0.0 : speed_of_sound(before, gamma);
    p_const = before.p * ((gamma - 1) / 2);
    g_p1_o2 = (gamma + 1) / 2; g_p1_o4 = (gamma + 1) / 4;
    e_val = ((gamma - 1) / 2) / gamma; e_prime = -((gamma + 1) / 2) / gamma;
    gm1_u2 = 2 / (gamma - 1);
  }
  MCFD_HD double P(double p_after) const { return g_p1_o2 * p_after + p_const; }
  MCFD_HD double value(double p_after) const {
    double v;
    if (p_after >= p_before) {
      double divisor = sqrt(rho_before * P(p_after));
      v = (p_after - p_before) / divisor;
    } else {
      v = pow(p_after / p_before, e_val) - 1;
      v *= gm1_u2 * a_before;
    }
    return v;
  }
  MCFD_HD double prime(double p_after) const {
    double v;
    if (p_after >= p_before) {
      double p = P(p_after);
      v = g_p1_o4 * (p_before - p_after);
      double divisor = sqrt(rho_before * p);
      v = (1 + v / p) / divisor;
    } else {
      double divisor = rho_before * a_before;
      v = pow(p_after / p_before, e_prime) / divisor;
    }
    return v;
  }
};

MCFD_HD Prim prim_zero() { return Prim{ 0, 0, 0, 0, 0 }; }

MCFD_HD Prim inside_expansion(double gri_1, double gri_2, double gamma) {  // exact.hpp:275-287
  double r = (gamma - 1) / (gamma + 1);
  double a = r * gri_2;
  double a_square = a * a;
  double rho = pow(a_square / gri_1 * (1 / gamma), 1 / (gamma - 1));
  Prim axis = prim_zero();
  axis.rho = rho; axis.u = a; axis.p = a_square * rho * (1 / gamma);
  return axis;
}

template <bool kFast>
MCFD_HD double div_sel(double a, double b) { return kFast ? fast_div(a, b) : a / b; }

template <int kField, bool kFast = false>
MCFD_HD Prim near_shock(const Prim &before, Prim after) {  // exact.hpp:224-258
  double shock_u = before.u;
  double divisor = (after.u - before.u) * before.rho;
  shock_u += (divisor != 0 ? div_sel<kFast>(after.p - before.p, divisor) : before.u);
  bool after_shock = kField == 1 ? (shock_u < 0) : (shock_u > 0);
  if (after_shock) {
    after.rho = before.rho;
    double d = after.u - shock_u;
    after.rho *= div_sel<kFast>(before.u - shock_u, d);
    return after;
  }
  return before;
}

template <int kField, bool kFast = false>
MCFD_HD Prim near_expansion(const Prim &before, Prim after, double a_before, double gamma) {  // exact.hpp:259-314
  double gm1u2 = 2 / (gamma - 1), gm1o2 = (gamma - 1) / 2;
  double a_after = kField == 1 ? a_before + gm1o2 * (before.u - after.u) : a_before - gm1o2 * (before.u - after.u);
  bool axis_after = kField == 1 ? (after.u - a_after < 0) : (after.u + a_after > 0);
  bool axis_before = kField == 1 ? (before.u - a_before > 0) : (before.u + a_before < 0);
  if (axis_after) {
    double d = a_after * a_after;
    after.rho = div_sel<kFast>(gamma * after.p, d);
    return after;
  } else if (axis_before) {
    return before;
  }
  double gri_1 = before.p / pow(before.rho, gamma);
  double gri_2 = kField == 1 ? before.u + gm1u2 * a_before : before.u - gm1u2 * a_before;
  return inside_expansion(gri_1, gri_2, gamma);
}
template <int kField>
MCFD_HD Prim near_expansion(const Prim &before, Prim after, double gamma) {
  return near_expansion<kField, false>(before, after, speed_of_sound(before, gamma), gamma);
}

MCFD_HD_NOINLINE Prim near_vacuum(Prim left, Prim right, double gamma) {  // exact.hpp:315-342
  double gm1u2 = 2 / (gamma - 1);
  double left_a = speed_of_sound(left, gamma);
  if (left.u > left_a) return left;
  double right_a = speed_of_sound(right, gamma);
  if (right.u + right_a < 0) return right;
  double gri_1 = left.p / pow(left.rho, gamma);
  double gri_2 = left.u + left_a * gm1u2;
  if (gri_2 >= 0) return inside_expansion(gri_1, gri_2, gamma);
  gri_1 = right.p / pow(right.rho, gamma);
  gri_2 = right.u - right_a * gm1u2;
  if (gri_2 < 0) return inside_expansion(gri_1, gri_2, gamma);
  return prim_zero();
}

// one-sided / two-sided vacuum in the data, exact.hpp:87-126 (rare; kept out of line)
MCFD_HD_NOINLINE Prim data_vacuum(Prim left, Prim right, double gamma) {
  double gm1u2 = 2 / (gamma - 1);
  if (left.rho > 0) {  // only right vacuum
    double a_left = speed_of_sound(left, gamma);
    if (left.u >= a_left) return left;
    double u_star = left.u + a_left * gm1u2;
    if (u_star <= 0) return prim_zero();
    return inside_expansion(left.p / pow(left.rho, gamma), u_star, gamma);
  }
  if (right.rho > 0) {  // only left vacuum
    double a_right = speed_of_sound(right, gamma);
    if (right.u + a_right <= 0) return right;
    double u_star = right.u - a_right * gm1u2;
    if (u_star >= 0) return prim_zero();
    return inside_expansion(right.p / pow(right.rho, gamma), u_star, gamma);
  }
  return prim_zero();
}

// exact.hpp:66-160 : state on the t-axis, straight transcription (every pow / sqrt / division where the
// reference has one).  Kept out of line as the fallback of exact_primitive_upwind() below for inputs
// its fast path does not take (non-positive pressure, pressure ratios below 1e-30).
// `star_u` is the contact speed the reference keeps in its Riemann object (exact.hpp:64); it is
// rewritten only in the ordinary branch and read by PassiveScalars (exact.hpp:37-46, 69-76) after EVERY solve, so a
// vacuum branch upwinds the tangential velocities by the contact speed of the last ordinary solve of the same face
// point (0 for a fresh object).  `star_u_io` is that member: one slot per face point, kept by the caller across solves
// (nullptr = a fresh object each time, the standalone known-answer entry points).
MCFD_HD_NOINLINE Prim exact_primitive_upwind_plain(Prim left, Prim right, double gamma, double *star_u_io = nullptr) {
  Prim result;
  double star_u = star_u_io ? *star_u_io : 0.0;
  if (!(left.rho > 0 && right.rho > 0)) {
    result = data_vacuum(left, right, gamma);
  } else {
    double du = right.u - left.u;
    SpeedChange cl, cr;
    cl.init(left, gamma); cr.init(right, gamma);
    if (cl.value(0.0) + cr.value(0.0) + du < 0) {
      double x = left.p;  // FindRoot, exact.hpp:163-174
      double fx = cl.value(x) + cr.value(x) + du;
      while (fx > 0) { x *= 0.5; fx = cl.value(x) + cr.value(x) + du; }
      while (fx < -1e-8) {
        double divisor = cl.prime(x) + cr.prime(x);
        x -= fx / divisor;
        fx = cl.value(x) + cr.value(x) + du;
      }
      Prim star = prim_zero();
      star.p = x;
      star.u = 0.5 * (right.u + cr.value(x) + left.u - cl.value(x));
      star_u = star.u;
      if (star_u_io) *star_u_io = star_u;
      if (0 < star.u) {
        result = star.p >= left.p ? near_shock<1>(left, star) : near_expansion<1>(left, star, gamma);
      } else {
        result = star.p >= right.p ? near_shock<3>(right, star) : near_expansion<3>(right, star, gamma);
      }
    } else {
      result = near_vacuum(left, right, gamma);
    }
  }
  result.v = star_u > 0 ? left.v : right.v;  // exact.hpp:37-46 PassiveScalars<3>
  result.w = star_u > 0 ? left.w : right.w;
  return result;
}

// ---- fast path of the pressure function -------------------------------------------------------
// The velocity-change function of one side and its derivative (exact.hpp:184-214) are evaluated
// TOGETHER at each iterate: both expansion-branch powers share one root,
//     r^e_prime = r^e_val / r      (e_prime = e_val - 1),
// and for gamma = 1.4 (IdealGas<Scalar, 1.4>, the only gas the reference's demos instantiate,
// demo/euler/sourceless.hpp:30) e_val = 1/7, so z = r^(-1/7) is obtained from a 22-bit SFU seed and
// one cubically convergent step z <- z (1 + e/7 + 4 e^2/49), e = 1 - r z^7,
// giving r^(1/7) = r z^6 and r^(-6/7) = z^6 in ~12 FP64 instructions instead of two generic pow().
// The shock branch needs one rsqrt and no division: with q = 1 / sqrt(rho P),
//     value = (p - p0) q,   prime = (1 + g (p0 - p) / P) q = (1 + g (p0 - p) rho q^2) q.
// The ITERATION is the reference's (same start p_L, same halving, same Newton updates, same 1e-8
// stop); only roundings differ (a few ulp), which is what the 1e-10 parity bar allows.
struct SideFn {            // kept small: it is live across the whole iteration (registers, not FP64 issue
  double rho, p;           // slots, are what limits the face kernel); state before the wave
  double inv_p;            // 1 / p
  double inv_rho_a;        // 1 / (rho a);  a = (gamma p) / (rho a) is recomputed where needed
  MCFD_HD double a(double gamma) const { return (gamma * p) * inv_rho_a; }
};

MCFD_HD float mufu_lg2(float x) {
#ifdef __CUDA_ARCH__
  float y; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y;
#else
  return std::log2(x);
#endif
}
MCFD_HD float mufu_ex2(float x) {
#ifdef __CUDA_ARCH__
  float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y;
#else
  return std::exp2(x) * (1.0f + 1.5e-6f);  // host stand-in made as coarse as the SFU approximation chain
#endif
}

// t = r^e_val, s = r^(e_val - 1) for r in [1e-30, 1)
template <bool kGamma14>
MCFD_HD void expansion_powers(double r, double e_val, double &t, double &s) {
  if (kGamma14) {
    // z ~ r^(-1/7) from the SFU (relative error < 2^-19), then ONE cubically convergent step
    //   z <- z (1 + e/7 + 4 e^2/49),  e = 1 - r z^7     (error ~ 20 e^3 < 2^-52)
    double z = double(mufu_ex2(mufu_lg2(float(r)) * (-1.0f / 7.0f)));
    double z2 = z * z, z4 = z2 * z2, z7 = (z4 * z2) * z;
    double e = mcfd_fma(-r, z7, 1.0);
    z = mcfd_fma(z, e * mcfd_fma(4.0 / 49.0, e, 1.0 / 7.0), z);
    z2 = z * z;
    s = (z2 * z2) * z2;
    t = r * s;
  } else {
    t = pow(r, e_val);
    s = t / r;
  }
}

// MCFD_SIDE_SELECT: both branches evaluated and selected (no divergence between lanes whose sides are
// shocks and lanes whose sides are expansions; ~45 instructions per side and iterate).  Default: a real
// branch -- the iterates of the halving path stay below both pressures (expansions only), and the first
// evaluation at p_L is a shock on the left for every lane, so warps are mostly uniform and skip the other
// branch's work.  The selected values are the same operations either way.
template <bool kGamma14>
MCFD_HD void side_eval(const SideFn &f, double x, double gamma, double g_p1_o2, double g_p1_o4, double e_val,
                       double gm1_o2, double gm1_u2, double &val, double &der) {
#ifdef MCFD_SIDE_SELECT
  // shock, exact.hpp:186-190,200-206
  const double P = mcfd_fma(g_p1_o2, x, f.p * gm1_o2);
  const double q = fast_rsqrt(f.rho * P);
  const double dp = x - f.p;
  const double v_s = dp * q;
  const double d_s = mcfd_fma(-g_p1_o4 * dp, f.rho * (q * q), 1.0) * q;
  // expansion, exact.hpp:191-195,207-212 (ratio clamped so that the unused branch stays finite)
  double t, s;
  expansion_powers<kGamma14>(fmin(x * f.inv_p, 1.0), e_val, t, s);
  const double v_e = (t - 1.0) * (gm1_u2 * f.a(gamma));
  const double d_e = s * f.inv_rho_a;
  const bool shock = x >= f.p;
  val = shock ? v_s : v_e;
  der = shock ? d_s : d_e;
#else
  if (x >= f.p) {  // shock, exact.hpp:186-190,200-206
    const double P = mcfd_fma(g_p1_o2, x, f.p * gm1_o2);
    const double q = fast_rsqrt(f.rho * P);
    const double dp = x - f.p;
    val = dp * q;
    der = mcfd_fma(-g_p1_o4 * dp, f.rho * (q * q), 1.0) * q;
  } else {         // expansion, exact.hpp:191-195,207-212
    double t, s;
    expansion_powers<kGamma14>(fmin(x * f.inv_p, 1.0), e_val, t, s);
    val = (t - 1.0) * (gm1_u2 * f.a(gamma));
    der = s * f.inv_rho_a;
  }
#endif
}

// u_change_right(p_L) as the reference rounds it when p_L/p_R is within 2^-20 of 1 (exact.hpp:191-195):
// pow(1 + d, e) = 1 + e d (1 + (e-1)/2 d (1 + (e-2)/3 d)) + O(d^4), rounded ONCE at "1 + s" like a
// correctly rounded pow, then "- 1" exactly.  Elsewhere (shock side, or a ratio further from 1) the
// fast value is already faithful: there |f| is far above rounding noise or its numerator p_L - p_R is exact.
MCFD_HD double first_value_right(const SideFn &R, double x, double gamma, double fast) {
  const double d = x * R.inv_p - 1.0;
  if (x >= R.p || d < -0x1p-20) return fast;
  const double dd = rn_div(x, R.p, R.inv_p) - 1.0;  // RN(p_L / p_R) - 1, exact subtraction
  const double e = ((gamma - 1) / 2) / gamma;
  const double sres = e * dd * (1.0 + (e - 1.0) * 0.5 * dd * (1.0 + (e - 2.0) * (1.0 / 3.0) * dd));
  const double t = MCFD_ADD(1.0, sres);
  return MCFD_MUL(MCFD_SUB(t, 1.0), (2 / (gamma - 1)) * R.a(gamma));
}

// Core of exact.hpp:84-160 on (rho, u, p) only; the tangential velocities are passive (exact.hpp:37-46):
// `from_left` tells the caller which side's (v, w) the result carries.  left / right may hold anything in
// v, w.  In the kGamma14 instance gamma is the literal 1.4, so every derived constant is an immediate.
template <bool kGamma14>
MCFD_HD Prim exact_star_state(const Prim &left, const Prim &right, double gamma_arg, bool &from_left,
                              double *star_u_io = nullptr) {
  const double gamma = kGamma14 ? 1.4 : gamma_arg;
  from_left = false;
  // inputs outside the fast path's domain take the transcription above
  if (!(mid_range(left.rho) && mid_range(right.rho) && mid_range(left.p) && mid_range(right.p))) {
    Prim l = left, r = right;
    l.v = 1.0; l.w = 0.0; r.v = 0.0; r.w = 0.0;
    Prim res = exact_primitive_upwind_plain(l, r, gamma, star_u_io);
    from_left = res.v != 0.0;
    return res;
  }
  const double gm1_u2 = 2 / (gamma - 1), g_p1_o2 = (gamma + 1) / 2, g_p1_o4 = (gamma + 1) / 4;
  const double e_val = ((gamma - 1) / 2) / gamma, gm1_o2 = (gamma - 1) / 2;
  SideFn L, R;
  L.rho = left.rho; L.p = left.p; R.rho = right.rho; R.p = right.p;
  L.inv_p = rn_rcp(left.p); R.inv_p = rn_rcp(right.p);
  {
    double ql = fast_rsqrt(gamma * left.p * left.rho), qr = fast_rsqrt(gamma * right.p * right.rho);  // 1 / (rho a)
    L.inv_rho_a = ql; R.inv_rho_a = qr;
  }
  const double du = MCFD_SUB(right.u, left.u);
  Prim result;
  // f(0): pow(0, e_val) - 1 = -1 on both sides (exact.hpp:191-195)
  if ((-(gm1_u2 * L.a(gamma))) + (-(gm1_u2 * R.a(gamma))) + du < 0) {
#define MCFD_X_MIN (1e-30 * fmin(L.p, R.p))
    double x = left.p, vl, dl, vr, dr, fx;
    // FindRoot (exact.hpp:163-174) as ONE loop, so that lanes still halving and lanes already in
    // Newton's iteration run the same evaluation code together
    side_eval<kGamma14>(L, x, gamma, g_p1_o2, g_p1_o4, e_val, gm1_o2, gm1_u2, vl, dl);
    side_eval<kGamma14>(R, x, gamma, g_p1_o2, g_p1_o4, e_val, gm1_o2, gm1_u2, vr, dr);
    vr = first_value_right(R, x, gamma, vr);  // vl = 0 exactly at p_L
    fx = MCFD_ADD(MCFD_ADD(vl, vr), du);
    bool halving = fx > 0;
#ifndef MCFD_NO_SHARED_ROOT
    // Shared root (gamma = 1.4): while an iterate lies below BOTH pressures -- every iterate of the halving path on
    // smooth data -- both sides are expansions and their roots differ by a constant factor,
    //     (x / p_R)^(-1/7) = (x / p_L)^(-1/7) (p_L / p_R)^(-1/7),
    // so one SFU seed + cubic step serves both sides (13 of ~74 instructions per iterate).  The factor is formed once,
    // when the loop is entered.
    double root_ratio = 1.0;
    if (kGamma14 && (halving || fx < -1e-8)) {
      const double rr = L.p * R.inv_p;
      double z = double(mufu_ex2(mufu_lg2(float(rr)) * (-1.0f / 7.0f)));
      double z2 = z * z, z4 = z2 * z2, z7 = (z4 * z2) * z;
      double e = mcfd_fma(-rr, z7, 1.0);
      root_ratio = mcfd_fma(z, e * mcfd_fma(4.0 / 49.0, e, 1.0 / 7.0), z);
    }
#endif
    while (halving || fx < -1e-8) {
      x = halving ? x * 0.5 : x - fast_div(fx, dl + dr);
      if (!(x >= MCFD_X_MIN)) break;
#ifndef MCFD_NO_SHARED_ROOT
      if (kGamma14 && x < L.p && x < R.p && mid_range(L.p * R.inv_p)) {
        const double rl = x * L.inv_p, rr = x * R.inv_p;
        double z = double(mufu_ex2(mufu_lg2(float(rl)) * (-1.0f / 7.0f)));
        double z2 = z * z, z4 = z2 * z2, z7 = (z4 * z2) * z;
        const double e = mcfd_fma(-rl, z7, 1.0);
        z = mcfd_fma(z, e * mcfd_fma(4.0 / 49.0, e, 1.0 / 7.0), z);   // (x / p_L)^(-1/7)
        const double zr = z * root_ratio;                             // (x / p_R)^(-1/7)
        z2 = z * z;
        const double sl = (z2 * z2) * z2, tl = rl * sl;
        const double zr2 = zr * zr;
        const double sr = (zr2 * zr2) * zr2, tr = rr * sr;
        vl = (tl - 1.0) * (gm1_u2 * L.a(gamma)); dl = sl * L.inv_rho_a;
        vr = (tr - 1.0) * (gm1_u2 * R.a(gamma)); dr = sr * R.inv_rho_a;
      } else
#endif
      {
        side_eval<kGamma14>(L, x, gamma, g_p1_o2, g_p1_o4, e_val, gm1_o2, gm1_u2, vl, dl);
        side_eval<kGamma14>(R, x, gamma, g_p1_o2, g_p1_o4, e_val, gm1_o2, gm1_u2, vr, dr);
      }
      fx = vl + vr + du;
      halving = halving && fx > 0;
    }
    if (!(x >= MCFD_X_MIN)) {
#undef MCFD_X_MIN
      Prim l = left, r = right;
      l.v = 1.0; l.w = 0.0; r.v = 0.0; r.w = 0.0;
      Prim res = exact_primitive_upwind_plain(l, r, gamma, star_u_io);
      from_left = res.v != 0.0;
      return res;
    }
    Prim star = prim_zero();
    star.p = x;
    star.u = 0.5 * (right.u + vr + left.u - vl);
    from_left = star.u > 0;
    if (star_u_io) *star_u_io = star.u;  // exact.hpp:129 : the member the next vacuum solve of this point reads
    if (0 < star.u) {
      result = star.p >= left.p ? near_shock<1, true>(left, star) : near_expansion<1, true>(left, star, L.a(gamma), gamma);
    } else {
      result = star.p >= right.p ? near_shock<3, true>(right, star) : near_expansion<3, true>(right, star, R.a(gamma), gamma);
    }
  } else {
    result = near_vacuum(left, right, gamma);
    from_left = star_u_io ? *star_u_io > 0 : false;  // the last ordinary solve's contact speed (0: fresh object)
  }
  return result;
}

template <bool kGamma14>
MCFD_HD Prim exact_primitive_upwind(const Prim &left, const Prim &right, double gamma, double *star_u_io = nullptr) {
  bool from_left;
  Prim result = exact_star_state<kGamma14>(left, right, gamma, from_left, star_u_io);
  result.v = from_left ? left.v : right.v;  // exact.hpp:37-46 PassiveScalars<3>
  result.w = from_left ? left.w : right.w;
  return result;
}

MCFD_HD Cons exact_flux(const Prim &left, const Prim &right, double gamma, double *star_u_io = nullptr) {  // exact.hpp:78-80
  // warp-uniform choice: gamma is a kernel argument
  Prim s = (gamma == 1.4) ? exact_primitive_upwind<true>(left, right, gamma, star_u_io)
                          : exact_primitive_upwind<false>(left, right, gamma, star_u_io);
  return prim_to_flux(s, gamma);
}

// ---------------------------------------------------------------------------------------------
// New 3-D approximate solvers in the normal frame (left/right primitives already rotated).
MCFD_HD Cons prim_to_cons(const Prim &s, double gamma) {
  Cons c;
  c.c[0] = s.rho; c.c[1] = s.rho * s.u; c.c[2] = s.rho * s.v; c.c[3] = s.rho * s.w;
  c.c[4] = s.p / (gamma - 1) + 0.5 * s.rho * (s.u * s.u + s.v * s.v + s.w * s.w);
  return c;
}

// HLLC: riemann/euler/hllc.hpp:138-180 (PVRS pressure guess, q-factor wave speeds, star flux)
// with the second tangential velocity w carried like v.
MCFD_HD double hllc_q(double p_estimate, double p_k, double gamma) {
  if (p_estimate <= p_k) return 1.0;
  double temp = 1 + ((gamma + 1) / 2) * (p_estimate / p_k - 1) / gamma;
  return sqrt(temp);
}
MCFD_HD Cons hllc_star_flux(const Prim &state, double wave_k, double wave_star, double gamma) {
  Cons flux = prim_to_flux(state, gamma);
  double energy = state.p / (gamma - 1) + state.rho * (state.u * state.u + state.v * state.v + state.w * state.w) * 0.5;
  double temp = state.rho * (wave_k - state.u) / (wave_k - wave_star);
  Cons u_k = prim_to_cons(state, gamma), u_star_k;
  u_star_k.c[0] = temp;
  u_star_k.c[1] = wave_star * temp;
  u_star_k.c[2] = state.v * temp;
  u_star_k.c[3] = state.w * temp;
  u_star_k.c[4] = energy / state.rho + (wave_star - state.u) * (wave_star + state.p / (state.rho * (wave_k - state.u)));
  u_star_k.c[4] *= temp;
  for (int i = 0; i < 5; ++i) flux.c[i] += (u_star_k.c[i] - u_k.c[i]) * wave_k;
  return flux;
}
MCFD_HD Cons hllc_flux(const Prim &left, const Prim &right, double gamma) {
  double rho_average = (left.rho + right.rho) / 2;
  double a_left = speed_of_sound(left, gamma), a_right = speed_of_sound(right, gamma);
  double a_average = (a_left + a_right) / 2;
  double p_pvrs = (left.p + right.p - (right.u - left.u) * rho_average * a_average) / 2;
  double p_estimate = fmax(0.0, p_pvrs);
  double wave_left = left.u - a_left * hllc_q(p_estimate, left.p, gamma);
  double wave_right = right.u + a_right * hllc_q(p_estimate, right.p, gamma);
  double wave_star = (right.p - left.p + left.rho * left.u * (wave_left - left.u)
                      - right.rho * right.u * (wave_right - right.u))
                   / (left.rho * (wave_left - left.u) - right.rho * (wave_right - right.u));
  if (0.0 <= wave_left) return prim_to_flux(left, gamma);
  if (wave_right <= 0.0) return prim_to_flux(right, gamma);
  if (0.0 <= wave_star) return hllc_star_flux(left, wave_left, wave_star, gamma);
  return hllc_star_flux(right, wave_right, wave_star, gamma);
}

// Roe: python/riemann.py:402-440 (Roe.get_upwind_flux, the 1-D form the reference's authors use: Roe averages,
// three waves, NO entropy fix) carried to 3-D -- the shear waves (eigenvalue u) take the jumps of the two
// tangential velocities.  With v = w = 0 on both sides it is the reference's 1-D flux.
MCFD_HD Cons roe_flux(const Prim &l, const Prim &r, double gamma) {
  const Cons fl = prim_to_flux(l, gamma), fr = prim_to_flux(r, gamma);
  const double g_o_gm1 = gamma / (gamma - 1);
  const double hl = l.p / l.rho * g_o_gm1 + (l.u * l.u + l.v * l.v + l.w * l.w) / 2;  // primitive_to_total_enthalpy
  const double hr = r.p / r.rho * g_o_gm1 + (r.u * r.u + r.v * r.v + r.w * r.w) / 2;
  const double sl = sqrt(l.rho), sr = sqrt(r.rho), sum = sl + sr;
  const double rho = sl * sr;
  const double u = (sl * l.u + sr * r.u) / sum, v = (sl * l.v + sr * r.v) / sum, w = (sl * l.w + sr * r.w) / sum;
  const double ke = (u * u + v * v + w * w) / 2;
  const double h0 = (sl * hl + sr * hr) / sum;
  const double aa = (gamma - 1) * (h0 - ke);
  const double a = sqrt(aa), ua = u * a;
  const double dp = r.p - l.p, du = r.u - l.u, dv = r.v - l.v, dw = r.w - l.w;
  const double c1 = 0.5 * ((dp - rho * a * du) / (2 * aa)) * fabs(u - a);   // 1st wave
  const double c2 = 0.5 * ((r.rho - l.rho) - dp / aa) * fabs(u);            // 2nd wave (entropy)
  const double cs = 0.5 * fabs(u) * rho;                                    // shear waves
  const double c3 = 0.5 * ((dp + rho * a * du) / (2 * aa)) * fabs(u + a);   // 3rd wave
  Cons f;
  f.c[0] = (fl.c[0] + fr.c[0]) / 2 - c1 - c2 - c3;
  f.c[1] = (fl.c[1] + fr.c[1]) / 2 - c1 * (u - a) - c2 * u - c3 * (u + a);
  f.c[2] = (fl.c[2] + fr.c[2]) / 2 - c1 * v - c2 * v - cs * dv - c3 * v;
  f.c[3] = (fl.c[3] + fr.c[3]) / 2 - c1 * w - c2 * w - cs * dw - c3 * w;
  f.c[4] = (fl.c[4] + fr.c[4]) / 2 - c1 * (h0 - ua) - c2 * ke - cs * (v * dv + w * dw) - c3 * (h0 + ua);
  return f;
}

// AUSM: riemann/euler/ausm.hpp:166-217 (the reference's 3-D class; its GetFluxUpwind is
// non-const, so it cannot be reached through rotated::Euler there -- the formulas are the same).
MCFD_HD Cons ausm_split(const Prim &state, double gamma, bool positive) {
  double p_split = state.p;
  double a = speed_of_sound(state, gamma);
  double mach = state.u / a;
  double mach_split = mach;
  double h = a * a / (gamma - 1) + state.u * state.u * 0.5 + state.v * state.v * 0.5 + state.w * state.w * 0.5;
  if (mach >= -1 && mach <= 1) {
    if (positive) { mach_split = (mach + 1) * (mach + 1) * 0.25; p_split = state.p * (mach + 1) * 0.5; }
    else { mach_split = -(mach - 1) * (mach - 1) * 0.25; p_split = -state.p * (mach - 1) * 0.5; }
  } else if (positive ? (mach < -1) : (mach > 1)) {
    mach_split = 0.0; p_split = 0.0;
  }
  double temp = state.rho * a * mach_split;
  Cons flux{ { 1 * temp, state.u * temp, state.v * temp, state.w * temp, h * temp } };
  flux.c[1] += p_split;
  return flux;
}
MCFD_HD Cons ausm_flux(const Prim &l, const Prim &r, double gamma) {
  Cons fp = ausm_split(l, gamma, true), fm = ausm_split(r, gamma, false);
  for (int i = 0; i < 5; ++i) fp.c[i] += fm.c[i];
  return fp;
}

// ---------------------------------------------------------------------------------------------
// rotated/euler.hpp:111-120 : GetFluxUpwind in the frame of the face point.  kRiemann is an
// mcfd_riemann value.
template <int kRiemann>
MCFD_HD Cons rotated_flux_upwind(const Frame &fr, const Cons &ul, const Cons &ur, double gamma, double *star_u_io = nullptr) {
  Prim l = cons_to_prim(ul, gamma), r = cons_to_prim(ur, gamma);
  global_to_normal(fr, l.u, l.v, l.w);
  global_to_normal(fr, r.u, r.v, r.w);
  Cons f;
  if (kRiemann == 0) f = exact_flux(l, r, gamma, star_u_io);
  else if (kRiemann == 1) f = roe_flux(l, r, gamma);
  else if (kRiemann == 2) f = hllc_flux(l, r, gamma);
  else f = ausm_flux(l, r, gamma);
  normal_to_global(fr, f.c[1], f.c[2], f.c[3]);
  return f;
}

MCFD_HD Cons flux_of_state(const Frame &fr, const Cons &u, double gamma) {  // :104-110,129-135
  Prim s = cons_to_prim(u, gamma);
  global_to_normal(fr, s.u, s.v, s.w);
  Cons f = prim_to_flux(s, gamma);
  normal_to_global(fr, f.c[1], f.c[2], f.c[3]);
  return f;
}
MCFD_HD Cons flux_solid_wall(const Frame &fr, const Cons &u, double gamma) {  // :122-128
  Prim s = cons_to_prim(u, gamma);
  Cons f{ { 0, s.p, 0, 0, 0 } };
  normal_to_global(fr, f.c[1], f.c[2], f.c[3]);
  return f;
}
MCFD_HD Cons flux_subsonic_inlet(const Frame &fr, const Cons &ui, const Cons &uo, double gamma) {  // :136-155
  Prim pi = cons_to_prim(ui, gamma), po = cons_to_prim(uo, gamma), pb = po;
  double u_nu_o = po.u * fr.n[0] + po.v * fr.n[1] + po.w * fr.n[2];
  double u_nu_i = pi.u * fr.n[0] + pi.v * fr.n[1] + pi.w * fr.n[2];
  double u_nu_jump = u_nu_o - u_nu_i;
  double a_i = speed_of_sound(pi, gamma);
  double rho_a_i = pi.rho * (u_nu_o > 0 ? a_i : -a_i);
  pb.p = (pi.p + po.p + rho_a_i * u_nu_jump) * 0.5;
  double p_jump = po.p - pb.p;
  pb.rho -= p_jump / (a_i * a_i);
  double s = p_jump / rho_a_i;
  pb.u += s * fr.n[0]; pb.v += s * fr.n[1]; pb.w += s * fr.n[2];
  global_to_normal(fr, pb.u, pb.v, pb.w);
  Cons f = prim_to_flux(pb, gamma);
  normal_to_global(fr, f.c[1], f.c[2], f.c[3]);
  return f;
}
MCFD_HD Cons flux_subsonic_outlet(const Frame &fr, const Cons &ui, const Cons &uo, double gamma) {  // :156-172
  Prim pi = cons_to_prim(ui, gamma), po = cons_to_prim(uo, gamma), pb = pi;
  pb.p = po.p;
  double p_jump = pi.p - pb.p;
  double a_i = speed_of_sound(pi, gamma);
  pb.rho -= p_jump / (a_i * a_i);
  double u_nu_i = pi.u * fr.n[0] + pi.v * fr.n[1] + pi.w * fr.n[2];
  double rho_a_i = pi.rho * (u_nu_i > 0 ? a_i : -a_i);
  double s = p_jump / rho_a_i;
  pb.u += s * fr.n[0]; pb.v += s * fr.n[1]; pb.w += s * fr.n[2];
  global_to_normal(fr, pb.u, pb.v, pb.w);
  Cons f = prim_to_flux(pb, gamma);
  normal_to_global(fr, f.c[1], f.c[2], f.c[3]);
  return f;
}
// kind = mcfd_bc_kind; rotated/euler.hpp:173-193 for SMART
MCFD_HD Cons boundary_flux(int kind, const Frame &fr, const Cons &ui, const Cons &uo, double gamma) {
  if (kind == 5) {
    Prim s = cons_to_prim(ui, gamma);
    double a = speed_of_sound(s, gamma);
    double u_nu = s.u * fr.n[0] + s.v * fr.n[1] + s.w * fr.n[2];
    if (u_nu < 0) kind = (u_nu + a < 0) ? 1 : 3;
    else kind = (u_nu - a > 0) ? 2 : 4;
  }
  switch (kind) {
    case 0: return flux_solid_wall(fr, ui, gamma);
    case 1: return flux_of_state(fr, uo, gamma);
    case 2: return flux_of_state(fr, ui, gamma);
    case 3: return flux_subsonic_inlet(fr, ui, uo, gamma);
    default: return flux_subsonic_outlet(fr, ui, uo, gamma);
  }
}

}  // namespace mcfd
