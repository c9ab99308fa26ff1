/* oracle/euler_oracle.c -- TEST INFRASTRUCTURE ONLY: CPU restatement of the hot path.
 *
 * A plain-C restatement of miniCFD's Euler residual + Runge-Kutta path on hexahedra: nodal
 * (FR = spatial::fr::Lobatto, DGSEM = spatial::dg::Lobatto) and modal (spatial::dg::General /
 * WithLimiterAndSource with the weno::Lazy / weno::Eigen limiters, in modal_oracle.h), exact Riemann
 * solver, the six boundary kinds, RungeKutta<1..3>.  Every function cites the reference file:line it follows
 * (paths relative to /root/reference/include/mini).  It is the checker the CUDA path is
 * compared with on the GPU box (where /root/reference does not exist); it is itself pinned
 * against the reference's own templates (oracle/_ref, built by oracle/Makefile) in
 * tests/test_oracle_vs_reference.py and against the committed fixtures in tests/golden/.
 * Parity status: PINNED (max relative difference vs oracle/_ref recorded in DESIGN.md).
 *
 * Nothing under minicfd_b200/ may include, link or call this file.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/minicfd_b200.h"

#define K 5
#define QMAX 4
#define NMAX (QMAX * QMAX * QMAX)
#define FQMAX (QMAX * QMAX)

typedef struct { double v[K]; } Val;

/* ------------------------------------------------------------------------------------------ */
/* gauss/lobatto.hpp:25-97 : Lobatto points and weights, Q = 2, 3, 4                            */
static void lobatto(int q, double *x, double *w) {
  if (q == 2) { x[0] = -1; x[1] = 1; w[0] = 1; w[1] = 1; }
  else if (q == 3) { x[0] = -1; x[1] = 0; x[2] = 1; w[0] = 1.0 / 3; w[1] = 4.0 / 3; w[2] = 1.0 / 3; }
  else { x[0] = -1; x[1] = -sqrt(1.0 / 5); x[2] = sqrt(1.0 / 5); x[3] = 1;
         w[0] = 1.0 / 6; w[1] = 5.0 / 6; w[2] = 5.0 / 6; w[3] = 1.0 / 6; }
}

/* basis/lagrange.hpp (Line::GetDerivatives, 1st order): D[a][m] = l_m'(x_a) */
static void lagrange_derivatives(int q, const double *x, double D[QMAX][QMAX]) {
  for (int a = 0; a < q; ++a)
    for (int m = 0; m < q; ++m) {
      double sum = 0;
      for (int s = 0; s < q; ++s) {
        if (s == m) continue;
        double prod = 1.0 / (x[m] - x[s]);
        for (int r = 0; r < q; ++r) {
          if (r == m || r == s) continue;
          prod *= (x[a] - x[r]) / (x[m] - x[r]);
        }
        sum += prod;
      }
      D[a][m] = sum;
    }
}

/* basis/vincent.hpp:27-29,71-87 : g_right'(1) with c_next = p/(2p+1) (Huynh's lumping Lobatto) */
static double legendre(int n, double x) {
  if (n == 0) return 1; if (n == 1) return x;
  double p0 = 1, p1 = x;
  for (int k = 2; k <= n; ++k) { double p2 = ((2 * k - 1) * x * p1 - (k - 1) * p0) / k; p0 = p1; p1 = p2; }
  return p1;
}
static double vincent_right_derivative(int degree, double local) {
  double c_next = degree / (2.0 * degree + 1), c_prev = 1 - c_next;
  double d_prev = 0, d_curr = 0, d_next = 0;
  for (int k_curr = 0; k_curr <= degree; ++k_curr) {
    if (k_curr > 0) { d_prev = d_curr; d_curr = d_next; }
    int k_next = k_curr + 1;
    d_next = k_next * legendre(k_curr, local) + local * d_curr;
  }
  return 0.5 * (d_curr + c_prev * d_prev + c_next * d_next);
}

/* ------------------------------------------------------------------------------------------ */
/* riemann/euler/types.hpp:197-204,240-258 */
typedef struct { double rho, u, v, w, p; } Prim;

static void set_zero_if_negative(Prim *s) {
  if (s->rho < 0) { s->rho = s->u = s->v = s->w = s->p = 0; }
  if (s->p < 0) { s->rho = s->u = s->v = s->w = s->p = 0; }
}
static Prim cons_to_prim(const double *c, double gamma) {
  Prim s = { c[0], c[1], c[2], c[3], c[4] };
  set_zero_if_negative(&s);
  if (s.rho) {
    s.u /= s.rho; s.v /= s.rho; s.w /= s.rho;
    double e_k = s.u * s.u + (s.v * s.v + (s.w * s.w));
    e_k *= 0.5;
    s.p -= s.rho * e_k;
    s.p *= (gamma - 1);
    set_zero_if_negative(&s);
  }
  return s;
}
static double speed_of_sound(const Prim *s, double gamma) {
  return (s->rho <= 0 || s->p <= 0) ? 0 : sqrt(gamma * s->p / s->rho);
}
/* types.hpp:259-278 : PrimitiveToConservative, PrimitiveToFlux (normal-frame flux) */
static void prim_to_flux(const Prim *s, double gamma, double *f) {
  double e_k = s->u * s->u + (s->v * s->v + (s->w * s->w));
  e_k *= 0.5;
  double cons[K] = { s->rho, s->u * s->rho, s->v * s->rho, s->w * s->rho, 0 };
  cons[4] = s->p * (1 / (gamma - 1));
  cons[4] += s->rho * e_k;
  for (int k = 0; k < K; ++k) f[k] = cons[k] * s->u;
  f[1] += s->p;
  f[4] += s->p * s->u;
}
/* types.hpp:279-290 : GetFluxMatrix, column-major 5x3 : F[d][k] */
static void flux_matrix(const double *cv, double gamma, double F[3][K]) {
  Prim s = cons_to_prim(cv, gamma);
  double vel[3] = { s.u, s.v, s.w };
  double rho_h0 = cv[4] + s.p;
  for (int d = 0; d < 3; ++d) {
    F[d][0] = cv[1 + d];
    F[d][1] = cv[1] * vel[d];
    F[d][2] = cv[2] * vel[d];
    F[d][3] = cv[3] * vel[d];
    F[d][1 + d] += s.p;
    F[d][4] = rho_h0 * vel[d];
  }
}

/* ------------------------------------------------------------------------------------------ */
/* riemann/euler/exact.hpp */
typedef struct { double rho_before, p_before, a_before, p_const, gamma; } SpeedChange;

static SpeedChange speed_change_init(const Prim *before, double gamma) { /* exact.hpp:177-183 */
  SpeedChange s;
  s.rho_before = before->rho; s.p_before = before->p;
  s.a_before = before->rho == 0 ? 0 : speed_of_sound(before, gamma);
  s.p_const = before->p * ((gamma - 1) / 2);
  s.gamma = gamma;
  return s;
}
static double sc_P(const SpeedChange *s, double p_after) { return ((s->gamma + 1) / 2) * p_after + s->p_const; }
static double sc_value(const SpeedChange *s, double p_after) { /* exact.hpp:184-197 */
  double value;
  if (p_after >= s->p_before) {
    double divisor = sqrt(s->rho_before * sc_P(s, p_after));
    value = (p_after - s->p_before) / divisor;
  } else {
    double e = ((s->gamma - 1) / 2) / s->gamma;
    value = pow(p_after / s->p_before, e) - 1;
    value *= (2 / (s->gamma - 1)) * s->a_before;
  }
  return value;
}
static double sc_prime(const SpeedChange *s, double p_after) { /* exact.hpp:198-214 */
  double value;
  if (p_after >= s->p_before) {
    double p = sc_P(s, p_after);
    value = ((s->gamma + 1) / 4) * (s->p_before - p_after);
    double divisor = sqrt(s->rho_before * p);
    value = (1 + value / p) / divisor;
  } else {
    double e = -((s->gamma + 1) / 2) / s->gamma;
    double divisor = s->rho_before * s->a_before;
    value = pow(p_after / s->p_before, e) / divisor;
  }
  return value;
}
static Prim prim_zero(void) { Prim z = { 0, 0, 0, 0, 0 }; return z; }

static Prim inside_expansion(double gri_1, double gri_2, double gamma) { /* exact.hpp:275-287 */
  double r = (gamma - 1) / (gamma + 1);
  double a = r * gri_2;
  double a_square = a * a;
  double rho = pow(a_square / gri_1 * (1 / gamma), 1 / (gamma - 1));
  Prim axis = prim_zero();  /* v, w are overwritten by the caller (PassiveScalars) */
  axis.rho = rho; axis.u = a; axis.p = a_square * rho * (1 / gamma);
  return axis;
}
/* exact.hpp:224-258 : field = 1 or 3 */
static Prim near_shock(int field, const Prim *before, Prim *after) {
  double shock_u = before->u;
  double divisor = (after->u - before->u) * before->rho;
  shock_u += (divisor ? (after->p - before->p) / divisor : before->u);
  int after_shock = field == 1 ? (shock_u < 0) : (shock_u > 0);
  if (after_shock) {
    after->rho = before->rho;
    double d = after->u - shock_u;
    after->rho *= (before->u - shock_u) / d;
    return *after;
  }
  return *before;
}
/* exact.hpp:259-274,288-314 */
static Prim near_expansion(int field, const Prim *before, Prim *after, double gamma) {
  double a_before = speed_of_sound(before, gamma);
  double gri_1 = before->p / pow(before->rho, gamma);
  double gm1u2 = 2 / (gamma - 1), gm1o2 = (gamma - 1) / 2;
  double gri_2 = field == 1 ? before->u + gm1u2 * a_before : before->u - gm1u2 * a_before;
  double a_after = field == 1 ? a_before + gm1o2 * (before->u - after->u)
                              : a_before - gm1o2 * (before->u - after->u);
  int axis_after = field == 1 ? (after->u - a_after < 0) : (after->u + a_after > 0);
  int axis_before = field == 1 ? (before->u - a_before > 0) : (before->u + a_before < 0);
  if (axis_after) {
    double d = a_after * a_after;
    after->rho = gamma * after->p / d;
    return *after;
  } else if (axis_before) {
    return *before;
  }
  return inside_expansion(gri_1, gri_2, gamma);
}
static Prim near_vacuum(const Prim *left, const Prim *right, double gamma) { /* exact.hpp:315-342 */
  double gm1u2 = 2 / (gamma - 1);
  double left_a = speed_of_sound(left, gamma);
  if (left->u > left_a) return *left;
  double right_a = speed_of_sound(right, gamma);
  if (right->u + right_a < 0) return *right;
  double gri_1 = left->p / pow(left->rho, gamma);
  double gri_2 = left->u + left_a * gm1u2;
  if (gri_2 >= 0) return inside_expansion(gri_1, gri_2, gamma);
  gri_1 = right->p / pow(right->rho, gamma);
  gri_2 = right->u - right_a * gm1u2;
  if (gri_2 < 0) return inside_expansion(gri_1, gri_2, gamma);
  return prim_zero();
}
/* exact.hpp:84-160 (private GetPrimitiveUpwind); *star_u persists between calls on the same
 * Riemann object in the reference (exact.hpp:64,69-76) -- it is only rewritten in the ordinary
 * branch, so a vacuum case reuses the value of the previous solve at that face point. */
static Prim exact_primitive_upwind(const Prim *left, const Prim *right, double gamma, double *star_u) {
  Prim result;
  double gm1u2 = 2 / (gamma - 1);
  if (left->rho > 0) {
    if (right->rho > 0) {
    } else {
      double a_left = speed_of_sound(left, gamma);
      if (left->u >= a_left) return *left;
      double u_star = left->u + a_left * gm1u2;
      if (u_star <= 0) return prim_zero();
      double gri_1 = left->p / pow(left->rho, gamma);
      return inside_expansion(gri_1, u_star, gamma);
    }
  } else {
    if (right->rho > 0) {
      double a_right = speed_of_sound(right, gamma);
      if (right->u + a_right <= 0) return *right;
      double u_star = right->u - a_right * gm1u2;
      if (u_star >= 0) return prim_zero();
      double gri_1 = right->p / pow(right->rho, gamma);
      return inside_expansion(gri_1, u_star, gamma);
    }
    return prim_zero();
  }
  double du = right->u - left->u;
  SpeedChange cl = speed_change_init(left, gamma), cr = speed_change_init(right, gamma);
#define F_OF(p) (sc_value(&cl, (p)) + sc_value(&cr, (p)) + du)
#define FP_OF(p) (sc_prime(&cl, (p)) + sc_prime(&cr, (p)))
  if (F_OF(0.0) < 0) {
    Prim star = prim_zero();
    double x = left->p;  /* FindRoot, exact.hpp:163-174 */
    while (F_OF(x) > 0) x *= 0.5;
    while (F_OF(x) < -1e-8) { double divisor = FP_OF(x); x -= F_OF(x) / divisor; }
    star.p = x;
    star.u = 0.5 * (right->u + sc_value(&cr, star.p) + left->u - sc_value(&cl, star.p));
    *star_u = star.u;
    if (0 < star.u) {
      result = star.p >= left->p ? near_shock(1, left, &star) : near_expansion(1, left, &star, gamma);
    } else {
      result = star.p >= right->p ? near_shock(3, right, &star) : near_expansion(3, right, &star, gamma);
    }
  } else {
    result = near_vacuum(left, right, gamma);
  }
#undef F_OF
#undef FP_OF
  return result;
}
/* exact.hpp:37-46,66-80 : public GetPrimitiveUpwind + PassiveScalars<3> + GetFluxUpwind */
static void exact_flux_upwind(const Prim *left, const Prim *right, double gamma, double *star_u, double *flux) {
  Prim state = exact_primitive_upwind(left, right, gamma, star_u);
  state.v = *star_u > 0 ? left->v : right->v;
  state.w = *star_u > 0 ? left->w : right->w;
  prim_to_flux(&state, gamma, flux);
}

/* ------------------------------------------------------------------------------------------ */
/* Approximate solvers in the normal frame.  The reference has Ausm<Gas,3> (usable directly, not
 * through rotated::Euler), Hllc<Gas,1|2> and a 1-D Roe in python/riemann.py; the 3-D forms below carry the
 * second tangential velocity w exactly like v.  Pinned by tests/test_riemann_approx.py against the compiled
 * Ausm<Gas,3> / Hllc<Gas,2> (oracle/_ref) and fixtures generated from python/riemann.py. */
static void prim_to_cons(const Prim *s, double gamma, double *c) { /* types.hpp:259-268 */
  double e_k = s->u * s->u + (s->v * s->v + (s->w * s->w));
  e_k *= 0.5;
  c[0] = s->rho; c[1] = s->u * s->rho; c[2] = s->v * s->rho; c[3] = s->w * s->rho;
  c[4] = s->p * (1 / (gamma - 1));
  c[4] += s->rho * e_k;
}
/* riemann/euler/ausm.hpp:178-216 : GetPositiveFlux / GetNegativeFlux of Ausm<Gas,3> */
static void ausm_split(const Prim *s, double gamma, int positive, double *flux) {
  double p_split = s->p;
  double a = speed_of_sound(s, gamma);
  double mach = s->u / a;
  double mach_split = mach;
  double h = a * a / (gamma - 1) + s->u * s->u * 0.5 + s->v * s->v * 0.5 + s->w * s->w * 0.5;
  if (mach >= -1 && mach <= 1) {
    if (positive) { mach_split = (mach + 1) * (mach + 1) * 0.25; p_split = s->p * (mach + 1) * 0.5; }
    else { mach_split = -(mach - 1) * (mach - 1) * 0.25; p_split = -s->p * (mach - 1) * 0.5; }
  } else if (positive ? (mach < -1) : (mach > 1)) {
    mach_split = 0.0; p_split = 0.0;
  }
  double temp = s->rho * a * mach_split;
  flux[0] = 1 * temp; flux[1] = s->u * temp; flux[2] = s->v * temp; flux[3] = s->w * temp; flux[4] = h * temp;
  flux[1] += p_split;
}
static void ausm_flux_upwind(const Prim *l, const Prim *r, double gamma, double *flux) { /* ausm.hpp:166-171 */
  double fm[K];
  ausm_split(l, gamma, 1, flux);
  ausm_split(r, gamma, 0, fm);
  for (int k = 0; k < K; ++k) flux[k] += fm[k];
}
/* riemann/euler/hllc.hpp:138-178 (Hllc<Gas,2>) + w */
static double hllc_q(double p_estimate, double p_k, double gamma) { /* :154-161 */
  if (p_estimate <= p_k) return 1.0;
  double temp = 1 + ((gamma + 1) / 2) * (p_estimate / p_k - 1) / gamma;
  return sqrt(temp);
}
static void hllc_star_flux(const Prim *s, double wave_k, double wave_star, double gamma, double *flux) { /* :162-178 */
  prim_to_flux(s, gamma, flux);
  double energy = s->p / (gamma - 1) + s->rho * (s->u * s->u + s->v * s->v + s->w * s->w) * 0.5;
  double temp = s->rho * (wave_k - s->u) / (wave_k - wave_star);
  double u_k[K], u_star[K];
  prim_to_cons(s, gamma, u_k);
  u_star[0] = temp;
  u_star[1] = wave_star * temp;
  u_star[2] = s->v * temp;
  u_star[3] = s->w * temp;
  u_star[4] = energy / s->rho + (wave_star - s->u) * (wave_star + s->p / (s->rho * (wave_k - s->u)));
  u_star[4] *= temp;
  for (int k = 0; k < K; ++k) { double d = u_star[k] - u_k[k]; d *= wave_k; flux[k] += d; }
}
static void hllc_flux_upwind(const Prim *l, const Prim *r, double gamma, double *flux) { /* :112-126,138-153 */
  double rho_average = (l->rho + r->rho) / 2;
  double a_left = speed_of_sound(l, gamma), a_right = speed_of_sound(r, gamma);
  double a_average = (a_left + a_right) / 2;
  double p_pvrs = (l->p + r->p - (r->u - l->u) * rho_average * a_average) / 2;
  double p_estimate = p_pvrs > 0.0 ? p_pvrs : 0.0;
  double wave_left = l->u - a_left * hllc_q(p_estimate, l->p, gamma);
  double wave_right = r->u + a_right * hllc_q(p_estimate, r->p, gamma);
  double wave_star = (r->p - l->p + l->rho * l->u * (wave_left - l->u) - r->rho * r->u * (wave_right - r->u)) /
                     (l->rho * (wave_left - l->u) - r->rho * (wave_right - r->u));
  if (0.0 <= wave_left) prim_to_flux(l, gamma, flux);
  else if (wave_right <= 0.0) prim_to_flux(r, gamma, flux);
  else if (0.0 <= wave_star) hllc_star_flux(l, wave_left, wave_star, gamma, flux);
  else if (wave_star < 0.0) hllc_star_flux(r, wave_right, wave_star, gamma, flux);
  else for (int k = 0; k < K; ++k) flux[k] = NAN;  /* the reference leaves `flux` unset when wave_star is NaN */
}
/* python/riemann.py:402-440 (Roe.get_upwind_flux, 1-D, no entropy fix), extended to 3-D: the shear waves
 * (eigenvalue u) carry the jumps of the tangential velocities; with v = w = 0 on both sides it is the 1-D form. */
static void roe_flux_upwind(const Prim *l, const Prim *r, double gamma, double *flux) {
  double fl[K], fr[K];
  prim_to_flux(l, gamma, fl); prim_to_flux(r, gamma, fr);
  for (int k = 0; k < K; ++k) flux[k] = (fl[k] + fr[k]) / 2;
  double ke_l = (l->u * l->u + l->v * l->v + l->w * l->w) / 2, ke_r = (r->u * r->u + r->v * r->v + r->w * r->w) / 2;
  double h0_l = l->p / l->rho * (gamma / (gamma - 1)) + ke_l, h0_r = r->p / r->rho * (gamma / (gamma - 1)) + ke_r;
  double sl = sqrt(l->rho), sr = sqrt(r->rho), sum = sl + sr;
  double rho = sl * sr;
  double u = (sl * l->u + sr * r->u) / sum, v = (sl * l->v + sr * r->v) / sum, w = (sl * l->w + sr * r->w) / sum;
  double ke = (u * u + v * v + w * w) / 2;
  double h0 = (sl * h0_l + sr * h0_r) / sum;
  double aa = (gamma - 1) * (h0 - ke);
  double a = sqrt(aa), ua = u * a;
  double p_jump = r->p - l->p, u_jump = r->u - l->u, v_jump = r->v - l->v, w_jump = r->w - l->w;
  double alpha, lam;
  alpha = (p_jump - rho * a * u_jump) / (2 * aa); lam = fabs(u - a);   /* 1st wave */
  { double e[K] = { 1, u - a, v, w, h0 - ua }; for (int k = 0; k < K; ++k) flux[k] -= 0.5 * alpha * lam * e[k]; }
  alpha = (r->rho - l->rho) - p_jump / aa; lam = fabs(u);              /* 2nd wave (entropy) */
  { double e[K] = { 1, u, v, w, ke }; for (int k = 0; k < K; ++k) flux[k] -= 0.5 * alpha * lam * e[k]; }
  { double e[K] = { 0, 0, rho * v_jump, rho * w_jump, rho * (v * v_jump + w * w_jump) };  /* shear waves */
    for (int k = 0; k < K; ++k) flux[k] -= 0.5 * lam * e[k]; }
  alpha = (p_jump + rho * a * u_jump) / (2 * aa); lam = fabs(u + a);   /* 3rd wave */
  { double e[K] = { 1, u + a, v, w, h0 + ua }; for (int k = 0; k < K; ++k) flux[k] -= 0.5 * alpha * lam * e[k]; }
}

/* ------------------------------------------------------------------------------------------ */
/* riemann/rotated/euler.hpp : frame = {a (normal), b, c}, each a 3-vector, fr[3*axis + xyz] */
static void global_to_normal(const double *fr, double *m) { /* :80-88 on elements 1..3 */
  double a = m[0] * fr[0] + m[1] * fr[1] + m[2] * fr[2];
  double b = m[0] * fr[3] + m[1] * fr[4] + m[2] * fr[5];
  double c = m[0] * fr[6] + m[1] * fr[7] + m[2] * fr[8];
  m[0] = a; m[1] = b; m[2] = c;
}
static void normal_to_global(const double *fr, double *m) { /* :89-97 */
  double x = m[0] * fr[0] + m[1] * fr[3] + m[2] * fr[6];
  double y = m[0] * fr[1] + m[1] * fr[4] + m[2] * fr[7];
  double z = m[0] * fr[2] + m[1] * fr[5] + m[2] * fr[8];
  m[0] = x; m[1] = y; m[2] = z;
}
static void prim_rotate(const double *fr, Prim *s) { global_to_normal(fr, &s->u); }

static void rotated_flux_upwind_kind(int riemann, const double *fr, const double *ul, const double *ur, double gamma,
                                     double *star_u, double *flux) { /* :111-120 */
  Prim l = cons_to_prim(ul, gamma), r = cons_to_prim(ur, gamma);
  prim_rotate(fr, &l); prim_rotate(fr, &r);
  switch (riemann) {
    case MCFD_RIEMANN_ROE: roe_flux_upwind(&l, &r, gamma, flux); break;
    case MCFD_RIEMANN_HLLC: hllc_flux_upwind(&l, &r, gamma, flux); break;
    case MCFD_RIEMANN_AUSM: ausm_flux_upwind(&l, &r, gamma, flux); break;
    default: exact_flux_upwind(&l, &r, gamma, star_u, flux); break;
  }
  normal_to_global(fr, flux + 1);
}
static void rotated_flux_upwind(const double *fr, const double *ul, const double *ur, double gamma,
                                double *star_u, double *flux) {
  rotated_flux_upwind_kind(MCFD_RIEMANN_EXACT, fr, ul, ur, gamma, star_u, flux);
}
static void flux_of_state(const double *fr, const double *u, double gamma, double *flux) { /* :104-110,129-135 */
  Prim s = cons_to_prim(u, gamma);
  prim_rotate(fr, &s);
  prim_to_flux(&s, gamma, flux);
  normal_to_global(fr, flux + 1);
}
static void flux_solid_wall(const double *fr, const double *u, double gamma, double *flux) { /* :122-128 */
  Prim s = cons_to_prim(u, gamma);
  flux[0] = 0; flux[1] = s.p; flux[2] = 0; flux[3] = 0; flux[4] = 0;
  normal_to_global(fr, flux + 1);
}
static void flux_subsonic_inlet(const double *fr, const double *ui, const double *uo, double gamma, double *flux) {
  /* :136-155 */
  Prim pi = cons_to_prim(ui, gamma), po = cons_to_prim(uo, gamma), pb = po;
  double u_nu_o = po.u * fr[0] + po.v * fr[1] + po.w * fr[2];
  double u_nu_i = pi.u * fr[0] + pi.v * fr[1] + pi.w * fr[2];
  double u_nu_jump = u_nu_o - u_nu_i;
  double a_i = speed_of_sound(&pi, gamma);
  double rho_a_i = pi.rho * (u_nu_o > 0 ? a_i : -a_i);
  pb.p = (pi.p + po.p + rho_a_i * u_nu_jump) * 0.5;
  double p_jump = po.p - pb.p;
  pb.rho -= p_jump / (a_i * a_i);
  double s = p_jump / rho_a_i;
  pb.u += s * fr[0]; pb.v += s * fr[1]; pb.w += s * fr[2];
  prim_rotate(fr, &pb);
  prim_to_flux(&pb, gamma, flux);
  normal_to_global(fr, flux + 1);
}
static void flux_subsonic_outlet(const double *fr, const double *ui, const double *uo, double gamma, double *flux) {
  /* :156-172 */
  Prim pi = cons_to_prim(ui, gamma), po = cons_to_prim(uo, gamma), pb = pi;
  pb.p = po.p;
  double p_jump = pi.p - pb.p;
  double a_i = speed_of_sound(&pi, gamma);
  pb.rho -= p_jump / (a_i * a_i);
  double u_nu_i = pi.u * fr[0] + pi.v * fr[1] + pi.w * fr[2];
  double rho_a_i = pi.rho * (u_nu_i > 0 ? a_i : -a_i);
  double s = p_jump / rho_a_i;
  pb.u += s * fr[0]; pb.v += s * fr[1]; pb.w += s * fr[2];
  prim_rotate(fr, &pb);
  prim_to_flux(&pb, gamma, flux);
  normal_to_global(fr, flux + 1);
}
static void flux_smart(const double *fr, const double *ui, const double *uo, double gamma, double *flux) {
  /* :173-193 */
  Prim s = cons_to_prim(ui, gamma);
  double a = speed_of_sound(&s, gamma);
  double u_nu = s.u * fr[0] + s.v * fr[1] + s.w * fr[2];
  if (u_nu < 0) {
    if (u_nu + a < 0) flux_of_state(fr, uo, gamma, flux);
    else flux_subsonic_inlet(fr, ui, uo, gamma, flux);
  } else {
    if (u_nu - a > 0) flux_of_state(fr, ui, gamma, flux);
    else flux_subsonic_outlet(fr, ui, uo, gamma, flux);
  }
}
static void boundary_flux(int kind, const double *fr, const double *ui, const double *uo, double gamma, double *flux) {
  switch (kind) {
    case MCFD_BC_SOLID_WALL: flux_solid_wall(fr, ui, gamma, flux); break;
    case MCFD_BC_SUPERSONIC_INLET: flux_of_state(fr, uo, gamma, flux); break;
    case MCFD_BC_SUPERSONIC_OUTLET: flux_of_state(fr, ui, gamma, flux); break;
    case MCFD_BC_SUBSONIC_INLET: flux_subsonic_inlet(fr, ui, uo, gamma, flux); break;
    case MCFD_BC_SUBSONIC_OUTLET: flux_subsonic_outlet(fr, ui, uo, gamma, flux); break;
    default: flux_smart(fr, ui, uo, gamma, flux); break;
  }
}

/* ------------------------------------------------------------------------------------------ */
/* Given-state functions (inputs of the comparison; same closed forms as oracle/ref/ref_functions.hpp
 * and the product's device code). */
typedef struct { int func; double p[MCFD_MAX_FUNC_PARAMS]; } StateFunc;

static void eval_state(const StateFunc *f, const double *xyz, double t, double *val) {
  const double *p = f->p;
  switch (f->func) {
    case MCFD_FUNC_CONSTANT:
      for (int k = 0; k < K; ++k) val[k] = p[k];
      break;
    case MCFD_FUNC_VORTEX: {
      const double kPi = 3.14159265358979323846;
      double beta = p[0], gamma = p[5];
      double x = xyz[0] - p[1] - p[3] * t, y = xyz[1] - p[2] - p[4] * t;
      double r2 = x * x + y * y;
      double e = exp(0.5 * (1.0 - r2));
      double c = beta / (2.0 * kPi) * e;
      double u = p[3] - c * y, v = p[4] + c * x;
      double temp = 1.0 - (gamma - 1.0) * beta * beta / (8.0 * gamma * kPi * kPi) * e * e;
      double rho = pow(temp, 1.0 / (gamma - 1.0));
      double pres = rho * temp;
      val[0] = rho; val[1] = rho * u; val[2] = rho * v; val[3] = 0.0;
      val[4] = pres / (gamma - 1.0) + 0.5 * rho * (u * u + v * v);
      break;
    }
    case MCFD_FUNC_PLANE_SPLIT: {
      double s = p[0] * xyz[0] + p[1] * xyz[1] + p[2] * xyz[2];
      const double *src = (s < p[3] + p[4] * t) ? &p[5] : &p[10];
      for (int k = 0; k < K; ++k) val[k] = src[k];
      break;
    }
    default:
      for (int k = 0; k < K; ++k) val[k] = 0;
  }
}

/* ------------------------------------------------------------------------------------------ */
/* Per-part tables: what Part + polynomial::Hexahedron + gauss::Hexahedron/Quadrangle +
 * fr::Lobatto::CacheCorrectionGradients / dg::Lobatto::MatchGaussianPoints cache. */
typedef struct {
  int64_t n_local, n_ghost, n_faces, n_local_faces, n_ghost_faces, n_boundary_faces;
  double *xyz;    /* [cell][node][3]     gauss/element.hpp:87-100 global coords                 */
  double *det;    /* [cell][node]        polynomial/hexahedron.hpp:172-173 det J                */
  double *jstar;  /* [cell][node][3][3]  hexahedron.hpp:174 det*J^{-1}, [global b][local a]     */
  double *wglob;  /* [cell][node]        global quadrature weight = w_local * |det J|           */
  int64_t *holder, *sharer;
  int *f_group;   /* [n_boundary_faces] */
  double *f_xyz;  /* [face][fq][3]  */
  double *f_frame;/* [face][fq][9]   geometry/face.hpp:98-110                                    */
  double *f_w;    /* [face][fq]      global face weight                                          */
  int *f_hnode, *f_snode;  /* [face][fq] collocated solution node of holder / sharer            */
  int *f_hside, *f_sside;  /* [face] local face id 0..5 (hexahedron.hpp:580-594)                 */
  double *f_star_u;        /* [face][fq] Exact::star_u, persistent per face point (exact.hpp:64) */
  double *u;      /* [(n_local + n_ghost)][N][K] coefficients (column layout per cell)          */
  /* halo */
  int n_nbr; int *nbr_rank; int64_t *send_ptr, *send_cells, *recv_ptr;
  /* modal DG (modal_oracle.h): basis centre, lower-triangular orthonormalisation matrix, volume per cell;
   * neighbours through interior faces in mesh order (part.hpp:1019-1032,1066-1078) */
  int64_t n_inner, n_troubled;
  double *m_center, *m_S, *m_volume;
  int64_t *adj_ptr, *adj_cell, *adj_face;
} PartTab;

typedef struct oracle_sim {
  int scheme, degree, rk_order, q, n, fq, n_parts;
  int nq;       /* quadrature points per cell: q^3 (= n for the nodal schemes; n = Taylor basis size for modal DG) */
  int limiter;  /* mcfd_limiter, modal DG only */
  int elem;     /* mcfd_cell_type of every cell: 0 HEXA_8, 1 TETRA_4 (modal DG only) */
  int riemann;  /* mcfd_riemann: which solver the interior / ghost faces use (boundaries never use one) */
  double weno_w0, weno_eps;
  double gamma;
  double x1[QMAX], w1[QMAX], D[QMAX][QMAX], g_prime;
  PartTab *parts;
  int n_groups;
  int *group_kind;      /* -1 = not set */
  StateFunc *group_func;
  double t_curr;
} oracle_sim;

/* geometry/hexahedron.hpp:345-352 local coords of the 8 corners; :371-380 face table */
static const double kCorner[8][3] = { {-1,-1,-1},{1,-1,-1},{1,1,-1},{-1,1,-1},{-1,-1,1},{1,-1,1},{1,1,1},{-1,1,1} };
static const int kFaceTable[6][4] = { {0,3,2,1}, {0,1,5,4}, {1,2,6,5}, {2,3,7,6}, {0,4,7,3}, {4,5,6,7} };

static void hexa_map(const double xn[8][3], const double *loc, double *glob, double J[3][3]) {
  /* geometry/hexahedron.hpp:419-455 : shape functions, gradients; cell.hpp LocalToJacobian:
   * J[a][b] = sum_i dN_i/dxi_a * x_i[b] */
  for (int b = 0; b < 3; ++b) glob[b] = 0;
  memset(J, 0, 9 * sizeof(double));
  for (int i = 0; i < 8; ++i) {
    double fx = 1 + kCorner[i][0] * loc[0], fy = 1 + kCorner[i][1] * loc[1], fz = 1 + kCorner[i][2] * loc[2];
    double N = fx * fy * fz / 8;
    double g[3] = { kCorner[i][0] * (fy * fz / 8), kCorner[i][1] * (fz * fx / 8), kCorner[i][2] * (fx * fy / 8) };
    for (int b = 0; b < 3; ++b) {
      glob[b] += xn[i][b] * N;
      for (int a = 0; a < 3; ++a) J[a][b] += g[a] * xn[i][b];
    }
  }
}
static double det3(const double J[3][3]) {
  return J[0][0] * (J[1][1] * J[2][2] - J[1][2] * J[2][1]) - J[0][1] * (J[1][0] * J[2][2] - J[1][2] * J[2][0])
       + J[0][2] * (J[1][0] * J[2][1] - J[1][1] * J[2][0]);
}
static void adjugate3(const double J[3][3], double A[3][3]) { /* A = det(J) * J^{-1} */
  A[0][0] = J[1][1] * J[2][2] - J[1][2] * J[2][1];
  A[0][1] = J[0][2] * J[2][1] - J[0][1] * J[2][2];
  A[0][2] = J[0][1] * J[1][2] - J[0][2] * J[1][1];
  A[1][0] = J[1][2] * J[2][0] - J[1][0] * J[2][2];
  A[1][1] = J[0][0] * J[2][2] - J[0][2] * J[2][0];
  A[1][2] = J[0][2] * J[1][0] - J[0][0] * J[1][2];
  A[2][0] = J[1][0] * J[2][1] - J[1][1] * J[2][0];
  A[2][1] = J[0][1] * J[2][0] - J[0][0] * J[2][1];
  A[2][2] = J[0][0] * J[1][1] - J[0][1] * J[1][0];
}

static int find_side(const int64_t *cell_nodes, const int64_t *face_nodes) {
  for (int s = 0; s < 6; ++s) {
    int cnt = 0;
    for (int v = 0; v < 4; ++v)
      for (int w = 0; w < 4; ++w)
        if (cell_nodes[kFaceTable[s][v]] == face_nodes[w]) ++cnt;
    if (cnt == 4) return s;
  }
  return -1;
}
/* the solution node of `cell` collocated with a point (dg/lobatto.hpp:55-75, fr/lobatto.hpp:95-100) */
static int match_node(const PartTab *pt, int n, int64_t cell, const double *x, double tol) {
  int best = -1; double best_d = 1e300;
  for (int h = 0; h < n; ++h) {
    const double *y = pt->xyz + (cell * n + h) * 3;
    double d = sqrt((x[0] - y[0]) * (x[0] - y[0]) + (x[1] - y[1]) * (x[1] - y[1]) + (x[2] - y[2]) * (x[2] - y[2]));
    if (d < best_d) { best_d = d; best = h; }
  }
  return best_d < tol ? best : -1;
}

static int modal_build(oracle_sim *sim, PartTab *pt, const mcfd_mesh *m);
static int build_part_tetra(oracle_sim *sim, PartTab *pt, const mcfd_mesh *m);
static int build_part(oracle_sim *sim, PartTab *pt, const mcfd_mesh *m) {
  if (sim->elem == MCFD_CELL_TETRA_4) return build_part_tetra(sim, pt, m);
  int q = sim->q, n = sim->nq, fq = sim->fq;  /* n: quadrature points */
  memset(pt, 0, sizeof(*pt));
  pt->n_local = m->n_local_cells; pt->n_ghost = m->n_ghost_cells; pt->n_inner = m->n_inner_cells;
  pt->n_local_faces = m->n_local_faces; pt->n_ghost_faces = m->n_ghost_faces; pt->n_boundary_faces = m->n_boundary_faces;
  pt->n_faces = m->n_local_faces + m->n_ghost_faces + m->n_boundary_faces;
  int64_t nc = pt->n_local + pt->n_ghost;
  pt->xyz = malloc(sizeof(double) * nc * n * 3);
  pt->det = malloc(sizeof(double) * nc * n);
  pt->jstar = malloc(sizeof(double) * nc * n * 9);
  pt->wglob = malloc(sizeof(double) * nc * n);
  pt->u = calloc(nc * sim->n * K, sizeof(double));
  double h_min = 1e300;
  for (int64_t c = 0; c < nc; ++c) {
    double xn[8][3];
    for (int v = 0; v < 8; ++v) memcpy(xn[v], m->node_xyz + 3 * m->cell_nodes[8 * c + v], 3 * sizeof(double));
    for (int v = 1; v < 8; ++v) {
      double d = sqrt(pow(xn[v][0] - xn[0][0], 2) + pow(xn[v][1] - xn[0][1], 2) + pow(xn[v][2] - xn[0][2], 2));
      if (d < h_min) h_min = d;
    }
    for (int i = 0; i < q; ++i) for (int j = 0; j < q; ++j) for (int k = 0; k < q; ++k) {
      int ijk = (i * q + j) * q + k;  /* basis/lagrange.hpp:236-238, gauss/hexahedron.hpp:66-80 */
      double loc[3] = { sim->x1[i], sim->x1[j], sim->x1[k] }, J[3][3], A[3][3];
      hexa_map(xn, loc, pt->xyz + (c * n + ijk) * 3, J);
      double det = det3(J);
      adjugate3(J, A);
      pt->det[c * n + ijk] = det;
      memcpy(pt->jstar + (c * n + ijk) * 9, A, sizeof(A));
      pt->wglob[c * n + ijk] = sim->w1[i] * sim->w1[j] * sim->w1[k] * fabs(det);
    }
  }
  double tol = 1e-8 * h_min;
  int64_t nf = pt->n_faces;
  pt->holder = malloc(sizeof(int64_t) * nf); pt->sharer = malloc(sizeof(int64_t) * nf);
  pt->f_group = malloc(sizeof(int) * (pt->n_boundary_faces + 1));
  pt->f_xyz = malloc(sizeof(double) * nf * fq * 3);
  pt->f_frame = malloc(sizeof(double) * nf * fq * 9);
  pt->f_w = malloc(sizeof(double) * nf * fq);
  pt->f_hnode = malloc(sizeof(int) * nf * fq); pt->f_snode = malloc(sizeof(int) * nf * fq);
  pt->f_hside = malloc(sizeof(int) * nf); pt->f_sside = malloc(sizeof(int) * nf);
  pt->f_star_u = calloc(nf * fq, sizeof(double));
  for (int64_t b = 0; b < pt->n_boundary_faces; ++b) pt->f_group[b] = m->bface_group[b];
  for (int64_t f = 0; f < nf; ++f) {
    int64_t h = m->face_holder[f], s = m->face_sharer[f];
    pt->holder[f] = h; pt->sharer[f] = s;
    int side = find_side(m->cell_nodes + 8 * h, m->face_nodes + 4 * f);
    if (side < 0) return 1;
    pt->f_hside[f] = side;
    pt->f_sside[f] = s >= 0 ? find_side(m->cell_nodes + 8 * s, m->face_nodes + 4 * f) : -1;
    if (s >= 0 && pt->f_sside[f] < 0) return 1;
    /* geometry/hexahedron.hpp:406-416 SortNodesOnFace: face nodes in the holder's table order */
    double xf[4][3];
    for (int v = 0; v < 4; ++v)
      memcpy(xf[v], m->node_xyz + 3 * m->cell_nodes[8 * h + kFaceTable[side][v]], 3 * sizeof(double));
    static const double kQuad[4][2] = { {-1,-1},{1,-1},{1,1},{-1,1} };  /* geometry/quadrangle.hpp */
    for (int i = 0; i < q; ++i) for (int j = 0; j < q; ++j) {
      int p = i * q + j;  /* gauss/quadrangle.hpp:62-73 */
      double lx = sim->x1[i], ly = sim->x1[j];
      double g[3] = { 0, 0, 0 }, Jf[2][3] = { { 0, 0, 0 }, { 0, 0, 0 } };
      for (int v = 0; v < 4; ++v) {
        double N = (1 + kQuad[v][0] * lx) * (1 + kQuad[v][1] * ly) / 4;
        double dx = kQuad[v][0] * (1 + kQuad[v][1] * ly) / 4, dy = kQuad[v][1] * (1 + kQuad[v][0] * lx) / 4;
        for (int b = 0; b < 3; ++b) { g[b] += xf[v][b] * N; Jf[0][b] += dx * xf[v][b]; Jf[1][b] += dy * xf[v][b]; }
      }
      double *fr = pt->f_frame + (f * fq + p) * 9;
      double nrm[3] = { Jf[0][1] * Jf[1][2] - Jf[0][2] * Jf[1][1], Jf[0][2] * Jf[1][0] - Jf[0][0] * Jf[1][2],
                        Jf[0][0] * Jf[1][1] - Jf[0][1] * Jf[1][0] };
      double ln = sqrt(nrm[0] * nrm[0] + nrm[1] * nrm[1] + nrm[2] * nrm[2]);
      double lt = sqrt(Jf[0][0] * Jf[0][0] + Jf[0][1] * Jf[0][1] + Jf[0][2] * Jf[0][2]);
      for (int b = 0; b < 3; ++b) { fr[b] = nrm[b] / ln; fr[3 + b] = Jf[0][b] / lt; }
      fr[6] = fr[1] * fr[5] - fr[2] * fr[4]; fr[7] = fr[2] * fr[3] - fr[0] * fr[5]; fr[8] = fr[0] * fr[4] - fr[1] * fr[3];
      /* gauss/element.hpp:92-96 : sqrt(det(J J^T)) */
      double a11 = lt * lt, a22 = Jf[1][0] * Jf[1][0] + Jf[1][1] * Jf[1][1] + Jf[1][2] * Jf[1][2];
      double a12 = Jf[0][0] * Jf[1][0] + Jf[0][1] * Jf[1][1] + Jf[0][2] * Jf[1][2];
      pt->f_w[f * fq + p] = sim->w1[i] * sim->w1[j] * sqrt(a11 * a22 - a12 * a12);
      memcpy(pt->f_xyz + (f * fq + p) * 3, g, sizeof(g));
      if (sim->scheme == MCFD_SCHEME_DG_MODAL) continue;  /* no collocated nodes */
      pt->f_hnode[f * fq + p] = match_node(pt, n, h, g, tol);
      pt->f_snode[f * fq + p] = s >= 0 ? match_node(pt, n, s, g, tol) : -1;
      if (pt->f_hnode[f * fq + p] < 0 || (s >= 0 && pt->f_snode[f * fq + p] < 0)) return 2;
    }
  }
  pt->n_nbr = m->n_neighbours;
  if (pt->n_nbr) {
    pt->nbr_rank = malloc(sizeof(int) * pt->n_nbr);
    pt->send_ptr = malloc(sizeof(int64_t) * (pt->n_nbr + 1)); pt->recv_ptr = malloc(sizeof(int64_t) * (pt->n_nbr + 1));
    memcpy(pt->send_ptr, m->send_ptr, sizeof(int64_t) * (pt->n_nbr + 1));
    memcpy(pt->recv_ptr, m->recv_ptr, sizeof(int64_t) * (pt->n_nbr + 1));
    for (int i = 0; i < pt->n_nbr; ++i) pt->nbr_rank[i] = m->nbr_rank[i];
    int64_t ns = pt->send_ptr[pt->n_nbr];
    pt->send_cells = malloc(sizeof(int64_t) * (ns + 1));
    memcpy(pt->send_cells, m->send_cells, sizeof(int64_t) * ns);
  }
  if (sim->scheme == MCFD_SCHEME_DG_MODAL) return modal_build(sim, pt, m);
  return 0;
}


/* ------------------------------------------------------------------------------------------ */
/* TETRA_4 cells (modal DG only): geometry/tetrahedron.hpp:96-150, gauss/tetrahedron.hpp:112-310,
 * geometry/triangle.hpp:78-112, gauss/triangle.hpp:122-246, gauss/element.hpp:87-100            */
static const int kTetFace[4][3] = { {0,2,1}, {0,1,3}, {0,3,2}, {1,2,3} };

static int tetra_rule(int degree, double abc[][3], double *w) {
  int q = 0;
#define S31(A, W) do { double a_ = (A), c_ = 1 - 3 * a_; \
    double p_[4][3] = { {a_, a_, a_}, {a_, a_, c_}, {a_, c_, a_}, {c_, a_, a_} }; \
    for (int i_ = 0; i_ < 4; ++i_) { memcpy(abc[q], p_[i_], sizeof(p_[i_])); w[q++] = (W) / 6.0; } } while (0)
  if (degree == 1) {
    S31(0.13819660112501051517954131656343619, 0.25);
  } else if (degree == 2) {
    S31(0.31088591926330060979734573376345783, 0.11268792571801585079918565233328633);
    S31(0.09273525031089122640232391373703061, 0.07349304311636194954371020548632750);
    double a = 0.04550370412564964949188052627933943, c = (1 - 2 * a) * .5;
    double p[6][3] = { {a, a, c}, {a, c, a}, {a, c, c}, {c, a, a}, {c, a, c}, {c, c, a} };
    for (int i = 0; i < 6; ++i) { memcpy(abc[q], p[i], sizeof(p[i])); w[q++] = 0.04254602077708146643806942812025744 / 6.0; }
  } else {
    S31(0.21460287125915202928883921938628499, 0.03992275025816749209969062755747998);
    S31(0.04067395853461135311557944895641006, 0.01007721105532064294801323744593686);
    S31(0.32233789014227551034399447076249213, 0.05535718154365472209515327785372602);
    double a = 0.06366100187501752529923552760572698, b = 0.60300566479164914136743113906093969, c = 1 - a - a - b;
    double p[12][3] = { {a, a, b}, {a, a, c}, {a, b, a}, {a, b, c}, {a, c, a}, {a, c, b},
                        {b, a, a}, {b, a, c}, {b, c, a}, {c, a, a}, {c, a, b}, {c, b, a} };
    for (int i = 0; i < 12; ++i) { memcpy(abc[q], p[i], sizeof(p[i])); w[q++] = (27. / 560.) / 6.0; }
  }
#undef S31
  return q;
}

static int triangle_rule(int degree, double ab[][2], double *w) {
  int q = 0;
#define S21(A, W) do { double a_ = (A), b_ = 1 - a_ - a_; \
    double p_[3][2] = { {a_, a_}, {a_, b_}, {b_, a_} }; \
    for (int i_ = 0; i_ < 3; ++i_) { ab[q][0] = p_[i_][0]; ab[q][1] = p_[i_][1]; w[q++] = (W) / 2.0; } } while (0)
  if (degree == 1) {
    S21(1. / 6., 1. / 3.);
  } else if (degree == 2) {
    S21(.44594849091596488631832925388305199, .22338158967801146569500700843312280);
    S21(.09157621350977074345957146340220151, .10995174365532186763832632490021053);
  } else {
    S21(.06308901449150222834033160287081916, .05084490637020681692093680910686898);
    S21(.24928674517091042129163855310701908, .11678627572637936602528961138557944);
    double a = .05314504984481694735324967163139815, b = .31035245103378440541660773395655215, c = 1 - a - b;
    double p[6][2] = { {a, b}, {a, c}, {b, a}, {b, c}, {c, a}, {c, b} };
    for (int i = 0; i < 6; ++i) { ab[q][0] = p[i][0]; ab[q][1] = p[i][1]; w[q++] = .08285107561837357519355345642044245 / 2.0; }
  }
#undef S21
  return q;
}

static void tetra_map(const double *p[4], double a, double b, double c, double *glob) {
  double shape[4] = { a, b, c, 1.0 - a - b - c };  /* geometry/tetrahedron.hpp:96-101 */
  for (int d = 0; d < 3; ++d) {
    double sum = p[0][d] * shape[0];
    for (int v = 1; v < 4; ++v) sum += p[v][d] * shape[v];
    glob[d] = sum;
  }
}

static int find_tet_side(const int64_t *cell_nodes, const int64_t *face_nodes) {
  for (int s = 0; s < 4; ++s) {
    int cnt = 0;
    for (int v = 0; v < 3; ++v) for (int w = 0; w < 3; ++w)
      if (cell_nodes[kTetFace[s][v]] == face_nodes[w]) ++cnt;
    if (cnt == 3) return s;
  }
  return -1;
}

static void copy_halo_lists(PartTab *pt, const mcfd_mesh *m) {
  pt->n_nbr = m->n_neighbours;
  if (!pt->n_nbr) return;
  pt->nbr_rank = malloc(sizeof(int) * pt->n_nbr);
  pt->send_ptr = malloc(sizeof(int64_t) * (pt->n_nbr + 1)); pt->recv_ptr = malloc(sizeof(int64_t) * (pt->n_nbr + 1));
  memcpy(pt->send_ptr, m->send_ptr, sizeof(int64_t) * (pt->n_nbr + 1));
  memcpy(pt->recv_ptr, m->recv_ptr, sizeof(int64_t) * (pt->n_nbr + 1));
  for (int i = 0; i < pt->n_nbr; ++i) pt->nbr_rank[i] = m->nbr_rank[i];
  int64_t ns = pt->send_ptr[pt->n_nbr];
  pt->send_cells = malloc(sizeof(int64_t) * (ns + 1));
  memcpy(pt->send_cells, m->send_cells, sizeof(int64_t) * ns);
}

static int build_part_tetra(oracle_sim *sim, PartTab *pt, const mcfd_mesh *m) {
  int n = sim->nq, fq = sim->fq;
  memset(pt, 0, sizeof(*pt));
  pt->n_local = m->n_local_cells; pt->n_ghost = m->n_ghost_cells; pt->n_inner = m->n_inner_cells;
  pt->n_local_faces = m->n_local_faces; pt->n_ghost_faces = m->n_ghost_faces; pt->n_boundary_faces = m->n_boundary_faces;
  pt->n_faces = m->n_local_faces + m->n_ghost_faces + m->n_boundary_faces;
  int64_t nc = pt->n_local + pt->n_ghost;
  pt->xyz = malloc(sizeof(double) * nc * n * 3);
  pt->wglob = malloc(sizeof(double) * nc * n);
  pt->u = calloc(nc * sim->n * K, sizeof(double));
  double abc[24][3], wl[24], ab[12][2], wf[12];
  if (tetra_rule(sim->degree, abc, wl) != n || triangle_rule(sim->degree, ab, wf) != fq) return 3;
  for (int64_t c = 0; c < nc; ++c) {
    const double *p[4];
    for (int v = 0; v < 4; ++v) p[v] = m->node_xyz + 3 * m->cell_nodes[4 * c + v];
    double J[3][3];  /* geometry/cell.hpp LocalToJacobian: rows p0 - p3, p1 - p3, p2 - p3 */
    for (int a = 0; a < 3; ++a) for (int d = 0; d < 3; ++d) J[a][d] = p[a][d] - p[3][d];
    double det = det3(J);
    for (int q = 0; q < n; ++q) {  /* gauss/element.hpp:87-100 */
      tetra_map(p, abc[q][0], abc[q][1], abc[q][2], pt->xyz + (c * n + q) * 3);
      pt->wglob[c * n + q] = wl[q] * fabs(det);
    }
  }
  int64_t nf = pt->n_faces;
  pt->holder = malloc(sizeof(int64_t) * nf); pt->sharer = malloc(sizeof(int64_t) * nf);
  pt->f_group = malloc(sizeof(int) * (pt->n_boundary_faces + 1));
  pt->f_xyz = malloc(sizeof(double) * nf * fq * 3);
  pt->f_frame = malloc(sizeof(double) * nf * fq * 9);
  pt->f_w = malloc(sizeof(double) * nf * fq);
  pt->f_hside = malloc(sizeof(int) * nf); pt->f_sside = malloc(sizeof(int) * nf);
  pt->f_star_u = calloc(nf * fq, sizeof(double));
  for (int64_t b = 0; b < pt->n_boundary_faces; ++b) pt->f_group[b] = m->bface_group[b];
  for (int64_t f = 0; f < nf; ++f) {
    int64_t h = m->face_holder[f], s = m->face_sharer[f];
    pt->holder[f] = h; pt->sharer[f] = s;
    int side = find_tet_side(m->cell_nodes + 4 * h, m->face_nodes + 3 * f);
    if (side < 0) return 1;
    pt->f_hside[f] = side;
    pt->f_sside[f] = s >= 0 ? find_tet_side(m->cell_nodes + 4 * s, m->face_nodes + 3 * f) : -1;
    if (s >= 0 && pt->f_sside[f] < 0) return 1;
    /* geometry/tetrahedron.hpp:76-95 SortNodesOnFace: the Triangle3 on the holder's table order */
    const double *qn[3];
    for (int v = 0; v < 3; ++v) qn[v] = m->node_xyz + 3 * m->cell_nodes[4 * h + kTetFace[side][v]];
    double r0[3], r1[3];  /* geometry/face.hpp LocalToJacobian: rows q0 - q2, q1 - q2 */
    for (int d = 0; d < 3; ++d) { r0[d] = qn[0][d] - qn[2][d]; r1[d] = qn[1][d] - qn[2][d]; }
    double nrm[3] = { r0[1] * r1[2] - r0[2] * r1[1], r0[2] * r1[0] - r0[0] * r1[2], r0[0] * r1[1] - r0[1] * r1[0] };
    double ln = sqrt(nrm[0] * nrm[0] + nrm[1] * nrm[1] + nrm[2] * nrm[2]);
    double lt = sqrt(r0[0] * r0[0] + r0[1] * r0[1] + r0[2] * r0[2]);
    double a11 = lt * lt, a22 = r1[0] * r1[0] + r1[1] * r1[1] + r1[2] * r1[2], a12 = r0[0] * r1[0] + r0[1] * r1[1] + r0[2] * r1[2];
    double area2 = sqrt(a11 * a22 - a12 * a12);
    for (int p = 0; p < fq; ++p) {
      double shape[3] = { ab[p][0], ab[p][1], 1.0 - ab[p][0] - ab[p][1] };  /* geometry/triangle.hpp:78-83 */
      double *g = pt->f_xyz + (f * fq + p) * 3;
      for (int d = 0; d < 3; ++d) g[d] = qn[0][d] * shape[0] + qn[1][d] * shape[1] + qn[2][d] * shape[2];
      double *fr = pt->f_frame + (f * fq + p) * 9;  /* geometry/face.hpp:98-110 */
      for (int d = 0; d < 3; ++d) { fr[d] = nrm[d] / ln; fr[3 + d] = r0[d] / lt; }
      fr[6] = fr[1] * fr[5] - fr[2] * fr[4]; fr[7] = fr[2] * fr[3] - fr[0] * fr[5]; fr[8] = fr[0] * fr[4] - fr[1] * fr[3];
      pt->f_w[f * fq + p] = wf[p] * area2;
    }
  }
  copy_halo_lists(pt, m);
  return modal_build(sim, pt, m);
}

/* ------------------------------------------------------------------------------------------ */
/* polynomial/hexahedron.hpp:257-270 : GetValue (FR: coeff / det J ; DGSEM: coeff) */
static void get_value(const oracle_sim *sim, const PartTab *pt, int64_t cell, int node, double *u) {
  const double *c = pt->u + (cell * sim->n + node) * K;
  if (sim->scheme == MCFD_SCHEME_FR) {
    double det = pt->det[cell * sim->n + node];
    for (int k = 0; k < K; ++k) u[k] = c[k] / det;
  } else {
    for (int k = 0; k < K; ++k) u[k] = c[k];
  }
}

/* fr/general.hpp:232-248 (FR) and dg/lobatto.hpp:86-98,142-151 (DGSEM) for one cell.
 * The reference contracts with the dense N x N gradient table; the Lagrange tensor product makes
 * all but 3Q entries per node exact zeros, which are skipped here. */
static void add_flux_divergence(const oracle_sim *sim, const PartTab *pt, int64_t cell, double *res) {
  int q = sim->q, n = sim->n;
  static _Thread_local double G[NMAX][3][K];  /* contravariant flux per node and local direction */
  for (int node = 0; node < n; ++node) {
    double u[K], F[3][K];
    get_value(sim, pt, cell, node, u);
    flux_matrix(u, sim->gamma, F);
    const double *A = pt->jstar + (cell * n + node) * 9;  /* A[b][a] = det * dxi_a/dx_b */
    double scale = 1.0;
    if (sim->scheme == MCFD_SCHEME_DGSEM)  /* w_q * J^{-1} = w_q * A / det */
      scale = pt->wglob[cell * n + node] / pt->det[cell * n + node];
    for (int a = 0; a < 3; ++a)
      for (int k = 0; k < K; ++k)
        G[node][a][k] = (F[0][k] * A[0 * 3 + a] + F[1][k] * A[1 * 3 + a] + F[2][k] * A[2 * 3 + a]) * scale;
  }
  for (int i = 0; i < q; ++i) for (int j = 0; j < q; ++j) for (int k = 0; k < q; ++k) {
    int node = (i * q + j) * q + k;
    double val[K] = { 0, 0, 0, 0, 0 };
    for (int m = 0; m < q; ++m) {
      /* FR : sum_m D[i][m] G(m,j,k)   (derivative of basis m at node i)
       * DG : sum_m D[m][i] G(m,j,k)   (derivative of basis i at node m) */
      double dx = sim->scheme == MCFD_SCHEME_FR ? sim->D[i][m] : sim->D[m][i];
      double dy = sim->scheme == MCFD_SCHEME_FR ? sim->D[j][m] : sim->D[m][j];
      double dz = sim->scheme == MCFD_SCHEME_FR ? sim->D[k][m] : sim->D[m][k];
      const double *gx = G[(m * q + j) * q + k][0], *gy = G[(i * q + m) * q + k][1], *gz = G[(i * q + j) * q + m][2];
      for (int c = 0; c < K; ++c) val[c] += dx * gx[c] + dy * gy[c] + dz * gz[c];
    }
    double *r = res + (cell * n + node) * K;
    if (sim->scheme == MCFD_SCHEME_FR) for (int c = 0; c < K; ++c) r[c] -= val[c];
    else for (int c = 0; c < K; ++c) r[c] += val[c];
  }
}

/* Sign of the outward direction of local face `side`: faces 0,1,4 are xi/eta/zeta = -1
 * (fr/lobatto.hpp:96-156) and which local direction they are normal to. */
static const int kSideDir[6] = { 2, 1, 0, 1, 0, 2 };
static const double kSideSign[6] = { -1, -1, +1, +1, -1, +1 };

/* fr/general.hpp:249-264 + fr/lobatto.hpp:228-237 : correction of one side of one flux point */
static void fr_correct(const oracle_sim *sim, const PartTab *pt, int64_t cell, int node, int side,
                       const double *u, const double *f_upwind, double upwind_sign, double *res) {
  const double *A = pt->jstar + (cell * sim->n + node) * 9;
  int d = kSideDir[side];
  double normal[3] = { A[0 * 3 + d], A[1 * 3 + d], A[2 * 3 + d] };  /* J*.col(d) */
  double scale = kSideSign[side] * sqrt(normal[0] * normal[0] + normal[1] * normal[1] + normal[2] * normal[2]);
  double g_prime = kSideSign[side] * sim->g_prime;
  double F[3][K];
  flux_matrix(u, sim->gamma, F);
  double *r = res + (cell * sim->n + node) * K;
  for (int k = 0; k < K; ++k) {
    double f = f_upwind[k] * (upwind_sign * scale);
    f -= F[0][k] * normal[0] + F[1][k] * normal[1] + F[2][k] * normal[2];
    f *= g_prime;
    r[k] -= f;
  }
}

/* fr/lobatto.hpp:219-259 / dg/lobatto.hpp:156-195 : interior (local or ghost) faces */
static void add_flux_on_faces(oracle_sim *sim, PartTab *pt, int64_t f_begin, int64_t f_end, int ghost, double *res) {
  int fq = sim->fq, n = sim->n;
  for (int64_t f = f_begin; f < f_end; ++f) {
    int64_t h = pt->holder[f], s = pt->sharer[f];
    for (int p = 0; p < fq; ++p) {
      int hn = pt->f_hnode[f * fq + p], sn = pt->f_snode[f * fq + p];
      double uh[K], us[K], flux[K];
      get_value(sim, pt, h, hn, uh);
      get_value(sim, pt, s, sn, us);
      rotated_flux_upwind_kind(sim->riemann, pt->f_frame + (f * fq + p) * 9, uh, us, sim->gamma, pt->f_star_u + f * fq + p, flux);
      if (sim->scheme == MCFD_SCHEME_FR) {
        fr_correct(sim, pt, h, hn, pt->f_hside[f], uh, flux, +1.0, res);
        if (!ghost) fr_correct(sim, pt, s, sn, pt->f_sside[f], us, flux, -1.0, res);
      } else {
        double w = pt->f_w[f * fq + p];
        for (int k = 0; k < K; ++k) {
          double fw = flux[k] * w;
          res[(h * n + hn) * K + k] -= fw;
          if (!ghost) res[(s * n + sn) * K + k] += fw;
        }
      }
    }
  }
}

/* fem.hpp:185-226 order of kinds; fr/lobatto.hpp:260-391, dg/lobatto.hpp:196-300 */
static void add_flux_on_boundaries(oracle_sim *sim, PartTab *pt, double *res) {
  static const int kOrder[6] = { MCFD_BC_SOLID_WALL, MCFD_BC_SUPERSONIC_INLET, MCFD_BC_SUPERSONIC_OUTLET,
                                 MCFD_BC_SUBSONIC_INLET, MCFD_BC_SUBSONIC_OUTLET, MCFD_BC_SMART };
  int fq = sim->fq, n = sim->n;
  int64_t f0 = pt->n_local_faces + pt->n_ghost_faces;
  for (int o = 0; o < 6; ++o) {
    for (int g = 0; g < sim->n_groups; ++g) {
      if (sim->group_kind[g] != kOrder[o]) continue;
      for (int64_t b = 0; b < pt->n_boundary_faces; ++b) {
        if (pt->f_group[b] != g) continue;
        int64_t f = f0 + b, h = pt->holder[f];
        for (int p = 0; p < fq; ++p) {
          int hn = pt->f_hnode[f * fq + p];
          double uh[K], given[K] = { 0, 0, 0, 0, 0 }, flux[K];
          get_value(sim, pt, h, hn, uh);
          if (kOrder[o] != MCFD_BC_SOLID_WALL && kOrder[o] != MCFD_BC_SUPERSONIC_OUTLET)
            eval_state(&sim->group_func[g], pt->f_xyz + (f * fq + p) * 3, sim->t_curr, given);
          boundary_flux(kOrder[o], pt->f_frame + (f * fq + p) * 9, uh, given, sim->gamma, flux);
          if (sim->scheme == MCFD_SCHEME_FR) {
            fr_correct(sim, pt, h, hn, pt->f_hside[f], uh, flux, +1.0, res);
          } else {
            double w = pt->f_w[f * fq + p];
            for (int k = 0; k < K; ++k) res[(h * n + hn) * K + k] -= flux[k] * w;
          }
        }
      }
    }
  }
}

/* part.hpp:1258-1302 : ghost cell coefficients <- the peers' inter cells */
static void exchange_ghosts(oracle_sim *sim) {
  int fields = sim->n * K;
  for (int r = 0; r < sim->n_parts; ++r) {
    PartTab *pt = &sim->parts[r];
    for (int i = 0; i < pt->n_nbr; ++i) {
      PartTab *peer = &sim->parts[pt->nbr_rank[i]];
      int j = 0;
      while (peer->nbr_rank[j] != r) ++j;
      int64_t cnt = pt->recv_ptr[i + 1] - pt->recv_ptr[i];
      for (int64_t c = 0; c < cnt; ++c) {
        int64_t src = peer->send_cells[peer->send_ptr[j] + c];
        int64_t dst = pt->n_local + pt->recv_ptr[i] + c;
        memcpy(pt->u + dst * fields, peer->u + src * fields, sizeof(double) * fields);
      }
    }
  }
}

#include "modal_oracle.h"

/* fem.hpp:173-183 (+ dg/lobatto.hpp:124-139 mass division) for one part; ghosts already exchanged */
static void residual_of_part(oracle_sim *sim, PartTab *pt, double *res) {
  int n = sim->n;
  if (sim->scheme == MCFD_SCHEME_DG_MODAL) { modal_residual_of_part(sim, pt, res); return; }
  memset(res, 0, sizeof(double) * pt->n_local * n * K);
  for (int64_t c = 0; c < pt->n_local; ++c) add_flux_divergence(sim, pt, c, res);
  add_flux_on_faces(sim, pt, 0, pt->n_local_faces, 0, res);
  add_flux_on_boundaries(sim, pt, res);
  add_flux_on_faces(sim, pt, pt->n_local_faces, pt->n_local_faces + pt->n_ghost_faces, 1, res);
  if (sim->scheme == MCFD_SCHEME_DGSEM) {
    for (int64_t i = 0; i < pt->n_local * n; ++i) {
      double scale = 1.0 / pt->wglob[i];
      for (int k = 0; k < K; ++k) res[i * K + k] *= scale;
    }
  }
}

/* ------------------------------------------------------------------------------------------ */
/* public API (loaded with ctypes by tests/, smoke() and bench.py's cpu_baseline leg only)      */

oracle_sim *oracle_create(int n_parts, const mcfd_mesh *meshes, int scheme, int degree, int rk_order, double gamma) {
  if (degree < 1 || degree > 3 || (scheme != MCFD_SCHEME_FR && scheme != MCFD_SCHEME_DGSEM && scheme != MCFD_SCHEME_DG_MODAL))
    return NULL;
  oracle_sim *sim = calloc(1, sizeof(*sim));
  sim->scheme = scheme; sim->degree = degree; sim->rk_order = rk_order; sim->gamma = gamma;
  sim->q = degree + 1; sim->nq = sim->q * sim->q * sim->q; sim->fq = sim->q * sim->q;
  sim->n = scheme == MCFD_SCHEME_DG_MODAL ? taylor_size(degree) : sim->nq;
  sim->elem = meshes[0].cell_type;
  if (sim->elem == MCFD_CELL_TETRA_4) {  /* part.hpp:439-452 : gauss::Tetrahedron<4|14|24>, gauss::Triangle<3|6|12> */
    if (scheme != MCFD_SCHEME_DG_MODAL) { free(sim); return NULL; }
    sim->nq = degree == 1 ? 4 : degree == 2 ? 14 : 24;
    sim->fq = degree == 1 ? 3 : degree == 2 ? 6 : 12;
  }
  sim->limiter = MCFD_LIMITER_NONE; sim->weno_w0 = 1e-3; sim->weno_eps = 1e-6;
  if (scheme == MCFD_SCHEME_DG_MODAL) legendre_rule(sim->q, sim->x1, sim->w1);  /* Projection::GaussOnLine, projection.hpp:84 */
  else lobatto(sim->q, sim->x1, sim->w1);
  lagrange_derivatives(sim->q, sim->x1, sim->D);
  sim->g_prime = vincent_right_derivative(degree, 1.0);
  sim->n_parts = n_parts;
  sim->parts = calloc(n_parts, sizeof(PartTab));
  sim->n_groups = 0;
  for (int r = 0; r < n_parts; ++r) {
    if (build_part(sim, &sim->parts[r], &meshes[r])) return NULL;
    if (meshes[r].n_bc_groups > sim->n_groups) sim->n_groups = meshes[r].n_bc_groups;
  }
  sim->group_kind = malloc(sizeof(int) * (sim->n_groups + 1));
  sim->group_func = calloc(sim->n_groups + 1, sizeof(StateFunc));
  for (int g = 0; g < sim->n_groups; ++g) sim->group_kind[g] = -1;
  return sim;
}

void oracle_destroy(oracle_sim *sim) {
  if (!sim) return;
  for (int r = 0; r < sim->n_parts; ++r) {
    PartTab *pt = &sim->parts[r];
    free(pt->xyz); free(pt->det); free(pt->jstar); free(pt->wglob); free(pt->u); free(pt->holder); free(pt->sharer);
    free(pt->f_group); free(pt->f_xyz); free(pt->f_frame); free(pt->f_w); free(pt->f_hnode); free(pt->f_snode);
    free(pt->f_hside); free(pt->f_sside); free(pt->f_star_u); free(pt->nbr_rank); free(pt->send_ptr);
    free(pt->send_cells); free(pt->recv_ptr);
    free(pt->m_center); free(pt->m_S); free(pt->m_volume); free(pt->adj_ptr); free(pt->adj_cell); free(pt->adj_face);
  }
  free(sim->parts); free(sim->group_kind); free(sim->group_func); free(sim);
}

int64_t oracle_column_size(const oracle_sim *sim, int part) { return sim->parts[part].n_local * sim->n * K; }

int oracle_set_boundary(oracle_sim *sim, int group, int kind, int func, const double *params, int n_params) {
  if (group < 0 || group >= sim->n_groups || n_params > MCFD_MAX_FUNC_PARAMS) return 1;
  sim->group_kind[group] = kind;
  sim->group_func[group].func = func;
  for (int i = 0; i < n_params; ++i) sim->group_func[group].p[i] = params[i];
  return 0;
}

/* polynomial/hexahedron.hpp:499-513 : Approximate */
int oracle_approximate(oracle_sim *sim, int func, const double *params, int n_params, double t) {
  StateFunc sf; memset(&sf, 0, sizeof(sf)); sf.func = func;
  for (int i = 0; i < n_params && i < MCFD_MAX_FUNC_PARAMS; ++i) sf.p[i] = params[i];
  for (int r = 0; r < sim->n_parts; ++r) {
    PartTab *pt = &sim->parts[r];
    if (sim->scheme == MCFD_SCHEME_DG_MODAL) {  /* polynomial/projection.hpp:159-162 */
      for (int64_t c = 0; c < pt->n_local; ++c) modal_project_function(sim, pt, c, &sf, t);
      continue;
    }
    for (int64_t i = 0; i < pt->n_local * sim->n; ++i) {
      double *c = pt->u + i * K;
      eval_state(&sf, pt->xyz + i * 3, t, c);
      if (sim->scheme == MCFD_SCHEME_FR) for (int k = 0; k < K; ++k) c[k] *= pt->det[i];
    }
  }
  return 0;
}

/* mesh/part.hpp:1082-1096 : Part::MeasureL1Error of one part; l1[K] */
int oracle_measure_l1_error(const oracle_sim *sim, int part, int func, const double *params, int n_params, double t, double *l1) {
  StateFunc sf; memset(&sf, 0, sizeof(sf)); sf.func = func;
  for (int i = 0; i < n_params && i < MCFD_MAX_FUNC_PARAMS; ++i) sf.p[i] = params[i];
  const PartTab *pt = &sim->parts[part];
  for (int k = 0; k < K; ++k) l1[k] = 0;
  for (int64_t c = 0; c < pt->n_local; ++c) {
    for (int q = 0; q < sim->nq; ++q) {
      const double *x = pt->xyz + (c * sim->nq + q) * 3;
      double value[K], exact[K];
      if (sim->scheme == MCFD_SCHEME_DG_MODAL) modal_value(sim, pt, c, x, value);
      else get_value(sim, pt, c, q, value);
      eval_state(&sf, x, t, exact);
      for (int k = 0; k < K; ++k) { double v = value[k]; v -= exact[k]; v = fabs(v) * pt->wglob[c * sim->nq + q]; l1[k] += v; }
    }
  }
  return 0;
}

int oracle_get_solution(const oracle_sim *sim, int part, double *column) {
  memcpy(column, sim->parts[part].u, sizeof(double) * oracle_column_size(sim, part));
  return 0;
}
int oracle_set_solution(oracle_sim *sim, int part, const double *column) {
  memcpy(sim->parts[part].u, column, sizeof(double) * oracle_column_size(sim, part));
  return 0;
}
/* dg/general.hpp:247-250 : WithLimiterAndSource::SetSolutionColumn also runs Part::Reconstruct; call once
 * after every part has been set (the limiter exchanges halos) */
int oracle_limit(oracle_sim *sim) { modal_reconstruct(sim); return 0; }
int oracle_set_limiter(oracle_sim *sim, int limiter, double w0, double eps) {
  if (sim->scheme != MCFD_SCHEME_DG_MODAL && limiter != MCFD_LIMITER_NONE) return 1;
  sim->limiter = limiter; sim->weno_w0 = w0; sim->weno_eps = eps;
  return 0;
}
int oracle_set_riemann(oracle_sim *sim, int riemann) {
  if (riemann < MCFD_RIEMANN_EXACT || riemann > MCFD_RIEMANN_AUSM) return 1;
  sim->riemann = riemann;
  return 0;
}
int64_t oracle_troubled_count(const oracle_sim *sim) {
  int64_t n = 0;
  for (int r = 0; r < sim->n_parts; ++r) n += sim->parts[r].n_troubled;
  return n;
}

int oracle_get_residual(oracle_sim *sim, double t, double *const *columns) {
  sim->t_curr = t;
  exchange_ghosts(sim);
  for (int r = 0; r < sim->n_parts; ++r) residual_of_part(sim, &sim->parts[r], columns[r]);
  return 0;
}

/* temporal/ode.hpp:119-131 : u += dt * R(u, t), all parts in lock-step */
static void euler_next(oracle_sim *sim, double t, double dt, double **res) {
  oracle_get_residual(sim, t, res);
  for (int r = 0; r < sim->n_parts; ++r) {
    int64_t n = oracle_column_size(sim, r);
    double *u = sim->parts[r].u;
    for (int64_t i = 0; i < n; ++i) { double rr = res[r][i]; rr *= dt; u[i] += rr; }
  }
}

/* temporal/rk.hpp:22-24, 36-43, 55-72 */
int oracle_rk_update(oracle_sim *sim, double t, double dt, int n_steps) {
  int P = sim->n_parts;
  double **res = malloc(sizeof(double *) * P), **u_curr = malloc(sizeof(double *) * P);
  for (int r = 0; r < P; ++r) {
    res[r] = malloc(sizeof(double) * oracle_column_size(sim, r));
    u_curr[r] = malloc(sizeof(double) * oracle_column_size(sim, r));
  }
  for (int s = 0; s < n_steps; ++s) {
    double t_curr = t + dt * s;
    for (int r = 0; r < P; ++r) memcpy(u_curr[r], sim->parts[r].u, sizeof(double) * oracle_column_size(sim, r));
    euler_next(sim, t_curr, dt, res);
    modal_reconstruct(sim);  /* every SetSolutionColumn of WithLimiterAndSource limits (dg/general.hpp:247-250) */
    if (sim->rk_order == 2) {
      euler_next(sim, t_curr + dt, dt, res);
      for (int r = 0; r < P; ++r) {
        int64_t n = oracle_column_size(sim, r); double *u = sim->parts[r].u;
        for (int64_t i = 0; i < n; ++i) { double v = u[i]; v += u_curr[r][i]; v *= 0.5; u[i] = v; }
      }
      modal_reconstruct(sim);
    } else if (sim->rk_order >= 3) {
      euler_next(sim, t_curr + dt, dt, res);
      for (int r = 0; r < P; ++r) {
        int64_t n = oracle_column_size(sim, r); double *u = sim->parts[r].u;
        for (int64_t i = 0; i < n; ++i) { double v = u_curr[r][i]; v *= 3; v += u[i]; v /= 4; u[i] = v; }
      }
      modal_reconstruct(sim);
      euler_next(sim, t_curr + dt / 2, dt, res);
      for (int r = 0; r < P; ++r) {
        int64_t n = oracle_column_size(sim, r); double *u = sim->parts[r].u;
        for (int64_t i = 0; i < n; ++i) { double v = u[i]; v *= 2; v += u_curr[r][i]; v /= 3; u[i] = v; }
      }
      modal_reconstruct(sim);
    }
  }
  for (int r = 0; r < P; ++r) { free(res[r]); free(u_curr[r]); }
  free(res); free(u_curr);
  return 0;
}

/* standalone pieces, frames [n][9]; each call uses a fresh Riemann object (star_u = 0) */
int oracle_riemann_flux(double gamma, int64_t n, const double *frames, const double *ul, const double *ur, double *flux) {
  for (int64_t i = 0; i < n; ++i) {
    double star_u = 0.0;
    rotated_flux_upwind(frames + 9 * i, ul + 5 * i, ur + 5 * i, gamma, &star_u, flux + 5 * i);
  }
  return 0;
}
int oracle_riemann_flux_kind(int riemann, double gamma, int64_t n, const double *frames, const double *ul, const double *ur,
                              double *flux) {
  for (int64_t i = 0; i < n; ++i) {
    double star_u = 0.0;
    rotated_flux_upwind_kind(riemann, frames + 9 * i, ul + 5 * i, ur + 5 * i, gamma, &star_u, flux + 5 * i);
  }
  return 0;
}
int oracle_boundary_flux(int kind, double gamma, int64_t n, const double *frames, const double *ui, const double *uo,
                         double *flux) {
  for (int64_t i = 0; i < n; ++i) boundary_flux(kind, frames + 9 * i, ui + 5 * i, uo + 5 * i, gamma, flux + 5 * i);
  return 0;
}
