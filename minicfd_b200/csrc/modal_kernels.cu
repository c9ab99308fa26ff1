// modal_kernels.cu -- the MODAL discontinuous Galerkin scheme and the WENO limiters on the device.
//
// Replaces (paths relative to /root/reference/include/mini):
//   spatial/dg/general.hpp:50-61     volume term      coeff += (w_q F(u(x_q))) (K x 3) . grad phi(x_q) (3 x N)
//   spatial/dg/general.hpp:69-105    interior / ghost faces   coeff -/+= (w_q f*) phi(x_q)
//   spatial/dg/general.hpp:108-212   the six boundary kinds
//   spatial/dg/general.hpp:247-250 + mesh/part.hpp:1304-1334   SetSolutionColumn -> Part::Reconstruct(limiter)
//   limiter/weno.hpp:36-264          weno::Lazy, weno::Eigen (IsNotSmooth, Borrow, ReconstructOnFace)
//   polynomial/projection.hpp:25-39  Project (Cell::Approximate for the modal basis)
//
// Layout: the column is the reference's, coeff[cell][n][5] (Eigen column-major K x N, projection.hpp:163-170).
// Face kernel: one thread per face quadrature point; it evaluates both polynomials at the point, runs the
// Riemann solver and stores w_q f* once (same flux buffer and deterministic gather as the nodal scheme).
// Cell kernel: ONE WARP PER CELL.  Lanes first own quadrature points (basis values / gradients, flux), park
// the per-point factors in the warp's shared memory, then own output entries (n, k) and contract over the
// points -- no atomics, no cross-warp traffic, a fixed summation order.
#include <algorithm>

#include "modal_device.cuh"
#include "solver.hpp"
#include "state_functions.cuh"

namespace mcfd {

namespace {

constexpr unsigned kFull = 0xffffffffu;
constexpr int kWarpsPerBlock = 4;

__device__ __forceinline__ Frame load_frame(const DeviceTables &tab, int64_t f, int fq, int p) {
  const double *fp = tab.face_frame + (tab.per_point_frames ? f * fq + p : f) * 9;
  Frame fr;
#pragma unroll
  for (int d = 0; d < 3; ++d) { fr.n[d] = __ldg(fp + d); fr.t[d] = __ldg(fp + 3 + d); fr.b[d] = __ldg(fp + 6 + d); }
  return fr;
}

__device__ __forceinline__ double stage_combine_modal(const StageOp &op, double u, double r, double u0) {
  if (op.kind == 4) return r;
  double next = u + r * op.dt;  // ode.hpp:123-124 : r *= dt; u += r
  switch (op.kind) {
    case 1: return (u0 * 3 + next) / 4;    // rk.hpp:59-63
    case 2: return (next * 2 + u0) / 3;    // rk.hpp:65-70
    case 3: return (next + u0) * 0.5;      // rk.hpp:40-42
    default: return next;
  }
}

// ---- face quadrature points ---------------------------------------------------------------------------
template <int P, int kRiemann, bool kBoundary>
__global__ void __launch_bounds__(128) modal_face_kernel(DeviceTables tab, double gamma, double t,
                                                         const double *__restrict__ u, double *__restrict__ fstar,
                                                         int64_t point_begin, int64_t n_points) {
  using M = ModalShape<P>;
  constexpr int N = M::N, FQ = M::FQ;
  const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n_points) return;
  const int64_t gp = point_begin + i, f = gp / FQ;
  const int p = int(gp - f * FQ);
  const double *xw = tab.f_xyzw + gp * 4;
  const double x = __ldg(xw), y = __ldg(xw + 1), z = __ldg(xw + 2), w = __ldg(xw + 3);
  const int64_t h = tab.face_holder[f];
  double phi[N];
  modal_basis<P>(tab.m_S + h * N * N, tab.m_center + h * 3, x, y, z, phi);
  const Cons uh = modal_value<P>(u + h * N * 5, phi);
  const Frame fr = load_frame(tab, f, FQ, p);
  Cons flux;
  if (!kBoundary) {
    const int64_t s = tab.face_sharer[f];
    modal_basis<P>(tab.m_S + s * N * N, tab.m_center + s * 3, x, y, z, phi);
    const Cons us = modal_value<P>(u + s * N * 5, phi);
    flux = rotated_flux_upwind<kRiemann>(fr, uh, us, gamma);
  } else {
    const int64_t bf = f - (tab.n_local_faces + tab.n_ghost_faces);
    const BcGroup &grp = tab.groups[tab.bface_group[bf]];
    const int kind = grp.kind;
    if (kind < 0) return;  // boundary name never registered: no flux (fem.hpp:185-226)
    Cons given;
#pragma unroll
    for (int k = 0; k < 5; ++k) given.c[k] = 0.0;
    if (kind != MCFD_BC_SOLID_WALL && kind != MCFD_BC_SUPERSONIC_OUTLET) {
      if (grp.func == MCFD_FUNC_HOST_VALUES) {
        const double *v = grp.host_values + (int64_t(tab.bface_local[bf]) * FQ + p) * 5;
#pragma unroll
        for (int k = 0; k < 5; ++k) given.c[k] = v[k];
      } else {
        given = eval_state(grp.func, grp.params, x, y, z, t);
      }
    }
    flux = boundary_flux(kind, fr, uh, given, gamma);
  }
  double *out = fstar + ((gp >> 5) * (32 * MCFD_K) + (gp & 31));  // FstarIndex()
#pragma unroll
  for (int k = 0; k < 5; ++k) out[k * 32] = flux.c[k] * w;  // dg/general.hpp:81 : flux *= GetGlobalWeight(q)
}

// ---- the warp-level contraction shared by every modal kernel -------------------------------------------
// Each lane has parked, for "its" point q, a factor A[q][d][k] (d < D) and a factor B[q][n][d]; lane o then
// owns output entries (n, k) = (o / 5, o % 5), o = lane + 32 r, and adds  sum_q sum_d A[q][d][k] B[q][n][d].
template <int N> struct WarpScratch {
  static constexpr int kA = 32 * 3 * 5, kB = 32 * N * 3, kDoubles = kA + kB;
  static constexpr int R = (N * 5 + 31) / 32;  // output entries per lane
};

template <int N, int D>
__device__ __forceinline__ void warp_contract(const double *sA, const double *sB, int lane, double (&acc)[WarpScratch<N>::R]) {
#pragma unroll
  for (int r = 0; r < WarpScratch<N>::R; ++r) {
    const int o = lane + 32 * r;
    if (o >= N * 5) continue;
    const int n = o / 5, k = o - n * 5;
    double s = acc[r];
    for (int q = 0; q < 32; ++q)
#pragma unroll
      for (int d = 0; d < D; ++d) s += sA[(q * 3 + d) * 5 + k] * sB[(q * N + n) * 3 + d];
    acc[r] = s;
  }
}

template <int N>
__device__ __forceinline__ void park_zero(double *sA, double *sB, int lane) {
  for (int i = 0; i < 15; ++i) sA[lane * 15 + i] = 0.0;
  for (int i = 0; i < N * 3; ++i) sB[lane * N * 3 + i] = 0.0;
}

// ---- fused residual + Runge-Kutta update, one warp per cell --------------------------------------------
template <int P>
__global__ void __launch_bounds__(kWarpsPerBlock * 32) modal_cell_kernel(DeviceTables tab, double gamma,
                                                                          const double *__restrict__ u,
                                                                          const double *__restrict__ fstar,
                                                                          const double *__restrict__ u0,
                                                                          double *__restrict__ out, StageOp op) {
  using M = ModalShape<P>;
  using W = WarpScratch<M::N>;
  constexpr int N = M::N, NQ = M::NQ, FQ = M::FQ;
  extern __shared__ __align__(16) double smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t cell = int64_t(blockIdx.x) * kWarpsPerBlock + warp;
  if (cell >= tab.n_local) return;  // warp-uniform
  double *sA = smem + warp * W::kDoubles, *sB = sA + W::kA;
  const double *S = tab.m_S + cell * N * N, *center = tab.m_center + cell * 3, *coeff = u + cell * N * 5;
  const int64_t f_bnd = tab.n_local_faces + tab.n_ghost_faces;
  double acc[W::R];
#pragma unroll
  for (int r = 0; r < W::R; ++r) acc[r] = 0.0;

  // volume term, dg/general.hpp:50-61
  for (int base = 0; base < NQ; base += 32) {
    const int q = base + lane;
    park_zero<N>(sA, sB, lane);
    if (q < NQ) {
      const double *xw = tab.m_qxyzw + (cell * NQ + q) * 4;
      const double dx = __ldg(xw) - center[0], dy = __ldg(xw + 1) - center[1], dz = __ldg(xw + 2) - center[2];
      const double w = __ldg(xw + 3);
      double T[N], GT[N][3];
      taylor_value<P>(dx, dy, dz, T);
      taylor_grad<P>(dx, dy, dz, GT);
      Cons uq;
#pragma unroll
      for (int k = 0; k < 5; ++k) uq.c[k] = 0.0;
#pragma unroll
      for (int n = 0; n < N; ++n) {
        double phi = 0, g0 = 0, g1 = 0, g2 = 0;
#pragma unroll
        for (int m = 0; m <= n; ++m) {
          const double s = __ldg(S + n * N + m);
          phi += s * T[m]; g0 += s * GT[m][0]; g1 += s * GT[m][1]; g2 += s * GT[m][2];
        }
#pragma unroll
        for (int k = 0; k < 5; ++k) uq.c[k] += __ldg(coeff + n * 5 + k) * phi;
        sB[(lane * N + n) * 3 + 0] = g0; sB[(lane * N + n) * 3 + 1] = g1; sB[(lane * N + n) * 3 + 2] = g2;
      }
      // fem.hpp:264-278 -> types.hpp:279-290 : F(u) (3 x 5), times the global weight
      const Prim s = cons_to_prim(uq, gamma);
      const double vel[3] = { s.u, s.v, s.w };
      const double rho_h0 = uq.c[4] + s.p;
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        double F[5] = { uq.c[1 + d], uq.c[1] * vel[d], uq.c[2] * vel[d], uq.c[3] * vel[d], rho_h0 * vel[d] };
        F[1 + d] += s.p;
#pragma unroll
        for (int k = 0; k < 5; ++k) sA[(lane * 3 + d) * 5 + k] = F[k] * w;
      }
    }
    __syncwarp();
    warp_contract<N, 3>(sA, sB, lane, acc);
    __syncwarp();
  }

  // faces of this cell, dg/general.hpp:69-212 : the flux w_q f* was stored once per point by modal_face_kernel
  for (int base = 0; base < 6 * FQ; base += 32) {
    const int j = base + lane;
    park_zero<N>(sA, sB, lane);
    if (j < 6 * FQ) {
      const int side = j / FQ, p = j - side * FQ;
      const int f = __ldg(tab.cell_face + cell * 6 + side);
      const bool flux = f >= 0 && !(f >= f_bnd && tab.groups[tab.bface_group[f - f_bnd]].kind < 0);
      if (flux) {
        const double sign = (__ldg(tab.cell_code + cell * 6 + side) & 1) ? 1.0 : -1.0;  // holder -=, sharer +=
        const int64_t gp = int64_t(f) * FQ + p;
        const double *fs = fstar + ((gp >> 5) * (32 * MCFD_K) + (gp & 31));
        const double *xw = tab.f_xyzw + gp * 4;
        double phi[N];
        modal_basis<P>(S, center, __ldg(xw), __ldg(xw + 1), __ldg(xw + 2), phi);
#pragma unroll
        for (int n = 0; n < N; ++n) sB[(lane * N + n) * 3] = phi[n];
#pragma unroll
        for (int k = 0; k < 5; ++k) sA[(lane * 3) * 5 + k] = sign * __ldg(fs + k * 32);
      }
    }
    __syncwarp();
    warp_contract<N, 1>(sA, sB, lane, acc);
    __syncwarp();
  }

  // temporal/ode.hpp:119-131, rk.hpp : the stage combination
  const bool need_u0 = op.kind >= 1 && op.kind <= 3;
#pragma unroll
  for (int r = 0; r < W::R; ++r) {
    const int o = lane + 32 * r;
    if (o >= N * 5) continue;
    const int64_t idx = cell * N * 5 + o;
    out[idx] = stage_combine_modal(op, u[idx], acc[r], need_u0 ? u0[idx] : 0.0);
  }
}

// ---- Cell::Approximate for the modal basis: coeff = int f(x) phi(x)^t  (projection.hpp:25-39) ----------
template <int P>
__global__ void __launch_bounds__(kWarpsPerBlock * 32) modal_approximate_kernel(DeviceTables tab, int func,
                                                                                 FuncParams params, double t,
                                                                                 double *__restrict__ u) {
  using M = ModalShape<P>;
  using W = WarpScratch<M::N>;
  constexpr int N = M::N, NQ = M::NQ;
  extern __shared__ __align__(16) double smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t cell = int64_t(blockIdx.x) * kWarpsPerBlock + warp;
  if (cell >= tab.n_local) return;
  double *sA = smem + warp * W::kDoubles, *sB = sA + W::kA;
  double acc[W::R];
#pragma unroll
  for (int r = 0; r < W::R; ++r) acc[r] = 0.0;
  for (int base = 0; base < NQ; base += 32) {
    const int q = base + lane;
    park_zero<N>(sA, sB, lane);
    if (q < NQ) {
      const double *xw = tab.m_qxyzw + (cell * NQ + q) * 4;
      const double x = xw[0], y = xw[1], z = xw[2], w = xw[3];
      double phi[N];
      modal_basis<P>(tab.m_S + cell * N * N, tab.m_center + cell * 3, x, y, z, phi);
      const Cons f = eval_state(func, params.p, x, y, z, t);
#pragma unroll
      for (int n = 0; n < N; ++n) sB[(lane * N + n) * 3] = phi[n];
#pragma unroll
      for (int k = 0; k < 5; ++k) sA[(lane * 3) * 5 + k] = f.c[k] * w;
    }
    __syncwarp();
    warp_contract<N, 1>(sA, sB, lane, acc);
    __syncwarp();
  }
#pragma unroll
  for (int r = 0; r < W::R; ++r) {
    const int o = lane + 32 * r;
    if (o < N * 5) u[cell * N * 5 + o] = acc[r];
  }
}

// ---- limiter/weno.hpp on the device: one warp per cell ------------------------------------------------
template <int N> struct LimiterScratch {
  static constexpr int kCand = 7 * N * 5;  // (<= 6 neighbours + me) projections on MY basis
  static constexpr int kDoubles = WarpScratch<N>::kDoubles + 2 * kCand + 2 * N * 5 + 50 + 40;
};

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) v += __shfl_xor_sync(kFull, v, s);
  return v;
}

// weno::GetSmoothness (weno.hpp:20-34) of a projection given by its Taylor coefficients ct[m][5] (shared
// memory): lanes integrate over the quadrature points of `cell`, every lane gets beta[0..4]
template <int P>
__device__ __forceinline__ void warp_smoothness(const DeviceTables &tab, int64_t cell, const double *center,
                                                const double *ct, double w1, double w2, double w3, int lane,
                                                double (&beta)[5]) {
  constexpr int N = ModalShape<P>::N, NQ = ModalShape<P>::NQ;
  double part[5] = { 0, 0, 0, 0, 0 };
  for (int base = 0; base < NQ; base += 32) {
    const int q = base + lane;
    if (q < NQ) {
      const double *xw = tab.m_qxyzw + (cell * NQ + q) * 4;
      const double dx = __ldg(xw) - center[0], dy = __ldg(xw + 1) - center[1], dz = __ldg(xw + 2) - center[2];
      const double w = __ldg(xw + 3);
#pragma unroll
      for (int k = 0; k < 5; ++k) {
        double c[N];
#pragma unroll
        for (int m = 0; m < N; ++m) c[m] = ct[m * 5 + k];
        part[k] += smoothness_integrand<P>(dx, dy, dz, c, w1, w2, w3) * w;
      }
    }
  }
#pragma unroll
  for (int k = 0; k < 5; ++k) beta[k] = warp_sum(part[k]);
}

template <int P>
__global__ void __launch_bounds__(kWarpsPerBlock * 32) modal_limit_kernel(DeviceTables tab, LimiterOp lim, double gamma,
                                                                           const double *__restrict__ u,
                                                                           double *__restrict__ fresh,
                                                                           int32_t *__restrict__ flags,
                                                                           int64_t cell_begin, int64_t cell_end) {
  using M = ModalShape<P>;
  using W = WarpScratch<M::N>;
  constexpr int N = M::N, NQ = M::NQ, FQ = M::FQ, NK = N * 5;
  extern __shared__ __align__(16) double smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t cell = cell_begin + int64_t(blockIdx.x) * kWarpsPerBlock + warp;
  if (cell >= cell_end) return;  // warp-uniform
  double *sA = smem + warp * LimiterScratch<N>::kDoubles, *sB = sA + W::kA;
  double *s_cand = sB + W::kB, *s_tay = s_cand + LimiterScratch<N>::kCand;
  double *s_rot = s_tay + LimiterScratch<N>::kCand, *s_mix = s_rot + NK, *s_L = s_mix + NK, *s_R = s_L + 25;
  double *s_wgt = s_R + 25;  // [7][5]
  const int a0 = tab.adj_ptr[cell], n_adj = tab.adj_ptr[cell + 1] - a0;
  const double *S = tab.m_S + cell * N * N, *center = tab.m_center + cell * 3, *coeff = u + cell * NK;
  const double s00 = S[0];
  double my_avg[5];  // projection.hpp:41-47
#pragma unroll
  for (int k = 0; k < 5; ++k) my_avg[k] = coeff[k] * s00;

  // ---- weno::Eigen::IsNotSmooth (weno.hpp:152-180); weno::Lazy limits every cell (weno.hpp:58-60)
  if (lim.kind == MCFD_LIMITER_EIGEN) {
    double mine[2] = { 0, 0 };  // components rho and E of my polynomial at my centre: T(0) = e_0
    for (int n = 0; n < N; ++n) { mine[0] += coeff[n * 5] * S[n * N]; mine[1] += coeff[n * 5 + 4] * S[n * N]; }
    double diff[2] = { 0, 0 }, their_avg[2] = { 0, 0 };
    if (lane < n_adj) {
      const int64_t adj = tab.adj_cell[a0 + lane];
      double phi[N];
      modal_basis<P>(tab.m_S + adj * N * N, tab.m_center + adj * 3, center[0], center[1], center[2], phi);
      const Cons theirs = modal_value<P>(u + adj * NK, phi);
      diff[0] = fabs(mine[0] - theirs.c[0]); diff[1] = fabs(mine[1] - theirs.c[4]);
      const double t00 = tab.m_S[adj * N * N];
      their_avg[0] = fabs(u[adj * NK] * t00); their_avg[1] = fabs(u[adj * NK + 4] * t00);
    }
    double sum[2] = { 0, 0 }, big[2] = { fmax(1e-9, fabs(my_avg[0])), fmax(1e-9, fabs(my_avg[4])) };
    for (int a = 0; a < n_adj; ++a)  // the reference's order of summation
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        sum[c] += __shfl_sync(kFull, diff[c], a);
        big[c] = fmax(big[c], __shfl_sync(kFull, their_avg[c], a));
      }
    double divisor = pow(tab.m_volume[cell], (P + 1.0) / 2.0 / 3);
    divisor *= n_adj;
    const double reference = P < 3 ? 1.0 : 3.0;
    const bool troubled = sum[0] / big[0] / divisor > reference || sum[1] / big[1] / divisor > reference;
    if (!troubled) {
      if (lane == 0) flags[cell] = 0;
      return;
    }
  }

  // ---- Borrow (weno.hpp:193-203): neighbours re-projected over MY quadrature points, shifted to my average
  for (int a = 0; a < n_adj; ++a) {
    const int64_t adj = tab.adj_cell[a0 + a];
    double acc[W::R];
#pragma unroll
    for (int r = 0; r < W::R; ++r) acc[r] = 0.0;
    for (int base = 0; base < NQ; base += 32) {
      const int q = base + lane;
      park_zero<N>(sA, sB, lane);
      if (q < NQ) {
        const double *xw = tab.m_qxyzw + (cell * NQ + q) * 4;
        const double x = __ldg(xw), y = __ldg(xw + 1), z = __ldg(xw + 2), w = __ldg(xw + 3);
        double phi[N];
        modal_basis<P>(tab.m_S + adj * N * N, tab.m_center + adj * 3, x, y, z, phi);
        const Cons f = modal_value<P>(u + adj * NK, phi);
        modal_basis<P>(S, center, x, y, z, phi);
#pragma unroll
        for (int n = 0; n < N; ++n) sB[(lane * N + n) * 3] = phi[n];
#pragma unroll
        for (int k = 0; k < 5; ++k) sA[(lane * 3) * 5 + k] = f.c[k] * w;
      }
      __syncwarp();
      warp_contract<N, 1>(sA, sB, lane, acc);
      __syncwarp();
    }
    if (lane < 5) acc[0] += (my_avg[lane] - acc[0] * s00) / s00;  // projection.hpp:282-286
#pragma unroll
    for (int r = 0; r < W::R; ++r) {
      const int o = lane + 32 * r;
      if (o < NK) s_cand[a * NK + o] = acc[r];
    }
  }
  for (int o = lane; o < NK; o += 32) s_cand[n_adj * NK + o] = coeff[o];
  __syncwarp();
  // Taylor coefficients of every candidate: coeff * S (projection.hpp:131-133)
  for (int io = lane; io < (n_adj + 1) * NK; io += 32) {
    const int i = io / NK, o = io - i * NK, m = o / 5, k = o - m * 5;
    double s = 0;
    for (int n = m; n < N; ++n) s += s_cand[i * NK + n * 5 + k] * S[n * N + m];
    s_tay[io] = s;
  }
  __syncwarp();
  const double volume = 1 / (s00 * s00);  // OrthoNormal::Measure, linear.hpp:122-125
  const double w1 = pow(volume, 2. / 3 - 1), w2 = pow(volume, 4. / 3 - 1), w3 = volume;
  // nonlinear weights of the candidates ct[i] -> s_wgt[i][k], normalised (weno.hpp:216-235)
  auto nonlinear_weights = [&](auto taylor_of) {
    for (int i = 0; i <= n_adj; ++i) {
      double beta[5];
      warp_smoothness<P>(tab, cell, center, taylor_of(i), w1, w2, w3, lane, beta);
      if (lane < 5) {
        double w = lim.w0;
        if (i == n_adj) { w *= -n_adj; w += 1.0; }
        double b = beta[lane] + lim.eps;
        b *= b;
        s_wgt[i * 5 + lane] = w / b;
      }
      __syncwarp();
    }
    if (lane < 5) {
      double sum = 0;
      for (int i = 0; i <= n_adj; ++i) sum += s_wgt[i * 5 + lane];
      for (int i = 0; i <= n_adj; ++i) s_wgt[i * 5 + lane] /= sum;
    }
    __syncwarp();
  };

  if (lim.kind == MCFD_LIMITER_LAZY) {  // weno.hpp:97-126
    nonlinear_weights([&](int i) { return static_cast<const double *>(s_tay + i * NK); });
    for (int o = lane; o < NK; o += 32) {
      const int k = o % 5;
      double v = s_cand[n_adj * NK + o] * s_wgt[n_adj * 5 + k];
      for (int i = 0; i < n_adj; ++i) v += s_cand[i * NK + o] * s_wgt[i * 5 + k];
      fresh[cell * NK + o] = v;
    }
    if (lane == 0) flags[cell] = 1;
    return;
  }

  // weno::Eigen (weno.hpp:204-263): per adjacent face, the same weighting in characteristic variables
  double out[W::R];
#pragma unroll
  for (int r = 0; r < W::R; ++r) out[r] = 0.0;
  double total_volume = 0.0;
  for (int a = 0; a < n_adj; ++a) {
    const int packed = tab.adj_face[a0 + a];
    const int64_t f = packed / FQ;
    if (lane == 0) {  // riemann->UpdateEigenMatrices(big_u) in the frame of adj_face.riemann(0) (weno.hpp:211-214)
      const Frame fr = load_frame(tab, f, FQ, packed - int(f) * FQ);
      eigen_matrices(my_avg, fr, gamma, s_L, s_R);
    }
    __syncwarp();
    nonlinear_weights([&](int i) {
      for (int o = lane; o < NK; o += 32) {  // Taylor coefficients of L * candidate_i
        const int m = o / 5, k = o - m * 5;
        double s = 0;
#pragma unroll
        for (int j = 0; j < 5; ++j) s += s_L[k * 5 + j] * s_tay[i * NK + m * 5 + j];
        s_rot[o] = s;
      }
      __syncwarp();
      return static_cast<const double *>(s_rot);
    });
    // weighted sum of the rotated candidates (weno.hpp:236-241), then back with R (:243)
    for (int o = lane; o < NK; o += 32) {
      const int n = o / 5, k = o - n * 5;
      double v = 0;
      for (int ii = 0; ii <= n_adj; ++ii) {
        const int i = ii == 0 ? n_adj : ii - 1;  // my own projection first, as the reference accumulates
        double rot = 0;
#pragma unroll
        for (int j = 0; j < 5; ++j) rot += s_L[k * 5 + j] * s_cand[i * NK + n * 5 + j];
        v += rot * s_wgt[i * 5 + k];
      }
      s_mix[o] = v;
    }
    __syncwarp();
    const double adj_volume = tab.m_volume[tab.adj_cell[a0 + a]];  // weno.hpp:246-249
#pragma unroll
    for (int r = 0; r < W::R; ++r) {
      const int o = lane + 32 * r;
      if (o >= NK) continue;
      const int n = o / 5, k = o - n * 5;
      double back = 0;
#pragma unroll
      for (int j = 0; j < 5; ++j) back += s_R[k * 5 + j] * s_mix[n * 5 + j];
      out[r] += back * adj_volume;
    }
    total_volume += adj_volume;
    __syncwarp();  // s_L, s_R, s_mix are rewritten for the next face
  }
#pragma unroll
  for (int r = 0; r < W::R; ++r) {
    const int o = lane + 32 * r;
    if (o < NK) fresh[cell * NK + o] = out[r] / total_volume;  // weno.hpp:262
  }
  if (lane == 0) flags[cell] = 1;
}

// part.hpp:1329-1332 : only after every troubled cell of the set has been rebuilt from the OLD data
__global__ void modal_writeback_kernel(const double *__restrict__ fresh, const int32_t *__restrict__ flags, int fields,
                                       double *__restrict__ u, int64_t cell_begin, int64_t cell_end) {
  const int64_t idx = cell_begin * fields + int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= cell_end * fields) return;
  if (flags[idx / fields]) u[idx] = fresh[idx];
}

template <class Kernel>
cudaError_t set_smem(Kernel kernel, size_t bytes) {
  return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(bytes));
}

}  // namespace

// ---- launchers ----------------------------------------------------------------------------------------
template <int P>
static cudaError_t launch_modal_face(const DeviceTables &tab, int riemann, double gamma, double t, const double *u,
                                     double *fstar, int64_t face_begin, int64_t face_end, cudaStream_t stream) {
  constexpr int FQ = ModalShape<P>::FQ;
  const int64_t n_points = (face_end - face_begin) * FQ, point_begin = face_begin * FQ;
  const int threads = 128;
  const unsigned blocks = unsigned((n_points + threads - 1) / threads);
  if (face_begin >= tab.n_local_faces + tab.n_ghost_faces) {
    modal_face_kernel<P, 0, true><<<blocks, threads, 0, stream>>>(tab, gamma, t, u, fstar, point_begin, n_points);
    return cudaGetLastError();
  }
  switch (riemann) {
    case MCFD_RIEMANN_EXACT: modal_face_kernel<P, 0, false><<<blocks, threads, 0, stream>>>(tab, gamma, t, u, fstar, point_begin, n_points); break;
    case MCFD_RIEMANN_ROE: modal_face_kernel<P, 1, false><<<blocks, threads, 0, stream>>>(tab, gamma, t, u, fstar, point_begin, n_points); break;
    case MCFD_RIEMANN_HLLC: modal_face_kernel<P, 2, false><<<blocks, threads, 0, stream>>>(tab, gamma, t, u, fstar, point_begin, n_points); break;
    case MCFD_RIEMANN_AUSM: modal_face_kernel<P, 3, false><<<blocks, threads, 0, stream>>>(tab, gamma, t, u, fstar, point_begin, n_points); break;
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

cudaError_t LaunchModalFaceFlux(const DeviceTables &tab, int degree, int riemann, double gamma, double t, const double *u,
                                double *fstar, int64_t face_begin, int64_t face_end, cudaStream_t stream) {
  if (face_end <= face_begin) return cudaSuccess;
  switch (degree) {
    case 1: return launch_modal_face<1>(tab, riemann, gamma, t, u, fstar, face_begin, face_end, stream);
    case 2: return launch_modal_face<2>(tab, riemann, gamma, t, u, fstar, face_begin, face_end, stream);
    case 3: return launch_modal_face<3>(tab, riemann, gamma, t, u, fstar, face_begin, face_end, stream);
    default: return cudaErrorInvalidValue;
  }
}

template <int P>
static cudaError_t launch_modal_cell(const DeviceTables &tab, double gamma, const double *u, const double *fstar,
                                     const double *u0, double *out, StageOp op, cudaStream_t stream) {
  const size_t smem = size_t(kWarpsPerBlock) * WarpScratch<ModalShape<P>::N>::kDoubles * sizeof(double);
  cudaError_t e = set_smem(modal_cell_kernel<P>, smem);
  if (e != cudaSuccess) return e;
  const unsigned blocks = unsigned((tab.n_local + kWarpsPerBlock - 1) / kWarpsPerBlock);
  modal_cell_kernel<P><<<blocks, kWarpsPerBlock * 32, smem, stream>>>(tab, gamma, u, fstar, u0, out, op);
  return cudaGetLastError();
}

cudaError_t LaunchModalCellUpdate(const DeviceTables &tab, int degree, double gamma, const double *u, const double *fstar,
                                  const double *u0, double *out, StageOp op, cudaStream_t stream) {
  if (tab.n_local <= 0) return cudaSuccess;
  switch (degree) {
    case 1: return launch_modal_cell<1>(tab, gamma, u, fstar, u0, out, op, stream);
    case 2: return launch_modal_cell<2>(tab, gamma, u, fstar, u0, out, op, stream);
    case 3: return launch_modal_cell<3>(tab, gamma, u, fstar, u0, out, op, stream);
    default: return cudaErrorInvalidValue;
  }
}

template <int P>
static cudaError_t launch_modal_approximate(const DeviceTables &tab, int func, const FuncParams &fp, double t, double *u,
                                            cudaStream_t stream) {
  const size_t smem = size_t(kWarpsPerBlock) * WarpScratch<ModalShape<P>::N>::kDoubles * sizeof(double);
  cudaError_t e = set_smem(modal_approximate_kernel<P>, smem);
  if (e != cudaSuccess) return e;
  const unsigned blocks = unsigned((tab.n_local + kWarpsPerBlock - 1) / kWarpsPerBlock);
  modal_approximate_kernel<P><<<blocks, kWarpsPerBlock * 32, smem, stream>>>(tab, func, fp, t, u);
  return cudaGetLastError();
}

cudaError_t LaunchModalApproximate(const DeviceTables &tab, int degree, int func, const double *params, double t, double *u,
                                   cudaStream_t stream) {
  FuncParams fp;
  for (int i = 0; i < MCFD_MAX_FUNC_PARAMS; ++i) fp.p[i] = params[i];
  switch (degree) {
    case 1: return launch_modal_approximate<1>(tab, func, fp, t, u, stream);
    case 2: return launch_modal_approximate<2>(tab, func, fp, t, u, stream);
    case 3: return launch_modal_approximate<3>(tab, func, fp, t, u, stream);
    default: return cudaErrorInvalidValue;
  }
}

template <int P>
static cudaError_t launch_modal_limit(const DeviceTables &tab, LimiterOp lim, double gamma, double *u, double *fresh,
                                      int32_t *flags, int64_t cell_begin, int64_t cell_end, cudaStream_t stream) {
  constexpr int N = ModalShape<P>::N;
  const size_t smem = size_t(kWarpsPerBlock) * LimiterScratch<N>::kDoubles * sizeof(double);
  cudaError_t e = set_smem(modal_limit_kernel<P>, smem);
  if (e != cudaSuccess) return e;
  const int64_t n = cell_end - cell_begin;
  const unsigned blocks = unsigned((n + kWarpsPerBlock - 1) / kWarpsPerBlock);
  modal_limit_kernel<P><<<blocks, kWarpsPerBlock * 32, smem, stream>>>(tab, lim, gamma, u, fresh, flags, cell_begin, cell_end);
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  const int threads = 256;
  const int64_t total = n * N * 5;
  modal_writeback_kernel<<<unsigned((total + threads - 1) / threads), threads, 0, stream>>>(fresh, flags, N * 5, u, cell_begin, cell_end);
  return cudaGetLastError();
}

cudaError_t LaunchModalLimit(const DeviceTables &tab, int degree, LimiterOp lim, double gamma, double *u, double *fresh,
                             int32_t *flags, int64_t cell_begin, int64_t cell_end, cudaStream_t stream) {
  if (cell_end <= cell_begin || lim.kind == MCFD_LIMITER_NONE) return cudaSuccess;
  switch (degree) {
    case 1: return launch_modal_limit<1>(tab, lim, gamma, u, fresh, flags, cell_begin, cell_end, stream);
    case 2: return launch_modal_limit<2>(tab, lim, gamma, u, fresh, flags, cell_begin, cell_end, stream);
    case 3: return launch_modal_limit<3>(tab, lim, gamma, u, fresh, flags, cell_begin, cell_end, stream);
    default: return cudaErrorInvalidValue;
  }
}

}  // namespace mcfd
