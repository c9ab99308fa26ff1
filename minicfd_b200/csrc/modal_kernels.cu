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
#include <atomic>
#include <cstdlib>

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

// ---- Taylor form of the cell polynomials ----------------------------------------------------------------
// u_h(x) = coeff . phi(x) = coeff . S T(x - c) = (S^t coeff) . T(x - c)  (basis/linear.hpp:116-121): the face
// kernel evaluates both neighbours at every face point, so the product S^t coeff (N x 5 per cell) is formed once
// per cell and stage here, and a face point reads 50 numbers per side instead of 150 (S and coeff).
template <int P>
__global__ void __launch_bounds__(256) modal_taylor_kernel(DeviceTables tab, const double *__restrict__ u,
                                                           double *__restrict__ tay, int64_t cell_begin,
                                                           int64_t n_entries) {
  constexpr int N = ModalShape<P>::N, NO = N * 5;
  const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n_entries) return;
  const int64_t cell = cell_begin + i / NO;
  const int o = int(i % NO), m = o / 5, k = o - m * 5;
  const double *S = tab.m_S + cell * N * N, *c = u + cell * NO;
  double s = 0;
  for (int n = m; n < N; ++n) s += __ldg(S + n * N + m) * c[n * 5 + k];
  tay[cell * NO + o] = s;
}

template <int P>
__device__ __forceinline__ Cons taylor_eval(const double *__restrict__ tay, const double *__restrict__ center, double x,
                                            double y, double z) {
  constexpr int N = ModalShape<P>::N;
  double T[N];
  taylor_value<P>(x - __ldg(center), y - __ldg(center + 1), z - __ldg(center + 2), T);
  Cons u;
#pragma unroll
  for (int k = 0; k < 5; ++k) u.c[k] = 0;
#pragma unroll
  for (int m = 0; m < N; ++m)
#pragma unroll
    for (int k = 0; k < 5; ++k) u.c[k] += __ldg(tay + m * 5 + k) * T[m];
  return u;
}

// ---- face quadrature points ---------------------------------------------------------------------------
// MCFD_MODAL_FACE_THREADS (build-time tuning): one-warp blocks, as in the nodal face kernel (face_kernels.cu, r01w): a
// warp's slot is handed on when ITS Newton loops end; 0.457 against 0.464 ms per launch at 128 threads (r02h)
#ifndef MCFD_MODAL_FACE_THREADS
#define MCFD_MODAL_FACE_THREADS 32
#endif
#ifndef MCFD_MODAL_FACE_MINB
#define MCFD_MODAL_FACE_MINB (1024 / MCFD_MODAL_FACE_THREADS)
#endif
template <int PE, int kRiemann, bool kBoundary>
__global__ void __launch_bounds__(MCFD_MODAL_FACE_THREADS, MCFD_MODAL_FACE_MINB) modal_face_kernel(DeviceTables tab, double gamma, double t,
                                                         const double *__restrict__ tay, double *__restrict__ fstar,
                                                         int64_t point_begin, int64_t n_points) {
  using M = ModalShape<PE>;
  constexpr int P = M::P, N = M::N, FQ = M::FQ;
  const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n_points) return;
  const int64_t gp = point_begin + i, f = gp / FQ;
  const int p = int(gp - f * FQ);
  const double *xw = tab.f_xyzw + gp * 4;
  const double x = __ldg(xw), y = __ldg(xw + 1), z = __ldg(xw + 2), w = __ldg(xw + 3);
  const int64_t h = tab.face_holder[f];
  const Cons uh = taylor_eval<P>(tay + h * N * 5, tab.m_center + h * 3, x, y, z);
  const Frame fr = load_frame(tab, f, FQ, p);
  Cons flux;
  if (!kBoundary) {
    const int64_t s = tab.face_sharer[f];
    const Cons us = taylor_eval<P>(tay + s * N * 5, tab.m_center + s * 3, x, y, z);
    flux = rotated_flux_upwind<kRiemann>(fr, uh, us, gamma, tab.star_u + gp);
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
template <int PE>
__global__ void __launch_bounds__(kWarpsPerBlock * 32) modal_cell_kernel(DeviceTables tab, double gamma,
                                                                          const double *__restrict__ u,
                                                                          const double *__restrict__ fstar,
                                                                          const double *__restrict__ u0,
                                                                          double *__restrict__ out, StageOp op) {
  using M = ModalShape<PE>;
  using W = WarpScratch<M::N>;
  constexpr int P = M::P, N = M::N, NQ = M::NQ, FQ = M::FQ, NFP = M::NF * M::FQ;
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
  for (int base = 0; base < NFP; base += 32) {
    const int j = base + lane;
    park_zero<N>(sA, sB, lane);
    if (j < NFP) {
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

// ---- the same fused residual + Runge-Kutta update for P <= 2, without the shared-memory contraction --------
// Every lane owns at most one volume quadrature point and two face points (NQ <= 32, 6 FQ <= 64) and keeps the
// N x 5 products  (w_q F) . grad phi_n  /  -+(w_q f*) phi_n  of ITS points in registers; a butterfly reduce-scatter
// over the warp (31 shuffles per 32 outputs instead of 5 per output) leaves output entry o = lane + 32 b fully
// summed in lane `lane`.  The cell's S matrix and coefficients are staged once per warp in shared memory (950 B),
// so 8+ warps per SM sub-partition are resident instead of 4 and no lane ever loops over another lane's zeros.
template <int P, int n>
__device__ __forceinline__ void ortho_grad_row(const double *__restrict__ Sn, double x, double y, double z, double &g0,
                                               double &g1, double &g2) {
  // sum_{m <= n} S[n][m] grad T_m : only the non-zero entries of basis/taylor.hpp:183-186, 261-277
  g0 = 0; g1 = 0; g2 = 0;
  if (n >= 1) g0 += Sn[1];
  if (n >= 2) g1 += Sn[2];
  if (n >= 3) g2 += Sn[3];
  if (P >= 2) {
    if (n >= 4) g0 += Sn[4] * (2 * x);
    if (n >= 5) { g0 += Sn[5] * y; g1 += Sn[5] * x; }
    if (n >= 6) { g0 += Sn[6] * z; g2 += Sn[6] * x; }
    if (n >= 7) g1 += Sn[7] * (2 * y);
    if (n >= 8) { g1 += Sn[8] * z; g2 += Sn[8] * y; }
    if (n >= 9) g2 += Sn[9] * (2 * z);
  }
}

template <int P, int n>
struct VolumeRows {
  template <int NO>
  static __device__ __forceinline__ void run(const double *__restrict__ S, double x, double y, double z,
                                             const double (&Fw)[3][5], double (&part)[NO]) {
    constexpr int N = ModalShape<P>::N;
    double g0, g1, g2;
    ortho_grad_row<P, n>(S + n * N, x, y, z, g0, g1, g2);
#pragma unroll
    for (int k = 0; k < 5; ++k) part[n * 5 + k] = fma(Fw[2][k], g2, fma(Fw[1][k], g1, Fw[0][k] * g0));
    VolumeRows<P, n + 1>::run(S, x, y, z, Fw, part);
  }
};
template <> struct VolumeRows<1, 4> {
  template <int NO> static __device__ __forceinline__ void run(const double *, double, double, double,
                                                               const double (&)[3][5], double (&)[NO]) {}
};
template <> struct VolumeRows<2, 10> {
  template <int NO> static __device__ __forceinline__ void run(const double *, double, double, double,
                                                               const double (&)[3][5], double (&)[NO]) {}
};

// one L1 prefetch per 128-byte line of [p, p + bytes), spread over the lanes of a warp
__device__ __forceinline__ void prefetch_range(const void *p, int bytes, int lane) {
  const int off = lane * 128;
  if (off < bytes + 128) asm volatile("prefetch.global.L1 [%0];" ::"l"(static_cast<const char *>(p) + min(off, bytes - 1)));
}

template <int H>
__device__ __forceinline__ void butterfly_step(double *v, int lane) {
  const bool up = lane & H;
#pragma unroll
  for (int i = 0; i < H; ++i) {
    const double send = up ? v[i] : v[i + H], keep = up ? v[i + H] : v[i];
    v[i] = keep + __shfl_xor_sync(kFull, send, H);
  }
}

// Per-warp staging buffer of ONE cell's inputs, filled by cp.async (LDGSTS: global -> shared without registers).
template <int PE> struct CellStage {
  using M = ModalShape<PE>;
  static constexpr int N = M::N, NO = N * 5, kSlots = (M::NF * M::FQ + 31) / 32;
  static constexpr int oS = 0, oC = oS + N * N, oU0 = oC + NO, oVq = oU0 + NO, oFq = oVq + 32 * 4,
                       oFf = oFq + kSlots * 32 * 4, kDoubles = oFf + kSlots * 5 * 32;
  static_assert((N * N) % 2 == 0 && NO % 2 == 0, "16-byte cp.async chunks");
};

__device__ __forceinline__ void cp_async16(void *dst_smem, const void *src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(uint32_t(__cvta_generic_to_shared(dst_smem))), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async8(void *dst_smem, const void *src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(uint32_t(__cvta_generic_to_shared(dst_smem))), "l"(src) : "memory");
}

// Persistent warps, two staging buffers per warp: while a warp works on one cell, the copies of the next cell's S,
// coefficients, u0, quadrature points, face points and common fluxes are in flight (ncu r01f: with one cell per
// warp and 3 warps per scheduler, 35 % of the samples sat on the first use of those loads).
#ifndef MCFD_MODAL_CELL_MINB
#define MCFD_MODAL_CELL_MINB 2
#endif
template <int PE>
__global__ void __launch_bounds__(kWarpsPerBlock * 32, MCFD_MODAL_CELL_MINB) modal_cell_kernel_reg(DeviceTables tab, double gamma,
                                                                                 const double *__restrict__ u,
                                                                                 const double *__restrict__ fstar,
                                                                                 const double *__restrict__ u0,
                                                                                 double *__restrict__ out,
                                                                                 double *__restrict__ tay_out, StageOp op) {
  using M = ModalShape<PE>;
  using St = CellStage<PE>;
  constexpr int P = M::P, N = M::N, NQ = M::NQ, FQ = M::FQ, NFP = M::NF * M::FQ, NO = N * 5, kSlots = St::kSlots;
  static_assert(NQ <= 32 && NFP <= 64, "one volume point and two face points per lane");
  extern __shared__ __align__(16) double smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int RP = NO <= 32 ? NO : NO / 2;  // rows of the lane-sum transpose per pass
  static_assert(NO % RP == 0 && RP <= 32, "whole passes");
  constexpr int kRed = (RP * 33 + 1) / 2 * 2;  // even: the next warp's staging buffers stay 16-byte aligned
  double *stage = smem + warp * (2 * St::kDoubles + kRed);
  double *red = stage + 2 * St::kDoubles;
  const int64_t f_bnd = tab.n_local_faces + tab.n_ghost_faces;
  const bool need_u0 = op.kind >= 1 && op.kind <= 3;
  const int64_t stride = int64_t(gridDim.x) * kWarpsPerBlock;

  // my face points: slot -> (side, point); the face ids of a cell are fetched one iteration before its copies
  int side_of[kSlots], pt_of[kSlots];
#pragma unroll
  for (int slot = 0; slot < kSlots; ++slot) {
    const int j = lane + 32 * slot;
    side_of[slot] = j < NFP ? j / FQ : -1;
    pt_of[slot] = j < NFP ? j % FQ : 0;
  }
  auto load_faces = [&](int64_t c, int (&fid)[kSlots], int (&code)[kSlots]) {
#pragma unroll
    for (int slot = 0; slot < kSlots; ++slot) {
      fid[slot] = -1; code[slot] = 0;
      if (c < tab.n_local && side_of[slot] >= 0) {
        fid[slot] = __ldg(tab.cell_face + c * 6 + side_of[slot]);
        code[slot] = __ldg(tab.cell_code + c * 6 + side_of[slot]);
      }
    }
  };
  // flux point of my face point in each slot (-1: none, or a boundary whose name was never registered)
  auto flux_points = [&](const int (&fid)[kSlots], int (&gp)[kSlots]) {
#pragma unroll
    for (int slot = 0; slot < kSlots; ++slot) {
      const int f = fid[slot];
      const bool flux = f >= 0 && !(f >= f_bnd && tab.groups[tab.bface_group[f - f_bnd]].kind < 0);
      gp[slot] = flux ? f * FQ + pt_of[slot] : -1;
    }
  };
  auto issue = [&](int64_t c, const int (&gp)[kSlots], double *buf) {
    for (int i = lane; i < N * N / 2; i += 32) cp_async16(buf + St::oS + 2 * i, tab.m_S + c * N * N + 2 * i);
    for (int i = lane; i < NO / 2; i += 32) cp_async16(buf + St::oC + 2 * i, u + c * NO + 2 * i);
    if (need_u0)
      for (int i = lane; i < NO / 2; i += 32) cp_async16(buf + St::oU0 + 2 * i, u0 + c * NO + 2 * i);
    if (lane < NQ) {
      const double *xw = tab.m_qxyzw + (c * NQ + lane) * 4;
      cp_async16(buf + St::oVq + lane * 4, xw);
      cp_async16(buf + St::oVq + lane * 4 + 2, xw + 2);
    }
#pragma unroll
    for (int slot = 0; slot < kSlots; ++slot) {
      if (gp[slot] < 0) continue;
      const int64_t g = gp[slot];
      const double *xw = tab.f_xyzw + g * 4;
      cp_async16(buf + St::oFq + (slot * 32 + lane) * 4, xw);
      cp_async16(buf + St::oFq + (slot * 32 + lane) * 4 + 2, xw + 2);
      const double *fs = fstar + ((g >> 5) * (32 * MCFD_K) + (g & 31));
#pragma unroll
      for (int k = 0; k < 5; ++k) cp_async8(buf + St::oFf + (slot * 5 + k) * 32 + lane, fs + k * 32);
    }
  };

  int64_t cell = int64_t(blockIdx.x) * kWarpsPerBlock + warp;
  if (cell >= tab.n_local) return;  // warp-uniform
  int fid[kSlots], code[kSlots], gp[kSlots], sharer[kSlots];
  load_faces(cell, fid, code);
  flux_points(fid, gp);
#pragma unroll
  for (int slot = 0; slot < kSlots; ++slot) sharer[slot] = code[slot] & 1;
  issue(cell, gp, stage);
  asm volatile("cp.async.commit_group;" ::: "memory");
  load_faces(cell + stride, fid, code);  // consumed when the next cell's copies are issued
  const double cxyz0 = 0;
  (void)cxyz0;

  for (int it = 0; cell < tab.n_local; cell += stride, ++it) {
    double *buf = stage + (it & 1) * St::kDoubles;
    const int64_t next = cell + stride;
    int gp_next[kSlots], sharer_next[kSlots];
#pragma unroll
    for (int slot = 0; slot < kSlots; ++slot) { gp_next[slot] = -1; sharer_next[slot] = code[slot] & 1; }
    if (next < tab.n_local) {
      flux_points(fid, gp_next);
      issue(next, gp_next, stage + ((it + 1) & 1) * St::kDoubles);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    load_faces(next + stride, fid, code);
    const double cx = __ldg(tab.m_center + cell * 3), cy = __ldg(tab.m_center + cell * 3 + 1),
                 cz = __ldg(tab.m_center + cell * 3 + 2);
    const double center[3] = { cx, cy, cz };
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    __syncwarp();
    const double *S = buf + St::oS, *cf = buf + St::oC;

    double part[NO];
#pragma unroll
    for (int o = 0; o < NO; ++o) part[o] = 0.0;

    // volume term, dg/general.hpp:50-61
    if (lane < NQ) {
      const double *vq = buf + St::oVq + lane * 4;
      const double dx = vq[0] - cx, dy = vq[1] - cy, dz = vq[2] - cz, w = vq[3];
      double T[N];
      taylor_value<P>(dx, dy, dz, T);
      Cons uq;
#pragma unroll
      for (int k = 0; k < 5; ++k) uq.c[k] = 0.0;
#pragma unroll
      for (int n = 0; n < N; ++n) {
        double phi = 0;
#pragma unroll
        for (int m = 0; m <= n; ++m) phi += S[n * N + m] * T[m];
#pragma unroll
        for (int k = 0; k < 5; ++k) uq.c[k] += cf[n * 5 + k] * phi;
      }
      // fem.hpp:264-278 -> types.hpp:279-290 : F(u) (3 x 5), times the global weight
      const Prim s = cons_to_prim(uq, gamma);
      const double vel[3] = { s.u, s.v, s.w };
      const double rho_h0 = uq.c[4] + s.p;
      double Fw[3][5];
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        double F[5] = { uq.c[1 + d], uq.c[1] * vel[d], uq.c[2] * vel[d], uq.c[3] * vel[d], rho_h0 * vel[d] };
        F[1 + d] += s.p;
#pragma unroll
        for (int k = 0; k < 5; ++k) Fw[d][k] = F[k] * w;
      }
      VolumeRows<P, 0>::run(S, dx, dy, dz, Fw, part);
    }

    // faces of this cell, dg/general.hpp:69-212 : the flux w_q f* was stored once per point by modal_face_kernel
#pragma unroll
    for (int slot = 0; slot < kSlots; ++slot) {
      if (gp[slot] >= 0) {
        const double sign = sharer[slot] ? 1.0 : -1.0;  // holder -=, sharer +=
        const double *xq = buf + St::oFq + (slot * 32 + lane) * 4;
        double fsv[5];
#pragma unroll
        for (int k = 0; k < 5; ++k) fsv[k] = sign * buf[St::oFf + (slot * 5 + k) * 32 + lane];
        double phi[N];
        modal_basis<P>(S, center, xq[0], xq[1], xq[2], phi);
#pragma unroll
        for (int n = 0; n < N; ++n)
#pragma unroll
          for (int k = 0; k < 5; ++k) part[n * 5 + k] = fma(fsv[k], phi[n], part[n * 5 + k]);
      }
    }

    // sum over the lanes through a [row][lane] transpose in shared memory (RP rows per pass, row stride 33: no
    // bank conflicts either way), then temporal/ode.hpp:119-131, rk.hpp : the stage combination.  A butterfly of
    // shuffles does the same in registers but costs 4x the instructions (selects and moves; ncu r01i: 40 % of the
    // kernel's instructions).
    double fresh[NO / RP];
#pragma unroll
    for (int h = 0; h < NO / RP; ++h) {
#pragma unroll
      for (int o = 0; o < RP; ++o) red[o * 33 + lane] = part[h * RP + o];
      __syncwarp();
      fresh[h] = 0.0;
      if (lane < RP) {
        const double *row = red + lane * 33;
        double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
#pragma unroll
        for (int j = 0; j < 32; j += 4) { s0 += row[j]; s1 += row[j + 1]; s2 += row[j + 2]; s3 += row[j + 3]; }
        const int o = h * RP + lane;
        fresh[h] = stage_combine_modal(op, cf[o], (s0 + s1) + (s2 + s3), need_u0 ? buf[St::oU0 + o] : 0.0);
        out[cell * NO + o] = fresh[h];
      }
      __syncwarp();
    }
    // The next stage's face kernel evaluates the Taylor form S^t coeff of the NEW coefficients (modal_taylor_kernel):
    // S is still staged and the coefficients are in this warp's registers, so it is formed here and the separate
    // pass over S and the column is skipped.
    if (tay_out) {
#pragma unroll
      for (int h = 0; h < NO / RP; ++h)
        if (lane < RP) red[h * RP + lane] = fresh[h];
      __syncwarp();
      // S is stored with exact zeros above its diagonal: the sums run over all n, fully unrolled and in four
      // independent chains (this tail is pure latency for a warp that has nothing else in flight)
#pragma unroll
      for (int r = 0; r < (NO + 31) / 32; ++r) {
        const int o = lane + 32 * r, oo = o < NO ? o : 0, m = oo / 5, k = oo - m * 5;
        double t0 = 0, t1 = 0;
#pragma unroll
        for (int n = 0; n < N; n += 2) {
          t0 += S[n * N + m] * red[n * 5 + k];
          t1 += S[(n + 1) * N + m] * red[(n + 1) * 5 + k];
        }
        if (o < NO) tay_out[cell * NO + o] = t0 + t1;
      }
    }
#pragma unroll
    for (int slot = 0; slot < kSlots; ++slot) { gp[slot] = gp_next[slot]; sharer[slot] = sharer_next[slot]; }
    __syncwarp();  // this buffer is refilled two iterations from now; all lanes must be done reading it
  }
}

// ---- the fused residual + Runge-Kutta update on shape-class tables, contractions on the FP64 tensor cores -------------
// On a mesh whose cells repeat up to a translation (HostTables::n_classes > 0: every undistorted box, the Kuhn
// tetrahedra of one) phi_n(x_q), w_q grad phi_n(x_q) and phi_n at the face points are numbers of the cell's CLASS, so
// the volume and face terms of dg/general.hpp:50-105 are three small dense products against constant matrices,
//   U_q   = A1 (points x modes)        . C   (modes x columns)            u at the quadrature points
//   out  += A2 (modes x (point, d))    . F   ((point, d) x columns)       sum_q w_q F_d(u_q) d phi_n / d x_d
//   out  += A3 (modes x face points)   . F*  (face points x columns)      sum_j (-+ w f*)_j phi_n(x_j)
// whose columns are (component k, cell b) of a BATCH of 8 cells of one class: one column tile of an m8n8k4 DMMA per
// component.  north_star asks for tensor cores "only if the small dense contraction is better on DMMA than on the FP64
// pipe": the DFMA form of the same tables (one lane per (mode, slice), operands from shared memory) is bound by the
// shared-memory pipe -- every broadcast operand costs a wavefront -- at 0.997 ms per launch for 64^3 p = 2, slower than
// the register kernel above (0.867 ms); an MMA fragment is one DISTINCT double per lane, so the same contraction needs
// a twentieth of the wavefronts.  One warp per batch, persistent warps, no block barrier, fixed summation order.
__device__ __forceinline__ void dmma(double (&c)[2], double a, double b) {
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}

#ifndef MCFD_MODAL_MMA_WARPS
#define MCFD_MODAL_MMA_WARPS 4
#endif
#ifndef MCFD_MODAL_MMA_MINB
#define MCFD_MODAL_MMA_MINB 4
#endif
constexpr int kMmaWarps = MCFD_MODAL_MMA_WARPS;
template <int PE> struct MmaShape {   // the padded operand shapes; setup.cpp (BuildShapeClasses) lays the tables out the same way
  using M = ModalShape<PE>;
  static constexpr int N = M::N, NQ = M::NQ, FQ = M::FQ, NF = M::NF, NO = N * 5;
  static constexpr int QT = (NQ + 7) / 8;            // tiles of 8 quadrature points
  static constexpr int K1 = (N + 3) / 4 * 4;         // modes, padded to whole k-steps
  static constexpr int MT = (N + 7) / 8;             // tiles of 8 modes
  static constexpr int R = QT * 24;                  // (point, direction) rows of the volume product
  static constexpr int KC = (2 * FQ + 3) / 4 * 4;    // face points of two sides, padded to whole k-steps
  static constexpr int NCH = NF / 2, J = NCH * KC;
  static constexpr int kRow = 44;                    // row stride of the exchange buffer: 5 components x 8 cells + 4 (bank spread)
  static constexpr int kBufRows = KC > 24 ? KC : 24;
  static constexpr int kBuf = kBufRows * kRow > 8 * NO ? kBufRows * kRow : 8 * NO;
  static constexpr int oC = 0, oBuf = 8 * NO, oI = oBuf + kBuf, kDoubles = oI + 4 + 4 * NF;   // ints: 8 cell ids + 8 NF packed faces
};

template <int PE>
__global__ void __launch_bounds__(kMmaWarps * 32, MCFD_MODAL_MMA_MINB) modal_cell_mma_kernel(DeviceTables tab, double gamma,
                                                                                            const double *__restrict__ u,
                                                                                            const double *__restrict__ fstar,
                                                                                            const double *__restrict__ u0,
                                                                                            double *__restrict__ out,
                                                                                            double *__restrict__ tay_out, StageOp op) {
  using Sh = MmaShape<PE>;
  constexpr int N = Sh::N, NQ = Sh::NQ, FQ = Sh::FQ, NF = Sh::NF, NO = Sh::NO, QT = Sh::QT, K1 = Sh::K1, MT = Sh::MT,
                R = Sh::R, KC = Sh::KC, NCH = Sh::NCH, J = Sh::J, kRow = Sh::kRow;
  extern __shared__ __align__(16) double smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;   // the fragment coordinates of an m8n8k4 DMMA: group id, thread in group
  double *sC = smem + warp * Sh::kDoubles + Sh::oC, *buf = smem + warp * Sh::kDoubles + Sh::oBuf;
  int *sId = reinterpret_cast<int *>(smem + warp * Sh::kDoubles + Sh::oI), *sGp0 = sId + 8;
  const int64_t f_bnd = tab.n_local_faces + tab.n_ghost_faces;
  const bool need_u0 = op.kind >= 1 && op.kind <= 3;
  const int stride = int(gridDim.x) * kMmaWarps;
  for (int bid = int(blockIdx.x) * kMmaWarps + warp; bid < tab.n_batches; bid += stride) {
    // ---- which class, which 8 cells ----
    int cls = 0, first = bid;
    for (;; ++cls) {
      const int nb = (__ldg(tab.class_ptr + cls + 1) - __ldg(tab.class_ptr + cls) + 7) >> 3;
      if (first < nb) break;
      first -= nb;
    }
    const int c_begin = __ldg(tab.class_ptr + cls) + first * 8, c_end = __ldg(tab.class_ptr + cls + 1);
    if (lane < 8) sId[lane] = c_begin + lane < c_end ? __ldg(tab.class_cells + c_begin + lane) : -1;
    __syncwarp();
    for (int i = lane; i < 8 * NF; i += 32) {
      const int b = i / NF, side = i - b * NF, cell = sId[b];
      int packed = -1;
      if (cell >= 0) {
        const int f = __ldg(tab.cell_face + int64_t(cell) * 6 + side);
        // a boundary name that was never registered contributes nothing (fem.hpp:185-226)
        const bool flux = f >= 0 && !(f >= f_bnd && tab.groups[tab.bface_group[f - f_bnd]].kind < 0);
        if (flux) packed = ((f * FQ) << 1) | (__ldg(tab.cell_code + int64_t(cell) * 6 + side) & 1);
      }
      sGp0[i] = packed;
    }
    // the 8 coefficient blocks -> shared memory: every load of the batch in flight before the first store
    constexpr int RU = (8 * NO + 31) / 32;
    {
      double tmp[RU];
#pragma unroll
      for (int r = 0; r < RU; ++r) {
        const int idx = lane + 32 * r, b = idx / NO, o = idx - b * NO;
        const int cell = idx < 8 * NO ? sId[b] : -1;
        tmp[r] = cell >= 0 ? u[int64_t(cell) * NO + o] : 0.0;
      }
#pragma unroll
      for (int r = 0; r < RU; ++r)
        if (lane + 32 * r < 8 * NO) sC[lane + 32 * r] = tmp[r];
    }
    // the common fluxes of the batch are needed after the volume term: ask for their lines now (L2)
    for (int i = lane; i < 8 * NF * 5; i += 32) {
      const int bs = i / 5, k = i - bs * 5, packed = sGp0[bs];
      if (packed >= 0) {
        const int64_t gp = int64_t(packed >> 1);
        const double *fs = fstar + ((gp >> 5) * (32 * MCFD_K) + (gp & 31)) + k * 32;
        asm volatile("prefetch.global.L2 [%0];" ::"l"(fs));
        if (((gp & 31) + FQ) > 32) asm volatile("prefetch.global.L2 [%0];" ::"l"(fs + (32 * MCFD_K - (gp & 31))));
      }
    }
    __syncwarp();
    const double *A1 = tab.mm_A1 + size_t(cls) * QT * 8 * K1, *A2 = tab.mm_A2 + size_t(cls) * MT * 8 * R,
                 *A3 = tab.mm_A3 + size_t(cls) * MT * 8 * J;
    double acc[MT][5][2];
#pragma unroll
    for (int i = 0; i < MT; ++i)
#pragma unroll
      for (int k = 0; k < 5; ++k) acc[i][k][0] = acc[i][k][1] = 0.0;

    // ---- volume term, 8 quadrature points at a time (dg/general.hpp:50-61) ----
#pragma unroll 1
    for (int qt = 0; qt < QT; ++qt) {
      double uq[5][2];
#pragma unroll
      for (int k = 0; k < 5; ++k) uq[k][0] = uq[k][1] = 0.0;
#pragma unroll
      for (int s = 0; s < K1 / 4; ++s) {   // U_q = A1 . C : row q = qt * 8 + g, column (k, cell g)
        const double a = __ldg(A1 + (qt * 8 + g) * K1 + 4 * s + t);
        const int n = 4 * s + t;
#pragma unroll
        for (int k = 0; k < 5; ++k) dmma(uq[k], a, n < N ? sC[g * NO + n * 5 + k] : 0.0);
      }
      // this lane now holds u at point q = qt * 8 + g of the cells 2 t and 2 t + 1: F(u) (fem.hpp:264-278 -> types.hpp:279-290)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int b = 2 * t + e;
        double F[3][5];
        if (qt * 8 + g < NQ && sId[b] >= 0) {
          const Cons c{ { uq[0][e], uq[1][e], uq[2][e], uq[3][e], uq[4][e] } };
          const Prim s = cons_to_prim(c, gamma);
          const double vel[3] = { s.u, s.v, s.w };
          const double rho_h0 = c.c[4] + s.p;
#pragma unroll
          for (int d = 0; d < 3; ++d) {
            F[d][0] = c.c[1 + d]; F[d][1] = c.c[1] * vel[d]; F[d][2] = c.c[2] * vel[d]; F[d][3] = c.c[3] * vel[d];
            F[d][4] = rho_h0 * vel[d];
            F[d][1 + d] += s.p;
          }
        } else {
#pragma unroll
          for (int d = 0; d < 3; ++d)
#pragma unroll
            for (int k = 0; k < 5; ++k) F[d][k] = 0.0;
        }
#pragma unroll
        for (int d = 0; d < 3; ++d)
#pragma unroll
          for (int k = 0; k < 5; ++k) buf[(g * 3 + d) * kRow + k * 8 + b] = F[d][k];
      }
      __syncwarp();
#pragma unroll
      for (int s = 0; s < 6; ++s) {        // out += A2 . F over the 24 (point, direction) rows of this tile
        double a2[MT];
#pragma unroll
        for (int i = 0; i < MT; ++i) a2[i] = __ldg(A2 + (i * 8 + g) * R + qt * 24 + 4 * s + t);
#pragma unroll
        for (int k = 0; k < 5; ++k) {
          const double f = buf[(4 * s + t) * kRow + k * 8 + g];
#pragma unroll
          for (int i = 0; i < MT; ++i) dmma(acc[i][k], a2[i], f);
        }
      }
      __syncwarp();
    }

    // ---- faces, two sides at a time (dg/general.hpp:69-212): the flux w_q f* was stored once per point by modal_face_kernel ----
#pragma unroll 1
    for (int ch = 0; ch < NCH; ++ch) {
      {
        constexpr int RF = KC / 4;   // KC * 8 entries, 32 lanes: every gather of the chunk in flight before the first store
        double v[RF][5];
#pragma unroll
        for (int r = 0; r < RF; ++r) {
          const int idx = lane + 32 * r, jl = idx >> 3, b = idx & 7;
#pragma unroll
          for (int k = 0; k < 5; ++k) v[r][k] = 0.0;
          if (jl < 2 * FQ) {
            const int sl = jl / FQ, p = jl - sl * FQ, packed = sGp0[b * NF + ch * 2 + sl];
            if (packed >= 0) {
              const double sign = (packed & 1) ? 1.0 : -1.0;  // dg/general.hpp:82-83 : holder -=, sharer +=
              const int64_t gp = int64_t(packed >> 1) + p;
              const double *fs = fstar + ((gp >> 5) * (32 * MCFD_K) + (gp & 31));  // FstarIndex()
#pragma unroll
              for (int k = 0; k < 5; ++k) v[r][k] = sign * __ldg(fs + k * 32);
            }
          }
        }
#pragma unroll
        for (int r = 0; r < RF; ++r) {
          const int idx = lane + 32 * r, jl = idx >> 3, b = idx & 7;
#pragma unroll
          for (int k = 0; k < 5; ++k) buf[jl * kRow + k * 8 + b] = v[r][k];
        }
      }
      __syncwarp();
#pragma unroll
      for (int s = 0; s < KC / 4; ++s) {
        double a3[MT];
#pragma unroll
        for (int i = 0; i < MT; ++i) a3[i] = __ldg(A3 + (i * 8 + g) * J + ch * KC + 4 * s + t);
#pragma unroll
        for (int k = 0; k < 5; ++k) {
          const double f = buf[(4 * s + t) * kRow + k * 8 + g];
#pragma unroll
          for (int i = 0; i < MT; ++i) dmma(acc[i][k], a3[i], f);
        }
      }
      __syncwarp();
    }

    // ---- back to the column layout: acc[i][k][e] = residual of mode i * 8 + g, component k, cell 2 t + e ----
#pragma unroll
    for (int i = 0; i < MT; ++i) {
      const int n = i * 8 + g;
      if (n < N) {
#pragma unroll
        for (int k = 0; k < 5; ++k) {
          buf[(2 * t) * NO + n * 5 + k] = acc[i][k][0];
          buf[(2 * t + 1) * NO + n * 5 + k] = acc[i][k][1];
        }
      }
    }
    __syncwarp();
    // temporal/ode.hpp:119-131, rk.hpp : the stage combination
    {
      double v0[RU];
#pragma unroll
      for (int r = 0; r < RU; ++r) {
        const int idx = lane + 32 * r, b = idx / NO, o = idx - b * NO;
        const int cell = idx < 8 * NO ? sId[b] : -1;
        v0[r] = (need_u0 && cell >= 0) ? u0[int64_t(cell) * NO + o] : 0.0;
      }
#pragma unroll
      for (int r = 0; r < RU; ++r) {
        const int idx = lane + 32 * r, b = idx / NO, o = idx - b * NO;
        const int cell = idx < 8 * NO ? sId[b] : -1;
        if (cell >= 0) {
          const double fresh = stage_combine_modal(op, sC[idx], buf[idx], v0[r]);
          out[int64_t(cell) * NO + o] = fresh;
          sC[idx] = fresh;
        }
      }
    }
    __syncwarp();
    // the Taylor form S^t coeff of the NEW coefficients for the next stage's face kernel (see modal_taylor_kernel)
    if (tay_out) {
      const double *S = tab.cls_S + size_t(cls) * N * N;
      for (int b = 0; b < 8; ++b) {
        const int cell = sId[b];
        if (cell < 0) continue;
        for (int o = lane; o < NO; o += 32) {
          const int m = o / 5, k = o - m * 5;
          double t0 = 0;
          for (int n = m; n < N; ++n) t0 = fma(__ldg(S + n * N + m), sC[b * NO + n * 5 + k], t0);
          tay_out[int64_t(cell) * NO + o] = t0;
        }
      }
    }
    __syncwarp();  // the scratch is rewritten for the warp's next batch
  }
}

// ---- Cell::Approximate for the modal basis: coeff = int f(x) phi(x)^t  (projection.hpp:25-39) ----------
template <int PE>
__global__ void __launch_bounds__(kWarpsPerBlock * 32) modal_approximate_kernel(DeviceTables tab, int func,
                                                                                 FuncParams params, double t,
                                                                                 double *__restrict__ u) {
  using M = ModalShape<PE>;
  using W = WarpScratch<M::N>;
  constexpr int P = M::P, N = M::N, NQ = M::NQ;
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
template <int PE>
__device__ __forceinline__ void warp_smoothness(const DeviceTables &tab, int64_t cell, const double *center,
                                                const double *ct, double w1, double w2, double w3, int lane,
                                                double (&beta)[5]) {
  constexpr int P = ModalShape<PE>::P, N = ModalShape<PE>::N, NQ = ModalShape<PE>::NQ;
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

// The smoothness indicator is a quadratic form of the polynomial (weno.hpp:20-34: an integral of squared
// derivatives), so for a projection with components k = 0..4 every LINEAR COMBINATION  sum_k l_k u_k  has
//   beta = sum_{a,b} l_a l_b G[a][b],   G[a][b] = int sum_m wgt_m  d_m u_a  d_m u_b .
// weno::Eigen asks for beta of the five characteristic combinations on each of the (<= 6) adjacent faces
// (weno.hpp:216-235); with G formed once per candidate those are 15-term sums instead of fresh quadratures.
// G is stored packed (a <= b, row by row; 15 entries, stride 16).  P <= 2 only (NQ <= 32, 5 N registers).
template <int PE>
__device__ __forceinline__ void warp_gram(const DeviceTables &tab, int64_t cell, const double *center, const double *ct,
                                          double w1, double w2, double w3, int lane, double *G) {
  constexpr int P = ModalShape<PE>::P, N = ModalShape<PE>::N, NQ = ModalShape<PE>::NQ;
  static_assert(NQ <= 32 && P <= 2, "one quadrature point per lane");
  double g[16];
#pragma unroll
  for (int e = 0; e < 16; ++e) g[e] = 0.0;
  if (lane < NQ) {
    const double *xw = tab.m_qxyzw + (cell * NQ + lane) * 4;
    const double dx = __ldg(xw) - center[0], dy = __ldg(xw + 1) - center[1], dz = __ldg(xw + 2) - center[2];
    const double w = __ldg(xw + 3);
    double r[5][N];
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      double c[N];
#pragma unroll
      for (int m = 0; m < N; ++m) c[m] = ct[m * 5 + k];
      taylor_pdv<P>(dx, dy, dz, c, r[k]);
    }
    int e = 0;
#pragma unroll
    for (int a = 0; a < 5; ++a)
#pragma unroll
      for (int b = a; b < 5; ++b, ++e) {
        double s1 = r[a][1] * r[b][1] + r[a][2] * r[b][2] + r[a][3] * r[b][3], s2 = 0;
        if (P >= 2) {
#pragma unroll
          for (int m = 4; m < 10; ++m) s2 += r[a][m] * r[b][m];
        }
        g[e] = (P == 1 ? s1 : s1 * w1 + s2 * w2) * w;
      }
  }
  butterfly_step<8>(g, lane); butterfly_step<4>(g, lane); butterfly_step<2>(g, lane); butterfly_step<1>(g, lane);
  g[0] += __shfl_xor_sync(kFull, g[0], 16);
  if (lane < 15) G[lane] = g[0];
}

// ---- weno::Eigen::IsNotSmooth (weno.hpp:152-180) for every cell of [cell_begin, cell_end) -----------------
// Eight lanes per cell: lane `sub` evaluates neighbour `sub`'s polynomial at my centre -- from its Taylor form
// tay = S^t coeff (see modal_taylor_kernel; 20 numbers per neighbour instead of S and the coefficients) -- and the
// jumps are summed in the reference's (adjacency) order.  No shared memory and few registers, so the detector runs at full occupancy
// instead of inside the limiter warp (ncu r01e: 1.86 ms per pass at 12 % occupancy with 6 of 32 lanes busy).
template <int P>
__global__ void __launch_bounds__(256) modal_detect_kernel(DeviceTables tab, const double *__restrict__ u,
                                                           const double *__restrict__ tay,
                                                           int32_t *__restrict__ flags, int32_t *__restrict__ list,
                                                           int32_t *__restrict__ count,
                                                           unsigned long long *__restrict__ total, int64_t cell_begin,
                                                           int64_t cell_end) {
  constexpr int N = ModalShape<P>::N, NK = N * 5;
  const int64_t t = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  const int sub = threadIdx.x & 7;
  const int64_t wanted = cell_begin + (t >> 3);
  const bool live = wanted < cell_end;
  const int64_t cell = live ? wanted : cell_end - 1;  // every lane takes part in the shuffles
  const int a0 = tab.adj_ptr[cell], n_adj = tab.adj_ptr[cell + 1] - a0;
  const double *center = tab.m_center + cell * 3, *coeff = u + cell * NK;
  const double s00 = tab.m_S[cell * N * N];
  const double my_avg0 = coeff[0] * s00, my_avg4 = coeff[4] * s00;  // projection.hpp:41-47
  // my polynomial at my own centre: T(0) = e_0, so it is row 0 of the Taylor form (components rho and E)
  const double mine[2] = { tay[cell * NK], tay[cell * NK + 4] };
  double diff[2] = { 0, 0 }, their_avg[2] = { 0, 0 };
  if (sub < n_adj) {
    const int64_t adj = tab.adj_cell[a0 + sub];
    const double *ac = tab.m_center + adj * 3, *at = tay + adj * NK;
    double T[N];
    taylor_value<P>(center[0] - ac[0], center[1] - ac[1], center[2] - ac[2], T);
    double theirs0 = 0, theirs4 = 0;
#pragma unroll
    for (int m = 0; m < N; ++m) { theirs0 += at[m * 5] * T[m]; theirs4 += at[m * 5 + 4] * T[m]; }
    diff[0] = fabs(mine[0] - theirs0); diff[1] = fabs(mine[1] - theirs4);
    const double t00 = tab.m_S[adj * N * N];
    their_avg[0] = fabs(u[adj * NK] * t00); their_avg[1] = fabs(u[adj * NK + 4] * t00);
  }
  double sum[2] = { 0, 0 }, big[2] = { fmax(1e-9, fabs(my_avg0)), fmax(1e-9, fabs(my_avg4)) };
#pragma unroll
  for (int a = 0; a < 8; ++a)  // the reference's order of summation
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const double d = __shfl_sync(kFull, diff[c], a, 8), v = __shfl_sync(kFull, their_avg[c], a, 8);
      if (a < n_adj) { sum[c] += d; big[c] = fmax(big[c], v); }
    }
  double divisor = pow(tab.m_volume[cell], (P + 1.0) / 2.0 / 3);
  divisor *= n_adj;
  const double reference = P < 3 ? 1.0 : 3.0;
  const bool troubled = sum[0] / big[0] / divisor > reference || sum[1] / big[1] / divisor > reference;
  if (live && sub == 0) {
    flags[cell] = troubled ? 1 : 0;
    // the order of the list does not reach the results: every listed cell is rebuilt from the old data into `fresh`
    if (troubled) {
      list[atomicAdd(count, 1)] = int32_t(cell);
      atomicAdd(total, 1ULL);  // running number of rebuilds (what the reference harness counts per Reconstruct)
    }
  }
}

// One warp rebuilds one cell (weno.hpp:61-126 Lazy, :181-263 Eigen) into `fresh`.
template <int PE>
__device__ __forceinline__ void limit_one_cell(const DeviceTables &tab, const LimiterOp &lim, double gamma,
                                               const double *__restrict__ u, double *__restrict__ fresh,
                                               int32_t *__restrict__ flags, int64_t cell, double *smem, int warp,
                                               int lane) {
  using M = ModalShape<PE>;
  using W = WarpScratch<M::N>;
  constexpr int P = M::P, N = M::N, NQ = M::NQ, FQ = M::FQ, NK = N * 5;
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

  // On a repetitive mesh (shape classes, setup.cpp) the re-projection of the neighbour across a side is one constant
  // N x N matrix per (class, side) and the smoothness a quadratic form of the Taylor coefficients: no quadrature here.
  const bool tables = tab.n_classes > 0 && tab.cls_P != nullptr;
  const int cls = tables ? __ldg(tab.cell_class + cell) : 0;
  // ---- Borrow (weno.hpp:193-203): neighbours re-projected over MY quadrature points, shifted to my average
  for (int a = 0; a < n_adj; ++a) {
    const int64_t adj = tab.adj_cell[a0 + a];
    if (tables) {
      const int f = tab.adj_face[a0 + a] / FQ;
      int side = 0;
      while (side < M::NF - 1 && __ldg(tab.cell_face + cell * 6 + side) != f) ++side;
      const double *Pm = tab.cls_P + (size_t(cls) * 6 + side) * N * N, *ca = u + adj * NK;
      for (int o = lane; o < NK; o += 32) {
        const int n = o / 5, k = o - n * 5;
        double v = 0;
#pragma unroll
        for (int m = 0; m < N; ++m) v = fma(__ldg(Pm + n * N + m), ca[m * 5 + k], v);
        if (n == 0) v += (my_avg[k] - v * s00) / s00;  // projection.hpp:282-286
        s_cand[a * NK + o] = v;
      }
      continue;
    }
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
      warp_smoothness<PE>(tab, cell, center, taylor_of(i), w1, w2, w3, lane, beta);
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

  // P <= 2: Gram matrices of the candidates once (see warp_gram), then beta = l^t G l per combination
  double *s_G = sA;  // [7][16], the contraction scratch is free after Borrow
  if constexpr (P <= 2) {
    if (tables) {  // G[a][b] = ct_a^t W ct_b : y = W ct (N x 5), then the 15 products
      const double *Wm = tab.cls_W + size_t(cls) * N * N;
      double *sY = sB;   // free after Borrow
      for (int i = 0; i <= n_adj; ++i) {
        const double *ct = s_tay + i * NK;
        for (int o = lane; o < NK; o += 32) {
          const int j = o / 5, k = o - j * 5;
          double v = 0;
#pragma unroll
          for (int m = 0; m < N; ++m) v = fma(__ldg(Wm + j * N + m), ct[m * 5 + k], v);
          sY[o] = v;
        }
        __syncwarp();
        if (lane < 15) {
          int a = 0, b = lane;   // packed (a <= b), row by row
          while (b >= 5 - a) { b -= 5 - a; ++a; }
          b += a;
          double v = 0;
#pragma unroll
          for (int j = 0; j < N; ++j) v = fma(ct[j * 5 + a], sY[j * 5 + b], v);
          s_G[i * 16 + lane] = v;
        }
        __syncwarp();
      }
    } else {
      for (int i = 0; i <= n_adj; ++i) warp_gram<PE>(tab, cell, center, s_tay + i * NK, w1, w2, w3, lane, s_G + i * 16);
      __syncwarp();
    }
  }
  auto weights_from_gram = [&](const double *L) {  // L == nullptr: the components themselves (weno::Lazy)
    for (int p = lane; p < (n_adj + 1) * 5; p += 32) {
      const int i = p / 5, k = p - i * 5;
      const double *G = s_G + i * 16;
      double beta = 0;
      int e = 0;
#pragma unroll
      for (int a = 0; a < 5; ++a)
#pragma unroll
        for (int b = a; b < 5; ++b, ++e) {
          const double la = L ? L[k * 5 + a] : (a == k ? 1.0 : 0.0), lb = L ? L[k * 5 + b] : (b == k ? 1.0 : 0.0);
          const double term = la * lb * G[e];
          beta += a == b ? term : 2 * term;
        }
      double w = lim.w0;
      if (i == n_adj) { w *= -n_adj; w += 1.0; }
      double b2 = beta + lim.eps;
      b2 *= b2;
      s_wgt[p] = w / b2;
    }
    __syncwarp();
    if (lane < 5) {
      double sum = 0;
      for (int i = 0; i <= n_adj; ++i) sum += s_wgt[i * 5 + lane];
      for (int i = 0; i <= n_adj; ++i) s_wgt[i * 5 + lane] /= sum;
    }
    __syncwarp();
  };

  if (lim.kind == MCFD_LIMITER_LAZY) {  // weno.hpp:97-126
    if constexpr (P <= 2) weights_from_gram(nullptr);
    else nonlinear_weights([&](int i) { return static_cast<const double *>(s_tay + i * NK); });
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
    if constexpr (P <= 2) {
      weights_from_gram(s_L);
    } else {
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
    }
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

// weno::Lazy limits every cell of [cell_begin, cell_end) (weno.hpp:58-60).  weno::Eigen rebuilds the cells
// modal_detect_kernel listed: the warps of a fixed-size grid take list entries in turn, so every resident warp has
// a troubled cell to work on (with one warp per cell of the range and an early exit, ncu showed ~1 busy warp per
// scheduler: 5 % of the cells are troubled and the big per-warp scratch allows only 3 blocks per SM).
template <int PE>
__global__ void __launch_bounds__(kWarpsPerBlock * 32) modal_limit_kernel(DeviceTables tab, LimiterOp lim, double gamma,
                                                                           const double *__restrict__ u,
                                                                           double *__restrict__ fresh,
                                                                           int32_t *__restrict__ flags,
                                                                           const int32_t *__restrict__ list,
                                                                           const int32_t *__restrict__ count,
                                                                           int64_t cell_begin, int64_t cell_end) {
  extern __shared__ __align__(16) double smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t first = int64_t(blockIdx.x) * kWarpsPerBlock + warp, stride = int64_t(gridDim.x) * kWarpsPerBlock;
  const int64_t n = list ? int64_t(*count) : cell_end - cell_begin;
  for (int64_t i = first; i < n; i += stride) {
    const int64_t cell = list ? int64_t(list[i]) : cell_begin + i;
    limit_one_cell<PE>(tab, lim, gamma, u, fresh, flags, cell, smem, warp, lane);
    __syncwarp();  // the warp's scratch is reused by its next cell
  }
}

// part.hpp:1329-1332 : only after every troubled cell of the set has been rebuilt from the OLD data
__global__ void modal_writeback_kernel(const double *__restrict__ fresh, const int32_t *__restrict__ flags, int n_basis,
                                       const double *__restrict__ S, double *__restrict__ u, double *__restrict__ tay,
                                       int64_t cell_begin, int64_t cell_end) {
  const int fields = n_basis * 5;
  const int64_t idx = cell_begin * fields + int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= cell_end * fields) return;
  const int64_t cell = idx / fields;
  if (!flags[cell]) return;
  u[idx] = fresh[idx];
  // keep the Taylor form S^t coeff of the rebuilt cell current (it is read by the detector of the next set and by
  // the next stage's face kernel)
  const int o = int(idx - cell * fields), m = o / 5, k = o - m * 5;
  const double *Sc = S + cell * n_basis * n_basis, *c = fresh + cell * fields;
  double t = 0;
  for (int n = m; n < n_basis; ++n) t += Sc[n * n_basis + m] * c[n * 5 + k];
  tay[idx] = t;
}

// the same for weno::Eigen, over the compact list of rebuilt cells instead of a flag test per coefficient of the mesh
// (5 % of the cells are troubled in the shock workloads: the flag pass was 0.055 ms per launch at 64^3 for nothing)
__global__ void modal_writeback_list_kernel(const double *__restrict__ fresh, const int32_t *__restrict__ list,
                                            const int32_t *__restrict__ count, int n_basis, const double *__restrict__ S,
                                            double *__restrict__ u, double *__restrict__ tay) {
  const int fields = n_basis * 5;
  const int64_t total = int64_t(*count) * fields;
  for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += int64_t(gridDim.x) * blockDim.x) {
    const int64_t cell = list[i / fields];
    const int o = int(i % fields), m = o / 5, k = o - m * 5;
    const int64_t idx = cell * fields + o;
    u[idx] = fresh[idx];
    const double *Sc = S + cell * n_basis * n_basis, *c = fresh + cell * fields;
    double t = 0;
    for (int n = m; n < n_basis; ++n) t += Sc[n * n_basis + m] * c[n * 5 + k];
    tay[idx] = t;
  }
}

template <class Kernel>
cudaError_t set_smem(Kernel kernel, size_t bytes) {
  return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(bytes));
}

}  // namespace

// ---- launchers ----------------------------------------------------------------------------------------
// shape code of a part: degree + 10 * element type (ModalShape); one instantiation of the shape-dependent kernels each
#define MCFD_SHAPE_SWITCH(tab, degree, CALL)                                  \
  switch ((degree) + 10 * (tab).elem) {                                        \
    case 1: return CALL(1);                                                    \
    case 2: return CALL(2);                                                    \
    case 3: return CALL(3);                                                    \
    case 11: return CALL(11);                                                  \
    case 12: return CALL(12);                                                  \
    case 13: return CALL(13);                                                  \
    default: return cudaErrorInvalidValue;                                     \
  }

template <int P>
static cudaError_t launch_modal_face(const DeviceTables &tab, int riemann, double gamma, double t, const double *tay,
                                     double *fstar, int64_t face_begin, int64_t face_end, cudaStream_t stream) {
  constexpr int FQ = ModalShape<P>::FQ;
  const int64_t n_points = (face_end - face_begin) * FQ, point_begin = face_begin * FQ;
  const int threads = MCFD_MODAL_FACE_THREADS;
  const unsigned blocks = unsigned((n_points + threads - 1) / threads);
  if (face_begin >= tab.n_local_faces + tab.n_ghost_faces) {
    modal_face_kernel<P, 0, true><<<blocks, threads, 0, stream>>>(tab, gamma, t, tay, fstar, point_begin, n_points);
    return cudaGetLastError();
  }
  switch (riemann) {
    case MCFD_RIEMANN_EXACT: modal_face_kernel<P, 0, false><<<blocks, threads, 0, stream>>>(tab, gamma, t, tay, fstar, point_begin, n_points); break;
    case MCFD_RIEMANN_ROE: modal_face_kernel<P, 1, false><<<blocks, threads, 0, stream>>>(tab, gamma, t, tay, fstar, point_begin, n_points); break;
    case MCFD_RIEMANN_HLLC: modal_face_kernel<P, 2, false><<<blocks, threads, 0, stream>>>(tab, gamma, t, tay, fstar, point_begin, n_points); break;
    case MCFD_RIEMANN_AUSM: modal_face_kernel<P, 3, false><<<blocks, threads, 0, stream>>>(tab, gamma, t, tay, fstar, point_begin, n_points); break;
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

cudaError_t LaunchModalTaylor(const DeviceTables &tab, int degree, const double *u, double *tay, int64_t cell_begin,
                              int64_t cell_end, cudaStream_t stream) {
  if (cell_end <= cell_begin) return cudaSuccess;
  const int n = degree == 1 ? 4 : degree == 2 ? 10 : 20;
  const int64_t entries = (cell_end - cell_begin) * n * 5;
  const unsigned blocks = unsigned((entries + 255) / 256);
  switch (degree) {
    case 1: modal_taylor_kernel<1><<<blocks, 256, 0, stream>>>(tab, u, tay, cell_begin, entries); break;
    case 2: modal_taylor_kernel<2><<<blocks, 256, 0, stream>>>(tab, u, tay, cell_begin, entries); break;
    case 3: modal_taylor_kernel<3><<<blocks, 256, 0, stream>>>(tab, u, tay, cell_begin, entries); break;
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

cudaError_t LaunchModalFaceFlux(const DeviceTables &tab, int degree, int riemann, double gamma, double t, const double *tay,
                                double *fstar, int64_t face_begin, int64_t face_end, cudaStream_t stream) {
  if (face_end <= face_begin) return cudaSuccess;
#define MCFD_CALL(PE) launch_modal_face<PE>(tab, riemann, gamma, t, tay, fstar, face_begin, face_end, stream)
  MCFD_SHAPE_SWITCH(tab, degree, MCFD_CALL)
#undef MCFD_CALL
}

template <int P>
static cudaError_t launch_modal_cell(const DeviceTables &tab, double gamma, const double *u, const double *fstar,
                                     const double *u0, double *out, double *tay_out, StageOp op, cudaStream_t stream) {
  const unsigned blocks = unsigned((tab.n_local + kWarpsPerBlock - 1) / kWarpsPerBlock);
  // MCFD_MODAL_TABLES=0 (tuning / A-B tests): never use the shape-class kernel
  static const bool tables_on = !(std::getenv("MCFD_MODAL_TABLES") && std::atoi(std::getenv("MCFD_MODAL_TABLES")) == 0);
  if (tab.n_classes > 0 && tables_on) {  // repetitive mesh: contractions against per-class tables on the FP64 tensor cores
    static std::atomic<int> resident_of[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    int resident = dev >= 0 && dev < 64 ? resident_of[dev].load(std::memory_order_acquire) : 0;
    constexpr size_t bytes = size_t(kMmaWarps) * MmaShape<P>::kDoubles * sizeof(double);
    if (!resident) {
      int sms = 0, per_sm = 0;
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
      cudaFuncSetAttribute(modal_cell_mma_kernel<P>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(bytes));
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, modal_cell_mma_kernel<P>, kMmaWarps * 32, bytes);
      resident = std::max(1, sms * std::max(1, per_sm));
      if (dev >= 0 && dev < 64) resident_of[dev].store(resident, std::memory_order_release);
    }
    const unsigned want = unsigned((tab.n_batches + kMmaWarps - 1) / kMmaWarps);
    modal_cell_mma_kernel<P><<<std::min<unsigned>(want, unsigned(resident)), kMmaWarps * 32, bytes, stream>>>(tab, gamma, u, fstar, u0, out,
                                                                                                             tay_out, op);
    return cudaGetLastError();
  }
  if constexpr (ModalShape<P>::P <= 2) {  // register-resident products + warp butterfly, persistent warps
    // the attribute and the occupancy belong to (function, device): one slot per device, for hosts that drive several
    // GPUs from one process
    static std::atomic<int> resident_of[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    int resident = dev >= 0 && dev < 64 ? resident_of[dev].load(std::memory_order_acquire) : 0;
    constexpr int NO = ModalShape<P>::N * 5, RP = NO <= 32 ? NO : NO / 2;
    constexpr size_t stage_bytes = size_t(kWarpsPerBlock) * (2 * CellStage<P>::kDoubles + (RP * 33 + 1) / 2 * 2) * sizeof(double);
    if (!resident) {
      int sms = 0, per_sm = 0;
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
      cudaFuncSetAttribute(modal_cell_kernel_reg<P>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(stage_bytes));
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, modal_cell_kernel_reg<P>, kWarpsPerBlock * 32, stage_bytes);
      resident = std::max(1, sms * std::max(1, per_sm));
      if (dev >= 0 && dev < 64) resident_of[dev].store(resident, std::memory_order_release);
    }
    const unsigned grid = std::min<unsigned>(blocks, unsigned(resident));
    modal_cell_kernel_reg<P><<<grid, kWarpsPerBlock * 32, stage_bytes, stream>>>(tab, gamma, u, fstar, u0, out, tay_out, op);
    return cudaGetLastError();
  }
  const size_t smem = size_t(kWarpsPerBlock) * WarpScratch<ModalShape<P>::N>::kDoubles * sizeof(double);
  cudaError_t e = set_smem(modal_cell_kernel<P>, smem);
  if (e != cudaSuccess) return e;
  modal_cell_kernel<P><<<blocks, kWarpsPerBlock * 32, smem, stream>>>(tab, gamma, u, fstar, u0, out, op);
  return cudaGetLastError();
}

cudaError_t LaunchModalCellUpdate(const DeviceTables &tab, int degree, double gamma, const double *u, const double *fstar,
                                  const double *u0, double *out, double *tay_out, StageOp op, cudaStream_t stream) {
  if (tab.n_local <= 0) return cudaSuccess;
#define MCFD_CALL(PE) launch_modal_cell<PE>(tab, gamma, u, fstar, u0, out, tay_out, op, stream)
  MCFD_SHAPE_SWITCH(tab, degree, MCFD_CALL)
#undef MCFD_CALL
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
#define MCFD_CALL(PE) launch_modal_approximate<PE>(tab, func, fp, t, u, stream)
  MCFD_SHAPE_SWITCH(tab, degree, MCFD_CALL)
#undef MCFD_CALL
}

template <int P>
static cudaError_t launch_modal_limit(const DeviceTables &tab, LimiterOp lim, double gamma, double *u, double *tay,
                                      double *fresh, int32_t *flags, int64_t cell_begin, int64_t cell_end, cudaStream_t stream) {
  constexpr int N = ModalShape<P>::N;
  const size_t smem = size_t(kWarpsPerBlock) * LimiterScratch<N>::kDoubles * sizeof(double);
  cudaError_t e = set_smem(modal_limit_kernel<P>, smem);
  if (e != cudaSuccess) return e;
  const int64_t n = cell_end - cell_begin;
  static std::atomic<int> resident_of[64] = {};  // per device, as in launch_modal_cell
  int dev = 0;
  cudaGetDevice(&dev);
  int resident = dev >= 0 && dev < 64 ? resident_of[dev].load(std::memory_order_acquire) : 0;
  if (!resident) {
    int sms = 0, per_sm = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, modal_limit_kernel<P>, kWarpsPerBlock * 32, smem);
    resident = std::max(1, sms * std::max(1, per_sm));
    if (dev >= 0 && dev < 64) resident_of[dev].store(resident, std::memory_order_release);
  }
  const unsigned blocks = unsigned(std::min<int64_t>((n + kWarpsPerBlock - 1) / kWarpsPerBlock, resident));
  // flags[n_local] is followed by the troubled-cell list [n_local], its length [1], a pad [1] and the running
  // 64-bit total of rebuilds (api.cpp allocates them together)
  int32_t *list = nullptr, *count = nullptr;
  if (lim.kind == MCFD_LIMITER_EIGEN) {
    list = flags + tab.n_local;
    count = flags + 2 * tab.n_local;
    e = cudaMemsetAsync(count, 0, sizeof(int32_t), stream);
    if (e != cudaSuccess) return e;
    auto *total = reinterpret_cast<unsigned long long *>(flags + 2 * tab.n_local + 2);
    modal_detect_kernel<ModalShape<P>::P><<<unsigned((n * 8 + 255) / 256), 256, 0, stream>>>(tab, u, tay, flags, list, count, total, cell_begin,
                                                                              cell_end);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
  }
  modal_limit_kernel<P><<<blocks, kWarpsPerBlock * 32, smem, stream>>>(tab, lim, gamma, u, fresh, flags, list, count,
                                                                         cell_begin, cell_end);
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  const int threads = 256;
  const int64_t total = n * N * 5;
  if (list) {
    const unsigned grid = unsigned(std::min<int64_t>((total + threads - 1) / threads, int64_t(resident) * 8));
    modal_writeback_list_kernel<<<grid, threads, 0, stream>>>(fresh, list, count, N, tab.m_S, u, tay);
  } else {
    modal_writeback_kernel<<<unsigned((total + threads - 1) / threads), threads, 0, stream>>>(fresh, flags, N, tab.m_S, u, tay, cell_begin,
                                                                                              cell_end);
  }
  return cudaGetLastError();
}

cudaError_t LaunchModalLimit(const DeviceTables &tab, int degree, LimiterOp lim, double gamma, double *u, double *tay,
                             double *fresh, int32_t *flags, int64_t cell_begin, int64_t cell_end, cudaStream_t stream) {
  if (cell_end <= cell_begin || lim.kind == MCFD_LIMITER_NONE) return cudaSuccess;
#define MCFD_CALL(PE) launch_modal_limit<PE>(tab, lim, gamma, u, tay, fresh, flags, cell_begin, cell_end, stream)
  MCFD_SHAPE_SWITCH(tab, degree, MCFD_CALL)
#undef MCFD_CALL
}

}  // namespace mcfd
