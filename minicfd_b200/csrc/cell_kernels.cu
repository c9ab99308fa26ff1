// cell_kernels.cu -- per-element residual assembly fused with the Runge-Kutta stage update.
//
// One thread per (cell, solution node); a block owns TC consecutive cells whose coefficients
// form one contiguous chunk of the column, moved HBM <-> shared memory by TMA bulk copies.
// Replaces (paths relative to /root/reference/include/mini):
//   spatial/fr/general.hpp:232-248      FR volume term: F^ = F(u) J*, R_q -= sum_k F^_k . grad phi_q[:,k]
//   spatial/dg/lobatto.hpp:86-151       DGSEM volume term + division by the (diagonal) mass matrix
//   spatial/fr/lobatto.hpp:228-237      FR face correction (f* scale - F(u).n^) g'
//   spatial/dg/lobatto.hpp:166-176      DGSEM face term -/+ f* w_f
//   temporal/ode.hpp:119-131, rk.hpp    u + dt R and the RK2/RK3 stage combinations
// The dense N x N gradient contraction of the reference has only 3Q non-zeros per node (tensor
// product of 1-D Lagrange bases); only those are evaluated.  The common flux f* comes from
// fstar[face][point][5] written by face_kernels.cu, gathered here for holder and sharer alike.
#include <algorithm>
#include <atomic>
#include <cstdlib>

#include "riemann.cuh"
#include "solver.hpp"

namespace mcfd {

template <int Q>
__device__ __forceinline__ void apply_perm(int code, int a, int b, int &a2, int &b2) {
  int x = (code & 2) ? Q - 1 - a : a, y = (code & 4) ? Q - 1 - b : b;
  a2 = (code & 1) ? y : x;
  b2 = (code & 1) ? x : y;
}

// riemann/euler/types.hpp:246-258,279-290 : inviscid flux matrix F[d][c] of a conservative state.
// ConservativeToPrimitive comes from riemann.cuh (ordinary states branch-free, others out of line).
__device__ __forceinline__ void flux_matrix(const double (&cv)[5], double gamma, double (&F)[3][5]) {
  const Cons c{ { cv[0], cv[1], cv[2], cv[3], cv[4] } };
  const Prim s = cons_to_prim(c, gamma);
  const double vel[3] = { s.u, s.v, s.w };
  const double rho_h0 = cv[4] + s.p;
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    F[d][0] = cv[1 + d];
    F[d][1] = cv[1] * vel[d];
    F[d][2] = cv[2] * vel[d];
    F[d][3] = cv[3] * vel[d];
    F[d][4] = rho_h0 * vel[d];
  }
  F[0][1] += s.p; F[1][2] += s.p; F[2][3] += s.p;
}

// The same contravariant fluxes F^_d = F(u) J*[:, d] * scale without forming F first: with the contravariant velocity
// v^_d = v . J*[:, d],   F^_d = (m . J*[:, d],  rho u v^_d + p J*[0][d],  rho v v^_d + p J*[1][d],  rho w v^_d + p J*[2][d],
// (E + p) v^_d), the mass flux from the momentum itself: 28 FP64 operations instead of 61 (flux matrix 16 + 5 x 3 x 3 product 45).  Same mathematics as
// fr/general.hpp:236-239 / polynomial/hexahedron.hpp:461-466, different rounding (a few ulp of the largest term).
__device__ __forceinline__ void contravariant_fluxes(const double (&cv)[5], double gamma, const double (&A)[9], double scale,
                                                     double (&G)[3][5]) {
  const Cons c{ { cv[0], cv[1], cv[2], cv[3], cv[4] } };
  const Prim s = cons_to_prim(c, gamma);
  const double rho_h0 = cv[4] + s.p;
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    const double vh = (s.u * A[0 * 3 + d] + s.v * A[1 * 3 + d] + s.w * A[2 * 3 + d]) * scale;
    const double ps = s.p * scale;
    // the mass flux is the conservative momentum itself (types.hpp:284-286), NOT rho v: the two differ when
    // ConservativeToPrimitive has zeroed a state with negative pressure (types.hpp:197-204)
    G[d][0] = (cv[1] * A[0 * 3 + d] + cv[2] * A[1 * 3 + d] + cv[3] * A[2 * 3 + d]) * scale;
    G[d][1] = cv[1] * vh + ps * A[0 * 3 + d];
    G[d][2] = cv[2] * vh + ps * A[1 * 3 + d];
    G[d][3] = cv[3] * vh + ps * A[2 * 3 + d];
    G[d][4] = rho_h0 * vh;
  }
}

__device__ __forceinline__ double stage_combine(const StageOp &op, double u, double r, double u0) {
  if (op.kind == 4) return r;
  double next = u + r * op.dt;  // ode.hpp:123-124 : r *= dt; u += r
  switch (op.kind) {
    case 1: return (u0 * 3 + next) / 4;    // rk.hpp:59-63
    case 2: {                              // rk.hpp:65-70 : (...) / 3, correctly rounded without a division
      const double x = next * 2 + u0, q = x * (1.0 / 3.0);
      return fma(fma(-q, 3.0, x), 1.0 / 3.0, q);
    }
    case 3: return (next + u0) * 0.5;      // rk.hpp:40-42
    default: return next;
  }
}

// ---- TMA (bulk async copy) + mbarrier helpers: cp.async.bulk moves a contiguous tile between HBM
// and shared memory without passing through registers (SASS: UBLKCP) ----
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return uint32_t(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "MCFD_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra MCFD_DONE;\n"
      "bra MCFD_WAIT;\n"
      "MCFD_DONE:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst_smem)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void *dst, const void *src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src_smem)), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// Tile = TC consecutive cells = one contiguous chunk of the column.  Data flow per block:
//   HBM --cp.async.bulk--> s_u (u tile), s_o (u0 tile)          one elected thread, mbarrier-tracked
//   registers: flux matrix, contravariant fluxes -> s_g         one thread per (cell, node)
//   registers: sparse divergence + face corrections (f* gathered from fstar) + RK combination -> s_o
//   s_o --cp.async.bulk--> HBM (out tile)                        one elected thread
// Several blocks are resident per SM (MINB), so one block's bulk copies overlap the others' FP64 work.
template <int Q, bool kFR, int TC, int MINB>
__global__ void __launch_bounds__(TC *Q *Q *Q, MINB) cell_update_kernel(DeviceTables tab, Basis basis, double gamma,
                                                                          const double *__restrict__ u,
                                                                          const double *__restrict__ fstar,
                                                                          const double *__restrict__ u0,
                                                                          double *__restrict__ out, StageOp op,
                                                                          int64_t cell_begin, int64_t cell_end) {
  constexpr int N = Q * Q * Q, FQ = Q * Q, CELL = N * 5, TILE = TC * CELL;
  extern __shared__ __align__(128) double smem[];
  double *s_u = smem;             // [TILE] coefficients of the tile
  double *s_o = smem + TILE;      // [TILE] u0 tile in, new coefficients out
  double *s_g = smem + 2 * TILE;  // [3][5][TC * N] contravariant fluxes, component-major: conflict-free line reads
  __shared__ double s_D[Q * Q], s_w[Q];
  __shared__ __align__(8) uint64_t s_bar;
  __shared__ int s_gp0[TC * 6];      // first flux point of the face on each side, -1: no flux there
  __shared__ uint8_t s_code[TC * 6];

  const int tid = threadIdx.x;
#ifdef MCFD_CELL_REVERSE
  // last cells first: the face kernel has just written the common fluxes in ascending order, so the tail of the
  // flux buffer (and of the column it read) is what the L2 still holds when this kernel starts
  const int64_t c0 = cell_begin + int64_t(gridDim.x - 1 - blockIdx.x) * TC;
#else
  const int64_t c0 = cell_begin + int64_t(blockIdx.x) * TC;
#endif
  const int ncell = int(min(int64_t(TC), cell_end - c0));
  const int64_t g0 = c0 * CELL;
  const bool need_u0 = op.kind >= 1 && op.kind <= 3;
  const uint32_t tile_bytes = uint32_t(ncell) * CELL * 8;
  // cp.async.bulk needs 16-byte aligned addresses and sizes (an odd number of p=2 cells is 8 mod 16)
  const bool bulk = ((g0 * 8) % 16 == 0) && (tile_bytes % 16 == 0);
  if (bulk) {
    if (tid == 0) {
      mbar_init(&s_bar, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      mbar_expect_tx(&s_bar, need_u0 ? 2 * tile_bytes : tile_bytes);
      bulk_g2s(s_u, u + g0, tile_bytes, &s_bar);
      if (need_u0) bulk_g2s(s_o, u0 + g0, tile_bytes, &s_bar);
    }
  } else {
    for (int idx = tid; idx < ncell * CELL; idx += TC * N) {
      s_u[idx] = u[g0 + idx];
      if (need_u0) s_o[idx] = u0[g0 + idx];
    }
  }
  if (tid < Q * Q) s_D[tid] = basis.D[tid / Q][tid % Q];
  if (tid < Q) s_w[tid] = basis.w[tid];
  if (tid < ncell * 6) {
    const int64_t f_bnd = tab.n_local_faces + tab.n_ghost_faces;
    const int f = __ldg(tab.cell_face + c0 * 6 + tid);
    // a boundary name that was never registered contributes nothing (fem.hpp:185-226)
    const bool flux = f >= 0 && !(f >= f_bnd && tab.groups[tab.bface_group[f - f_bnd]].kind < 0);
    s_gp0[tid] = flux ? f * FQ : -1;
    s_code[tid] = __ldg(tab.cell_code + c0 * 6 + tid);
  }

  const int cl = tid / N, node = tid % N;
  const int ijk[3] = { node / (Q * Q), (node / Q) % Q, node % Q };
  const bool active = cl < ncell;
  const int64_t cell = c0 + cl;
  double A[9], det = 1.0;
  if (active) {  // metric loads overlap the bulk copy
    const double *mp = tab.metrics + (tab.per_node_metrics ? (cell * N + node) * 10 : cell * 10);
#pragma unroll
    for (int e = 0; e < 9; ++e) A[e] = __ldg(mp + e);
    det = __ldg(mp + 9);
  }
  __syncthreads();  // s_bar initialised / plain copies done / s_D, s_w visible
  // The common fluxes this node will need are the kernel's longest stall (ncu r01d: 25 % of the samples sit on
  // their loads): ask for them now, into L1, while the tile is still in flight and the volume term not begun.
  if (active) {
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      const bool hi = ijk[d] == Q - 1;
      if (!hi && ijk[d] != 0) continue;
      const int side = hi ? (d == 0 ? 2 : d == 1 ? 3 : 5) : (d == 0 ? 4 : d == 1 ? 1 : 0);
      const int gp0 = s_gp0[cl * 6 + side];
      if (gp0 < 0) continue;
      const int code = s_code[cl * 6 + side];
      int a = d == 0 ? ijk[1] : ijk[0], b = d == 2 ? ijk[1] : ijk[2];
      if (code & 1) { int a2, b2; apply_perm<Q>(code >> 1, a, b, a2, b2); a = a2; b = b2; }
      const int64_t gp = int64_t(gp0) + a * Q + b;
      const double *fs = fstar + ((gp >> 5) * (32 * MCFD_K) + (gp & 31));
#pragma unroll
      for (int c = 0; c < 5; ++c) asm volatile("prefetch.global.L1 [%0];" ::"l"(fs + c * 32));
    }
  }
  if (bulk) mbar_wait(&s_bar, 0);

  double res[5];
  if (active) {
    double uu[5];
    const double rdet = fast_rcp(det);  // det J > 0 is checked at setup
#pragma unroll
    for (int c = 0; c < 5; ++c) {
      const double mine = s_u[tid * 5 + c];
      uu[c] = kFR ? mine * rdet : mine;
    }
    // FR: F^ = F J* (polynomial/hexahedron.hpp:461-466).  DGSEM: w_q F J^{-1} (dg/lobatto.hpp:86-98)
    const double scale = kFR ? 1.0 : (s_w[ijk[0]] * s_w[ijk[1]] * s_w[ijk[2]] * fabs(det)) * rdet;
    double G[3][5];
    contravariant_fluxes(uu, gamma, A, scale, G);
#pragma unroll
    for (int d = 0; d < 3; ++d)
#pragma unroll
      for (int c = 0; c < 5; ++c) s_g[d * TILE + c * (TC * N) + tid] = G[d][c];
  }
  __syncthreads();

  if (active) {
    double val[5] = { 0, 0, 0, 0, 0 };
    constexpr int NT = TC * N;
    const double *gx = s_g + cl * N, *gy = gx + TILE, *gz = gy + TILE;
#pragma unroll
    for (int m = 0; m < Q; ++m) {
      // FR  : derivative of basis m at my node    D[my][m]   (fr/general.hpp:240-246)
      // DG  : derivative of my basis at node m    D[m][my]   (dg/lobatto.hpp:146-149)
      // (read from the kernel parameter: a constant-bank load, not a shared-memory one)
      const double dx = kFR ? basis.D[ijk[0]][m] : basis.D[m][ijk[0]];
      const double dy = kFR ? basis.D[ijk[1]][m] : basis.D[m][ijk[1]];
      const double dz = kFR ? basis.D[ijk[2]][m] : basis.D[m][ijk[2]];
      const double *px = gx + ((m * Q + ijk[1]) * Q + ijk[2]);
      const double *py = gy + ((ijk[0] * Q + m) * Q + ijk[2]);
      const double *pz = gz + ((ijk[0] * Q + ijk[1]) * Q + m);
#pragma unroll
      for (int c = 0; c < 5; ++c) val[c] += dx * px[c * NT] + dy * py[c * NT] + dz * pz[c * NT];
    }
#pragma unroll
    for (int c = 0; c < 5; ++c) res[c] = kFR ? -val[c] : val[c];

    // faces this node lies on: direction d, low end -> sides {4,1,0}[d], high end -> {2,3,5}[d].
    // One pass per direction serves the low and the high face alike (2/3 of the lanes are busy at
    // p = 2); s_gp0 / s_code were filled once per block.
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      const bool hi = ijk[d] == Q - 1;
      if (!hi && ijk[d] != 0) continue;
      const int side = hi ? (d == 0 ? 2 : d == 1 ? 3 : 5) : (d == 0 ? 4 : d == 1 ? 1 : 0);
      const int gp0 = s_gp0[cl * 6 + side];
      if (gp0 < 0) continue;
      const int code = s_code[cl * 6 + side];
      int a = d == 0 ? ijk[1] : ijk[0], b = d == 2 ? ijk[1] : ijk[2];
      const bool sharer = code & 1;
      if (sharer) { int a2, b2; apply_perm<Q>(code >> 1, a, b, a2, b2); a = a2; b = b2; }
      const int64_t gp = int64_t(gp0) + a * Q + b;
      const double *fs = fstar + ((gp >> 5) * (32 * MCFD_K) + (gp & 31));
      const double nx = A[0 * 3 + d], ny = A[1 * 3 + d], nz = A[2 * 3 + d];  // J*.col(d)
      const double l2 = nx * nx + ny * ny + nz * nz;
      const double len = l2 * fast_rsqrt(l2);
      const double sgn = hi ? 1.0 : -1.0;
      if (kFR) {
        // fr/general.hpp:259-262 : f = f* (+-scale) - F(u) n^ ; fr/lobatto.hpp:234-237 : R -= f g'
        const double up = sharer ? -(sgn * len) : (sgn * len);
        const double gprime = sgn * basis.g_prime;
        const double *own = s_g + d * TILE + tid;  // my own F^ along d
#pragma unroll
        for (int c = 0; c < 5; ++c) res[c] -= (__ldg(fs + c * 32) * up - own[c * (TC * N)]) * gprime;
      } else {
        // dg/lobatto.hpp:171-175 : flux *= w_f ; holder -= ; sharer +=
        const int fa = d == 0 ? ijk[1] : ijk[0], fb = d == 2 ? ijk[1] : ijk[2];
        const double wf = (sharer ? 1.0 : -1.0) * (s_w[fa] * s_w[fb] * len);
#pragma unroll
        for (int c = 0; c < 5; ++c) res[c] += __ldg(fs + c * 32) * wf;
      }
    }
    if (!kFR) {  // dg/lobatto.hpp:124-139 : divide by the mass matrix
      const double inv = fast_rcp(s_w[ijk[0]] * s_w[ijk[1]] * s_w[ijk[2]] * fabs(det));
#pragma unroll
      for (int c = 0; c < 5; ++c) res[c] *= inv;
    }
    // RK combination in place: each thread reads and rewrites only its own 5 entries of s_o
#pragma unroll
    for (int c = 0; c < 5; ++c) {
      const int idx = tid * 5 + c;
      const double v0 = need_u0 ? s_o[idx] : 0.0;
      s_o[idx] = stage_combine(op, s_u[idx], res[c], v0);
    }
  }
  if (bulk) {
    fence_async_smem();  // generic-proxy writes to s_o -> visible to the async proxy (TMA)
    __syncthreads();
    if (tid == 0) {
      bulk_s2g(out + g0, s_o, tile_bytes);
      bulk_wait_read();  // smem must stay intact until the copy engine has read it
    }
  } else {
    __syncthreads();
    for (int idx = tid; idx < ncell * CELL; idx += TC * N) out[g0 + idx] = s_o[idx];
  }
}

template <int Q, int TC, int MINB>
static cudaError_t launch_cells(const DeviceTables &tab, const Basis &basis, int scheme, double gamma, const double *u,
                                const double *fstar, const double *u0, double *out, StageOp op, int64_t cell_begin,
                                int64_t cell_end, cudaStream_t stream) {
  constexpr int N = Q * Q * Q;
  const int64_t blocks = (cell_end - cell_begin + TC - 1) / TC;
  const size_t smem = size_t(5) * TC * N * 5 * sizeof(double);
  auto fr = cell_update_kernel<Q, true, TC, MINB>;
  auto dg = cell_update_kernel<Q, false, TC, MINB>;
  // the attribute belongs to (function, device): one flag per device, for hosts that drive several GPUs from one process
  static std::atomic<bool> configured[64] = {};  // handles on different devices may be driven from different host threads
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64 || !configured[dev].load(std::memory_order_acquire)) {
    cudaFuncSetAttribute(fr, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    cudaFuncSetAttribute(dg, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    if (dev >= 0 && dev < 64) configured[dev].store(true, std::memory_order_release);
  }
  if (scheme == MCFD_SCHEME_FR) {
    fr<<<blocks, TC * N, smem, stream>>>(tab, basis, gamma, u, fstar, u0, out, op, cell_begin, cell_end);
  } else {
    dg<<<blocks, TC * N, smem, stream>>>(tab, basis, gamma, u, fstar, u0, out, op, cell_begin, cell_end);
  }
  return cudaGetLastError();
}


// ---- line-per-thread variant ---------------------------------------------------------------------------
// Same arithmetic per node as cell_update_kernel, different ownership: a thread owns the Q nodes (i, j, 0..Q-1)
// of one cell.  ncu (profiles/r01o) shows the node-per-thread kernel sitting on the LSU data pipe -- 45 shared-
// memory loads per node for the divergence alone.  Here the z-derivative never leaves the registers, the x- and
// y-derivatives read Q lines of Q contiguous values, the node's own F^ along z stays in registers for the face
// correction, and only two of the three contravariant-flux tiles exist in shared memory: 30 loads and 10 stores
// per node instead of 55 and 15, and Q x 5 independent accumulation chains per thread instead of 5.
template <int Q, bool kFR, int TC, int MINB>
__global__ void __launch_bounds__(TC *Q *Q, MINB) cell_update_line_kernel(DeviceTables tab, Basis basis, double gamma,
                                                                           const double *__restrict__ u,
                                                                           const double *__restrict__ fstar,
                                                                           const double *__restrict__ u0,
                                                                           double *__restrict__ out, StageOp op,
                                                                           int64_t cell_begin, int64_t cell_end) {
  constexpr int N = Q * Q * Q, FQ = Q * Q, CELL = N * 5, TILE = TC * CELL, NT = TC * N;
  extern __shared__ __align__(128) double smem[];
  double *s_u = smem;             // [TILE] coefficients of the tile
  double *s_o = smem + TILE;      // [TILE] u0 tile in, new coefficients out
  double *s_g = smem + 2 * TILE;  // [2][5][TC * N] contravariant fluxes along x and y, component-major
  __shared__ double s_w[Q];
  __shared__ __align__(8) uint64_t s_bar;
  __shared__ int s_gp0[TC * 6];
  __shared__ uint8_t s_code[TC * 6];

  const int tid = threadIdx.x;
#ifdef MCFD_CELL_REVERSE
  // last cells first: the face kernel has just written the common fluxes in ascending order, so the tail of the
  // flux buffer (and of the column it read) is what the L2 still holds when this kernel starts
  const int64_t c0 = cell_begin + int64_t(gridDim.x - 1 - blockIdx.x) * TC;
#else
  const int64_t c0 = cell_begin + int64_t(blockIdx.x) * TC;
#endif
  const int ncell = int(min(int64_t(TC), cell_end - c0));
  const int64_t g0 = c0 * CELL;
  const bool need_u0 = op.kind >= 1 && op.kind <= 3;
  const uint32_t tile_bytes = uint32_t(ncell) * CELL * 8;
  const bool bulk = ((g0 * 8) % 16 == 0) && (tile_bytes % 16 == 0);
  if (bulk) {
    if (tid == 0) {
      mbar_init(&s_bar, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      mbar_expect_tx(&s_bar, need_u0 ? 2 * tile_bytes : tile_bytes);
      bulk_g2s(s_u, u + g0, tile_bytes, &s_bar);
      if (need_u0) bulk_g2s(s_o, u0 + g0, tile_bytes, &s_bar);
    }
  } else {
    for (int idx = tid; idx < ncell * CELL; idx += TC * FQ) {
      s_u[idx] = u[g0 + idx];
      if (need_u0) s_o[idx] = u0[g0 + idx];
    }
  }
  if (tid < Q) s_w[tid] = basis.w[tid];
  for (int t = tid; t < ncell * 6; t += TC * FQ) {  // at p = 1 a block has fewer threads (4 per cell) than face slots
    const int64_t f_bnd = tab.n_local_faces + tab.n_ghost_faces;
    const int f = __ldg(tab.cell_face + c0 * 6 + t);
    const bool flux = f >= 0 && !(f >= f_bnd && tab.groups[tab.bface_group[f - f_bnd]].kind < 0);
    s_gp0[t] = flux ? f * FQ : -1;
    s_code[t] = __ldg(tab.cell_code + c0 * 6 + t);
  }
  const int cl = tid / FQ, ij = tid - cl * FQ;
  const int i = ij / Q, j = ij - i * Q;
  const bool active = cl < ncell;
  const int64_t cell = c0 + cl;
  const int node0 = ij * Q;  // node (i, j, 0); the line is node0 .. node0 + Q - 1
  double A[9], det = 1.0;
  auto load_metrics = [&](int k) {
    const double *mp = tab.metrics + (tab.per_node_metrics ? (cell * N + node0 + k) * 10 : cell * 10);
#pragma unroll
    for (int e = 0; e < 9; ++e) A[e] = __ldg(mp + e);
    det = __ldg(mp + 9);
  };
  if (active && !tab.per_node_metrics) load_metrics(0);
  __syncthreads();
  // the common fluxes of this line's face points: asked for now (into L1), needed after the volume term
  if (active) {
    auto prefetch = [&](int d, bool hi, int a, int b) {
      const int side = hi ? (d == 0 ? 2 : d == 1 ? 3 : 5) : (d == 0 ? 4 : d == 1 ? 1 : 0);
      const int gp0 = s_gp0[cl * 6 + side];
      if (gp0 < 0) return;
      const int code = s_code[cl * 6 + side];
      if (code & 1) { int a2, b2; apply_perm<Q>(code >> 1, a, b, a2, b2); a = a2; b = b2; }
      const int64_t gp = int64_t(gp0) + a * Q + b;
      const double *fs = fstar + ((gp >> 5) * (32 * MCFD_K) + (gp & 31));
#pragma unroll
      for (int c = 0; c < 5; ++c) asm volatile("prefetch.global.L1 [%0];" ::"l"(fs + c * 32));
    };
    prefetch(2, false, i, j);
    prefetch(2, true, i, j);
    // the Q points of a line on an x- or y-face are neighbours in f*: the first and the last cover their lines
    if (i == 0 || i == Q - 1) { prefetch(0, i == Q - 1, j, 0); prefetch(0, i == Q - 1, j, Q - 1); }
    if (j == 0 || j == Q - 1) { prefetch(1, j == Q - 1, i, 0); prefetch(1, j == Q - 1, i, Q - 1); }
  }
  if (bulk) mbar_wait(&s_bar, 0);

  double gz[Q][5];  // this line's F^ along z
  if (active) {
#pragma unroll
    for (int k = 0; k < Q; ++k) {
      if (tab.per_node_metrics) load_metrics(k);
      double uu[5];
      const double rdet = fast_rcp(det);
#pragma unroll
      for (int c = 0; c < 5; ++c) {
        const double mine = s_u[(tid * Q + k) * 5 + c];
        uu[c] = kFR ? mine * rdet : mine;
      }
      const double scale = kFR ? 1.0 : (s_w[i] * s_w[j] * s_w[k] * fabs(det)) * rdet;
#ifdef MCFD_CELL_V1
      double F[3][5];
      flux_matrix(uu, gamma, F);
#pragma unroll
      for (int c = 0; c < 5; ++c) {
        s_g[c * NT + tid * Q + k] = (F[0][c] * A[0] + F[1][c] * A[3] + F[2][c] * A[6]) * scale;
        s_g[TILE + c * NT + tid * Q + k] = (F[0][c] * A[1] + F[1][c] * A[4] + F[2][c] * A[7]) * scale;
        gz[k][c] = (F[0][c] * A[2] + F[1][c] * A[5] + F[2][c] * A[8]) * scale;
      }
#else
      double G[3][5];
      contravariant_fluxes(uu, gamma, A, scale, G);
#pragma unroll
      for (int c = 0; c < 5; ++c) {
        s_g[c * NT + tid * Q + k] = G[0][c];
        s_g[TILE + c * NT + tid * Q + k] = G[1][c];
        gz[k][c] = G[2][c];
      }
#endif
    }
  }
  __syncthreads();

  if (active) {
    double res[5];  // of the node the k loop below is at
    const double *gx = s_g + cl * N, *gy = gx + TILE;
    double flen[3];
    auto face_lengths = [&]() {
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        const double l2 = A[0 * 3 + d] * A[0 * 3 + d] + A[1 * 3 + d] * A[1 * 3 + d] + A[2 * 3 + d] * A[2 * 3 + d];
        flen[d] = l2 * fast_rsqrt(l2);
      }
    };
    if (!tab.per_node_metrics) face_lengths();

    // face terms: (direction d, low/high end) of node (i, j, k); point (a, b) on the holder's lattice
    auto face_term = [&](int k, int d, bool hi, int a, int b, const double (&own)[5]) {
      const int side = hi ? (d == 0 ? 2 : d == 1 ? 3 : 5) : (d == 0 ? 4 : d == 1 ? 1 : 0);
      const int gp0 = s_gp0[cl * 6 + side];
      if (gp0 < 0) return;
      const int code = s_code[cl * 6 + side];
      const bool sharer = code & 1;
      if (sharer) { int a2, b2; apply_perm<Q>(code >> 1, a, b, a2, b2); a = a2; b = b2; }
      const int64_t gp = int64_t(gp0) + a * Q + b;
      const double *fs = fstar + ((gp >> 5) * (32 * MCFD_K) + (gp & 31));
#ifdef MCFD_CELL_V1
      const double nx = A[0 * 3 + d], ny = A[1 * 3 + d], nz = A[2 * 3 + d];
      const double l2 = nx * nx + ny * ny + nz * nz;
      const double len = l2 * fast_rsqrt(l2);
#else
      const double len = flen[d];   // |J*.col(d)|, once per thread (per node when the metrics are per node)
#endif
      const double sgn = hi ? 1.0 : -1.0;
      if (kFR) {
        const double up = sharer ? -(sgn * len) : (sgn * len);
        const double gprime = sgn * basis.g_prime;
#pragma unroll
        for (int c = 0; c < 5; ++c) res[c] -= (__ldg(fs + c * 32) * up - own[c]) * gprime;
      } else {
        const int fa = d == 0 ? j : i, fb = d == 2 ? j : k;
        const double wf = (sharer ? 1.0 : -1.0) * (s_w[fa] * s_w[fb] * len);
#pragma unroll
        for (int c = 0; c < 5; ++c) res[c] += __ldg(fs + c * 32) * wf;
      }
    };
#pragma unroll
    for (int k = 0; k < Q; ++k) {
      if (tab.per_node_metrics) { load_metrics(k); face_lengths(); }
      {
        double val[5] = { 0, 0, 0, 0, 0 };
#pragma unroll
        for (int m = 0; m < Q; ++m) {
          const double dx = kFR ? basis.D[i][m] : basis.D[m][i];
          const double dy = kFR ? basis.D[j][m] : basis.D[m][j];
          const double dz = kFR ? basis.D[k][m] : basis.D[m][k];
          const double *px = gx + ((m * Q + j) * Q + k);
          const double *py = gy + ((i * Q + m) * Q + k);
#pragma unroll
          for (int c = 0; c < 5; ++c) val[c] += dx * px[c * NT] + dy * py[c * NT] + dz * gz[m][c];
        }
#pragma unroll
        for (int c = 0; c < 5; ++c) res[c] = kFR ? -val[c] : val[c];
      }
      if (i == 0 || i == Q - 1) {
        double own[5];
#pragma unroll
        for (int c = 0; c < 5; ++c) own[c] = kFR ? s_g[c * NT + tid * Q + k] : 0.0;
        face_term(k, 0, i == Q - 1, j, k, own);
      }
      if (j == 0 || j == Q - 1) {
        double own[5];
#pragma unroll
        for (int c = 0; c < 5; ++c) own[c] = kFR ? s_g[TILE + c * NT + tid * Q + k] : 0.0;
        face_term(k, 1, j == Q - 1, i, k, own);
      }
      if (k == 0 || k == Q - 1) face_term(k, 2, k == Q - 1, i, j, gz[k]);
      double inv = 1.0;
      if (!kFR) inv = fast_rcp(s_w[i] * s_w[j] * s_w[k] * fabs(det));  // dg/lobatto.hpp:124-139
#pragma unroll
      for (int c = 0; c < 5; ++c) {
        const int idx = (tid * Q + k) * 5 + c;
        const double v0 = need_u0 ? s_o[idx] : 0.0;
        s_o[idx] = stage_combine(op, s_u[idx], kFR ? res[c] : res[c] * inv, v0);
      }
    }
  }
  if (bulk) {
    fence_async_smem();
    __syncthreads();
    if (tid == 0) {
      bulk_s2g(out + g0, s_o, tile_bytes);
      bulk_wait_read();
    }
  } else {
    __syncthreads();
    for (int idx = tid; idx < ncell * CELL; idx += TC * FQ) out[g0 + idx] = s_o[idx];
  }
}

template <int Q, int TC, int MINB>
static cudaError_t launch_cell_lines(const DeviceTables &tab, const Basis &basis, int scheme, double gamma, const double *u,
                                     const double *fstar, const double *u0, double *out, StageOp op, int64_t cell_begin,
                                     int64_t cell_end, cudaStream_t stream) {
  constexpr int N = Q * Q * Q;
  const int64_t blocks = (cell_end - cell_begin + TC - 1) / TC;
  const size_t smem = size_t(4) * TC * N * 5 * sizeof(double);
  auto fr = cell_update_line_kernel<Q, true, TC, MINB>;
  auto dg = cell_update_line_kernel<Q, false, TC, MINB>;
  // the attribute belongs to (function, device): one flag per device, for hosts that drive several GPUs from one process
  static std::atomic<bool> configured[64] = {};  // handles on different devices may be driven from different host threads
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64 || !configured[dev].load(std::memory_order_acquire)) {
    cudaFuncSetAttribute(fr, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    cudaFuncSetAttribute(dg, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    if (dev >= 0 && dev < 64) configured[dev].store(true, std::memory_order_release);
  }
  if (scheme == MCFD_SCHEME_FR) {
    fr<<<blocks, TC * Q * Q, smem, stream>>>(tab, basis, gamma, u, fstar, u0, out, op, cell_begin, cell_end);
  } else {
    dg<<<blocks, TC * Q * Q, smem, stream>>>(tab, basis, gamma, u, fstar, u0, out, op, cell_begin, cell_end);
  }
  return cudaGetLastError();
}

#ifndef MCFD_CELL_MINB
#define MCFD_CELL_MINB 4
#endif

cudaError_t LaunchCellUpdate(const DeviceTables &tab, const Basis &basis, int scheme, double gamma, const double *u,
                             const double *fstar, const double *u0, double *out, StageOp op, int64_t cell_begin,
                             int64_t cell_end, cudaStream_t stream) {
  if (cell_end <= cell_begin) return cudaSuccess;
  // MCFD_CELL_LINES (tuning): 0 = node-per-thread kernel everywhere, 1 = line-per-thread kernel everywhere;
  // default: whichever measured faster on B200 -- lines for FR p = 2 (0.343 against 0.365 ms), nodes for
  // DGSEM p = 2 (0.374 against 0.400), p = 1 and p = 3 (0.32 against 0.54: four nodes per thread cost 150 registers)
  static const int lines = std::getenv("MCFD_CELL_LINES") ? std::atoi(std::getenv("MCFD_CELL_LINES")) : -1;
  const bool use_lines = lines < 0 ? (basis.q == 3 && scheme == MCFD_SCHEME_FR) : lines != 0;
  if (use_lines) {
    switch (basis.q) {
      case 2: return launch_cell_lines<2, 32, 5>(tab, basis, scheme, gamma, u, fstar, u0, out, op, cell_begin, cell_end, stream);
      case 3: return launch_cell_lines<3, 10, 4>(tab, basis, scheme, gamma, u, fstar, u0, out, op, cell_begin, cell_end, stream);
      case 4: return lines == 6 ? launch_cell_lines<4, 6, 3>(tab, basis, scheme, gamma, u, fstar, u0, out, op, cell_begin, cell_end, stream)
                                : launch_cell_lines<4, 4, 4>(tab, basis, scheme, gamma, u, fstar, u0, out, op, cell_begin, cell_end, stream);
      default: return cudaErrorInvalidValue;
    }
  }
  switch (basis.q) {
    case 2: return launch_cells<2, 32, MCFD_CELL_MINB>(tab, basis, scheme, gamma, u, fstar, u0, out, op, cell_begin, cell_end, stream);
    case 3: return launch_cells<3, 8, MCFD_CELL_MINB>(tab, basis, scheme, gamma, u, fstar, u0, out, op, cell_begin, cell_end, stream);
    case 4: return launch_cells<4, 4, MCFD_CELL_MINB>(tab, basis, scheme, gamma, u, fstar, u0, out, op, cell_begin, cell_end, stream);
    default: return cudaErrorInvalidValue;
  }
}

// Part::ShareGhostCellCoeffs packing (mesh/part.hpp:1258-1270): gather whole coefficient blocks
__global__ void pack_cells_kernel(const double *__restrict__ u, const int32_t *__restrict__ cells, int64_t n_cells,
                                  int fields, double *__restrict__ buffer) {
  int64_t idx = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= n_cells * fields) return;
  int64_t i = idx / fields;
  int k = int(idx % fields);
  buffer[idx] = u[int64_t(cells[i]) * fields + k];
}

cudaError_t LaunchPackCells(const double *u, const int32_t *cells, int64_t n_cells, int fields, double *buffer,
                            cudaStream_t stream) {
  if (n_cells <= 0) return cudaSuccess;
  const int threads = 256;
  const int64_t total = n_cells * fields;
  pack_cells_kernel<<<(total + threads - 1) / threads, threads, 0, stream>>>(u, cells, n_cells, fields, buffer);
  return cudaGetLastError();
}

// ---- halo exchange over peer memory (NVLink): pack + remote store in one kernel -------------------------
// Part::ShareGhostCellCoeffs + the MPI_Isend/Irecv pair (mesh/part.hpp:1258-1302) without a library call: the
// coefficient blocks of the send cells are gathered and stored straight into the ghost tail of each neighbour's
// column (peer-mapped device memory), 8-byte stores coalesced over the K N fields of a cell.
__global__ void pack_to_peers_kernel(const double *__restrict__ u, const int32_t *__restrict__ cells,
                                     const int32_t *__restrict__ cell_nbr, const int64_t *__restrict__ dst_offset,
                                     double *const *__restrict__ peer_u, int64_t n_cells, int fields,
                                     unsigned long long *const *__restrict__ peer_flag, int n_neighbours,
                                     unsigned long long seq, unsigned int *__restrict__ blocks_done) {
  // a few fat blocks (grid-stride), not one thin block per 256 entries: every block ends on a system-wide fence that
  // waits for its remote stores to be acknowledged over NVLink, and next to the face kernel only a handful of blocks of
  // this kernel are resident at a time -- 2160 thin blocks took the whole face kernel to drain (the ghost-face kernel
  // then polled for ~17 us per stage at every N, profiles/r02_multi/r02z_pack_*)
  const int64_t total = n_cells * fields;
  for (int64_t idx = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; idx < total; idx += int64_t(gridDim.x) * blockDim.x) {
    const int64_t i = idx / fields;
    const int k = int(idx - i * fields);
    double *dst = peer_u[cell_nbr[i]] + dst_offset[i] + k;
    *dst = u[int64_t(cells[i]) * fields + k];
  }
  // The block that finishes last raises the neighbours' arrival counters (monotone: one exchange = one increment on
  // every rank): every block makes its remote stores visible system-wide before it checks in, so the release store of
  // the last one orders all of them.  No separate signalling launch.
  __threadfence_system();
  __syncthreads();
  __shared__ bool s_last;
  if (threadIdx.x == 0) s_last = atomicInc(blocks_done, gridDim.x - 1) == gridDim.x - 1;  // wraps to 0: ready for the next launch
  __syncthreads();
  if (s_last) {
    __threadfence_system();
    for (int n = threadIdx.x; n < n_neighbours; n += blockDim.x)   // any number of neighbours, not just blockDim.x
      asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(peer_flag[n]), "l"(seq) : "memory");
  }
}

// a rank with neighbours but nothing to send still has to check in
__global__ void signal_peers_kernel(unsigned long long *const *__restrict__ peer_flag, int n_neighbours,
                                    unsigned long long seq) {
  __threadfence_system();
  for (int n = threadIdx.x; n < n_neighbours; n += blockDim.x)
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(peer_flag[n]), "l"(seq) : "memory");
}

// UpdateGhostCellCoeffs' MPI_Waitall: one thread per neighbour spins on this rank's arrival counters.  A peer that
// never arrives must not hang the GPU: after `timeout_cycles` the kernel gives up and raises *error.
__global__ void wait_peers_kernel(const unsigned long long *__restrict__ flags, int n_neighbours, unsigned long long seq,
                                  long long timeout_cycles, int *__restrict__ error) {
  for (int n = threadIdx.x; n < n_neighbours; n += blockDim.x) {
    const long long start = clock64();
    unsigned long long seen = 0;
    for (;;) {
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(seen) : "l"(flags + n) : "memory");
      if (seen >= seq) break;
      if (clock64() - start > timeout_cycles) { atomicExch(error, 1 + n); break; }
      __nanosleep(200);
    }
  }
}

cudaError_t LaunchPackToPeers(const double *u, const int32_t *cells, const int32_t *cell_nbr, const int64_t *dst_offset,
                              double *const *peer_u, int64_t n_cells, int fields, unsigned long long *const *peer_flag,
                              int n_neighbours, unsigned long long seq, unsigned int *blocks_done, cudaStream_t stream) {
  if (n_neighbours <= 0) return cudaSuccess;
  if (n_cells > 0) {   // the last block signals every neighbour (its lanes loop over them)
    const int threads = 256;
    const int64_t total = n_cells * fields;
    // per call, for the device that is current: handles of one process may live on different GPUs
    int dev = 0, sms = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return e;
    // MCFD_PACK_BLOCKS_PER_SM (tuning): how many blocks per SM the pack kernel asks for
    static const int per_sm = std::getenv("MCFD_PACK_BLOCKS_PER_SM") ? std::max(1, std::atoi(std::getenv("MCFD_PACK_BLOCKS_PER_SM"))) : 2;
    const int64_t blocks = std::min<int64_t>((total + threads - 1) / threads, int64_t(sms) * per_sm);
    pack_to_peers_kernel<<<unsigned(blocks), threads, 0, stream>>>(
        u, cells, cell_nbr, dst_offset, peer_u, n_cells, fields, peer_flag, n_neighbours, seq, blocks_done);
  } else {
    signal_peers_kernel<<<1, std::min(1024, 32 * ((n_neighbours + 31) / 32)), 0, stream>>>(peer_flag, n_neighbours, seq);
  }
  return cudaGetLastError();
}

cudaError_t LaunchWaitPeers(const unsigned long long *flags, int n_neighbours, unsigned long long seq, int *error,
                            cudaStream_t stream) {
  if (n_neighbours <= 0) return cudaSuccess;
  const long long timeout_cycles = 20LL * 1000 * 1000 * 1000;  // ~10 s at 2 GHz
  wait_peers_kernel<<<1, std::min(1024, 32 * ((n_neighbours + 31) / 32)), 0, stream>>>(flags, n_neighbours, seq, timeout_cycles, error);
  return cudaGetLastError();
}

}  // namespace mcfd
