// face_kernels.cu -- common (upwind) flux at every face quadrature point, one thread per point.
//
// Replaces the per-face loops of the reference (paths relative to /root/reference/include/mini):
//   spatial/fr/lobatto.hpp:219-259, spatial/dg/lobatto.hpp:156-195   interior / ghost faces
//   spatial/fr/lobatto.hpp:260-391, spatial/dg/lobatto.hpp:196-300   the six boundary kinds
// The reference scatters each flux straight into the residual of both cells (a data race on a
// GPU).  Here the flux f*(x_p) in global coordinates is stored ONCE per face point in
// fstar[face][p][5]; the cell kernel (cell_kernels.cu) gathers it for both neighbours, so the
// result does not depend on scheduling and no FP64 atomics are needed.
//
// HBM layout: u is the reference's column ([cell][node][5]), ghosts appended after the local
// cells.  Gathers are 40-byte node records (L1/L2 absorb the reuse between the points of a face);
// fluxes are written in the chunked component-major layout of FstarIndex() (solver.hpp), 256
// contiguous bytes per component and warp.
#include <algorithm>
#include <cstdlib>

#include "riemann.cuh"
#include "solver.hpp"
#include "state_functions.cuh"

namespace mcfd {

// kPeerWritten: the record may be stored by a neighbour GPU while this kernel is resident (the ghost tail, filled
// during wait_for_peers' spin).  ld.global.nc is only defined for data that stays read-only for the whole kernel and
// the acquire on the arrival counter does not order it, so those records are read with ld.global.cg (L2, the point
// of coherence for peer stores) instead.
template <bool kPeerWritten = false>
__device__ __forceinline__ Cons load_state(const double *__restrict__ u, int64_t record) {
  Cons c;
  const double *p = u + record * 5;
#pragma unroll
  for (int k = 0; k < 5; ++k) c.c[k] = kPeerWritten ? __ldcg(p + k) : __ldg(p + k);
  return c;
}

// One warp per block, 32 blocks/SM: the same 32 resident warps as 128 x 8, but a warp's slot is handed on the moment
// its own Newton loops end instead of when the slowest warp of its block does (0.277 ms per launch; 64 x 16: 0.282,
// 128 x 8: 0.288, 256 x 4: 0.309)
#ifndef MCFD_FACE_THREADS
#define MCFD_FACE_THREADS 32
#endif
#ifndef MCFD_FACE_MINB
#define MCFD_FACE_MINB (1024 / MCFD_FACE_THREADS)
#endif

// One thread per face point gp = face * FQ + p, all in 32-bit index arithmetic.  Which solution node
// of which cell sits on either side of the point comes from the tables pt_holder / pt_sharer built at
// setup (setup.cpp), so the kernel does no face-orientation logic.
// kBoundary selects the boundary-face variant (given state + one of the six boundary fluxes); the
// interior variant (local and ghost faces) carries none of that code.
// kGhostSharer: the sharer side may be a ghost record that arrives while the kernel runs (see load_state).
template <int Q, bool kFR, int kRiemann, bool kBoundary, bool kGhostSharer = false>
__device__ __forceinline__ void face_point(const DeviceTables &tab, double gamma, double t, const double *__restrict__ u,
                                           double *__restrict__ fstar, const int gp) {
  constexpr int FQ = Q * Q, N = Q * Q * Q;
  const int rec_h = __ldg(tab.pt_holder + gp), rec_s = __ldg(tab.pt_sharer + gp);
  const int f = gp / FQ;
  Cons uh = load_state(u, rec_h);
  if (kFR) {  // polynomial/hexahedron.hpp:257-259 : u = U^ / det J
    const double det = __ldg(tab.metrics + int64_t(tab.per_node_metrics ? rec_h : rec_h / N) * 10 + 9);
    const double rdet = rn_rcp(det);  // RN(U^ / det J) without the division's special-case branches
#pragma unroll
    for (int k = 0; k < 5; ++k) uh.c[k] = rn_div(uh.c[k], det, rdet);
  }
  Frame fr;
  {
    const double *fp = tab.face_frame + int64_t(tab.per_point_frames ? gp : f) * 9;
#pragma unroll
    for (int d = 0; d < 3; ++d) { fr.n[d] = __ldg(fp + d); fr.t[d] = __ldg(fp + 3 + d); fr.b[d] = __ldg(fp + 6 + d); }
  }
  Cons flux;
  if (!kBoundary) {
    Cons us = load_state<kGhostSharer>(u, rec_s);
    if (kFR) {
      const double det = __ldg(tab.metrics + int64_t(tab.per_node_metrics ? rec_s : rec_s / N) * 10 + 9);
      const double rdet = rn_rcp(det);
#pragma unroll
      for (int k = 0; k < 5; ++k) us.c[k] = rn_div(us.c[k], det, rdet);
    }
    if (kRiemann == MCFD_RIEMANN_EXACT) {
      // rotated_flux_upwind<0> (riemann.cuh) with the register budget managed by hand: the iteration needs
      // only (rho, u, p) of either side, so the passive tangential velocities wait in shared memory and the
      // frame is dropped and fetched again (L1/L2 hit) for the rotation back.  Same operations, same results.
      __shared__ double s_vw[4][MCFD_FACE_THREADS];
      Prim l = cons_to_prim(uh, gamma), r = cons_to_prim(us, gamma);
      global_to_normal(fr, l.u, l.v, l.w);
      global_to_normal(fr, r.u, r.v, r.w);
      s_vw[0][threadIdx.x] = l.v; s_vw[1][threadIdx.x] = l.w; s_vw[2][threadIdx.x] = r.v; s_vw[3][threadIdx.x] = r.w;
      l.v = l.w = r.v = r.w = 0.0;
      bool from_left;
      Prim star = (gamma == 1.4) ? exact_star_state<true>(l, r, gamma, from_left, tab.star_u + gp)
                                 : exact_star_state<false>(l, r, gamma, from_left, tab.star_u + gp);
      star.v = s_vw[from_left ? 0 : 2][threadIdx.x];  // exact.hpp:37-46 PassiveScalars<3>
      star.w = s_vw[from_left ? 1 : 3][threadIdx.x];
      flux = prim_to_flux(star, gamma);
      const double *fp = tab.face_frame + int64_t(tab.per_point_frames ? gp : f) * 9;
      asm volatile("" : "+l"(fp));  // keep the compiler from holding the first copy of the frame across the solve
      Frame fr2;
#pragma unroll
      for (int d = 0; d < 3; ++d) { fr2.n[d] = __ldg(fp + d); fr2.t[d] = __ldg(fp + 3 + d); fr2.b[d] = __ldg(fp + 6 + d); }
      normal_to_global(fr2, flux.c[1], flux.c[2], flux.c[3]);
    } else {
      flux = rotated_flux_upwind<kRiemann>(fr, uh, us, gamma, tab.star_u + gp);
    }
  } else {
    const int p = gp - f * FQ;
    const int bf = f - int(tab.n_local_faces + tab.n_ghost_faces);
    const int g = tab.bface_group[bf];
    const BcGroup &grp = tab.groups[g];
    int kind = grp.kind;
    if (kind < 0) return;  // boundary name never registered: no flux (fem.hpp:185-226 loops over registered names)
    Cons given;
#pragma unroll
    for (int k = 0; k < 5; ++k) given.c[k] = 0.0;
    if (kind != MCFD_BC_SOLID_WALL && kind != MCFD_BC_SUPERSONIC_OUTLET) {
      if (grp.func == MCFD_FUNC_HOST_VALUES) {
        const double *v = grp.host_values + (int64_t(tab.bface_local[bf]) * FQ + p) * 5;
#pragma unroll
        for (int k = 0; k < 5; ++k) given.c[k] = v[k];
      } else {
        const double *x = tab.bface_xyz + (int64_t(bf) * FQ + p) * 3;
        given = eval_state(grp.func, grp.params, x[0], x[1], x[2], t);
      }
    }
    flux = boundary_flux(kind, fr, uh, given, gamma);
  }
  double *out = fstar + (int64_t(gp >> 5) * (32 * MCFD_K) + (gp & 31));
#pragma unroll
  for (int k = 0; k < 5; ++k) out[k * 32] = flux.c[k];
}

// UpdateGhostCellCoeffs' MPI_Waitall inside the ghost-face kernel: lane n of every warp polls this rank's arrival
// counter for neighbour n until it has reached `seq` (the neighbours' stores into the ghost tail are then visible:
// st.release.sys on their side, ld.acquire.sys here).  A peer that never arrives must not hang the GPU: after
// `timeout_cycles` the warp gives up, raises *error and goes on with whatever is in the ghost tail.
__device__ __forceinline__ void wait_for_peers(const PeerWait &w) {
  if (!w.flags) return;
  const int lane = threadIdx.x & 31;
  for (int n = lane; n < w.n_neighbours; n += 32) {
    const long long start = clock64();
    unsigned long long seen = 0;
    for (;;) {
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(seen) : "l"(w.flags + n) : "memory");
      if (seen >= w.seq) break;
      if (clock64() - start > w.timeout_cycles) { atomicExch(w.error, 1 + n); break; }
      __nanosleep(100);
    }
  }
  __syncwarp();
}

template <int Q, bool kFR, int kRiemann, bool kBoundary>
__global__ void __launch_bounds__(MCFD_FACE_THREADS, MCFD_FACE_MINB) face_flux_kernel(DeviceTables tab, double gamma, double t,
                                                                       const double *__restrict__ u,
                                                                       double *__restrict__ fstar, int point_begin,
                                                                       int n_points, PeerWait wait) {
  if (!kBoundary) wait_for_peers(wait);
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_points) return;
  face_point<Q, kFR, kRiemann, kBoundary, !kBoundary>(tab, gamma, t, u, fstar, point_begin + i);
}

// Interior (local) faces and boundary faces in ONE launch: neither waits for the halo, and at 32^3 cells per GPU the
// boundary launch alone is a 10 us tail.  Blocks [0, blocks_a) take the interior range, the rest the boundary range.
template <int Q, bool kFR, int kRiemann>
__global__ void __launch_bounds__(MCFD_FACE_THREADS, MCFD_FACE_MINB) face_flux_merged_kernel(DeviceTables tab, double gamma, double t,
                                                                              const double *__restrict__ u,
                                                                              double *__restrict__ fstar, int begin_a, int n_a,
                                                                              int blocks_a, int begin_b, int n_b) {
  if (int(blockIdx.x) < blocks_a) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_a) face_point<Q, kFR, kRiemann, false>(tab, gamma, t, u, fstar, begin_a + i);
  } else {
    const int i = (blockIdx.x - blocks_a) * blockDim.x + threadIdx.x;
    if (i < n_b) face_point<Q, kFR, 0, true>(tab, gamma, t, u, fstar, begin_b + i);
  }
}

template <int Q, bool kFR>
static cudaError_t launch_face(int riemann, const DeviceTables &tab, double gamma, double t, const double *u,
                               double *fstar, int64_t face_begin, int64_t n_points64, const PeerWait &wait,
                               cudaStream_t stream) {
  const int threads = MCFD_FACE_THREADS;
  const int n_points = int(n_points64), blocks = (n_points + threads - 1) / threads;
  constexpr int FQ = Q * Q;
  const int point_begin = int(face_begin * FQ);
  if (face_begin >= tab.n_local_faces + tab.n_ghost_faces) {  // boundary faces: the Riemann solver is not involved
    face_flux_kernel<Q, kFR, 0, true><<<blocks, threads, 0, stream>>>(tab, gamma, t, u, fstar, point_begin, n_points, wait);
    return cudaGetLastError();
  }
  switch (riemann) {
    case MCFD_RIEMANN_EXACT: face_flux_kernel<Q, kFR, 0, false><<<blocks, threads, 0, stream>>>(tab, gamma, t, u, fstar, point_begin, n_points, wait); break;
    case MCFD_RIEMANN_ROE: face_flux_kernel<Q, kFR, 1, false><<<blocks, threads, 0, stream>>>(tab, gamma, t, u, fstar, point_begin, n_points, wait); break;
    case MCFD_RIEMANN_HLLC: face_flux_kernel<Q, kFR, 2, false><<<blocks, threads, 0, stream>>>(tab, gamma, t, u, fstar, point_begin, n_points, wait); break;
    case MCFD_RIEMANN_AUSM: face_flux_kernel<Q, kFR, 3, false><<<blocks, threads, 0, stream>>>(tab, gamma, t, u, fstar, point_begin, n_points, wait); break;
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

template <int Q, bool kFR>
static cudaError_t launch_face_merged(int riemann, const DeviceTables &tab, double gamma, double t, const double *u,
                                      double *fstar, cudaStream_t stream) {
  const int threads = MCFD_FACE_THREADS;
  constexpr int FQ = Q * Q;
  const int n_a = int(tab.n_local_faces * FQ), n_b = int(tab.n_boundary_faces * FQ);
  const int begin_b = int((tab.n_local_faces + tab.n_ghost_faces) * FQ);
  const int blocks_a = (n_a + threads - 1) / threads, blocks_b = (n_b + threads - 1) / threads;
  if (blocks_a + blocks_b == 0) return cudaSuccess;
#define MCFD_MERGED(R) face_flux_merged_kernel<Q, kFR, R><<<blocks_a + blocks_b, threads, 0, stream>>>(tab, gamma, t, u, fstar, 0, n_a, blocks_a, begin_b, n_b)
  switch (riemann) {
    case MCFD_RIEMANN_EXACT: MCFD_MERGED(0); break;
    case MCFD_RIEMANN_ROE: MCFD_MERGED(1); break;
    case MCFD_RIEMANN_HLLC: MCFD_MERGED(2); break;
    case MCFD_RIEMANN_AUSM: MCFD_MERGED(3); break;
    default: return cudaErrorInvalidValue;
  }
#undef MCFD_MERGED
  return cudaGetLastError();
}

cudaError_t LaunchFaceFlux(const DeviceTables &tab, const Basis &basis, int scheme, int riemann, double gamma, double t,
                           const double *u, double *fstar, int64_t face_begin, int64_t face_end, cudaStream_t stream,
                           const PeerWait *wait) {
  if (face_end <= face_begin) return cudaSuccess;
  const int64_t n_points = (face_end - face_begin) * tab.fq;
  const bool fr = scheme == MCFD_SCHEME_FR;
  const PeerWait none{ nullptr, 0, 0, 0, nullptr };
  const PeerWait &w = wait ? *wait : none;
  switch (basis.q) {
    case 2: return fr ? launch_face<2, true>(riemann, tab, gamma, t, u, fstar, face_begin, n_points, w, stream)
                      : launch_face<2, false>(riemann, tab, gamma, t, u, fstar, face_begin, n_points, w, stream);
    case 3: return fr ? launch_face<3, true>(riemann, tab, gamma, t, u, fstar, face_begin, n_points, w, stream)
                      : launch_face<3, false>(riemann, tab, gamma, t, u, fstar, face_begin, n_points, w, stream);
    case 4: return fr ? launch_face<4, true>(riemann, tab, gamma, t, u, fstar, face_begin, n_points, w, stream)
                      : launch_face<4, false>(riemann, tab, gamma, t, u, fstar, face_begin, n_points, w, stream);
    default: return cudaErrorInvalidValue;
  }
}

// local + boundary faces of a nodal scheme in one launch (see face_flux_merged_kernel)
cudaError_t LaunchFaceFluxLocalAndBoundary(const DeviceTables &tab, const Basis &basis, int scheme, int riemann, double gamma,
                                           double t, const double *u, double *fstar, cudaStream_t stream) {
  const bool fr = scheme == MCFD_SCHEME_FR;
  switch (basis.q) {
    case 2: return fr ? launch_face_merged<2, true>(riemann, tab, gamma, t, u, fstar, stream)
                      : launch_face_merged<2, false>(riemann, tab, gamma, t, u, fstar, stream);
    case 3: return fr ? launch_face_merged<3, true>(riemann, tab, gamma, t, u, fstar, stream)
                      : launch_face_merged<3, false>(riemann, tab, gamma, t, u, fstar, stream);
    case 4: return fr ? launch_face_merged<4, true>(riemann, tab, gamma, t, u, fstar, stream)
                      : launch_face_merged<4, false>(riemann, tab, gamma, t, u, fstar, stream);
    default: return cudaErrorInvalidValue;
  }
}

// ---- standalone entry points for the known-answer tests (test/riemann/euler/exact.cpp, rotated/euler.cpp) ----
template <int kRiemann>
__global__ void riemann_flux_kernel(double gamma, int64_t n, const double *frames, const double *ul, const double *ur,
                                    double *flux) {
  int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Frame fr;
  for (int d = 0; d < 3; ++d) { fr.n[d] = frames[9 * i + d]; fr.t[d] = frames[9 * i + 3 + d]; fr.b[d] = frames[9 * i + 6 + d]; }
  Cons l = load_state(ul, i), r = load_state(ur, i);
  Cons f = rotated_flux_upwind<kRiemann>(fr, l, r, gamma);
  for (int k = 0; k < 5; ++k) flux[5 * i + k] = f.c[k];
}

cudaError_t LaunchRiemannFlux(int riemann, double gamma, int64_t n, const double *frames, const double *ul,
                              const double *ur, double *flux, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  const int threads = 128;
  const int64_t blocks = (n + threads - 1) / threads;
  switch (riemann) {
    case 0: riemann_flux_kernel<0><<<blocks, threads, 0, stream>>>(gamma, n, frames, ul, ur, flux); break;
    case 1: riemann_flux_kernel<1><<<blocks, threads, 0, stream>>>(gamma, n, frames, ul, ur, flux); break;
    case 2: riemann_flux_kernel<2><<<blocks, threads, 0, stream>>>(gamma, n, frames, ul, ur, flux); break;
    case 3: riemann_flux_kernel<3><<<blocks, threads, 0, stream>>>(gamma, n, frames, ul, ur, flux); break;
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

__global__ void boundary_flux_kernel(int kind, double gamma, int64_t n, const double *frames, const double *ui,
                                     const double *uo, double *flux) {
  int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Frame fr;
  for (int d = 0; d < 3; ++d) { fr.n[d] = frames[9 * i + d]; fr.t[d] = frames[9 * i + 3 + d]; fr.b[d] = frames[9 * i + 6 + d]; }
  Cons a = load_state(ui, i), b = load_state(uo, i);
  Cons f = boundary_flux(kind, fr, a, b, gamma);
  for (int k = 0; k < 5; ++k) flux[5 * i + k] = f.c[k];
}

cudaError_t LaunchBoundaryFlux(int kind, double gamma, int64_t n, const double *frames, const double *ui,
                               const double *uo, double *flux, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  if (kind < 0 || kind > 5) return cudaErrorInvalidValue;
  const int threads = 128;
  boundary_flux_kernel<<<(n + threads - 1) / threads, threads, 0, stream>>>(kind, gamma, n, frames, ui, uo, flux);
  return cudaGetLastError();
}

// ---- Cell::Approximate (polynomial/hexahedron.hpp:499-513) with a built-in function ----
template <int Q, bool kFR>
__global__ void approximate_kernel(DeviceTables tab, Basis basis, const double *__restrict__ corner_xyz, int func,
                                   FuncParams params, double t, double *__restrict__ u) {
  constexpr int N = Q * Q * Q;
  int64_t idx = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= tab.n_local * N) return;
  int64_t cell = idx / N;
  int node = int(idx % N), i = node / (Q * Q), j = (node / Q) % Q, k = node % Q;
  const double lx = basis.x[i], ly = basis.x[j], lz = basis.x[k];
  const double cx[8] = { -1, 1, 1, -1, -1, 1, 1, -1 }, cy[8] = { -1, -1, 1, 1, -1, -1, 1, 1 }, cz[8] = { -1, -1, -1, -1, 1, 1, 1, 1 };
  const double *xn = corner_xyz + cell * 24;
  double x[3] = { 0, 0, 0 };
  for (int v = 0; v < 8; ++v) {  // geometry/hexahedron.hpp:419-428 + LocalToGlobal, same operation order
    double val = (1 + cx[v] * lx);
    val *= (1 + cy[v] * ly);
    val *= (1 + cz[v] * lz);
    double shape = val / 8;
    for (int d = 0; d < 3; ++d) x[d] = v == 0 ? xn[d] * shape : x[d] + xn[3 * v + d] * shape;
  }
  Cons val = eval_state(func, params.p, x[0], x[1], x[2], t);
  if (kFR) {
    double det = tab.per_node_metrics ? tab.metrics[idx * 10 + 9] : tab.metrics[cell * 10 + 9];
    for (int c = 0; c < 5; ++c) val.c[c] *= det;
  }
  for (int c = 0; c < 5; ++c) u[idx * 5 + c] = val.c[c];
}

cudaError_t LaunchApproximate(const DeviceTables &tab, const Basis &basis, int scheme, const double *corner_xyz,
                              int func, const double *params, double t, double *u, cudaStream_t stream) {
  FuncParams fp;
  for (int i = 0; i < MCFD_MAX_FUNC_PARAMS; ++i) fp.p[i] = params[i];
  const int threads = 128;
  const int64_t total = tab.n_local * tab.n;
  const int64_t blocks = (total + threads - 1) / threads;
  const bool fr = scheme == MCFD_SCHEME_FR;
#define MCFD_LAUNCH(QQ) \
  if (fr) approximate_kernel<QQ, true><<<blocks, threads, 0, stream>>>(tab, basis, corner_xyz, func, fp, t, u); \
  else approximate_kernel<QQ, false><<<blocks, threads, 0, stream>>>(tab, basis, corner_xyz, func, fp, t, u)
  switch (basis.q) {
    case 2: MCFD_LAUNCH(2); break;
    case 3: MCFD_LAUNCH(3); break;
    case 4: MCFD_LAUNCH(4); break;
    default: return cudaErrorInvalidValue;
  }
#undef MCFD_LAUNCH
  return cudaGetLastError();
}

__global__ void eval_points_kernel(int func, FuncParams params, double t, int64_t n, const double *xyz, double *values) {
  int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Cons v = eval_state(func, params.p, xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], t);
  for (int k = 0; k < 5; ++k) values[5 * i + k] = v.c[k];
}

cudaError_t LaunchEvalBcPoints(int func, const double *params, double t, int64_t n, const double *xyz, double *values,
                               cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  FuncParams fp;
  for (int i = 0; i < MCFD_MAX_FUNC_PARAMS; ++i) fp.p[i] = params[i];
  const int threads = 128;
  eval_points_kernel<<<(n + threads - 1) / threads, threads, 0, stream>>>(func, fp, t, n, xyz, values);
  return cudaGetLastError();
}

}  // namespace mcfd
