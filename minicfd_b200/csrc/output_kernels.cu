// output_kernels.cu -- the step after the hot path in the reference's Main(): turning the coefficient
// column into what its writers consume.
//
// Replaces (paths relative to /root/reference/include/mini):
//   mesh/vtk.hpp:341-377, 86-95         vtk::Writer::Prepare: every hexahedron is sampled at the 27 nodes of a
//                                       VTK tri-quadratic hexahedron (LocalToGlobalAndValue per node)
//   polynomial/hexahedron.hpp:230-243   nodal LocalToValue (FR: interpolate U^, divide by det J at the point)
//   polynomial/projection.hpp:104-124   modal GlobalToValue at the image of the local point
//   mesh/part.hpp:370-387               Section::GatherFields / ScatterFields: coefficient blocks <-> the
//                                       field-major arrays fields_[i_field][i_cell] the CGNS writer stores
// One thread per (cell, sample point); HBM-bound: reads the cell's coefficients (L1-shared by the 27 threads
// of a cell), writes 8 doubles per sample.
#include <mutex>

#include "modal_device.cuh"
#include "riemann.cuh"
#include "solver.hpp"
#include "state_functions.cuh"

namespace mcfd {

// mesh/vtk.hpp:229-246 : Hexahedron27<Local>::locals (corners, edge mid-points, face centres, body centre)
__constant__ signed char kVtkHex27[27][3] = {
  { -1, -1, -1 }, { +1, -1, -1 }, { +1, +1, -1 }, { -1, +1, -1 }, { -1, -1, +1 }, { +1, -1, +1 }, { +1, +1, +1 }, { -1, +1, +1 },
  { 0, -1, -1 }, { +1, 0, -1 }, { 0, +1, -1 }, { -1, 0, -1 }, { 0, -1, +1 }, { +1, 0, +1 }, { 0, +1, +1 }, { -1, 0, +1 },
  { -1, -1, 0 }, { +1, -1, 0 }, { +1, +1, 0 }, { -1, +1, 0 },
  { -1, 0, 0 }, { +1, 0, 0 }, { 0, -1, 0 }, { 0, +1, 0 }, { 0, 0, -1 }, { 0, 0, +1 },
  { 0, 0, 0 },
};

__device__ __forceinline__ void hex8_corner_signs(int v, double &cx, double &cy, double &cz) {
  // geometry/hexahedron.hpp:345-352 : CGNS HEXA_8 corner order
  cx = (v == 1 || v == 2 || v == 5 || v == 6) ? 1.0 : -1.0;
  cy = (v == 2 || v == 3 || v == 6 || v == 7) ? 1.0 : -1.0;
  cz = v >= 4 ? 1.0 : -1.0;
}

// geometry/hexahedron.hpp:419-428 + geometry/cell.hpp:67-75, same operation order (the reference's base64
// point block is reproduced bit for bit)
__device__ __forceinline__ void hex8_local_to_global(const double *__restrict__ xn, double lx, double ly, double lz,
                                                     double (&x)[3]) {
#pragma unroll
  for (int v = 0; v < 8; ++v) {
    double cx, cy, cz;
    hex8_corner_signs(v, cx, cy, cz);
    double val = (1 + cx * lx);
    val *= (1 + cy * ly);
    val *= (1 + cz * lz);
    const double shape = val / 8;
#pragma unroll
    for (int d = 0; d < 3; ++d) x[d] = v == 0 ? xn[d] * shape : x[d] + xn[3 * v + d] * shape;
  }
}

// geometry/hexahedron.hpp:436-455 + geometry/cell.hpp:80-88 : det of J[a][b] = sum_v dN_v/dxi_a x_v[b]
__device__ __forceinline__ double hex8_jacobian_det(const double *__restrict__ xn, double lx, double ly, double lz) {
  double J[3][3] = { { 0, 0, 0 }, { 0, 0, 0 }, { 0, 0, 0 } };
#pragma unroll
  for (int v = 0; v < 8; ++v) {
    double cx, cy, cz;
    hex8_corner_signs(v, cx, cy, cz);
    const double fx = 1 + cx * lx, fy = 1 + cy * ly, fz = 1 + cz * lz;
    const double g[3] = { cx * (fy * fz / 8), cy * (fz * fx / 8), cz * (fx * fy / 8) };
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
      for (int b = 0; b < 3; ++b) J[a][b] += g[a] * xn[3 * v + b];
  }
  return J[0][0] * (J[1][1] * J[2][2] - J[1][2] * J[2][1]) - J[0][1] * (J[1][0] * J[2][2] - J[1][2] * J[2][0]) +
         J[0][2] * (J[1][0] * J[2][1] - J[1][1] * J[2][0]);
}

// basis/lagrange.hpp:108-120 : values of the Q 1-D Lagrange polynomials at the three abscissae a VTK tri-quadratic
// hexahedron uses (-1, 0, +1), same operation order; computed once on the host (the kernel would otherwise spend its
// time on 54 FP64 divisions per sample)
struct VtkLineValues { double v[3][kMaxQ]; };

static VtkLineValues vtk_line_values(const Basis &basis) {
  VtkLineValues t{};
  for (int a = 0; a < 3; ++a) {
    const double x = double(a - 1);
    for (int i = 0; i < basis.q; ++i) {
      volatile double v = 1.0;  // volatile: one rounding per operation, as the reference's scalar loop
      for (int j = 0; j < basis.q; ++j)
        if (i != j) { volatile double q = (x - basis.x[j]) / (basis.x[i] - basis.x[j]); v = v * q; }
      t.v[a][i] = v;
    }
  }
  return t;
}

template <int Q, bool kFR>
__global__ void vtk_sample_nodal_kernel(DeviceTables tab, VtkLineValues line, const double *__restrict__ corner_xyz,
                                        const double *__restrict__ u, double *__restrict__ coords,
                                        double *__restrict__ values) {
  constexpr int N = Q * Q * Q;
  const int64_t idx = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= tab.n_local * 27) return;
  const int64_t cell = idx / 27;
  const int pt = int(idx - cell * 27);
  const int ax = kVtkHex27[pt][0], ay = kVtkHex27[pt][1], az = kVtkHex27[pt][2];
  const double lx = ax, ly = ay, lz = az;
  const double *xn = corner_xyz + cell * 24;
  double x[3];
  hex8_local_to_global(xn, lx, ly, lz, x);
  double vx[Q], vy[Q], vz[Q];
#pragma unroll
  for (int q = 0; q < Q; ++q) { vx[q] = line.v[ax + 1][q]; vy[q] = line.v[ay + 1][q]; vz[q] = line.v[az + 1][q]; }
  double val[5] = { 0, 0, 0, 0, 0 };
  const double *c = u + cell * (N * 5);
  int node = 0;
#pragma unroll
  for (int i = 0; i < Q; ++i)
#pragma unroll
    for (int j = 0; j < Q; ++j) {
      const double vxy = vx[i] * vy[j];  // basis/lagrange.hpp:272-288
#pragma unroll
      for (int k = 0; k < Q; ++k, ++node) {
        const double phi = vxy * vz[k];
#pragma unroll
        for (int m = 0; m < 5; ++m) val[m] += c[node * 5 + m] * phi;
      }
    }
  if (kFR) {  // polynomial/hexahedron.hpp:230-234
    const double rdet = 1.0 / hex8_jacobian_det(xn, lx, ly, lz);
#pragma unroll
    for (int m = 0; m < 5; ++m) val[m] *= rdet;
  }
#pragma unroll
  for (int d = 0; d < 3; ++d) coords[idx * 3 + d] = x[d];
#pragma unroll
  for (int m = 0; m < 5; ++m) values[idx * 5 + m] = val[m];
}

template <int P>
__global__ void vtk_sample_modal_kernel(DeviceTables tab, const double *__restrict__ corner_xyz,
                                        const double *__restrict__ u, double *__restrict__ coords,
                                        double *__restrict__ values) {
  constexpr int N = ModalShape<P>::N;
  const int64_t idx = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= tab.n_local * 27) return;
  const int64_t cell = idx / 27;
  const int pt = int(idx - cell * 27);
  double x[3];
  hex8_local_to_global(corner_xyz + cell * 24, kVtkHex27[pt][0], kVtkHex27[pt][1], kVtkHex27[pt][2], x);
  double phi[N];
  modal_basis<P>(tab.m_S + cell * N * N, tab.m_center + cell * 3, x[0], x[1], x[2], phi);
  const Cons v = modal_value<P>(u + cell * (N * 5), phi);
#pragma unroll
  for (int d = 0; d < 3; ++d) coords[idx * 3 + d] = x[d];
#pragma unroll
  for (int m = 0; m < 5; ++m) values[idx * 5 + m] = v.c[m];
}

cudaError_t LaunchVtkSample(const DeviceTables &tab, const Basis &basis, int scheme, int degree,
                            const double *corner_xyz, const double *u, double *coords, double *values,
                            cudaStream_t stream) {
  const int64_t total = tab.n_local * 27;
  if (total <= 0) return cudaSuccess;
  const int threads = 108;  // four cells per block
  const unsigned blocks = unsigned((total + threads - 1) / threads);
  const VtkLineValues line = vtk_line_values(basis);
#define MCFD_NODAL(QQ) \
  if (scheme == MCFD_SCHEME_FR) vtk_sample_nodal_kernel<QQ, true><<<blocks, threads, 0, stream>>>(tab, line, corner_xyz, u, coords, values); \
  else vtk_sample_nodal_kernel<QQ, false><<<blocks, threads, 0, stream>>>(tab, line, corner_xyz, u, coords, values)
  if (scheme == MCFD_SCHEME_DG_MODAL) {
    switch (degree) {
      case 1: vtk_sample_modal_kernel<1><<<blocks, threads, 0, stream>>>(tab, corner_xyz, u, coords, values); break;
      case 2: vtk_sample_modal_kernel<2><<<blocks, threads, 0, stream>>>(tab, corner_xyz, u, coords, values); break;
      case 3: vtk_sample_modal_kernel<3><<<blocks, threads, 0, stream>>>(tab, corner_xyz, u, coords, values); break;
      default: return cudaErrorInvalidValue;
    }
  } else {
    switch (degree) {
      case 1: MCFD_NODAL(2); break;
      case 2: MCFD_NODAL(3); break;
      case 3: MCFD_NODAL(4); break;
      default: return cudaErrorInvalidValue;
    }
  }
#undef MCFD_NODAL
  return cudaGetLastError();
}

// Section::GatherFields (part.hpp:370-378): fields[f][cell] = column[cell][f]; ScatterFields is the inverse.
// A 32 x 32 shared-memory tile keeps both the reads and the writes coalesced.
template <bool kGather>
__global__ void transpose_fields_kernel(int64_t n_cells, int n_fields, const double *__restrict__ in,
                                        double *__restrict__ out) {
  __shared__ double tile[32][33];
  const int64_t c0 = int64_t(blockIdx.x) * 32;
  const int f0 = blockIdx.y * 32;
  // column side: [cell][field] (field fastest); field side: [field][cell] (cell fastest)
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    if (kGather) {
      const int64_t c = c0 + r; const int f = f0 + threadIdx.x;
      if (c < n_cells && f < n_fields) tile[r][threadIdx.x] = in[c * n_fields + f];
    } else {
      const int f = f0 + r; const int64_t c = c0 + threadIdx.x;
      if (c < n_cells && f < n_fields) tile[threadIdx.x][r] = in[int64_t(f) * n_cells + c];
    }
  }
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    if (kGather) {
      const int f = f0 + r; const int64_t c = c0 + threadIdx.x;
      if (c < n_cells && f < n_fields) out[int64_t(f) * n_cells + c] = tile[threadIdx.x][r];
    } else {
      const int64_t c = c0 + r; const int f = f0 + threadIdx.x;
      if (c < n_cells && f < n_fields) out[c * n_fields + f] = tile[r][threadIdx.x];
    }
  }
}

cudaError_t LaunchTransposeFields(bool gather, int64_t n_cells, int n_fields, const double *in, double *out,
                                  cudaStream_t stream) {
  if (n_cells <= 0) return cudaSuccess;
  const dim3 block(32, 8), grid(unsigned((n_cells + 31) / 32), unsigned((n_fields + 31) / 32));
  if (gather) transpose_fields_kernel<true><<<grid, block, 0, stream>>>(n_cells, n_fields, in, out);
  else transpose_fields_kernel<false><<<grid, block, 0, stream>>>(n_cells, n_fields, in, out);
  return cudaGetLastError();
}

// ---- Part::MeasureL1Error (mesh/part.hpp:1082-1096): sum over cells and quadrature points of
//   |u_h(x_q) - exact(x_q, t)| * global weight(q),   per component -- the number demo/euler/sourceless.cpp prints and the
// one check of a run that does not depend on any oracle (convergence order of a smooth exact solution).
// One thread per (cell, point); a block leaves one 5-vector of partial sums, added up on the host in block order
// (the result does not depend on scheduling).
constexpr int kL1Threads = 256;

__device__ __forceinline__ void l1_block_reduce(double (&v)[5], double *__restrict__ partial) {
  __shared__ double s_part[kL1Threads / 32][5];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 5; ++k) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v[k] += __shfl_down_sync(0xffffffffu, v[k], off);
    if (lane == 0) s_part[warp][k] = v[k];
  }
  __syncthreads();
  if (threadIdx.x < 5) {
    double sum = 0;
    for (int w = 0; w < kL1Threads / 32; ++w) sum += s_part[w][threadIdx.x];
    partial[int64_t(blockIdx.x) * 5 + threadIdx.x] = sum;
  }
}

template <int Q, bool kFR>
__global__ void __launch_bounds__(kL1Threads) l1_error_nodal_kernel(DeviceTables tab, Basis basis,
                                                                    const double *__restrict__ corner_xyz, int func,
                                                                    FuncParams params, double t,
                                                                    const double *__restrict__ u, double *__restrict__ partial) {
  constexpr int N = Q * Q * Q;
  const int64_t idx = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  double err[5] = { 0, 0, 0, 0, 0 };
  if (idx < tab.n_local * N) {
    const int64_t cell = idx / N;
    const int node = int(idx - cell * N), i = node / (Q * Q), j = (node / Q) % Q, k = node % Q;
    const double lx = basis.x[i], ly = basis.x[j], lz = basis.x[k];
    const double *xn = corner_xyz + cell * 24;
    double x[3] = { 0, 0, 0 };
    for (int v = 0; v < 8; ++v) {  // geometry/hexahedron.hpp:419-428 + LocalToGlobal, the reference's operation order
      double cx, cy, cz;
      hex8_corner_signs(v, cx, cy, cz);
      double val = (1 + cx * lx);
      val *= (1 + cy * ly);
      val *= (1 + cz * lz);
      const double shape = val / 8;
      for (int d = 0; d < 3; ++d) x[d] = v == 0 ? xn[d] * shape : x[d] + xn[3 * v + d] * shape;
    }
    const double det = tab.per_node_metrics ? tab.metrics[idx * 10 + 9] : tab.metrics[cell * 10 + 9];
    const double weight = basis.w[i] * basis.w[j] * basis.w[k] * fabs(det);  // gauss/hexahedron.hpp: local weight x |det J|
    const Cons exact = eval_state(func, params.p, x[0], x[1], x[2], t);
#pragma unroll
    for (int c = 0; c < 5; ++c) {
      const double value = kFR ? u[idx * 5 + c] / det : u[idx * 5 + c];   // projection.GetValue(q)
      err[c] = fabs(value - exact.c[c]) * weight;
    }
  }
  l1_block_reduce(err, partial);
}

template <int PE>
__global__ void __launch_bounds__(kL1Threads) l1_error_modal_kernel(DeviceTables tab, int func, FuncParams params, double t,
                                                                    const double *__restrict__ u, double *__restrict__ partial) {
  using M = ModalShape<PE>;
  constexpr int P = M::P, N = M::N, NQ = M::NQ;
  const int64_t idx = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  double err[5] = { 0, 0, 0, 0, 0 };
  if (idx < tab.n_local * NQ) {
    const int64_t cell = idx / NQ;
    const double *xw = tab.m_qxyzw + idx * 4;
    double phi[N];
    modal_basis<P>(tab.m_S + cell * N * N, tab.m_center + cell * 3, xw[0], xw[1], xw[2], phi);
    const Cons value = modal_value<P>(u + cell * N * 5, phi);
    const Cons exact = eval_state(func, params.p, xw[0], xw[1], xw[2], t);
#pragma unroll
    for (int c = 0; c < 5; ++c) err[c] = fabs(value.c[c] - exact.c[c]) * xw[3];
  }
  l1_block_reduce(err, partial);
}

cudaError_t LaunchL1Error(const DeviceTables &tab, const Basis &basis, int scheme, int degree, const double *corner_xyz,
                          int func, const double *params, double t, const double *u, double *partial, int64_t *n_blocks,
                          cudaStream_t stream) {
  FuncParams fp;
  for (int i = 0; i < MCFD_MAX_FUNC_PARAMS; ++i) fp.p[i] = params[i];
  const bool modal = scheme == MCFD_SCHEME_DG_MODAL;
  const int64_t total = tab.n_local * (modal ? tab.nq : tab.n);
  const int64_t blocks = (total + kL1Threads - 1) / kL1Threads;
  *n_blocks = blocks;
  if (!partial) return cudaSuccess;   // size query
  if (modal) {
    switch (degree + 10 * tab.elem) {  // ModalShape's shape code: degree + 10 * element type
      case 1: l1_error_modal_kernel<1><<<unsigned(blocks), kL1Threads, 0, stream>>>(tab, func, fp, t, u, partial); break;
      case 2: l1_error_modal_kernel<2><<<unsigned(blocks), kL1Threads, 0, stream>>>(tab, func, fp, t, u, partial); break;
      case 3: l1_error_modal_kernel<3><<<unsigned(blocks), kL1Threads, 0, stream>>>(tab, func, fp, t, u, partial); break;
      case 11: l1_error_modal_kernel<11><<<unsigned(blocks), kL1Threads, 0, stream>>>(tab, func, fp, t, u, partial); break;
      case 12: l1_error_modal_kernel<12><<<unsigned(blocks), kL1Threads, 0, stream>>>(tab, func, fp, t, u, partial); break;
      case 13: l1_error_modal_kernel<13><<<unsigned(blocks), kL1Threads, 0, stream>>>(tab, func, fp, t, u, partial); break;
      default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
  }
  const bool fr = scheme == MCFD_SCHEME_FR;
#define MCFD_L1(QQ) \
  if (fr) l1_error_nodal_kernel<QQ, true><<<unsigned(blocks), kL1Threads, 0, stream>>>(tab, basis, corner_xyz, func, fp, t, u, partial); \
  else l1_error_nodal_kernel<QQ, false><<<unsigned(blocks), kL1Threads, 0, stream>>>(tab, basis, corner_xyz, func, fp, t, u, partial)
  switch (basis.q) {
    case 2: MCFD_L1(2); break;
    case 3: MCFD_L1(3); break;
    case 4: MCFD_L1(4); break;
    default: return cudaErrorInvalidValue;
  }
#undef MCFD_L1
  return cudaGetLastError();
}

// ---- mcfd_measure_l1_error_static: the same sum for a caller that must not allocate ------------------------------
// (No counterpart in the reference.)  The per-block partial sums of the kernels above are scratch memory: the ordinary
// entry point borrows an idle Runge-Kutta stage column for them and adds them up in a std::vector on the host.  A
// monitor that runs WHILE the caller's own integrator owns all three columns, or inside a process whose memory is
// fully handed out (another arena holds the rest of the HBM, so cudaMalloc fails) or in which cudaMalloc / cudaFree are
// forbidden because they synchronise the whole device against peers that spin on this GPU's halo counters, has neither.
// For it the partial sums live in this file-scope array, which every GPU that loads the library holds until exit:
//   kL1StaticMaxBlocks blocks x 5 components, then 5 doubles for the total = 2 621 480 bytes.
// g_l1_scratch is only a host-side handle in host code; the device address comes from cudaGetSymbolAddress, asked once
// per device with that device current (L1StaticScratch below).
static_assert(kL1StaticMaxPoints == kL1StaticMaxBlocks * kL1Threads, "the documented limit counts points, the array counts blocks");
static __device__ __align__(256) double g_l1_scratch[(kL1StaticMaxBlocks + 1) * 5];

// Adds the partial sums in block order, one thread per component: the same additions in the same order as the host
// loop of mcfd_measure_l1_error, hence the same bits.
__global__ void l1_sum_partials_kernel(const double *__restrict__ partial, int64_t n_blocks, double *__restrict__ total) {
  if (threadIdx.x < 5) {
    double sum = 0.0;
    for (int64_t b = 0; b < n_blocks; ++b) sum += partial[b * 5 + threadIdx.x];
    total[threadIdx.x] = sum;
  }
}

namespace {
constexpr int kL1StaticMaxDevices = 64;
std::mutex g_l1_scratch_mutex;
double *g_l1_scratch_on_device[kL1StaticMaxDevices] = {};

// Device address of g_l1_scratch on the CURRENT device (looked up once per device, then cached).
cudaError_t L1StaticScratch(double **scratch) {
  int device = -1;
  cudaError_t e = cudaGetDevice(&device);
  if (e != cudaSuccess) return e;
  if (device < 0 || device >= kL1StaticMaxDevices) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(g_l1_scratch_mutex);
  if (!g_l1_scratch_on_device[device]) {
    double *address = nullptr;
    e = cudaGetSymbolAddress(reinterpret_cast<void **>(&address), g_l1_scratch);
    if (e != cudaSuccess) return e;
    if (!address) return cudaErrorInvalidSymbol;
    g_l1_scratch_on_device[device] = address;
  }
  *scratch = g_l1_scratch_on_device[device];
  return cudaSuccess;
}
}  // namespace

cudaError_t MeasureL1ErrorStatic(const DeviceTables &tab, const Basis &basis, int scheme, int degree, const double *corner_xyz,
                                 int func, const double *params, double t, const double *u, double *l1_error_host,
                                 cudaStream_t stream) {
  int64_t n_blocks = 0;
  LaunchL1Error(tab, basis, scheme, degree, corner_xyz, func, params, t, u, nullptr, &n_blocks, stream);  // size only
  if (n_blocks > kL1StaticMaxBlocks) return cudaErrorInvalidValue;   // the caller has checked this and named the limit
  if (n_blocks <= 0) {
    for (int c = 0; c < 5; ++c) l1_error_host[c] = 0.0;
    return cudaSuccess;
  }
  double *scratch = nullptr;
  cudaError_t e = L1StaticScratch(&scratch);
  if (e != cudaSuccess) return e;
  // Nothing in the array has to start at zero: every block stores its five sums, the five totals are stored, not
  // added to, and only [0, n_blocks) is read back -- what an earlier, larger call left behind is never looked at.
  e = LaunchL1Error(tab, basis, scheme, degree, corner_xyz, func, params, t, u, scratch, &n_blocks, stream);
  if (e != cudaSuccess) return e;
  double *total = scratch + kL1StaticMaxBlocks * 5;
  l1_sum_partials_kernel<<<1, 32, 0, stream>>>(scratch, n_blocks, total);
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  e = cudaMemcpyAsync(l1_error_host, total, 5 * sizeof(double), cudaMemcpyDeviceToHost, stream);
  if (e != cudaSuccess) return e;
  return cudaStreamSynchronize(stream);
}


}  // namespace mcfd
