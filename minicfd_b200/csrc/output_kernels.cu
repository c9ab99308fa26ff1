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
#include "modal_device.cuh"
#include "riemann.cuh"
#include "solver.hpp"

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

}  // namespace mcfd
