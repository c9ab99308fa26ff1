// solver.hpp -- host-side state of one mcfd_solver (= one mesh part on one B200) and the
// kernel launch interface between the C-ABI (api.cpp), the table builder (setup.cpp) and the
// CUDA translation units (face_kernels.cu, cell_kernels.cu).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <string>
#include <vector>

#include "minicfd_b200.h"

namespace mcfd {

constexpr int kMaxQ = 4;

// 1-D nodal basis data shared by all cells (gauss/lobatto.hpp, basis/lagrange.hpp, basis/vincent.hpp)
struct Basis {
  int q;                     // points per line = degree + 1
  double x[kMaxQ];           // Lobatto points
  double w[kMaxQ];           // Lobatto weights
  double D[kMaxQ][kMaxQ];    // D[a][m] = l_m'(x_a)
  double g_prime;            // Vincent g_right'(1) with Huynh's lumping constant
};

// One boundary group = one CGNS boundary-section name of the reference (fem.hpp:126-147)
struct BcGroup {
  int32_t kind;              // mcfd_bc_kind, -1 = never set (contributes no flux)
  int32_t func;              // mcfd_bc_func
  double params[MCFD_MAX_FUNC_PARAMS];
  const double *host_values; // device copy of user-evaluated states [n_points][5] (FUNC_HOST_VALUES)
};

// Tables built on the host by setup.cpp (what Part/Projection/Gauss/Face and the FR/DG face
// caches hold in the reference), then uploaded once.
struct HostTables {
  int q, n, fq;
  int64_t n_local, n_ghost, n_inner, n_faces, n_local_faces, n_ghost_faces, n_boundary_faces;
  bool per_node_metrics, per_point_frames;
  std::vector<double> metrics;        // [(n_local+n_ghost)][per_node ? n : 1][10] : J*[b][a] (9), det J
  std::vector<int32_t> cell_face;     // [n_local][6] face id or -1
  std::vector<uint8_t> cell_code;     // [n_local][6] bit0 = sharer side, bits1-3 = lattice perm -> face point
  std::vector<int32_t> face_holder, face_sharer;  // [n_faces]; sharer -1 on boundary faces
  std::vector<uint16_t> face_code;    // [n_faces] hside | sside<<3 | perm(holder point -> sharer lattice)<<6
  std::vector<double> face_frame;     // [n_faces][per_point ? fq : 1][9] : n, t, b
  std::vector<int32_t> pt_holder, pt_sharer;  // [n_faces][fq] node record (cell * n + node) on either side; sharer -1 on boundaries
  std::vector<int32_t> bface_group, bface_local;  // [n_boundary_faces] group, index of the face within its group
  std::vector<double> bface_xyz;      // [n_boundary_faces][fq][3] in face-point order
  std::vector<int64_t> group_faces;   // faces per group
  std::vector<double> corner_xyz;     // [n_local][8][3] (for mcfd_approximate / node coordinates)
  int n_groups;
  // modal DG (scheme MCFD_SCHEME_DG_MODAL): n = Taylor basis size, nq = quadrature points per cell (q^3 Legendre
  // points of a hexahedron; 4 / 14 / 24 of a tetrahedron), fq per face (q^2; 3 / 6 / 12 of a triangle)
  bool modal = false;
  int nq = 0;
  int elem = 0;    // mcfd_cell_type of every cell
  int sides = 6;   // faces per cell (cell_face / cell_code keep their stride of 6)
  std::vector<double> m_center;      // [(n_local+n_ghost)][3]   basis centre = image of the local origin
  std::vector<double> m_S;           // [(n_local+n_ghost)][n][n] lower-triangular orthonormalisation matrix
  std::vector<double> m_volume;      // [(n_local+n_ghost)]
  std::vector<double> m_qxyzw;       // [(n_local+n_ghost)][nq][4] quadrature point and global weight
  std::vector<double> f_xyzw;        // [n_faces][fq][4]         face quadrature point and global weight
  std::vector<int32_t> adj_ptr, adj_cell, adj_face;  // neighbours through interior faces in MESH order (limiter)
  // Shape classes (modal DG on meshes whose cells repeat up to a translation: every box mesh, and the six Kuhn
  // tetrahedra of a box): the basis values and weighted gradients at the quadrature points depend on the class only,
  // so they are tabulated once per class instead of being rebuilt from S and the points for every cell and stage.
  // n_classes = 0: no such structure (distorted / unstructured meshes) -- the general kernels run.
  int n_classes = 0, n_patterns = 0, n_shapes = 0;   // class = (shape, arrangement of the face points on every side)
  std::vector<int32_t> cell_class;    // [n_local]
  std::vector<int32_t> class_cells;   // [n_local] cells grouped by class
  std::vector<int32_t> class_ptr;     // [n_classes + 1] into class_cells
  std::vector<double> mm_A1;          // [n_classes][QT * 8][K1]   phi_n(x_q), row q, zero padded
  std::vector<double> mm_A2;          // [n_classes][MT * 8][R]    w_q d phi_n / d x_d (x_q), row n, column q * 3 + d
  std::vector<double> mm_A3;          // [n_classes][MT * 8][J]    phi_n at the face points, row n, column (side pair, point)
  std::vector<double> cls_S;          // [n_classes][n][n]
  std::vector<double> cls_P;          // [n_classes][6][n][n]  limiter: the neighbour across a side re-projected on my basis
  std::vector<double> cls_W;          // [n_classes][n][n]     limiter: smoothness as a quadratic form of the Taylor coefficients (p <= 2)
  // halo
  std::vector<int32_t> nbr_rank;
  std::vector<int64_t> send_ptr, recv_ptr;
  std::vector<int32_t> send_cells;
};

int BuildBasis(int degree, int scheme, Basis *basis);
// returns 0 or an mcfd_status; `error` receives a message
int BuildTables(const mcfd_config &cfg, const mcfd_mesh &mesh, const Basis &basis, HostTables *out, std::string *error);

// Device view handed to kernels (plain pointers)
struct DeviceTables {
  int q, n, fq;
  int64_t n_local, n_ghost, n_faces, n_local_faces, n_ghost_faces, n_boundary_faces;
  int per_node_metrics, per_point_frames;
  const double *metrics;
  const int32_t *cell_face;
  const uint8_t *cell_code;
  const int32_t *face_holder, *face_sharer;
  const uint16_t *face_code;
  const double *face_frame;
  const int32_t *pt_holder, *pt_sharer;
  const int32_t *bface_group, *bface_local;
  const double *bface_xyz;
  const BcGroup *groups;
  int n_groups;
  // modal DG
  int nq;
  int elem;        // mcfd_cell_type
  int64_t n_inner;
  const double *m_center, *m_S, *m_volume, *m_qxyzw, *f_xyzw;
  const int32_t *adj_ptr, *adj_cell, *adj_face;
  int n_classes;   // > 0: shape-class tables exist (see HostTables)
  int n_batches;   // sum over classes of ceil(cells of the class / 8)
  const int32_t *class_cells, *class_ptr;
  const double *mm_A1, *mm_A2, *mm_A3, *cls_S, *cls_P, *cls_W;
  const int32_t *cell_class;
  // Exact::star_u of the Riemann object of every interior / ghost face point (exact.hpp:64), carried from solve to
  // solve: [n_faces * fq], zero at creation like a fresh object.  Written by every ordinary solve, read by vacuum ones.
  double *star_u;
};

// Common-flux buffer layout: face points (gp = face * fq + point) in chunks of 32, component-major
// inside a chunk.  The face kernel (one thread per point) then writes 256 contiguous bytes per
// component and warp; the cell kernel reads the fq points of one face as one contiguous run.
#ifdef __CUDACC__
__host__ __device__
#endif
inline int64_t FstarIndex(int64_t gp, int k) { return (gp >> 5) * (32 * MCFD_K) + k * 32 + (gp & 31); }
inline int64_t FstarSize(int64_t n_points) { return ((n_points + 31) / 32) * (32 * MCFD_K); }

// RK stage combination, mirroring the in-place column arithmetic of temporal/rk.hpp:
//   kind 0: out = u + (r*dt)                        (Euler::Update, ode.hpp:119-131)
//   kind 1: out = (u0*3 + (u + r*dt)) / 4           (rk.hpp:59-63)
//   kind 2: out = ((u + r*dt)*2 + u0) / 3           (rk.hpp:65-70)
//   kind 3: out = ((u + r*dt) + u0) * 0.5           (rk.hpp:40-42)
//   kind 4: out = r                                 (GetResidualColumn only)
struct StageOp { int kind; double dt; };

// UpdateGhostCellCoeffs' MPI_Waitall done inside the ghost-face kernel (face_kernels.cu: wait_for_peers)
struct PeerWait {
  const unsigned long long *flags;  // this rank's arrival counters; nullptr: nothing to wait for
  int n_neighbours;
  unsigned long long seq;
  long long timeout_cycles;
  int *error;
};

// face_kernels.cu
cudaError_t LaunchFaceFlux(const DeviceTables &tab, const Basis &basis, int scheme, int riemann, double gamma, double t,
                           const double *u, double *fstar, int64_t face_begin, int64_t face_end, cudaStream_t stream,
                           const PeerWait *wait = nullptr);
cudaError_t LaunchFaceFluxLocalAndBoundary(const DeviceTables &tab, const Basis &basis, int scheme, int riemann, double gamma,
                                           double t, const double *u, double *fstar, cudaStream_t stream);
cudaError_t LaunchRiemannFlux(int riemann, double gamma, int64_t n, const double *frames, const double *ul,
                              const double *ur, double *flux, cudaStream_t stream);
cudaError_t LaunchBoundaryFlux(int kind, double gamma, int64_t n, const double *frames, const double *ui,
                               const double *uo, double *flux, cudaStream_t stream);
cudaError_t LaunchApproximate(const DeviceTables &tab, const Basis &basis, int scheme, const double *corner_xyz,
                              int func, const double *params, double t, double *u, cudaStream_t stream);
cudaError_t LaunchEvalBcPoints(int func, const double *params, double t, int64_t n, const double *xyz, double *values,
                               cudaStream_t stream);
// cell_kernels.cu
cudaError_t LaunchCellUpdate(const DeviceTables &tab, const Basis &basis, int scheme, double gamma, const double *u,
                             const double *fstar, const double *u0, double *out, StageOp op, int64_t cell_begin,
                             int64_t cell_end, cudaStream_t stream);
// modal_kernels.cu : spatial::dg::General + limiter::weno on the device
struct LimiterOp { int kind; double w0, eps; };
// tay[cell][m][5] = S^t coeff of cells [cell_begin, cell_end): what the modal face kernel evaluates
cudaError_t LaunchModalTaylor(const DeviceTables &tab, int degree, const double *u, double *tay, int64_t cell_begin,
                              int64_t cell_end, cudaStream_t stream);
cudaError_t LaunchModalFaceFlux(const DeviceTables &tab, int degree, int riemann, double gamma, double t, const double *tay,
                                double *fstar, int64_t face_begin, int64_t face_end, cudaStream_t stream);
// tay_out != nullptr (degree <= 2 only): also writes the Taylor form of the new coefficients of the local cells
cudaError_t LaunchModalCellUpdate(const DeviceTables &tab, int degree, double gamma, const double *u, const double *fstar,
                                  const double *u0, double *out, double *tay_out, StageOp op, cudaStream_t stream);
cudaError_t LaunchModalApproximate(const DeviceTables &tab, int degree, int func, const double *params, double t, double *u,
                                   cudaStream_t stream);
// Part::Reconstruct on cells [cell_begin, cell_end): detect + rebuild into `fresh` / `flags`, then write back
// `tay` must hold the Taylor form of `u` for the cells of the range and their neighbours; it is kept current
cudaError_t LaunchModalLimit(const DeviceTables &tab, int degree, LimiterOp lim, double gamma, double *u, double *tay,
                             double *fresh, int32_t *flags, int64_t cell_begin, int64_t cell_end, cudaStream_t stream);
// vtk_writer.cpp : the text of vtk::Writer::WriteSolutions (mesh/vtk.hpp:385-487) from sampled points
int WriteVtuFiles(const char *directory, const char *soln_name, const char *const *field_names, int rank, int world,
                  int64_t n_cells, const double *coords, const double *values, std::string *error);
// output_kernels.cu : vtk::Writer::Prepare sampling and Section::Gather/ScatterFields
cudaError_t LaunchVtkSample(const DeviceTables &tab, const Basis &basis, int scheme, int degree, const double *corner_xyz,
                            const double *u, double *coords, double *values, cudaStream_t stream);
// Part::MeasureL1Error: per-block partial sums [n_blocks][5]; partial == nullptr only reports n_blocks
cudaError_t LaunchL1Error(const DeviceTables &tab, const Basis &basis, int scheme, int degree, const double *corner_xyz,
                          int func, const double *params, double t, const double *u, double *partial, int64_t *n_blocks,
                          cudaStream_t stream);
// mcfd_measure_l1_error_static: the same sum with the partial sums in a file-scope device array of output_kernels.cu
// (at most kL1StaticMaxBlocks blocks of 256 (cell, quadrature point) pairs); allocates nothing, writes l1_error_host[5]
// and returns after the stream has drained.
constexpr int64_t kL1StaticMaxBlocks = 65536;
constexpr int64_t kL1StaticMaxPoints = kL1StaticMaxBlocks * 256;
cudaError_t MeasureL1ErrorStatic(const DeviceTables &tab, const Basis &basis, int scheme, int degree, const double *corner_xyz,
                                 int func, const double *params, double t, const double *u, double *l1_error_host,
                                 cudaStream_t stream);
cudaError_t LaunchTransposeFields(bool gather, int64_t n_cells, int n_fields, const double *in, double *out,
                                  cudaStream_t stream);
cudaError_t LaunchPackCells(const double *u, const int32_t *cells, int64_t n_cells, int fields, double *buffer,
                            cudaStream_t stream);

// halo exchange over peer-mapped memory (cell_kernels.cu): gather + remote store + arrival counters
cudaError_t LaunchPackToPeers(const double *u, const int32_t *cells, const int32_t *cell_nbr, const int64_t *dst_offset,
                              double *const *peer_u, int64_t n_cells, int fields, unsigned long long *const *peer_flag,
                              int n_neighbours, unsigned long long seq, unsigned int *blocks_done, cudaStream_t stream);
cudaError_t LaunchWaitPeers(const unsigned long long *flags, int n_neighbours, unsigned long long seq, int *error,
                            cudaStream_t stream);

}  // namespace mcfd
