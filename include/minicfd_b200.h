/* minicfd_b200.h -- C-ABI of the B200-native Euler residual + Runge-Kutta hot path.
 *
 * This is the drop-in boundary for miniCFD's operator interface.  The reference has
 * no FFI layer: its operator API is the C++ virtual interface
 *   mini::temporal::System<Scalar>   (include/mini/temporal/ode.hpp:15-46)
 *   mini::temporal::Solver<Scalar>   (include/mini/temporal/ode.hpp:91-103)
 *   mini::temporal::RungeKutta<N>    (include/mini/temporal/rk.hpp:13-73)
 * implemented by mini::spatial::FiniteElement<Part> (include/mini/spatial/fem.hpp:38-279)
 * and its subclasses fr::Lobatto / dg::Lobatto / dg::General.  Every entry point below
 * cites the reference method it replaces.  Plain pointers and sizes only; all pointer
 * arguments are borrowed for the duration of the call; a handle owns all device memory.
 * Every function returns 0 on success and a non-zero mcfd_status otherwise (never
 * aborts, unlike the reference's assert / cgp_error_exit()).
 *
 * Threading: one handle per GPU; calls on one handle must be serialised by the caller
 * (the reference is single-threaded per MPI rank and not re-entrant either).
 */
#ifndef MINICFD_B200_H_
#define MINICFD_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MCFD_K 5 /* conservative components: rho, rho*u, rho*v, rho*w, E */

typedef enum mcfd_status {
  MCFD_OK = 0,
  MCFD_ERR_ARG = 1,      /* bad argument / inconsistent mesh */
  MCFD_ERR_CUDA = 2,     /* CUDA runtime error (see mcfd_last_error) */
  MCFD_ERR_UNSUPPORTED = 3,
  MCFD_ERR_NCCL = 4,
  MCFD_ERR_STATE = 5
} mcfd_status;

/* Spatial scheme.  FR = spatial::fr::Lobatto (fr/lobatto.hpp:33), parametric-space
 * unknowns U^ = detJ * u (polynomial/hexahedron.hpp:257-259); DGSEM = spatial::dg::Lobatto
 * with polynomial::Hexahedron<...,kLocal=false> (dg/lobatto.hpp:29), physical unknowns. */
typedef enum mcfd_scheme { MCFD_SCHEME_FR = 2, MCFD_SCHEME_DGSEM = 1, MCFD_SCHEME_DG_MODAL = 0 } mcfd_scheme;

/* Riemann solver evaluated in the face-normal frame (riemann/rotated/euler.hpp:111-120).
 * EXACT follows riemann/euler/exact.hpp.  ROE/HLLC/AUSM are new 3-D solvers (the
 * reference's Hllc/Ausm do not compile through rotated::Euler in 3-D). */
typedef enum mcfd_riemann { MCFD_RIEMANN_EXACT = 0, MCFD_RIEMANN_ROE = 1, MCFD_RIEMANN_HLLC = 2, MCFD_RIEMANN_AUSM = 3 } mcfd_riemann;

/* MCFD_LIMITER_DUMMY is weno::Dummy (limiter/weno.hpp:266-281): every cell counts as troubled and is replaced by a copy of
 * itself -- the column does not change, but Part::Reconstruct still exchanges the halo on every SetSolutionColumn and
 * mcfd_troubled_cells reports every cell. */
typedef enum mcfd_limiter { MCFD_LIMITER_NONE = 0, MCFD_LIMITER_LAZY = 1, MCFD_LIMITER_EIGEN = 2, MCFD_LIMITER_DUMMY = 3 } mcfd_limiter;

/* Boundary-condition kinds, in the order spatial::FiniteElement::AddFluxOnBoundaries
 * applies them (fem.hpp:185-226). */
typedef enum mcfd_bc_kind {
  MCFD_BC_SOLID_WALL = 0,        /* rotated::Euler::GetFluxOnSolidWall        (rotated/euler.hpp:122-128) */
  MCFD_BC_SUPERSONIC_INLET = 1,  /* GetFluxOnSupersonicInlet                   (rotated/euler.hpp:104-110) */
  MCFD_BC_SUPERSONIC_OUTLET = 2, /* GetFluxOnSupersonicOutlet                  (rotated/euler.hpp:129-135) */
  MCFD_BC_SUBSONIC_INLET = 3,    /* GetFluxOnSubsonicInlet                     (rotated/euler.hpp:136-155) */
  MCFD_BC_SUBSONIC_OUTLET = 4,   /* GetFluxOnSubsonicOutlet                    (rotated/euler.hpp:156-172) */
  MCFD_BC_SMART = 5              /* GetFluxOnSmartBoundary                     (rotated/euler.hpp:173-193) */
} mcfd_bc_kind;

/* Built-in "given state" functions u(x, t) evaluated ON THE DEVICE for a BC group.
 * They replace the host std::function<Value(Global const&, double)> callbacks the
 * reference stores per boundary name (fem.hpp:56-58,126-141).  Arbitrary host
 * callbacks remain possible through mcfd_bc_points() + mcfd_set_bc_values(). */
typedef enum mcfd_bc_func {
  MCFD_FUNC_HOST_VALUES = 0, /* values uploaded by mcfd_set_bc_values()                         */
  MCFD_FUNC_CONSTANT = 1,    /* params[0..4] = conservative state                                */
  MCFD_FUNC_VORTEX = 2,      /* isentropic vortex advected in the xy-plane:
                                params = {beta, xc, yc, u0, v0, gamma}                           */
  MCFD_FUNC_PLANE_SPLIT = 3, /* params = {nx, ny, nz, offset, speed, left[5], right[5]}: left where
                                n.x < offset + speed*t, else right (conservative states).  Covers
                                demo/euler/shock_tube_sod.cpp:10-13 and the moving shock of
                                demo/euler/double_mach.cpp:14-31                                 */
} mcfd_bc_func;
#define MCFD_MAX_FUNC_PARAMS 16

/* One mesh part (= one MPI rank of the reference = one GPU here), hexahedra (HEXA_8,
 * CGNS node order, geometry/hexahedron.hpp:345-352).
 *  - cells:  local cells first (inner cells, then inter cells: part.hpp:818-835), then
 *            ghost cells (grouped by neighbour part, ascending global id: part.hpp:966-986);
 *            the solution column holds local cells only, cell i at offset i*K*N
 *            (part.hpp:844-857).
 *  - faces:  local | ghost | boundary (ids as fr/lobatto.hpp:192-207 assumes).  The
 *            face normal points from holder to sharer / outward (part.hpp:1521-1522).
 *            face_nodes lists the 4 corner nodes in any order; they are re-sorted by the
 *            holder's face table as geometry::SortNodesOnFace does (geometry/hexahedron.hpp:406-416).
 */
typedef struct mcfd_mesh {
  int64_t n_nodes;
  const double *node_xyz;      /* [n_nodes][3] */
  int64_t n_local_cells;       /* inner + inter */
  int64_t n_inner_cells;       /* local cells [0, n_inner) touch no ghost face */
  int64_t n_ghost_cells;
  const int64_t *cell_nodes;   /* [(n_local+n_ghost)][8] */
  int64_t n_local_faces, n_ghost_faces, n_boundary_faces;
  const int64_t *face_holder;  /* [n_faces] local cell index */
  const int64_t *face_sharer;  /* [n_faces] local or ghost (>= n_local_cells) cell index; -1 on boundary */
  const int64_t *face_nodes;   /* [n_faces][4] */
  const int32_t *bface_group;  /* [n_boundary_faces] BC group id in [0, n_bc_groups) (= one CGNS section name) */
  int32_t n_bc_groups;
  /* halo exchange lists (part.hpp:1258-1302); all NULL / 0 on a single part */
  int32_t n_neighbours;
  const int32_t *nbr_rank;     /* [n_neighbours] */
  const int64_t *send_ptr;     /* [n_neighbours+1] into send_cells */
  const int64_t *send_cells;   /* local (inter) cell indices, ascending global id per neighbour */
  const int64_t *recv_ptr;     /* [n_neighbours+1]; ghost cell g of neighbour n is n_local_cells + recv_ptr[n] + g */
  /* Element type of EVERY cell of the part (mcfd_cell_type; 0 = HEXA_8, so a zero-initialised struct means what it
   * always meant).  TETRA_4 (CGNS node order, geometry/tetrahedron.hpp:139-150): cell_nodes is [..][4], face_nodes
   * [..][3]; modal DG only, as in the reference (part.hpp:437-452 picks gauss::Tetrahedron<1|4|14|24> and
   * gauss::Triangle<1|3|6|12> by degree; FR and DGSEM are hexahedral, polynomial/hexahedron.hpp). */
  int32_t cell_type;
  int32_t reserved;
} mcfd_mesh;
typedef enum mcfd_cell_type { MCFD_CELL_HEXA_8 = 0, MCFD_CELL_TETRA_4 = 1 } mcfd_cell_type;

/* Synthetic partitioned box meshes (host only, no GPU needed).  Stands in for gmsh +
 * mesh::Shuffler::PartitionAndShuffle + Part's constructor (demo/euler/sourceless.cpp:61-71) with a
 * deterministic block split parts[0] x parts[1] x parts[2]; rank = bx + parts[0]*(by + parts[1]*bz).
 * BC groups 0..5 = x-min, x-max, y-min, y-max, z-min, z-max.  distortion > 0 moves interior nodes by
 * distortion*h*hash(global node id) so that cells are no longer affine (same on every part). */
typedef struct mcfd_box_mesh mcfd_box_mesh;
int mcfd_box_mesh_create(const int64_t n[3], const double lo[3], const double hi[3], const int32_t parts[3],
                         int32_t rank, double distortion, mcfd_box_mesh **out);
/* The same generator with a block of cells left out (cell index ranges hole_lo[d] <= i_d < hole_hi[d]; NULL = none):
 * the faces around the hole become boundary group 6.  The forward-facing step of demo/euler/forward_step.cpp
 * (input/forward_step.geo: a 3 x 1 channel with a 0.2-high step from x = 0.6) is a box with such a hole. */
int mcfd_box_mesh_create_ex(const int64_t n[3], const double lo[3], const double hi[3], const int32_t parts[3],
                            int32_t rank, double distortion, const int64_t hole_lo[3], const int64_t hole_hi[3],
                            mcfd_box_mesh **out);
/* Every hexahedron of `hexa` cut into six TETRA_4 along its main diagonal (Kuhn's split: the two triangles on a shared
 * quadrilateral agree between neighbours whose local axes agree, as in a box mesh), CGNS orientation, same part
 * layout: tetrahedra 6c .. 6c+5 of hexahedron c stay in c's category (inner | inter | ghost), send and receive
 * lists list all six of every listed hexahedron, boundary triangles inherit the group of their quadrilateral.
 * Stands in for gmsh's tetrahedral mesher (absent here) for BASELINE configs[3]. */
int mcfd_box_mesh_split_tetra(const mcfd_box_mesh *hexa, mcfd_box_mesh **out);
const mcfd_mesh *mcfd_box_mesh_view(const mcfd_box_mesh *bm);
const int64_t *mcfd_box_mesh_cell_global_ids(const mcfd_box_mesh *bm); /* [n_local + n_ghost] */
void mcfd_box_mesh_destroy(mcfd_box_mesh *bm);

/* ---- Partition + ghost-index pipeline for general hexahedral meshes ----
 * Stands where the reference has mesh::Shuffler::PartitionAndShuffle + Part's constructor.  METIS itself is absent from
 * the reference tree (un-vendored submodule pvc1989/METIS @ 3d8fc59): the cell -> part vector is an INPUT here (the
 * caller's METIS, or any other partitioner); everything downstream of it follows the reference:
 *   dual graph        METIS_MeshToDual as called at mesh/metis.hpp:191-197 with n_common_nodes = 3 (shuffler.hpp:525-526):
 *                     METIS 5.x libmetis/mesh.c CreateGraphDual / FindCommonElements restated (neighbour order = order
 *                     of first discovery over the cell's nodes, elements ascending per node);
 *   node parts        metis::GetNodeParts (mesh/metis.hpp:316-337): a node belongs to the lowest part using it;
 *   new order         mesh::GetNewOrder / ShuffleData / ShuffleConnectivity (mesh/shuffler.hpp:36-104);
 *   partition info    the per-part text of Shuffler::WritePartitionInfo (mesh/shuffler.hpp:314-506);
 *   part mesh         Part's build order (mesh/part.hpp:761-1079,1440-1533): local cells inner | inter by ascending
 *                     metis id, ghost cells by neighbour part then ascending metis id, local faces in inner-adjacency
 *                     order (holder = the smaller metis id), ghost faces in interpart-adjacency order, boundary faces
 *                     by section then part order, send / recv lists ascending metis id. */
typedef struct mcfd_global_mesh {
  int64_t n_nodes;
  const double *node_xyz;        /* [n_nodes][3] */
  int64_t n_cells;
  const int64_t *cell_nodes;     /* [n_cells][8] 0-based, HEXA_8 in CGNS order; the cell index is the "metis id" */
  int64_t n_boundary_faces;
  const int64_t *bface_nodes;    /* [n_boundary_faces][4] 0-based, any order */
  const int32_t *bface_group;    /* [n_boundary_faces] in [0, n_bc_groups): one CGNS boundary section each */
  int32_t n_bc_groups;
} mcfd_global_mesh;
typedef struct mcfd_partitioner mcfd_partitioner;
int mcfd_partitioner_create(const mcfd_global_mesh *mesh, const int32_t *cell_parts, int32_t n_parts,
                            mcfd_partitioner **out);
void mcfd_partitioner_destroy(mcfd_partitioner *p);
const int64_t *mcfd_partitioner_dual_range(const mcfd_partitioner *p);   /* [n_cells + 1] */
const int64_t *mcfd_partitioner_dual_index(const mcfd_partitioner *p);   /* [range[n_cells]] */
const int32_t *mcfd_partitioner_node_parts(const mcfd_partitioner *p);   /* [n_nodes] */
const int32_t *mcfd_partitioner_face_parts(const mcfd_partitioner *p);   /* [n_boundary_faces]: Shuffler::FillFaceParts */
const int32_t *mcfd_partitioner_cells_new_to_old(const mcfd_partitioner *p); /* [n_cells] */
const int32_t *mcfd_partitioner_nodes_new_to_old(const mcfd_partitioner *p); /* [n_nodes] */
/* partition/<part>.txt as the reference writes it (zone 1; section 1 = the cells, section 2 + g = boundary group g);
 * returns the length needed in *needed, fills buf when capacity suffices */
int mcfd_partitioner_info_text(const mcfd_partitioner *p, int32_t part, char *buf, int64_t capacity, int64_t *needed);
/* the mcfd_mesh of one part; owner type and accessors shared with the box generator */
int mcfd_partitioner_part_mesh(const mcfd_partitioner *p, int32_t part, struct mcfd_box_mesh **out);
/* the reference's free helpers under their own signatures (int index type, 1-based connectivity) */
int mcfd_get_new_order(const int32_t *parts, int32_t n, int32_t *new_to_old, int32_t *old_to_new);
int mcfd_shuffle_connectivity(const int32_t *old_to_new_for_nodes, int32_t n_nodes, const int32_t *new_to_old_for_cells,
                              int32_t n_cells, int32_t npe, int32_t *cell_nodes_1based);
int mcfd_get_node_parts(int32_t n_cells, int32_t n_nodes, const int32_t *range, const int32_t *index,
                        const int32_t *cell_parts, int32_t n_parts, int32_t *node_parts);

typedef struct mcfd_config {
  int32_t scheme;    /* mcfd_scheme */
  int32_t degree;    /* polynomial degree p: 1, 2 or 3 */
  int32_t riemann;   /* mcfd_riemann */
  int32_t limiter;   /* mcfd_limiter (modal DG only) */
  int32_t rk_order;  /* 1, 2 or 3 (temporal::RungeKutta<kOrders>) */
  int32_t device;    /* CUDA device ordinal */
  int32_t rank, world;
  double gamma;      /* ratio of specific heats (IdealGas<Scalar, kGamma>) */
  double weno_w0, weno_eps;
  int32_t flags;     /* MCFD_FLAG_* */
  int32_t reserved;
} mcfd_config;

#define MCFD_FLAG_FORCE_PER_NODE_METRICS 1 /* never compress affine cells to per-cell constants */
#define MCFD_FLAG_NO_SHAPE_CLASSES 8      /* modal DG: never tabulate the basis per shape class (general kernels only) */
#define MCFD_FLAG_KEEP_FACE_ORDER 4       /* tuning: do not re-sort faces by holder side inside each category */

typedef struct mcfd_solver mcfd_solver; /* opaque */

/* Build per-part device tables (what Part + Projection + Gauss + fr::Lobatto's caches hold:
 * part.hpp:470-494, polynomial/hexahedron.hpp:166-214, fr/lobatto.hpp:76-207). */
int mcfd_create(const mcfd_config *cfg, const mcfd_mesh *mesh, mcfd_solver **out);
int mcfd_destroy(mcfd_solver *s);
const char *mcfd_last_error(const mcfd_solver *s); /* s may be NULL: error of the last failed create */

/* Sizes: column length = K*N*n_local_cells (Part::GetCellDataSize, part.hpp:852-857). */
int64_t mcfd_column_size(const mcfd_solver *s);
int32_t mcfd_fields_per_cell(const mcfd_solver *s); /* Cell::kFields = K*N */

/* temporal::System::SetTime (fem.hpp:150-152). */
int mcfd_set_time(mcfd_solver *s, double t);
/* temporal::System::SetSolutionColumn / GetSolutionColumn (fem.hpp:153-172): host columns. */
int mcfd_set_solution(mcfd_solver *s, const double *column, int64_t n);
int mcfd_get_solution(mcfd_solver *s, double *column, int64_t n);
/* temporal::System::GetResidualColumn (fem.hpp:173-183; dg/lobatto.hpp:124-139). */
int mcfd_get_residual(mcfd_solver *s, double *column, int64_t n);
/* Cell::Approximate for all local cells (polynomial/hexahedron.hpp:499-513) with a built-in
 * function evaluated on the device at time t. */
int mcfd_approximate(mcfd_solver *s, int32_t func, const double *params, int32_t n_params, double t);
/* Part::MeasureL1Error (mesh/part.hpp:1082-1096) of this part against a built-in exact solution at time t:
 * l1_error[c] = sum over local cells and quadrature points of |u_h(x_q) - exact(x_q, t)|_c * global weight(q).
 * (The quadrature points are the solution nodes for FR/DGSEM and the Legendre points for modal DG; the parts of a
 * partitioned mesh add up, as the ranks of the reference do.) */
int mcfd_measure_l1_error(mcfd_solver *s, int32_t func, const double *params, int32_t n_params, double t,
                          double *l1_error /* [5] */);
/* No counterpart in the reference.  The same five sums, bit for bit, for a caller that cannot allocate: an in-run error
 * monitor inside a host code whose own arena already holds the rest of the HBM (cudaMalloc would fail), or one that
 * runs next to peers spinning on this GPU's halo counters (cudaMalloc / cudaFree stall the whole device), or one whose
 * integrator owns all three stage columns -- mcfd_measure_l1_error keeps its per-block partial sums in an idle stage
 * column and adds them in a host vector; this variant keeps them in a fixed array inside the library and adds them
 * on the device.  It calls no cudaMalloc, cudaFree or allocating library routine; the host heap is touched only when
 * the call fails and leaves its message for mcfd_last_error().
 *   Limit: n_local * (quadrature points per cell) <= 16 777 216 (65 536 blocks of 256; 64^3 hexahedra at p = 2 use
 *   7.1 M).  Beyond it: MCFD_ERR_UNSUPPORTED, with the count in mcfd_last_error().  The array is 2.6 MB and every GPU
 *   that loads the library holds it until the process exits.
 *   Nodal schemes read the cell-corner table that mcfd_approximate / mcfd_measure_l1_error / mcfd_vtk_sample upload
 *   on their first call; if none of them has run, this call refuses (MCFD_ERR_UNSUPPORTED) rather than allocate it.
 *   The first call on each device looks up the array's device address; make that call during set-up, before any
 *   stream capture.
 *   One array per GPU, shared by every solver and every call: calls that overlap on the same GPU (two solvers, two
 *   host threads) race on it -- serialise them.  Nothing in it needs clearing between calls; it keeps the last
 *   call's partial sums, which the next call overwrites before it reads. */
int mcfd_measure_l1_error_static(mcfd_solver *s, int32_t func, const double *params, int32_t n_params, double t,
                                 double *l1_error /* [5] */);

/* Boundary conditions (fem.hpp:126-147).  A group is one boundary name; this one call is
 * SetSolidWall / SetSupersonicOutlet (func ignored) or SetSupersonicInlet / SetSubsonicInlet /
 * SetSubsonicOutlet / SetSmartBoundary (kind + given-state function).  Groups that were never
 * set contribute no flux, exactly like an unregistered name in the reference. */
int mcfd_set_boundary(mcfd_solver *s, int32_t group, int32_t kind, int32_t func, const double *params,
                      int32_t n_params);
int64_t mcfd_bc_point_count(const mcfd_solver *s, int32_t group);
int mcfd_bc_points(const mcfd_solver *s, int32_t group, double *xyz /* [n][3] */, int64_t n);
int mcfd_set_bc_values(mcfd_solver *s, int32_t group, const double *values /* [n][5] */, int64_t n);

/* temporal::RungeKutta<rk_order>::Update applied n_steps times from t (rk.hpp:22-24,36-43,
 * 55-72), entirely device-resident: t_k = t + k*dt. */
int mcfd_rk_update(mcfd_solver *s, double t, double dt, int32_t n_steps);

/* Multi-GPU: the NCCL communicator used for the halo exchange that replaces
 * Part::ShareGhostCellCoeffs / UpdateGhostCellCoeffs (part.hpp:1258-1302).
 * unique_id is the 128-byte ncclUniqueId produced by mcfd_nccl_unique_id on rank 0 and
 * distributed by the caller (e.g. torch.distributed broadcast). */
int mcfd_nccl_unique_id(void *unique_id_128);
int mcfd_attach_nccl(mcfd_solver *s, const void *unique_id_128, int32_t rank, int32_t world);
/* How mcfd_rk_update exchanges the ghost-cell coefficients after mcfd_attach_nccl: 0 = no communicator,
 * 1 = NCCL send/recv, 2 = direct stores into the neighbours' columns over peer-mapped memory (NVLink) with
 * arrival counters -- chosen at attach time when every rank could map all of its neighbours and no limiter is
 * configured (MCFD_HALO_P2P=0 in the environment keeps NCCL). */
int mcfd_halo_transport(const mcfd_solver *s);
/* COLLECTIVE over the communicator: every rank closes the peer mappings it imported, all ranks meet, the communicator
 * goes (MPI_Finalize's place in the reference's Main).  Call it on every rank before mcfd_destroy: a solver that still
 * has peer links when it is destroyed keeps the buffers its neighbours mapped until the process exits (freeing an
 * exported allocation under an importer is undefined in CUDA), because mcfd_destroy itself is not collective. */
int mcfd_detach_nccl(mcfd_solver *s);
/* Test hook: copy ghost-cell coefficients in by hand (what UpdateGhostCellCoeffs unpacks). */
int mcfd_set_ghost_solution(mcfd_solver *s, const double *ghost_column, int64_t n);

/* Measurement hooks (bench.py): device pointers and per-kernel timing on the library's stream. */
typedef struct mcfd_timing {
  double ms_total;      /* CUDA-event time of the last mcfd_rk_update call */
  double ms_face;       /* sum over launches of the Riemann face kernels (after mcfd_set_profiling(s, 1)) */
  double ms_cell;       /* sum over launches of the fused volume+correction+RK kernels, ditto */
  int64_t n_face_launches, n_cell_launches, n_other_launches;
} mcfd_timing;
int mcfd_get_timing(const mcfd_solver *s, mcfd_timing *out);
int mcfd_set_profiling(mcfd_solver *s, int32_t per_kernel_events);
/* Troubled cells (limiter/weno.hpp:152-180, what Part::Reconstruct branches on at mesh/part.hpp:1316-1326):
 * flags[n_local] (if not NULL) = 1 for the cells the most recent limiter pass rebuilt; *count (if not NULL) = the
 * running number of rebuilds over every pass since mcfd_create (weno::Lazy rebuilds every cell of every pass). */
int mcfd_troubled_cells(mcfd_solver *s, int32_t *flags, int64_t *count);
/* Shape classes of a modal part: how many distinct cell shapes (up to a translation) the mesh has; in *n_classes (if
 * not NULL) how many classes the cell kernel batches by (shape x arrangement of the face points on every side).
 * > 0: the cell kernel contracts batches of 8 same-class cells against per-class tables of phi_n(x_q),
 * w_q grad phi_n(x_q) and phi_n at the face points on the FP64 tensor cores (every undistorted box mesh: 1 shape; its
 * Kuhn tetrahedra: 6); 0: the mesh does not repeat (or MCFD_FLAG_NO_SHAPE_CLASSES, or a nodal scheme) and the general
 * kernels run.  A data-layout choice of this library, not a concept of the reference: results agree to rounding. */
int32_t mcfd_shape_classes(const mcfd_solver *s, int32_t *n_classes);
/* The same answer from the host tables alone (no device needed; what mcfd_create would find), plus whether the face
 * frames had to be kept per quadrature point (curved faces) instead of per face. */
int mcfd_mesh_shape_classes(const mcfd_config *cfg, const mcfd_mesh *mesh, int32_t *n_shapes, int32_t *n_classes,
                            int32_t *per_point_frames);
/* ---- After the path: what Main() does with the column between steps ---- */
/* Section::GatherFields / ScatterFields (mesh/part.hpp:370-387): the current column as the field-major arrays the
 * reference's CGNS writer stores, fields[i_field][i_cell] with i_field = node * 5 + component (0-based here,
 * `coeff.reshaped()[i_field]`), i_cell = local cell id; n = mcfd_column_size().  Scatter is the restart direction
 * and, like SetSolutionColumn on a limited scheme, runs the limiter on what it read. */
int mcfd_gather_fields(mcfd_solver *s, double *fields, int64_t n);
int mcfd_scatter_fields(mcfd_solver *s, const double *fields, int64_t n);
/* vtk::Writer::Prepare (mesh/vtk.hpp:341-377): every local hexahedron sampled at the 27 nodes of a VTK tri-quadratic
 * hexahedron (order of Hexahedron27<Local>::locals, vtk.hpp:229-246) with LocalToGlobalAndValue
 * (polynomial/hexahedron.hpp:239-243, polynomial/projection.hpp:120-124).
 * coords[n_points][3], values[n_points][5], n_points = 27 * n_local_cells. */
int mcfd_vtk_sample(mcfd_solver *s, double *coords, double *values, int64_t n_points);
/* vtk::Writer::WriteSolutions (mesh/vtk.hpp:385-487): writes <directory>/<soln_name>/<rank>.vtu for this part
 * (ascii point data, base64 points with a UInt64 byte count, VTK cell type 29) and, on rank 0, the
 * <directory>/<soln_name>.pvtu that lists the pieces of all `world` parts.  field_names: the 5 names
 * Part::SetFieldNames was given (demo/euler/sourceless.cpp:72); NULL = those of the Euler demos. */
int mcfd_write_vtu(mcfd_solver *s, const char *directory, const char *soln_name, const char *const *field_names);
/* The host half of mcfd_write_vtu on its own (no device involved): the same files from points and values the caller
 * holds, coords[27 * n_cells][3], values[27 * n_cells][5]. */
int mcfd_write_vtu_points(const char *directory, const char *soln_name, const char *const *field_names, int32_t rank,
                          int32_t world, int64_t n_cells, const double *coords, const double *values);
void *mcfd_device_solution(mcfd_solver *s); /* device pointer of the current solution column */
void *mcfd_stream(mcfd_solver *s);          /* cudaStream_t the solver launches on */

/* sizeof() of the ABI structs as this library was compiled (0 mcfd_mesh, 1 mcfd_config, 2 mcfd_timing):
 * lets a foreign-language binding check its own struct mirrors. */
int64_t mcfd_struct_size(int32_t which);

/* Standalone device entry points used by the unit tests that replay the reference's
 * known-answer tests (test/riemann/euler/exact.cpp, test/riemann/rotated/euler.cpp). */
int mcfd_riemann_flux(int32_t riemann, double gamma, int64_t n, const double *frames /* [n][9] n,t,b */,
                      const double *u_left /* [n][5] */, const double *u_right /* [n][5] */, double *flux /* [n][5] */);
int mcfd_boundary_flux(int32_t kind, double gamma, int64_t n, const double *frames, const double *u_inner,
                       const double *u_given, double *flux);

#ifdef __cplusplus
}
#endif
#endif /* MINICFD_B200_H_ */
