/* oracle/ref/harness.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Compiles the UNMODIFIED reference templates from /root/reference/include (read-only, found
 * through -I at build time, see oracle/Makefile) into oracle/_ref/libref_<scheme>_p<P>.so and
 * drives them through the reference's own operator interface:
 *   mini::spatial::fr::Lobatto / dg::Lobatto / dg::General / dg::WithLimiterAndSource
 *   (include/mini/spatial/...), mini::temporal::RungeKutta<N> (include/mini/temporal/rk.hpp).
 * The only thing that is not reference code is `ListPart`, a stand-in for
 * mini::mesh::part::Part (include/mini/mesh/part.hpp:398) that offers the same duck-typed
 * members the spatial operators use, but is filled from the flat `mcfd_mesh` arrays instead of
 * parallel-CGNS files (CGNS/HDF5/MPI/METIS are not installed here).  Cells and
 * faces are the reference's own part::Cell / part::Face built exactly as Part does
 * (part.hpp:698-708, 739-748, 990-1032): Hexahedron8 + gauss::Hexahedron, faces sorted by the
 * holder's face table via geometry::SortNodesOnFace.  The MPI ghost exchange
 * (part.hpp:1258-1302) becomes an in-process copy between ListParts run on one thread each.
 *
 * Select the operator at compile time: -DREF_SCHEME={0 modal DG,1 DGSEM,2 FR} -DREF_DEGREE=p
 *   [-DREF_LIMITER={0 none,1 lazy,2 eigen}]  (modal only)
 */
#include <algorithm>
#include <array>
#include <barrier>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <functional>
#include <memory>
#include <ranges>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "mini/riemann/euler/types.hpp"
#include "mini/riemann/euler/eigen.hpp"
#include "mini/riemann/euler/exact.hpp"
#include "mini/riemann/euler/ausm.hpp"
#include "mini/riemann/euler/hllc.hpp"
#include "mini/riemann/rotated/euler.hpp"
#include "mini/polynomial/projection.hpp"
#include "mini/polynomial/hexahedron.hpp"
#include "mini/gauss/tetrahedron.hpp"
#include "mini/gauss/triangle.hpp"
#include "mini/mesh/part.hpp"
#include "mini/limiter/weno.hpp"
#include "mini/temporal/rk.hpp"
#include "mini/spatial/dg/general.hpp"
#include "mini/spatial/dg/lobatto.hpp"
#include "mini/spatial/fr/lobatto.hpp"
#include "mini/mesh/vtk.hpp"

#include "minicfd_b200.h"  // only for the plain mesh struct + enums shared with the product
#include "ref_functions.hpp"

#ifndef REF_SCHEME
#define REF_SCHEME 2
#endif
#ifndef REF_DEGREE
#define REF_DEGREE 2
#endif
#ifndef REF_LIMITER
#define REF_LIMITER 0
#endif
#ifndef REF_CELL  /* 0 HEXA_8, 1 TETRA_4 (modal DG only): the element types Part builds, mesh/part.hpp:437-452, 663-670, 726-736 */
#define REF_CELL 0
#endif

/* mesh/cgns.hpp defines non-template inline functions that name cg_npe; nothing calls them. */
extern "C" int cg_npe(CGNS_ENUMT(ElementType_t), int *npe) { *npe = 0; return 1; }

namespace {

using Scalar = double;
constexpr int kDimensions = 3;
constexpr int kDegrees = REF_DEGREE;
/* typedefs as demo/euler/sourceless.hpp:24-70 */
using Primitive = mini::riemann::euler::Primitives<Scalar, kDimensions>;
using Conservative = mini::riemann::euler::Conservatives<Scalar, kDimensions>;
using Gas = mini::riemann::euler::IdealGas<Scalar, 1.4>;
using Unrotated = mini::riemann::euler::Exact<Gas, kDimensions>;
using Riemann = mini::riemann::rotated::Euler<Unrotated>;
#if REF_SCHEME == 0
using Projection = mini::polynomial::Projection<Scalar, kDimensions, kDegrees, 5>;
using Gx = typename Projection::GaussOnLine;
#elif REF_SCHEME == 1
using Gx = mini::gauss::Lobatto<Scalar, kDegrees + 1>;
using Projection = mini::polynomial::Hexahedron<Gx, Gx, Gx, 5, false>;
#else
using Gx = mini::gauss::Lobatto<Scalar, kDegrees + 1>;
using Projection = mini::polynomial::Hexahedron<Gx, Gx, Gx, 5, true>;
#endif

/** Stand-in for mini::mesh::part::Part (same member names; part.hpp:398-1400). */
class ListPart {
 public:
  using Index = long;
  using Riemann = ::Riemann;
  using Projection = ::Projection;
  using Scalar = double;
  using Cell = mini::mesh::part::Cell<Index, Riemann, Projection>;
  using Face = mini::mesh::part::Face<Index, Riemann, Projection>;
  using Global = typename Cell::Global;
  using Value = typename Cell::Value;
  static constexpr int kDegrees = ::kDegrees;
  static constexpr int kComponents = 5;
  static constexpr int kFields = Cell::kFields;

#if REF_CELL == 0
  static constexpr int kCellNodes = 8, kFaceNodes = 4;
  using LagrangeOnCell = mini::geometry::Hexahedron8<Scalar>;
  using GaussOnCell = mini::gauss::Hexahedron<Gx, Gx, Gx>;
  using LagrangeOnFace = mini::geometry::Quadrangle4<Scalar, 3>;
  using GaussOnFace = mini::gauss::Quadrangle<3, Gx, Gx>;
#else  /* part.hpp:439-452 : select_t<kDegrees, ...> picks entry kDegrees of { 1, 3, 6, 12 } and { 1, 4, 14, 24 } */
  static constexpr int kCellNodes = 4, kFaceNodes = 3;
  static constexpr int kTetraPoints = ::kDegrees == 1 ? 4 : ::kDegrees == 2 ? 14 : 24;
  static constexpr int kTrianglePoints = ::kDegrees == 1 ? 3 : ::kDegrees == 2 ? 6 : 12;
  using LagrangeOnCell = mini::geometry::Tetrahedron4<Scalar>;
  using GaussOnCell = mini::gauss::Tetrahedron<Scalar, kTetraPoints>;
  using LagrangeOnFace = mini::geometry::Triangle3<Scalar, 3>;
  using GaussOnFace = mini::gauss::Triangle<Scalar, 3, kTrianglePoints>;
#endif

  std::vector<Cell> cells_;  // local (inner | inter) then ghost
  Index n_local_{0}, n_inner_{0}, n_ghost_{0};
  std::vector<Cell *> local_ptrs_;
  std::array<std::vector<Cell *>, 2> inner_and_inter_cells_;
  std::vector<std::unique_ptr<Face>> local_faces_, ghost_faces_, bound_faces_;
  std::vector<std::vector<Face *>> group_faces_;
  int rank_{0};
  // halo
  std::vector<int> nbr_rank_;
  std::vector<std::vector<Cell *>> send_cell_ptrs_, recv_cell_ptrs_;
  std::array<std::vector<std::vector<Scalar>>, 2> send_coeffs_;  // double-buffered
  int share_count_{0};
  std::vector<ListPart *> const *peers_{nullptr};
  std::barrier<> *barrier_{nullptr};

  ListPart(const mcfd_mesh &m, int rank) : rank_(rank) {
    n_local_ = m.n_local_cells; n_inner_ = m.n_inner_cells; n_ghost_ = m.n_ghost_cells;
    auto coord = [&](int64_t i) {
      return Global(m.node_xyz[3 * i], m.node_xyz[3 * i + 1], m.node_xyz[3 * i + 2]);
    };
    cells_.reserve(n_local_ + n_ghost_);
    for (Index c = 0; c < n_local_ + n_ghost_; ++c) {
      const int64_t *n = m.cell_nodes + kCellNodes * c;
#if REF_CELL == 0
      auto coords = { coord(n[0]), coord(n[1]), coord(n[2]), coord(n[3]),
                      coord(n[4]), coord(n[5]), coord(n[6]), coord(n[7]) };
#else
      auto coords = { coord(n[0]), coord(n[1]), coord(n[2]), coord(n[3]) };
#endif
      auto lagrange = std::make_unique<LagrangeOnCell>(coords);
      auto gauss = std::make_unique<GaussOnCell>(*lagrange);
      cells_.emplace_back(std::move(lagrange), std::move(gauss), c);
      cells_.back().id_ = c;
      cells_.back().inner_ = (c < n_inner_);
    }
    for (Index c = 0; c < n_local_; ++c) {
      local_ptrs_.push_back(&cells_[c]);
      inner_and_inter_cells_[c < n_inner_ ? 0 : 1].push_back(&cells_[c]);
    }
    Index n_faces = m.n_local_faces + m.n_ghost_faces + m.n_boundary_faces;
    group_faces_.resize(m.n_bc_groups);
    for (Index f = 0; f < n_faces; ++f) {
      Cell *holder = &cells_[m.face_holder[f]];
      Cell *sharer = m.face_sharer[f] >= 0 ? &cells_[m.face_sharer[f]] : nullptr;
      int64_t face_nodes[4];
      std::copy_n(m.face_nodes + kFaceNodes * f, kFaceNodes, face_nodes);
      mini::geometry::SortNodesOnFace(holder->lagrange(), m.cell_nodes + kCellNodes * m.face_holder[f],
          face_nodes, kFaceNodes);
#if REF_CELL == 0
      auto coords = { coord(face_nodes[0]), coord(face_nodes[1]), coord(face_nodes[2]),
                      coord(face_nodes[3]) };
#else
      auto coords = { coord(face_nodes[0]), coord(face_nodes[1]), coord(face_nodes[2]) };
#endif
      auto lagrange = std::make_unique<LagrangeOnFace>(coords);
      auto gauss = std::make_unique<GaussOnFace>(*lagrange);
      auto face = std::make_unique<Face>(std::move(lagrange), std::move(gauss), holder, sharer, f);
      if (f < m.n_local_faces) {  // part.hpp:1019-1032
        holder->adj_cells_.emplace_back(sharer); sharer->adj_cells_.emplace_back(holder);
        holder->adj_faces_.emplace_back(face.get()); sharer->adj_faces_.emplace_back(face.get());
        local_faces_.emplace_back(std::move(face));
      } else if (f < m.n_local_faces + m.n_ghost_faces) {  // part.hpp:1066-1078
        holder->adj_cells_.emplace_back(sharer);
        holder->adj_faces_.emplace_back(face.get());
        ghost_faces_.emplace_back(std::move(face));
      } else {
        Index b = f - m.n_local_faces - m.n_ghost_faces;
        group_faces_.at(m.bface_group[b]).push_back(face.get());
        bound_faces_.emplace_back(std::move(face));
      }
    }
    for (int n = 0; n < m.n_neighbours; ++n) {
      nbr_rank_.push_back(m.nbr_rank[n]);
      auto &send = send_cell_ptrs_.emplace_back();
      for (auto i = m.send_ptr[n]; i < m.send_ptr[n + 1]; ++i) send.push_back(&cells_[m.send_cells[i]]);
      auto &recv = recv_cell_ptrs_.emplace_back();
      for (auto i = m.recv_ptr[n]; i < m.recv_ptr[n + 1]; ++i) recv.push_back(&cells_[n_local_ + i]);
      send_coeffs_[0].emplace_back(send.size() * kFields);
      send_coeffs_[1].emplace_back(send.size() * kFields);
    }
  }

  int mpi_rank() const { return rank_; }
  /* what mesh::vtk::Writer asks of a Part (part.hpp:495-509, 1549-1563) */
  int size_{1};
  std::string directory_;
  std::array<std::string, kComponents> field_names_{ "Density", "MomentumX", "MomentumY", "MomentumZ",
      "EnergyStagnationDensity" };  /* demo/euler/sourceless.cpp:72-73 */
  int mpi_size() const { return size_; }
  const std::string &GetDirectoryName() const { return directory_; }
  const std::string &GetFieldName(int i) const { return field_names_.at(i); }
  std::ofstream GetFileStream(std::string const &soln_name, bool binary, std::string const &suffix) const {
    char temp[1024];
    std::snprintf(temp, sizeof(temp), "mkdir -p %s/%s", directory_.c_str(), soln_name.c_str());
    if (std::system(temp)) throw std::runtime_error(temp + std::string(" failed."));
    std::snprintf(temp, sizeof(temp), "%s/%s/%d.%s", directory_.c_str(), soln_name.c_str(), rank_, suffix.c_str());
    return std::ofstream(temp, std::ios::out | (binary ? (std::ios::binary) : std::ios::out));
  }
  Index CountLocalCells() const { return n_local_; }
  Index GetCellDataOffset(Index i_cell) const { return i_cell * kFields; }
  Index GetCellDataSize() const { return n_local_ * kFields; }

  std::ranges::input_range auto GetLocalCells() const {
    return std::ranges::subrange(cells_.begin(), cells_.begin() + n_local_);
  }
  std::vector<Cell *> const &GetLocalCellPointers() { return local_ptrs_; }
  std::ranges::input_range auto GetLocalFaces() const {
    auto t = [](auto &uptr) -> const Face & { return *uptr; };
    return local_faces_ | std::views::transform(t);
  }
  std::ranges::input_range auto GetGhostFaces() const {
    auto t = [](auto &uptr) -> const Face & { return *uptr; };
    return ghost_faces_ | std::views::transform(t);
  }
  std::ranges::input_range auto GetBoundaryFaces() const {
    auto t = [](auto &uptr) -> const Face & { return *uptr; };
    return bound_faces_ | std::views::transform(t);
  }
  std::ranges::input_range auto GetBoundaryFaces(std::string const &name) const {
    auto const &faces = group_faces_.at(std::stoi(name));
    auto t = [](Face *ptr) -> const Face & { return *ptr; };
    return faces | std::views::transform(t);
  }

  /* part.hpp:1258-1284, with the send buffer left in place for the peer to read. */
  void ShareGhostCellCoeffs() {
    auto &bufs = send_coeffs_[share_count_ & 1];
    for (size_t n = 0; n < send_cell_ptrs_.size(); ++n) {
      Scalar *data = bufs[n].data();
      for (auto *cell_ptr : send_cell_ptrs_[n]) data = cell_ptr->projection().WriteCoeffTo(data);
    }
  }
  /* part.hpp:1285-1302: "MPI_Waitall" = barrier, then unpack from the peers' buffers. */
  void UpdateGhostCellCoeffs() {
    if (barrier_) barrier_->arrive_and_wait();
    for (size_t n = 0; n < recv_cell_ptrs_.size(); ++n) {
      ListPart *peer = (*peers_)[nbr_rank_[n]];
      size_t k = std::find(peer->nbr_rank_.begin(), peer->nbr_rank_.end(), rank_) - peer->nbr_rank_.begin();
      Scalar const *data = peer->send_coeffs_[share_count_ & 1][k].data();
      for (auto *cell_ptr : recv_cell_ptrs_[n]) data = cell_ptr->projection().GetCoeffFrom(data);
    }
    ++share_count_;
  }
  /* part.hpp:1304-1334 verbatim in structure. */
  template <typename Callable>
  void Reconstruct(Callable &&limiter) {
    if (kDegrees == 0) return;
    ShareGhostCellCoeffs();
    Reconstruct(limiter, inner_and_inter_cells_[0]);
    UpdateGhostCellCoeffs();
    Reconstruct(limiter, inner_and_inter_cells_[1]);
  }
  template <typename Limiter>
  void Reconstruct(Limiter &&limiter, std::vector<Cell *> const &cell_ptrs) {
#if REF_SCHEME == 0
    auto troubled_cells = std::vector<Cell *>();
    for (Cell *cell_ptr : cell_ptrs) {
      if (limiter.IsNotSmooth(*cell_ptr)) troubled_cells.push_back(cell_ptr);
    }
    n_troubled_ += troubled_cells.size();
    auto new_projections = std::vector<typename Projection::Wrapper>();
    for (Cell *cell_ptr : troubled_cells) new_projections.emplace_back(limiter(*cell_ptr));
    int i = 0;
    for (Cell *cell_ptr : troubled_cells) cell_ptr->projection().coeff() = new_projections[i++].coeff();
#endif
  }
  long n_troubled_{0};
};

using Cell = ListPart::Cell;
using Global = ListPart::Global;
using Value = ListPart::Value;

#if REF_SCHEME == 0
#if REF_LIMITER == 2
using Limiter = mini::limiter::weno::Eigen<Cell>;
#elif REF_LIMITER == 1
using Limiter = mini::limiter::weno::Lazy<Cell>;
#else
using Limiter = mini::limiter::weno::Dummy<Cell>;
#endif
using Spatial = mini::spatial::dg::WithLimiterAndSource<ListPart, Limiter>;
#elif REF_SCHEME == 1
using Spatial = mini::spatial::dg::Lobatto<ListPart>;
#else
using Spatial = mini::spatial::fr::Lobatto<ListPart>;
#endif

struct Sim {
  std::vector<std::unique_ptr<ListPart>> parts;
  std::vector<ListPart *> part_ptrs;
  std::vector<std::unique_ptr<Spatial>> spatials;
  std::unique_ptr<std::barrier<>> barrier;
  int rk_order{3};
};

template <class F>
void ForEachPart(Sim *sim, F &&f) {
  int n = sim->parts.size();
  if (n == 1) { f(0); return; }
  std::vector<std::thread> threads;
  for (int r = 0; r < n; ++r) threads.emplace_back([&, r] { f(r); });
  for (auto &t : threads) t.join();
}

}  // namespace

extern "C" {

int ref_scheme() { return REF_SCHEME; }
int ref_degree() { return REF_DEGREE; }
int ref_limiter() { return REF_LIMITER; }
int ref_fields_per_cell() { return Cell::kFields; }

void *ref_create(int n_parts, const mcfd_mesh *meshes, int rk_order, double weno_w0, double weno_eps) {
  auto *sim = new Sim();
  sim->rk_order = rk_order;
  sim->parts.resize(n_parts);
  {  /* the parts are independent: build them on one thread each (a 64^3 mesh is a minute of set-up on one) */
    std::vector<std::thread> builders;
    for (int r = 0; r < n_parts; ++r)
      builders.emplace_back([&, r] { sim->parts[r] = std::make_unique<ListPart>(meshes[r], r); });
    for (auto &t : builders) t.join();
  }
  for (auto &p : sim->parts) sim->part_ptrs.push_back(p.get());
  if (n_parts > 1) sim->barrier = std::make_unique<std::barrier<>>(n_parts);
  for (auto &p : sim->parts) {
    p->peers_ = &sim->part_ptrs;
    p->barrier_ = sim->barrier.get();
#if REF_SCHEME == 0
    sim->spatials.emplace_back(std::make_unique<Spatial>(p.get(), Limiter(weno_w0, weno_eps)));
#else
    sim->spatials.emplace_back(std::make_unique<Spatial>(p.get()));
#endif
  }
  return sim;
}

void ref_destroy(void *h) { delete static_cast<Sim *>(h); }

int64_t ref_column_size(void *h, int part) { return static_cast<Sim *>(h)->parts[part]->GetCellDataSize(); }

/* fem.hpp:126-147 */
int ref_set_boundary(void *h, int group, int kind, int func, const double *params, int n_params) {
  auto *sim = static_cast<Sim *>(h);
  auto state = ref_functions::Make<Value, Global>(func, params, n_params);
  std::string name = std::to_string(group);
  for (auto &sp : sim->spatials) {
    switch (kind) {
      case MCFD_BC_SOLID_WALL: sp->SetSolidWall(name); break;
      case MCFD_BC_SUPERSONIC_OUTLET: sp->SetSupersonicOutlet(name); break;
      case MCFD_BC_SUPERSONIC_INLET: sp->SetSupersonicInlet(name, state); break;
      case MCFD_BC_SUBSONIC_INLET: sp->SetSubsonicInlet(name, state); break;
      case MCFD_BC_SUBSONIC_OUTLET: sp->SetSubsonicOutlet(name, state); break;
      case MCFD_BC_SMART: sp->SetSmartBoundary(name, state); break;
      default: return 1;
    }
  }
  return 0;
}

/* demo/euler/sourceless.cpp:82-84 : cell.Approximate(ic) for all local cells. */
int ref_approximate(void *h, int func, const double *params, int n_params, double t) {
  auto *sim = static_cast<Sim *>(h);
  auto state = ref_functions::Make<Value, Global>(func, params, n_params);
  for (auto &p : sim->parts) {
    for (Cell *cell : p->GetLocalCellPointers()) {
      cell->Approximate([&](Global const &xyz) { return state(xyz, t); });
    }
  }
  return 0;
}

int ref_get_solution(void *h, int part, double *column) {
  auto *sim = static_cast<Sim *>(h);
  auto col = sim->spatials[part]->GetSolutionColumn();
  std::memcpy(column, col.data(), sizeof(double) * col.size());
  return 0;
}

/* Goes through Spatial::SetSolutionColumn, i.e. also runs the limiter when there is one
 * (dg/general.hpp:247-250); all parts must be set together because Reconstruct exchanges halos. */
int ref_set_solution(void *h, double const *const *columns) {
  auto *sim = static_cast<Sim *>(h);
  ForEachPart(sim, [&](int r) {
    auto n = sim->parts[r]->GetCellDataSize();
    Eigen::VectorXd col = Eigen::Map<const Eigen::VectorXd>(columns[r], n);
    sim->spatials[r]->SetSolutionColumn(col);
  });
  return 0;
}

int ref_get_residual(void *h, double t, double *const *columns) {
  auto *sim = static_cast<Sim *>(h);
  ForEachPart(sim, [&](int r) {
    sim->spatials[r]->SetTime(t);
    auto col = sim->spatials[r]->GetResidualColumn();
    std::memcpy(columns[r], col.data(), sizeof(double) * col.size());
  });
  return 0;
}

/* The time loop of demo/euler/sourceless.cpp:132-142; returns wall seconds. */
double ref_rk_update(void *h, double t, double dt, int n_steps) {
  auto *sim = static_cast<Sim *>(h);
  auto wall_0 = std::chrono::steady_clock::now();
  ForEachPart(sim, [&](int r) {
    auto *sys = sim->spatials[r].get();
    mini::temporal::RungeKutta<1, Scalar> rk1;
    mini::temporal::RungeKutta<2, Scalar> rk2;
    mini::temporal::RungeKutta<3, Scalar> rk3;
    for (int i = 0; i < n_steps; ++i) {
      double t_curr = t + dt * i;
      switch (sim->rk_order) {
        case 1: rk1.Update(sys, t_curr, dt); break;
        case 2: rk2.Update(sys, t_curr, dt); break;
        default: rk3.Update(sys, t_curr, dt); break;
      }
    }
  });
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - wall_0).count();
}

/* mesh::vtk::Writer<Part>::WriteSolutions (mesh/vtk.hpp:385-487) for every part, unmodified */
int ref_write_vtu(void *h, const char *directory, const char *soln_name) {
  auto *sim = static_cast<Sim *>(h);
  for (auto &part : sim->parts) {
    part->size_ = int(sim->parts.size());
    part->directory_ = directory;
    mini::mesh::vtk::Writer<ListPart>::WriteSolutions(*part, soln_name);
  }
  return 0;
}

/* What vtk::Writer::Prepare (mesh/vtk.hpp:341-377, 86-95) collects before it prints 6 digits of it: the reference's
 * own LocalToGlobalAndValue at the 27 nodes of vtk::Hexahedron27, in full precision. */
int ref_vtk_sample(void *h, int part, double *coords, double *values) {
  auto *sim = static_cast<Sim *>(h);
  auto &p = *sim->parts[part];
  long i = 0;
  for (const Cell &cell : p.GetLocalCells()) {
    for (int k = 0; k < 27; ++k, ++i) {
      Global global; Value value;
      cell.projection().LocalToGlobalAndValue(mini::mesh::vtk::Hexahedron27<Global>::locals[k], &global, &value);
      for (int d = 0; d < 3; ++d) coords[3 * i + d] = global[d];
      for (int c = 0; c < 5; ++c) values[5 * i + c] = value[c];
    }
  }
  return 0;
}

/* Section::GatherFields (mesh/part.hpp:370-378) over the local cells of one part: fields[i_field][i_cell] */
int ref_gather_fields(void *h, int part, double *fields) {
  auto *sim = static_cast<Sim *>(h);
  auto &p = *sim->parts[part];
  long n = p.n_local_;
  for (long i_cell = 0; i_cell < n; ++i_cell) {
    const auto &coeff = p.cells_[i_cell].projection().coeff();
    for (int i_field = 1; i_field <= Cell::kFields; ++i_field)
      fields[(i_field - 1) * n + i_cell] = coeff.reshaped()[i_field - 1];
  }
  return 0;
}

/* Part::MeasureL1Error (mesh/part.hpp:1082-1096), the same loop over the reference's own cell.gauss() /
 * cell.projection() objects (ListPart stands in for Part, so the member itself cannot be called). */
int ref_measure_l1_error(void *h, int part, int func, const double *params, int n_params, double t, double *l1) {
  auto *sim = static_cast<Sim *>(h);
  auto exact_solution = ref_functions::Make<Value, Global>(func, params, n_params);
  Value l1_error; l1_error.setZero();
  for (Cell const &cell : sim->parts[part]->GetLocalCells()) {
    auto &gauss = cell.gauss();
    auto &projection = cell.projection();
    for (int q = 0, n = gauss.CountPoints(); q < n; ++q) {
#if REF_SCHEME == 0
      Value value = projection.GlobalToValue(gauss.GetGlobalCoord(q));
#else
      Value value = projection.GetValue(q);
#endif
      value -= exact_solution(gauss.GetGlobalCoord(q), t);
      value = value.cwiseAbs() * gauss.GetGlobalWeight(q);
      l1_error += value;
    }
  }
  for (int k = 0; k < 5; ++k) l1[k] = l1_error[k];
  return 0;
}

int64_t ref_troubled_count(void *h) {
  auto *sim = static_cast<Sim *>(h);
  int64_t n = 0;
  for (auto &p : sim->parts) n += p->n_troubled_;
  return n;
}

/* Standalone pieces for unit parity: rotated Riemann flux and boundary fluxes
 * (riemann/rotated/euler.hpp:104-193). frames: [n][9] = normal, tangent, bitangent. */
int ref_riemann_flux(int64_t n, const double *frames, const double *u_left, const double *u_right, double *flux) {
  for (int64_t i = 0; i < n; ++i) {
    Riemann::Frame frame;
    for (int a = 0; a < 3; ++a) frame[a] = Global(frames[9 * i + 3 * a], frames[9 * i + 3 * a + 1], frames[9 * i + 3 * a + 2]);
    Riemann riemann; riemann.Rotate(frame);
    Conservative l, r;
    for (int k = 0; k < 5; ++k) { l[k] = u_left[5 * i + k]; r[k] = u_right[5 * i + k]; }
    auto f = riemann.GetFluxUpwind(l, r);
    for (int k = 0; k < 5; ++k) flux[5 * i + k] = f[k];
  }
  return 0;
}

int ref_boundary_flux(int kind, int64_t n, const double *frames, const double *u_inner, const double *u_given,
    double *flux) {
  for (int64_t i = 0; i < n; ++i) {
    Riemann::Frame frame;
    for (int a = 0; a < 3; ++a) frame[a] = Global(frames[9 * i + 3 * a], frames[9 * i + 3 * a + 1], frames[9 * i + 3 * a + 2]);
    Riemann riemann; riemann.Rotate(frame);
    Conservative ui, uo;
    for (int k = 0; k < 5; ++k) { ui[k] = u_inner[5 * i + k]; uo[k] = u_given[5 * i + k]; }
    Riemann::Flux f;
    switch (kind) {
      case MCFD_BC_SOLID_WALL: f = riemann.GetFluxOnSolidWall(ui); break;
      case MCFD_BC_SUPERSONIC_INLET: f = riemann.GetFluxOnSupersonicInlet(uo); break;
      case MCFD_BC_SUPERSONIC_OUTLET: f = riemann.GetFluxOnSupersonicOutlet(ui); break;
      case MCFD_BC_SUBSONIC_INLET: f = riemann.GetFluxOnSubsonicInlet(ui, uo); break;
      case MCFD_BC_SUBSONIC_OUTLET: f = riemann.GetFluxOnSubsonicOutlet(ui, uo); break;
      case MCFD_BC_SMART: f = riemann.GetFluxOnSmartBoundary(ui, uo); break;
      default: return 1;
    }
    for (int k = 0; k < 5; ++k) flux[5 * i + k] = f[k];
  }
  return 0;
}

/* the reference's approximate solvers called directly in the normal frame (they cannot be reached
 * through rotated::Euler in 3-D: GetFluxUpwind is non-const).  prim_*: [n][5] = rho, u, v, w, p.
 * riemann/euler/ausm.hpp:153-217 : Ausm<Gas,3>, all five components. */
int ref_ausm3_flux(int64_t n, const double *prim_left, const double *prim_right, double *flux) {
  using Solver = mini::riemann::euler::Ausm<Gas, 3>;
  Solver solver;
  for (int64_t i = 0; i < n; ++i) {
    const double *l = prim_left + 5 * i, *r = prim_right + 5 * i;
    Solver::Primitive left{ l[0], l[1], l[2], l[3], l[4] }, right{ r[0], r[1], r[2], r[3], r[4] };
    auto f = solver.GetFluxUpwind(left, right);
    for (int k = 0; k < 5; ++k) flux[5 * i + k] = f[k];
  }
  return 0;
}
/* riemann/euler/hllc.hpp:100-180 : Hllc<Gas,2> on (rho, u, v, p); w is ignored, flux[3] = 0. */
int ref_hllc2_flux(int64_t n, const double *prim_left, const double *prim_right, double *flux) {
  using Solver = mini::riemann::euler::Hllc<Gas, 2>;
  Solver solver;
  for (int64_t i = 0; i < n; ++i) {
    const double *l = prim_left + 5 * i, *r = prim_right + 5 * i;
    Solver::Primitive left{ l[0], l[1], l[2], l[4] }, right{ r[0], r[1], r[2], r[4] };
    auto f = solver.GetFluxUpwind(left, right);
    flux[5 * i + 0] = f[0]; flux[5 * i + 1] = f[1]; flux[5 * i + 2] = f[2]; flux[5 * i + 3] = 0.0; flux[5 * i + 4] = f[3];
  }
  return 0;
}

}  // extern "C"
