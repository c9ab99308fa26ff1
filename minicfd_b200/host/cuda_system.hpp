// cuda_system.hpp -- C++ host side above the C-ABI (include/minicfd_b200.h): the reference's operator
// interface for this path, implemented by the B200 library.  Header-only, C++17, no dependency on the
// reference tree or on Eigen.
//
// What it mirrors (paths relative to /root/reference/include/mini):
//   temporal/ode.hpp:15-46     System<Scalar>: SetTime / SetSolutionColumn / GetSolutionColumn / GetResidualColumn
//   spatial/fem.hpp:126-147    SetSolidWall / SetSupersonicOutlet / SetSupersonicInlet / SetSubsonicInlet /
//                              SetSubsonicOutlet / SetSmartBoundary (name [, std::function<Value(Global const&, double)>])
//   temporal/rk.hpp:13-73      RungeKutta<kOrders>::Update(System*, t_curr, dt)
//
// Two ways to use it in the reference's driver (demo/euler/sourceless.cpp:122-135):
//   (a) drop-in:  `mini_b200::CudaSystem<mini::temporal::System<double>> spatial(cfg, mesh, names);`
//       then the reference's own `temporal.Update(&spatial, t, dt)` drives the GPU through the four
//       virtuals, columns crossing PCIe every call exactly where the reference memcpy's them
//       (fem.hpp:153-183).  Arbitrary host BC callbacks work: they are evaluated at the boundary
//       flux points at every SetTime, as the reference evaluates them per stage.
//   (b) fast path: `spatial.Update(t, dt, n_steps)` = RungeKutta<rk_order>::Update n_steps times, fully
//       device-resident (mcfd_rk_update).  Boundary states must then be device functions
//       (mcfd_bc_func) because no host code runs between the stages.
#pragma once
#include <array>
#include <cstdint>
#include <functional>
#include <map>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "minicfd_b200.h"

namespace mini_b200 {

// Stand-alone copy of the reference's abstract ODE system for builds without the reference tree.
template <class ColumnT>
struct SystemInterface {
  using Column = ColumnT;
  virtual ~SystemInterface() noexcept = default;
  virtual void SetTime(double t_curr) = 0;
  virtual void SetSolutionColumn(Column const &column) = 0;
  virtual Column GetSolutionColumn() const = 0;
  virtual Column GetResidualColumn() const = 0;
};

using Global = std::array<double, 3>;   // mini::spatial::FiniteElement::Global  (x, y, z)
using Value = std::array<double, MCFD_K>;  // ... ::Value  (rho, rho u, rho v, rho w, E)
using Function = std::function<Value(Global const &, double)>;  // fem.hpp:56-58

class Error : public std::runtime_error {
 public:
  Error(int status, std::string const &what) : std::runtime_error(what), status_(status) {}
  int status() const { return status_; }
 private:
  int status_;
};

// Base = mini::temporal::System<double> in the reference's build, SystemInterface<std::vector<double>>
// elsewhere.  Base::Column needs `Column(n)`, `.data()`, `.size()` (Eigen::VectorXd and std::vector do).
template <class Base = SystemInterface<std::vector<double>>>
class CudaSystem : public Base {
 public:
  using Column = typename Base::Column;

  // `names[g]` is the boundary-section name of mesh BC group g (the CGNS section names the
  // reference's setters take, e.g. "4_S_27" in demo/euler/shock_tube_sod.cpp:19-38).
  CudaSystem(mcfd_config const &cfg, mcfd_mesh const &mesh, std::vector<std::string> const &names = {})
      : cfg_(cfg) {
    Check(nullptr, mcfd_create(&cfg, &mesh, &solver_));
    for (int g = 0; g < mesh.n_bc_groups; ++g) {
      name_to_group_[g < int(names.size()) ? names[g] : std::to_string(g)] = g;
    }
  }
  CudaSystem(CudaSystem const &) = delete;
  CudaSystem &operator=(CudaSystem const &) = delete;
  ~CudaSystem() noexcept { mcfd_destroy(solver_); }  // (the reference's System has no virtual destructor)

  // ---- boundary conditions, fem.hpp:126-147 ----
  void SetSolidWall(std::string const &name) { SetKind(name, MCFD_BC_SOLID_WALL, nullptr); }
  void SetSupersonicOutlet(std::string const &name) { SetKind(name, MCFD_BC_SUPERSONIC_OUTLET, nullptr); }
  void SetSupersonicInlet(std::string const &name, Function f) { SetKind(name, MCFD_BC_SUPERSONIC_INLET, std::move(f)); }
  void SetSubsonicInlet(std::string const &name, Function f) { SetKind(name, MCFD_BC_SUBSONIC_INLET, std::move(f)); }
  void SetSubsonicOutlet(std::string const &name, Function f) { SetKind(name, MCFD_BC_SUBSONIC_OUTLET, std::move(f)); }
  void SetSmartBoundary(std::string const &name, Function f) { SetKind(name, MCFD_BC_SMART, std::move(f)); }
  // the same with a state function evaluated on the device (needed by the fast path)
  void SetDeviceBoundary(std::string const &name, mcfd_bc_kind kind, mcfd_bc_func func,
                         std::vector<double> const &params) {
    int g = Group(name);
    callbacks_.erase(g);
    Check(solver_, mcfd_set_boundary(solver_, g, kind, func, params.data(), int(params.size())));
  }

  // ---- temporal::System, fem.hpp:150-183 ----
  void SetTime(double t_curr) override {
    Check(solver_, mcfd_set_time(solver_, t_curr));
    // the reference calls the std::function per face point per stage (fr/lobatto.hpp:312,
    // dg/lobatto.hpp:239); here once per SetTime, then one upload per boundary name
    for (auto &[g, cb] : callbacks_) {
      auto &pts = points_[g];
      auto &vals = values_[g];
      for (size_t i = 0; i < pts.size() / 3; ++i) {
        Value v = cb(Global{ pts[3 * i], pts[3 * i + 1], pts[3 * i + 2] }, t_curr);
        for (int k = 0; k < MCFD_K; ++k) vals[MCFD_K * i + k] = v[k];
      }
      if (!pts.empty()) Check(solver_, mcfd_set_bc_values(solver_, g, vals.data(), int64_t(pts.size() / 3)));
    }
  }
  void SetSolutionColumn(Column const &column) override {
    Check(solver_, mcfd_set_solution(solver_, column.data(), int64_t(column.size())));
  }
  Column GetSolutionColumn() const override {
    Column column(mcfd_column_size(solver_));
    Check(solver_, mcfd_get_solution(solver_, column.data(), int64_t(column.size())));
    return column;
  }
  Column GetResidualColumn() const override {
    Column column(mcfd_column_size(solver_));
    Check(solver_, mcfd_get_residual(solver_, column.data(), int64_t(column.size())));
    return column;
  }

  // ---- fast path: temporal::RungeKutta<cfg.rk_order>::Update, n_steps times, on the device ----
  void Update(double t_curr, double dt, int n_steps = 1) {
    if (!callbacks_.empty()) {
      throw Error(MCFD_ERR_STATE, "device-resident Update needs device boundary functions (SetDeviceBoundary): "
                                  "host callbacks cannot run between the stages");
    }
    Check(solver_, mcfd_rk_update(solver_, t_curr, dt, n_steps));
  }
  // Cell::Approximate over all local cells with a device function (demo/euler/sourceless.cpp:82-84)
  void Approximate(mcfd_bc_func func, std::vector<double> const &params, double t = 0.0) {
    Check(solver_, mcfd_approximate(solver_, func, params.data(), int(params.size()), t));
  }

  // ---- between steps: what Main() does with the column when a frame is due (demo/euler/sourceless.cpp:104-150) ----
  // mesh::vtk::Writer<Part>::WriteSolutions(part, soln_name) (mesh/vtk.hpp:385-487)
  void WriteVtu(std::string const &directory, std::string const &soln_name) {
    Check(solver_, mcfd_write_vtu(solver_, directory.c_str(), soln_name.c_str(), nullptr));
  }
  // Part::GatherSolutions / ScatterSolutions (mesh/part.hpp:1101-1118): fields[i_field][i_cell], what cgp_field_write_data
  // is handed per field (part.hpp:1216-1240)
  std::vector<double> GatherFields() {
    std::vector<double> fields(static_cast<size_t>(size()));
    Check(solver_, mcfd_gather_fields(solver_, fields.data(), size()));
    return fields;
  }
  void ScatterFields(std::vector<double> const &fields) {
    Check(solver_, mcfd_scatter_fields(solver_, fields.data(), int64_t(fields.size())));
  }

  int64_t size() const { return mcfd_column_size(solver_); }
  mcfd_solver *handle() const { return solver_; }

 private:
  int Group(std::string const &name) const {
    auto it = name_to_group_.find(name);
    if (it == name_to_group_.end()) throw Error(MCFD_ERR_ARG, "unknown boundary name: " + name);
    return it->second;
  }
  void SetKind(std::string const &name, mcfd_bc_kind kind, Function f) {
    int g = Group(name);
    Check(solver_, mcfd_set_boundary(solver_, g, kind, MCFD_FUNC_HOST_VALUES, nullptr, 0));
    if (!f) { callbacks_.erase(g); return; }
    int64_t n = mcfd_bc_point_count(solver_, g);
    points_[g].assign(size_t(n) * 3, 0.0);
    values_[g].assign(size_t(n) * MCFD_K, 0.0);
    if (n > 0) Check(solver_, mcfd_bc_points(solver_, g, points_[g].data(), n));
    callbacks_[g] = std::move(f);
  }
  static void Check(mcfd_solver const *s, int status) {
    if (status != MCFD_OK) throw Error(status, mcfd_last_error(s));
  }

  mcfd_config cfg_;
  mcfd_solver *solver_ = nullptr;
  std::map<std::string, int> name_to_group_;
  std::map<int, Function> callbacks_;
  std::map<int, std::vector<double>> points_, values_;
};

// temporal::RungeKutta<kOrders> (rk.hpp:13-73) over the abstract interface: the slow, call-for-call
// equivalent of the reference's solver for builds without the reference tree.  Column needs
// `*=`, `+=`, `/=` like Eigen; std::vector<double> goes through the helpers below.
namespace detail {
template <class C> void Scale(C &c, double a) { for (auto &x : c) x *= a; }
template <class C> void Add(C &c, C const &d) { auto it = d.begin(); for (auto &x : c) x += *it++; }
template <class C> void Div(C &c, double a) { for (auto &x : c) x /= a; }
}  // namespace detail

template <int kOrders, class System>
void RungeKuttaUpdate(System *sys, double t_curr, double dt) {
  using Column = typename System::Column;
  auto next = [&](double t) {  // Euler::NextSolution, ode.hpp:119-126
    sys->SetTime(t);
    Column u = sys->GetSolutionColumn();
    Column r = sys->GetResidualColumn();
    detail::Scale(r, dt);
    detail::Add(u, r);
    return u;
  };
  static_assert(kOrders >= 1 && kOrders <= 3, "RungeKutta<1|2|3>");
  if (kOrders == 1) {
    sys->SetSolutionColumn(next(t_curr));
  } else if (kOrders == 2) {  // rk.hpp:36-43
    Column u_curr = sys->GetSolutionColumn();
    sys->SetSolutionColumn(next(t_curr));
    Column u_next = next(t_curr + dt);
    detail::Add(u_next, u_curr);
    detail::Scale(u_next, 0.5);
    sys->SetSolutionColumn(u_next);
  } else {                    // rk.hpp:55-72
    Column u_curr = sys->GetSolutionColumn();
    sys->SetSolutionColumn(next(t_curr));
    Column u_next = u_curr;
    detail::Scale(u_next, 3.0);
    detail::Add(u_next, next(t_curr + dt));
    detail::Div(u_next, 4.0);
    sys->SetSolutionColumn(u_next);
    u_next = next(t_curr + dt / 2);
    detail::Scale(u_next, 2.0);
    detail::Add(u_next, u_curr);
    detail::Div(u_next, 3.0);
    sys->SetSolutionColumn(u_next);
  }
}

}  // namespace mini_b200
