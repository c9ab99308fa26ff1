/* oracle/ref/dropin_test.cpp -- TEST INFRASTRUCTURE ONLY.
 *
 * The drop-in claim, executed: the reference's OWN time integrator
 *   mini::temporal::RungeKutta<3, double>::Update(System*, t, dt)     (include/mini/temporal/rk.hpp:46-73)
 * drives (1) the reference's own spatial operator (spatial::fr::Lobatto / dg::Lobatto on the harness
 * Part of harness.cpp) and (2) mini_b200::CudaSystem<mini::temporal::System<double>>
 * (minicfd_b200/host/cuda_system.hpp over the C-ABI) through the same four virtuals, with the same
 * std::function boundary callbacks, on the same mesh; the solution columns are then compared.
 * A third run uses the device-resident fast path (CudaSystem::Update = mcfd_rk_update).
 *
 * Built by oracle/Makefile into oracle/_ref/ (needs /root/reference at build time); run on the GPU box
 * by tests/test_dropin.py.  usage: dropin_test nx ny nz distortion steps dt
 */
#include <cstdio>
#include <cstdlib>

#include "harness.cpp"
#include "../../minicfd_b200/host/cuda_system.hpp"

namespace {

double RelErr(const Eigen::VectorXd &a, const Eigen::VectorXd &b) {
  return (a - b).cwiseAbs().maxCoeff() / b.cwiseAbs().maxCoeff();
}

}  // namespace

int main(int argc, char **argv) {
  int64_t n[3] = { argc > 1 ? atol(argv[1]) : 5, argc > 2 ? atol(argv[2]) : 4, argc > 3 ? atol(argv[3]) : 3 };
  double distortion = argc > 4 ? atof(argv[4]) : 0.1;
  int steps = argc > 5 ? atoi(argv[5]) : 5;
  double dt = argc > 6 ? atof(argv[6]) : 5e-3;
  double lo[3] = { 0, 0, 0 }, hi[3] = { 10, 10, 10 };
  int32_t parts[3] = { 1, 1, 1 };
  mcfd_box_mesh *bm = nullptr;
  if (mcfd_box_mesh_create(n, lo, hi, parts, 0, distortion, &bm) != 0) { std::printf("mesh failed\n"); return 2; }
  const mcfd_mesh *mesh = mcfd_box_mesh_view(bm);
  const double vortex[6] = { 5.0, 5.0, 5.0, 1.0, 1.0, 1.4 };

  // (1) the reference
  auto *sim = static_cast<Sim *>(ref_create(1, mesh, 3, 1e-3, 1e-6));
  for (int g = 0; g < 4; ++g) ref_set_boundary(sim, g, MCFD_BC_SMART, MCFD_FUNC_VORTEX, vortex, 6);
  for (int g = 4; g < 6; ++g) ref_set_boundary(sim, g, MCFD_BC_SOLID_WALL, 0, nullptr, 0);
  ref_approximate(sim, MCFD_FUNC_VORTEX, vortex, 6, 0.0);
  Spatial *ref_sys = sim->spatials[0].get();
  Eigen::VectorXd u0 = ref_sys->GetSolutionColumn();

  // (2) the GPU behind the same interface, with std::function callbacks like the reference's
  mcfd_config cfg{};
  cfg.scheme = REF_SCHEME; cfg.degree = REF_DEGREE; cfg.riemann = MCFD_RIEMANN_EXACT; cfg.limiter = 0;
  cfg.rk_order = 3; cfg.device = 0; cfg.rank = 0; cfg.world = 1; cfg.gamma = 1.4;
  using GpuSystem = mini_b200::CudaSystem<mini::temporal::System<double>>;
  auto state = ref_functions::Make<Value, Global>(MCFD_FUNC_VORTEX, vortex, 6);
  mini_b200::Function callback = [state](mini_b200::Global const &x, double t) {
    Value v = state(Global(x[0], x[1], x[2]), t);
    return mini_b200::Value{ v[0], v[1], v[2], v[3], v[4] };
  };
  try {
    GpuSystem gpu(cfg, *mesh), fast(cfg, *mesh);
    for (int g = 0; g < 4; ++g) {
      gpu.SetSmartBoundary(std::to_string(g), callback);
      fast.SetDeviceBoundary(std::to_string(g), MCFD_BC_SMART, MCFD_FUNC_VORTEX, std::vector<double>(vortex, vortex + 6));
    }
    for (int g = 4; g < 6; ++g) { gpu.SetSolidWall(std::to_string(g)); fast.SetSolidWall(std::to_string(g)); }
    gpu.SetSolutionColumn(u0);
    fast.SetSolutionColumn(u0);

    mini::temporal::RungeKutta<3, double> rk;  // the reference's solver, unmodified
    for (int i = 0; i < steps; ++i) {
      rk.Update(ref_sys, dt * i, dt);
      rk.Update(&gpu, dt * i, dt);
    }
    fast.Update(0.0, dt, steps);
    Eigen::VectorXd u_ref = ref_sys->GetSolutionColumn();
    double e_dropin = RelErr(gpu.GetSolutionColumn(), u_ref), e_fast = RelErr(fast.GetSolutionColumn(), u_ref);
    std::printf("{\"scheme\": %d, \"degree\": %d, \"dof\": %ld, \"steps\": %d, \"dropin_rel_err\": %.3e, \"fast_rel_err\": %.3e}\n",
                REF_SCHEME, REF_DEGREE, long(u_ref.size()), steps, e_dropin, e_fast);
    ref_destroy(sim);
    mcfd_box_mesh_destroy(bm);
    return (e_dropin < 1e-10 && e_fast < 1e-10) ? 0 : 1;
  } catch (mini_b200::Error const &e) {
    std::printf("mini_b200::Error %d: %s\n", e.status(), e.what());
    return 3;
  }
}
