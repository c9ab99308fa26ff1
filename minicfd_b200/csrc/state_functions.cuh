// state_functions.cuh -- built-in "given state" functions u(x, t) evaluated on the device.
// They take the place of the host std::function<Value(Global const&, double)> callbacks the
// reference stores per boundary name and calls per face point per stage
// (spatial/fem.hpp:56-58, fr/lobatto.hpp:312, dg/lobatto.hpp:239) and of the IC passed to
// Cell::Approximate (demo/euler/sourceless.cpp:82-84).  ids = mcfd_bc_func.
#pragma once
#include "minicfd_b200.h"
#include "riemann.cuh"

namespace mcfd {

struct FuncParams { double p[MCFD_MAX_FUNC_PARAMS]; };

__device__ __forceinline__ Cons eval_state(int func, const double *p, double x, double y, double z, double t) {
  Cons val;
  if (func == MCFD_FUNC_CONSTANT) {
    for (int k = 0; k < 5; ++k) val.c[k] = p[k];
  } else if (func == MCFD_FUNC_VORTEX) {
    // isentropic vortex advected with (u0, v0): {beta, xc, yc, u0, v0, gamma}
    const double kPi = 3.14159265358979323846;
    double beta = p[0], gamma = p[5];
    double xr = x - p[1] - p[3] * t, yr = y - p[2] - p[4] * t;
    double r2 = xr * xr + yr * yr;
    double e = exp(0.5 * (1.0 - r2));
    double c = beta / (2.0 * kPi) * e;
    double u = p[3] - c * yr, v = p[4] + c * xr;
    double temp = 1.0 - (gamma - 1.0) * beta * beta / (8.0 * gamma * kPi * kPi) * e * e;
    double rho = pow(temp, 1.0 / (gamma - 1.0));
    double pres = rho * temp;
    val.c[0] = rho; val.c[1] = rho * u; val.c[2] = rho * v; val.c[3] = 0.0;
    val.c[4] = pres / (gamma - 1.0) + 0.5 * rho * (u * u + v * v);
  } else if (func == MCFD_FUNC_PLANE_SPLIT) {
    // {nx, ny, nz, offset, speed, left[5], right[5]} (shock_tube_sod.cpp:10-13, double_mach.cpp:14-31)
    double s = p[0] * x + p[1] * y + p[2] * z;
    const double *src = (s < p[3] + p[4] * t) ? p + 5 : p + 10;
    for (int k = 0; k < 5; ++k) val.c[k] = src[k];
  } else {
    for (int k = 0; k < 5; ++k) val.c[k] = 0.0;
  }
  return val;
}

}  // namespace mcfd
