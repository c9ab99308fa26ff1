/* oracle/ref/ref_functions.hpp -- TEST INFRASTRUCTURE ONLY.
 * The "given state" callbacks u(x, t) handed to the reference's Set*Boundary / Approximate
 * (std::function<Value(Global const&, double)>, include/mini/spatial/fem.hpp:56-58).
 * They are inputs of the comparison, not part of the path: the same closed forms are
 * evaluated on the device by the product (minicfd_b200/csrc/state_functions.cuh), selected
 * by the mcfd_bc_func ids of include/minicfd_b200.h.
 */
#ifndef ORACLE_REF_FUNCTIONS_HPP_
#define ORACLE_REF_FUNCTIONS_HPP_

#include <array>
#include <cmath>
#include <functional>
#include <stdexcept>

namespace ref_functions {

template <class Value, class Global>
std::function<Value(Global const &, double)> Make(int func, const double *params, int n_params) {
  std::array<double, 16> p{};
  for (int i = 0; i < n_params && i < 16; ++i) p[i] = params[i];
  switch (func) {
    case 1:  // MCFD_FUNC_CONSTANT: conservative state
      return [p](Global const &, double) {
        Value v; for (int k = 0; k < 5; ++k) v[k] = p[k];
        return v;
      };
    case 2:  // MCFD_FUNC_VORTEX: {beta, xc, yc, u0, v0, gamma}
      return [p](Global const &xyz, double t) {
        const double kPi = 3.14159265358979323846;
        double beta = p[0], gamma = p[5];
        double x = xyz[0] - p[1] - p[3] * t, y = xyz[1] - p[2] - p[4] * t;
        double r2 = x * x + y * y;
        double e = std::exp(0.5 * (1.0 - r2));
        double c = beta / (2.0 * kPi) * e;
        double u = p[3] - c * y, v = p[4] + c * x;
        double temp = 1.0 - (gamma - 1.0) * beta * beta / (8.0 * gamma * kPi * kPi) * e * e;
        double rho = std::pow(temp, 1.0 / (gamma - 1.0));
        double pres = rho * temp;
        Value val;
        val[0] = rho; val[1] = rho * u; val[2] = rho * v; val[3] = 0.0;
        val[4] = pres / (gamma - 1.0) + 0.5 * rho * (u * u + v * v);
        return val;
      };
    case 3:  // MCFD_FUNC_PLANE_SPLIT: {nx, ny, nz, offset, speed, left[5], right[5]}
      return [p](Global const &xyz, double t) {
        double s = p[0] * xyz[0] + p[1] * xyz[1] + p[2] * xyz[2];
        const double *src = (s < p[3] + p[4] * t) ? &p[5] : &p[10];
        Value v; for (int k = 0; k < 5; ++k) v[k] = src[k];
        return v;
      };
    default:
      return [](Global const &, double) { Value v; v.setZero(); return v; };
  }
}

}  // namespace ref_functions
#endif  // ORACLE_REF_FUNCTIONS_HPP_
