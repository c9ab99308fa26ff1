// modal_device.cuh -- device-side pieces of the MODAL discontinuous Galerkin path and of the WENO limiters:
// Taylor monomials about the cell centre, their gradients and all partial derivatives, the orthonormal
// basis phi = S T (S lower triangular, one per cell), the smoothness indicator and the eigenvector matrices.
//
// Replaces (paths relative to /root/reference/include/mini):
//   basis/taylor.hpp:163-507        Taylor<Scalar, 3, 1|2|3>::GetValue / GetGradValue / GetPdvValue / GetSmoothness
//   basis/linear.hpp:87-134         OrthoNormal::operator(), GetGradValue, Measure
//   polynomial/projection.hpp:41-47 GetAverage
//   riemann/euler/eigen.hpp:27-56   EigenMatrices
#pragma once
#include "riemann.cuh"

namespace mcfd {

// Shape code PE = degree + 10 * mcfd_cell_type: 1, 2, 3 = hexahedra (what every ModalShape<P> with the plain degree
// means), 11, 12, 13 = tetrahedra.  The basis (N, Taylor monomials, S) depends on the degree only; the element type
// sets the quadrature sizes and the number of faces (part.hpp:437-452).
template <int PE> struct ModalShape {
  static constexpr int P = PE % 10, E = PE / 10;
  static constexpr int N = P == 1 ? 4 : P == 2 ? 10 : 20;  // Taylor<3, P>::N
  static constexpr int Q = P + 1;                          // Legendre points per line (projection.hpp:84)
  static constexpr int NQ = E == 0 ? Q * Q * Q : (P == 1 ? 4 : P == 2 ? 14 : 24);  // gauss::Tetrahedron<4|14|24>
  static constexpr int FQ = E == 0 ? Q * Q : (P == 1 ? 3 : P == 2 ? 6 : 12);       // gauss::Triangle<3|6|12>
  static constexpr int NF = E == 0 ? 6 : 4;                                         // faces per cell
};

// basis/taylor.hpp:171-175, 213-217, 336-343
template <int P>
__device__ __forceinline__ void taylor_value(double x, double y, double z, double (&T)[ModalShape<P>::N]) {
  T[0] = 1; T[1] = x; T[2] = y; T[3] = z;
  if (P >= 2) { T[4] = x * x; T[5] = x * y; T[6] = x * z; T[7] = y * y; T[8] = y * z; T[9] = z * z; }
  if (P >= 3) {
    const double xx = x * x, xy = x * y, xz = x * z, yy = y * y, yz = y * z, zz = z * z;
    T[10] = x * xx; T[11] = x * xy; T[12] = x * xz; T[13] = x * yy; T[14] = x * yz; T[15] = x * zz;
    T[16] = y * yy; T[17] = y * yz; T[18] = y * zz; T[19] = z * zz;
  }
}

// gradient of monomial m (compile-time m after unrolling); Taylor::GetGradValue is coeff * this table
template <int P>
__device__ __forceinline__ void taylor_grad(double x, double y, double z, double (&G)[ModalShape<P>::N][3]) {
#pragma unroll
  for (int m = 0; m < ModalShape<P>::N; ++m) { G[m][0] = 0; G[m][1] = 0; G[m][2] = 0; }
  G[1][0] = 1; G[2][1] = 1; G[3][2] = 1;
  if (P >= 2) {
    G[4][0] = 2 * x; G[5][0] = y; G[5][1] = x; G[6][0] = z; G[6][2] = x;
    G[7][1] = 2 * y; G[8][1] = z; G[8][2] = y; G[9][2] = 2 * z;
  }
  if (P >= 3) {
    const double xx = x * x, xy = x * y, xz = x * z, yy = y * y, yz = y * z, zz = z * z;
    G[10][0] = 3 * xx;
    G[11][0] = 2 * xy; G[11][1] = xx;
    G[12][0] = 2 * xz; G[12][2] = xx;
    G[13][0] = yy; G[13][1] = 2 * xy;
    G[14][0] = yz; G[14][1] = xz; G[14][2] = xy;
    G[15][0] = zz; G[15][2] = 2 * xz;
    G[16][1] = 3 * yy;
    G[17][1] = 2 * yz; G[17][2] = yy;
    G[18][1] = zz; G[18][2] = 2 * yz;
    G[19][2] = 3 * zz;
  }
}

// phi_n(x) = sum_{m <= n} S[n][m] T_m(x - c)   (basis/linear.hpp:116-121); S row-major [N][N] in any memory space
template <int P>
__device__ __forceinline__ void modal_basis(const double *__restrict__ S, const double *__restrict__ center, double x,
                                            double y, double z, double (&phi)[ModalShape<P>::N]) {
  constexpr int N = ModalShape<P>::N;
  double T[N];
  taylor_value<P>(x - center[0], y - center[1], z - center[2], T);
#pragma unroll
  for (int n = 0; n < N; ++n) {
    double s = 0;
#pragma unroll
    for (int m = 0; m <= n; ++m) s += S[n * N + m] * T[m];
    phi[n] = s;
  }
}

// u(x) = coeff * phi(x)   (polynomial/projection.hpp:104-106); coeff is the column layout [N][5]
template <int P>
__device__ __forceinline__ Cons modal_value(const double *__restrict__ coeff, const double (&phi)[ModalShape<P>::N]) {
  Cons u;
#pragma unroll
  for (int k = 0; k < 5; ++k) u.c[k] = 0;
#pragma unroll
  for (int n = 0; n < ModalShape<P>::N; ++n)
#pragma unroll
    for (int k = 0; k < 5; ++k) u.c[k] += coeff[n * 5 + k] * phi[n];
  return u;
}

// basis/taylor.hpp:176-182 (P=1), 218-260 (P=2), 322-429 (P=3): every partial derivative of sum_m c[m] T_m at
// (x, y, z), one component at a time.  P=3 keeps `res.col(YY) += res.col(XX)` of taylor.hpp:383 as it is.
template <int P>
__device__ __forceinline__ void taylor_pdv(double x, double y, double z, const double (&c)[ModalShape<P>::N],
                                           double (&r)[ModalShape<P>::N]) {
  enum { X = 1, Y, Z, XX, XY, XZ, YY, YZ, ZZ, XXX, XXY, XXZ, XYY, XYZ, XZZ, YYY, YYZ, YZZ, ZZZ };
#pragma unroll
  for (int m = 0; m < ModalShape<P>::N; ++m) r[m] = m == 0 ? 0.0 : c[m];
  if (P == 2) {
    r[X] += c[XX] * (2 * x); r[X] += c[XY] * y; r[X] += c[XZ] * z;
    r[Y] += c[XY] * x; r[Y] += c[YY] * (2 * y); r[Y] += c[YZ] * z;
    r[Z] += c[XZ] * x; r[Z] += c[YZ] * y; r[Z] += c[ZZ] * (2 * z);
    r[XX] += c[XX]; r[YY] += c[YY]; r[ZZ] += c[ZZ];
  } else if (P == 3) {
    const double xx = x * x, xy = x * y, xz = x * z, yy = y * y, yz = y * z, zz = z * z;
    r[X] += c[XX] * (2 * x); r[X] += c[XY] * y; r[X] += c[XZ] * z;
    r[X] += c[XXX] * (3 * xx); r[X] += c[XXY] * (2 * xy); r[X] += c[XXZ] * (2 * xz);
    r[X] += c[XYY] * yy; r[X] += c[XYZ] * yz; r[X] += c[XZZ] * zz;
    r[Y] += c[XY] * x; r[Y] += c[YY] * (2 * y); r[Y] += c[YZ] * z;
    r[Y] += c[XXY] * xx; r[Y] += c[XYZ] * xz; r[Y] += c[XYY] * (2 * xy);
    r[Y] += c[YYY] * (3 * yy); r[Y] += c[YYZ] * (2 * yz); r[Y] += c[YZZ] * zz;
    r[Z] += c[XZ] * x; r[Z] += c[YZ] * y; r[Z] += c[ZZ] * (2 * z);
    r[Z] += c[XXZ] * xx; r[Z] += c[XYZ] * xy; r[Z] += c[XZZ] * (2 * xz);
    r[Z] += c[YYZ] * yy; r[Z] += c[YZZ] * (2 * yz); r[Z] += c[ZZZ] * (3 * zz);
    r[XX] += c[XXY] * y; r[XX] += c[XXZ] * z; r[XX] += c[XXX] * x * 3; r[XX] += r[XX];
    r[XY] += c[XXY] * x * 2; r[XY] += c[XYY] * y * 2; r[XY] += c[XYZ] * z;
    r[XZ] += c[XXZ] * x * 2; r[XZ] += c[XZZ] * z * 2; r[XZ] += c[XYZ] * y;
    r[YY] += c[XYY] * x; r[YY] += c[YYZ] * z; r[YY] += c[YYY] * y * 3; r[YY] += r[XX];  // sic
    r[YZ] += c[XYZ] * x; r[YZ] += c[YYZ] * y * 2; r[YZ] += c[YZZ] * z * 2;
    r[ZZ] += c[XZZ] * x; r[ZZ] += c[YZZ] * y; r[ZZ] += c[ZZZ] * z * 3; r[ZZ] += r[ZZ];
    r[XXX] *= 6; r[XXY] += r[XXY]; r[XXZ] += r[XXZ]; r[XYY] += r[XYY]; r[XZZ] += r[XZZ];
    r[YYY] *= 6; r[YYZ] += r[YYZ]; r[YZZ] += r[YZZ]; r[ZZZ] *= 6;
  }
}

// The integrand of weno::GetSmoothness (limiter/weno.hpp:20-34) at one quadrature point, already combined with
// Taylor::GetSmoothness' derivative-order weights (taylor.hpp:284-302, 478-507; linear in the integral):
//   sum_m wgt(order of m) * pdv_m(x)^2   for one component with Taylor coefficients c
template <int P>
__device__ __forceinline__ double smoothness_integrand(double x, double y, double z, const double (&c)[ModalShape<P>::N],
                                                        double w1, double w2, double w3) {
  constexpr int N = ModalShape<P>::N;
  double r[N];
  taylor_pdv<P>(x, y, z, c, r);
  double s1 = r[1] * r[1] + r[2] * r[2] + r[3] * r[3], s2 = 0, s3 = 0;
  if (P >= 2) {
#pragma unroll
    for (int m = 4; m < 10; ++m) s2 += r[m] * r[m];
  }
  if (P >= 3) {
#pragma unroll
    for (int m = 10; m < N; ++m) s3 += r[m] * r[m];
  }
  return P == 1 ? s1 : s1 * w1 + s2 * w2 + s3 * w3;
}

// riemann/euler/eigen.hpp:27-56 : L and R (row-major 5 x 5) at the conservative state big_u, frame (nu, mu, pi)
__device__ inline void eigen_matrices(const double *big_u, const Frame &fr, double gamma, double *L, double *R) {
  Cons c{ { big_u[0], big_u[1], big_u[2], big_u[3], big_u[4] } };
  const Prim s = cons_to_prim(c, gamma);
  const double vel[3] = { s.u, s.v, s.w };
  const double u_nu = s.u * fr.n[0] + s.v * fr.n[1] + s.w * fr.n[2];
  const double u_mu = s.u * fr.t[0] + s.v * fr.t[1] + s.w * fr.t[2];
  const double u_pi = s.u * fr.b[0] + s.v * fr.b[1] + s.w * fr.b[2];
  const double a = speed_of_sound(s, gamma);
  const double ek = (s.u * s.u + s.v * s.v + s.w * s.w) / 2, e0 = a * a * (1 / (gamma - 1)), h0 = e0 + ek;
  R[0] = 1; R[1] = 1; R[2] = 0; R[3] = 0; R[4] = 1;
  for (int d = 0; d < 3; ++d) {
    double *row = R + (1 + d) * 5;
    row[0] = vel[d] - a * fr.n[d]; row[1] = vel[d]; row[2] = fr.t[d]; row[3] = fr.b[d]; row[4] = vel[d] + a * fr.n[d];
  }
  R[20] = h0 - a * u_nu; R[21] = ek; R[22] = u_mu; R[23] = u_pi; R[24] = h0 + a * u_nu;
  const double b1 = 1 / e0, b2 = b1 * ek;
  L[0] = (b2 + u_nu / a) / 2; L[5] = 1 - b2; L[10] = -u_mu; L[15] = -u_pi; L[20] = (b2 - u_nu / a) / 2;
  for (int d = 0; d < 3; ++d) {
    L[0 + 1 + d] = -(b1 * vel[d] + fr.n[d] / a) / 2; L[5 + 1 + d] = b1 * vel[d]; L[10 + 1 + d] = fr.t[d];
    L[15 + 1 + d] = fr.b[d]; L[20 + 1 + d] = -(b1 * vel[d] - fr.n[d] / a) / 2;
  }
  L[4] = b1 / 2; L[9] = -b1; L[14] = 0; L[19] = 0; L[24] = b1 / 2;
}

}  // namespace mcfd
