// setup.cpp -- host-side construction of the per-part tables (run once per mcfd_create).
//
// Replaces, for HEXA_8 cells with Lobatto nodes (paths relative to /root/reference/include/mini):
//   polynomial/hexahedron.hpp:166-214   per-node det J and J* = det(J) J^{-1}
//   gauss/quadrangle.hpp + geometry/face.hpp:98-110   face points and normal frames (n, t, b)
//   spatial/fr/lobatto.hpp:76-162, spatial/dg/lobatto.hpp:55-75   face point <-> solution node maps
//   mesh/part.hpp:990-1078              holder/sharer bookkeeping, face node sorting
// The reference matches face points to solution nodes by comparing coordinates; here the match is
// purely combinatorial (node ids of the face corners), so it cannot fail on distorted meshes.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <unordered_map>

#include "solver.hpp"

namespace mcfd {

namespace {

// geometry/hexahedron.hpp:345-352 : local coordinates of the HEXA_8 corners
const double kCorner[8][3] = { {-1,-1,-1},{1,-1,-1},{1,1,-1},{-1,1,-1},{-1,-1,1},{1,-1,1},{1,1,1},{-1,1,1} };
// geometry/hexahedron.hpp:371-380 : corners of the six faces, outward orientation
const int kFaceTable[6][4] = { {0,3,2,1}, {0,1,5,4}, {1,2,6,5}, {2,3,7,6}, {0,4,7,3}, {4,5,6,7} };
// which local direction each face is normal to, and on which end it lies (fr/lobatto.hpp:96-156)
const int kSideDir[6] = { 2, 1, 0, 1, 0, 2 };
const int kSideHigh[6] = { 0, 0, 1, 1, 0, 1 };
// the two free local directions of a face in increasing node-index order (i slowest, k fastest)
const int kSideFree[6][2] = { {0,1}, {0,2}, {1,2}, {0,2}, {1,2}, {0,1} };

double Legendre(int n, double x) {
  if (n == 0) return 1;
  if (n == 1) return x;
  double p0 = 1, p1 = x;
  for (int k = 2; k <= n; ++k) { double p2 = ((2 * k - 1) * x * p1 - (k - 1) * p0) / k; p0 = p1; p1 = p2; }
  return p1;
}

// corner index (0..7) of the lattice corner (a01, b01) of face `side`
int CornerOfSide(int side, int a01, int b01) {
  int ijk[3];
  ijk[kSideDir[side]] = kSideHigh[side];
  ijk[kSideFree[side][0]] = a01;
  ijk[kSideFree[side][1]] = b01;
  static const int vmap[2][2] = { {0, 3}, {1, 2} };
  return vmap[ijk[0]][ijk[1]] + 4 * ijk[2];
}

void ApplyPerm(int code, int q, int a, int b, int *a2, int *b2) {
  int x = (code & 2) ? q - 1 - a : a, y = (code & 4) ? q - 1 - b : b;
  if (code & 1) std::swap(x, y);
  *a2 = x; *b2 = y;
}

// node = i*q*q + j*q + k (basis/lagrange.hpp:236-238) of lattice point (a, b) on face `side`
int SideNode(int q, int side, int a, int b) {
  int ijk[3];
  ijk[kSideDir[side]] = kSideHigh[side] ? q - 1 : 0;
  ijk[kSideFree[side][0]] = a;
  ijk[kSideFree[side][1]] = b;
  return (ijk[0] * q + ijk[1]) * q + ijk[2];
}

int FindSide(const int64_t *cell_nodes, const int64_t *face_nodes) {
  for (int s = 0; s < 6; ++s) {
    int cnt = 0;
    for (int v = 0; v < 4; ++v)
      for (int w = 0; w < 4; ++w)
        if (cell_nodes[kFaceTable[s][v]] == face_nodes[w]) ++cnt;
    if (cnt == 4) return s;
  }
  return -1;
}

// code c with: lattice corner (a,b) of (cell_from, side_from) == lattice corner Apply(c,a,b) of (cell_to, side_to)
int FindPerm(const int64_t *from_nodes, int side_from, const int64_t *to_nodes, int side_to) {
  for (int c = 0; c < 8; ++c) {
    bool ok = true;
    for (int a = 0; a < 2 && ok; ++a)
      for (int b = 0; b < 2 && ok; ++b) {
        int a2, b2;
        ApplyPerm(c, 2, a, b, &a2, &b2);
        if (from_nodes[CornerOfSide(side_from, a, b)] != to_nodes[CornerOfSide(side_to, a2, b2)]) ok = false;
      }
    if (ok) return c;
  }
  return -1;
}

// geometry/hexahedron.hpp:419-455 + geometry/cell.hpp LocalToJacobian : J[a][b] = d x_b / d xi_a
void HexaJacobian(const double xn[8][3], const double loc[3], double J[3][3]) {
  std::memset(J, 0, 9 * sizeof(double));
  for (int i = 0; i < 8; ++i) {
    double fx = 1 + kCorner[i][0] * loc[0], fy = 1 + kCorner[i][1] * loc[1], fz = 1 + kCorner[i][2] * loc[2];
    double g[3] = { kCorner[i][0] * (fy * fz / 8), kCorner[i][1] * (fz * fx / 8), kCorner[i][2] * (fx * fy / 8) };
    for (int a = 0; a < 3; ++a)
      for (int b = 0; b < 3; ++b) J[a][b] += g[a] * xn[i][b];
  }
}

// m[0..8] = det(J) * J^{-1} laid out [global b][local a]; m[9] = det(J)
void Metrics(const double J[3][3], double *m) {
  double A[3][3];
  A[0][0] = J[1][1] * J[2][2] - J[1][2] * J[2][1];
  A[0][1] = J[0][2] * J[2][1] - J[0][1] * J[2][2];
  A[0][2] = J[0][1] * J[1][2] - J[0][2] * J[1][1];
  A[1][0] = J[1][2] * J[2][0] - J[1][0] * J[2][2];
  A[1][1] = J[0][0] * J[2][2] - J[0][2] * J[2][0];
  A[1][2] = J[0][2] * J[1][0] - J[0][0] * J[1][2];
  A[2][0] = J[1][0] * J[2][1] - J[1][1] * J[2][0];
  A[2][1] = J[0][1] * J[2][0] - J[0][0] * J[2][1];
  A[2][2] = J[0][0] * J[1][1] - J[0][1] * J[1][0];
  std::memcpy(m, A, sizeof(A));
  m[9] = J[0][0] * A[0][0] + J[0][1] * A[1][0] + J[0][2] * A[2][0];
}

// ---- TETRA_4 cells and their triangles (modal DG only) ---------------------------------------------------------
// geometry/tetrahedron.hpp:139-150 : local coordinates of the corners (1,0,0), (0,1,0), (0,0,1), (0,0,0), i.e.
// x = a p0 + b p1 + c p2 + (1 - a - b - c) p3, and the faces in outward orientation
const int kTetFace[4][3] = { {0,2,1}, {0,1,3}, {0,3,2}, {1,2,3} };

// gauss/tetrahedron.hpp:112-310 : the symmetric rules Part picks by degree (part.hpp:448-452): 4, 14 or 24 points.
// Local coordinates (a, b, c) and local weights (they sum to 1/6), orbit by orbit in the reference's order.
int TetraRule(int degree, double (*abc)[3], double *w) {
  int q = 0;
  auto s31 = [&](double a, double weight) {
    const double c = 1 - 3 * a;
    const double pts[4][3] = { {a, a, a}, {a, a, c}, {a, c, a}, {c, a, a} };
    for (auto &p : pts) { std::copy_n(p, 3, abc[q]); w[q++] = weight / 6.0; }
  };
  if (degree == 1) {
    s31(0.13819660112501051517954131656343619, 0.25);
  } else if (degree == 2) {
    s31(0.31088591926330060979734573376345783, 0.11268792571801585079918565233328633);
    s31(0.09273525031089122640232391373703061, 0.07349304311636194954371020548632750);
    const double a = 0.04550370412564964949188052627933943, c = (1 - 2 * a) * .5;
    const double pts[6][3] = { {a, a, c}, {a, c, a}, {a, c, c}, {c, a, a}, {c, a, c}, {c, c, a} };
    for (auto &p : pts) { std::copy_n(p, 3, abc[q]); w[q++] = 0.04254602077708146643806942812025744 / 6.0; }
  } else {
    s31(0.21460287125915202928883921938628499, 0.03992275025816749209969062755747998);
    s31(0.04067395853461135311557944895641006, 0.01007721105532064294801323744593686);
    s31(0.32233789014227551034399447076249213, 0.05535718154365472209515327785372602);
    const double a = 0.06366100187501752529923552760572698, b = 0.60300566479164914136743113906093969, c = 1 - a - a - b;
    const double pts[12][3] = { {a, a, b}, {a, a, c}, {a, b, a}, {a, b, c}, {a, c, a}, {a, c, b},
                                {b, a, a}, {b, a, c}, {b, c, a}, {c, a, a}, {c, a, b}, {c, b, a} };
    for (auto &p : pts) { std::copy_n(p, 3, abc[q]); w[q++] = (27. / 560.) / 6.0; }
  }
  return q;
}

// gauss/triangle.hpp:122-246 : 3, 6 or 12 points (part.hpp:439-443); local (a, b) on x = a q0 + b q1 + (1 - a - b) q2
// (geometry/triangle.hpp:78-89), local weights summing to 1/2
int TriangleRule(int degree, double (*ab)[2], double *w) {
  int q = 0;
  auto s21 = [&](double a, double weight) {
    const double b = 1 - a - a;
    const double pts[3][2] = { {a, a}, {a, b}, {b, a} };
    for (auto &p : pts) { ab[q][0] = p[0]; ab[q][1] = p[1]; w[q++] = weight / 2.0; }
  };
  if (degree == 1) {
    s21(1. / 6., 1. / 3.);
  } else if (degree == 2) {
    s21(.44594849091596488631832925388305199, .22338158967801146569500700843312280);
    s21(.09157621350977074345957146340220151, .10995174365532186763832632490021053);
  } else {
    s21(.06308901449150222834033160287081916, .05084490637020681692093680910686898);
    s21(.24928674517091042129163855310701908, .11678627572637936602528961138557944);
    const double a = .05314504984481694735324967163139815, b = .31035245103378440541660773395655215, c = 1 - a - b;
    const double pts[6][2] = { {a, b}, {a, c}, {b, a}, {b, c}, {c, a}, {c, b} };
    for (auto &p : pts) { ab[q][0] = p[0]; ab[q][1] = p[1]; w[q++] = .08285107561837357519355345642044245 / 2.0; }
  }
  return q;
}

int FindTetSide(const int64_t *cell_nodes, const int64_t *face_nodes) {  // geometry/tetrahedron.hpp:76-95
  for (int s = 0; s < 4; ++s) {
    int cnt = 0;
    for (int v = 0; v < 3; ++v)
      for (int w = 0; w < 3; ++w)
        if (cell_nodes[kTetFace[s][v]] == face_nodes[w]) ++cnt;
    if (cnt == 3) return s;
  }
  return -1;
}

// gauss/function.hpp:109-141 : Gram-Schmidt on the Taylor monomials, A = int T T^t over the cell's quadrature
void OrthoNormalize(int N, const std::vector<double> &A, std::vector<double> &S) {
  std::fill(S.begin(), S.end(), 0.0);
  for (int i = 0; i < N; ++i) S[i * N + i] = 1;
  S[0] = 1 / std::sqrt(A[0]);
  for (int i = 1; i < N; ++i) {
    for (int j = 0; j < i; ++j) {
      double temp = 0;
      for (int k = 0; k <= j; ++k) temp += S[j * N + k] * A[k * N + i];
      for (int l = 0; l <= j; ++l) S[i * N + l] -= temp * S[j * N + l];
    }
    double norm_sq = 0;
    for (int j = 0; j <= i; ++j) {
      double sum = 0, Sij = S[i * N + j];
      for (int k = 0; k < j; ++k) sum += 2 * S[i * N + k] * A[k * N + j];
      norm_sq += Sij * (Sij * A[j * N + j] + sum);
    }
    double nrm = std::sqrt(norm_sq);
    for (int j = 0; j < N; ++j) S[i * N + j] /= nrm;
  }
}

void TaylorValue(int degree, const double *d, double *T);

// Cell + gauss::Tetrahedron<4|14|24> + basis::OrthoNormal (part.hpp:663-670, gauss/element.hpp:87-100)
int BuildModalTetra(const mcfd_mesh &m, int degree, HostTables *out, std::string *error) {
  HostTables &t = *out;
  const int nq = t.nq, N = t.n;
  const int64_t nc = t.n_local + t.n_ghost;
  t.m_center.resize(size_t(nc) * 3); t.m_S.assign(size_t(nc) * N * N, 0.0); t.m_volume.resize(nc);
  t.m_qxyzw.resize(size_t(nc) * nq * 4);
  double abc[24][3], wl[24];
  if (TetraRule(degree, abc, wl) != nq) { *error = "tetrahedral rule size"; return MCFD_ERR_ARG; }
  std::vector<double> A(N * N), S(N * N), T(N);
  for (int64_t c = 0; c < nc; ++c) {
    const double *p[4];
    for (int v = 0; v < 4; ++v) p[v] = m.node_xyz + 3 * m.cell_nodes[4 * c + v];
    auto map = [&](double a, double b, double cc, double *glob) {  // geometry/cell.hpp LocalToGlobal: p0 N0 first, then +=
      const double shape[4] = { a, b, cc, 1.0 - a - b - cc };
      for (int d = 0; d < 3; ++d) {
        double sum = p[0][d] * shape[0];
        for (int v = 1; v < 4; ++v) sum += p[v][d] * shape[v];
        glob[d] = sum;
      }
    };
    double J[3][3];  // rows d x / d(a, b, c) = p0 - p3, p1 - p3, p2 - p3
    for (int a = 0; a < 3; ++a) for (int d = 0; d < 3; ++d) J[a][d] = p[a][d] - p[3][d];
    const double det = J[0][0] * (J[1][1] * J[2][2] - J[1][2] * J[2][1]) - J[0][1] * (J[1][0] * J[2][2] - J[1][2] * J[2][0]) +
                       J[0][2] * (J[1][0] * J[2][1] - J[1][1] * J[2][0]);
    if (det == 0) { *error = "degenerate tetrahedron"; return MCFD_ERR_ARG; }
    double *center = &t.m_center[size_t(c) * 3];
    map(0.25, 0.25, 0.25, center);  // geometry/tetrahedron.hpp:43-46
    double volume = 0;
    std::fill(A.begin(), A.end(), 0.0);
    for (int q = 0; q < nq; ++q) {
      double *xw = &t.m_qxyzw[(size_t(c) * nq + q) * 4];
      map(abc[q][0], abc[q][1], abc[q][2], xw);
      xw[3] = wl[q] * std::fabs(det);
      volume += xw[3];
      const double d[3] = { xw[0] - center[0], xw[1] - center[1], xw[2] - center[2] };
      TaylorValue(degree, d, T.data());
      for (int a = 0; a < N; ++a) for (int b = 0; b < N; ++b) A[a * N + b] += T[a] * T[b] * xw[3];
    }
    t.m_volume[c] = volume;
    OrthoNormalize(N, A, S);
    std::copy(S.begin(), S.end(), t.m_S.begin() + size_t(c) * N * N);
  }
  return MCFD_OK;
}

// Cell + gauss::Hexahedron<Legendre> + basis::OrthoNormal for the modal scheme: quadrature points and global
// weights (gauss/element.hpp:87-100), centre (geometry/hexahedron.hpp:77-80), Gram-Schmidt on the Taylor
// monomials (gauss/function.hpp:109-141), volume.
void TaylorValue(int degree, const double *d, double *T) {  // basis/taylor.hpp:171-175, 213-217, 336-343
  double x = d[0], y = d[1], z = d[2];
  T[0] = 1; T[1] = x; T[2] = y; T[3] = z;
  if (degree >= 2) { T[4] = x * x; T[5] = x * y; T[6] = x * z; T[7] = y * y; T[8] = y * z; T[9] = z * z; }
  if (degree >= 3) {
    double xx = x * x, xy = x * y, xz = x * z, yy = y * y, yz = y * z, zz = z * z;
    T[10] = x * xx; T[11] = x * xy; T[12] = x * xz; T[13] = x * yy; T[14] = x * yz; T[15] = x * zz;
    T[16] = y * yy; T[17] = y * yz; T[18] = y * zz; T[19] = z * zz;
  }
}

void BuildModalCells(const mcfd_mesh &m, const Basis &bs, HostTables *out) {
  HostTables &t = *out;
  const int q = bs.q, nq = t.nq, N = t.n, degree = q - 1;
  const int64_t nc = t.n_local + t.n_ghost;
  t.m_center.resize(size_t(nc) * 3); t.m_S.assign(size_t(nc) * N * N, 0.0); t.m_volume.resize(nc);
  t.m_qxyzw.resize(size_t(nc) * nq * 4);
  std::vector<double> A(N * N), S(N * N), T(N);
  for (int64_t c = 0; c < nc; ++c) {
    double xn[8][3];
    for (int v = 0; v < 8; ++v)
      for (int b = 0; b < 3; ++b) xn[v][b] = m.node_xyz[3 * m.cell_nodes[8 * c + v] + b];
    auto map = [&](const double loc[3], double glob[3]) {
      glob[0] = glob[1] = glob[2] = 0;
      for (int i = 0; i < 8; ++i) {
        double shape = (1 + kCorner[i][0] * loc[0]) * (1 + kCorner[i][1] * loc[1]) * (1 + kCorner[i][2] * loc[2]) / 8;
        for (int b = 0; b < 3; ++b) glob[b] += xn[i][b] * shape;
      }
    };
    const double zero[3] = { 0, 0, 0 };
    double *center = &t.m_center[size_t(c) * 3];
    map(zero, center);
    double volume = 0;
    std::fill(A.begin(), A.end(), 0.0);
    for (int i = 0; i < q; ++i) for (int j = 0; j < q; ++j) for (int k = 0; k < q; ++k) {
      double loc[3] = { bs.x[i], bs.x[j], bs.x[k] }, J[3][3], mm[10];
      double *xw = &t.m_qxyzw[(size_t(c) * nq + (i * q + j) * q + k) * 4];
      map(loc, xw);
      HexaJacobian(xn, loc, J);
      Metrics(J, mm);
      xw[3] = bs.w[i] * bs.w[j] * bs.w[k] * std::fabs(mm[9]);
      volume += xw[3];
      double d[3] = { xw[0] - center[0], xw[1] - center[1], xw[2] - center[2] };
      TaylorValue(degree, d, T.data());
      for (int a = 0; a < N; ++a) for (int b = 0; b < N; ++b) A[a * N + b] += T[a] * T[b] * xw[3];
    }
    t.m_volume[c] = volume;
    std::fill(S.begin(), S.end(), 0.0);
    for (int i = 0; i < N; ++i) S[i * N + i] = 1;
    S[0] = 1 / std::sqrt(A[0]);
    for (int i = 1; i < N; ++i) {
      for (int j = 0; j < i; ++j) {
        double temp = 0;
        for (int k = 0; k <= j; ++k) temp += S[j * N + k] * A[k * N + i];
        for (int l = 0; l <= j; ++l) S[i * N + l] -= temp * S[j * N + l];
      }
      double norm_sq = 0;
      for (int j = 0; j <= i; ++j) {
        double sum = 0, Sij = S[i * N + j];
        for (int k = 0; k < j; ++k) sum += 2 * S[i * N + k] * A[k * N + j];
        norm_sq += Sij * (Sij * A[j * N + j] + sum);
      }
      double nrm = std::sqrt(norm_sq);
      for (int j = 0; j < N; ++j) S[i * N + j] /= nrm;
    }
    std::copy(S.begin(), S.end(), t.m_S.begin() + size_t(c) * N * N);
  }
}


// ---- shape classes of a modal part (HostTables::n_classes) ------------------------------------------------------
// Two cells are of one class when their quadrature points, taken relative to the cell centre, agree to 1e-12 of the
// cell size -- or to the rounding the coordinates themselves carry, 64 eps |x| / h, where that is larger (a fine mesh
// far from the origin) -- and so do their weights: S, phi_n(x_q) and grad phi_n(x_q) of the representative then differ
// from the cell's own by about that much, two orders inside the 1e-10 parity bar (phi and w grad phi move by O(delta)
// for a displacement delta: a 1e-9 match would not be).  Meshes that are not exactly repetitive (any distortion above
// the tolerance) simply produce too many classes and keep the general kernels.  The same for the points of a face
// relative to the centre of the cell on either side ("patterns").
void TaylorGradHost(int degree, const double *d, double (*G)[3]) {  // basis/taylor.hpp:183-186, 261-277, 431-476
  const double x = d[0], y = d[1], z = d[2];
  const int N = degree == 1 ? 4 : degree == 2 ? 10 : 20;
  for (int m = 0; m < N; ++m) G[m][0] = G[m][1] = G[m][2] = 0;
  G[1][0] = 1; G[2][1] = 1; G[3][2] = 1;
  if (degree >= 2) {
    G[4][0] = 2 * x; G[5][0] = y; G[5][1] = x; G[6][0] = z; G[6][2] = x;
    G[7][1] = 2 * y; G[8][1] = z; G[8][2] = y; G[9][2] = 2 * z;
  }
  if (degree >= 3) {
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

struct QuantKey {
  std::vector<int64_t> v;
  bool operator==(const QuantKey &o) const { return v == o.v; }
};
struct QuantKeyHash {
  size_t operator()(const QuantKey &k) const {
    uint64_t h = 0x9E3779B97F4A7C15ull;
    for (int64_t x : k.v) { h ^= uint64_t(x) + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2); h *= 0xBF58476D1CE4E5B9ull; }
    return size_t(h ^ (h >> 31));
  }
};

void BuildShapeClasses(int degree, HostTables *out) {
  HostTables &t = *out;
  constexpr int kMaxClasses = 256;
  const int nq = t.nq, fq = t.fq, N = t.n, nf = t.sides;
  t.n_classes = 0; t.n_patterns = 0; t.n_shapes = 0;
  if (t.per_point_frames && t.elem == MCFD_CELL_HEXA_8) return;  // curved faces: certainly not repetitive
  // modal_cell_mma_kernel packs (face point id << 1 | side bit) into an int and reads a negative value as "no flux on
  // this side": ids from 2^30 up would be dropped silently, so such a part keeps the general kernels
  if (t.n_faces * int64_t(fq) >= (int64_t(1) << 30)) return;
  // A cell's signature: its size, its quadrature points and the points of the face on every side, all relative to the
  // cell centre, and which sides it holds.  Two cells are of one class when every entry agrees to `tol` (see above).
  // Candidates come from a hash of the signature rounded to 1e-3 of the size (a rounding that falls differently for two
  // equal cells -- about once in 1e10 entries -- only makes a duplicate class); the comparison itself is tolerant.
  // ... and, for the limiter's tables, where the neighbour across every side sits and what shape it has (its quadrature
  // points relative to MY centre; zeros on a boundary side)
  const int n_vol = 1 + 3 * nq, n_side = 3 * fq + 1 + 1 + 3 * nq, L = n_vol + nf * n_side;
  std::vector<std::vector<double>> class_sig, shape_sig;
  std::vector<int64_t> class_rep;
  std::unordered_map<QuantKey, std::vector<int>, QuantKeyHash> class_bucket, shape_bucket;
  t.cell_class.assign(t.n_local, 0);
  auto give_up = [&]() { t.n_classes = 0; t.n_shapes = 0; t.n_patterns = 0; t.cell_class.clear(); };
  std::vector<double> sig(L);
  QuantKey key;
  auto match = [&](std::unordered_map<QuantKey, std::vector<int>, QuantKeyHash> &buckets, std::vector<std::vector<double>> &sigs,
                   int len, double size, double tol) -> int {   // class of sig[0, len), -1 when the cap is reached
    key.v.resize(len);
    for (int i = 0; i < len; ++i) key.v[i] = std::llround(sig[i] / (1e-3 * size));
    std::vector<int> &cands = buckets[key];
    for (int id : cands) {
      bool same = true;
      for (int i = 0; i < len && same; ++i) same = std::fabs(sigs[id][i] - sig[i]) <= tol;
      if (same) return id;
    }
    if (int(sigs.size()) >= kMaxClasses) return -1;
    cands.push_back(int(sigs.size()));
    sigs.emplace_back(sig.begin(), sig.begin() + len);
    return int(sigs.size()) - 1;
  };
  for (int64_t c = 0; c < t.n_local; ++c) {
    const double *center = &t.m_center[size_t(c) * 3];
    const double size = std::cbrt(t.m_volume[c]);
    sig[0] = size;
    const double reach = std::max({ std::fabs(center[0]), std::fabs(center[1]), std::fabs(center[2]) }) + size;
    const double tol = std::max(1e-12 * size, 64 * 2.2e-16 * reach);
    for (int q = 0; q < nq; ++q)
      for (int d = 0; d < 3; ++d) sig[1 + 3 * q + d] = t.m_qxyzw[(size_t(c) * nq + q) * 4 + d] - center[d];
    for (int s = 0; s < nf; ++s) {
      const int32_t f = t.cell_face[size_t(c) * 6 + s];
      if (f < 0) return give_up();   // an open side: not a mesh this layout was made for
      double *fs = &sig[n_vol + s * n_side];
      for (int p = 0; p < fq; ++p)
        for (int d = 0; d < 3; ++d) fs[3 * p + d] = t.f_xyzw[(size_t(f) * fq + p) * 4 + d] - center[d];
      fs[3 * fq] = (t.cell_code[size_t(c) * 6 + s] & 1) ? size : 0.0;   // holder / sharer of that face
      const int64_t other = t.face_holder[f] == c ? t.face_sharer[f] : t.face_holder[f];   // -1 across a boundary face
      fs[3 * fq + 1] = other >= 0 ? size : 0.0;
      for (int q = 0; q < nq; ++q)
        for (int d = 0; d < 3; ++d)
          fs[3 * fq + 2 + 3 * q + d] = other >= 0 ? t.m_qxyzw[(size_t(other) * nq + q) * 4 + d] - center[d] : 0.0;
    }
    if (match(shape_bucket, shape_sig, n_vol, size, tol) < 0) return give_up();
    const size_t before = class_sig.size();
    const int cls = match(class_bucket, class_sig, L, size, tol);
    if (cls < 0) return give_up();
    if (class_sig.size() > before) class_rep.push_back(c);
    t.cell_class[c] = cls;
  }
  t.n_classes = int(class_sig.size()); t.n_patterns = 0; t.n_shapes = int(shape_sig.size());
  // cells grouped by class (a batch of the cell kernel takes 8 cells of ONE class), mesh order kept inside a class
  t.class_ptr.assign(t.n_classes + 1, 0);
  for (int64_t c = 0; c < t.n_local; ++c) t.class_ptr[t.cell_class[c] + 1]++;
  for (int k = 0; k < t.n_classes; ++k) t.class_ptr[k + 1] += t.class_ptr[k];
  t.class_cells.resize(t.n_local);
  {
    std::vector<int32_t> fill(t.class_ptr.begin(), t.class_ptr.end() - 1);
    for (int64_t c = 0; c < t.n_local; ++c) t.class_cells[fill[t.cell_class[c]]++] = int32_t(c);
  }
  // operand tables in the padded shapes the tensor-core kernel loads its fragments from (modal_kernels.cu: MmaShape)
  const int QT = (nq + 7) / 8, K1 = (N + 3) / 4 * 4, MT = (N + 7) / 8, R = QT * 24;
  const int KC = (2 * fq + 3) / 4 * 4, NCH = nf / 2, J = NCH * KC;
  t.mm_A1.assign(size_t(t.n_classes) * QT * 8 * K1, 0.0);
  t.mm_A2.assign(size_t(t.n_classes) * MT * 8 * R, 0.0);
  t.mm_A3.assign(size_t(t.n_classes) * MT * 8 * J, 0.0);
  t.cls_S.assign(size_t(t.n_classes) * N * N, 0.0);
  // the limiter's tables (limiter/weno.hpp): Borrow (:193-203) re-projects the neighbour across a side onto my basis over
  // MY quadrature points -- a constant N x N matrix per (class, side),  P[n][m] = sum_q w_q phi_n^me(x_q) phi_m^adj(x_q);
  // GetSmoothness (:20-34, basis/taylor.hpp:284-302) is a quadratic form of the Taylor coefficients,
  // beta = ct^t W ct,  W[i][j] = sum_q w_q sum_m wgt(order of m) d_m T_i(x_q) d_m T_j(x_q)   (degree <= 2)
  t.cls_P.assign(size_t(t.n_classes) * 6 * N * N, 0.0);
  t.cls_W.assign(size_t(t.n_classes) * N * N, 0.0);
  std::vector<double> T(N), phi(N), phi_adj(N);
  double G[20][3];
  for (int cls = 0; cls < t.n_classes; ++cls) {
    const int64_t c = class_rep[cls];
    const double *S = &t.m_S[size_t(c) * N * N], *center = &t.m_center[size_t(c) * 3];
    std::copy_n(S, N * N, &t.cls_S[size_t(cls) * N * N]);
    double *A1 = &t.mm_A1[size_t(cls) * QT * 8 * K1], *A2 = &t.mm_A2[size_t(cls) * MT * 8 * R], *A3 = &t.mm_A3[size_t(cls) * MT * 8 * J];
    for (int q = 0; q < nq; ++q) {
      const double *xw = &t.m_qxyzw[(size_t(c) * nq + q) * 4];
      const double d[3] = { xw[0] - center[0], xw[1] - center[1], xw[2] - center[2] };
      TaylorValue(degree, d, T.data());
      TaylorGradHost(degree, d, G);
      for (int n = 0; n < N; ++n) {
        double value = 0, g[3] = { 0, 0, 0 };
        for (int m = 0; m <= n; ++m) {
          value += S[n * N + m] * T[m];
          for (int a = 0; a < 3; ++a) g[a] += S[n * N + m] * G[m][a];
        }
        A1[size_t(q) * K1 + n] = value;                                     // phi_n(x_q)
        for (int a = 0; a < 3; ++a) A2[size_t(n) * R + q * 3 + a] = g[a] * xw[3];  // w_q d phi_n / d x_a
      }
    }
    for (int s = 0; s < nf; ++s) {   // Borrow matrices
      const int32_t f = t.cell_face[size_t(c) * 6 + s];
      const int64_t other = t.face_holder[f] == c ? t.face_sharer[f] : t.face_holder[f];
      if (other < 0) continue;
      const double *So = &t.m_S[size_t(other) * N * N], *co = &t.m_center[size_t(other) * 3];
      double *P = &t.cls_P[(size_t(cls) * 6 + s) * N * N];
      for (int q = 0; q < nq; ++q) {
        const double *xw = &t.m_qxyzw[(size_t(c) * nq + q) * 4];
        const double dm[3] = { xw[0] - center[0], xw[1] - center[1], xw[2] - center[2] };
        const double da[3] = { xw[0] - co[0], xw[1] - co[1], xw[2] - co[2] };
        TaylorValue(degree, dm, T.data());
        for (int n = 0; n < N; ++n) { phi[n] = 0; for (int m = 0; m <= n; ++m) phi[n] += S[n * N + m] * T[m]; }
        TaylorValue(degree, da, T.data());
        for (int n = 0; n < N; ++n) { phi_adj[n] = 0; for (int m = 0; m <= n; ++m) phi_adj[n] += So[n * N + m] * T[m]; }
        for (int n = 0; n < N; ++n)
          for (int m = 0; m < N; ++m) P[n * N + m] += xw[3] * phi[n] * phi_adj[m];
      }
    }
    if (degree <= 2) {               // smoothness form; weights as limit_one_cell forms them from 1 / S00^2
      const double volume = 1 / (S[0] * S[0]);
      const double w1 = degree == 1 ? 1.0 : std::pow(volume, 2. / 3 - 1), w2 = std::pow(volume, 4. / 3 - 1);
      double *W = &t.cls_W[size_t(cls) * N * N];
      std::vector<double> D(size_t(N) * N);   // D[m][j] = d_m T_j at the point (basis/taylor.hpp:176-182, 218-260)
      for (int q = 0; q < nq; ++q) {
        const double *xw = &t.m_qxyzw[(size_t(c) * nq + q) * 4];
        const double x = xw[0] - center[0], y = xw[1] - center[1], z = xw[2] - center[2];
        std::fill(D.begin(), D.end(), 0.0);
        for (int j = 1; j < N; ++j) D[j * N + j] = 1;      // r[m] = c[m], m >= 1
        if (degree == 2) {
          enum { X = 1, Y, Z, XX, XY, XZ, YY, YZ, ZZ };
          D[X * N + XX] += 2 * x; D[X * N + XY] += y; D[X * N + XZ] += z;
          D[Y * N + XY] += x; D[Y * N + YY] += 2 * y; D[Y * N + YZ] += z;
          D[Z * N + XZ] += x; D[Z * N + YZ] += y; D[Z * N + ZZ] += 2 * z;
          D[XX * N + XX] += 1; D[YY * N + YY] += 1; D[ZZ * N + ZZ] += 1;   // second derivatives of the squares: 2 c
        }
        for (int m = 1; m < N; ++m) {
          const double wm = (m <= 3 ? w1 : w2) * xw[3];
          for (int i = 0; i < N; ++i)
            for (int j = 0; j < N; ++j) W[i * N + j] += wm * D[m * N + i] * D[m * N + j];
        }
      }
    }
    for (int s = 0; s < nf; ++s) {
      const int32_t f = t.cell_face[size_t(c) * 6 + s];
      for (int p = 0; p < fq; ++p) {
        const double *xw = &t.f_xyzw[(size_t(f) * fq + p) * 4];
        const double d[3] = { xw[0] - center[0], xw[1] - center[1], xw[2] - center[2] };
        TaylorValue(degree, d, T.data());
        const int jj = (s / 2) * KC + (s % 2) * fq + p;
        for (int n = 0; n < N; ++n) {
          double value = 0;
          for (int m = 0; m <= n; ++m) value += S[n * N + m] * T[m];
          A3[size_t(n) * J + jj] = value;                                   // phi_n at the face point
        }
      }
    }
  }
}

}  // namespace

int BuildBasis(int degree, int scheme, Basis *basis) {
  if (degree < 1 || degree > 3) return MCFD_ERR_UNSUPPORTED;
  Basis &bs = *basis;
  std::memset(&bs, 0, sizeof(bs));
  int q = degree + 1;
  bs.q = q;
  if (scheme == MCFD_SCHEME_DG_MODAL) {  // Projection::GaussOnLine = gauss::Legendre<p+1> (projection.hpp:84, legendre.hpp:46-123)
    if (q == 2) {
      bs.x[0] = -std::sqrt(1.0 / 3.0); bs.x[1] = std::sqrt(1.0 / 3.0); bs.w[0] = bs.w[1] = 1.0;
    } else if (q == 3) {
      bs.x[0] = -std::sqrt(0.6); bs.x[1] = 0.0; bs.x[2] = std::sqrt(0.6); bs.w[0] = bs.w[2] = 5.0 / 9.0; bs.w[1] = 8.0 / 9.0;
    } else {
      bs.x[0] = -std::sqrt((3 + 2 * std::sqrt(1.2)) / 7); bs.x[1] = -std::sqrt((3 - 2 * std::sqrt(1.2)) / 7);
      bs.x[2] = -bs.x[1]; bs.x[3] = -bs.x[0];
      bs.w[0] = bs.w[3] = (18 - std::sqrt(30.0)) / 36; bs.w[1] = bs.w[2] = (18 + std::sqrt(30.0)) / 36;
    }
    return MCFD_OK;  // no nodal derivative matrix, no correction function
  }
  if (q == 2) {  // gauss/lobatto.hpp:25-97
    bs.x[0] = -1; bs.x[1] = 1; bs.w[0] = 1; bs.w[1] = 1;
  } else if (q == 3) {
    bs.x[0] = -1; bs.x[1] = 0; bs.x[2] = 1; bs.w[0] = 1.0 / 3; bs.w[1] = 4.0 / 3; bs.w[2] = 1.0 / 3;
  } else {
    bs.x[0] = -1; bs.x[1] = -std::sqrt(1.0 / 5); bs.x[2] = std::sqrt(1.0 / 5); bs.x[3] = 1;
    bs.w[0] = 1.0 / 6; bs.w[1] = 5.0 / 6; bs.w[2] = 5.0 / 6; bs.w[3] = 1.0 / 6;
  }
  for (int a = 0; a < q; ++a)  // derivative of the Lagrange basis m at node a
    for (int m = 0; m < q; ++m) {
      double sum = 0;
      for (int s = 0; s < q; ++s) {
        if (s == m) continue;
        double prod = 1.0 / (bs.x[m] - bs.x[s]);
        for (int r = 0; r < q; ++r) {
          if (r == m || r == s) continue;
          prod *= (bs.x[a] - bs.x[r]) / (bs.x[m] - bs.x[r]);
        }
        sum += prod;
      }
      bs.D[a][m] = sum;
    }
  // basis/vincent.hpp:27-29,71-87 at xi = 1 with c_next = p / (2p + 1)
  double c_next = degree / (2.0 * degree + 1), c_prev = 1 - c_next;
  double d_prev = 0, d_curr = 0, d_next = 0;
  for (int k_curr = 0; k_curr <= degree; ++k_curr) {
    if (k_curr > 0) { d_prev = d_curr; d_curr = d_next; }
    d_next = (k_curr + 1) * Legendre(k_curr, 1.0) + 1.0 * d_curr;
  }
  bs.g_prime = 0.5 * (d_curr + c_prev * d_prev + c_next * d_next);
  return MCFD_OK;
}

// ---- halo lists (part.hpp:964-986) ----
static int CopyHaloLists(const mcfd_mesh &m, HostTables *out, std::string *error) {
  HostTables &t = *out;
  t.nbr_rank.assign(m.nbr_rank, m.nbr_rank + m.n_neighbours);
  if (m.n_neighbours > 0) {
    t.send_ptr.assign(m.send_ptr, m.send_ptr + m.n_neighbours + 1);
    t.recv_ptr.assign(m.recv_ptr, m.recv_ptr + m.n_neighbours + 1);
    t.send_cells.resize(t.send_ptr.back());
    for (int64_t i = 0; i < t.send_ptr.back(); ++i) {
      if (m.send_cells[i] < t.n_inner || m.send_cells[i] >= t.n_local) { *error = "send cell is not an inter cell"; return MCFD_ERR_ARG; }
      t.send_cells[i] = int32_t(m.send_cells[i]);
    }
    if (t.recv_ptr.back() != t.n_ghost) { *error = "recv_ptr does not cover the ghost cells"; return MCFD_ERR_ARG; }
  } else {
    t.send_ptr.assign(1, 0); t.recv_ptr.assign(1, 0);
    if (t.n_ghost != 0) { *error = "ghost cells without neighbours"; return MCFD_ERR_ARG; }
  }
  return MCFD_OK;
}

// TETRA_4 parts (modal DG): the same tables as below from gauss::Tetrahedron / gauss::Triangle.  A face is the
// Triangle3 on the holder's sorted corners (geometry/tetrahedron.hpp:76-95), its frame constant over the face
// (geometry/face.hpp:98-110: n = (q0 - q2) x (q1 - q2) normalised, t = (q0 - q2) normalised, b = n x t).
static int BuildTetraTables(const mcfd_config &cfg, const mcfd_mesh &m, HostTables *out, std::string *error) {
  HostTables &t = *out;
  const int degree = cfg.degree;
  const int fq = t.fq;
  const int64_t nc = t.n_local + t.n_ghost;
  if (int rc = BuildModalTetra(m, degree, &t, error)) return rc;
  t.per_node_metrics = false;
  t.corner_xyz.assign(size_t(t.n_local) * 24, 0.0);  // 4 corners of 8 slots used (Approximate is quadrature based)
  for (int64_t c = 0; c < t.n_local; ++c)
    for (int v = 0; v < 4; ++v) std::copy_n(m.node_xyz + 3 * m.cell_nodes[4 * c + v], 3, &t.corner_xyz[size_t(c) * 24 + 3 * v]);
  t.cell_face.assign(size_t(t.n_local) * 6, -1);
  t.cell_code.assign(size_t(t.n_local) * 6, 0);
  t.face_holder.resize(t.n_faces); t.face_sharer.resize(t.n_faces); t.face_code.assign(t.n_faces, 0);
  if (nc * t.n >= (int64_t(1) << 31) || t.n_faces * fq >= (int64_t(1) << 31)) { *error = "mesh part too large for 32-bit point ids"; return MCFD_ERR_ARG; }
  t.f_xyzw.resize(size_t(t.n_faces) * fq * 4);
  t.face_frame.resize(size_t(t.n_faces) * 9);
  t.per_point_frames = false;
  t.bface_group.resize(t.n_boundary_faces); t.bface_local.resize(t.n_boundary_faces);
  t.bface_xyz.resize(size_t(t.n_boundary_faces) * fq * 3);
  t.group_faces.assign(std::max(t.n_groups, 1), 0);
  const int64_t f_bnd = t.n_local_faces + t.n_ghost_faces;
  double ab[12][2], wl[12];
  if (TriangleRule(degree, ab, wl) != fq) { *error = "triangular rule size"; return MCFD_ERR_ARG; }
  for (int64_t f = 0; f < t.n_faces; ++f) {  // faces keep the caller's order
    const int64_t h = m.face_holder[f], s = m.face_sharer[f];
    const bool boundary = f >= f_bnd;
    if (h < 0 || h >= t.n_local || (boundary ? s >= 0 : (s < 0 || s >= nc))) { *error = "bad face holder/sharer"; return MCFD_ERR_ARG; }
    if (!boundary && (f < t.n_local_faces) != (s < t.n_local)) { *error = "face list is not ordered local|ghost|boundary"; return MCFD_ERR_ARG; }
    const int64_t *hn = m.cell_nodes + 4 * h;
    const int hs = FindTetSide(hn, m.face_nodes + 3 * f);
    if (hs < 0) { *error = "face nodes do not form a face of the holder"; return MCFD_ERR_ARG; }
    if (!boundary) {
      const int ss = FindTetSide(m.cell_nodes + 4 * s, m.face_nodes + 3 * f);
      if (ss < 0) { *error = "face nodes do not form a face of the sharer"; return MCFD_ERR_ARG; }
      if (s < t.n_local) {
        if (t.cell_face[s * 6 + ss] >= 0) { *error = "two faces on one side of a cell"; return MCFD_ERR_ARG; }
        t.cell_face[s * 6 + ss] = int32_t(f);
        t.cell_code[s * 6 + ss] = 1;
      }
    }
    if (t.cell_face[h * 6 + hs] >= 0) { *error = "two faces on one side of a cell"; return MCFD_ERR_ARG; }
    t.cell_face[h * 6 + hs] = int32_t(f);
    t.face_holder[f] = int32_t(h); t.face_sharer[f] = boundary ? -1 : int32_t(s);
    const double *q[3];
    for (int v = 0; v < 3; ++v) q[v] = m.node_xyz + 3 * hn[kTetFace[hs][v]];
    double r0[3], r1[3];  // rows of the face Jacobian: d x / d a = q0 - q2, d x / d b = q1 - q2
    for (int d = 0; d < 3; ++d) { r0[d] = q[0][d] - q[2][d]; r1[d] = q[1][d] - q[2][d]; }
    const double nrm[3] = { r0[1] * r1[2] - r0[2] * r1[1], r0[2] * r1[0] - r0[0] * r1[2], r0[0] * r1[1] - r0[1] * r1[0] };
    const double ln = std::sqrt(nrm[0] * nrm[0] + nrm[1] * nrm[1] + nrm[2] * nrm[2]);
    const double lt = std::sqrt(r0[0] * r0[0] + r0[1] * r0[1] + r0[2] * r0[2]);
    if (!(ln > 0)) { *error = "degenerate triangle"; return MCFD_ERR_ARG; }
    double *fr = &t.face_frame[size_t(f) * 9];
    for (int d = 0; d < 3; ++d) { fr[d] = nrm[d] / ln; fr[3 + d] = r0[d] / lt; }
    fr[6] = fr[1] * fr[5] - fr[2] * fr[4]; fr[7] = fr[2] * fr[3] - fr[0] * fr[5]; fr[8] = fr[0] * fr[4] - fr[1] * fr[3];
    // gauss/element.hpp:92-96 : global weight = local weight x sqrt(det(J J^t))
    const double a11 = lt * lt, a22 = r1[0] * r1[0] + r1[1] * r1[1] + r1[2] * r1[2], a12 = r0[0] * r1[0] + r0[1] * r1[1] + r0[2] * r1[2];
    const double area2 = std::sqrt(a11 * a22 - a12 * a12);
    for (int p = 0; p < fq; ++p) {
      const double shape[3] = { ab[p][0], ab[p][1], 1.0 - ab[p][0] - ab[p][1] };
      double *xw = &t.f_xyzw[(size_t(f) * fq + p) * 4];
      for (int d = 0; d < 3; ++d) xw[d] = q[0][d] * shape[0] + q[1][d] * shape[1] + q[2][d] * shape[2];
      xw[3] = wl[p] * area2;
      if (boundary) std::copy_n(xw, 3, &t.bface_xyz[(size_t(f - f_bnd) * fq + p) * 3]);
    }
    if (boundary) {
      const int g = m.bface_group ? m.bface_group[f - f_bnd] : 0;
      if (g < 0 || g >= t.n_groups) { *error = "boundary group id out of range"; return MCFD_ERR_ARG; }
      t.bface_group[f - f_bnd] = g;
      t.bface_local[f - f_bnd] = int32_t(t.group_faces[g]++);
    }
  }
  // neighbours through interior faces in the caller's face order (part.hpp:1019-1032, 1066-1078); the limiter's frame
  // is that of the face's quadrature point 0 (limiter/weno.hpp:212-214)
  t.adj_ptr.assign(t.n_local + 1, 0);
  for (int64_t f = 0; f < f_bnd; ++f) {
    t.adj_ptr[m.face_holder[f] + 1]++;
    if (f < t.n_local_faces) t.adj_ptr[m.face_sharer[f] + 1]++;
  }
  for (int64_t c = 0; c < t.n_local; ++c) t.adj_ptr[c + 1] += t.adj_ptr[c];
  t.adj_cell.resize(t.adj_ptr[t.n_local]); t.adj_face.resize(t.adj_ptr[t.n_local]);
  std::vector<int32_t> fill(t.n_local, 0);
  for (int64_t f = 0; f < f_bnd; ++f) {
    const int64_t h = m.face_holder[f], s = m.face_sharer[f];
    const int32_t packed = int32_t(f * fq);
    t.adj_cell[t.adj_ptr[h] + fill[h]] = int32_t(s); t.adj_face[t.adj_ptr[h] + fill[h]++] = packed;
    if (f < t.n_local_faces) { t.adj_cell[t.adj_ptr[s] + fill[s]] = int32_t(h); t.adj_face[t.adj_ptr[s] + fill[s]++] = packed; }
  }
  if (!(cfg.flags & MCFD_FLAG_NO_SHAPE_CLASSES)) BuildShapeClasses(degree, &t);
  return CopyHaloLists(m, out, error);
}

int BuildTables(const mcfd_config &cfg, const mcfd_mesh &m, const Basis &bs, HostTables *out, std::string *error) {
  HostTables &t = *out;
  const bool modal = cfg.scheme == MCFD_SCHEME_DG_MODAL;
  const int q = bs.q, nq = q * q * q, fq = q * q;
  const int n = modal ? (cfg.degree == 1 ? 4 : cfg.degree == 2 ? 10 : 20) : nq;  // fields per cell / 5
  t.q = q; t.n = n; t.fq = fq; t.nq = nq; t.modal = modal;
  t.n_local = m.n_local_cells; t.n_ghost = m.n_ghost_cells; t.n_inner = m.n_inner_cells;
  t.n_local_faces = m.n_local_faces; t.n_ghost_faces = m.n_ghost_faces; t.n_boundary_faces = m.n_boundary_faces;
  t.n_faces = m.n_local_faces + m.n_ghost_faces + m.n_boundary_faces;
  t.n_groups = m.n_bc_groups;
  const int64_t nc = t.n_local + t.n_ghost;
  if (nc >= (int64_t(1) << 31) || t.n_faces >= (int64_t(1) << 31)) { *error = "mesh part too large for 32-bit ids"; return MCFD_ERR_ARG; }
  if (t.n_local <= 0 || !m.node_xyz || !m.cell_nodes) { *error = "empty mesh"; return MCFD_ERR_ARG; }
  if (t.n_faces > 0 && (!m.face_holder || !m.face_sharer || !m.face_nodes)) { *error = "missing face arrays"; return MCFD_ERR_ARG; }
  t.elem = m.cell_type; t.sides = 6;
  if (m.cell_type == MCFD_CELL_TETRA_4) {
    // FR and DGSEM live on polynomial::Hexahedron (tensor-product Lobatto nodes); tetrahedra exist for the modal basis only
    if (!modal) { *error = "tetrahedra are supported by the modal DG scheme only (as in the reference)"; return MCFD_ERR_UNSUPPORTED; }
    t.sides = 4;
    t.nq = cfg.degree == 1 ? 4 : cfg.degree == 2 ? 14 : 24;    // part.hpp:448-452
    t.fq = cfg.degree == 1 ? 3 : cfg.degree == 2 ? 6 : 12;     // part.hpp:439-443
    return BuildTetraTables(cfg, m, out, error);
  } else if (m.cell_type != MCFD_CELL_HEXA_8) {
    *error = "unknown cell type"; return MCFD_ERR_UNSUPPORTED;
  }

  // ---- metrics: are all cells affine (constant Jacobian)? ----
  auto corners = [&](int64_t c, double xn[8][3]) {
    for (int v = 0; v < 8; ++v) {
      int64_t id = m.cell_nodes[8 * c + v];
      for (int b = 0; b < 3; ++b) xn[v][b] = m.node_xyz[3 * id + b];
    }
  };
  bool affine = !(cfg.flags & MCFD_FLAG_FORCE_PER_NODE_METRICS) && !modal;
  for (int64_t c = 0; c < nc && affine; ++c) {
    double xn[8][3];
    corners(c, xn);
    // affine <=> the Jacobian of the trilinear map is the same at all 8 corners.  (Two opposite corners are not
    // enough: the xi*eta*zeta term of the map adds a7 * eta * zeta etc. to J, and those products are +1 at both.)
    static const double kC[8][3] = { {-1,-1,-1},{1,-1,-1},{1,1,-1},{-1,1,-1},{-1,-1,1},{1,-1,1},{1,1,1},{-1,1,1} };
    double J0[3][3];
    HexaJacobian(xn, kC[0], J0);
    double scale = 0, diff = 0;
    for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) scale = std::max(scale, std::fabs(J0[a][b]));
    for (int v = 1; v < 8; ++v) {
      double J1[3][3];
      HexaJacobian(xn, kC[v], J1);
      for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) diff = std::max(diff, std::fabs(J0[a][b] - J1[a][b]));
    }
    if (diff > 1e-13 * scale) affine = false;
  }
  t.per_node_metrics = !affine;
  const int mstride = t.per_node_metrics ? n : 1;
  t.metrics.resize(modal ? 0 : size_t(nc) * mstride * 10);
  if (modal) BuildModalCells(m, bs, &t);
  for (int64_t c = 0; c < nc && !modal; ++c) {
    double xn[8][3], J[3][3];
    corners(c, xn);
    if (!t.per_node_metrics) {
      double l[3] = { 0, 0, 0 };
      HexaJacobian(xn, l, J);
      Metrics(J, &t.metrics[size_t(c) * 10]);
      if (!(t.metrics[size_t(c) * 10 + 9] > 0)) { *error = "non-positive Jacobian determinant"; return MCFD_ERR_ARG; }
    } else {
      for (int i = 0; i < q; ++i) for (int j = 0; j < q; ++j) for (int k = 0; k < q; ++k) {
        double l[3] = { bs.x[i], bs.x[j], bs.x[k] };
        HexaJacobian(xn, l, J);
        double *mm = &t.metrics[(size_t(c) * n + (i * q + j) * q + k) * 10];
        Metrics(J, mm);
        if (!(mm[9] > 0)) { *error = "non-positive Jacobian determinant"; return MCFD_ERR_ARG; }
      }
    }
  }
  t.corner_xyz.resize(size_t(t.n_local) * 24);
  for (int64_t c = 0; c < t.n_local; ++c) {
    double xn[8][3];
    corners(c, xn);
    std::memcpy(&t.corner_xyz[size_t(c) * 24], xn, sizeof(xn));
  }

  // ---- faces ----
  t.cell_face.assign(size_t(t.n_local) * 6, -1);
  t.cell_code.assign(size_t(t.n_local) * 6, 0);
  t.face_holder.resize(t.n_faces); t.face_sharer.resize(t.n_faces); t.face_code.resize(t.n_faces);
  // node records and face points are 32-bit ids; every offset derived from them (x 5 components, x 10 metric entries)
  // is formed in 64 bits by the kernels
  if (nc * n >= (int64_t(1) << 31) || t.n_faces * fq >= (int64_t(1) << 31)) { *error = "mesh part too large for 32-bit point ids"; return MCFD_ERR_ARG; }
  if (!modal) { t.pt_holder.resize(size_t(t.n_faces) * fq); t.pt_sharer.resize(size_t(t.n_faces) * fq); }
  std::vector<int32_t> face_p0(modal ? t.n_faces : 0, 0);
  if (modal) t.f_xyzw.resize(size_t(t.n_faces) * fq * 4);
  std::vector<double> frames(size_t(t.n_faces) * fq * 9);
  t.bface_group.resize(t.n_boundary_faces); t.bface_local.resize(t.n_boundary_faces);
  t.bface_xyz.resize(size_t(t.n_boundary_faces) * fq * 3);
  t.group_faces.assign(std::max(t.n_groups, 1), 0);
  bool constant_frames = true;
  const int64_t f_bnd = t.n_local_faces + t.n_ghost_faces;
  // Internal face order: inside each category (local | ghost | boundary) faces are sorted by (chunk of
  // kFaceChunk holder cells, side of the holder they lie on), mesh order kept otherwise.  A warp of the face
  // kernel covers 32 consecutive flux points (3.5 faces at p = 2): faces of one orientation see the same
  // frame and similar jumps, so the trip count of the Riemann iteration varies less inside a warp than with
  // interleaved orientations (a flow aligned with a coordinate plane needs 1 evaluation on the faces of that
  // plane and 2-6 elsewhere; ncu: 21 -> 28 active lanes of 32).  The chunking keeps the three orientation
  // passes over a group of cells close in time, so their node data is re-read from L2, not from HBM.
  constexpr int64_t kFaceChunk = 512;
  std::vector<int64_t> order(t.n_faces);
  std::vector<int32_t> local_in_group(t.n_boundary_faces, 0);
  {
    std::vector<int64_t> key(t.n_faces, 0);
    for (int64_t f = 0; f < t.n_faces; ++f) {
      int64_t h = m.face_holder[f];
      if (h < 0 || h >= t.n_local) { *error = "bad face holder/sharer"; return MCFD_ERR_ARG; }
      int hs = FindSide(m.cell_nodes + 8 * h, m.face_nodes + 4 * f);
      if (hs < 0) { *error = "face nodes do not form a face of the holder"; return MCFD_ERR_ARG; }
      key[f] = (h / kFaceChunk) * 8 + hs;
      order[f] = f;
    }
    const int64_t bounds[4] = { 0, t.n_local_faces, f_bnd, t.n_faces };
    for (int c = 0; c < 3 && !(cfg.flags & MCFD_FLAG_KEEP_FACE_ORDER); ++c)
      std::stable_sort(order.begin() + bounds[c], order.begin() + bounds[c + 1],
                       [&](int64_t a, int64_t b) { return key[a] < key[b]; });
    for (int64_t b = 0; b < t.n_boundary_faces; ++b) {  // index inside the group follows the MESH order (mcfd_bc_points)
      int g = m.bface_group ? m.bface_group[b] : 0;
      if (g < 0 || g >= t.n_groups) { *error = "boundary group id out of range"; return MCFD_ERR_ARG; }
      local_in_group[b] = int32_t(t.group_faces[g]++);
    }
  }
  for (int64_t f = 0; f < t.n_faces; ++f) {
    const int64_t fo = order[f];  // face id in the caller's mesh
    int64_t h = m.face_holder[fo], s = m.face_sharer[fo];
    bool boundary = f >= f_bnd;
    if (h < 0 || h >= t.n_local || (boundary ? s >= 0 : (s < 0 || s >= nc))) { *error = "bad face holder/sharer"; return MCFD_ERR_ARG; }
    if (!boundary && (f < t.n_local_faces) != (s < t.n_local)) { *error = "face list is not ordered local|ghost|boundary"; return MCFD_ERR_ARG; }
    const int64_t *hn = m.cell_nodes + 8 * h;
    int hs = FindSide(hn, m.face_nodes + 4 * fo);
    if (hs < 0) { *error = "face nodes do not form a face of the holder"; return MCFD_ERR_ARG; }
    int ss = 0, h2s = 0;
    if (!boundary) {
      const int64_t *sn = m.cell_nodes + 8 * s;
      ss = FindSide(sn, m.face_nodes + 4 * fo);
      if (ss < 0) { *error = "face nodes do not form a face of the sharer"; return MCFD_ERR_ARG; }
      h2s = FindPerm(hn, hs, sn, ss);
      int s2h = FindPerm(sn, ss, hn, hs);
      if (h2s < 0 || s2h < 0) { *error = "cannot orient the sharer's face"; return MCFD_ERR_ARG; }
      if (s < t.n_local) {
        if (t.cell_face[s * 6 + ss] >= 0) { *error = "two faces on one side of a cell"; return MCFD_ERR_ARG; }
        t.cell_face[s * 6 + ss] = int32_t(f);
        t.cell_code[s * 6 + ss] = uint8_t(1 | (s2h << 1));
      }
    }
    if (t.cell_face[h * 6 + hs] >= 0) { *error = "two faces on one side of a cell"; return MCFD_ERR_ARG; }
    t.cell_face[h * 6 + hs] = int32_t(f);
    t.cell_code[h * 6 + hs] = 0;
    t.face_holder[f] = int32_t(h); t.face_sharer[f] = boundary ? -1 : int32_t(s);
    t.face_code[f] = uint16_t(hs | (ss << 3) | (h2s << 6));
    // face point p = (a, b) of the holder's lattice <-> solution node on either side
    // (fr/lobatto.hpp:76-162 and dg/lobatto.hpp:55-75 find these by comparing coordinates)
    for (int a = 0; a < q && !modal; ++a) for (int b = 0; b < q; ++b) {
      t.pt_holder[size_t(f) * fq + a * q + b] = int32_t(h * n + SideNode(q, hs, a, b));
      int a2, b2;
      ApplyPerm(h2s, q, a, b, &a2, &b2);
      t.pt_sharer[size_t(f) * fq + a * q + b] = boundary ? -1 : int32_t(s * n + SideNode(q, ss, a2, b2));
    }

    // the face as the reference sees it: a Quadrangle4 on the holder's sorted corners
    // (geometry/hexahedron.hpp:406-416); face-local x runs corner0->corner1, y runs corner0->corner3
    double xf[4][3];
    for (int v = 0; v < 4; ++v)
      for (int b = 0; b < 3; ++b) xf[v][b] = m.node_xyz[3 * hn[kFaceTable[hs][v]] + b];
    const double *c0 = kCorner[kFaceTable[hs][0]], *c1 = kCorner[kFaceTable[hs][1]], *c3 = kCorner[kFaceTable[hs][3]];
    int dim_x = 0, dim_y = 0; double sign_x = 1, sign_y = 1;
    for (int d = 0; d < 3; ++d) {
      if (c1[d] != c0[d]) { dim_x = d; sign_x = (c1[d] - c0[d]) / 2; }
      if (c3[d] != c0[d]) { dim_y = d; sign_y = (c3[d] - c0[d]) / 2; }
    }
    static const double kQuad[4][2] = { {-1,-1},{1,-1},{1,1},{-1,1} };
    for (int a = 0; a < q; ++a) for (int b = 0; b < q; ++b) {
      int p = a * q + b;
      double loc[3];
      loc[kSideDir[hs]] = kSideHigh[hs] ? 1.0 : -1.0;
      loc[kSideFree[hs][0]] = bs.x[a];
      loc[kSideFree[hs][1]] = bs.x[b];
      double lx = sign_x * loc[dim_x], ly = sign_y * loc[dim_y];
      double g[3] = { 0, 0, 0 }, Jf[2][3] = { { 0, 0, 0 }, { 0, 0, 0 } };
      for (int v = 0; v < 4; ++v) {
        double N = (1 + kQuad[v][0] * lx) * (1 + kQuad[v][1] * ly) / 4;
        double dx = kQuad[v][0] * (1 + kQuad[v][1] * ly) / 4, dy = kQuad[v][1] * (1 + kQuad[v][0] * lx) / 4;
        for (int d = 0; d < 3; ++d) { g[d] += xf[v][d] * N; Jf[0][d] += dx * xf[v][d]; Jf[1][d] += dy * xf[v][d]; }
      }
      double *fr = &frames[(size_t(f) * fq + p) * 9];  // geometry/face.hpp:98-110
      double nrm[3] = { Jf[0][1] * Jf[1][2] - Jf[0][2] * Jf[1][1], Jf[0][2] * Jf[1][0] - Jf[0][0] * Jf[1][2],
                        Jf[0][0] * Jf[1][1] - Jf[0][1] * Jf[1][0] };
      double ln = std::sqrt(nrm[0] * nrm[0] + nrm[1] * nrm[1] + nrm[2] * nrm[2]);
      double lt = std::sqrt(Jf[0][0] * Jf[0][0] + Jf[0][1] * Jf[0][1] + Jf[0][2] * Jf[0][2]);
      for (int d = 0; d < 3; ++d) { fr[d] = nrm[d] / ln; fr[3 + d] = Jf[0][d] / lt; }
      fr[6] = fr[1] * fr[5] - fr[2] * fr[4]; fr[7] = fr[2] * fr[3] - fr[0] * fr[5]; fr[8] = fr[0] * fr[4] - fr[1] * fr[3];
      if (p > 0) {
        const double *f0 = &frames[size_t(f) * fq * 9];
        // a flat face far from the origin: its edge vectors are differences of corner coordinates and carry
        // ~eps |x| / h of relative rounding, which is what two points of the same flat face may differ by
        double xmax = 0;
        for (int v = 0; v < 4; ++v) for (int d = 0; d < 3; ++d) xmax = std::max(xmax, std::fabs(xf[v][d]));
        const double frame_tol = std::max(1e-13, 64 * 2.2e-16 * xmax / std::min(lt, ln / lt));
        for (int d = 0; d < 9; ++d) if (std::fabs(fr[d] - f0[d]) > frame_tol) constant_frames = false;
      }
      if (boundary) std::memcpy(&t.bface_xyz[(size_t(f - f_bnd) * fq + p) * 3], g, sizeof(g));
      if (modal) {  // gauss/element.hpp:92-96 : global weight = w_local * sqrt(det(J J^T))
        double a11 = lt * lt, a22 = Jf[1][0] * Jf[1][0] + Jf[1][1] * Jf[1][1] + Jf[1][2] * Jf[1][2];
        double a12 = Jf[0][0] * Jf[1][0] + Jf[0][1] * Jf[1][1] + Jf[0][2] * Jf[1][2];
        double *xw = &t.f_xyzw[(size_t(f) * fq + p) * 4];
        xw[0] = g[0]; xw[1] = g[1]; xw[2] = g[2];
        xw[3] = bs.w[a] * bs.w[b] * std::sqrt(a11 * a22 - a12 * a12);
        // the reference's quadrature point 0 of this face (gauss/quadrangle.hpp:62-73: face-local (x_0, x_0)), whose
        // frame weno::Eigen takes through adj_face.riemann(0) (limiter/weno.hpp:212-214)
        if (std::fabs(lx - bs.x[0]) < 1e-12 && std::fabs(ly - bs.x[0]) < 1e-12) face_p0[f] = p;
      }
    }
    if (boundary) {
      t.bface_group[f - f_bnd] = m.bface_group ? m.bface_group[fo - f_bnd] : 0;
      t.bface_local[f - f_bnd] = local_in_group[fo - f_bnd];
    }
  }
  t.per_point_frames = !constant_frames;
  if (modal) {
    // neighbours through interior faces, in the caller's face order (part.hpp:1019-1032, 1066-1078): the order in
    // which weno::Lazy / weno::Eigen walk adj_cells_ / adj_faces_.  adj_face packs (internal face id, point whose
    // frame the limiter uses) as face * fq + p0.
    std::vector<int64_t> internal_of(t.n_faces);
    for (int64_t f = 0; f < t.n_faces; ++f) internal_of[order[f]] = f;
    t.adj_ptr.assign(t.n_local + 1, 0);
    for (int64_t fo = 0; fo < f_bnd; ++fo) {
      t.adj_ptr[m.face_holder[fo] + 1]++;
      if (fo < t.n_local_faces) t.adj_ptr[m.face_sharer[fo] + 1]++;
    }
    for (int64_t c = 0; c < t.n_local; ++c) t.adj_ptr[c + 1] += t.adj_ptr[c];
    t.adj_cell.resize(t.adj_ptr[t.n_local]); t.adj_face.resize(t.adj_ptr[t.n_local]);
    std::vector<int32_t> fill(t.n_local, 0);
    for (int64_t fo = 0; fo < f_bnd; ++fo) {
      const int64_t h = m.face_holder[fo], s = m.face_sharer[fo], f = internal_of[fo];
      const int32_t packed = int32_t(f * fq + face_p0[f]);
      t.adj_cell[t.adj_ptr[h] + fill[h]] = int32_t(s); t.adj_face[t.adj_ptr[h] + fill[h]++] = packed;
      if (fo < t.n_local_faces) { t.adj_cell[t.adj_ptr[s] + fill[s]] = int32_t(h); t.adj_face[t.adj_ptr[s] + fill[s]++] = packed; }
    }
    for (int64_t c = 0; c < t.n_local; ++c)
      if (t.adj_ptr[c + 1] - t.adj_ptr[c] > 6) { *error = "a hexahedron with more than 6 neighbours"; return MCFD_ERR_ARG; }
  }
  if (t.per_point_frames) {
    t.face_frame.swap(frames);
  } else {
    t.face_frame.resize(size_t(t.n_faces) * 9);
    for (int64_t f = 0; f < t.n_faces; ++f) std::memcpy(&t.face_frame[size_t(f) * 9], &frames[size_t(f) * fq * 9], 9 * sizeof(double));
  }

  if (modal && !(cfg.flags & MCFD_FLAG_NO_SHAPE_CLASSES)) BuildShapeClasses(cfg.degree, &t);
  return CopyHaloLists(m, out, error);
}

}  // namespace mcfd
