/* oracle/modal_oracle.h -- TEST INFRASTRUCTURE ONLY, included by euler_oracle.c.
 *
 * CPU restatement of the MODAL discontinuous Galerkin path and of the WENO limiters:
 *   spatial/dg/general.hpp:50-212        volume, interior / ghost face and boundary terms
 *   polynomial/projection.hpp:25-47,63-292   Projection, ProjectionWrapper, Project, GetAverage
 *   basis/linear.hpp:87-134, gauss/function.hpp:109-141   OrthoNormal basis (Gram-Schmidt on the Taylor basis)
 *   basis/taylor.hpp:163-507             Taylor<3, 1|2|3>: values, gradients, all partial derivatives, smoothness
 *   limiter/weno.hpp:20-264              GetSmoothness, Lazy, Eigen (IsNotSmooth, Borrow, ReconstructOnFace)
 *   riemann/euler/eigen.hpp:27-56        left / right eigenvector matrices
 *   mesh/part.hpp:1304-1334              Part::Reconstruct (inner cells, halo update, inter cells)
 * Paths are relative to /root/reference/include/mini.  Pinned against oracle/_ref/libref_dg_*.so in
 * tests/test_oracle_vs_reference.py and against tests/golden/dg_*.npz.
 */

#define MN_MAX 20 /* Taylor<3,3>::N */

/* gauss/legendre.hpp:46-123 */
static void legendre_rule(int q, double *x, double *w) {
  if (q == 2) { x[0] = -sqrt(1.0 / 3.0); x[1] = sqrt(1.0 / 3.0); w[0] = w[1] = 1.0; }
  else if (q == 3) { x[0] = -sqrt(0.6); x[1] = 0.0; x[2] = sqrt(0.6); w[0] = w[2] = 5.0 / 9.0; w[1] = 8.0 / 9.0; }
  else {
    x[0] = -sqrt((3 + 2 * sqrt(1.2)) / 7); x[1] = -sqrt((3 - 2 * sqrt(1.2)) / 7); x[2] = -x[1]; x[3] = -x[0];
    w[0] = w[3] = (18 - sqrt(30.0)) / 36; w[1] = w[2] = (18 + sqrt(30.0)) / 36;
  }
}

static int taylor_size(int degree) { return degree == 1 ? 4 : degree == 2 ? 10 : 20; }

/* basis/taylor.hpp:171-175, 213-217, 336-343 : GetValue */
static void taylor_value(int degree, const double *d, double *T) {
  double x = d[0], y = d[1], z = d[2];
  T[0] = 1; T[1] = x; T[2] = y; T[3] = z;
  if (degree >= 2) { T[4] = x * x; T[5] = x * y; T[6] = x * z; T[7] = y * y; T[8] = y * z; T[9] = z * z; }
  if (degree >= 3) {
    double xx = x * x, xy = x * y, xz = x * z, yy = y * y, yz = y * z, zz = z * z;
    T[10] = x * xx; T[11] = x * xy; T[12] = x * xz; T[13] = x * yy; T[14] = x * yz; T[15] = x * zz;
    T[16] = y * yy; T[17] = y * yz; T[18] = y * zz; T[19] = z * zz;
  }
}

/* gradients of the Taylor monomials; Taylor::GetGradValue(local, coeff) (taylor.hpp:183-194, 262-282, 431-476)
 * is coeff * G with this G[m][dim] */
static void taylor_grad(int degree, const double *d, double G[MN_MAX][3]) {
  double x = d[0], y = d[1], z = d[2];
  memset(G, 0, sizeof(double) * MN_MAX * 3);
  G[1][0] = 1; G[2][1] = 1; G[3][2] = 1;
  if (degree >= 2) {
    G[4][0] = 2 * x; G[5][0] = y; G[5][1] = x; G[6][0] = z; G[6][2] = x;
    G[7][1] = 2 * y; G[8][1] = z; G[8][2] = y; G[9][2] = 2 * z;
  }
  if (degree >= 3) {
    double xx = x * x, xy = x * y, xz = x * z, yy = y * y, yz = y * z, zz = z * z;
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

/* basis/taylor.hpp:176-182 (P=1), 218-260 (P=2), 322-429 (P=3) : GetPdvValue -- every partial derivative of
 * sum_m c[m] T_m at d; c, res are [N][K].  P=3 reproduces `res.col(YY) += res.col(XX)` (taylor.hpp:383). */
static void taylor_pdv(int degree, const double *d, double c[MN_MAX][K], double res[MN_MAX][K]) {
  enum { X = 1, Y, Z, XX, XY, XZ, YY, YZ, ZZ, XXX, XXY, XXZ, XYY, XYZ, XZZ, YYY, YYZ, YZZ, ZZZ };
  int N = taylor_size(degree);
  double x = d[0], y = d[1], z = d[2];
  for (int m = 0; m < N; ++m) for (int k = 0; k < K; ++k) res[m][k] = m == 0 ? 0.0 : c[m][k];
  if (degree == 1) return;
  double xx = x * x, xy = x * y, xz = x * z, yy = y * y, yz = y * z, zz = z * z;
  for (int k = 0; k < K; ++k) {
    if (degree == 2) {
      res[X][k] += c[XX][k] * (2 * x); res[X][k] += c[XY][k] * y; res[X][k] += c[XZ][k] * z;
      res[Y][k] += c[XY][k] * x; res[Y][k] += c[YY][k] * (2 * y); res[Y][k] += c[YZ][k] * z;
      res[Z][k] += c[XZ][k] * x; res[Z][k] += c[YZ][k] * y; res[Z][k] += c[ZZ][k] * (2 * z);
      res[XX][k] += c[XX][k]; res[YY][k] += c[YY][k]; res[ZZ][k] += c[ZZ][k];
    } else {
      res[X][k] += c[XX][k] * (2 * x); res[X][k] += c[XY][k] * y; res[X][k] += c[XZ][k] * z;
      res[X][k] += c[XXX][k] * (3 * xx); res[X][k] += c[XXY][k] * (2 * xy); res[X][k] += c[XXZ][k] * (2 * xz);
      res[X][k] += c[XYY][k] * yy; res[X][k] += c[XYZ][k] * yz; res[X][k] += c[XZZ][k] * zz;
      res[Y][k] += c[XY][k] * x; res[Y][k] += c[YY][k] * (2 * y); res[Y][k] += c[YZ][k] * z;
      res[Y][k] += c[XXY][k] * xx; res[Y][k] += c[XYZ][k] * xz; res[Y][k] += c[XYY][k] * (2 * xy);
      res[Y][k] += c[YYY][k] * (3 * yy); res[Y][k] += c[YYZ][k] * (2 * yz); res[Y][k] += c[YZZ][k] * zz;
      res[Z][k] += c[XZ][k] * x; res[Z][k] += c[YZ][k] * y; res[Z][k] += c[ZZ][k] * (2 * z);
      res[Z][k] += c[XXZ][k] * xx; res[Z][k] += c[XYZ][k] * xy; res[Z][k] += c[XZZ][k] * (2 * xz);
      res[Z][k] += c[YYZ][k] * yy; res[Z][k] += c[YZZ][k] * (2 * yz); res[Z][k] += c[ZZZ][k] * (3 * zz);
      res[XX][k] += c[XXY][k] * y; res[XX][k] += c[XXZ][k] * z; res[XX][k] += c[XXX][k] * x * 3;
      res[XX][k] += res[XX][k];
      res[XY][k] += c[XXY][k] * x * 2; res[XY][k] += c[XYY][k] * y * 2; res[XY][k] += c[XYZ][k] * z;
      res[XZ][k] += c[XXZ][k] * x * 2; res[XZ][k] += c[XZZ][k] * z * 2; res[XZ][k] += c[XYZ][k] * y;
      res[YY][k] += c[XYY][k] * x; res[YY][k] += c[YYZ][k] * z; res[YY][k] += c[YYY][k] * y * 3;
      res[YY][k] += res[XX][k];  /* sic, taylor.hpp:383 */
      res[YZ][k] += c[XYZ][k] * x; res[YZ][k] += c[YYZ][k] * y * 2; res[YZ][k] += c[YZZ][k] * z * 2;
      res[ZZ][k] += c[XZZ][k] * x; res[ZZ][k] += c[YZZ][k] * y; res[ZZ][k] += c[ZZZ][k] * z * 3;
      res[ZZ][k] += res[ZZ][k];
      res[XXX][k] *= 6; res[XXY][k] += res[XXY][k]; res[XXZ][k] += res[XXZ][k]; res[XYY][k] += res[XYY][k];
      res[XZZ][k] += res[XZZ][k]; res[YYY][k] *= 6; res[YYZ][k] += res[YYZ][k]; res[YZZ][k] += res[YZZ][k];
      res[ZZZ][k] *= 6;
    }
  }
}

/* basis/taylor.hpp:196-205 (P=1), 284-302 (P=2), 478-507 (P=3) : Taylor::GetSmoothness */
static void taylor_smoothness(int degree, double integral[MN_MAX][K], double volume, double *beta) {
  for (int k = 0; k < K; ++k) {
    double s = integral[1][k];
    s += integral[2][k]; s += integral[3][k];
    if (degree >= 2) {
      double w1 = pow(volume, 2. / 3 - 1), w2 = pow(volume, 4. / 3 - 1);
      s *= w1;
      for (int m = 4; m < 10; ++m) s += integral[m][k] * w2;
      if (degree >= 3) for (int m = 10; m < 20; ++m) s += integral[m][k] * volume;
    }
    beta[k] = s;
  }
}

/* gauss/function.hpp:109-141 : OrthoNormalize.  A = int T T^t over the cell's quadrature; S lower triangular. */
static void ortho_normalize(int N, double A[MN_MAX][MN_MAX], double S[MN_MAX][MN_MAX]) {
  memset(S, 0, sizeof(double) * MN_MAX * MN_MAX);
  for (int i = 0; i < N; ++i) S[i][i] = 1;
  S[0][0] = 1 / sqrt(A[0][0]);
  for (int i = 1; i < N; ++i) {
    for (int j = 0; j < i; ++j) {
      double temp = 0;
      for (int k = 0; k <= j; ++k) temp += S[j][k] * A[k][i];
      for (int l = 0; l <= j; ++l) S[i][l] -= temp * S[j][l];
    }
    double norm_sq = 0;
    for (int j = 0; j <= i; ++j) {
      double sum = 0, Sij = S[i][j];
      for (int k = 0; k < j; ++k) sum += 2 * S[i][k] * A[k][j];
      norm_sq += Sij * (Sij * A[j][j] + sum);
    }
    double nrm = sqrt(norm_sq);
    for (int j = 0; j < N; ++j) S[i][j] /= nrm;
  }
}

/* ------------------------------------------------------------------------------------------ */
/* basis/linear.hpp:116-134 : OrthoNormal::operator() and GetGradValue for one cell             */
static void modal_basis(const oracle_sim *sim, const PartTab *pt, int64_t cell, const double *x, double *phi) {
  int N = sim->n;
  double d[3], T[MN_MAX];
  for (int b = 0; b < 3; ++b) d[b] = x[b] - pt->m_center[cell * 3 + b];
  taylor_value(sim->degree, d, T);
  const double *S = pt->m_S + cell * N * N;
  for (int n = 0; n < N; ++n) {
    double s = 0;
    for (int m = 0; m <= n; ++m) s += S[n * N + m] * T[m];
    phi[n] = s;
  }
}
static void modal_basis_grad(const oracle_sim *sim, const PartTab *pt, int64_t cell, const double *x, double G[MN_MAX][3]) {
  int N = sim->n;
  double d[3], GT[MN_MAX][3];
  for (int b = 0; b < 3; ++b) d[b] = x[b] - pt->m_center[cell * 3 + b];
  taylor_grad(sim->degree, d, GT);
  const double *S = pt->m_S + cell * N * N;
  for (int n = 0; n < N; ++n)
    for (int b = 0; b < 3; ++b) {
      double s = 0;
      for (int m = 1; m <= n; ++m) s += S[n * N + m] * GT[m][b];
      G[n][b] = s;
    }
}
/* polynomial/projection.hpp:104-106 : GlobalToValue = coeff * basis(x); coeff is [N][K] (column layout) */
static void modal_value_with(const oracle_sim *sim, const double *coeff, const double *phi, double *u) {
  for (int k = 0; k < K; ++k) u[k] = 0;
  for (int n = 0; n < sim->n; ++n)
    for (int k = 0; k < K; ++k) u[k] += coeff[n * K + k] * phi[n];
}
static void modal_value(const oracle_sim *sim, const PartTab *pt, int64_t cell, const double *x, double *u) {
  double phi[MN_MAX];
  modal_basis(sim, pt, cell, x, phi);
  modal_value_with(sim, pt->u + cell * sim->n * K, phi, u);
}

/* Cell + Projection + OrthoNormal construction (part.hpp:215-220, basis/linear.hpp:99-103) and the
 * adjacency lists the limiter walks (part.hpp:1019-1032, 1066-1078: interior faces only, mesh order) */
static int modal_build(oracle_sim *sim, PartTab *pt, const mcfd_mesh *m) {
  int N = sim->n, nq = sim->nq;
  int64_t nc = pt->n_local + pt->n_ghost;
  pt->m_center = malloc(sizeof(double) * nc * 3);
  pt->m_S = calloc(nc * N * N, sizeof(double));
  pt->m_volume = malloc(sizeof(double) * nc);
  for (int64_t c = 0; c < nc; ++c) {
    if (sim->elem == MCFD_CELL_TETRA_4) {  /* geometry/tetrahedron.hpp:43-46 : LocalToGlobal(1/4, 1/4, 1/4) */
      const double *p[4];
      for (int v = 0; v < 4; ++v) p[v] = m->node_xyz + 3 * m->cell_nodes[4 * c + v];
      tetra_map(p, 0.25, 0.25, 0.25, pt->m_center + c * 3);
    } else {
      double xn[8][3], J[3][3], zero[3] = { 0, 0, 0 };
      for (int v = 0; v < 8; ++v) memcpy(xn[v], m->node_xyz + 3 * m->cell_nodes[8 * c + v], 3 * sizeof(double));
      hexa_map(xn, zero, pt->m_center + c * 3, J);  /* geometry/hexahedron.hpp:77-80 */
    }
    double A[MN_MAX][MN_MAX], S[MN_MAX][MN_MAX], vol = 0;
    memset(A, 0, sizeof(A));
    for (int q = 0; q < nq; ++q) {
      double d[3], T[MN_MAX], w = pt->wglob[c * nq + q];
      for (int b = 0; b < 3; ++b) d[b] = pt->xyz[(c * nq + q) * 3 + b] - pt->m_center[c * 3 + b];
      taylor_value(sim->degree, d, T);
      for (int i = 0; i < N; ++i) for (int j = 0; j < N; ++j) A[i][j] += T[i] * T[j] * w;
      vol += w;
    }
    ortho_normalize(N, A, S);
    for (int i = 0; i < N; ++i) for (int j = 0; j <= i; ++j) pt->m_S[c * N * N + i * N + j] = S[i][j];
    pt->m_volume[c] = vol;
  }
  int64_t n_int = pt->n_local_faces + pt->n_ghost_faces;
  pt->adj_ptr = calloc(pt->n_local + 2, sizeof(int64_t));
  for (int64_t f = 0; f < n_int; ++f) {
    pt->adj_ptr[pt->holder[f] + 1]++;
    if (f < pt->n_local_faces) pt->adj_ptr[pt->sharer[f] + 1]++;
  }
  for (int64_t c = 0; c < pt->n_local; ++c) pt->adj_ptr[c + 1] += pt->adj_ptr[c];
  pt->adj_cell = malloc(sizeof(int64_t) * (pt->adj_ptr[pt->n_local] + 1));
  pt->adj_face = malloc(sizeof(int64_t) * (pt->adj_ptr[pt->n_local] + 1));
  int64_t *fill = calloc(pt->n_local + 1, sizeof(int64_t));
  for (int64_t f = 0; f < n_int; ++f) {
    int64_t h = pt->holder[f], s = pt->sharer[f];
    pt->adj_cell[pt->adj_ptr[h] + fill[h]] = s; pt->adj_face[pt->adj_ptr[h] + fill[h]++] = f;
    if (f < pt->n_local_faces) { pt->adj_cell[pt->adj_ptr[s] + fill[s]] = h; pt->adj_face[pt->adj_ptr[s] + fill[s]++] = f; }
  }
  free(fill);
  return 0;
}

/* polynomial/projection.hpp:25-39 : Project -- coeff = int f(x) phi(x)^t */
static void modal_project_function(const oracle_sim *sim, PartTab *pt, int64_t cell, const StateFunc *sf, double t) {
  int N = sim->n, nq = sim->nq;
  double *coeff = pt->u + cell * N * K;
  memset(coeff, 0, sizeof(double) * N * K);
  for (int q = 0; q < nq; ++q) {
    const double *x = pt->xyz + (cell * nq + q) * 3;
    double f[K], phi[MN_MAX], w = pt->wglob[cell * nq + q];
    eval_state(sf, x, t, f);
    modal_basis(sim, pt, cell, x, phi);
    for (int n = 0; n < N; ++n) for (int k = 0; k < K; ++k) { double prod = f[k] * phi[n]; prod *= w; coeff[n * K + k] += prod; }
  }
}

/* dg/general.hpp:50-61 volume term, :69-105 interior / ghost faces, :108-212 boundaries (fem.hpp:185-226 order) */
static void modal_residual_of_part(oracle_sim *sim, PartTab *pt, double *res) {
  int N = sim->n, nq = sim->nq, fq = sim->fq;
  memset(res, 0, sizeof(double) * pt->n_local * N * K);
  if (sim->degree > 0) {
    for (int64_t c = 0; c < pt->n_local; ++c)
      for (int q = 0; q < nq; ++q) {
        const double *x = pt->xyz + (c * nq + q) * 3;
        double u[K], F[3][K], G[MN_MAX][3], w = pt->wglob[c * nq + q];
        modal_value(sim, pt, c, x, u);
        flux_matrix(u, sim->gamma, F);
        modal_basis_grad(sim, pt, c, x, G);
        for (int n = 0; n < N; ++n)
          for (int k = 0; k < K; ++k)
            res[(c * N + n) * K + k] += (F[0][k] * w) * G[n][0] + (F[1][k] * w) * G[n][1] + (F[2][k] * w) * G[n][2];
      }
  }
  int64_t n_int = pt->n_local_faces + pt->n_ghost_faces;
  for (int pass = 0; pass < 3; ++pass) {  /* local faces, boundaries, ghost faces (fem.hpp:173-183) */
    if (pass == 1) {
      static const int kOrder[6] = { MCFD_BC_SOLID_WALL, MCFD_BC_SUPERSONIC_INLET, MCFD_BC_SUPERSONIC_OUTLET,
                                     MCFD_BC_SUBSONIC_INLET, MCFD_BC_SUBSONIC_OUTLET, MCFD_BC_SMART };
      for (int o = 0; o < 6; ++o) for (int g = 0; g < sim->n_groups; ++g) {
        if (sim->group_kind[g] != kOrder[o]) continue;
        for (int64_t b = 0; b < pt->n_boundary_faces; ++b) {
          if (pt->f_group[b] != g) continue;
          int64_t f = n_int + b, h = pt->holder[f];
          for (int p = 0; p < fq; ++p) {
            const double *x = pt->f_xyz + (f * fq + p) * 3;
            double phi[MN_MAX], uh[K], given[K] = { 0, 0, 0, 0, 0 }, flux[K], w = pt->f_w[f * fq + p];
            modal_basis(sim, pt, h, x, phi);
            modal_value_with(sim, pt->u + h * N * K, phi, uh);
            if (kOrder[o] != MCFD_BC_SOLID_WALL && kOrder[o] != MCFD_BC_SUPERSONIC_OUTLET)
              eval_state(&sim->group_func[g], x, sim->t_curr, given);
            boundary_flux(kOrder[o], pt->f_frame + (f * fq + p) * 9, uh, given, sim->gamma, flux);
            for (int n = 0; n < N; ++n) for (int k = 0; k < K; ++k) res[(h * N + n) * K + k] -= (flux[k] * w) * phi[n];
          }
        }
      }
      continue;
    }
    int64_t f0 = pass == 0 ? 0 : pt->n_local_faces, f1 = pass == 0 ? pt->n_local_faces : n_int;
    for (int64_t f = f0; f < f1; ++f) {
      int64_t h = pt->holder[f], s = pt->sharer[f];
      for (int p = 0; p < fq; ++p) {
        const double *x = pt->f_xyz + (f * fq + p) * 3;
        double ph[MN_MAX], ps[MN_MAX], uh[K], us[K], flux[K], w = pt->f_w[f * fq + p];
        modal_basis(sim, pt, h, x, ph); modal_basis(sim, pt, s, x, ps);
        modal_value_with(sim, pt->u + h * N * K, ph, uh);
        modal_value_with(sim, pt->u + s * N * K, ps, us);
        rotated_flux_upwind_kind(sim->riemann, pt->f_frame + (f * fq + p) * 9, uh, us, sim->gamma, pt->f_star_u + f * fq + p, flux);
        for (int n = 0; n < N; ++n) for (int k = 0; k < K; ++k) {
          res[(h * N + n) * K + k] -= (flux[k] * w) * ph[n];
          if (pass == 0) res[(s * N + n) * K + k] += (flux[k] * w) * ps[n];
        }
      }
    }
  }
}

/* ------------------------------------------------------------------------------------------ */
/* limiter/weno.hpp                                                                            */
#define ADJ_MAX 6
/* polynomial/projection.hpp:41-47 : GetAverage on the basis of `cell` */
static void modal_average(const oracle_sim *sim, const PartTab *pt, int64_t cell, const double *coeff, double *avg) {
  double s00 = pt->m_S[cell * sim->n * sim->n];
  for (int k = 0; k < K; ++k) { double v = coeff[k]; v *= s00; avg[k] = v; }
}
/* weno.hpp:20-34 : GetSmoothness of a projection expressed on the basis of cell `me` */
static void modal_smoothness(const oracle_sim *sim, const PartTab *pt, int64_t me, const double *coeff, double *beta) {
  int N = sim->n, nq = sim->nq;
  const double *S = pt->m_S + me * N * N;
  double C[MN_MAX][K], pdv[MN_MAX][K], integral[MN_MAX][K];
  for (int m = 0; m < N; ++m) for (int k = 0; k < K; ++k) {  /* GetCoeffOnTaylorBasis = coeff * S */
    double s = 0;
    for (int n = m; n < N; ++n) s += coeff[n * K + k] * S[n * N + m];
    C[m][k] = s; integral[m][k] = 0;
  }
  for (int q = 0; q < nq; ++q) {
    double d[3], w = pt->wglob[me * nq + q];
    for (int b = 0; b < 3; ++b) d[b] = pt->xyz[(me * nq + q) * 3 + b] - pt->m_center[me * 3 + b];
    taylor_pdv(sim->degree, d, C, pdv);
    for (int m = 0; m < N; ++m) for (int k = 0; k < K; ++k) { double v = pdv[m][k] * pdv[m][k]; v *= w; integral[m][k] += v; }
  }
  double s00 = S[0];
  taylor_smoothness(sim->degree, integral, 1 / (s00 * s00), beta);  /* OrthoNormal::Measure, linear.hpp:122-125 */
}
/* weno.hpp:193-203 (Eigen) = :70-91 (Lazy) : candidates on MY basis: neighbours re-projected over my quadrature
 * points and shifted to my average, then my own projection last.  cand is [(n_adj + 1)][N][K]. */
static int modal_borrow(const oracle_sim *sim, const PartTab *pt, int64_t me, double *cand) {
  int N = sim->n, nq = sim->nq;
  int n_adj = (int)(pt->adj_ptr[me + 1] - pt->adj_ptr[me]);
  double my_avg[K];
  modal_average(sim, pt, me, pt->u + me * N * K, my_avg);
  for (int a = 0; a < n_adj; ++a) {
    int64_t adj = pt->adj_cell[pt->adj_ptr[me] + a];
    double *c = cand + a * N * K;
    memset(c, 0, sizeof(double) * N * K);
    for (int q = 0; q < nq; ++q) {  /* projection.hpp:25-39 with func = the neighbour's polynomial */
      const double *x = pt->xyz + (me * nq + q) * 3;
      double f[K], phi[MN_MAX], w = pt->wglob[me * nq + q];
      modal_value(sim, pt, adj, x, f);
      modal_basis(sim, pt, me, x, phi);
      for (int n = 0; n < N; ++n) for (int k = 0; k < K; ++k) { double prod = f[k] * phi[n]; prod *= w; c[n * K + k] += prod; }
    }
    double avg[K], s00 = pt->m_S[me * N * N];
    modal_average(sim, pt, me, c, avg);
    for (int k = 0; k < K; ++k) { double off = my_avg[k] - avg[k]; off /= s00; c[k] += off; }  /* projection.hpp:282-286 */
  }
  memcpy(cand + n_adj * N * K, pt->u + me * N * K, sizeof(double) * N * K);
  return n_adj;
}
/* weno.hpp:216-241 shared by Lazy and Eigen: nonlinear weights from the smoothness of each candidate, then the
 * weighted sum into out[N][K] */
static void modal_weighted_sum(const oracle_sim *sim, const PartTab *pt, int64_t me, int n_adj, const double *cand,
                               double *out) {
  int N = sim->n;
  double weights[ADJ_MAX + 1][K], sum[K] = { 0, 0, 0, 0, 0 };
  for (int i = 0; i <= n_adj; ++i) {
    double beta[K];
    modal_smoothness(sim, pt, me, cand + i * N * K, beta);
    for (int k = 0; k < K; ++k) {
      double w = sim->weno_w0;
      if (i == n_adj) { w *= -n_adj; w += 1.0; }
      double b = beta[k] + sim->weno_eps;
      b *= b;
      weights[i][k] = w / b;
    }
  }
  for (int i = 0; i <= n_adj; ++i) for (int k = 0; k < K; ++k) sum[k] += weights[i][k];
  for (int i = 0; i <= n_adj; ++i) for (int k = 0; k < K; ++k) weights[i][k] /= sum[k];
  for (int n = 0; n < N; ++n) for (int k = 0; k < K; ++k) out[n * K + k] = cand[(n_adj * N + n) * K + k] * weights[n_adj][k];
  for (int i = 0; i < n_adj; ++i)
    for (int n = 0; n < N; ++n) for (int k = 0; k < K; ++k) out[n * K + k] += cand[(i * N + n) * K + k] * weights[i][k];
}
/* riemann/euler/eigen.hpp:27-56 : L and R at the conservative state `big_u` in the frame (nu, mu, pi) */
static void eigen_matrices(const double *big_u, const double *fr, double gamma, double L[K][K], double R[K][K]) {
  Prim s = cons_to_prim(big_u, gamma);
  const double *nu = fr, *mu = fr + 3, *pi = fr + 6;
  double ux = s.u, uy = s.v, uz = s.w;
  double u_nu = ux * nu[0] + uy * nu[1] + uz * nu[2], u_mu = ux * mu[0] + uy * mu[1] + uz * mu[2];
  double u_pi = ux * pi[0] + uy * pi[1] + uz * pi[2];
  double a = speed_of_sound(&s, gamma);
  double ek = (ux * ux + uy * uy + uz * uz) / 2, e0 = a * a * (1 / (gamma - 1)), h0 = e0 + ek;
  double vel[3] = { ux, uy, uz };
  R[0][0] = 1; R[0][1] = 1; R[0][2] = 0; R[0][3] = 0; R[0][4] = 1;
  for (int d = 0; d < 3; ++d) {
    R[1 + d][0] = vel[d] - a * nu[d]; R[1 + d][1] = vel[d]; R[1 + d][2] = mu[d]; R[1 + d][3] = pi[d];
    R[1 + d][4] = vel[d] + a * nu[d];
  }
  R[4][0] = h0 - a * u_nu; R[4][1] = ek; R[4][2] = u_mu; R[4][3] = u_pi; R[4][4] = h0 + a * u_nu;
  double b1 = 1 / e0, b2 = b1 * ek;
  L[0][0] = (b2 + u_nu / a) / 2; L[1][0] = 1 - b2; L[2][0] = -u_mu; L[3][0] = -u_pi; L[4][0] = (b2 - u_nu / a) / 2;
  for (int d = 0; d < 3; ++d) {
    L[0][1 + d] = -(b1 * vel[d] + nu[d] / a) / 2; L[1][1 + d] = b1 * vel[d]; L[2][1 + d] = mu[d]; L[3][1 + d] = pi[d];
    L[4][1 + d] = -(b1 * vel[d] - nu[d] / a) / 2;
  }
  L[0][4] = b1 / 2; L[1][4] = -b1; L[2][4] = 0; L[3][4] = 0; L[4][4] = b1 / 2;
}
static void left_multiply(int N, double M[K][K], const double *in, double *out) {  /* projection.hpp:252-256 */
  for (int n = 0; n < N; ++n)
    for (int r = 0; r < K; ++r) {
      double s = 0;
      for (int k = 0; k < K; ++k) s += M[r][k] * in[n * K + k];
      out[n * K + r] = s;
    }
}
/* weno.hpp:152-180 : Eigen::IsNotSmooth (Lazy: always true, :58-60) */
static int modal_is_not_smooth(const oracle_sim *sim, const PartTab *pt, int64_t me) {
  if (sim->limiter != MCFD_LIMITER_EIGEN) return 1;
  int N = sim->n;
  static const int comps[2] = { 0, K - 1 };
  double max_abs[K], sum_abs[K] = { 0, 0, 0, 0, 0 }, mine[K];
  modal_average(sim, pt, me, pt->u + me * N * K, max_abs);
  for (int i = 0; i < 2; ++i) max_abs[comps[i]] = fmax(1e-9, fabs(max_abs[comps[i]]));
  const double *center = pt->m_center + me * 3;
  modal_value(sim, pt, me, center, mine);
  int n_adj = (int)(pt->adj_ptr[me + 1] - pt->adj_ptr[me]);
  for (int a = 0; a < n_adj; ++a) {
    int64_t adj = pt->adj_cell[pt->adj_ptr[me] + a];
    double theirs[K], avg[K];
    modal_value(sim, pt, adj, center, theirs);
    modal_average(sim, pt, adj, pt->u + adj * N * K, avg);
    for (int i = 0; i < 2; ++i) {
      int c = comps[i];
      sum_abs[c] += fabs(mine[c] - theirs[c]);
      max_abs[c] = fmax(max_abs[c], fabs(avg[c]));
    }
  }
  double divisor = pow(pt->m_volume[me], (sim->degree + 1.0) / 2.0 / 3);
  divisor *= n_adj;
  double reference = sim->degree < 3 ? 1.0 : 3.0;
  for (int i = 0; i < 2; ++i) {
    int c = comps[i];
    double smoothness = sum_abs[c] / max_abs[c] / divisor;
    if (smoothness > reference) return 1;
  }
  return 0;
}
/* weno.hpp:61-68 (Lazy) and :181-186, 204-263 (Eigen) : the limited projection of one cell -> out[N][K] */
static void modal_limit_cell(const oracle_sim *sim, const PartTab *pt, int64_t me, double *out) {
  int N = sim->n;
  double cand[(ADJ_MAX + 1) * MN_MAX * K];
  int n_adj = modal_borrow(sim, pt, me, cand);
  if (sim->limiter == MCFD_LIMITER_LAZY) {
    modal_weighted_sum(sim, pt, me, n_adj, cand, out);
    return;
  }
  double big_u[K], total_volume = 0.0;
  modal_average(sim, pt, me, pt->u + me * N * K, big_u);
  memset(out, 0, sizeof(double) * N * K);
  for (int a = 0; a < n_adj; ++a) {  /* ReconstructOnFace, weno.hpp:208-250 */
    int64_t f = pt->adj_face[pt->adj_ptr[me] + a], adj = pt->adj_cell[pt->adj_ptr[me] + a];
    double L[K][K], R[K][K], rotated[(ADJ_MAX + 1) * MN_MAX * K], mixed[MN_MAX * K], back[MN_MAX * K];
    eigen_matrices(big_u, pt->f_frame + (f * sim->fq) * 9, sim->gamma, L, R);  /* adj_face.riemann(0) */
    for (int i = 0; i <= n_adj; ++i) left_multiply(N, L, cand + i * N * K, rotated + i * N * K);
    modal_weighted_sum(sim, pt, me, n_adj, rotated, mixed);
    left_multiply(N, R, mixed, back);
    double adj_volume = pt->m_volume[adj];
    for (int i = 0; i < N * K; ++i) { double v = back[i] * adj_volume; out[i] += v; }
    total_volume += adj_volume;
  }
  for (int i = 0; i < N * K; ++i) out[i] /= total_volume;
}
/* part.hpp:1318-1334 : detect, rebuild all troubled cells of the set from the OLD data, then write back */
static void modal_reconstruct_set(oracle_sim *sim, PartTab *pt, int64_t c0, int64_t c1) {
  int N = sim->n;
  int64_t n_troubled = 0, *list = malloc(sizeof(int64_t) * (c1 - c0 + 1));
  for (int64_t c = c0; c < c1; ++c) if (modal_is_not_smooth(sim, pt, c)) list[n_troubled++] = c;
  double *fresh = malloc(sizeof(double) * (n_troubled + 1) * N * K);
  for (int64_t i = 0; i < n_troubled; ++i) modal_limit_cell(sim, pt, list[i], fresh + i * N * K);
  for (int64_t i = 0; i < n_troubled; ++i) memcpy(pt->u + list[i] * N * K, fresh + i * N * K, sizeof(double) * N * K);
  pt->n_troubled += n_troubled;
  free(list); free(fresh);
}
/* part.hpp:1304-1315 : Part::Reconstruct on every part; the halo carries the coefficients packed BEFORE any
 * cell of this pass was limited (ShareGhostCellCoeffs comes first on every rank) */
static void modal_reconstruct(oracle_sim *sim) {
  if (sim->scheme != MCFD_SCHEME_DG_MODAL || sim->limiter == MCFD_LIMITER_NONE) return;
  exchange_ghosts(sim);
  for (int r = 0; r < sim->n_parts; ++r) {
    PartTab *pt = &sim->parts[r];
    modal_reconstruct_set(sim, pt, 0, pt->n_inner);
    modal_reconstruct_set(sim, pt, pt->n_inner, pt->n_local);
  }
}
