// box_mesh.cpp -- host-side generator of partitioned hexahedral box meshes in the layout of
// `mcfd_mesh` (include/minicfd_b200.h).
//
// It stands where the reference has gmsh + mesh::Shuffler::PartitionAndShuffle + Part's
// constructor (demo/euler/sourceless.cpp:61-71): the named configs are "synthetic box meshes",
// and METIS/gmsh/CGNS do not exist on this box, so the partition is a
// deterministic block split.  What IS kept from the reference is the per-part ordering contract:
//   cells : inner cells, then inter cells (part.hpp:818-835), each ascending global id;
//           ghost cells grouped by neighbour part, ascending global id (part.hpp:966-986);
//   faces : local | ghost | boundary, the normal pointing holder -> sharer / outward
//           (part.hpp:1019-1078, 1521-1522); ghost faces are held by the local cell;
//   nodes : HEXA_8 in CGNS order (geometry/hexahedron.hpp:345-352).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <new>
#include <unordered_map>
#include <vector>

#include "minicfd_b200.h"
#include "part_storage.hpp"

namespace {

struct Range { int64_t lo, hi; };

Range BlockRange(int64_t n, int parts, int b) {
  // contiguous split, remainder spread over the first blocks
  int64_t q = n / parts, r = n % parts;
  int64_t lo = b * q + std::min<int64_t>(b, r);
  return { lo, lo + q + (b < r ? 1 : 0) };
}

uint64_t SplitMix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

double Jitter(uint64_t node_gid, int axis) {
  uint64_t h = SplitMix64(node_gid * 3 + axis + 0x1234567ull);
  return (double)(h >> 11) / (double)(1ull << 53) * 2.0 - 1.0;  // [-1, 1)
}

}  // namespace

extern "C" {

int mcfd_box_mesh_create(const int64_t n[3], const double lo[3], const double hi[3], const int32_t parts[3],
                         int32_t rank, double distortion, mcfd_box_mesh **out) {
  return mcfd_box_mesh_create_ex(n, lo, hi, parts, rank, distortion, nullptr, nullptr, out);
}

int mcfd_box_mesh_create_ex(const int64_t n[3], const double lo[3], const double hi[3], const int32_t parts[3],
                            int32_t rank, double distortion, const int64_t hole_lo[3], const int64_t hole_hi[3],
                            mcfd_box_mesh **out) {
  if (!n || !lo || !hi || !parts || !out) return MCFD_ERR_ARG;
  const bool has_hole = hole_lo && hole_hi;
  for (int d = 0; d < 3; ++d) {
    if (n[d] < 1 || parts[d] < 1 || parts[d] > n[d] || !(hi[d] > lo[d])) return MCFD_ERR_ARG;
  }
  int world = parts[0] * parts[1] * parts[2];
  if (rank < 0 || rank >= world) return MCFD_ERR_ARG;
  auto *bm = new (std::nothrow) mcfd_box_mesh();
  if (!bm) return MCFD_ERR_ARG;

  const int64_t nx = n[0], ny = n[1], nz = n[2];
  int b[3] = { rank % parts[0], (rank / parts[0]) % parts[1], rank / (parts[0] * parts[1]) };
  Range r[3];
  for (int d = 0; d < 3; ++d) r[d] = BlockRange(n[d], parts[d], b[d]);
  auto cell_gid = [&](int64_t i, int64_t j, int64_t k) { return i + nx * (j + ny * k); };
  auto owner = [&](int64_t i, int64_t j, int64_t k) -> int {
    int64_t ijk[3] = { i, j, k };
    int bb[3];
    for (int d = 0; d < 3; ++d) {
      int lo_b = 0;
      for (int p = 0; p < parts[d]; ++p) {
        Range rr = BlockRange(n[d], parts[d], p);
        if (ijk[d] >= rr.lo && ijk[d] < rr.hi) { lo_b = p; break; }
      }
      bb[d] = lo_b;
    }
    return bb[0] + parts[0] * (bb[1] + parts[1] * bb[2]);
  };
  auto is_local = [&](int64_t i, int64_t j, int64_t k) {
    return i >= r[0].lo && i < r[0].hi && j >= r[1].lo && j < r[1].hi && k >= r[2].lo && k < r[2].hi;
  };
  auto in_hole = [&](int64_t i, int64_t j, int64_t k) {
    return has_hole && i >= hole_lo[0] && i < hole_hi[0] && j >= hole_lo[1] && j < hole_hi[1] && k >= hole_lo[2] && k < hole_hi[2];
  };
  auto in_domain = [&](int64_t i, int64_t j, int64_t k) {
    return i >= 0 && i < nx && j >= 0 && j < ny && k >= 0 && k < nz && !in_hole(i, j, k);
  };
  static const int kDir[6][3] = { {-1,0,0}, {1,0,0}, {0,-1,0}, {0,1,0}, {0,0,-1}, {0,0,1} };

  // ---- cells: inner | inter, ascending global id ----
  std::vector<int64_t> inner, inter;
  for (int64_t k = r[2].lo; k < r[2].hi; ++k)
    for (int64_t j = r[1].lo; j < r[1].hi; ++j)
      for (int64_t i = r[0].lo; i < r[0].hi; ++i) {
        if (in_hole(i, j, k)) continue;
        bool touches = false;
        for (auto &d : kDir) {
          int64_t a = i + d[0], bq = j + d[1], c = k + d[2];
          if (in_domain(a, bq, c) && !is_local(a, bq, c)) touches = true;
        }
        (touches ? inter : inner).push_back(cell_gid(i, j, k));
      }
  // ---- ghosts grouped by owning part ----
  std::map<int, std::vector<int64_t>> ghosts_by_part, sends_by_part;
  for (int64_t g : inter) {
    int64_t i = g % nx, j = (g / nx) % ny, k = g / (nx * ny);
    for (auto &d : kDir) {
      int64_t a = i + d[0], bq = j + d[1], c = k + d[2];
      if (in_domain(a, bq, c) && !is_local(a, bq, c)) {
        int p = owner(a, bq, c);
        ghosts_by_part[p].push_back(cell_gid(a, bq, c));
        sends_by_part[p].push_back(g);
      }
    }
  }
  std::unordered_map<int64_t, int64_t> gid_to_index;
  gid_to_index.reserve((inner.size() + inter.size()) * 2);
  for (int64_t g : inner) { gid_to_index[g] = (int64_t)bm->cell_gid.size(); bm->cell_gid.push_back(g); }
  for (int64_t g : inter) { gid_to_index[g] = (int64_t)bm->cell_gid.size(); bm->cell_gid.push_back(g); }
  int64_t n_local = (int64_t)bm->cell_gid.size();
  bm->send_ptr.push_back(0); bm->recv_ptr.push_back(0);
  for (auto &[p, list] : ghosts_by_part) {
    std::sort(list.begin(), list.end());
    list.erase(std::unique(list.begin(), list.end()), list.end());
    auto &send = sends_by_part[p];
    std::sort(send.begin(), send.end());
    send.erase(std::unique(send.begin(), send.end()), send.end());
    bm->nbr_rank.push_back(p);
    for (int64_t g : list) { gid_to_index[g] = (int64_t)bm->cell_gid.size(); bm->cell_gid.push_back(g); }
    for (int64_t g : send) bm->send_cells.push_back(gid_to_index.at(g));
    bm->recv_ptr.push_back((int64_t)bm->cell_gid.size() - n_local);
    bm->send_ptr.push_back((int64_t)bm->send_cells.size());
  }
  int64_t n_cells = (int64_t)bm->cell_gid.size();

  // ---- nodes: numbered in order of first use ----
  const int64_t mx = nx + 1, my = ny + 1;
  std::unordered_map<int64_t, int64_t> node_index;
  node_index.reserve(n_cells * 2);
  double h[3] = { (hi[0] - lo[0]) / nx, (hi[1] - lo[1]) / ny, (hi[2] - lo[2]) / nz };
  auto node_of = [&](int64_t i, int64_t j, int64_t k) -> int64_t {
    int64_t gid = i + mx * (j + my * k);
    auto it = node_index.find(gid);
    if (it != node_index.end()) return it->second;
    int64_t id = (int64_t)node_index.size();
    node_index.emplace(gid, id);
    double x[3] = { lo[0] + h[0] * i, lo[1] + h[1] * j, lo[2] + h[2] * k };
    if (i == nx) x[0] = hi[0];
    if (j == ny) x[1] = hi[1];
    if (k == nz) x[2] = hi[2];
    bool interior = i > 0 && i < nx && j > 0 && j < ny && k > 0 && k < nz;
    if (distortion != 0.0 && interior) {
      for (int d = 0; d < 3; ++d) x[d] += distortion * h[d] * Jitter((uint64_t)gid, d);
    }
    bm->node_xyz.insert(bm->node_xyz.end(), x, x + 3);
    return id;
  };
  static const int kCorner[8][3] = { {0,0,0},{1,0,0},{1,1,0},{0,1,0},{0,0,1},{1,0,1},{1,1,1},{0,1,1} };
  bm->cell_nodes.resize(n_cells * 8);
  for (int64_t c = 0; c < n_cells; ++c) {
    int64_t g = bm->cell_gid[c];
    int64_t i = g % nx, j = (g / nx) % ny, k = g / (nx * ny);
    for (int v = 0; v < 8; ++v) bm->cell_nodes[8 * c + v] = node_of(i + kCorner[v][0], j + kCorner[v][1], k + kCorner[v][2]);
  }

  // ---- faces ----
  // corner ids (into the cell's 8 nodes) of the face on side s of a cell, s as in kDir
  static const int kFaceCorners[6][4] = { {0,4,7,3}, {1,2,6,5}, {0,1,5,4}, {2,3,7,6}, {0,3,2,1}, {4,5,6,7} };
  struct FaceRec { int64_t holder, sharer; int side; int group; };
  std::vector<FaceRec> local_faces, ghost_faces, bound_faces;
  for (int64_t c = 0; c < n_local; ++c) {
    int64_t g = bm->cell_gid[c];
    int64_t i = g % nx, j = (g / nx) % ny, k = g / (nx * ny);
    for (int s = 0; s < 6; ++s) {
      int64_t a = i + kDir[s][0], bq = j + kDir[s][1], cc = k + kDir[s][2];
      if (!in_domain(a, bq, cc)) {
        bound_faces.push_back({ c, -1, s, in_hole(a, bq, cc) ? 6 : s });
      } else if (is_local(a, bq, cc)) {
        if (s & 1) local_faces.push_back({ c, gid_to_index.at(cell_gid(a, bq, cc)), s, -1 });  // + side only: holder = lower gid
      } else {
        ghost_faces.push_back({ c, gid_to_index.at(cell_gid(a, bq, cc)), s, -1 });
      }
    }
  }
  std::stable_sort(bound_faces.begin(), bound_faces.end(), [](auto &x, auto &y) { return x.group < y.group; });
  auto emit = [&](std::vector<FaceRec> const &faces) {
    for (auto &f : faces) {
      bm->face_holder.push_back(f.holder);
      bm->face_sharer.push_back(f.sharer);
      for (int v = 0; v < 4; ++v) bm->face_nodes.push_back(bm->cell_nodes[8 * f.holder + kFaceCorners[f.side][v]]);
      if (f.group >= 0) bm->bface_group.push_back(f.group);
    }
  };
  emit(local_faces); emit(ghost_faces); emit(bound_faces);

  mcfd_mesh &m = bm->view;
  m.n_nodes = (int64_t)bm->node_xyz.size() / 3;
  m.node_xyz = bm->node_xyz.data();
  m.n_local_cells = n_local;
  m.n_inner_cells = (int64_t)inner.size();
  m.n_ghost_cells = n_cells - n_local;
  m.cell_nodes = bm->cell_nodes.data();
  m.n_local_faces = (int64_t)local_faces.size();
  m.n_ghost_faces = (int64_t)ghost_faces.size();
  m.n_boundary_faces = (int64_t)bound_faces.size();
  m.face_holder = bm->face_holder.data();
  m.face_sharer = bm->face_sharer.data();
  m.face_nodes = bm->face_nodes.data();
  m.bface_group = bm->bface_group.data();
  m.n_bc_groups = has_hole ? 7 : 6;
  m.n_neighbours = (int32_t)bm->nbr_rank.size();
  m.nbr_rank = bm->nbr_rank.data();
  m.send_ptr = bm->send_ptr.data();
  m.send_cells = bm->send_cells.data();
  m.recv_ptr = bm->recv_ptr.data();
  *out = bm;
  return MCFD_OK;
}

// ---- six tetrahedra per hexahedron ---------------------------------------------------------------------------
// Kuhn's split along the diagonal corner 0 -> corner 6: one tetrahedron per order in which the three local axes are
// walked from (0,0,0) to (1,1,1).  Every quadrilateral is cut along the diagonal between its lowest and highest
// corner in that lattice, so hexahedra whose local axes agree (every box mesh of this file) produce matching
// triangles.  Nodes are ordered for CGNS TETRA_4 (the base 0-1-2 seen counter-clockwise from OUTSIDE, i.e.
// (p1 - p0) x (p2 - p0) . (p3 - p0) > 0), which is what makes geometry::Tetrahedron4's face table
// (geometry/tetrahedron.hpp:145-150) produce outward normals.
int mcfd_box_mesh_split_tetra(const mcfd_box_mesh *hexa, mcfd_box_mesh **out) {
  if (!hexa || !out || hexa->view.cell_type != MCFD_CELL_HEXA_8) return MCFD_ERR_ARG;
  const mcfd_mesh &h = hexa->view;
  auto *bm = new (std::nothrow) mcfd_box_mesh();
  if (!bm) return MCFD_ERR_ARG;
  bm->node_xyz = hexa->node_xyz;
  // lattice corner (x, y, z) -> HEXA_8 corner id
  static const int kCornerOf[2][2][2] = { { {0, 4}, {3, 7} }, { {1, 5}, {2, 6} } };
  static const int kPerm[6][3] = { {0,1,2}, {0,2,1}, {1,0,2}, {1,2,0}, {2,0,1}, {2,1,0} };
  const int64_t n_hex = h.n_local_cells + h.n_ghost_cells;
  bm->cell_nodes.resize(size_t(n_hex) * 24);
  bm->cell_gid.resize(size_t(n_hex) * 6);
  for (int64_t c = 0; c < n_hex; ++c) {
    const int64_t *hn = h.cell_nodes + 8 * c;
    for (int t = 0; t < 6; ++t) {
      int at[3] = { 0, 0, 0 };
      int64_t v[4];
      v[0] = hn[kCornerOf[0][0][0]];
      for (int step = 0; step < 3; ++step) { at[kPerm[t][step]] = 1; v[step + 1] = hn[kCornerOf[at[0]][at[1]][at[2]]]; }
      const double *p0 = &bm->node_xyz[3 * v[0]], *p1 = &bm->node_xyz[3 * v[1]], *p2 = &bm->node_xyz[3 * v[2]], *p3 = &bm->node_xyz[3 * v[3]];
      double a[3], b[3], d[3];
      for (int k = 0; k < 3; ++k) { a[k] = p1[k] - p0[k]; b[k] = p2[k] - p0[k]; d[k] = p3[k] - p0[k]; }
      const double vol6 = (a[1] * b[2] - a[2] * b[1]) * d[0] + (a[2] * b[0] - a[0] * b[2]) * d[1] + (a[0] * b[1] - a[1] * b[0]) * d[2];
      if (vol6 < 0) std::swap(v[1], v[2]);
      for (int k = 0; k < 4; ++k) bm->cell_nodes[(size_t(c) * 6 + t) * 4 + k] = v[k];
      bm->cell_gid[size_t(c) * 6 + t] = hexa->cell_gid[c] * 6 + t;
    }
  }
  const int64_t n_local = h.n_local_cells * 6, n_cells = n_hex * 6;
  // triangles by their sorted node ids -> the (at most two) tetrahedra on them
  struct Key { int64_t a, b, c; bool operator==(const Key &o) const { return a == o.a && b == o.b && c == o.c; } };
  struct KeyHash { size_t operator()(const Key &k) const { return size_t(SplitMix64(uint64_t(k.a) * 0x9E3779B1ull ^ SplitMix64(uint64_t(k.b) ^ SplitMix64(uint64_t(k.c))))); } };
  static const int kTetFace[4][3] = { {0,2,1}, {0,1,3}, {0,3,2}, {1,2,3} };  // geometry/tetrahedron.hpp:145-150
  auto key_of = [&](int64_t cell, int side) {
    int64_t v[3];
    for (int k = 0; k < 3; ++k) v[k] = bm->cell_nodes[size_t(cell) * 4 + kTetFace[side][k]];
    std::sort(v, v + 3);
    return Key{ v[0], v[1], v[2] };
  };
  std::unordered_map<Key, std::pair<int64_t, int64_t>, KeyHash> tri;
  tri.reserve(size_t(n_cells) * 2);
  for (int64_t c = 0; c < n_cells; ++c)
    for (int s = 0; s < 4; ++s) {
      auto it = tri.find(key_of(c, s));
      if (it == tri.end()) tri.emplace(key_of(c, s), std::make_pair(c, int64_t(-1)));
      else it->second.second = c;
    }
  // group of a boundary triangle = group of the boundary quadrilateral of its hexahedron that holds its nodes
  struct Key4 { int64_t v[4]; bool operator==(const Key4 &o) const { return !std::memcmp(v, o.v, sizeof(v)); } };
  struct Key4Hash { size_t operator()(const Key4 &k) const { return size_t(SplitMix64(uint64_t(k.v[0]) ^ SplitMix64(uint64_t(k.v[1]) ^ SplitMix64(uint64_t(k.v[2]) ^ SplitMix64(uint64_t(k.v[3])))))); } };
  std::unordered_map<Key4, int, Key4Hash> quad_group;
  const int64_t f_bnd = h.n_local_faces + h.n_ghost_faces;
  for (int64_t b = 0; b < h.n_boundary_faces; ++b) {
    Key4 k;
    std::copy_n(h.face_nodes + 4 * (f_bnd + b), 4, k.v);
    std::sort(k.v, k.v + 4);
    quad_group[k] = h.bface_group ? h.bface_group[b] : 0;
  }
  static const int kHexFace[6][4] = { {0,3,2,1}, {0,1,5,4}, {1,2,6,5}, {2,3,7,6}, {0,4,7,3}, {4,5,6,7} };
  struct FaceRec { int64_t holder, sharer; int side, group; };
  std::vector<FaceRec> local_faces, ghost_faces, bound_faces;
  for (int64_t c = 0; c < n_local; ++c)
    for (int s = 0; s < 4; ++s) {
      const Key k = key_of(c, s);
      const auto &pr = tri.at(k);
      const int64_t other = pr.first == c ? pr.second : pr.first;
      if (other < 0) {
        const int64_t *hn = h.cell_nodes + 8 * (c / 6);
        int group = -1;
        for (int f = 0; f < 6 && group < 0; ++f) {
          int hit = 0;
          for (int v = 0; v < 4; ++v) { const int64_t id = hn[kHexFace[f][v]]; if (id == k.a || id == k.b || id == k.c) ++hit; }
          if (hit == 3) {
            Key4 q;
            for (int v = 0; v < 4; ++v) q.v[v] = hn[kHexFace[f][v]];
            std::sort(q.v, q.v + 4);
            auto it = quad_group.find(q);
            if (it != quad_group.end()) group = it->second;
          }
        }
        if (group < 0) { delete bm; return MCFD_ERR_ARG; }
        bound_faces.push_back({ c, -1, s, group });
      } else if (other < n_local) {
        if (other > c) local_faces.push_back({ c, other, s, -1 });
      } else {
        ghost_faces.push_back({ c, other, s, -1 });
      }
    }
  std::stable_sort(bound_faces.begin(), bound_faces.end(), [](auto &x, auto &y) { return x.group < y.group; });
  auto emit = [&](std::vector<FaceRec> const &faces) {
    for (auto &f : faces) {
      bm->face_holder.push_back(f.holder);
      bm->face_sharer.push_back(f.sharer);
      for (int v = 0; v < 3; ++v) bm->face_nodes.push_back(bm->cell_nodes[size_t(f.holder) * 4 + kTetFace[f.side][v]]);
      if (f.group >= 0) bm->bface_group.push_back(f.group);
    }
  };
  emit(local_faces); emit(ghost_faces); emit(bound_faces);
  bm->nbr_rank = hexa->nbr_rank;
  bm->send_ptr.push_back(0); bm->recv_ptr.push_back(0);
  for (int nb = 0; nb < h.n_neighbours; ++nb) {
    for (int64_t i = h.send_ptr[nb]; i < h.send_ptr[nb + 1]; ++i)
      for (int t = 0; t < 6; ++t) bm->send_cells.push_back(h.send_cells[i] * 6 + t);
    bm->send_ptr.push_back(int64_t(bm->send_cells.size()));
    bm->recv_ptr.push_back(h.recv_ptr[nb + 1] * 6);
  }
  mcfd_mesh &m = bm->view;
  m = h;
  m.cell_type = MCFD_CELL_TETRA_4;
  m.node_xyz = bm->node_xyz.data();
  m.n_local_cells = n_local; m.n_inner_cells = h.n_inner_cells * 6; m.n_ghost_cells = h.n_ghost_cells * 6;
  m.cell_nodes = bm->cell_nodes.data();
  m.n_local_faces = int64_t(local_faces.size()); m.n_ghost_faces = int64_t(ghost_faces.size());
  m.n_boundary_faces = int64_t(bound_faces.size());
  m.face_holder = bm->face_holder.data(); m.face_sharer = bm->face_sharer.data(); m.face_nodes = bm->face_nodes.data();
  m.bface_group = bm->bface_group.data();
  m.nbr_rank = bm->nbr_rank.data();
  m.send_ptr = bm->send_ptr.data(); m.send_cells = bm->send_cells.data(); m.recv_ptr = bm->recv_ptr.data();
  *out = bm;
  return MCFD_OK;
}

const mcfd_mesh *mcfd_box_mesh_view(const mcfd_box_mesh *bm) { return bm ? &bm->view : nullptr; }
const int64_t *mcfd_box_mesh_cell_global_ids(const mcfd_box_mesh *bm) { return bm ? bm->cell_gid.data() : nullptr; }
void mcfd_box_mesh_destroy(mcfd_box_mesh *bm) { delete bm; }

}  // extern "C"
