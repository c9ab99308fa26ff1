// partition.cpp -- the partition + ghost-index pipeline for general hexahedral meshes (host only).
//
// Replaces, for a given cell -> part vector (paths relative to /root/reference/include/mini):
//   mesh/metis.hpp:191-197      Mesh::GetDualGraph = METIS_MeshToDual (METIS 5.x libmetis/mesh.c, restated; the
//                               library is an un-vendored submodule, parity of the neighbour ORDER is unpinned)
//   mesh/metis.hpp:316-337      GetNodeParts
//   mesh/shuffler.hpp:36-104    GetNewOrder, ShuffleData, ShuffleConnectivity
//   mesh/shuffler.hpp:150-236   Shuffler::FillFaceParts
//   mesh/shuffler.hpp:314-506   Shuffler::WritePartitionInfo (the per-part text)
//   mesh/part.hpp:761-1079, 1440-1533   Part's build order of cells, ghost cells, faces and send / recv lists
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <new>
#include <numeric>
#include <set>
#include <string>
#include <unordered_map>
#include <vector>

#include "minicfd_b200.h"
#include "part_storage.hpp"

struct mcfd_partitioner {
  // inputs (copied)
  int64_t n_nodes = 0, n_cells = 0, n_bfaces = 0;
  int32_t n_parts = 0, n_groups = 0;
  std::vector<double> xyz;
  std::vector<int64_t> cell_nodes, bface_nodes;
  std::vector<int32_t> bface_group, cell_parts;
  // derived
  std::vector<int64_t> range, index;          // dual graph (CSR)
  std::vector<int32_t> node_parts, face_parts;
  std::vector<int64_t> face_owner;            // cell that uses all 4 nodes of a boundary face
  std::vector<int32_t> cells_new_to_old, cells_old_to_new, nodes_new_to_old, nodes_old_to_new;
};

namespace {

// METIS_MeshToDual for one element type: node -> element lists in ascending element order, then per element the
// neighbours in order of first discovery, kept if they share at least n_common nodes (or all but one of either).
void MeshToDual(int64_t n_cells, int64_t n_nodes, int npe, const int64_t *eind, int n_common,
                std::vector<int64_t> *range, std::vector<int64_t> *index) {
  std::vector<int64_t> nptr(n_nodes + 1, 0), nind(size_t(n_cells) * npe);
  for (int64_t i = 0; i < n_cells * npe; ++i) nptr[eind[i] + 1]++;
  for (int64_t i = 0; i < n_nodes; ++i) nptr[i + 1] += nptr[i];
  {
    std::vector<int64_t> fill(nptr.begin(), nptr.end() - 1);
    for (int64_t e = 0; e < n_cells; ++e)
      for (int v = 0; v < npe; ++v) nind[fill[eind[e * npe + v]]++] = e;
  }
  std::vector<int32_t> marker(n_cells, 0);
  std::vector<int64_t> nbrs;
  range->assign(n_cells + 1, 0);
  index->clear();
  for (int64_t q = 0; q < n_cells; ++q) {
    nbrs.clear();
    for (int v = 0; v < npe; ++v) {
      const int64_t node = eind[q * npe + v];
      for (int64_t j = nptr[node]; j < nptr[node + 1]; ++j) {
        const int64_t e = nind[j];
        if (marker[e] == 0) nbrs.push_back(e);
        marker[e]++;
      }
    }
    marker[q] = 0;  // the element itself is not its neighbour
    for (int64_t e : nbrs) {
      const int overlap = marker[e];
      if (overlap >= n_common || overlap >= npe - 1) index->push_back(e);
      marker[e] = 0;
    }
    (*range)[q + 1] = int64_t(index->size());
  }
}

template <class T>
void NewOrder(const T *parts, int64_t n, std::vector<int32_t> *new_to_old, std::vector<int32_t> *old_to_new) {
  new_to_old->resize(n);
  old_to_new->resize(n);
  std::iota(new_to_old->begin(), new_to_old->end(), 0);
  // shuffler.hpp:41-44 : by part, ties by the old index
  std::sort(new_to_old->begin(), new_to_old->end(), [parts](int32_t l, int32_t r) {
    return parts[l] < parts[r] || (parts[l] == parts[r] && l < r);
  });
  for (int64_t i = 0; i < n; ++i) (*old_to_new)[(*new_to_old)[i]] = int32_t(i);
}

}  // namespace

extern "C" {

int mcfd_get_new_order(const int32_t *parts, int32_t n, int32_t *new_to_old, int32_t *old_to_new) {
  if (!parts || n < 0 || !new_to_old || !old_to_new) return MCFD_ERR_ARG;
  std::vector<int32_t> a, b;
  NewOrder(parts, n, &a, &b);
  std::copy(a.begin(), a.end(), new_to_old);
  std::copy(b.begin(), b.end(), old_to_new);
  return MCFD_OK;
}

// shuffler.hpp:82-104 : renumber the (1-based) node ids, then move whole cells to their new places
int mcfd_shuffle_connectivity(const int32_t *old_to_new_for_nodes, int32_t n_nodes, const int32_t *new_to_old_for_cells,
                              int32_t n_cells, int32_t npe, int32_t *conn) {
  if (!old_to_new_for_nodes || !new_to_old_for_cells || !conn || npe <= 0) return MCFD_ERR_ARG;
  std::vector<int32_t> renumbered(size_t(n_cells) * npe);
  for (int64_t i = 0; i < int64_t(n_cells) * npe; ++i) {
    const int32_t old_nid = conn[i];
    if (old_nid < 1 || old_nid > n_nodes) return MCFD_ERR_ARG;  // the reference throws from vector::at
    renumbered[i] = old_to_new_for_nodes[old_nid - 1] + 1;
  }
  for (int32_t c = 0; c < n_cells; ++c)
    std::memcpy(conn + int64_t(c) * npe, renumbered.data() + int64_t(new_to_old_for_cells[c]) * npe, sizeof(int32_t) * npe);
  return MCFD_OK;
}

// metis.hpp:316-337
int mcfd_get_node_parts(int32_t n_cells, int32_t n_nodes, const int32_t *range, const int32_t *index,
                        const int32_t *cell_parts, int32_t n_parts, int32_t *node_parts) {
  if (!range || !index || !cell_parts || !node_parts) return MCFD_ERR_ARG;
  std::fill(node_parts, node_parts + n_nodes, n_parts);
  for (int32_t c = 0; c < n_cells; ++c)
    for (int32_t j = range[c]; j < range[c + 1]; ++j)
      if (cell_parts[c] < node_parts[index[j]]) node_parts[index[j]] = cell_parts[c];
  return MCFD_OK;
}

int mcfd_partitioner_create(const mcfd_global_mesh *m, const int32_t *cell_parts, int32_t n_parts, mcfd_partitioner **out) {
  if (!m || !cell_parts || !out || n_parts < 1 || m->n_cells < 1 || !m->cell_nodes || !m->node_xyz) return MCFD_ERR_ARG;
  if (m->n_cells > INT32_MAX || m->n_nodes > INT32_MAX) return MCFD_ERR_ARG;  // the reference's orders are int
  auto *p = new (std::nothrow) mcfd_partitioner;
  if (!p) return MCFD_ERR_STATE;
  p->n_nodes = m->n_nodes; p->n_cells = m->n_cells; p->n_bfaces = m->n_boundary_faces;
  p->n_parts = n_parts; p->n_groups = m->n_bc_groups;
  p->xyz.assign(m->node_xyz, m->node_xyz + 3 * m->n_nodes);
  p->cell_nodes.assign(m->cell_nodes, m->cell_nodes + 8 * m->n_cells);
  p->cell_parts.assign(cell_parts, cell_parts + m->n_cells);
  if (m->n_boundary_faces > 0) {
    if (!m->bface_nodes || !m->bface_group) { delete p; return MCFD_ERR_ARG; }
    p->bface_nodes.assign(m->bface_nodes, m->bface_nodes + 4 * m->n_boundary_faces);
    p->bface_group.assign(m->bface_group, m->bface_group + m->n_boundary_faces);
  }
  for (int64_t c = 0; c < p->n_cells; ++c)
    if (cell_parts[c] < 0 || cell_parts[c] >= n_parts) { delete p; return MCFD_ERR_ARG; }
  for (int64_t i = 0; i < 8 * p->n_cells; ++i)
    if (p->cell_nodes[i] < 0 || p->cell_nodes[i] >= p->n_nodes) { delete p; return MCFD_ERR_ARG; }

  MeshToDual(p->n_cells, p->n_nodes, 8, p->cell_nodes.data(), 3, &p->range, &p->index);
  // GetNodeParts
  p->node_parts.assign(p->n_nodes, n_parts);
  for (int64_t c = 0; c < p->n_cells; ++c)
    for (int v = 0; v < 8; ++v) {
      int32_t &np = p->node_parts[p->cell_nodes[8 * c + v]];
      if (cell_parts[c] < np) np = cell_parts[c];
    }
  // FillFaceParts: the one cell that uses all the nodes of the face owns it
  std::vector<std::vector<int64_t>> users(p->n_nodes);
  for (int64_t c = 0; c < p->n_cells; ++c)
    for (int v = 0; v < 8; ++v) users[p->cell_nodes[8 * c + v]].push_back(c);
  p->face_owner.assign(p->n_bfaces, -1);
  p->face_parts.assign(p->n_bfaces, 0);
  for (int64_t f = 0; f < p->n_bfaces; ++f) {
    std::unordered_map<int64_t, int> count;
    for (int v = 0; v < 4; ++v) {
      const int64_t node = p->bface_nodes[4 * f + v];
      if (node < 0 || node >= p->n_nodes) { delete p; return MCFD_ERR_ARG; }
      for (int64_t c : users[node]) count[c]++;
    }
    int owners = 0;
    for (auto &[c, n] : count)
      if (n == 4) { p->face_owner[f] = c; ++owners; }
    if (owners != 1) { delete p; return MCFD_ERR_ARG; }  // shuffler.hpp:224 asserts the same
    p->face_parts[f] = cell_parts[p->face_owner[f]];
  }
  NewOrder(p->cell_parts.data(), p->n_cells, &p->cells_new_to_old, &p->cells_old_to_new);
  NewOrder(p->node_parts.data(), p->n_nodes, &p->nodes_new_to_old, &p->nodes_old_to_new);
  *out = p;
  return MCFD_OK;
}

void mcfd_partitioner_destroy(mcfd_partitioner *p) { delete p; }
const int64_t *mcfd_partitioner_dual_range(const mcfd_partitioner *p) { return p ? p->range.data() : nullptr; }
const int64_t *mcfd_partitioner_dual_index(const mcfd_partitioner *p) { return p ? p->index.data() : nullptr; }
const int32_t *mcfd_partitioner_node_parts(const mcfd_partitioner *p) { return p ? p->node_parts.data() : nullptr; }
const int32_t *mcfd_partitioner_face_parts(const mcfd_partitioner *p) { return p ? p->face_parts.data() : nullptr; }
const int32_t *mcfd_partitioner_cells_new_to_old(const mcfd_partitioner *p) { return p ? p->cells_new_to_old.data() : nullptr; }
const int32_t *mcfd_partitioner_nodes_new_to_old(const mcfd_partitioner *p) { return p ? p->nodes_new_to_old.data() : nullptr; }

}  // extern "C"

// ---- one part as an mcfd_mesh: Part's build order -------------------------------------------------------------------
extern "C" int mcfd_partitioner_part_mesh(const mcfd_partitioner *p, int32_t me, mcfd_box_mesh **out) {
  if (!p || !out || me < 0 || me >= p->n_parts) return MCFD_ERR_ARG;
  auto *bm = new (std::nothrow) mcfd_box_mesh;
  if (!bm) return MCFD_ERR_STATE;
  const auto &parts = p->cell_parts;
  // local cells in ascending metis id (the shuffled section lists a part's cells in their old order, shuffler.hpp:41-44);
  // a cell that is sent to a neighbour is an "inter" cell (part.hpp:964-976), ids run over inner then inter cells (:818-835)
  std::vector<int64_t> mine, inner, inter;
  std::map<int32_t, std::set<int64_t>> send, recv;  // [neighbour part] -> metis ids, ascending (part.hpp:860-886 uses std::map)
  for (int64_t c = 0; c < p->n_cells; ++c) {
    if (parts[c] != me) continue;
    mine.push_back(c);
    bool sent = false;
    for (int64_t r = p->range[c]; r < p->range[c + 1]; ++r) {
      const int64_t j = p->index[r];
      if (parts[j] != me) { send[parts[j]].insert(c); recv[parts[j]].insert(j); sent = true; }
    }
    (sent ? inter : inner).push_back(c);
  }
  std::unordered_map<int64_t, int64_t> local_of;
  local_of.reserve(mine.size() * 2);
  for (int64_t c : inner) { local_of[c] = int64_t(bm->cell_gid.size()); bm->cell_gid.push_back(c); }
  for (int64_t c : inter) { local_of[c] = int64_t(bm->cell_gid.size()); bm->cell_gid.push_back(c); }
  const int64_t n_local = int64_t(bm->cell_gid.size());
  bm->send_ptr.push_back(0); bm->recv_ptr.push_back(0);
  for (auto &[q, cells] : recv) {
    bm->nbr_rank.push_back(q);
    for (int64_t j : cells) { local_of[j] = int64_t(bm->cell_gid.size()); bm->cell_gid.push_back(j); }
    for (int64_t i : send.at(q)) bm->send_cells.push_back(local_of.at(i));
    bm->recv_ptr.push_back(int64_t(bm->cell_gid.size()) - n_local);
    bm->send_ptr.push_back(int64_t(bm->send_cells.size()));
  }
  const int64_t n_cells = int64_t(bm->cell_gid.size());
  // nodes in order of first use
  std::unordered_map<int64_t, int64_t> node_of;
  node_of.reserve(size_t(n_cells) * 2);
  bm->cell_nodes.resize(size_t(n_cells) * 8);
  for (int64_t c = 0; c < n_cells; ++c)
    for (int v = 0; v < 8; ++v) {
      const int64_t g = p->cell_nodes[8 * bm->cell_gid[c] + v];
      auto it = node_of.find(g);
      if (it == node_of.end()) {
        it = node_of.emplace(g, int64_t(node_of.size())).first;
        bm->node_xyz.insert(bm->node_xyz.end(), &p->xyz[3 * g], &p->xyz[3 * g] + 3);
      }
      bm->cell_nodes[8 * c + v] = it->second;
    }
  auto emit_face = [&](int64_t holder, int64_t sharer) {  // metis ids; the common nodes in the holder's node order
    bm->face_holder.push_back(local_of.at(holder));
    bm->face_sharer.push_back(local_of.at(sharer));
    int n = 0;
    for (int v = 0; v < 8 && n < 4; ++v) {
      const int64_t g = p->cell_nodes[8 * holder + v];
      bool shared = false;
      for (int w = 0; w < 8; ++w) shared |= p->cell_nodes[8 * sharer + w] == g;
      if (shared) { bm->face_nodes.push_back(node_of.at(g)); ++n; }
    }
    return n == 4;
  };
  bool ok = true;
  // local faces: the inner adjacency list, `i < j` in the order the dual graph is walked (shuffler.hpp:395-407)
  int64_t n_local_faces = 0, n_ghost_faces = 0;
  for (int64_t i : mine)
    for (int64_t r = p->range[i]; r < p->range[i + 1]; ++r) {
      const int64_t j = p->index[r];
      if (parts[j] == me && i < j) { ok &= emit_face(i, j); ++n_local_faces; }
    }
  // ghost faces: the interpart adjacency list, by neighbour part, then as the graph is walked (shuffler.hpp:408, 477-489)
  for (auto &[q, cells] : recv)
    for (int64_t i : mine)
      for (int64_t r = p->range[i]; r < p->range[i + 1]; ++r) {
        const int64_t j = p->index[r];
        if (parts[j] == q) { ok &= emit_face(i, j); ++n_ghost_faces; }
      }
  // boundary faces: section by section, a part's faces in their old order (shuffler.hpp:270-279, part.hpp:1440-1533)
  int64_t n_bound = 0;
  for (int32_t g = 0; g < p->n_groups; ++g)
    for (int64_t f = 0; f < p->n_bfaces; ++f) {
      if (p->bface_group[f] != g || p->face_parts[f] != me) continue;
      bm->face_holder.push_back(local_of.at(p->face_owner[f]));
      bm->face_sharer.push_back(-1);
      for (int v = 0; v < 4; ++v) bm->face_nodes.push_back(node_of.at(p->bface_nodes[4 * f + v]));
      bm->bface_group.push_back(g);
      ++n_bound;
    }
  if (!ok) { delete bm; return MCFD_ERR_ARG; }  // two face-adjacent hexahedra share exactly 4 nodes
  mcfd_mesh &m = bm->view;
  m.n_nodes = int64_t(bm->node_xyz.size()) / 3;
  m.node_xyz = bm->node_xyz.data();
  m.n_local_cells = n_local;
  m.n_inner_cells = int64_t(inner.size());
  m.n_ghost_cells = n_cells - n_local;
  m.cell_nodes = bm->cell_nodes.data();
  m.n_local_faces = n_local_faces; m.n_ghost_faces = n_ghost_faces; m.n_boundary_faces = n_bound;
  m.face_holder = bm->face_holder.data();
  m.face_sharer = bm->face_sharer.data();
  m.face_nodes = bm->face_nodes.data();
  m.bface_group = bm->bface_group.data();
  m.n_bc_groups = p->n_groups;
  m.n_neighbours = int32_t(bm->nbr_rank.size());
  m.nbr_rank = bm->nbr_rank.data();
  m.send_ptr = bm->send_ptr.data();
  m.send_cells = bm->send_cells.data();
  m.recv_ptr = bm->recv_ptr.data();
  *out = bm;
  return MCFD_OK;
}

// ---- partition/<part>.txt : Shuffler::WritePartitionInfo (shuffler.hpp:314-506) --------------------------------------
// One zone; section 1 = the cells (elements 1 .. n_cells after the shuffle), section 2 + g = boundary group g, its
// elements numbered on from the previous section.  Ids of nodes, cells and faces in ranges are the SHUFFLED 1-based
// ones; "metis" ids are the 0-based ids of the input mesh.
extern "C" int mcfd_partitioner_info_text(const mcfd_partitioner *p, int32_t me, char *buf, int64_t capacity,
                                          int64_t *needed) {
  if (!p || !needed || me < 0 || me >= p->n_parts) return MCFD_ERR_ARG;
  std::string s;
  auto line = [&s](std::initializer_list<long long> v) {
    bool first = true;
    for (long long x : v) { if (!first) s += ' '; s += std::to_string(x); first = false; }
    s += '\n';
  };
  const auto &parts = p->cell_parts;
  // a part's nodes / cells are contiguous after the shuffle: [head, tail) of the new 1-based ids
  auto slice = [me](const std::vector<int32_t> &new_to_old, const int32_t *part_of) {
    long long head = 0, tail = 0;
    const long long n = (long long)new_to_old.size();
    for (long long i = 0; i < n; ++i)
      if (part_of[new_to_old[i]] == me) { if (!tail) head = i + 1; tail = i + 2; }
    return std::pair<long long, long long>(head, tail);
  };
  s += "# i_zone i_node_head i_node_tail\n";
  auto [nh, nt] = slice(p->nodes_new_to_old, p->node_parts.data());
  line({ 1, nh, nt });
  // nodes of other parts' cells (or of their foreign neighbours) that I own are sent; the mirror image is received
  std::map<int32_t, std::set<int64_t>> send_nodes, adj_nodes;
  auto visit = [&](int32_t part_i, int64_t cell) {
    for (int v = 0; v < 8; ++v) {
      const int64_t node = p->cell_nodes[8 * cell + v];
      const int32_t owner = p->node_parts[node];
      if (owner == part_i) continue;
      if (part_i == me) adj_nodes[owner].insert(node);
      if (owner == me) send_nodes[part_i].insert(node);
    }
  };
  std::vector<std::pair<int64_t, int64_t>> inner_adj;
  std::map<int32_t, std::vector<std::pair<int64_t, int64_t>>> inter_adj;
  for (int64_t i = 0; i < p->n_cells; ++i) {
    const int32_t part_i = parts[i];
    visit(part_i, i);
    for (int64_t r = p->range[i]; r < p->range[i + 1]; ++r) {
      const int64_t j = p->index[r];
      if (parts[j] == part_i) {
        if (part_i == me && i < j) inner_adj.emplace_back(i, j);
      } else {
        if (part_i == me) inter_adj[parts[j]].emplace_back(i, j);
        visit(part_i, j);
      }
    }
  }
  s += "# i_part i_node_metis\n";
  for (auto &[q, nodes] : send_nodes) for (int64_t n : nodes) line({ q, n });
  s += "# i_part i_node_metis i_zone i_node\n";
  for (auto &[q, nodes] : adj_nodes) for (int64_t n : nodes) line({ q, n, 1, p->nodes_old_to_new[n] + 1 });
  s += "# i_zone i_sect i_cell_head i_cell_tail\n";
  auto [ch, ct] = slice(p->cells_new_to_old, parts.data());
  line({ 1, 1, ch, ct });
  s += "# i_cell_metis j_cell_metis\n";
  for (auto &[i, j] : inner_adj) line({ i, j });
  s += "# i_part i_cell_metis j_cell_metis i_node_npe j_node_npe\n";
  for (auto &[q, pairs] : inter_adj) for (auto &[i, j] : pairs) line({ q, i, j, 8, 8 });
  s += "# i_zone i_sect i_face_head i_face_tail\n";
  long long first = p->n_cells + 1;  // element ids run on through the sections
  for (int32_t g = 0; g < p->n_groups; ++g) {
    long long before = 0, count = 0, total = 0;
    for (int64_t f = 0; f < p->n_bfaces; ++f) {
      if (p->bface_group[f] != g) continue;
      ++total;
      if (p->face_parts[f] < me) ++before;
      if (p->face_parts[f] == me) ++count;
    }
    // the reference writes the default-constructed {0, 0} for a part without faces in the section
    if (count) line({ 1, 2 + g, first + before, first + before + count }); else line({ 1, 2 + g, 0, 0 });
    first += total;
  }
  s += "#\n";
  *needed = int64_t(s.size());
  if (buf && capacity >= int64_t(s.size())) std::memcpy(buf, s.data(), s.size());
  return MCFD_OK;
}
