// TEST INFRASTRUCTURE ONLY -- the reference's own partition helpers (header templates under /root/reference,
// compiled where they lie): mesh::GetNewOrder, ShuffleData, ShuffleConnectivity (mesh/shuffler.hpp:36-104) and
// metis::GetNodeParts (mesh/metis.hpp:316-337), behind a C interface for tests/test_partition.py.
#include <cassert>
#include <cstring>
#include <vector>

#include "cgnslib.h"
extern "C" int cg_npe(CGNS_ENUMT(ElementType_t), int *npe) { *npe = 0; return 1; }  // named by mesh/cgns.hpp only, never called

#include "mini/mesh/metis.hpp"
#include "mini/mesh/shuffler.hpp"

extern "C" {

void ref_get_new_order(const int *parts, int n, int *new_to_old, int *old_to_new) {
  auto [a, b] = mini::mesh::GetNewOrder(parts, n);
  std::memcpy(new_to_old, a.data(), sizeof(int) * n);
  std::memcpy(old_to_new, b.data(), sizeof(int) * n);
}

void ref_shuffle_data(const int *new_to_old, int n, double *data) {
  std::vector<int> order(new_to_old, new_to_old + n);
  mini::mesh::ShuffleData(order, data);
}

void ref_shuffle_connectivity(const int *old_to_new_nodes, int n_nodes, const int *new_to_old_cells, int n_cells, int npe,
                              int *conn) {
  std::vector<int> nodes(old_to_new_nodes, old_to_new_nodes + n_nodes), cells(new_to_old_cells, new_to_old_cells + n_cells);
  mini::mesh::ShuffleConnectivity(nodes, cells, npe, conn);
}

void ref_get_node_parts(int n_cells, int n_nodes, const int *range, const int *index, const int *cell_parts, int n_parts,
                        int *node_parts) {
  std::vector<int> r(range, range + n_cells + 1), i(index, index + range[n_cells]), parts(cell_parts, cell_parts + n_cells);
  auto mesh = mini::mesh::metis::Mesh<int>(r, i, n_nodes);
  auto out = mini::mesh::metis::GetNodeParts(mesh, parts, n_parts);
  std::memcpy(node_parts, out.data(), sizeof(int) * n_nodes);
}

}  // extern "C"
