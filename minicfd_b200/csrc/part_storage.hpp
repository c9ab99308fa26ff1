// part_storage.hpp -- owner of the arrays behind one `mcfd_mesh` view (one mesh part).  Filled by the box-mesh
// generator (box_mesh.cpp) and by the general partition pipeline (partition.cpp).
#pragma once
#include <cstdint>
#include <vector>

#include "minicfd_b200.h"

struct mcfd_box_mesh {
  mcfd_mesh view{};
  std::vector<double> node_xyz;
  std::vector<int64_t> cell_nodes, cell_gid;
  std::vector<int64_t> face_holder, face_sharer, face_nodes;
  std::vector<int32_t> bface_group;
  std::vector<int32_t> nbr_rank;
  std::vector<int64_t> send_ptr, send_cells, recv_ptr;
};
