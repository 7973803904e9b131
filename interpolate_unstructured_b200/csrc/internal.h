// internal.h -- shared declarations of libiu_b200.so (not installed; the public ABI is include/iu_b200.h)
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace iu {

constexpr int LEAF = 8;           // centres per kd-tree leaf block (fixed trip count on the device)
constexpr int MAX_VARS = 32;      // variables per launch of the interpolation kernel

// cell types (m_interp_unstructured.f90:13-15) and the per-query "step cap hit" sentinel
constexpr int IU_TRIANGLE_CT = 1;
constexpr int IU_QUAD_CT = 2;
constexpr int IU_TETRA_CT = 3;
constexpr int IU_WALK_LIMIT = -2;

void set_error(const std::string& msg);

// ---- BINDA container (m_binda.f90:52-194; written by convert_to_binary.py:68-115) ----------
struct HostMesh {
  int cell_type = 0, ppc = 0, n_points = 0, n_cells = 0;
  std::vector<double> points;       // (3, Np)
  std::vector<int32_t> cells;       // (ppc, Nc) 1-based
  std::vector<int32_t> neighbors;   // (ppc, Nc) 1-based, < 1 none
  std::vector<double> point_data;   // (Np, n) column-major
  std::vector<double> cell_data;    // (Nc, n)
  std::vector<int32_t> icell_data;  // (Nc, n)
  std::vector<std::string> point_names, cell_names, icell_names;
};
// iu_read_grid's file part (m_interp_unstructured.f90:829-913). Returns false + set_error on failure.
bool read_binda_mesh(const char* filename, double coord_scale_factor, HostMesh& out);

// ---- face neighbours (convert_to_binary.py:118-162), 0-based in/out --------------------------
bool build_cell_neighbors(int cell_type, int32_t n_points, int32_t n_cells, const double* points,
                          const int32_t* cells, int32_t* neighbors);

// ---- iu_write_vtk (m_interp_unstructured.f90:930-985 through m_vtk.f90:45-221) ------------------
// Writes the same byte stream as the reference: XML header with appended raw data (int32 byte
// count before each block), points as Float32, connectivity 0-based.  cells: (ppc, Nc) 1-based.
struct VtuColumns {
  std::vector<std::string> names;
  const void* data = nullptr;   // column-major (n_rows, names.size()), double or int32
};
bool write_vtu(const char* filename, int cell_type, int32_t n_points, int32_t n_cells, const double* points,
               const int32_t* cells, const VtuColumns& cell_data, const VtuColumns& icell_data,
               const VtuColumns& point_data);

// ---- nearest-centre kd-tree ------------------------------------------------------------------
// Complete (left-balanced, heap-indexed) binary tree over the cell centres: n_leaves =
// ceil(Nc / LEAF) leaf blocks of exactly LEAF slots (the last one padded), n_leaves - 1 internal
// nodes {split dimension = widest extent of the node's exact bounding box, split value s with
// left <= s <= right}.  Node i has children 2i+1, 2i+2; nodes >= n_leaves-1 are leaves and leaf
// block b = node - (n_leaves-1) owns device slots [b*LEAF, (b+1)*LEAF).
struct KdTreeHost {
  int n_leaves = 0;
  std::vector<double> split;        // n_leaves - 1
  std::vector<uint8_t> sdim;        // n_leaves - 1
  std::vector<int32_t> orig_of_slot;  // n_leaves*LEAF, 1-based original cell id, 0 = padding
  // tight bounding box of the centres below every node (internal nodes and leaves, heap order),
  // 8 floats per node {lo x, y, z, hi x, y, z, 0, 0}, lo rounded down / hi rounded up to float
  std::vector<float> box;           // 8 * (2 * n_leaves - 1)
};
// centers: (3, Nc) as iu_get_cell_center computes them (:264-269)
void build_kdtree(const double* centers, int32_t n_cells, KdTreeHost& out);

}  // namespace iu
