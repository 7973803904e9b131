// host_mesh.cpp -- host-side (C++17) parts of libiu_b200.so that are not device work:
//   * BINDA reader            (replaces m_binda.f90 + the file part of iu_read_grid, :829-913)
//   * face-neighbour builder  (replaces convert_to_binary.py:118-162; sort-based, scales to 50M tets)
//   * kd-tree construction    (replaces kdtree2_create at m_interp_unstructured.f90:261)
// Compiled with -ffp-contract=off: the only floating-point work here is comparisons and the
// scale-factor multiply, but nothing may be contracted.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <future>
#include <limits>
#include <thread>

#include "internal.h"

namespace iu {

// ------------------------------------------------------------------------------------------------
// BINDA
// ------------------------------------------------------------------------------------------------
namespace {
struct Entry {
  std::string name, dtype, meta;
  int64_t ndim = 0, shape[8] = {0}, offset = 0;
};

std::string trim(const char* s, int n) {
  while (n > 0 && (s[n - 1] == ' ' || s[n - 1] == '\0')) --n;
  return std::string(s, n);
}

struct File {
  FILE* f = nullptr;
  ~File() { if (f) fclose(f); }
};

bool read_at(FILE* f, int64_t off, void* dst, size_t bytes) {
  if (fseeko(f, off, SEEK_SET) != 0) return false;
  return fread(dst, 1, bytes, f) == bytes;
}

// integer payload (int32 | int64) -> int32, as binda_read_alloc_int32_2d / binda_read_int32_1d
bool read_i32(FILE* f, const Entry& e, size_t n, int32_t* out) {
  if (e.dtype == "int32") return read_at(f, e.offset, out, n * 4);
  if (e.dtype == "int64") {
    std::vector<int64_t> tmp(n);
    if (!read_at(f, e.offset, tmp.data(), n * 8)) return false;
    for (size_t i = 0; i < n; ++i) out[i] = (int32_t)tmp[i];
    return true;
  }
  set_error("Unsupported data type " + e.dtype);
  return false;
}
// float payload (float64 | float32) -> float64, as binda_read_alloc_float64_2d / binda_read_float64_1d
bool read_f64(FILE* f, const Entry& e, size_t n, double* out) {
  if (e.dtype == "float64") return read_at(f, e.offset, out, n * 8);
  if (e.dtype == "float32") {
    std::vector<float> tmp(n);
    if (!read_at(f, e.offset, tmp.data(), n * 4)) return false;
    for (size_t i = 0; i < n; ++i) out[i] = (double)tmp[i];
    return true;
  }
  set_error("Unsupported data type " + e.dtype);
  return false;
}
}  // namespace

bool read_binda_mesh(const char* filename, double scale, HostMesh& m) {
  File file;
  file.f = fopen(filename, "rb");
  if (!file.f) { set_error(std::string("cannot open ") + filename); return false; }
  char hdr[24];
  if (fread(hdr, 1, 24, file.f) != 24 || trim(hdr, 8) != "BINDA") { set_error("Wrong file format"); return false; }
  int64_t n_entries;
  std::memcpy(&n_entries, hdr + 8, 8);
  if (n_entries < 0 || n_entries > (1 << 20)) { set_error("Wrong file format"); return false; }
  std::vector<Entry> ent((size_t)n_entries);
  std::vector<char> rec(464);
  for (auto& e : ent) {
    if (fread(rec.data(), 1, 464, file.f) != 464) { set_error("truncated BINDA header"); return false; }
    e.name = trim(&rec[0], 128); e.dtype = trim(&rec[128], 128); e.meta = trim(&rec[256], 128);
    std::memcpy(&e.ndim, &rec[384], 8);
    std::memcpy(e.shape, &rec[392], 64);
    std::memcpy(&e.offset, &rec[456], 8);
  }
  auto find = [&](const char* nm) -> const Entry* {
    for (auto& e : ent) if (e.name == nm) return &e;   // first match, binda_get_index
    return nullptr;
  };
  const Entry* e = find("cells");
  if (!e) { set_error("cells not found in binda file"); return false; }
  if (e->ndim != 2) { set_error("Invalid ndim"); return false; }
  if (e->meta == "triangle") m.cell_type = 1;
  else if (e->meta == "quad") m.cell_type = 2;
  else if (e->meta == "tetra") m.cell_type = 3;
  else { set_error("Unsupported cell type '" + e->meta + "'"); return false; }
  m.n_cells = (int)e->shape[0];
  m.ppc = (int)e->shape[1];
  if (m.ppc != (m.cell_type == 1 ? 3 : 4)) { set_error("cells array has wrong number of points per cell"); return false; }
  m.cells.resize((size_t)m.ppc * m.n_cells);
  if (!read_i32(file.f, *e, m.cells.size(), m.cells.data())) { set_error("cannot read cells"); return false; }

  e = find("points");
  if (!e) { set_error("points not found in binda file"); return false; }
  if (e->ndim != 2 || e->shape[1] != 3) { set_error("Invalid ndim"); return false; }
  m.n_points = (int)e->shape[0];
  m.points.resize((size_t)3 * m.n_points);
  if (!read_f64(file.f, *e, m.points.size(), m.points.data())) { set_error("cannot read points"); return false; }

  e = find("cell_neighbors");
  if (!e) { set_error("cell_neighbors not found in binda file"); return false; }
  if (e->ndim != 2 || e->shape[0] != m.n_cells || e->shape[1] != m.ppc) { set_error("Invalid ndim"); return false; }
  m.neighbors.resize(m.cells.size());
  if (!read_i32(file.f, *e, m.neighbors.size(), m.neighbors.data())) { set_error("cannot read cell_neighbors"); return false; }

  if (scale != 1.0) for (auto& x : m.points) x = x * scale;      // :858-860
  for (auto& c : m.cells) c += 1;                                  // :871
  for (auto& c : m.neighbors) c += 1;                              // :872

  for (auto& en : ent) {                                           // :894-911, file order
    if (en.name == "point_data" || en.name == "cell_data" || en.name == "icell_data") {
      size_t want = en.name == "point_data" ? m.n_points : m.n_cells;
      if (en.ndim != 1) { set_error("Invalid ndim"); return false; }
      if ((size_t)en.shape[0] != want) { set_error("Invalid array_size"); return false; }
      if (en.name == "point_data") {
        m.point_names.push_back(en.meta);
        m.point_data.resize(m.point_data.size() + want);
        if (!read_f64(file.f, en, want, &m.point_data[m.point_data.size() - want])) return false;
      } else if (en.name == "cell_data") {
        m.cell_names.push_back(en.meta);
        m.cell_data.resize(m.cell_data.size() + want);
        if (!read_f64(file.f, en, want, &m.cell_data[m.cell_data.size() - want])) return false;
      } else {
        m.icell_names.push_back(en.meta);
        m.icell_data.resize(m.icell_data.size() + want);
        if (!read_i32(file.f, en, want, &m.icell_data[m.icell_data.size() - want])) return false;
      }
    }
  }
  return true;
}

// ------------------------------------------------------------------------------------------------
// face neighbours
// ------------------------------------------------------------------------------------------------
namespace {
struct Face {
  uint32_t v[3];   // sorted unique-point ids (v[2] = 0xffffffff for 2-point faces)
  uint32_t cf;     // cell * ppc + face
};
inline bool face_less(const Face& a, const Face& b) {
  if (a.v[0] != b.v[0]) return a.v[0] < b.v[0];
  if (a.v[1] != b.v[1]) return a.v[1] < b.v[1];
  return a.v[2] < b.v[2];
}
inline bool face_eq(const Face& a, const Face& b) { return a.v[0] == b.v[0] && a.v[1] == b.v[1] && a.v[2] == b.v[2]; }

unsigned hw_threads() {
  unsigned t = std::thread::hardware_concurrency();
  return t == 0 ? 4 : std::min(t, 32u);
}
}  // namespace

bool build_cell_neighbors(int cell_type, int32_t n_points, int32_t n_cells, const double* points,
                          const int32_t* cells, int32_t* neighbors) {
  if (cell_type < 1 || cell_type > 3) { set_error("bad cell type"); return false; }
  const int ppc = cell_type == 1 ? 3 : 4;
  const int npf = cell_type == 3 ? 3 : 2;
  if ((int64_t)n_cells * ppc >= (int64_t)0xffffffffLL) { set_error("mesh too large"); return false; }

  // merge coincident points: np.unique(points, axis=0, return_inverse=True), :130-136
  std::vector<uint32_t> order((size_t)n_points), uniq((size_t)n_points);
  for (int32_t i = 0; i < n_points; ++i) order[i] = (uint32_t)i;
  auto pless = [points](uint32_t a, uint32_t b) {
    const double* pa = points + 3 * (size_t)a; const double* pb = points + 3 * (size_t)b;
    if (pa[0] != pb[0]) return pa[0] < pb[0];
    if (pa[1] != pb[1]) return pa[1] < pb[1];
    return pa[2] < pb[2];
  };
  std::sort(order.begin(), order.end(), pless);
  uint32_t uid = 0;
  for (int32_t i = 0; i < n_points; ++i) {
    if (i > 0) {
      const double* pa = points + 3 * (size_t)order[i - 1]; const double* pb = points + 3 * (size_t)order[i];
      if (!(pa[0] == pb[0] && pa[1] == pb[1] && pa[2] == pb[2])) ++uid;
    }
    uniq[order[i]] = uid;
  }
  order.clear(); order.shrink_to_fit();

  const size_t nf = (size_t)n_cells * ppc;
  std::vector<Face> faces(nf);
  for (int32_t c = 0; c < n_cells; ++c) {
    for (int i = 0; i < ppc; ++i) {
      Face f;
      f.v[2] = 0xffffffffu;
      for (int k = 0; k < npf; ++k) {
        int32_t p = cells[(size_t)c * ppc + (i + k) % ppc];
        if (p < 0 || p >= n_points) { set_error("cell references a point outside 0..n_points-1"); return false; }
        f.v[k] = uniq[p];
      }
      if (npf == 2) { if (f.v[0] > f.v[1]) std::swap(f.v[0], f.v[1]); }
      else {
        if (f.v[0] > f.v[1]) std::swap(f.v[0], f.v[1]);
        if (f.v[1] > f.v[2]) std::swap(f.v[1], f.v[2]);
        if (f.v[0] > f.v[1]) std::swap(f.v[0], f.v[1]);
      }
      f.cf = (uint32_t)((size_t)c * ppc + i);
      faces[(size_t)c * ppc + i] = f;
    }
  }
  // bucket by smallest vertex id so that buckets can be sorted in parallel
  const unsigned T = hw_threads();
  const size_t n_buckets = nf < (1u << 16) ? 1 : (size_t)T * 8;
  if (n_buckets > 1) {
    const uint64_t nu = (uint64_t)uid + 1;
    auto bucket_of = [&](const Face& f) { return (size_t)(((uint64_t)f.v[0] * n_buckets) / nu); };
    std::vector<size_t> count(n_buckets + 1, 0);
    for (auto& f : faces) ++count[bucket_of(f) + 1];
    for (size_t b = 0; b < n_buckets; ++b) count[b + 1] += count[b];
    std::vector<Face> tmp(nf);
    {
      std::vector<size_t> pos(count.begin(), count.end() - 1);
      for (auto& f : faces) tmp[pos[bucket_of(f)]++] = f;
    }
    faces.swap(tmp);
    tmp.clear(); tmp.shrink_to_fit();
    std::vector<std::thread> th;
    std::atomic<size_t> next{0};
    for (unsigned t = 0; t < T; ++t)
      th.emplace_back([&]() {
        for (;;) {
          size_t b = next.fetch_add(1);
          if (b >= n_buckets) break;
          std::sort(faces.begin() + count[b], faces.begin() + count[b + 1], face_less);
        }
      });
    for (auto& t : th) t.join();
  } else {
    std::sort(faces.begin(), faces.end(), face_less);
  }
  for (size_t i = 0; i < nf; ++i) neighbors[i] = -1;
  size_t i = 0;
  while (i < nf) {
    size_t j = i + 1;
    while (j < nf && face_eq(faces[i], faces[j])) ++j;
    if (j - i == 2) {   // "There can only be a neighbor if the face connects two cells", :156-160
      neighbors[faces[i].cf] = (int32_t)(faces[i + 1].cf / ppc);
      neighbors[faces[i + 1].cf] = (int32_t)(faces[i].cf / ppc);
    }
    i = j;
  }
  return true;
}

// ------------------------------------------------------------------------------------------------
// VTU writer
// ------------------------------------------------------------------------------------------------
bool write_vtu(const char* filename, int cell_type, int32_t n_points, int32_t n_cells, const double* points,
               const int32_t* cells, const VtuColumns& cell_data, const VtuColumns& icell_data,
               const VtuColumns& point_data) {
  const int ppc = cell_type == IU_TRIANGLE_CT ? 3 : 4;
  const int32_t vtk_type = cell_type == IU_TRIANGLE_CT ? 5 : (cell_type == IU_QUAD_CT ? 9 : 10);   // :941-950
  FILE* f = std::fopen(filename, "wb");
  if (!f) { set_error(std::string("cannot open ") + filename + " for writing"); return false; }
  std::string xml;            // the header part (funit); the raw blocks (sunit) follow in `raw`
  std::vector<char> raw;
  int indent = 0;
  long long offset = 0;
  auto line = [&](const std::string& t) { xml.append((size_t)indent, ' '); xml += t; xml += '\n'; };
  auto open_tag = [&](const char* name) { line(std::string("<") + name + ">"); indent += 2; };       // vtk_dat_xml
  auto close_tag = [&](const char* name) { indent -= 2; line(std::string("</") + name + ">"); };
  auto block = [&](const void* p, size_t bytes) {                                                    // "write(sunit) n_bytes, data"
    const int32_t nb = (int32_t)bytes;
    const char* h = reinterpret_cast<const char*>(&nb);
    raw.insert(raw.end(), h, h + 4);
    const char* d = reinterpret_cast<const char*>(p);
    raw.insert(raw.end(), d, d + bytes);
    offset += 4 + (long long)bytes;
  };
  xml += "<?xml version=\"1.0\"?>\n";                                                               // vtk_ini_xml
  xml += "<VTKFile type=\"UnstructuredGrid\" version=\"0.1\" byte_order=\"LittleEndian\">\n";
  indent = 2;
  open_tag("UnstructuredGrid");
  line("<Piece NumberOfPoints=\"" + std::to_string(n_points) + "\" NumberOfCells=\"" + std::to_string(n_cells) + "\">");
  indent += 2;
  open_tag("Points");                                                                               // vtk_unstr_geo_xml
  line("<DataArray type=\"Float32\" NumberOfComponents=\"3\" Name=\"Points\" format=\"appended\" offset=\"" +
       std::to_string(offset) + "\"/>");
  {
    std::vector<float> p32((size_t)3 * n_points);
    for (size_t i = 0; i < p32.size(); ++i) p32[i] = (float)points[i];
    block(p32.data(), p32.size() * 4);
  }
  close_tag("Points");
  open_tag("Cells");                                                                                // vtk_unstr_con_xml
  {
    std::vector<int32_t> tmp((size_t)ppc * n_cells);
    line("<DataArray type=\"Int32\" Name=\"connectivity\" format=\"appended\" offset=\"" + std::to_string(offset) + "\"/>");
    for (size_t i = 0; i < tmp.size(); ++i) tmp[i] = cells[i] - 1;                                  // :957
    block(tmp.data(), tmp.size() * 4);
    line("<DataArray type=\"Int32\" Name=\"offsets\" format=\"appended\" offset=\"" + std::to_string(offset) + "\"/>");
    tmp.resize((size_t)n_cells);
    for (int32_t n = 0; n < n_cells; ++n) tmp[n] = (n + 1) * ppc;                                   // :952-954
    block(tmp.data(), tmp.size() * 4);
    line("<DataArray type=\"Int32\" Name=\"types\" format=\"appended\" offset=\"" + std::to_string(offset) + "\"/>");
    for (int32_t n = 0; n < n_cells; ++n) tmp[n] = vtk_type;
    block(tmp.data(), tmp.size() * 4);
  }
  close_tag("Cells");
  auto columns = [&](const VtuColumns& c, size_t rows, bool is_int) {
    for (size_t v = 0; v < c.names.size(); ++v) {
      line(std::string("<DataArray type=\"") + (is_int ? "Int32" : "Float64") + "\" Name=\"" + c.names[v] +
           "\" NumberOfComponents=\"1\" format=\"appended\" offset=\"" + std::to_string(offset) + "\"/>");
      const size_t w = is_int ? 4 : 8;
      block(reinterpret_cast<const char*>(c.data) + v * rows * w, rows * w);
    }
  };
  open_tag("CellData");
  columns(cell_data, (size_t)n_cells, false);
  columns(icell_data, (size_t)n_cells, true);
  close_tag("CellData");
  open_tag("PointData");
  columns(point_data, (size_t)n_points, false);
  close_tag("PointData");
  close_tag("Piece");                                                                               // vtk_unstr_geo_xml_close
  close_tag("UnstructuredGrid");
  line("<AppendedData encoding=\"raw\">");                                                          // vtk_end_xml
  bool ok = std::fwrite(xml.data(), 1, xml.size(), f) == xml.size();
  ok = ok && std::fputc('_', f) != EOF;
  ok = ok && (raw.empty() || std::fwrite(raw.data(), 1, raw.size(), f) == raw.size());
  xml.clear();
  xml += '\n';
  line("</AppendedData>");
  close_tag("VTKFile");
  ok = ok && std::fwrite(xml.data(), 1, xml.size(), f) == xml.size();
  ok = (std::fclose(f) == 0) && ok;
  if (!ok) set_error(std::string("short write to ") + filename);
  return ok;
}

// ------------------------------------------------------------------------------------------------
// kd-tree
// ------------------------------------------------------------------------------------------------
namespace {
struct CRec {
  double c[3];
  int32_t id;   // 0-based original cell
  int32_t pad;
};

// leaves in the left subtree of a complete (left-balanced) binary tree with l >= 2 leaves
inline int leaves_left(int l) {
  if (l == 2) return 1;
  int d = 0;
  while ((1 << d) < l) ++d;                      // d = ceil(log2 l) >= 2
  return std::min(1 << (d - 1), l - (1 << (d - 2)));
}

// nearest floats on either side of a double (outward rounding of the node boxes)
inline float float_below(double x) {
  float f = (float)x;
  if ((double)f > x) f = std::nextafterf(f, -std::numeric_limits<float>::infinity());
  return f;
}
inline float float_above(double x) {
  float f = (float)x;
  if ((double)f < x) f = std::nextafterf(f, std::numeric_limits<float>::infinity());
  return f;
}

struct Builder {
  CRec* rec;
  KdTreeHost* out;
  int n_internal;

  void build(int node, int64_t lo, int64_t hi, int l, int depth) {
    if (l == 1) {
      int block = node - n_internal;
      for (int64_t j = lo; j < hi; ++j) out->orig_of_slot[(size_t)block * LEAF + (j - lo)] = rec[j].id + 1;
      float* b = out->box.data() + 8 * (size_t)node;
      for (int k = 0; k < 3; ++k) {
        double mn = rec[lo].c[k], mx = rec[lo].c[k];
        for (int64_t j = lo + 1; j < hi; ++j) { mn = std::min(mn, rec[j].c[k]); mx = std::max(mx, rec[j].c[k]); }
        b[k] = float_below(mn);
        b[3 + k] = float_above(mx);
      }
      return;
    }
    double mn[3], mx[3];
    for (int k = 0; k < 3; ++k) mn[k] = mx[k] = rec[lo].c[k];
    for (int64_t i = lo + 1; i < hi; ++i)
      for (int k = 0; k < 3; ++k) {
        double v = rec[i].c[k];
        mn[k] = v < mn[k] ? v : mn[k];
        mx[k] = v > mx[k] ? v : mx[k];
      }
    int d = 0;
    for (int k = 1; k < 3; ++k) if (mx[k] - mn[k] > mx[d] - mn[d]) d = k;
    const int ll = leaves_left(l);
    const int64_t mid = lo + (int64_t)ll * LEAF;
    std::nth_element(rec + lo, rec + mid, rec + hi, [d](const CRec& a, const CRec& b) { return a.c[d] < b.c[d]; });
    out->split[node] = rec[mid].c[d];   // all of [lo,mid) <= split <= all of [mid,hi)
    out->sdim[node] = (uint8_t)d;
    if (depth < 4 && hi - lo > (1 << 16)) {
      auto fut = std::async(std::launch::async, [=]() { build(2 * node + 1, lo, mid, ll, depth + 1); });
      build(2 * node + 2, mid, hi, l - ll, depth + 1);
      fut.get();
    } else {
      build(2 * node + 1, lo, mid, ll, depth + 1);
      build(2 * node + 2, mid, hi, l - ll, depth + 1);
    }
  }
};
}  // namespace

void build_kdtree(const double* centers, int32_t n_cells, KdTreeHost& out) {
  const int L = (n_cells + LEAF - 1) / LEAF;
  out.n_leaves = L;
  out.split.assign(L > 1 ? L - 1 : 0, 0.0);
  out.sdim.assign(L > 1 ? L - 1 : 0, 0);
  out.orig_of_slot.assign((size_t)L * LEAF, 0);
  std::vector<CRec> rec((size_t)n_cells);
  for (int32_t i = 0; i < n_cells; ++i) {
    rec[i].c[0] = centers[3 * (size_t)i]; rec[i].c[1] = centers[3 * (size_t)i + 1]; rec[i].c[2] = centers[3 * (size_t)i + 2];
    rec[i].id = i; rec[i].pad = 0;
  }
  out.box.assign(L > 0 ? 8 * (size_t)(2 * L - 1) : 0, 0.0f);
  Builder b{rec.data(), &out, L - 1};
  if (n_cells > 0) b.build(0, 0, n_cells, L, 0);
  for (int i = L - 2; i >= 0; --i) {   // internal nodes: union of the children, bottom-up
    const float* a = out.box.data() + 8 * (size_t)(2 * i + 1);
    const float* c = out.box.data() + 8 * (size_t)(2 * i + 2);
    float* o = out.box.data() + 8 * (size_t)i;
    for (int k = 0; k < 3; ++k) { o[k] = std::min(a[k], c[k]); o[3 + k] = std::max(a[3 + k], c[3 + k]); }
  }
}

}  // namespace iu
