// vtk_writer.cpp -- the files of mesh::vtk::Writer<Part>::WriteSolutions (/root/reference/include/mini/mesh/
// vtk.hpp:385-487) from the points and values the device sampled (output_kernels.cu): one
// <directory>/<soln_name>/<rank>.vtu per part and, on rank 0, <directory>/<soln_name>.pvtu naming the pieces.
// Same XML, same number formatting (ostream default, 6 significant digits) and the same base64 point block
// (a UInt64 byte count encoded on its own, then the raw little-endian doubles), so ParaView and a `diff`
// against the reference's files see the same thing.
#include <sys/stat.h>

#include <algorithm>
#include <cerrno>
#include <charconv>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <string>
#include <thread>
#include <vector>

#include "solver.hpp"

namespace mcfd {

namespace {

// RFC 4648 base64 with '=' padding (what vtk.hpp:16-57 produces)
std::string Base64(const unsigned char *in, size_t n) {
  static const char tab[] = "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789+/";
  std::string out;
  out.reserve((n + 2) / 3 * 4);
  size_t i = 0;
  for (; i + 3 <= n; i += 3) {
    const unsigned v = (unsigned(in[i]) << 16) | (unsigned(in[i + 1]) << 8) | unsigned(in[i + 2]);
    out.push_back(tab[v >> 18]); out.push_back(tab[(v >> 12) & 63]); out.push_back(tab[(v >> 6) & 63]); out.push_back(tab[v & 63]);
  }
  if (i < n) {
    const unsigned a = in[i], b = i + 1 < n ? in[i + 1] : 0u;
    const unsigned v = (a << 16) | (b << 8);
    out.push_back(tab[v >> 18]); out.push_back(tab[(v >> 12) & 63]);
    out.push_back(i + 1 < n ? tab[(v >> 6) & 63] : '=');
    out.push_back('=');
  }
  return out;
}

bool MakeDirs(const std::string &path) {  // `mkdir -p` (part.hpp:1553-1556 shells out for it)
  std::string cur;
  for (size_t i = 0; i <= path.size(); ++i) {
    if (i == path.size() || path[i] == '/') {
      if (!cur.empty() && mkdir(cur.c_str(), 0777) != 0 && errno != EEXIST) return false;
    }
    if (i < path.size()) cur.push_back(path[i]);
  }
  return true;
}

// Numbers as `ostream << double` prints them with the default precision: printf's %.6g, which is what std::to_chars with
// chars_format::general and precision 6 is specified to produce -- without the stream machinery per value (a 64^3 mesh has
// 35 M of them).
class NumberText {
 public:
  explicit NumberText(std::ofstream *out) : out_(out) { buffer_.reserve(kFlush + 64); }
  ~NumberText() { Flush(); }
  // the five point-data arrays are independent: each is formatted by its own host thread into one string
  static std::string OfColumn(const double *values, int64_t n, int stride) {
    std::string text;
    text.reserve(size_t(n) * 10);
    char tmp[48];
    for (int64_t i = 0; i < n; ++i) {
      auto res = std::to_chars(tmp, tmp + sizeof(tmp), values[i * stride], std::chars_format::general, 6);
      text.append(tmp, res.ptr);
      text.push_back(' ');
    }
    return text;
  }
  void Add(double v) {
    char tmp[48];
    auto res = std::to_chars(tmp, tmp + sizeof(tmp), v, std::chars_format::general, 6);
    buffer_.append(tmp, res.ptr);
    buffer_.push_back(' ');
    if (buffer_.size() >= kFlush) Flush();
  }
  void Add(int64_t v) {
    char tmp[24];
    auto res = std::to_chars(tmp, tmp + sizeof(tmp), v);
    buffer_.append(tmp, res.ptr);
    buffer_.push_back(' ');
    if (buffer_.size() >= kFlush) Flush();
  }
  void Flush() { out_->write(buffer_.data(), std::streamsize(buffer_.size())); buffer_.clear(); }

 private:
  static constexpr size_t kFlush = size_t(1) << 20;
  std::ofstream *out_;
  std::string buffer_;
};

const char *const kEulerFieldNames[5] = { "Density", "MomentumX", "MomentumY", "MomentumZ", "EnergyStagnationDensity" };

}  // namespace

int WriteVtuFiles(const char *directory, const char *soln_name, const char *const *field_names, int rank, int world,
                  int64_t n_cells, const double *coords, const double *values, std::string *error) {
  if (!field_names) field_names = kEulerFieldNames;  // demo/euler/sourceless.cpp:72-73
  const std::string dir = directory, name = soln_name;
  const char *endianness = "\"LittleEndian\"";
  if (!MakeDirs(dir + "/" + name)) { *error = "cannot create " + dir + "/" + name + ": " + std::strerror(errno); return MCFD_ERR_ARG; }
  if (rank == 0) {
    std::ofstream pvtu(dir + "/" + name + ".pvtu", std::ios::out);
    pvtu << "<VTKFile type=\"PUnstructuredGrid\" version=\"1.0\" "
         << "byte_order=" << endianness << " header_type=\"UInt64\">\n";
    pvtu << "  <PUnstructuredGrid GhostLevel=\"1\">\n";
    pvtu << "    <PPointData>\n";
    for (int k = 0; k < MCFD_K; ++k) pvtu << "      <PDataArray type=\"Float64\" Name=\"" << field_names[k] << "\"/>\n";
    pvtu << "    </PPointData>\n";
    pvtu << "    <PPoints>\n";
    pvtu << "      <PDataArray type=\"Float64\" Name=\"Points\" " << "NumberOfComponents=\"3\"/>\n";
    pvtu << "    </PPoints>\n";
    for (int i_part = 0; i_part < world; ++i_part) pvtu << "    <Piece Source=\"./" << name << '/' << i_part << ".vtu\"/>\n";
    pvtu << "  </PUnstructuredGrid>\n";
    pvtu << "</VTKFile>\n";
    if (!pvtu) { *error = "cannot write " + dir + "/" + name + ".pvtu"; return MCFD_ERR_ARG; }
  }
  const int64_t n_points = 27 * n_cells;
  const std::string path = dir + "/" + name + "/" + std::to_string(rank) + ".vtu";
  std::vector<char> buffer(1 << 22);
  std::ofstream vtu;
  vtu.rdbuf()->pubsetbuf(buffer.data(), std::streamsize(buffer.size()));
  vtu.open(path, std::ios::out);
  const char *format = "\"ascii\"";
  vtu << "<VTKFile type=\"UnstructuredGrid\" version=\"1.0\"" << " byte_order=" << endianness << " header_type=\"UInt64\">\n";
  vtu << "  <UnstructuredGrid>\n";
  vtu << "    <Piece NumberOfPoints=\"" << n_points << "\" NumberOfCells=\"" << n_cells << "\">\n";
  vtu << "      <PointData>\n";
  {
    std::string text[MCFD_K];
    std::vector<std::thread> workers;
    for (int k = 0; k < MCFD_K; ++k)
      workers.emplace_back([&, k] { text[k] = NumberText::OfColumn(values + k, n_points, MCFD_K); });
    for (auto &w : workers) w.join();
    for (int k = 0; k < MCFD_K; ++k) {
      vtu << "        <DataArray type=\"Float64\" Name=\"" << field_names[k] << "\" format=" << format << ">\n";
      vtu.write(text[k].data(), std::streamsize(text[k].size()));
      vtu << "\n        </DataArray>\n";
    }
  }
  vtu << "      </PointData>\n";
  vtu << "      <DataOnCells>\n";
  vtu << "      </DataOnCells>\n";
  vtu << "      <Points>\n";
  vtu << "        <DataArray type=\"Float64\" Name=\"Points\" " << "NumberOfComponents=\"3\" format=" << "\"binary\"" << ">\n";
  {
    const uint64_t n_char = uint64_t(n_points) * 3 * sizeof(double);
    std::string head = Base64(reinterpret_cast<const unsigned char *>(&n_char), sizeof(n_char));
    while (head.size() < 8) head.push_back('=');  // vtk.hpp:441-443
    vtu << head;
    // in pieces that are multiples of 3 bytes, so the concatenation equals one encoding of the whole block
    const unsigned char *bytes = reinterpret_cast<const unsigned char *>(coords);
    const size_t piece = size_t(3) << 20;
    for (size_t off = 0; off < n_char; off += piece) vtu << Base64(bytes + off, std::min(piece, size_t(n_char) - off));
  }
  vtu << "\n        </DataArray>\n";
  vtu << "      </Points>\n";
  vtu << "      <Cells>\n";
  vtu << "        <DataArray type=\"Int32\" Name=\"connectivity\" " << "format=\"ascii\">\n";
  { NumberText text(&vtu); for (int64_t i = 0; i < n_points; ++i) text.Add(i); }
  vtu << "\n        </DataArray>\n";
  vtu << "        <DataArray type=\"Int32\" Name=\"offsets\" " << "format=" << format << ">\n";
  { NumberText text(&vtu); for (int64_t c = 1; c <= n_cells; ++c) text.Add(27 * c); }
  vtu << "\n        </DataArray>\n";
  vtu << "        <DataArray type=\"UInt8\" Name=\"types\" " << "format=" << format << ">\n";
  { NumberText text(&vtu); for (int64_t c = 0; c < n_cells; ++c) text.Add(int64_t(29)); }  // CellType::kHexahedron27 (vtk.hpp:78)
  vtu << "\n        </DataArray>\n";
  vtu << "      </Cells>\n";
  vtu << "    </Piece>\n";
  vtu << "  </UnstructuredGrid>\n";
  vtu << "</VTKFile>\n";
  vtu.close();
  if (!vtu) { *error = "cannot write " + path; return MCFD_ERR_ARG; }
  return MCFD_OK;
}

}  // namespace mcfd
