// Host-side helpers of the drop-in command-line programs (C++11, no dependencies):
// option parsing with the reference's flag names (util/text_cmd.hxx:11-37 wraps
// boost::program_options there), image readers for the formats usable without ITK
// (.npy, MetaImage .mha/.mhd with raw data), and the reference's text formats
// (util/text_io.hxx:40-133, type/tuple.hxx:24-28).
#pragma once
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <fstream>
#include <iostream>
#include <map>
#include <sstream>
#include <string>
#include <vector>

namespace gh {

// glia_base.hxx:66-69 perr: message to stderr, exit(EXIT_FAILURE)
inline void perr(const std::string& msg) { std::cerr << msg << std::endl; exit(EXIT_FAILURE); }

// ---- options: "--name v [v ...]" / "-n v"; repeated options accumulate (vector semantics)
struct Args {
  std::map<std::string, std::vector<std::string>> kv;
  std::map<std::string, std::string> alias;  // short -> long
  void add_alias(const std::string& s, const std::string& l) { alias[s] = l; }
  bool parse(int argc, char** argv) {
    std::string cur;
    for (int i = 1; i < argc; ++i) {
      std::string a = argv[i];
      bool isNum = a.size() > 1 && a[0] == '-' && (isdigit((unsigned char)a[1]) || a[1] == '.');
      if (a.size() > 2 && a[0] == '-' && a[1] == '-') { cur = a.substr(2); kv[cur]; }
      else if (a.size() > 1 && a[0] == '-' && !isNum) {
        std::string s = a.substr(1);
        cur = alias.count(s) ? alias[s] : s;
        kv[cur];
      } else {
        if (cur.empty()) return false;
        kv[cur].push_back(a);
      }
    }
    return true;
  }
  bool has(const std::string& k) const { return kv.count(k) > 0; }
  std::vector<std::string> strs(const std::string& k) const { auto it = kv.find(k); return it == kv.end() ? std::vector<std::string>() : it->second; }
  std::string str(const std::string& k, const std::string& d = "") const { auto v = strs(k); return v.empty() ? d : v.back(); }
  std::vector<double> nums(const std::string& k) const { std::vector<double> r; for (auto& s : strs(k)) r.push_back(atof(s.c_str())); return r; }
  bool flag(const std::string& k, bool d = false) const {
    if (!has(k)) return d;
    auto v = strs(k);
    if (v.empty()) return true;
    return v.back() == "1" || v.back() == "true" || v.back() == "True";
  }
  void require(const std::string& k) const { if (!has(k) || strs(k).empty()) perr("Error: unable to parse input arguments (missing --" + k + ")"); }
};

// ---- images
struct Image {
  int ndim = 0;
  int dims[3] = {1, 1, 1};  // x (fastest), y, z
  std::vector<float> f32;
  std::vector<uint32_t> u32;
  size_t size() const { return (size_t)dims[0] * dims[1] * dims[2]; }
};

template <class T>
static void convert_raw(const std::string& type, const unsigned char* raw, size_t n, std::vector<T>& out, const std::string& file) {
  out.resize(n);
#define GH_CONV(tname, ctype) if (type == tname) { const ctype* p = (const ctype*)raw; for (size_t i = 0; i < n; ++i) out[i] = (T)p[i]; return; }
  GH_CONV("u1", uint8_t) GH_CONV("i1", int8_t) GH_CONV("u2", uint16_t) GH_CONV("i2", int16_t)
  GH_CONV("u4", uint32_t) GH_CONV("i4", int32_t) GH_CONV("u8", uint64_t) GH_CONV("i8", int64_t)
  GH_CONV("f4", float) GH_CONV("f8", double)
#undef GH_CONV
  perr("Error: unsupported pixel type '" + type + "' in " + file);
}
static size_t type_size(const std::string& t) { return (size_t)atoi(t.substr(1).c_str()); }

static std::vector<unsigned char> read_file(const std::string& file) {
  std::ifstream is(file, std::ios::binary);
  if (!is) perr("Error: cannot open file " + file);
  return std::vector<unsigned char>((std::istreambuf_iterator<char>(is)), std::istreambuf_iterator<char>());
}

// numpy .npy v1/v2, C order, little endian; shape ([z,] y, x)
template <class T>
static void read_npy(const std::string& file, Image& im, std::vector<T>& out) {
  std::vector<unsigned char> buf = read_file(file);
  if (buf.size() < 12 || memcmp(buf.data(), "\x93NUMPY", 6) != 0) perr("Error: not an npy file: " + file);
  size_t hlen, off;
  if (buf[6] == 1) { hlen = buf[8] | (buf[9] << 8); off = 10; } else { hlen = buf[8] | (buf[9] << 8) | (buf[10] << 16) | ((size_t)buf[11] << 24); off = 12; }
  std::string hdr((const char*)buf.data() + off, hlen);
  size_t d = hdr.find("'descr'"); size_t q1 = hdr.find('\'', d + 7); size_t q2 = hdr.find('\'', q1 + 1);
  std::string descr = hdr.substr(q1 + 1, q2 - q1 - 1);
  if (descr.size() < 3 || (descr[0] != '<' && descr[0] != '|')) perr("Error: unsupported npy dtype " + descr + " in " + file);
  if (hdr.find("'fortran_order': True") != std::string::npos) perr("Error: fortran-order npy unsupported: " + file);
  size_t s1 = hdr.find('(', hdr.find("'shape'")), s2 = hdr.find(')', s1);
  std::vector<int> shape; std::stringstream ss(hdr.substr(s1 + 1, s2 - s1 - 1)); std::string tok;
  while (std::getline(ss, tok, ',')) { if (tok.find_first_of("0123456789") != std::string::npos) shape.push_back(atoi(tok.c_str())); }
  if (shape.size() < 2 || shape.size() > 3) perr("Error: npy image must be 2D or 3D: " + file);
  im.ndim = (int)shape.size();
  for (int i = 0; i < im.ndim; ++i) im.dims[i] = shape[im.ndim - 1 - i];
  std::string t = descr.substr(1);
  size_t n = im.size();
  if (buf.size() < off + hlen + n * type_size(t)) perr("Error: truncated npy file " + file);
  convert_raw(t, buf.data() + off + hlen, n, out, file);
}

// MetaImage (.mha single file / .mhd + raw), uncompressed little-endian
template <class T>
static void read_meta(const std::string& file, Image& im, std::vector<T>& out) {
  std::vector<unsigned char> buf = read_file(file);
  size_t pos = 0; std::map<std::string, std::string> kv; std::string dataFile;
  while (pos < buf.size()) {
    size_t e = pos; while (e < buf.size() && buf[e] != '\n') ++e;
    std::string line((const char*)buf.data() + pos, e - pos); pos = e + 1;
    size_t eq = line.find('=');
    if (eq == std::string::npos) continue;
    auto trim = [](std::string s) { size_t a = s.find_first_not_of(" \t\r"), b = s.find_last_not_of(" \t\r"); return a == std::string::npos ? std::string() : s.substr(a, b - a + 1); };
    std::string k = trim(line.substr(0, eq)), v = trim(line.substr(eq + 1));
    kv[k] = v;
    if (k == "ElementDataFile") { dataFile = v; break; }
  }
  if (kv["CompressedData"] == "True") perr("Error: compressed MetaImage unsupported: " + file);
  if (kv.count("BinaryDataByteOrderMSB") && kv["BinaryDataByteOrderMSB"] == "True") perr("Error: big-endian MetaImage unsupported: " + file);
  im.ndim = atoi(kv["NDims"].c_str());
  if (im.ndim < 2 || im.ndim > 3) perr("Error: MetaImage must be 2D or 3D: " + file);
  std::stringstream ss(kv["DimSize"]);
  for (int i = 0; i < im.ndim; ++i) ss >> im.dims[i];
  static const std::map<std::string, std::string> types = {{"MET_UCHAR", "u1"}, {"MET_CHAR", "i1"}, {"MET_USHORT", "u2"}, {"MET_SHORT", "i2"}, {"MET_UINT", "u4"}, {"MET_INT", "i4"}, {"MET_ULONG", "u8"}, {"MET_LONG", "i8"}, {"MET_ULONG_LONG", "u8"}, {"MET_LONG_LONG", "i8"}, {"MET_FLOAT", "f4"}, {"MET_DOUBLE", "f8"}};
  auto it = types.find(kv["ElementType"]);
  if (it == types.end()) perr("Error: unsupported ElementType in " + file);
  size_t n = im.size();
  if (dataFile == "LOCAL") {
    if (buf.size() < pos + n * type_size(it->second)) perr("Error: truncated MetaImage " + file);
    convert_raw(it->second, buf.data() + pos, n, out, file);
  } else {
    size_t slash = file.find_last_of('/');
    std::string path = (slash == std::string::npos ? std::string() : file.substr(0, slash + 1)) + dataFile;
    std::vector<unsigned char> raw = read_file(path);
    if (raw.size() < n * type_size(it->second)) perr("Error: truncated raw file " + path);
    convert_raw(it->second, raw.data(), n, out, file);
  }
}

static bool ends_with(const std::string& s, const std::string& e) { return s.size() >= e.size() && s.compare(s.size() - e.size(), e.size(), e) == 0; }

inline Image read_label_image(const std::string& file) {
  Image im;
  if (ends_with(file, ".npy")) read_npy(file, im, im.u32);
  else if (ends_with(file, ".mha") || ends_with(file, ".mhd")) read_meta(file, im, im.u32);
  else perr("Error: unsupported image format (use .npy, .mha or .mhd): " + file);
  return im;
}
inline Image read_real_image(const std::string& file) {
  Image im;
  if (ends_with(file, ".npy")) read_npy(file, im, im.f32);
  else if (ends_with(file, ".mha") || ends_with(file, ".mhd")) read_meta(file, im, im.f32);
  else perr("Error: unsupported image format (use .npy, .mha or .mhd): " + file);
  return im;
}

// label image out: .npy (u4 / u2) or uncompressed .mha (MET_UINT / MET_USHORT), same axis conventions as the readers
inline void write_label_image(const std::string& file, const Image& like, const uint32_t* data, bool as16) {
  const size_t n = like.size();
  std::vector<uint16_t> v16;
  if (as16) { v16.resize(n); for (size_t i = 0; i < n; ++i) v16[i] = (uint16_t)data[i]; }
  const void* raw = as16 ? (const void*)v16.data() : (const void*)data;
  const size_t bytes = n * (as16 ? 2 : 4);
  std::ofstream os(file, std::ios::binary);
  if (!os) perr("Error: cannot write file " + file);
  if (ends_with(file, ".npy")) {
    std::stringstream shape;
    for (int i = like.ndim - 1; i >= 0; --i) shape << like.dims[i] << (i ? ", " : "");
    std::string hdr = std::string("{'descr': '") + (as16 ? "<u2" : "<u4") + "', 'fortran_order': False, 'shape': (" + shape.str() + "), }";
    while ((10 + hdr.size() + 1) % 64) hdr += ' ';
    hdr += '\n';
    const unsigned char pre[10] = {0x93, 'N', 'U', 'M', 'P', 'Y', 1, 0, (unsigned char)(hdr.size() & 255), (unsigned char)(hdr.size() >> 8)};
    os.write((const char*)pre, 10); os.write(hdr.data(), (std::streamsize)hdr.size());
  } else if (ends_with(file, ".mha")) {
    os << "ObjectType = Image\nNDims = " << like.ndim << "\nBinaryData = True\nBinaryDataByteOrderMSB = False\nCompressedData = False\nDimSize =";
    for (int i = 0; i < like.ndim; ++i) os << " " << like.dims[i];
    os << "\nElementType = " << (as16 ? "MET_USHORT" : "MET_UINT") << "\nElementDataFile = LOCAL\n";
  } else perr("Error: unsupported output image format (use .npy or .mha): " + file);
  os.write((const char*)raw, (std::streamsize)bytes);
}

// ---- text files
// order: one "r0 r1 r2" per line (TTriple operator<<, type/tuple.hxx:24-28)
inline std::vector<uint32_t> read_order(const std::string& file) {
  std::ifstream is(file);
  if (!is) perr("Error: cannot open file " + file);
  std::vector<uint32_t> v; unsigned long long x;
  while (is >> x) v.push_back((uint32_t)x);
  if (v.size() % 3) perr("Error: invalid data file dimension in " + file);
  return v;
}
inline std::vector<double> read_values(const std::string& file) {
  std::ifstream is(file);
  if (!is) perr("Error: cannot open file " + file);
  std::vector<double> v; double x;
  while (is >> x) v.push_back(x);
  return v;
}
// rows of a feature file: space/newline separated, every line one row (util/text_io.hxx:137-180)
inline std::vector<double> read_rows(const std::string& file, size_t& nrows, size_t& ncols) {
  std::ifstream is(file);
  if (!is) perr("Error: cannot open file " + file);
  std::vector<double> v; std::string line; nrows = 0; ncols = 0;
  while (std::getline(is, line)) {
    std::stringstream ss(line); double x; size_t c = 0;
    while (ss >> x) { v.push_back(x); ++c; }
    if (!c) continue;
    if (ncols && c != ncols) perr("Error: invalid data file dimension in " + file);
    ncols = c; ++nrows;
  }
  return v;
}
inline void write_order(const std::string& file, const uint32_t* o, size_t n) {
  std::ofstream os(file);
  if (!os) perr("Error: cannot create file " + file);
  for (size_t i = 0; i < n; ++i) os << o[3 * i] << " " << o[3 * i + 1] << " " << o[3 * i + 2] << "\n";
}
// one value per line; precision <= 0 keeps the stream default of 6 (main_merge_order_bc.cxx:149)
inline void write_values(const std::string& file, const double* v, size_t n, int precision) {
  std::ofstream os(file);
  if (!os) perr("Error: cannot create file " + file);
  if (precision > 0) os.precision(precision);
  for (size_t i = 0; i < n; ++i) os << v[i] << "\n";
}
// every value followed by " ", every row by "\n" (util/text_io.hxx:120-133)
inline void write_rows(const std::string& file, const double* v, size_t nrows, size_t ncols, int precision) {
  std::ofstream os(file);
  if (!os) perr("Error: cannot create file " + file);
  if (precision > 0) os.precision(precision);
  for (size_t i = 0; i < nrows; ++i) {
    for (size_t j = 0; j < ncols; ++j) os << v[i * ncols + j] << " ";
    os << "\n";
  }
}

}  // namespace gh
