// npy.hpp -- minimal NumPy .npy (format 1.0/2.0/3.0) reader/writer for the driver CLIs.
//
// File contract kept from the reference (nlsolvers/common/include/util.hpp:36-58, which wraps
// the vendored libnpy): C-order arrays of '<f8' or '<c16'; readers return the flat data and the
// shape; writers emit format 1.0 with fortran_order False.  Errors throw std::runtime_error,
// like libnpy does, so the drivers keep the reference's failure behaviour.
#pragma once
#include <complex>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace nls_host {

template <class T> struct npy_descr;
template <> struct npy_descr<double> { static const char *str() { return "<f8"; } };
template <> struct npy_descr<std::complex<double>> { static const char *str() { return "<c16"; } };

namespace detail {
inline std::string dict_value(const std::string &hdr, const std::string &key) {
  size_t p = hdr.find("'" + key + "'");
  if (p == std::string::npos) throw std::runtime_error("npy: header key missing: " + key);
  p = hdr.find(':', p);
  if (p == std::string::npos) throw std::runtime_error("npy: malformed header");
  ++p;
  while (p < hdr.size() && hdr[p] == ' ') ++p;
  size_t q = p;
  if (hdr[p] == '(') q = hdr.find(')', p) + 1;
  else if (hdr[p] == '\'') q = hdr.find('\'', p + 1) + 1;
  else while (q < hdr.size() && hdr[q] != ',' && hdr[q] != '}') ++q;
  return hdr.substr(p, q - p);
}
} // namespace detail

// reads a C-order array of T; shape returned through `shape`
template <class T>
std::vector<T> read_npy(const std::string &path, std::vector<uint64_t> &shape) {
  FILE *f = std::fopen(path.c_str(), "rb");
  if (!f) throw std::runtime_error("npy: cannot open " + path);
  struct Closer { FILE *f; ~Closer() { std::fclose(f); } } closer{f};
  unsigned char pre[10];
  if (std::fread(pre, 1, 10, f) != 10 || std::memcmp(pre, "\x93NUMPY", 6) != 0)
    throw std::runtime_error("npy: bad magic in " + path);
  size_t hlen;
  if (pre[6] == 1) hlen = pre[8] | (pre[9] << 8);
  else {
    unsigned char more[2];
    if (std::fread(more, 1, 2, f) != 2) throw std::runtime_error("npy: truncated header");
    hlen = pre[8] | (pre[9] << 8) | ((size_t)more[0] << 16) | ((size_t)more[1] << 24);
  }
  std::string hdr(hlen, '\0');
  if (std::fread(&hdr[0], 1, hlen, f) != hlen) throw std::runtime_error("npy: truncated header");
  const std::string descr = detail::dict_value(hdr, "descr");
  if (descr.find(npy_descr<T>::str() + 1) == std::string::npos || descr.find('>') != std::string::npos)
    throw std::runtime_error("npy: " + path + " has dtype " + descr + ", expected '" + npy_descr<T>::str() + "'");
  if (detail::dict_value(hdr, "fortran_order").find("True") != std::string::npos)
    throw std::runtime_error("npy: fortran_order arrays are not supported");
  const std::string shp = detail::dict_value(hdr, "shape");
  shape.clear();
  uint64_t cur = 0;
  bool have = false;
  for (char c : shp) {
    if (c >= '0' && c <= '9') { cur = cur * 10 + (uint64_t)(c - '0'); have = true; }
    else if (have) { shape.push_back(cur); cur = 0; have = false; }
  }
  uint64_t count = 1;
  for (uint64_t s : shape) count *= s;
  std::vector<T> data(count);
  if (count && std::fread(data.data(), sizeof(T), count, f) != count)
    throw std::runtime_error("npy: truncated data in " + path);
  return data;
}

template <class T>
void write_npy(const std::string &path, const T *data, const std::vector<uint64_t> &shape) {
  std::string dict = std::string("{'descr': '") + npy_descr<T>::str() + "', 'fortran_order': False, 'shape': (";
  for (size_t i = 0; i < shape.size(); ++i) dict += std::to_string(shape[i]) + (shape.size() == 1 || i + 1 < shape.size() ? ", " : "");
  dict += "), }";
  size_t total = 10 + dict.size() + 1;
  const size_t pad = (64 - total % 64) % 64;
  dict.append(pad, ' ');
  dict.push_back('\n');
  if (dict.size() > 65535) throw std::runtime_error("npy: header too long");
  FILE *f = std::fopen(path.c_str(), "wb");
  if (!f) throw std::runtime_error("npy: cannot create " + path);
  const unsigned char pre[10] = {0x93, 'N', 'U', 'M', 'P', 'Y', 1, 0, (unsigned char)(dict.size() & 0xff),
                                 (unsigned char)(dict.size() >> 8)};
  uint64_t count = 1;
  for (uint64_t s : shape) count *= s;
  bool ok = std::fwrite(pre, 1, 10, f) == 10 && std::fwrite(dict.data(), 1, dict.size(), f) == dict.size() &&
            (count == 0 || std::fwrite(data, sizeof(T), count, f) == count);
  ok = (std::fclose(f) == 0) && ok;
  if (!ok) throw std::runtime_error("npy: write failed for " + path);
}

} // namespace nls_host
