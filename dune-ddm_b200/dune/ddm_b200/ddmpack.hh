// Reader of the `.ddmpack` container of setup products (layout: ddm_b200/setup/pack.py) and the upload of a
// whole problem into one ddm_ctx.  Header-only, no dependencies beyond the C ABI.
#ifndef DUNE_DDM_B200_DDMPACK_HH
#define DUNE_DDM_B200_DDMPACK_HH

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace Dune::ddm_b200 {

struct PackArray {
  int dtype = 0; // 0 f64, 1 i32, 2 i64, 3 u8
  std::vector<std::int64_t> dims;
  std::vector<unsigned char> bytes;
  std::int64_t count() const
  {
    std::int64_t c = 1;
    for (auto d : dims) c *= d;
    return c;
  }
  template <class T>
  const T *as() const
  {
    return reinterpret_cast<const T *>(bytes.data());
  }
};

class DdmPack {
public:
  explicit DdmPack(const std::string &path)
  {
    std::FILE *fh = std::fopen(path.c_str(), "rb");
    if (!fh) throw std::runtime_error("cannot open " + path);
    char magic[8];
    if (std::fread(magic, 1, 8, fh) != 8 || std::memcmp(magic, "DDMPACK1", 8) != 0) {
      std::fclose(fh);
      throw std::runtime_error(path + " is not a .ddmpack file");
    }
    std::int64_t n = 0;
    rd(fh, &n, 8);
    static const int width[4] = {8, 4, 8, 1};
    for (std::int64_t a = 0; a < n; ++a) {
      std::int32_t ln = 0, code = 0, ndim = 0;
      rd(fh, &ln, 4);
      std::string name(ln, '\0');
      rd(fh, name.data(), ln);
      rd(fh, &code, 4);
      rd(fh, &ndim, 4);
      PackArray arr;
      arr.dtype = code;
      arr.dims.resize(ndim);
      rd(fh, arr.dims.data(), 8 * ndim);
      arr.bytes.resize((std::size_t)arr.count() * width[code]);
      rd(fh, arr.bytes.data(), arr.bytes.size());
      arrays_.emplace(std::move(name), std::move(arr));
    }
    std::fclose(fh);
  }

  const PackArray &operator[](const std::string &name) const
  {
    auto it = arrays_.find(name);
    if (it == arrays_.end()) throw std::runtime_error("ddmpack: no array '" + name + "'");
    return it->second;
  }
  bool has(const std::string &name) const { return arrays_.count(name) != 0; }
  std::int64_t scalar(const std::string &name) const { return (*this)[name].as<std::int64_t>()[0]; }

private:
  static void rd(std::FILE *fh, void *dst, std::size_t n)
  {
    if (n && std::fread(dst, 1, n, fh) != n) throw std::runtime_error("ddmpack: truncated file");
  }
  std::map<std::string, PackArray> arrays_;
};

} // namespace Dune::ddm_b200
#endif
