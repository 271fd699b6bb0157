// Reader of the `.ddmpack` container of setup products (layout: ddm_b200/setup/pack.py) and the upload of a
// whole problem into one ddm_ctx.  Header-only, no dependencies beyond the C ABI.
#ifndef DUNE_DDM_B200_DDMPACK_HH
#define DUNE_DDM_B200_DDMPACK_HH

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <memory>
#include <type_traits>
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
  // Typed view; the stored dtype must be the requested one (a foreign or damaged pack must not hand wrongly typed pointers
  // to the upload calls).
  template <class T>
  const T *as() const
  {
    const int want = std::is_same<T, double>::value ? 0 : std::is_same<T, std::int32_t>::value ? 1 : std::is_same<T, std::int64_t>::value ? 2
                     : (std::is_same<T, std::uint8_t>::value || std::is_same<T, unsigned char>::value || std::is_same<T, char>::value) ? 3 : -1;
    if (want != dtype) throw std::runtime_error("ddmpack: array read with the wrong element type");
    return reinterpret_cast<const T *>(bytes.data());
  }
};

class DdmPack {
public:
  explicit DdmPack(const std::string &path)
  {
    struct Closer {
      void operator()(std::FILE *f) const { if (f) std::fclose(f); }
    };
    std::unique_ptr<std::FILE, Closer> file(std::fopen(path.c_str(), "rb"));  // closed on every path, also when rd() throws
    std::FILE *fh = file.get();
    if (!fh) throw std::runtime_error("cannot open " + path);
    std::fseek(fh, 0, SEEK_END);
    const long long fsize = std::ftell(fh);
    std::fseek(fh, 0, SEEK_SET);
    char magic[8];
    if (std::fread(magic, 1, 8, fh) != 8 || std::memcmp(magic, "DDMPACK1", 8) != 0)
      throw std::runtime_error(path + " is not a .ddmpack file");
    std::int64_t n = 0;
    rd(fh, &n, 8);
    if (n < 0 || n > (1 << 24)) throw std::runtime_error("ddmpack: implausible array count");
    static const int width[4] = {8, 4, 8, 1};
    for (std::int64_t a = 0; a < n; ++a) {
      std::int32_t ln = 0, code = 0, ndim = 0;
      rd(fh, &ln, 4);
      if (ln < 0 || ln > 4096) throw std::runtime_error("ddmpack: implausible name length");
      std::string name(ln, '\0');
      rd(fh, name.data(), ln);
      rd(fh, &code, 4);
      rd(fh, &ndim, 4);
      if (code < 0 || code > 3) throw std::runtime_error("ddmpack: unknown element type in array '" + name + "'");
      if (ndim < 0 || ndim > 8) throw std::runtime_error("ddmpack: implausible rank of array '" + name + "'");
      PackArray arr;
      arr.dtype = code;
      arr.dims.resize(ndim);
      rd(fh, arr.dims.data(), 8 * (std::size_t)ndim);
      const long long left = fsize - std::ftell(fh);
      long long cnt = 1;
      for (auto d : arr.dims) {
        if (d < 0 || (d > 0 && cnt > left / d)) throw std::runtime_error("ddmpack: array '" + name + "' is larger than the file");
        cnt *= d;
      }
      if (cnt > left / width[code]) throw std::runtime_error("ddmpack: array '" + name + "' is larger than the file");
      arr.bytes.resize((std::size_t)cnt * width[code]);
      rd(fh, arr.bytes.data(), arr.bytes.size());
      arrays_.emplace(std::move(name), std::move(arr));
    }
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
