// Exporter for a real dune-ddm build: every MPI rank dumps what it holds after the preconditioner setup --
// its additive local matrix and index set, the overlapping index set, the partition of unity, the CHOLMOD factor of its
// subdomain matrix, its coarse basis vectors, the right-hand side -- into `<dir>/rank<r>.ddmpack`.  ddm_b200/setup/rankpacks.py
// merges the rank files into the unique-DOF products the GPU library is fed with, so that "setup reused from the reference,
// factors identical" holds literally: the device sweeps then run on CHOLMOD's own L.
//
// Header-only, templated on the Dune types so that it compiles against whatever the host build uses:
//   Mat        Dune::BCRSMatrix<Dune::FieldMatrix<double,b,b>>, b = 1 or 3   (problem.getA().storage(), A_dir; poisson.cc:229,307)
//   IndexSet   Dune::ParallelIndexSet<GlobalIndex, Dune::ParallelLocalIndex<OwnerOverlapCopyAttributeSet::AttributeSet>>
//              (novlp_comm->indexSet(), ovlp_comm->indexSet(); overlap_extension.hh:100-104,260-262)
//   Pou        anything with size() and operator[] convertible to double            (dune/ddm/pou.hh:196-199)
//   Factor     cholmod_factor (Dune::Cholmod::cholmodFactor(), extern/dune-istl/dune/istl/cholmod.hh:495-508)
//   Vec        Dune::BlockVector<Dune::FieldVector<double,b>>                        (basis: poisson.cc:281-284)
// Call site in the reference's driver: examples/poisson.cc after line 306 (schwarz, basis and coarse exist), see INTEGRATION.md.
#ifndef DUNE_DDM_B200_EXPORTER_HH
#define DUNE_DDM_B200_EXPORTER_HH

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <utility>
#include <vector>

namespace Dune::ddm_b200 {

/** Writer of the `.ddmpack` container (layout: ddm_b200/setup/pack.py; reader: ddmpack.hh). */
class DdmPackWriter {
public:
  template <class T>
  void add(const std::string &name, const T *data, std::vector<std::int64_t> dims)
  {
    static_assert(std::is_same_v<T, double> || std::is_same_v<T, std::int32_t> || std::is_same_v<T, std::int64_t> || std::is_same_v<T, std::uint8_t>,
                  "ddmpack holds f64, i32, i64 and u8 arrays");
    Entry e;
    e.name = name;
    e.code = std::is_same_v<T, double> ? 0 : std::is_same_v<T, std::int32_t> ? 1 : std::is_same_v<T, std::int64_t> ? 2 : 3;
    e.dims = std::move(dims);
    std::size_t count = 1;
    for (auto d : e.dims) count *= (std::size_t)d;
    e.bytes.resize(count * sizeof(T));
    if (count) std::memcpy(e.bytes.data(), data, count * sizeof(T));
    entries_.push_back(std::move(e));
  }
  template <class T>
  void add(const std::string &name, const std::vector<T> &v)
  {
    add(name, v.data(), {(std::int64_t)v.size()});
  }
  void addScalar(const std::string &name, std::int64_t v) { add(name, &v, {1}); }

  void write(const std::string &path) const
  {
    std::FILE *fh = std::fopen(path.c_str(), "wb");
    if (!fh) throw std::runtime_error("ddmpack: cannot create " + path);
    const std::int64_t n = (std::int64_t)entries_.size();
    bool ok = std::fwrite("DDMPACK1", 1, 8, fh) == 8 && std::fwrite(&n, 8, 1, fh) == 1;
    for (const auto &e : entries_) {
      const std::int32_t ln = (std::int32_t)e.name.size(), code = e.code, ndim = (std::int32_t)e.dims.size();
      ok = ok && std::fwrite(&ln, 4, 1, fh) == 1 && std::fwrite(e.name.data(), 1, (std::size_t)ln, fh) == (std::size_t)ln;
      ok = ok && std::fwrite(&code, 4, 1, fh) == 1 && std::fwrite(&ndim, 4, 1, fh) == 1;
      if (ndim) ok = ok && std::fwrite(e.dims.data(), 8, (std::size_t)ndim, fh) == (std::size_t)ndim;
      if (!e.bytes.empty()) ok = ok && std::fwrite(e.bytes.data(), 1, e.bytes.size(), fh) == e.bytes.size();
    }
    ok = (std::fclose(fh) == 0) && ok;
    if (!ok) throw std::runtime_error("ddmpack: short write to " + path);
  }

private:
  struct Entry {
    std::string name;
    std::int32_t code = 0;
    std::vector<std::int64_t> dims;
    std::vector<unsigned char> bytes;
  };
  std::vector<Entry> entries_;
};

/** BCRSMatrix<FieldMatrix<double,b,b>> -> scalar CSR (row b*i + r, column b*j + c; columns ascending as BCRSMatrix stores them). */
template <class Mat>
void addMatrix(DdmPackWriter &w, const std::string &prefix, const Mat &A)
{
  using Block = typename Mat::block_type;
  constexpr int b = Block::rows;
  static_assert(Block::rows == Block::cols, "square blocks expected");
  const std::int64_t n = (std::int64_t)A.N() * b;
  std::vector<std::int64_t> rowptr((std::size_t)n + 1, 0);
  std::vector<std::int32_t> col;
  std::vector<double> val;
  col.reserve((std::size_t)A.nonzeroes() * b * b);
  val.reserve((std::size_t)A.nonzeroes() * b * b);
  for (auto row = A.begin(); row != A.end(); ++row)
    for (int r = 0; r < b; ++r) {
      for (auto e = row->begin(); e != row->end(); ++e)
        for (int c = 0; c < b; ++c) {
          col.push_back((std::int32_t)(e.index() * b + c));
          val.push_back((*e)[r][c]);
        }
      rowptr[(std::size_t)(row.index() * b + r) + 1] = (std::int64_t)col.size();
    }
  for (std::size_t i = 1; i < rowptr.size(); ++i)  // rows without entries inherit the running count
    if (rowptr[i] < rowptr[i - 1]) rowptr[i] = rowptr[i - 1];
  w.add(prefix + "rowptr", rowptr);
  w.add(prefix + "col", col);
  w.add(prefix + "val", val);
  w.addScalar(prefix + "block", b);
}

/** Global ids and owner/overlap/copy attributes by local index (b scalar entries per index: id * b + c). */
template <class IndexSet>
void addIndexSet(DdmPackWriter &w, const std::string &prefix, const IndexSet &is, int b = 1)
{
  const std::size_t n = is.size();
  std::vector<std::int64_t> gid(n * (std::size_t)b, -1);
  std::vector<std::uint8_t> attr(n * (std::size_t)b, 0);
  for (const auto &ip : is) {
    const std::size_t l = ip.local().local();
    if (l >= n) throw std::runtime_error("ddm_b200 exporter: local index outside the index set's size");
    for (int c = 0; c < b; ++c) {
      gid[l * b + c] = (std::int64_t)ip.global() * b + c;
      attr[l * b + c] = (std::uint8_t)ip.local().attribute();
    }
  }
  for (auto g : gid)
    if (g < 0) throw std::runtime_error("ddm_b200 exporter: index set does not cover all local indices");
  w.add(prefix + "gid", gid);
  w.add(prefix + "attr", attr);
}

/** BlockVector<FieldVector<double,b>> (or anything with size() and operator[]) flattened. */
template <class Vec>
std::vector<double> flatten(const Vec &v)
{
  std::vector<double> out;
  out.reserve(v.size());
  for (std::size_t i = 0; i < v.size(); ++i) {
    if constexpr (std::is_convertible_v<decltype(v[i]), double>)
      out.push_back((double)v[i]);
    else
      for (std::size_t c = 0; c < v[i].size(); ++c) out.push_back(v[i][c]);
  }
  return out;
}

namespace detail {
template <class I>
std::vector<std::int64_t> widen(const void *p, std::size_t count)
{
  const I *q = static_cast<const I *>(p);
  return std::vector<std::int64_t>(q, q + count);
}
inline std::vector<std::int64_t> indices(const void *p, std::size_t count, int itype)
{
  // cholmod.h: CHOLMOD_INT 0 (int32 everywhere), CHOLMOD_INTLONG 1, CHOLMOD_LONG 2 (SuiteSparse_long; what Dune::Cholmod uses)
  return itype == 0 ? widen<std::int32_t>(p, count) : widen<std::int64_t>(p, count);
}
inline std::vector<std::int32_t> narrow(const std::vector<std::int64_t> &v)
{
  return std::vector<std::int32_t>(v.begin(), v.end());
}
} // namespace detail

/** cholmod_factor -> the supernodal LL^T arrays ddm_upload_factor takes (include/ddm_b200.h).
 *
 * Supernodal LL^T factors (what CHOLMOD's AUTO strategy produces for 3-D problems) are copied as they are.  Simplicial
 * factors (LL^T or LDL^T, packed or not) are regrouped into their fundamental supernodes -- consecutive columns whose row
 * pattern is nested -- and LDL^T is rescaled to LL^T (l_jj = sqrt(d_j)); a non-positive d_j is reported, the matrix was
 * not positive definite.  Real double precision only (xtype CHOLMOD_REAL = 1, dtype CHOLMOD_DOUBLE = 0). */
template <class Factor>
void addCholmodFactor(DdmPackWriter &w, const std::string &prefix, const Factor &L)
{
  if (L.xtype != 1 || L.dtype != 0) throw std::runtime_error("ddm_b200 exporter: the CHOLMOD factor must be real, double precision and numeric");
  const std::size_t n = (std::size_t)L.n;
  w.add(prefix + "perm", detail::narrow(detail::indices(L.Perm, n, L.itype)));
  const double *x = static_cast<const double *>(L.x);
  if (L.is_super) {
    if (!L.is_ll) throw std::runtime_error("ddm_b200 exporter: supernodal CHOLMOD factors are LL^T by construction");
    const std::size_t ns = (std::size_t)L.nsuper;
    w.add(prefix + "super", detail::narrow(detail::indices(L.super, ns + 1, L.itype)));
    w.add(prefix + "pi", detail::indices(L.pi, ns + 1, L.itype));
    w.add(prefix + "px", detail::indices(L.px, ns + 1, L.itype));
    w.add(prefix + "s", detail::narrow(detail::indices(L.s, (std::size_t)L.ssize, L.itype)));
    w.add(prefix + "x", x, {(std::int64_t)L.xsize});
    return;
  }
  // simplicial: column j holds nz[j] entries from p[j], diagonal first, rows ascending
  const auto p = detail::indices(L.p, n + 1, L.itype), nz = detail::indices(L.nz, n, L.itype);
  const auto rows = detail::indices(L.i, (std::size_t)L.nzmax, L.itype);
  std::vector<std::int32_t> super{0}, s;
  std::vector<std::int64_t> pi{0}, px{0};
  std::vector<double> xs;
  std::size_t j = 0;
  while (j < n) {
    // grow the supernode while column k+1's pattern is column k's pattern without its first row
    std::size_t k = j;
    while (k + 1 < n && nz[k + 1] + 1 == nz[k] && rows[(std::size_t)p[k] + 1] == (std::int64_t)(k + 1)) {
      bool nested = true;
      for (std::int64_t t = 0; t < nz[k + 1] && nested; ++t) nested = rows[(std::size_t)(p[k + 1] + t)] == rows[(std::size_t)(p[k] + 1 + t)];
      if (!nested) break;
      ++k;
    }
    const std::size_t c = k - j + 1, m = (std::size_t)nz[j];
    if (m == 0 || rows[(std::size_t)p[j]] != (std::int64_t)j) throw std::runtime_error("ddm_b200 exporter: simplicial factor column without a leading diagonal entry");
    for (std::size_t t = 0; t < m; ++t) s.push_back((std::int32_t)rows[(std::size_t)p[j] + t]);
    const std::size_t base = xs.size();
    xs.resize(base + m * c, 0.0);  // column-major m x c panel, zeros above the diagonal of the leading block
    for (std::size_t q = 0; q < c; ++q) {
      const std::size_t col = j + q;
      const double d = x[(std::size_t)p[col]];
      double scale = 1.0;
      if (!L.is_ll) {
        if (!(d > 0.0)) throw std::runtime_error("ddm_b200 exporter: LDL^T factor with a non-positive pivot (matrix not positive definite)");
        scale = std::sqrt(d);
      }
      xs[base + q * m + q] = L.is_ll ? d : scale;
      for (std::int64_t t = 1; t < nz[col]; ++t) xs[base + q * m + q + (std::size_t)t] = x[(std::size_t)(p[col] + t)] * scale;
    }
    super.push_back((std::int32_t)(k + 1));
    pi.push_back((std::int64_t)s.size());
    px.push_back((std::int64_t)xs.size());
    j = k + 1;
  }
  w.add(prefix + "super", super);
  w.add(prefix + "pi", pi);
  w.add(prefix + "px", px);
  w.add(prefix + "s", s);
  w.add(prefix + "x", xs);
}

/** Everything rank `rank` contributes.  Optional pieces may be null: pou (restricted Schwarz without weights, schwarz.hh:140),
 * factor (non-symmetric problems use UMFPack, whose row pivoting has no counterpart in the symmetric-structure layout: the
 * merger re-factorises A_ovlp), A_ovlp (then the merger forms R_i A R_i^T itself), rhs. */
template <class Mat, class IndexSet, class Pou, class Factor, class Vec>
void exportRank(const std::string &dir, int rank, int nranks, const Mat &A_local, const IndexSet &novlp_indices, const IndexSet &ovlp_indices,
                const Mat *A_ovlp, const Pou *pou, const Factor *factor, const std::vector<Vec> &basis, const Vec *rhs)
{
  constexpr int b = Mat::block_type::rows;
  DdmPackWriter w;
  w.addScalar("rank", rank);
  w.addScalar("nranks", nranks);
  addMatrix(w, "A_", A_local);
  addIndexSet(w, "novlp_", novlp_indices, b);
  addIndexSet(w, "ovlp_", ovlp_indices, b);
  if (A_local.N() != novlp_indices.size()) throw std::runtime_error("ddm_b200 exporter: local matrix and non-overlapping index set differ in size");
  const std::size_t ni = ovlp_indices.size() * (std::size_t)b;
  if (A_ovlp) {
    if (A_ovlp->N() * (std::size_t)b != ni) throw std::runtime_error("ddm_b200 exporter: overlapping matrix and index set differ in size");
    addMatrix(w, "Aovlp_", *A_ovlp);
  }
  if (pou) {
    const auto d = flatten(*pou);
    // a scalar partition of unity weighs all b components of a node alike
    std::vector<double> full;
    if (d.size() * (std::size_t)b == ni && b > 1)
      for (double v : d) full.insert(full.end(), (std::size_t)b, v);
    else
      full = d;
    if (full.size() != ni) throw std::runtime_error("ddm_b200 exporter: partition of unity and overlapping index set differ in size");
    w.add("pou", full);
  }
  if (factor) {
    if ((std::size_t)factor->n != ni) throw std::runtime_error("ddm_b200 exporter: factor and overlapping index set differ in size");
    addCholmodFactor(w, "F_", *factor);
  }
  std::vector<double> Z;
  Z.reserve(ni * basis.size());
  for (const auto &v : basis) {  // column-major n_i x k_i, the layout of ddm_upload_coarse
    const auto col = flatten(v);
    if (col.size() != ni) throw std::runtime_error("ddm_b200 exporter: basis vector and overlapping index set differ in size");
    Z.insert(Z.end(), col.begin(), col.end());
  }
  w.add("Z", Z.data(), {(std::int64_t)basis.size(), (std::int64_t)ni});
  if (rhs) {
    const auto r = flatten(*rhs);
    if (r.size() != A_local.N() * (std::size_t)b) throw std::runtime_error("ddm_b200 exporter: right-hand side and local matrix differ in size");
    w.add("rhs", r);
  }
  w.write(dir + "/rank" + std::to_string(rank) + ".ddmpack");
}

} // namespace Dune::ddm_b200
#endif
