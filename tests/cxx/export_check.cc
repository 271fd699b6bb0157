// Drives dune/ddm_b200/exporter.hh with the reference's own container types (Dune::BCRSMatrix, BlockVector,
// ParallelIndexSet with the OwnerOverlapCopy attributes) and a stand-in for cholmod_factor with cholmod.h's field names
// (SuiteSparse is not in this image).  Input: one rank's share as tests/test_exporter.py lays it out; output: rank<r>.ddmpack
// in the given directory, written by exportRank exactly as a dune-ddm build would.  Test infrastructure only.
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <memory>
#include <string>
#include <vector>

#include <dune/common/fmatrix.hh>
#include <dune/common/fvector.hh>
#include <dune/common/parallel/indexset.hh>
#include <dune/common/parallel/plocalindex.hh>
#include <dune/istl/bcrsmatrix.hh>
#include <dune/istl/bvector.hh>
#include <dune/istl/owneroverlapcopy.hh>

#include <dune/ddm_b200/ddmpack.hh>
#include <dune/ddm_b200/exporter.hh>

using namespace Dune::ddm_b200;

// field names and meanings of cholmod_factor (cholmod.h); index arrays are void* there too
struct MockCholmodFactor {
  std::size_t n = 0, nzmax = 0, nsuper = 0, ssize = 0, xsize = 0;
  void *Perm = nullptr, *p = nullptr, *i = nullptr, *x = nullptr, *nz = nullptr;
  void *super = nullptr, *pi = nullptr, *px = nullptr, *s = nullptr;
  int is_ll = 1, is_super = 1, itype = 2, xtype = 1, dtype = 0;
};

using Attr = Dune::OwnerOverlapCopyAttributeSet::AttributeSet;
using LocalIndex = Dune::ParallelLocalIndex<Attr>;
using IndexSet = Dune::ParallelIndexSet<std::size_t, LocalIndex>;

static void fillIndexSet(IndexSet &is, const PackArray &gid, const PackArray &attr, int b)
{
  const std::int64_t *g = gid.as<std::int64_t>();
  const std::uint8_t *a = attr.as<std::uint8_t>();
  is.beginResize();
  for (std::int64_t l = 0; l < gid.count() / b; ++l) is.add((std::size_t)(g[l * b] / b), LocalIndex((std::size_t)l, (Attr)a[l * b], true));
  is.endResize();
}

template <int b>
static std::shared_ptr<Dune::BCRSMatrix<Dune::FieldMatrix<double, b, b>>> buildMatrix(const DdmPack &p, const std::string &prefix)
{
  using Mat = Dune::BCRSMatrix<Dune::FieldMatrix<double, b, b>>;
  const auto &rpA = p[prefix + "rowptr"];
  const std::int64_t *rp = rpA.as<std::int64_t>();
  const std::int32_t *ci = p[prefix + "col"].as<std::int32_t>();
  const double *va = p[prefix + "val"].as<double>();
  const std::size_t n = (std::size_t)(rpA.count() - 1) / b;
  auto A = std::make_shared<Mat>(n, n, Mat::random);
  std::vector<std::vector<std::size_t>> pat(n);
  for (std::size_t I = 0; I < n; ++I) {
    for (int r = 0; r < b; ++r)
      for (std::int64_t k = rp[I * b + r]; k < rp[I * b + r + 1]; ++k) pat[I].push_back((std::size_t)ci[k] / b);
    std::sort(pat[I].begin(), pat[I].end());
    pat[I].erase(std::unique(pat[I].begin(), pat[I].end()), pat[I].end());
    A->setrowsize(I, pat[I].size());
  }
  A->endrowsizes();
  for (std::size_t I = 0; I < n; ++I)
    for (auto J : pat[I]) A->addindex(I, J);
  A->endindices();
  *A = 0.0;
  for (std::size_t I = 0; I < n; ++I)
    for (int r = 0; r < b; ++r)
      for (std::int64_t k = rp[I * b + r]; k < rp[I * b + r + 1]; ++k) (*A)[I][(std::size_t)ci[k] / b][r][ci[k] % b] = va[k];
  return A;
}

template <int b>
static int run(const DdmPack &p, const std::string &outdir)
{
  using Vec = Dune::BlockVector<Dune::FieldVector<double, b>>;
  auto A = buildMatrix<b>(p, "A_");
  auto Aov = p.has("Aovlp_rowptr") ? buildMatrix<b>(p, "Aovlp_") : nullptr;
  IndexSet novlp, ovlp;
  fillIndexSet(novlp, p["novlp_gid"], p["novlp_attr"], b);
  fillIndexSet(ovlp, p["ovlp_gid"], p["ovlp_attr"], b);
  const std::size_t ni = ovlp.size();

  auto toVec = [](const double *src, std::size_t nodes) {
    Vec v(nodes);
    for (std::size_t i = 0; i < nodes; ++i)
      for (int c = 0; c < b; ++c) v[i][c] = src[i * b + c];
    return v;
  };
  // the reference's partition of unity is a scalar vector over the index set (pou.hh:196-199)
  Dune::BlockVector<Dune::FieldVector<double, 1>> pou(ni);
  const bool have_pou = p.has("pou");
  if (have_pou)
    for (std::size_t i = 0; i < ni; ++i) pou[i] = p["pou"].as<double>()[i * b];
  std::vector<Vec> basis;
  const auto &Z = p["Z"];
  for (std::int64_t k = 0; k < Z.dims.at(0); ++k) basis.push_back(toVec(Z.as<double>() + k * (std::int64_t)ni * b, ni));
  Vec rhs = toVec(p["rhs"].as<double>(), novlp.size());

  MockCholmodFactor L;
  const bool have_factor = p.has("L_Perm");
  std::vector<std::int32_t> i32[8];  // CHOLMOD_INT variant: narrowed copies
  if (have_factor) {
    L.n = ni * b;
    L.is_super = (int)p.scalar("L_is_super");
    L.is_ll = (int)p.scalar("L_is_ll");
    L.itype = (int)p.scalar("L_itype");
    int slot = 0;
    auto idx = [&](const std::string &name) -> void * {
      const auto &a = p[name];
      if (L.itype != 0) return const_cast<std::int64_t *>(a.as<std::int64_t>());
      auto &v = i32[slot++];
      v.assign(a.as<std::int64_t>(), a.as<std::int64_t>() + a.count());
      return v.data();
    };
    L.Perm = idx("L_Perm");
    L.x = const_cast<double *>(p["L_x"].as<double>());
    if (L.is_super) {
      L.super = idx("L_super");
      L.pi = idx("L_pi");
      L.px = idx("L_px");
      L.s = idx("L_s");
      L.nsuper = (std::size_t)p["L_super"].count() - 1;
      L.ssize = (std::size_t)p["L_s"].count();
      L.xsize = (std::size_t)p["L_x"].count();
    } else {
      L.p = idx("L_p");
      L.i = idx("L_i");
      L.nz = idx("L_nz");
      L.nzmax = (std::size_t)p["L_i"].count();
    }
  }
  exportRank(outdir, (int)p.scalar("rank"), (int)p.scalar("nranks"), *A, novlp, ovlp, Aov.get(), have_pou ? &pou : nullptr,
             have_factor ? &L : nullptr, basis, &rhs);
  return 0;
}

int main(int argc, char **argv)
{
  if (argc < 3) {
    std::fprintf(stderr, "usage: export_check <rank input .ddmpack> <output directory>\n");
    return 2;
  }
  try {
    DdmPack p(argv[1]);
    const std::int64_t b = p.scalar("A_block");
    if (b == 1) return run<1>(p, argv[2]);
    if (b == 3) return run<3>(p, argv[2]);
    std::fprintf(stderr, "block size %lld not supported\n", (long long)b);
    return 2;
  } catch (std::exception &e) {
    std::fprintf(stderr, "error: %s\n", e.what());
    return 1;
  }
}
