// Host-logic check of Dune::ddm_b200::Context::fromPack without a GPU: this file DEFINES the C-ABI upload entry points as
// recorders (it is linked instead of libddm_b200.so), runs fromPack on a .ddmpack and writes what the adaptor handed to
// the library into a second .ddmpack-like dump that tests/test_host.py compares with scipy's view of the same problem.
// Test infrastructure only.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include <dune/ddm_b200/adaptors.hh>

struct ddm_ctx {
  int dummy;
};
struct ddm_vec {
  int dummy;
};

namespace {
struct Rec {
  std::string name;
  std::vector<unsigned char> bytes;
};
std::vector<Rec> g_rec;
ddm_ctx g_ctx;

template <class T>
void rec(const std::string &name, const T *p, std::size_t count)
{
  Rec r{name, {}};
  r.bytes.resize(count * sizeof(T));
  if (count) std::memcpy(r.bytes.data(), p, count * sizeof(T));
  g_rec.push_back(std::move(r));
}
void rec1(const std::string &name, std::int64_t v) { rec(name, &v, 1); }
} // namespace

extern "C" {
const char *ddm_last_error(void) { return "stub"; }
int ddm_ctx_create(int, ddm_ctx **out)
{
  *out = &g_ctx;
  return 0;
}
int ddm_ctx_destroy(ddm_ctx *) { return 0; }
int ddm_vec_create(ddm_ctx *, ddm_vec **) { return 1; }
int ddm_vec_destroy(ddm_vec *) { return 0; }
int ddm_upload_operator(ddm_ctx *, int64_t n_own, int64_t n_ghost, const int64_t *rowptr, const int32_t *col, const double *val)
{
  rec1("op_kind", 1);
  rec1("op_n_own", n_own);
  rec1("op_n_ghost", n_ghost);
  rec("op_rowptr", rowptr, (std::size_t)n_own + 1);
  rec("op_col", col, (std::size_t)rowptr[n_own]);
  rec("op_val", val, (std::size_t)rowptr[n_own]);
  return 0;
}
int ddm_upload_operator_bcsr3(ddm_ctx *, int64_t nb_own, int64_t nb_ghost, const int64_t *rowptr, const int32_t *col, const double *val)
{
  rec1("op_kind", 3);
  rec1("op_n_own", nb_own);
  rec1("op_n_ghost", nb_ghost);
  rec("op_rowptr", rowptr, (std::size_t)nb_own + 1);
  rec("op_col", col, (std::size_t)rowptr[nb_own]);
  rec("op_val", val, (std::size_t)rowptr[nb_own] * 9);
  return 0;
}
int ddm_upload_decomposition(ddm_ctx *, int32_t nsub, const int64_t *sub_ptr, const int32_t *sub_dofs, const double *pou)
{
  rec1("nsub", nsub);
  rec("sub_ptr", sub_ptr, (std::size_t)nsub + 1);
  rec("sub_dofs", sub_dofs, (std::size_t)sub_ptr[nsub]);
  rec("pou", pou, (std::size_t)sub_ptr[nsub]);
  return 0;
}
static int factor_common(const char *kind, int32_t sub, int32_t n, int32_t nsuper, const int32_t *super, const int64_t *pi, const int64_t *px,
                         const int32_t *s, const double *x, const double *xu, const int32_t *perm)
{
  const std::string f = "F" + std::to_string(sub) + "_";
  rec1(f + "lu", xu != nullptr);
  rec1(f + "n", n);
  rec(f + "super", super, (std::size_t)nsuper + 1);
  rec(f + "pi", pi, (std::size_t)nsuper + 1);
  rec(f + "px", px, (std::size_t)nsuper + 1);
  rec(f + "s", s, (std::size_t)pi[nsuper]);
  rec(f + "x", x, (std::size_t)px[nsuper]);
  if (xu) rec(f + "xu", xu, (std::size_t)px[nsuper]);
  rec(f + "perm", perm, (std::size_t)n);
  (void)kind;
  return 0;
}
int ddm_upload_factor(ddm_ctx *, int32_t sub, int32_t n, int32_t nsuper, const int32_t *super, const int64_t *pi, const int64_t *px,
                      const int32_t *s, const double *x, const int32_t *perm)
{
  return factor_common("chol", sub, n, nsuper, super, pi, px, s, x, nullptr, perm);
}
int ddm_upload_factor_lu(ddm_ctx *, int32_t sub, int32_t n, int32_t nsuper, const int32_t *super, const int64_t *pi, const int64_t *px,
                         const int32_t *s, const double *xl, const double *xu, const int32_t *perm)
{
  return factor_common("lu", sub, n, nsuper, super, pi, px, s, xl, xu, perm);
}
int ddm_upload_coarse(ddm_ctx *, const int32_t *k, const double *Z, int32_t N0, int32_t k_offset, const double *A0)
{
  rec1("N0", N0);
  rec1("k_offset", k_offset);
  rec("A0", A0, (std::size_t)N0 * N0);
  (void)k;
  (void)Z;
  return 0;
}
int ddm_finalize(ddm_ctx *)
{
  rec1("finalized", 1);
  return 0;
}
}

int main(int argc, char **argv)
{
  if (argc < 3) {
    std::fprintf(stderr, "usage: pack_upload_check <in.ddmpack> <out.dump>\n");
    return 2;
  }
  try {
    Dune::ddm_b200::DdmPack pack(argv[1]);
    auto ctx = Dune::ddm_b200::Context::fromPack(pack, 0, true);
    if (ctx->size() != pack.scalar("n")) return 3;
  } catch (Dune::Exception &e) {
    std::fprintf(stderr, "Dune::Exception: %s\n", e.what());
    return 1;
  } catch (std::exception &e) {  // the pack reader refuses damaged or foreign files with std::runtime_error
    std::fprintf(stderr, "error: %s\n", e.what());
    return 4;
  }
  // dump: [u64 name length][name][u64 byte count][bytes] ...
  std::FILE *fh = std::fopen(argv[2], "wb");
  for (const auto &r : g_rec) {
    const std::uint64_t nl = r.name.size(), nb = r.bytes.size();
    std::fwrite(&nl, 8, 1, fh);
    std::fwrite(r.name.data(), 1, nl, fh);
    std::fwrite(&nb, 8, 1, fh);
    if (nb) std::fwrite(r.bytes.data(), 1, nb, fh);
  }
  std::fclose(fh);
  return 0;
}
