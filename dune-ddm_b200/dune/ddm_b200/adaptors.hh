// Dune-side drop-in classes over the C ABI (include/ddm_b200.h).
//
// Each class mirrors the reference class of the same name and sits behind the same dune-istl interface, so
// Dune::CGSolver / Dune::RestartedGMResSolver / Dune::BiCGSTABSolver drive the GPU path unchanged:
//   NonOverlappingOperator<X>       : Dune::LinearOperator<X,X>   (dune/ddm/nonoverlapping_operator.hh:12-50)
//   NonOverlappingScalarProduct<X>  : Dune::ScalarProduct<X>      (dune/ddm/nonoverlapping_operator.hh:64-90)
//   SchwarzPreconditioner<X>        : Dune::Preconditioner<X,X>   (dune/ddm/schwarz.hh:55-220)
//   GalerkinPreconditioner<X>       : Dune::Preconditioner<X,X>   (dune/ddm/galerkin_preconditioner.hh:48-363)
//   CombinedPreconditioner<X>       : Dune::Preconditioner<X,X>   (dune/ddm/combined_preconditioner.hh:41-180)
// X is either DeviceVector (data stays in HBM; only the Krylov scalars cross to the host) or a host
// Dune::BlockVector<Dune::FieldVector<double,1>> (every call stages through the context's vectors: the literal
// "unchanged driver" mode).  Errors become DUNE_THROW(Dune::Exception), the reference's convention
// (dune/ddm/schwarz.hh:83,91,190).
#ifndef DUNE_DDM_B200_ADAPTORS_HH
#define DUNE_DDM_B200_ADAPTORS_HH

#include <algorithm>
#include <memory>
#include <string>
#include <vector>

#include <dune/common/exceptions.hh>
#include <dune/common/fvector.hh>
#include <dune/istl/bvector.hh>
#include <dune/istl/operators.hh>
#include <dune/istl/preconditioner.hh>
#include <dune/istl/scalarproducts.hh>
#include <dune/istl/solvercategory.hh>

#include "ddm_b200.h"
#include "ddmpack.hh"

namespace Dune::ddm_b200 {

inline void check(int status)
{
  if (status != 0) DUNE_THROW(Dune::Exception, "ddm_b200: " << ddm_last_error());
}

enum class SchwarzType { Standard = DDM_SCHWARZ_STANDARD, Restricted = DDM_SCHWARZ_RESTRICTED }; // schwarz.hh:35
enum class ApplyMode { Additive = DDM_COMBINE_ADDITIVE, Multiplicative = DDM_COMBINE_MULTIPLICATIVE }; // combined_preconditioner.hh:21

/** One GPU's share of the solve (RAII over ddm_ctx).  Setup products are uploaded once. */
class Context {
public:
  explicit Context(int device = 0) { check(ddm_ctx_create(device, &c_)); }
  ~Context()
  {
    for (auto *v : scratch_) ddm_vec_destroy(v);
    ddm_ctx_destroy(c_);
  }
  Context(const Context &) = delete;
  Context &operator=(const Context &) = delete;

  /** Whole problem on one GPU from a .ddmpack file. */
  static std::shared_ptr<Context> fromPack(const DdmPack &p, int device = 0, bool with_coarse = true)
  {
    auto ctx = std::make_shared<Context>(device);
    ddm_ctx *c = ctx->c_;
    const std::int64_t n = p.scalar("n");
    const std::int32_t nsub = (std::int32_t)p.scalar("nsub");
    if (p.has("block") && p.scalar("block") == 3) {
      // node-major 3x3 blocks (BCRSMatrix<FieldMatrix<double,3,3>>): scalar CSR of the pack -> block CSR
      const std::int64_t *rp = p["A_rowptr"].as<std::int64_t>();
      const std::int32_t *ci = p["A_col"].as<std::int32_t>();
      const double *va = p["A_val"].as<double>();
      const std::int64_t nb = n / 3;
      std::vector<std::int64_t> brp(nb + 1, 0);
      std::vector<std::int32_t> bci;
      std::vector<double> bva;
      std::vector<std::int64_t> slot(nb, -1);
      for (std::int64_t I = 0; I < nb; ++I) {
        const std::size_t first = bci.size();
        for (int r = 0; r < 3; ++r)
          for (std::int64_t k = rp[3 * I + r]; k < rp[3 * I + r + 1]; ++k) {
            const std::int32_t J = ci[k] / 3;
            if (slot[J] < (std::int64_t)first) {
              slot[J] = (std::int64_t)bci.size();
              bci.push_back(J);
              bva.insert(bva.end(), 9, 0.0);
            }
            bva[9 * slot[J] + 3 * r + ci[k] % 3] = va[k];
          }
        // ascending block columns, as BCRSMatrix stores them
        std::vector<std::size_t> ord(bci.size() - first);
        for (std::size_t q = 0; q < ord.size(); ++q) ord[q] = first + q;
        std::sort(ord.begin(), ord.end(), [&](std::size_t a, std::size_t b) { return bci[a] < bci[b]; });
        std::vector<std::int32_t> cs;
        std::vector<double> vs;
        for (std::size_t q : ord) {
          cs.push_back(bci[q]);
          vs.insert(vs.end(), bva.begin() + 9 * q, bva.begin() + 9 * q + 9);
        }
        std::copy(cs.begin(), cs.end(), bci.begin() + first);
        std::copy(vs.begin(), vs.end(), bva.begin() + 9 * first);
        for (std::size_t q = first; q < bci.size(); ++q) slot[bci[q]] = -1;
        brp[I + 1] = (std::int64_t)bci.size();
      }
      check(ddm_upload_operator_bcsr3(c, nb, 0, brp.data(), bci.data(), bva.data()));
    } else
      check(ddm_upload_operator(c, n, 0, p["A_rowptr"].as<std::int64_t>(), p["A_col"].as<std::int32_t>(), p["A_val"].as<double>()));
    check(ddm_upload_decomposition(c, nsub, p["sub_ptr"].as<std::int64_t>(), p["sub_dofs"].as<std::int32_t>(),
                                   p.has("pou") ? p["pou"].as<double>() : nullptr));  // no partition of unity in the pack: schwarz.hh:140
    const std::int64_t *sp = p["sub_ptr"].as<std::int64_t>();
    for (std::int32_t i = 0; i < nsub; ++i) {
      const std::string f = "F" + std::to_string(i) + "_";
      const auto &su = p[f + "super"];
      if (p.has(f + "xu"))  // LU factor of a non-symmetric subdomain matrix (the reference's UMFPack branch)
        check(ddm_upload_factor_lu(c, i, (std::int32_t)(sp[i + 1] - sp[i]), (std::int32_t)su.count() - 1, su.as<std::int32_t>(),
                                   p[f + "pi"].as<std::int64_t>(), p[f + "px"].as<std::int64_t>(), p[f + "s"].as<std::int32_t>(),
                                   p[f + "x"].as<double>(), p[f + "xu"].as<double>(), p[f + "perm"].as<std::int32_t>()));
      else
        check(ddm_upload_factor(c, i, (std::int32_t)(sp[i + 1] - sp[i]), (std::int32_t)su.count() - 1, su.as<std::int32_t>(),
                                p[f + "pi"].as<std::int64_t>(), p[f + "px"].as<std::int64_t>(), p[f + "s"].as<std::int32_t>(),
                                p[f + "x"].as<double>(), p[f + "perm"].as<std::int32_t>()));
    }
    const std::int32_t N0 = (std::int32_t)p["A0"].dims.at(0);
    if (with_coarse && N0 > 0)
      check(ddm_upload_coarse(c, p["k"].as<std::int32_t>(), p["Z"].as<double>(), N0, 0, p["A0"].as<double>()));
    check(ddm_finalize(c));
    ctx->n_ = n;
    return ctx;
  }

  ddm_ctx *handle() const { return c_; }
  std::int64_t size() const { return n_; }
  void setSize(std::int64_t n) { n_ = n; }
  /** Context-owned device vectors used to stage host vectors. */
  ddm_vec *scratch(std::size_t slot)
  {
    while (scratch_.size() <= slot) {
      ddm_vec *v = nullptr;
      check(ddm_vec_create(c_, &v));
      scratch_.push_back(v);
    }
    return scratch_[slot];
  }

private:
  ddm_ctx *c_ = nullptr;
  std::int64_t n_ = 0;
  std::vector<ddm_vec *> scratch_;
};

/** The vector type X for device-resident Krylov loops.  Provides exactly what dune-istl's solvers use
 *  (extern/dune-istl/dune/istl/solvers.hh:292-335,933-1012). */
class DeviceVector {
public:
  using field_type = double;
  using size_type = std::size_t;

  explicit DeviceVector(std::shared_ptr<Context> ctx) : ctx_(std::move(ctx)) { check(ddm_vec_create(ctx_->handle(), &v_)); }
  DeviceVector(const DeviceVector &o) : ctx_(o.ctx_)
  {
    check(ddm_vec_create(ctx_->handle(), &v_));
    check(ddm_vec_copy(v_, o.v_));
  }
  DeviceVector(DeviceVector &&o) noexcept : ctx_(std::move(o.ctx_)), v_(o.v_) { o.v_ = nullptr; }
  ~DeviceVector() { ddm_vec_destroy(v_); }

  DeviceVector &operator=(const DeviceVector &o)
  {
    if (this != &o) check(ddm_vec_copy(v_, o.v_));
    return *this;
  }
  DeviceVector &operator=(field_type a)
  {
    check(ddm_vec_fill(v_, a));
    return *this;
  }
  DeviceVector &axpy(field_type a, const DeviceVector &x)
  {
    check(ddm_vec_axpy(v_, a, x.v_));
    return *this;
  }
  DeviceVector &operator*=(field_type a)
  {
    check(ddm_vec_scale(v_, a));
    return *this;
  }
  DeviceVector &operator+=(const DeviceVector &x)
  {
    check(ddm_vec_add(v_, x.v_));
    return *this;
  }
  DeviceVector &operator-=(const DeviceVector &x) { return axpy(-1.0, x); }

  // Dune::ScalarProduct<X>'s default implementation asks the vector itself (scalarproducts.hh:63-74)
  field_type dot(const DeviceVector &y) const
  {
    double r = 0;
    check(ddm_dot(ctx_->handle(), v_, y.v_, &r));
    return r;
  }
  field_type two_norm() const
  {
    double r = 0;
    check(ddm_norm(ctx_->handle(), v_, &r));
    return r;
  }

  size_type N() const { return (size_type)ctx_->size(); }
  size_type dim() const { return N(); }
  template <class HostVector>
  void upload(const HostVector &h)
  {
    check(ddm_vec_upload(v_, &h[0][0]));
  }
  template <class HostVector>
  void download(HostVector &h) const
  {
    check(ddm_vec_download(v_, &h[0][0]));
  }
  ddm_vec *handle() const { return v_; }
  const std::shared_ptr<Context> &context() const { return ctx_; }

private:
  std::shared_ptr<Context> ctx_;
  ddm_vec *v_ = nullptr;
};

// ---- how an X reaches the C ABI -------------------------------------------------------------------------
template <class X>
struct Bridge { // host vectors: stage through context-owned device vectors
  static ddm_vec *in(Context &c, const X &x, std::size_t slot)
  {
    if ((std::int64_t)x.N() != c.size()) DUNE_THROW(Dune::Exception, "vector has " << x.N() << " entries, the context owns " << c.size());
    ddm_vec *v = c.scratch(slot);
    check(ddm_vec_upload(v, &x[0][0]));
    return v;
  }
  static ddm_vec *out(Context &c, X &, std::size_t slot) { return c.scratch(slot); }
  static void commit(Context &, X &x, ddm_vec *v) { check(ddm_vec_download(v, &x[0][0])); }
};
template <>
struct Bridge<DeviceVector> {
  static ddm_vec *in(Context &, const DeviceVector &x, std::size_t) { return x.handle(); }
  static ddm_vec *out(Context &, DeviceVector &x, std::size_t) { return x.handle(); }
  static void commit(Context &, DeviceVector &, ddm_vec *) {}
};

template <class X>
class NonOverlappingOperator : public Dune::LinearOperator<X, X> {
public:
  using field_type = typename X::field_type;
  explicit NonOverlappingOperator(std::shared_ptr<Context> ctx) : ctx_(std::move(ctx)) {}
  void apply(const X &x, X &y) const override
  {
    ddm_vec *xi = Bridge<X>::in(*ctx_, x, 0), *yo = Bridge<X>::out(*ctx_, y, 1);
    check(ddm_op_apply(ctx_->handle(), xi, yo));
    Bridge<X>::commit(*ctx_, y, yo);
  }
  void applyscaleadd(field_type alpha, const X &x, X &y) const override
  {
    ddm_vec *xi = Bridge<X>::in(*ctx_, x, 0), *yo = Bridge<X>::in(*ctx_, y, 1);
    check(ddm_op_applyscaleadd(ctx_->handle(), alpha, xi, yo));
    Bridge<X>::commit(*ctx_, y, yo);
  }
  Dune::SolverCategory::Category category() const override { return Dune::SolverCategory::nonoverlapping; }

private:
  std::shared_ptr<Context> ctx_;
};

template <class X>
class NonOverlappingScalarProduct : public Dune::ScalarProduct<X> {
public:
  using field_type = typename X::field_type;
  using real_type = typename Dune::FieldTraits<field_type>::real_type;
  explicit NonOverlappingScalarProduct(std::shared_ptr<Context> ctx) : ctx_(std::move(ctx)) {}
  field_type dot(const X &x, const X &y) const override
  {
    double r = 0;
    ddm_vec *xi = Bridge<X>::in(*ctx_, x, 0);
    ddm_vec *yi = (&x == &y) ? xi : Bridge<X>::in(*ctx_, y, 1);
    check(ddm_dot(ctx_->handle(), xi, yi, &r));
    return r;
  }
  real_type norm(const X &x) const override
  {
    double r = 0;
    check(ddm_norm(ctx_->handle(), Bridge<X>::in(*ctx_, x, 0), &r));
    return r;
  }
  Dune::SolverCategory::Category category() const override { return Dune::SolverCategory::nonoverlapping; }

private:
  std::shared_ptr<Context> ctx_;
};

template <class X>
class SchwarzPreconditioner : public Dune::Preconditioner<X, X> {
public:
  SchwarzPreconditioner(std::shared_ptr<Context> ctx, const std::string &type) : ctx_(std::move(ctx))
  {
    if (type == "standard") type_ = SchwarzType::Standard;
    else if (type == "restricted") type_ = SchwarzType::Restricted;
    else DUNE_THROW(Dune::Exception, "Unknown Schwarz type '" << type << "'"); // schwarz.hh:83
  }
  void pre(X &, X &) override {}
  void post(X &) override {}
  void apply(X &x, const X &d) override
  {
    ddm_vec *di = Bridge<X>::in(*ctx_, d, 0), *xo = Bridge<X>::out(*ctx_, x, 1);
    check(ddm_schwarz_apply(ctx_->handle(), (int)type_, di, xo));
    Bridge<X>::commit(*ctx_, x, xo);
  }
  Dune::SolverCategory::Category category() const override { return Dune::SolverCategory::nonoverlapping; }
  SchwarzType type() const { return type_; }
  const std::shared_ptr<Context> &context() const { return ctx_; }

private:
  std::shared_ptr<Context> ctx_;
  SchwarzType type_ = SchwarzType::Restricted;
};

template <class X>
class GalerkinPreconditioner : public Dune::Preconditioner<X, X> {
public:
  explicit GalerkinPreconditioner(std::shared_ptr<Context> ctx) : ctx_(std::move(ctx)) {}
  void pre(X &, X &) override {}
  void post(X &) override {}
  void apply(X &x, const X &d) override
  {
    ddm_vec *di = Bridge<X>::in(*ctx_, d, 0), *xo = Bridge<X>::out(*ctx_, x, 1);
    check(ddm_coarse_apply(ctx_->handle(), di, xo));
    Bridge<X>::commit(*ctx_, x, xo);
  }
  Dune::SolverCategory::Category category() const override { return Dune::SolverCategory::nonoverlapping; }
  const std::shared_ptr<Context> &context() const { return ctx_; }

private:
  std::shared_ptr<Context> ctx_;
};

/** Any list of preconditioners, applied in the order they were added (combined_preconditioner.hh:90-109,127-163).
 *  The reference's own pair -- Schwarz level first, Galerkin level second (examples/poisson.cc:305-306) -- on one
 *  context runs as one fused pass on the GPU (ddm_twolevel_apply); every other list (a third level, the coarse level
 *  first, a foreign Dune::Preconditioner) is applied level by level exactly as the reference does. */
template <class X>
class CombinedPreconditioner : public Dune::Preconditioner<X, X> {
public:
  explicit CombinedPreconditioner(const std::string &mode = "additive")
  {
    if (mode == "additive") mode_ = ApplyMode::Additive;
    else if (mode == "multiplicative") mode_ = ApplyMode::Multiplicative;
    else DUNE_THROW(Dune::NotImplemented, "Unknown apply mode '" << mode << "'"); // combined_preconditioner.hh:66
  }
  void set_op(std::shared_ptr<Dune::LinearOperator<X, X>> op) { op_ = std::move(op); } // :111
  void add(std::shared_ptr<Dune::Preconditioner<X, X>> prec)
  {
    if (!precs_.empty() && prec->category() != precs_[0]->category())
      DUNE_THROW(Dune::Exception, "ERROR: Categories of the new preconditioner does not match"); // :92-96
    precs_.push_back(std::move(prec));
    fine_ = precs_.size() <= 2 ? std::dynamic_pointer_cast<SchwarzPreconditioner<X>>(precs_[0]) : nullptr;
    coarse_ = precs_.size() == 2 ? std::dynamic_pointer_cast<GalerkinPreconditioner<X>>(precs_[1]) : nullptr;
    fused_ = fine_ && (precs_.size() == 1 || (coarse_ && coarse_->context() == fine_->context()));
  }
  void pre(X &x, X &b) override
  {
    for (auto &p : precs_) p->pre(x, b);
  }
  void post(X &x) override
  {
    for (auto &p : precs_) p->post(x);
  }
  void apply(X &x, const X &d) override
  {
    if (precs_.empty()) DUNE_THROW(Dune::Exception, "No preconditioners added");
    if (mode_ == ApplyMode::Multiplicative && precs_.size() > 1 && !op_)
      DUNE_THROW(Dune::Exception, "ERROR: ApplyMode is multiplicative but operator A is not provided. Set with `set_op`"); // :147
    if (fused_) {
      Context &c = *fine_->context();
      ddm_vec *di = Bridge<X>::in(c, d, 0), *xo = Bridge<X>::out(c, x, 1);
      check(ddm_twolevel_apply(c.handle(), (int)fine_->type(), (int)mode_, coarse_ ? 1 : 0, di, xo));
      Bridge<X>::commit(c, x, xo);
      return;
    }
    // generic list: combined_preconditioner.hh:132-158, level by level
    x = 0;
    precs_[0]->apply(x, d);
    X xnext(x);
    if (mode_ == ApplyMode::Additive) {
      for (std::size_t i = 1; i < precs_.size(); ++i) {
        xnext = 0;
        precs_[i]->apply(xnext, d);
        x += xnext;
      }
    }
    else {
      X dnext(d);
      for (std::size_t i = 1; i < precs_.size(); ++i) {
        op_->applyscaleadd(-1.0, x, dnext);
        xnext = 0;
        precs_[i]->apply(xnext, dnext);
        x += xnext;
      }
    }
  }
  Dune::SolverCategory::Category category() const override
  {
    return precs_.empty() ? Dune::SolverCategory::nonoverlapping : precs_[0]->category();
  }
  bool fused() const { return fused_; }

private:
  ApplyMode mode_ = ApplyMode::Additive;
  std::shared_ptr<Dune::LinearOperator<X, X>> op_;
  std::vector<std::shared_ptr<Dune::Preconditioner<X, X>>> precs_;
  std::shared_ptr<SchwarzPreconditioner<X>> fine_;
  std::shared_ptr<GalerkinPreconditioner<X>> coarse_;
  bool fused_ = false;
};

} // namespace Dune::ddm_b200
#endif
