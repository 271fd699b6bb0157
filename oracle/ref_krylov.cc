// TEST INFRASTRUCTURE (oracle/_ref): the reference's OWN Krylov layer -- Dune::CGSolver and
// Dune::RestartedGMResSolver, compiled unmodified from /root/reference/extern/dune-istl -- driving either
//   cpu       the CPU oracle (oracle.c) behind Dune's three interfaces, host BlockVector; this pins the oracle's
//             PCG loop against the real solvers.hh and is the CPU baseline's outer loop;
//   gpu       the product (libddm_b200.so) through dune/ddm_b200/adaptors.hh with X = DeviceVector;
//   gpu-host  the product with X = Dune::BlockVector<FieldVector<double,1>>: the unchanged-driver drop-in.
// Usage: dune_krylov <cpu|gpu|gpu-host> <problem.ddmpack> <cg|gmres|bicgstab> <reduction> <maxit> <restart>
//                    <standard|restricted> <additive|multiplicative> <with_coarse 0|1> <nthreads> [solution.bin]
// Prints one JSON line.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <memory>
#include <string>
#include <vector>

#include <dune/common/fvector.hh>
#include <dune/istl/bvector.hh>
#include <dune/istl/operators.hh>
#include <dune/istl/preconditioner.hh>
#include <dune/istl/scalarproducts.hh>
#include <dune/istl/solvers.hh>

#include <dune/ddm_b200/adaptors.hh>

#include "oracle.h"

using HostVector = Dune::BlockVector<Dune::FieldVector<double, 1>>;
using namespace Dune::ddm_b200;

// ---- oracle behind Dune's interfaces ---------------------------------------------------------------------
struct OracleProblem {
  const DdmPack &p;
  oracle_setup S{};
  std::vector<std::int32_t> nsuper, k_off;
  std::vector<const std::int32_t *> super, s, perm;
  std::vector<const std::int64_t *> pi, px;
  std::vector<const double *> x;
  std::vector<std::int64_t> z_ptr;
  std::vector<double> chol, a0lu;
  std::vector<const double *> xu;  // LU packs: panels of U^T per subdomain
  std::vector<std::int32_t> a0piv;
  int nthreads;

  OracleProblem(const DdmPack &pack, int nt) : p(pack), nthreads(nt)
  {
    const int ns = (int)p.scalar("nsub");
    S.nsub = ns;
    S.n = p.scalar("n");
    S.sub_ptr = p["sub_ptr"].as<std::int64_t>();
    S.sub_dofs = p["sub_dofs"].as<std::int32_t>();
    S.pou = p["pou"].as<double>();
    const std::int32_t *k = p["k"].as<std::int32_t>();
    k_off.assign(ns + 1, 0);
    z_ptr.assign(ns + 1, 0);
    for (int i = 0; i < ns; ++i) {
      const std::string f = "F" + std::to_string(i) + "_";
      nsuper.push_back((std::int32_t)p[f + "super"].count() - 1);
      super.push_back(p[f + "super"].as<std::int32_t>());
      pi.push_back(p[f + "pi"].as<std::int64_t>());
      px.push_back(p[f + "px"].as<std::int64_t>());
      s.push_back(p[f + "s"].as<std::int32_t>());
      x.push_back(p[f + "x"].as<double>());
      perm.push_back(p[f + "perm"].as<std::int32_t>());
      if (p.has(f + "xu")) xu.push_back(p[f + "xu"].as<double>());
      k_off[i + 1] = k_off[i] + k[i];
      z_ptr[i + 1] = z_ptr[i] + (std::int64_t)k[i] * (S.sub_ptr[i + 1] - S.sub_ptr[i]);
    }
    S.nsuper = nsuper.data();
    S.super = super.data();
    S.pi = pi.data();
    S.px = px.data();
    S.s = s.data();
    S.x = x.data();
    S.perm = perm.data();
    S.z_ptr = z_ptr.data();
    S.k_off = k_off.data();
    S.Z = p["Z"].as<double>();
    const int N0 = k_off[ns];
    chol.assign((std::size_t)std::max(1, N0) * std::max(1, N0), 0.0);
    const double *A0 = p["A0"].as<double>();
    if ((int)xu.size() == ns && ns > 0) {
      // non-symmetric problem: LU subdomain factors and a dense LU with row swaps of A0 (LAPACK getrf layout, row-major)
      S.xu = xu.data();
      a0lu.assign(A0, A0 + (std::size_t)N0 * N0);
      a0piv.assign(std::max(1, N0), 0);
      for (int k = 0; k < N0; ++k) {
        int piv = k;
        for (int i = k + 1; i < N0; ++i)
          if (std::fabs(a0lu[(std::size_t)i * N0 + k]) > std::fabs(a0lu[(std::size_t)piv * N0 + k])) piv = i;
        a0piv[k] = piv;
        if (piv != k)
          for (int j = 0; j < N0; ++j) std::swap(a0lu[(std::size_t)k * N0 + j], a0lu[(std::size_t)piv * N0 + j]);
        for (int i = k + 1; i < N0; ++i) {
          const double l = a0lu[(std::size_t)i * N0 + k] / a0lu[(std::size_t)k * N0 + k];
          a0lu[(std::size_t)i * N0 + k] = l;
          for (int j = k + 1; j < N0; ++j) a0lu[(std::size_t)i * N0 + j] -= l * a0lu[(std::size_t)k * N0 + j];
        }
      }
      if (N0 > 0) {
        S.A0lu = a0lu.data();
        S.A0piv = a0piv.data();
      }
    } else
    for (int i = 0; i < N0; ++i) // dense Cholesky, row-major lower
      for (int j = 0; j <= i; ++j) {
        double acc = A0[(std::size_t)i * N0 + j];
        for (int l = 0; l < j; ++l) acc -= chol[(std::size_t)i * N0 + l] * chol[(std::size_t)j * N0 + l];
        chol[(std::size_t)i * N0 + j] = (i == j) ? std::sqrt(acc) : acc / chol[(std::size_t)j * N0 + j];
      }
    S.A0chol = chol.data();
    S.rowptr = p["A_rowptr"].as<std::int64_t>();
    S.col = p["A_col"].as<std::int32_t>();
    S.val = p["A_val"].as<double>();
  }
};

struct OracleOperator : Dune::LinearOperator<HostVector, HostVector> {
  const OracleProblem &P;
  explicit OracleOperator(const OracleProblem &p) : P(p) {}
  void apply(const HostVector &x, HostVector &y) const override
  {
    oracle_spmv(P.S.n, P.S.rowptr, P.S.col, P.S.val, &x[0][0], &y[0][0], P.nthreads);
  }
  void applyscaleadd(double a, const HostVector &x, HostVector &y) const override
  {
    HostVector t(x.N());
    apply(x, t);
    y.axpy(a, t);
  }
  Dune::SolverCategory::Category category() const override { return Dune::SolverCategory::nonoverlapping; }
};
struct OracleScalarProduct : Dune::ScalarProduct<HostVector> {
  double dot(const HostVector &x, const HostVector &y) const override { return oracle_dot((std::int64_t)x.N(), &x[0][0], &y[0][0]); }
  double norm(const HostVector &x) const override { return std::sqrt(dot(x, x)); }
  Dune::SolverCategory::Category category() const override { return Dune::SolverCategory::nonoverlapping; }
};
struct OraclePreconditioner : Dune::Preconditioner<HostVector, HostVector> {
  const OracleProblem &P;
  int restricted, mode, with_coarse;
  OraclePreconditioner(const OracleProblem &p, int r, int m, int w) : P(p), restricted(r), mode(m), with_coarse(w) {}
  void pre(HostVector &, HostVector &) override {}
  void post(HostVector &) override {}
  void apply(HostVector &x, const HostVector &d) override
  {
    oracle_twolevel_apply(&P.S, restricted, mode, with_coarse, &d[0][0], &x[0][0], P.nthreads);
  }
  Dune::SolverCategory::Category category() const override { return Dune::SolverCategory::nonoverlapping; }
};

// records every defect norm the solver asks for (CG: exactly one per iteration, solvers.hh:286,322)
template <class X>
struct RecordingScalarProduct : Dune::ScalarProduct<X> {
  std::shared_ptr<Dune::ScalarProduct<X>> sp;
  mutable std::vector<double> norms;
  explicit RecordingScalarProduct(std::shared_ptr<Dune::ScalarProduct<X>> s) : sp(std::move(s)) {}
  double dot(const X &x, const X &y) const override { return sp->dot(x, y); }
  double norm(const X &x) const override
  {
    const double r = sp->norm(x);
    norms.push_back(r);
    return r;
  }
  Dune::SolverCategory::Category category() const override { return sp->category(); }
};

template <class X>
static Dune::InverseOperatorResult run(const std::string &solver, std::shared_ptr<Dune::LinearOperator<X, X>> op,
                                       std::shared_ptr<RecordingScalarProduct<X>> sp, std::shared_ptr<Dune::Preconditioner<X, X>> prec,
                                       double reduction, int maxit, int restart, X &x, X &b)
{
  Dune::InverseOperatorResult res;
  if (solver == "cg") {
    Dune::CGSolver<X> cg(op, sp, prec, reduction, maxit, 0);
    cg.apply(x, b, res);
  } else if (solver == "gmres") {
    Dune::RestartedGMResSolver<X> gm(op, sp, prec, reduction, restart, maxit, 0);
    gm.apply(x, b, res);
  } else if (solver == "bicgstab") {
    Dune::BiCGSTABSolver<X> bi(op, sp, prec, reduction, maxit, 0);
    bi.apply(x, b, res);
  } else
    DUNE_THROW(Dune::Exception, "unknown solver " << solver);
  return res;
}

int main(int argc, char **argv)
{
  if (argc < 11) {
    std::fprintf(stderr, "usage: %s <cpu|gpu|gpu-host> pack <cg|gmres|bicgstab> reduction maxit restart <standard|restricted> "
                         "<additive|multiplicative> with_coarse nthreads [solution.bin]\n", argv[0]);
    return 2;
  }
  try {
    const std::string arm = argv[1], solver = argv[3], stype = argv[7], mode = argv[8];
    const double reduction = std::atof(argv[4]);
    const int maxit = std::atoi(argv[5]), restart = std::atoi(argv[6]), with_coarse = std::atoi(argv[9]), nthreads = std::atoi(argv[10]);
    DdmPack pack(argv[2]);
    const std::int64_t n = pack.scalar("n");
    HostVector xh(n), bh(n);
    xh = 0.0;
    const double *b = pack["b"].as<double>();
    for (std::int64_t i = 0; i < n; ++i) bh[i] = b[i];
    Dune::InverseOperatorResult res;
    std::vector<double> hist;
    const auto t0 = std::chrono::steady_clock::now();
    if (arm == "cpu") {
      OracleProblem P(pack, nthreads);
      auto op = std::make_shared<OracleOperator>(P);
      auto sp = std::make_shared<RecordingScalarProduct<HostVector>>(std::make_shared<OracleScalarProduct>());
      auto prec = std::make_shared<OraclePreconditioner>(P, stype == "restricted", mode == "multiplicative", with_coarse);
      res = run<HostVector>(solver, op, sp, prec, reduction, maxit, restart, xh, bh);
      hist = sp->norms;
    } else if (arm == "gpu-host") {
      auto ctx = Context::fromPack(pack, 0, with_coarse);
      auto op = std::make_shared<NonOverlappingOperator<HostVector>>(ctx);
      auto sp = std::make_shared<RecordingScalarProduct<HostVector>>(std::make_shared<NonOverlappingScalarProduct<HostVector>>(ctx));
      auto prec = std::make_shared<CombinedPreconditioner<HostVector>>(mode);
      prec->set_op(op);
      prec->add(std::make_shared<SchwarzPreconditioner<HostVector>>(ctx, stype));
      if (with_coarse) prec->add(std::make_shared<GalerkinPreconditioner<HostVector>>(ctx));
      res = run<HostVector>(solver, op, sp, prec, reduction, maxit, restart, xh, bh);
      hist = sp->norms;
    } else if (arm == "gpu") {
      auto ctx = Context::fromPack(pack, 0, with_coarse);
      DeviceVector x(ctx), bd(ctx);
      x.upload(xh);
      bd.upload(bh);
      auto op = std::make_shared<NonOverlappingOperator<DeviceVector>>(ctx);
      auto sp = std::make_shared<RecordingScalarProduct<DeviceVector>>(std::make_shared<NonOverlappingScalarProduct<DeviceVector>>(ctx));
      auto prec = std::make_shared<CombinedPreconditioner<DeviceVector>>(mode);
      prec->set_op(op);
      prec->add(std::make_shared<SchwarzPreconditioner<DeviceVector>>(ctx, stype));
      if (with_coarse) prec->add(std::make_shared<GalerkinPreconditioner<DeviceVector>>(ctx));
      res = run<DeviceVector>(solver, op, sp, prec, reduction, maxit, restart, x, bd);
      hist = sp->norms;
      x.download(xh);
    } else if (arm == "gpu-list") {
      // a list the fused pass does not cover -- the coarse level added first, then the Schwarz level: applied level by
      // level as combined_preconditioner.hh:132-158 does (the generic branch of the drop-in CombinedPreconditioner)
      auto ctx = Context::fromPack(pack, 0, with_coarse);
      DeviceVector x(ctx), bd(ctx);
      x.upload(xh);
      bd.upload(bh);
      auto op = std::make_shared<NonOverlappingOperator<DeviceVector>>(ctx);
      auto sp = std::make_shared<RecordingScalarProduct<DeviceVector>>(std::make_shared<NonOverlappingScalarProduct<DeviceVector>>(ctx));
      auto prec = std::make_shared<CombinedPreconditioner<DeviceVector>>(mode);
      prec->set_op(op);
      if (with_coarse) prec->add(std::make_shared<GalerkinPreconditioner<DeviceVector>>(ctx));
      prec->add(std::make_shared<SchwarzPreconditioner<DeviceVector>>(ctx, stype));
      if (with_coarse && prec->fused()) DUNE_THROW(Dune::Exception, "the reversed list must not take the fused pass");
      res = run<DeviceVector>(solver, op, sp, prec, reduction, maxit, restart, x, bd);
      hist = sp->norms;
      x.download(xh);
    } else
      DUNE_THROW(Dune::Exception, "unknown arm " << arm);
    const double wall = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (argc > 11) {
      std::FILE *fh = std::fopen(argv[11], "wb");
      std::fwrite(&xh[0][0], sizeof(double), (std::size_t)n, fh);
      std::fclose(fh);
    }
    std::printf("{\"arm\": \"%s\", \"solver\": \"%s\", \"iterations\": %d, \"converged\": %s, \"reduction\": %.17g, "
                "\"conv_rate\": %.17g, \"elapsed\": %.6f, \"wall\": %.6f, \"norms\": [",
                arm.c_str(), solver.c_str(), res.iterations, res.converged ? "true" : "false", res.reduction, res.conv_rate, res.elapsed, wall);
    for (std::size_t i = 0; i < hist.size(); ++i) std::printf("%s%.17g", i ? ", " : "", hist[i]);
    std::printf("]}\n");
    return 0;
  } catch (Dune::Exception &e) {
    std::fprintf(stderr, "Dune::Exception: %s\n", e.what());
    return 1;
  } catch (std::exception &e) {
    std::fprintf(stderr, "error: %s\n", e.what());
    return 1;
  }
}
