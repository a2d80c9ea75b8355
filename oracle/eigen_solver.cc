/*
 * eigen_solver.cc -- ORACLE ONLY (test infrastructure).  Thin wrapper that hands the oracle's
 * triplets to the REFERENCE'S OWN vendored Eigen 3.3.4 (compiled from /root/reference/Eigen where
 * it lies; never copied into this repo).  Solver type and call sequence are fluid.cc:1352
 * (declaration), :540 (setFromTriplets), :1473 (cg.compute(A)), :1474 (p = cg.solve(b)).
 * Compiled WITHOUT -fopenmp: run.sh:3 passes none, so the reference's Eigen is single-threaded.
 */
#include <Eigen/Sparse>
#include <Eigen/IterativeLinearSolvers>
#include <vector>
#include <stdint.h>

typedef Eigen::Triplet<double> Triplet;                                             /* fluid.cc:21 */
typedef Eigen::ConjugateGradient<Eigen::SparseMatrix<double>, Eigen::Lower | Eigen::Upper,
                                 Eigen::IncompleteCholesky<double> > RefCG;         /* fluid.cc:1352 */

extern "C" {

void* oracle_eig_create(void) { return new RefCG(); }
void oracle_eig_destroy(void* h) { delete static_cast<RefCG*>(h); }

int oracle_eig_compute_solve(void* h, int n, int64_t nnz, const int32_t* rows, const int32_t* cols,
                             const double* vals, const double* b, double* x, int max_iter, int* iters,
                             double* err) {
  RefCG& cg = *static_cast<RefCG*>(h);
  std::vector<Triplet> tripletList;
  tripletList.reserve((size_t)nnz);
  for (int64_t t = 0; t < nnz; ++t) tripletList.push_back(Triplet(rows[t], cols[t], vals[t]));
  Eigen::SparseMatrix<double> A(n, n);                                              /* fluid.cc:1460 */
  A.setFromTriplets(tripletList.begin(), tripletList.end());                        /* fluid.cc:540 */
  Eigen::Map<const Eigen::VectorXd> bv(b, n);
  if (max_iter >= 0) cg.setMaxIterations(max_iter);   /* oracle option for capped timing runs only */
  cg.compute(A);                                                                    /* fluid.cc:1473 */
  Eigen::VectorXd p = cg.solve(bv);                                                 /* fluid.cc:1474 */
  for (int i = 0; i < n; ++i) x[i] = p[i];
  *iters = (int)cg.iterations();
  *err = cg.error();
  return (int)cg.info();
}

int oracle_eig_dense(int n, const double* A, const double* b, double* x) {
  std::vector<Triplet> t;
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j)
      if (A[i * n + j] != 0) t.push_back(Triplet(i, j, A[i * n + j]));
  Eigen::SparseMatrix<double> M(n, n);
  M.setFromTriplets(t.begin(), t.end());
  RefCG cg;
  cg.compute(M);
  Eigen::Map<const Eigen::VectorXd> bv(b, n);
  Eigen::VectorXd p = cg.solve(bv);
  for (int i = 0; i < n; ++i) x[i] = p[i];
  return (int)cg.iterations();
}
}
