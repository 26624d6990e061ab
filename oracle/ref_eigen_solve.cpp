// Cross-check helper (TEST INFRASTRUCTURE): solves the damped reduced pose system exactly the way
// the reference does in SparseBlock::solve (/root/reference/src/droid_kernels.cu:1192-1213):
// Eigen::SparseMatrix<double>, diag += ep + lm*diag, Eigen::SimplicialLLT, zeros on failure.
// Built against the reference's vendored Eigen (thirdparty/eigen) into oracle/_ref/.
#include <Eigen/Sparse>
#include <Eigen/SparseCholesky>
#include <vector>

extern "C" __attribute__((visibility("default")))
int ref_eigen_llt_solve(const double* A, const double* b, int n, float lm, float ep, double* x) {
  typedef Eigen::Triplet<double> T;
  std::vector<T> trip;
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++)
      if (A[size_t(i) * n + j] != 0.0) trip.push_back(T(i, j, A[size_t(i) * n + j]));
  Eigen::SparseMatrix<double> L(n, n);
  L.setFromTriplets(trip.begin(), trip.end());
  L.diagonal().array() += ep + lm * L.diagonal().array();
  Eigen::SimplicialLLT<Eigen::SparseMatrix<double>> solver;
  solver.compute(L);
  Eigen::VectorXd bv = Eigen::Map<const Eigen::VectorXd>(b, n);
  if (solver.info() == Eigen::Success) {
    Eigen::VectorXd xv = solver.solve(bv);
    for (int i = 0; i < n; i++) x[i] = xv[i];
    return 0;
  }
  for (int i = 0; i < n; i++) x[i] = 0.0;
  return 1;
}
