// stand-in: the one Eigen type graph.hpp uses (it reaches Eigen through GTSAM's headers) — a dynamic matrix of
// doubles with (rows, cols) construction, element access, rows()/cols() and `m.row(i) << a, b, c;`
#pragma once
#include <cstddef>
#include <vector>
namespace Eigen {
class MatrixXd {
 public:
  MatrixXd() = default;
  MatrixXd(int r, int c) : r_(r), c_(c), v_((std::size_t)r * c, 0.0) {}
  int rows() const { return r_; }
  int cols() const { return c_; }
  double& operator()(int i, int j) { return v_[(std::size_t)i * c_ + j]; }
  double operator()(int i, int j) const { return v_[(std::size_t)i * c_ + j]; }
  class RowFill {
   public:
    RowFill(MatrixXd& m, int i) : m_(m), i_(i) {}
    RowFill& operator<<(double x) { return put(x); }
    RowFill& operator,(double x) { return put(x); }
   private:
    RowFill& put(double x) {
      m_(i_, k_++) = x;
      return *this;
    }
    MatrixXd& m_;
    int i_, k_ = 0;
  };
  RowFill row(int i) { return RowFill(*this, i); }
 private:
  int r_ = 0, c_ = 0;
  std::vector<double> v_;
};
}  // namespace Eigen
