// C entry points over the reference's own graph.hpp (compiled from /root/reference where it lies, through the
// stand-in headers in oracle/ref_stubs/): the only arithmetic of the path that is not inside GTSAM.
// TEST INFRASTRUCTURE ONLY — pins gs_compose_reference / oracle compose_reference and the covariance layout
// against the reference's compiled code (tests/test_oracle_kat.py, tests/golden/make_golden.py).
#include <graph.hpp>   // reference src/graph/include/graph.hpp: compose :30-43, covariance_to_eigen :45-58, eigen_to_covariance :60-68

extern "C" {

// compose(input_1, input_2) -> (x, y, theta)
void ref_compose(const double a[3], const double b[3], double out[3]) {
  common::Pose2DWithCovariance p1, p2;
  p1.pose.x = a[0]; p1.pose.y = a[1]; p1.pose.theta = a[2];
  p2.pose.x = b[0]; p2.pose.y = b[1]; p2.pose.theta = b[2];
  const common::Pose2DWithCovariance o = compose(p1, p2);
  out[0] = o.pose.x; out[1] = o.pose.y; out[2] = o.pose.theta;
}

// covariance_to_eigen(factor) -> Q(i, j) written row-major into q9; eigen_to_covariance(pose, Q) -> back9
void ref_covariance_roundtrip(const double cov9[9], double q9[9], double back9[9]) {
  common::Factor f;
  for (int k = 0; k < 9; ++k) f.delta.covariance[k] = cov9[k];
  Eigen::MatrixXd Q = covariance_to_eigen(f);
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) q9[3 * i + j] = Q(i, j);
  common::Pose2DWithCovariance p;
  p = eigen_to_covariance(p, Q);
  for (int k = 0; k < 9; ++k) back9[k] = p.covariance[k];
}

}  // extern "C"
