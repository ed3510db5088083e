// stand-in for the generated message header: fields of src/common/msg/Pose2DWithCovariance.msg
// (geometry_msgs/Pose2D pose; float64[9] covariance)
#pragma once
namespace geometry_msgs {
struct Pose2D {
  double x = 0.0, y = 0.0, theta = 0.0;
};
}  // namespace geometry_msgs
namespace common {
struct Pose2DWithCovariance {
  geometry_msgs::Pose2D pose;
  double covariance[9] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
};
}  // namespace common
