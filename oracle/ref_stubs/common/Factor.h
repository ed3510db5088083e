// stand-in for the generated message header: fields of src/common/msg/Factor.msg (int8 id_1, id_2; delta)
#pragma once
#include <cstdint>
#include <common/Pose2DWithCovariance.h>
namespace common {
struct Factor {
  int8_t id_1 = 0, id_2 = 0;
  Pose2DWithCovariance delta;
};
}  // namespace common
