// graph.cpp with ROS stripped: replays a synthetic registration stream through the node's verbs
// (prior_factor / new_factor / loop_factor / solve, reference src/graph/src/graph.cpp:27-132).
//   g++ -std=c++17 -Iinclude examples/graph_node_noros.cpp -Lgraphslam_b200/lib -lgraphslam_cuda
//       -Wl,-rpath,$PWD/graphslam_b200/lib -o graph_node_noros   (one command line), then ./graph_node_noros 2000
#include <cstdio>
#include <cstdlib>
#include <random>

#include "graphslam_cuda.hpp"

int main(int argc, char** argv) {
  const int n = argc > 1 ? std::atoi(argv[1]) : 1000;
  const int width = 50;
  std::mt19937_64 rng(7);
  std::normal_distribution<double> nxy(0.0, 0.05), nth(0.0, 0.02);
  try {
    gs::GraphNode node;
    auto gt = [&](int k, double p[3]) {  // lawn-mower ground truth
      const int r = k / width, q = k % width, c = (r % 2 == 0) ? q : width - 1 - q;
      p[0] = c; p[1] = r; p[2] = (q == width - 1) ? M_PI / 2 : ((r % 2 == 0) ? 0.0 : M_PI);
    };
    auto between = [&](const double a[3], const double b[3], gs::Pose2DWithCovariance& d) {
      const double c = std::cos(a[2]), s = std::sin(a[2]), dx = b[0] - a[0], dy = b[1] - a[1];
      d.x = c * dx + s * dy + nxy(rng);
      d.y = -s * dx + c * dy + nxy(rng);
      d.theta = std::atan2(std::sin(b[2] - a[2]), std::cos(b[2] - a[2])) + nth(rng);
      d.covariance = {0.0025, 0, 0, 0, 0.0025, 0, 0, 0, 0.0004};
    };
    for (int k = 0; k < n; ++k) {
      gs::Registration reg;
      if (k == 0) {
        reg.first_frame_flag = true;
      } else {
        reg.keyframe_flag = true;
        double a[3], b[3];
        gt(k - 1, a);
        gt(k, b);
        reg.factor_new.id_1 = k;  // ids start at 1 (graph.cpp:31,68)
        between(a, b, reg.factor_new.delta);
        const int r = k / width;
        if (r > 0 && k % 3 == 0) {  // loop closure to the cell below: id_1 = last keyframe, id_2 = older (scanner.cpp:151-152)
          const int c = (r % 2 == 0) ? k % width : width - 1 - k % width;
          const int j = (r - 1) * width + (((r - 1) % 2 == 0) ? c : width - 1 - c);
          double o[3];
          gt(j, o);
          reg.loop_closure_flag = true;
          reg.factor_loop.id_1 = k + 1;
          reg.factor_loop.id_2 = j + 1;
          between(b, o, reg.factor_loop.delta);
        }
      }
      node.registration_callback(reg, /*do_solve=*/(k % 250 == 249));
    }
    gs_stats st = node.solve();
    std::printf("keyframes %zu factors %zu | LM iterations %d, PCG iterations %lld, chi2 %.6g -> %.6g, %.2f ms\n",
                node.keyframes.size(), node.graph.num_factors(), st.iterations, (long long)st.pcg_iterations_total,
                st.chi2_initial, st.chi2_final, st.ms_total);
    uint64_t id;
    double dist;
    if (node.graph.closest(node.keyframes.back().id, node.keyframes_to_skip_in_loop_closing, &id, &dist))
      std::printf("closest keyframe to %llu: %llu at %.3f m\n", (unsigned long long)node.keyframes.back().id,
                  (unsigned long long)id, dist);
  } catch (const gs::Error& e) {
    std::fprintf(stderr, "graphslam_cuda error %d: %s\n", (int)e.status, e.what());
    return 1;
  }
  return 0;
}
