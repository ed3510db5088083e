// graph.cpp with ROS stripped: replays a RECORDED registration stream (one common/Registration message per line, see
// graphslam_b200/regstream.py) through the node's verbs — registration_callback -> prior_factor / new_factor /
// loop_factor, solve (reference src/graph/src/graph.cpp:27-132,180-200) — and prints the Keyframes message the node
// would publish afterwards (graph.cpp:18-25), one keyframe per line at full precision.
//   g++ -std=c++17 -Iinclude examples/graph_node_noros.cpp -Lgraphslam_b200/lib -lgraphslam_cuda
//       -Wl,-rpath,$PWD/graphslam_b200/lib -o graph_node_noros   (one command line)
//   ./graph_node_noros tests/golden/registration_stream_manhattan400.reg [keyframes_out.txt]
#include <cstdio>
#include <cstdlib>

#include "graphslam_cuda.hpp"

int main(int argc, char** argv) {
  if (argc < 2) {
    std::fprintf(stderr, "usage: %s <registration stream file> [keyframes output file]\n", argv[0]);
    return 2;
  }
  try {
    const std::vector<gs::StreamEvent> events = gs::read_registration_stream(argv[1]);
    gs::GraphNode node;
    const gs_stats st = node.replay(events);
    const gs::Keyframes out = node.publish_keyframes();
    std::FILE* f = argc > 2 ? std::fopen(argv[2], "w") : stdout;
    if (!f) {
      std::fprintf(stderr, "cannot write %s\n", argv[2]);
      return 2;
    }
    for (const gs::Keyframe& kf : out.keyframes)
      std::fprintf(f, "KEYFRAME %llu %.17g %.17g %.17g\n", (unsigned long long)kf.id, kf.pose_opti.x, kf.pose_opti.y,
                   kf.pose_opti.theta);
    if (f != stdout) std::fclose(f);
    std::fprintf(stderr, "events %zu keyframes %zu factors %zu | last solve: LM iterations %d, PCG iterations %lld, chi2 %.17g -> %.17g\n",
                 events.size(), out.keyframes.size(), node.graph.num_factors(), st.iterations,
                 (long long)st.pcg_iterations_total, st.chi2_initial, st.chi2_final);
    uint64_t id;
    double dist;
    if (node.graph.closest(out.keyframes.back().id, node.keyframes_to_skip_in_loop_closing, &id, &dist))
      std::fprintf(stderr, "closest keyframe to %llu: %llu at %.17g m\n", (unsigned long long)out.keyframes.back().id,
                   (unsigned long long)id, dist);
  } catch (const gs::Error& e) {
    std::fprintf(stderr, "graphslam_cuda error %d: %s\n", (int)e.status, e.what());
    return 1;
  }
  return 0;
}
