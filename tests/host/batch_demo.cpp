// batch_demo.cpp -- drives include/ccicp2d/icpToolsBatch.h the way scan_registration / graph_slam style hosts would
// (ccicp2d/src/icpTools.cpp:248-279: Eigen::Vector2d vectors per class + a geometry_msgs::PoseStamped), with
// stand-in types of the same shape (no Eigen / ROS in this image).  Test scaffolding.
//
// usage: batch_demo <pairs.bin> [single]
// pairs.bin: int32 n_pairs; per pair: int32 nmGA,nmNGA,nsGA,nsNGA; double pose[7] (x,y,z,qx,qy,qz,qw); 4 xy arrays.
// prints one line per pair: x y z qx qy qz qw num_corr iters status
#include <cstdio>
#include <cstring>
#include <vector>

#include "ccicp2d/icpToolsBatch.h"

struct Vec2 {                       // shape of Eigen::Vector2d
  double v[2];
  double operator()(int i) const { return v[i]; }
};
struct Point { double x = 0, y = 0, z = 0; };
struct Quat { double x = 0, y = 0, z = 0, w = 0; };
struct Pose { Point position; Quat orientation; };
struct PoseStamped { Pose pose; };  // shape of geometry_msgs::PoseStamped

int main(int argc, char** argv) {
  if (argc < 2) return 2;
  FILE* f = std::fopen(argv[1], "rb");
  if (!f) return 3;
  const bool single = argc > 2 && !std::strcmp(argv[2], "single");
  int32_t n_pairs = 0;
  if (std::fread(&n_pairs, 4, 1, f) != 1) return 4;
  CCICPBatch batch(0);
  if (batch.status() != CCICP_OK) return 10;
  std::vector<PoseStamped> inits, results;
  for (int p = 0; p < n_pairs; p++) {
    int32_t n[4];
    double pose[7];
    if (std::fread(n, 4, 4, f) != 4 || std::fread(pose, 8, 7, f) != 7) return 5;
    std::vector<Vec2> a[4];
    for (int k = 0; k < 4; k++) {
      a[k].resize(n[k]);
      if (n[k] && std::fread(a[k].data(), 16, n[k], f) != (size_t)n[k]) return 6;
    }
    PoseStamped init;
    init.pose.position.x = pose[0]; init.pose.position.y = pose[1]; init.pose.position.z = pose[2];
    init.pose.orientation.x = pose[3]; init.pose.orientation.y = pose[4];
    init.pose.orientation.z = pose[5]; init.pose.orientation.w = pose[6];
    if (single) {
      results.push_back(batch.doICPMatch(a[0], a[1], a[2], a[3], init));
      const ccicp_result& r = batch.record(0);
      const Pose& o = results.back().pose;
      std::printf("%.17g %.17g %.17g %.17g %.17g %.17g %.17g %d %d %u\n", o.position.x, o.position.y, o.position.z,
                  o.orientation.x, o.orientation.y, o.orientation.z, o.orientation.w, r.num_corr, r.iters, r.status);
    } else {
      batch.add(a[0], a[1], a[2], a[3], init);
    }
  }
  std::fclose(f);
  if (!single) {
    if (batch.run() != CCICP_OK) return 11;
    for (int p = 0; p < n_pairs; p++) {
      PoseStamped out;
      batch.result(p, out);
      const ccicp_result& r = batch.record(p);
      const Pose& o = out.pose;
      std::printf("%.17g %.17g %.17g %.17g %.17g %.17g %.17g %d %d %u\n", o.position.x, o.position.y, o.position.z,
                  o.orientation.x, o.orientation.y, o.orientation.z, o.orientation.w, r.num_corr, r.iters, r.status);
    }
  }
  return 0;
}
