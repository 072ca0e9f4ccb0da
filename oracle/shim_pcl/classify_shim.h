// oracle/shim_pcl/classify_shim.h -- TEST INFRASTRUCTURE ONLY.
// The few declarations CCICP::classifyPoints (ccicp2d/src/icpTools.cpp:36-103) needs, so that the function body can be
// compiled UNMODIFIED from the reference file where it lies (oracle/Makefile extracts exactly those lines at build time;
// nothing of the reference is checked in).  Stand-ins for: pcl::PointCloud (pcl/point_cloud.h: `points`, clear(),
// push_back(), Ptr), PointXYZGD (ground_segmentation/include/ground_segmentation/PointcloudXYZGD.h:8-14), the three
// constants of ccicp2d/include/ccicp2d/icpTools.h:24-26 and its `using namespace std` (:28).
#pragma once
#include <cmath>
#include <cstdint>
#include <memory>
#include <vector>

#define NUMBINSGA 1200
#define RESOLUTION 0.5
#define GRD_ADJ_THRESH 2

using namespace std;

struct PointXYZGD {
  float x, y, z, pad_;        // PCL_ADD_POINT4D
  uint16_t ground_adj;
  uint16_t drivable;
};

namespace pcl {
template <class T>
struct PointCloud {
  typedef std::shared_ptr<PointCloud<T> > Ptr;
  std::vector<T> points;
  void clear() { points.clear(); }
  void push_back(const T& p) { points.push_back(p); }
};
}  // namespace pcl

class CCICP {
 public:
  void classifyPoints(pcl::PointCloud<PointXYZGD>::Ptr outcloud);
};
