#include "FrameOps.h"

#include <stdexcept>
#include <string>

#include "../include/orb_b200.h"

namespace ORB_SLAM2 {

static orb_camera make_camera(float fx, float fy, float cx, float cy, const std::vector<float>& d) {
  orb_camera c;
  c.fx = fx; c.fy = fy; c.cx = cx; c.cy = cy;
  for (int i = 0; i < 8; ++i) c.dist[i] = i < (int)d.size() ? d[i] : 0.f;
  c.n_dist = (int)d.size();
  return c;
}

void UndistortKeyPoints(const std::vector<cv::KeyPoint>& mvKeys, float fx, float fy, float cx, float cy, const std::vector<float>& distCoef,
                        std::vector<cv::KeyPoint>& mvKeysUn) {
  static_assert(sizeof(cv::KeyPoint) == sizeof(orb_keypoint), "cv::KeyPoint must be the 28-byte POD the C ABI reads");
  mvKeysUn.resize(mvKeys.size());
  if (mvKeys.empty()) return;
  const orb_camera c = make_camera(fx, fy, cx, cy, distCoef);
  const int32_t n = (int32_t)mvKeys.size();
  if (orb_undistort_keypoints(ORB_MEM_HOST, &c, reinterpret_cast<const orb_keypoint*>(mvKeys.data()), &n, n, 1,
                              reinterpret_cast<orb_keypoint*>(mvKeysUn.data()), nullptr) < 0)
    throw std::runtime_error(std::string("orb_undistort_keypoints: ") + orb_last_error());
}

void ComputeImageBounds(int cols, int rows, float fx, float fy, float cx, float cy, const std::vector<float>& distCoef, float& mnMinX,
                        float& mnMaxX, float& mnMinY, float& mnMaxY) {
  const orb_camera c = make_camera(fx, fy, cx, cy, distCoef);
  float b[4];
  if (orb_image_bounds(&c, cols, rows, b) < 0) throw std::runtime_error(std::string("orb_image_bounds: ") + orb_last_error());
  mnMinX = b[0]; mnMinY = b[1]; mnMaxX = b[2]; mnMaxY = b[3];
}

}  // namespace ORB_SLAM2
