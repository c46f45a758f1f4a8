// Bodies of Frame::UndistortKeyPoints and Frame::ComputeImageBounds (src/Frame.cc:770-806, 832-868) over the C ABI
// (orb_undistort_keypoints / orb_image_bounds): call them from the reference's member functions with the same members.
#ifndef ORB_B200_FRAMEOPS_H
#define ORB_B200_FRAMEOPS_H

#include <vector>

#ifdef ORB_B200_USE_OPENCV
#include <opencv2/core/core.hpp>
#include <opencv2/features2d/features2d.hpp>
#else
#include "cv_shim.h"
#endif

namespace ORB_SLAM2 {

// fx, fy, cx, cy = mK.at<float>(0,0), (1,1), (0,2), (1,2); distCoef = mDistCoef's 4, 5 or 8 floats (k1 k2 p1 p2 [k3 [k4 k5 k6]]).
void UndistortKeyPoints(const std::vector<cv::KeyPoint>& mvKeys, float fx, float fy, float cx, float cy, const std::vector<float>& distCoef,
                        std::vector<cv::KeyPoint>& mvKeysUn);
void ComputeImageBounds(int cols, int rows, float fx, float fy, float cx, float cy, const std::vector<float>& distCoef, float& mnMinX,
                        float& mnMaxX, float& mnMinY, float& mnMaxY);

}  // namespace ORB_SLAM2
#endif
