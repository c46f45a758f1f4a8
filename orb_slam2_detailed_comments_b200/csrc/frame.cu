// Frame::UndistortKeyPoints and Frame::ComputeImageBounds (src/Frame.cc:770-806, 832-868): cv::undistortPoints(mat, mat, mK, mDistCoef,
// cv::Mat(), mK) on the keypoint coordinates. OpenCV >= 4 runs a fixed 5 iterations of the inverse radial-tangential model in double
// precision (default TermCriteria(MAX_ITER, 5, 0.01)) and re-projects with P = mK; restated here operation by operation (no FMA:
// the library is compiled with --fmad=false), verified bit-exact against cv2 4.13 (tests/golden/undistort_*.npz).
#include "orb_common.cuh"

struct OrbCamera {   // == orb_camera
  float fx, fy, cx, cy;
  float dist[8];     // k1 k2 p1 p2 [k3 [k4 k5 k6]]
  int n_dist;
};

__device__ __forceinline__ void undistort_point(const OrbCamera& c, float px, float py, float& ox, float& oy) {
  const double fx = c.fx, fy = c.fy, cx = c.cx, cy = c.cy;
  double k[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) k[i] = i < c.n_dist ? (double)c.dist[i] : 0.0;
  const double ifx = 1. / fx, ify = 1. / fy;
  const double u = px, v = py;
  double x = (u - cx) * ifx, y = (v - cy) * ify;
  const double x0 = x, y0 = y;
  for (int j = 0; j < 5; ++j) {
    const double r2 = x * x + y * y;
    const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
    if (icdist < 0) {   // OpenCV's guard against a non-invertible model: fall back to the normalised input
      x = (u - cx) * ifx;
      y = (v - cy) * ify;
      break;
    }
    const double dx = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x);
    const double dy = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y;
    x = (x0 - dx) * icdist;
    y = (y0 - dy) * icdist;
  }
  // P = mK, R = I: xx = fx*x + 0*y + cx, ww = 1
  ox = (float)(fx * x + cx);
  oy = (float)(fy * y + cy);
}

__global__ void __launch_bounds__(256) orb_undistort_kernel(OrbCamera cam, const OrbKeypoint* __restrict__ in, const int* __restrict__ n_kps,
                                                            int cap, OrbKeypoint* __restrict__ out) {
  const int f = blockIdx.y;
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= min(n_kps[f], cap)) return;
  OrbKeypoint k = in[(size_t)f * cap + i];
  if (cam.dist[0] != 0.0f) undistort_point(cam, k.x, k.y, k.x, k.y);   // "mDistCoef.at<float>(0) == 0.0 -> mvKeysUn = mvKeys" (:772-776)
  out[(size_t)f * cap + i] = k;
}

void orb_launch_undistort(const void* cam, const OrbKeypoint* in, const int* n_kps, int cap, int frames, OrbKeypoint* out, cudaStream_t st) {
  orb_undistort_kernel<<<dim3((cap + 255) / 256, frames), 256, 0, st>>>(*static_cast<const OrbCamera*>(cam), in, n_kps, cap, out);
}
