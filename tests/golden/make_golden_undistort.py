"""Golden vectors for Frame::UndistortKeyPoints / ComputeImageBounds: the real cv2.undistortPoints(pts, K, dist, None, K)
(OpenCV 4.13 in this container) on seeded points for the cameras of the reference's example settings (TUM1, EuRoC) and an
8-coefficient rational model. Run from the repo root: python tests/golden/make_golden_undistort.py"""
import os

import cv2
import numpy as np

CAMERAS = {
    "tum1": (640, 480, 517.306408, 516.469215, 318.643040, 255.313989, [0.262383, -0.953104, -0.005358, 0.002628, 1.163314]),   # Examples/Monocular/TUM1.yaml
    "euroc": (752, 480, 458.654, 457.296, 367.215, 248.375, [-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05]),            # Examples/Monocular/EuRoC.yaml
    "rational8": (640, 480, 535.4, 539.2, 320.1, 247.6, [0.1, -0.2, 0.001, -0.002, 0.05, 0.01, -0.02, 0.003]),
    "strong": (640, 480, 300.0, 300.0, 320.0, 240.0, [-0.9, 0.1, 0.01, 0.01, 0.0]),   # drives icdist < 0 at the image corners
}
here = os.path.dirname(os.path.abspath(__file__))
out = {}
for name, (w, h, fx, fy, cx, cy, k) in CAMERAS.items():
    rng = np.random.default_rng(len(name))
    pts = np.stack([rng.uniform(-20, w + 20, 6000), rng.uniform(-20, h + 20, 6000)], 1).astype(np.float32)
    pts[:4] = [[0, 0], [w, 0], [0, h], [w, h]]   # the four corners of ComputeImageBounds
    pts[4:2000] = np.round(pts[4:2000])          # integer pixel positions (level-0 keypoints)
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
    und = cv2.undistortPoints(pts.reshape(-1, 1, 2), K, np.array(k, np.float32), None, K).reshape(-1, 2)
    out[name + "_cam"] = np.array([w, h, fx, fy, cx, cy] + k, np.float64)
    out[name + "_pts"] = pts
    out[name + "_und"] = und
np.savez_compressed(os.path.join(here, "undistort_cv2.npz"), **out)
print("wrote", len(out) // 3, "cameras, opencv", cv2.__version__)
