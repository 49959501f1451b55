// CPU oracle, part 4: SURVEY 8(f) row N4, the pre-/post-processing around extractAndCompute.  TEST INFRASTRUCTURE ONLY:
// imported by tests/ and bench tools' cpu legs, never by the product path.
//
// The reference calls OpenCV for all three steps (cv::cvtColor src/tracking.cpp:73-89,:97-107; cv::undistortPoints
// src/frame.cpp:204,:241; Mat::convertTo src/tracking.cpp:110).  OpenCV is un-vendored and unpinned in the reference, so
// these restate OpenCV 4.13.0's arithmetic; tests/test_oracle_preprocess.py pins them bit-exactly against cv2 4.13.0
// (cvtColor, undistortPoints) and tests/golden/preprocess.npz carries cv2-generated vectors.
#include <cmath>
#include <cstdint>
#include <cstring>

#include "ydorb_oracle.h"

extern "C" {

// cv::cvtColor 8-bit *2GRAY: RGB2Gray<uchar> with 15-bit coefficients (R 9798, G 19235, B 3735), round to nearest.
// code: 0 RGB2GRAY, 1 BGR2GRAY, 2 RGBA2GRAY, 3 BGRA2GRAY.
void orc_cvt_gray(const uint8_t* src, int width, int height, int src_stride, int code, uint8_t* dst, int dst_stride) {
  const int ch = code >= 2 ? 4 : 3;
  const bool swap = code == 1 || code == 3;
  for (int y = 0; y < height; y++) {
    const uint8_t* s = src + (size_t)y * src_stride;
    uint8_t* d = dst + (size_t)y * dst_stride;
    for (int x = 0; x < width; x++, s += ch) {
      const unsigned r = swap ? s[2] : s[0], g = s[1], b = swap ? s[0] : s[2];
      d[x] = (uint8_t)((r * 9798u + g * 19235u + b * 3735u + (1u << 14)) >> 15);
    }
  }
}

// cv::undistortPoints(src, dst, K, dist, noArray(), P) -> cvUndistortPointsInternal with TermCriteria(MAX_ITER, 5, 0.01)
// (no tilt model: invMatTilt = I).  K, P row-major 3x3 float; n_dist in {4, 5, 8, 12}.
void orc_undistort_points(const float* xy, int n, const float* K, const float* dist, int n_dist, const float* P, float* out) {
  double k[14] = {0};
  for (int i = 0; i < n_dist && i < 12; i++) k[i] = dist[i];
  const double fx = K[0], fy = K[4], cx = K[2], cy = K[5];
  const double ifx = 1. / fx, ify = 1. / fy;
  double RR[3][3];
  for (int i = 0; i < 9; i++) RR[i / 3][i % 3] = P[i];
  for (int i = 0; i < n; i++) {
    double x = xy[2 * i], y = xy[2 * i + 1];
    const double u = x, v = y;
    x = (x - cx) * ifx;
    y = (y - cy) * ify;
    const double x0 = x, y0 = y;
    for (int j = 0; j < 5; j++) {
      const double r2 = x * x + y * y;
      const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
      if (icdist < 0) {
        x = (u - cx) * ifx;
        y = (v - cy) * ify;
        break;
      }
      const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
      const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
      x = (x0 - deltaX) * icdist;
      y = (y0 - deltaY) * icdist;
    }
    const double xx = RR[0][0] * x + RR[0][1] * y + RR[0][2];
    const double yy = RR[1][0] * x + RR[1][1] * y + RR[1][2];
    const double ww = 1. / (RR[2][0] * x + RR[2][1] * y + RR[2][2]);
    out[2 * i] = (float)(xx * ww);
    out[2 * i + 1] = (float)(yy * ww);
  }
}

// Tracking::grabImageRGBD's convertTo (src/tracking.cpp:109-111) + Frame::computeStereoFromRGBD (src/frame.cpp:214-223)
// at the key points only.  depth_type 0 = CV_16U, 1 = CV_32F.  A key point outside the map (undefined behaviour in the
// reference) yields -1.
void orc_stereo_from_rgbd(const OrcKeyPoint* kps, int n, const void* depth_map, int depth_type, int width, int height,
                          int depth_stride, float factor, float bf, float* right_x, float* depth) {
  const bool convert = depth_type == 0 || std::fabs(factor - 1.0f) > 1e-5;  // :109
  for (int i = 0; i < n; i++) {
    right_x[i] = -1.f;  // :215-216
    depth[i] = -1.f;
    const float px = kps[i].x, py = kps[i].y;
    if (!(px > -1.f && py > -1.f)) continue;
    const int ix = (int)px, iy = (int)py;  // at<float>(float, float) truncates
    if (ix < 0 || iy < 0 || ix >= width || iy >= height) continue;
    const uint8_t* row = (const uint8_t*)depth_map + (size_t)iy * depth_stride;
    float d;
    if (depth_type == 0) {
      uint16_t raw;
      std::memcpy(&raw, row + 2 * (size_t)ix, 2);
      d = (float)raw * factor;  // cvt_32f<ushort, float>: float multiply (+ 0.f)
    } else {
      std::memcpy(&d, row + 4 * (size_t)ix, 4);
      if (convert) d = d * factor;
    }
    if (d > 0) {  // :219
      depth[i] = d;
      right_x[i] = px - bf / d;  // :221
    }
  }
}

}  // extern "C"
