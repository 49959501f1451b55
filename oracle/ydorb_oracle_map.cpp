// CPU oracle, part 2: the SURVEY 8(f) "next" rows (map-point side of the ORB front-end).  TEST INFRASTRUCTURE ONLY:
// imported by tests/, __graft_entry__.smoke() and bench.py's cpu legs, never by the product path.
//
// Parity status: these functions restate reference control flow line by line; the OpenCV primitives they lean on
// (3x3 * 3x1 CV_32F gemm, cv::norm, Mat::dot) were pinned against cv2 4.13.0 (tests/test_oracle_map.py regenerates the
// check when cv2 is importable).  No built reference binary exists in this image => "parity unpinned" beyond that.
#include <algorithm>
#include <climits>
#include <cmath>
#include <vector>

#include "ydorb_oracle.h"

extern "C" {

// N3  src/mapPoint.cpp:169-218
void orc_distinctive_descriptors(const uint8_t* desc, const int32_t* start, int n_points, int32_t* best) {
  for (int p = 0; p < n_points; p++) {
    const int n = start[p + 1] - start[p];
    if (n <= 0) { best[p] = -1; continue; }  // :179-181, :188-190: nothing to choose from
    const uint8_t* d = desc + (size_t)start[p] * 32;
    std::vector<float> dist((size_t)n * n);  // float distances[n][n]  (:193)
    for (int i = 0; i < n; i++) {
      dist[(size_t)i * n + i] = 0;
      for (int j = i + 1; j < n; j++) {
        const int dij = orc_hamming(d + (size_t)i * 32, d + (size_t)j * 32);  // :197
        dist[(size_t)i * n + j] = (float)dij;
        dist[(size_t)j * n + i] = (float)dij;
      }
    }
    int best_median = INT_MAX, best_idx = 0;  // :203-204
    for (int i = 0; i < n; i++) {
      std::vector<int> v(dist.begin() + (size_t)i * n, dist.begin() + (size_t)(i + 1) * n);  // :206
      std::sort(v.begin(), v.end());
      const int median = v[(size_t)(0.5 * n)];  // :208
      if (median < best_median) { best_median = median; best_idx = i; }
    }
    best[p] = best_idx;
  }
}

// 3x3 * 3x1 + 3x1 as cv::gemm's small-matrix path computes it for CV_32F (matmul: `float t0 = a[0]*b[0] + a[1]*b[1] +
// a[2]*b[2]; d = (float)(t0*alpha + c*beta)` with double alpha = beta = 1): pinned against cv2.gemm.
static inline float gemm3(const float* a, const float* b, float c) {
  const float t0 = a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
  return (float)((double)t0 * 1.0 + (double)c * 1.0);
}

// N2  src/frame.cpp:295-326
void orc_frustum(const OrcCamera* cam, const OrcMapPoint3d* mps, int n, float view_cos_limit, OrcMapPointProj* out) {
  for (int i = 0; i < n; i++) {
    const OrcMapPoint3d& mp = mps[i];
    OrcMapPointProj o = {0.f, 0.f, 0.f, 0.f, 0, 0};  // m_b_isTrackInView = false (:296)
    float pc[3];
    for (int r = 0; r < 3; r++) pc[r] = gemm3(cam->R + 3 * r, mp.pos, cam->t[r]);  // :300
    const float x = cam->fx * pc[0] / pc[2] + cam->cx;                              // :302
    const float y = cam->fy * pc[1] / pc[2] + cam->cy;                              // :303
    const float max_d = 1.2f * mp.max_distance, min_d = 0.8f * mp.min_distance;     // src/mapPoint.hpp:67-74
    float v[3];
    for (int r = 0; r < 3; r++) v[r] = mp.pos[r] - cam->origin[r];                  // :307
    double s = 0;
    for (int r = 0; r < 3; r++) s += (double)v[r] * (double)v[r];
    const float distance = (float)std::sqrt(s);                                     // :308 cv::norm -> double -> float
    double dot = 0;
    for (int r = 0; r < 3; r++) dot += (double)v[r] * (double)mp.normal[r];         // :311 Mat::dot -> double
    const float view_cos = (float)(dot / (double)distance);
    const bool in_image = x >= cam->min_x && x < cam->max_x && y >= cam->min_y && y < cam->max_y;  // :292-294
    if (pc[2] >= 0.0 && in_image && distance >= min_d && distance <= max_d && view_cos >= view_cos_limit) {  // :312
      // predictScaleLevel (src/mapPoint.cpp:265-277): unqualified log() on a float inside namespace YDORBSLAM with
      // only <cmath> in sight resolves to ::log(double); the quotient is then double / float -> double.
      const float ratio = mp.max_distance / distance;
      int level = (int)std::ceil(std::log((double)ratio) / (double)cam->log_scale_factor);
      if (level < 0) level = 0;
      else if (level >= cam->n_levels) level = cam->n_levels - 1;
      o.active = 1;
      o.proj_x = x;
      o.proj_right_x = x - cam->bf / pc[2];  // :318
      o.proj_y = y;
      o.level = level;
      o.view_cos = view_cos;
    }
    out[i] = o;
  }
}

}  // extern "C"
