/*
 * ref_extractor_driver.cpp -- extern "C" entry points around the REFERENCE's own OrbExtractor.  TEST INFRASTRUCTURE.
 *
 * This file is linked with /root/reference/src/orbExtractor.cpp compiled UNMODIFIED (oracle/refshim/Makefile) against
 * the OpenCV stand-in in oracle/refshim/opencv2/.  Nothing here restates the algorithm: every result below comes out of
 * YDORBSLAM::OrbExtractor.  A fresh extractor is constructed per image (the reference's pyramid goes stale
 * from the second call on; parity is defined on first-call semantics).
 */
#include <stdint.h>

#include <thread>
#include <vector>

#include "orbExtractor.hpp"

namespace {

struct RefKeyPoint {  // = cv::KeyPoint = OrcKeyPoint = YdKeyPoint (28 B)
  float x, y, size, angle, response;
  int32_t octave, class_id;
};
struct RefCandidate {  // = OrcCandidate
  int32_t x, y;
  float response;
};

/* protected members are reachable from a derived class; nothing is overridden */
struct Probe : YDORBSLAM::OrbExtractor {
  Probe(int n, float sf, int nl, int ini, int min) : YDORBSLAM::OrbExtractor(n, sf, nl, ini, min) {}
  const std::vector<int>& quotas() const { return m_v_keyPointsNumsPerLevel; }
  const std::vector<int>& umax() const { return m_v_maxXcords; }
  int iniTh() const { return m_int_initFastThd; }
  int minTh() const { return m_int_minFastThd; }
  void quadtree(const std::vector<cv::KeyPoint>& in, int minX, int maxX, int minY, int maxY, int want,
                std::vector<cv::KeyPoint>& out) {
    distributeQuadTree(in, minX, maxX, minY, maxY, want, out);
  }
};

}  // namespace

extern "C" {

/* E0: ctor tables as the reference's constructor computes them (src/orbExtractor.cpp:315-354).
 * umax gets m_v_maxXcords.size() entries (32 in the reference, F4); only [0,16) are ever read. */
int ref_ctor_tables(int n_features, float scale_factor, int n_levels, int ini_th, int min_th, float* sf, float* inv_sf,
                    float* sf2, float* inv_sf2, int32_t* quota, int32_t* umax, int umax_cap, int32_t* th_out) {
  Probe ex(n_features, scale_factor, n_levels, ini_th, min_th);
  std::vector<float> a = ex.getScaleFactors(), b = ex.getInvScaleFactors(), c = ex.getScaleFactorSquares(),
                     d = ex.getInvScaleFactorSquares();
  for (int i = 0; i < n_levels; i++) {
    sf[i] = a[i];
    inv_sf[i] = b[i];
    sf2[i] = c[i];
    inv_sf2[i] = d[i];
    quota[i] = ex.quotas()[i];
  }
  int n = (int)ex.umax().size();
  for (int i = 0; i < n && i < umax_cap; i++) umax[i] = ex.umax()[i];
  th_out[0] = ex.iniTh();
  th_out[1] = ex.minTh();
  return n;
}

/* E1: extractAndCompute on one image.  Optional outputs:
 *  per_level_n[L]                 keypoints per level
 *  pyr / pyr_off[L] / pyr_cap     the padded pyramid buffers (level l: (lw+38) x (lh+38), dense rows) -- E2
 *  cand / cand_level_start[L+1]   the raw FAST keypoints of every cell in call order, level coordinates relative to
 *                                 the 16 px border (what distributeQuadTree receives, :586-590) -- E3
 * Returns the keypoint count (outputs are truncated to cap). */
int ref_extract(int n_features, float scale_factor, int n_levels, int ini_th, int min_th, const uint8_t* img, int w, int h,
                int stride, RefKeyPoint* kps, uint8_t* desc, int cap, int32_t* per_level_n, uint8_t* pyr, int64_t* pyr_off,
                int64_t pyr_cap, RefCandidate* cand, int cand_cap, int32_t* cand_level_start) {
  Probe ex(n_features, scale_factor, n_levels, ini_th, min_th);
  cv::Mat image(h, w, CV_8UC1, const_cast<uint8_t*>(img), (size_t)stride);
  std::vector<cv::KeyPoint> keypoints;
  cv::Mat descriptors;
  cv::cvshim_trace trace;
  if (cand || cand_level_start) cv::cvshim_trace_ptr() = &trace;
  ex.extractAndCompute(image, keypoints, descriptors);
  cv::cvshim_trace_ptr() = nullptr;
  const int n = (int)keypoints.size();
  for (int i = 0; i < n && i < cap; i++) {
    const cv::KeyPoint& k = keypoints[i];
    RefKeyPoint r = {k.pt.x, k.pt.y, k.size, k.angle, k.response, k.octave, k.class_id};
    kps[i] = r;
    if (desc) memcpy(desc + (size_t)i * 32, descriptors.ptr(i), 32);
  }
  if (per_level_n) {
    for (int l = 0; l < n_levels; l++) per_level_n[l] = 0;
    for (int i = 0; i < n; i++) per_level_n[keypoints[i].octave]++;
  }
  const std::vector<cv::Mat>& P = ex.m_v_imagePyramid;
  if (pyr && pyr_off) {
    int64_t off = 0;
    for (int l = 0; l < n_levels && l < (int)P.size(); l++) {
      const cv::Mat& m = P[l];  // ROI view; the padded buffer starts 19 rows / 19 cols before it
      const int pw = m.cols + 38, ph = m.rows + 38;
      pyr_off[l] = off;
      if (off + (int64_t)pw * ph > pyr_cap) return -1;
      for (int y = 0; y < ph; y++) memcpy(pyr + off + (int64_t)y * pw, m.data + (ptrdiff_t)(y - 19) * (ptrdiff_t)m.step - 19, pw);
      off += (int64_t)pw * ph;
    }
  }
  if (cand_level_start) {
    /* map every traced cv::FAST call to its level through the data pointer, then undo the ROI origin */
    int out = 0, kp_i = 0, level = 0;
    for (int l = 0; l <= n_levels; l++) cand_level_start[l] = 0;
    for (size_t c = 0; c < trace.fast_calls.size(); c++) {
      const cv::cvshim_fast_call& fc = trace.fast_calls[c];
      int l = -1;
      for (int q = 0; q < (int)P.size() && q < n_levels; q++) {
        const uint8_t* lo = P[q].data;
        const uint8_t* hi = P[q].data + (size_t)P[q].step * (P[q].rows - 1) + P[q].cols;
        if (fc.data >= lo && fc.data < hi) l = q;
      }
      if (l < 0) return -2;
      while (level < l) cand_level_start[++level] = out;
      const ptrdiff_t d = fc.data - P[l].data;
      const int y0 = (int)(d / (ptrdiff_t)P[l].step), x0 = (int)(d % (ptrdiff_t)P[l].step);
      for (int i = 0; i < fc.count; i++, kp_i++) {
        if (cand && out < cand_cap) {
          const cv::KeyPoint& k = trace.fast_keypoints[kp_i];
          RefCandidate rc = {(int)k.pt.x + x0 - 16, (int)k.pt.y + y0 - 16, k.response};
          cand[out] = rc;
        }
        out++;
      }
    }
    while (level < n_levels) cand_level_start[++level] = out;
  }
  return n;
}

/* E4 alone: distributeQuadTree (src/orbExtractor.cpp:455-544) on a caller-supplied candidate list.
 * sel[i] = index into cand of the i-th output keypoint (recovered through class_id, which the function copies). */
int ref_quadtree(const RefCandidate* cand, int n, int rootw, int rooth, int want, int32_t* sel) {
  Probe ex(1000, 1.2f, 8, 20, 7);
  std::vector<cv::KeyPoint> in, out;
  in.reserve(n);
  for (int i = 0; i < n; i++) in.push_back(cv::KeyPoint((float)cand[i].x, (float)cand[i].y, 7.f, -1.f, cand[i].response, 0, i));
  ex.quadtree(in, 0, rootw, 0, rooth, want, out);
  for (size_t i = 0; i < out.size(); i++) sel[i] = out[i].class_id;
  return (int)out.size();
}

/* batch over host threads, one fresh extractor per frame -- the CPU baseline of bench.py (`kind: "reference"`) */
int ref_extract_batch(int n_features, float scale_factor, int n_levels, int ini_th, int min_th, const uint8_t* imgs, int w,
                      int h, int stride, int64_t frame_stride, int batch, RefKeyPoint* kps, uint8_t* desc,
                      int cap_per_frame, int32_t* n_out, int threads) {
  if (threads < 1) threads = 1;
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; t++) {
    pool.emplace_back([=]() {
      for (int b = t; b < batch; b += threads) {
        int n = ref_extract(n_features, scale_factor, n_levels, ini_th, min_th, imgs + (size_t)b * frame_stride, w, h,
                            stride, kps + (size_t)b * cap_per_frame, desc + (size_t)b * cap_per_frame * 32,
                            cap_per_frame, nullptr, nullptr, nullptr, 0, nullptr, 0, nullptr);
        n_out[b] = n < cap_per_frame ? n : cap_per_frame;
      }
    });
  }
  for (auto& th : pool) th.join();
  return 0;
}

}  // extern "C"
