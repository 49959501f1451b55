// ydorb.hpp -- header-only C++11 host side above the C ABI of libydorb.so (include/ydorb.h).
//
// The reference is compiled C++ on OpenCV; its OpenCV toolchain is absent from this image, so this shim mirrors the
// reference's class surface for the front-end hot path WITHOUT OpenCV types:
//   YDORBSLAM::OrbExtractor  (src/orbExtractor.hpp:31-73)   same ctor arguments, extractAndCompute, scale getters,
//                                                           pyramid access
//   YDORBSLAM::OrbMatcher    (src/orbMatcher.hpp:24-66)     ratio / orientation flags, thresholds, distance, the
//                                                           search functions on flattened inputs
//   YDORBSLAM::FrameFeatures (Frame hooks, src/frame.hpp:83,96,166)  assignKeyPointsToGrid, getKeyPointsInArea,
//                                                           computeStereoMatches, computeStereoFromRGBD, computeBoW
// cv::KeyPoint -> YDORBSLAM::KeyPoint (same 28-byte layout), cv::Mat image -> ImageView, descriptor cv::Mat ->
// std::vector<uint8_t> (N x 32, continuous).  With OpenCV present a maintainer swaps these three aliases back
// (INTEGRATION.md).  Error behaviour: the reference has no error codes of its own; OpenCV throws cv::Exception.  Here
// every failing C-ABI call throws YDORBSLAM::Error carrying ydorb_last_error().  There is no CPU fallback.
#ifndef YDORB_HPP
#define YDORB_HPP
#include <cstring>
#include <limits>
#include <stdexcept>
#include <string>
#include <vector>

#include "ydorb.h"

// The classes below carry the reference's names in the reference's namespace.  A translation unit that still sees the
// reference's own classes of the same names (tests, a staged migration) puts these elsewhere: -DYDORB_NAMESPACE=ydorb_abi
#ifndef YDORB_NAMESPACE
#define YDORB_NAMESPACE YDORBSLAM
#endif

namespace YDORB_NAMESPACE {

typedef YdKeyPoint KeyPoint;  // layout of cv::KeyPoint: pt.x, pt.y, size, angle, response, octave, class_id

struct ImageView {  // a CV_8UC1 cv::Mat header: no ownership
  const uint8_t* data;
  int width, height, stride;
  ImageView() : data(nullptr), width(0), height(0), stride(0) {}
  ImageView(const uint8_t* d, int w, int h, int s = 0) : data(d), width(w), height(h), stride(s ? s : w) {}
  bool empty() const { return !data || width <= 0 || height <= 0; }
};

class Error : public std::runtime_error {
 public:
  Error(int code_, const std::string& what_) : std::runtime_error(what_), code(code_) {}
  int code;
};

inline void check(int rc) {
  if (rc != YDORB_OK) throw Error(rc, std::string("libydorb: ") + ydorb_last_error());
}

// Page-locked host memory (ydorb_host_alloc).  Frames and outputs kept in it are moved by DMA without the driver's
// staging copy; a cv::Mat header can wrap data(): cv::Mat(rows, cols, CV_8UC1, buf.data(), step).
class PinnedBuffer {
 public:
  explicit PinnedBuffer(size_t bytes) : m_p(nullptr), m_n(bytes) { check(ydorb_host_alloc(bytes, &m_p)); }
  ~PinnedBuffer() { ydorb_host_free(m_p); }
  PinnedBuffer(const PinnedBuffer&) = delete;
  PinnedBuffer& operator=(const PinnedBuffer&) = delete;
  uint8_t* data() const { return static_cast<uint8_t*>(m_p); }
  size_t size() const { return m_n; }

 private:
  void* m_p;
  size_t m_n;
};

// ---------------------------------------------------------------------------------------------- OrbExtractor
class OrbExtractor {
 public:
  // src/orbExtractor.hpp:33-34.  maxWidth/maxHeight size the device buffers (the reference allocates per call).
  OrbExtractor(int _keyPointsNum, float _scaleFactor, int _levelsNum, int _initFastThd, int _minFastThd, int maxWidth = 1920,
               int maxHeight = 1080, int device = 0)
      : m_h(nullptr), m_levels(_levelsNum), m_scale(_scaleFactor) {
    YdExtractorParams p = {_keyPointsNum, _scaleFactor, _levelsNum, _initFastThd, _minFastThd,
                           1 /* the reference ctor initialises minFastThd with initFastThd, src/orbExtractor.cpp:318 */};
    check(ydorb_extractor_create(&p, maxWidth, maxHeight, 1, device, &m_h));
    std::vector<int32_t> q(m_levels);
    check(ydorb_extractor_quotas(m_h, q.data()));
    m_capacity = 0;
    for (int v : q) m_capacity += v;
    if (m_capacity < 1) m_capacity = 1;
  }
  ~OrbExtractor() { ydorb_extractor_destroy(m_h); }
  OrbExtractor(const OrbExtractor&) = delete;
  OrbExtractor& operator=(const OrbExtractor&) = delete;

  // src/orbExtractor.cpp:355-399: key points level-major, descriptors N x 32.  Empty image: silent return (:356-358).
  void extractAndCompute(const ImageView& _image, std::vector<KeyPoint>& _keyPoints, std::vector<uint8_t>& _descriptors) {
    _keyPoints.clear();
    _descriptors.clear();
    if (_image.empty()) return;
    _keyPoints.resize(m_capacity);
    _descriptors.resize((size_t)m_capacity * 32);
    int32_t n = 0;
    check(ydorb_extract(m_h, _image.data, _image.width, _image.height, _image.stride, _keyPoints.data(), _descriptors.data(),
                        m_capacity, &n));
    _keyPoints.resize(n);
    _descriptors.resize((size_t)n * 32);
  }

  int getLevelsNum() const { return m_levels; }          // :42
  float getScaleFactor() const { return m_scale; }       // :43
  std::vector<float> getScaleFactors() const { return table(0); }            // :44
  std::vector<float> getInvScaleFactors() const { return table(1); }         // :45
  std::vector<float> getScaleFactorSquares() const { return table(2); }      // :46
  std::vector<float> getInvScaleFactorSquares() const { return table(3); }   // :47

  // m_v_imagePyramid[level] (:49-50) of the last call, copied to the host; padded = with the 19 px border the
  // reference's ROI views sit in
  std::vector<uint8_t> getImagePyramidLevel(int level, int* width, int* height, bool padded = false) {
    const uint8_t* d;
    int w, h, st;
    check(ydorb_pyramid_level(m_h, 0, level, &d, &w, &h, &st));
    if (padded) { w += 38; h += 38; }
    std::vector<uint8_t> out((size_t)w * h);
    check(ydorb_pyramid_download(m_h, 0, level, padded ? 1 : 0, out.data(), w));
    if (width) *width = w;
    if (height) *height = h;
    return out;
  }

  YdExtractor* handle() const { return m_h; }
  int capacity() const { return m_capacity; }

 private:
  std::vector<float> table(int which) const {
    std::vector<float> t[4];
    for (int i = 0; i < 4; i++) t[i].resize(m_levels);
    check(ydorb_extractor_scale_tables(m_h, t[0].data(), t[1].data(), t[2].data(), t[3].data()));
    return t[which];
  }
  YdExtractor* m_h;
  int m_levels, m_capacity;
  float m_scale;
};

// ------------------------------------------------------------------------------------------------ OrbMatcher
// DBoW3::FeatureVector (std::map<NodeId, std::vector<unsigned>>) flattened to CSR, the form the BoW matchers take
struct FeatureVector {
  std::vector<int32_t> nodeIds, nodeStart, items;
  YdFeatureVector view() const {
    YdFeatureVector v = {nodeIds.data(), nodeStart.data(), items.data(), (int32_t)nodeIds.size()};
    return v;
  }
};

class FrameFeatures;

class OrbMatcher {
 public:
  static const int m_int_highThd = 100, m_int_lowThd = 50, m_int_histLen = 30;  // src/orbMatcher.cpp:7-9
  OrbMatcher(float _bestSecondRatio = 0.6f, bool _isToCheckOrientation = true, int device = 0)  // src/orbMatcher.hpp:26
      : m_h(nullptr), m_flt_bestSecondRatio(_bestSecondRatio), m_b_isToCheckOrientation(_isToCheckOrientation) {
    check(ydorb_matcher_create(device, &m_h));
  }
  ~OrbMatcher() { ydorb_matcher_destroy(m_h); }
  OrbMatcher(const OrbMatcher&) = delete;
  OrbMatcher& operator=(const OrbMatcher&) = delete;

  // computeDescriptorsDistance (src/orbMatcher.cpp:11-23) for n descriptor pairs at once
  std::vector<int32_t> computeDescriptorsDistance(const uint8_t* a, const uint8_t* b, int n) {
    std::vector<int32_t> d(n);
    check(ydorb_hamming_pairs(m_h, a, b, n, d.data()));
    return d;
  }
  int computeDescriptorsDistance(const uint8_t* a, const uint8_t* b) { return computeDescriptorsDistance(a, b, 1)[0]; }

  // the top-2 scan every brute-force matcher runs (src/orbMatcher.cpp:319-334) + the ratio rule of :336
  std::vector<YdTop2> top2(const uint8_t* queries, int nq, const uint8_t* db, int ndb, int idxBase = 0) {
    std::vector<YdTop2> out(nq);
    check(ydorb_hamming_top2(m_h, queries, nq, db, ndb, idxBase, out.data()));
    return out;
  }
  bool accept(const YdTop2& t, int thd = m_int_lowThd) const {
    return t.best_dist <= thd && (float)t.best_dist < m_flt_bestSecondRatio * (float)t.second_dist;
  }

  // searchByProjectionInFrameAndMapPoint (src/orbMatcher.cpp:24-64) on the flattened map points (the projections
  // Frame::isInCameraFrustum stored, see frustum()); matchOut[i] = map point index given to key point i or -1
  int searchByProjectionInFrameAndMapPoint(FrameFeatures& _frame, const std::vector<YdMapPointProj>& _mapPoints,
                                           const uint8_t* _mapPointDescriptors, const uint8_t* _mapPointObsPositive,
                                           const uint8_t* _keyPointOccupied, std::vector<int32_t>& matchOut, float _thd = 3.0f);

  // the window searches after M1 (src/orbMatcher.cpp:65-302, :566-807) in their flattened form, see ydorb.h
  int searchWindows(FrameFeatures& _frame, const std::vector<YdWindowQuery>& q, const uint8_t* qDescriptors, const YdWindowParams& prm,
                    const uint8_t* kpOccupied, std::vector<int32_t>* bestOf, std::vector<int32_t>* matchOut);

  // searchByBowInKeyFrameAndFrame (:303-379, mode 0) / searchByBowInTwoKeyFrames (:380-462, mode 1)
  int searchByBow(int mode, const std::vector<KeyPoint>& kps1, const uint8_t* desc1, const uint8_t* valid1, const FeatureVector& fv1,
                  const std::vector<KeyPoint>& kps2, const uint8_t* desc2, const uint8_t* valid2, const FeatureVector& fv2,
                  std::vector<int32_t>& matchOut) {
    const YdFeatureVector a = fv1.view(), b = fv2.view();
    matchOut.assign(mode == 0 ? kps2.size() : kps1.size(), -1);
    if (matchOut.empty()) return 0;
    int32_t n = 0;
    check(ydorb_search_by_bow(m_h, kps1.data(), desc1, valid1, (int)kps1.size(), &a, kps2.data(), desc2, valid2, (int)kps2.size(), &b,
                              mode, m_flt_bestSecondRatio, m_b_isToCheckOrientation ? 1 : 0, matchOut.data(), &n));
    return n;
  }

  // Frame::isInCameraFrustum + MapPoint::predictScaleLevel over all local map points (src/tracking.cpp:585-591)
  std::vector<YdMapPointProj> frustum(const YdCamera& cam, const std::vector<YdMapPoint3d>& mps, float viewCosLimit = 0.5f) {
    std::vector<YdMapPointProj> out(mps.size());
    check(ydorb_frustum(m_h, &cam, mps.data(), (int)mps.size(), viewCosLimit, out.data()));
    return out;
  }
  // MapPoint::computeDistinctiveDescriptors (src/mapPoint.cpp:169-218), batched: start = CSR offsets per map point
  std::vector<int32_t> computeDistinctiveDescriptors(const uint8_t* desc, const std::vector<int32_t>& start) {
    std::vector<int32_t> best(start.empty() ? 0 : start.size() - 1);
    check(ydorb_distinctive_descriptors(m_h, desc, start.data(), (int)best.size(), best.data()));
    return best;
  }

  YdMatcher* handle() const { return m_h; }
  float ratio() const { return m_flt_bestSecondRatio; }
  bool checkOrientation() const { return m_b_isToCheckOrientation; }

 private:
  YdMatcher* m_h;
  float m_flt_bestSecondRatio;
  bool m_b_isToCheckOrientation;
};

// --------------------------------------------------------------------------------------------- Frame hooks
// The feature-side state of a Frame (key points, descriptors, right x) on the device, with the three hooks the hot
// path owns: assignKeyPointsToGrid (src/frame.cpp:249-264, done at construction), getKeyPointsInArea (:337-361),
// computeStereoMatches (:362-472) / computeStereoFromRGBD (:214-223).
class FrameFeatures {
 public:
  FrameFeatures(OrbMatcher& m, const std::vector<KeyPoint>& kps, const std::vector<uint8_t>& desc, const std::vector<float>& rightX,
                float minX, float minY, float maxX, float maxY, const std::vector<float>& scaleFactors)
      : m_h(nullptr), m_n((int)kps.size()) {
    YdFrameGeom g;
    std::memset(&g, 0, sizeof g);
    g.min_x = minX; g.min_y = minY; g.max_x = maxX; g.max_y = maxY;
    g.n_levels = (int32_t)scaleFactors.size();
    for (size_t i = 0; i < scaleFactors.size() && i < YDORB_MAX_LEVELS; i++) g.scale_factors[i] = scaleFactors[i];
    check(ydorb_frame_create(m.handle(), &g, kps.data(), desc.data(), rightX.empty() ? nullptr : rightX.data(), m_n, &m_h));
  }
  ~FrameFeatures() { ydorb_frame_destroy(m_h); }
  FrameFeatures(const FrameFeatures&) = delete;
  FrameFeatures& operator=(const FrameFeatures&) = delete;

  std::vector<int> getKeyPointsInArea(const float& _xCord, const float& _yCord, const float& _radius, const int& _minLevel = -1,
                                      const int& _maxLevel = -1) const {
    std::vector<int> out(m_n > 0 ? m_n : 1);
    int32_t n = 0;
    check(ydorb_frame_keypoints_in_area(m_h, _xCord, _yCord, _radius, _minLevel, _maxLevel, out.data(), (int)out.size(), &n));
    out.resize(n < (int)out.size() ? n : (int)out.size());
    return out;
  }
  YdFrame* handle() const { return m_h; }
  int size() const { return m_n; }

 private:
  YdFrame* m_h;
  int m_n;
};

inline int OrbMatcher::searchByProjectionInFrameAndMapPoint(FrameFeatures& _frame, const std::vector<YdMapPointProj>& _mapPoints,
                                                            const uint8_t* _mapPointDescriptors, const uint8_t* _mapPointObsPositive,
                                                            const uint8_t* _keyPointOccupied, std::vector<int32_t>& matchOut,
                                                            float _thd) {
  matchOut.assign(_frame.size(), -1);
  int32_t n = 0;
  check(ydorb_search_by_projection(_frame.handle(), _mapPoints.data(), _mapPointDescriptors, _mapPointObsPositive,
                                   (int)_mapPoints.size(), _keyPointOccupied, _thd, m_flt_bestSecondRatio, matchOut.data(), &n));
  return n;
}

inline int OrbMatcher::searchWindows(FrameFeatures& _frame, const std::vector<YdWindowQuery>& q, const uint8_t* qDescriptors,
                                     const YdWindowParams& prm, const uint8_t* kpOccupied, std::vector<int32_t>* bestOf,
                                     std::vector<int32_t>* matchOut) {
  if (bestOf) bestOf->assign(q.size(), -1);
  if (matchOut) matchOut->assign(_frame.size(), -1);
  int32_t n = 0;
  check(ydorb_search_windows(_frame.handle(), q.data(), qDescriptors, (int)q.size(), &prm, kpOccupied,
                             bestOf ? bestOf->data() : nullptr, matchOut ? matchOut->data() : nullptr, &n));
  return n;
}

// EXTRA searchForInitialization: `north_star` names it, the reference has no such function (src/orbMatcher.hpp:24-66).
// ORB-SLAM2's published algorithm on this reference's Frame hooks (ydorb_match_windowed_top2, no reference parity claim):
// vbPrevMatched in/out as (x, y) pairs, vnMatches12[i1] = matched key point of _frame2 or -1; returns the match count.
// areaMode 0 = the reference's getKeyPointsInArea predicate, 1 = ORB-SLAM2's own.
inline int searchForInitialization(OrbMatcher& /*matcher of _frame2*/, const std::vector<KeyPoint>& kps1, const std::vector<uint8_t>& desc1,
                                   FrameFeatures& _frame2, std::vector<float>& vbPrevMatched, std::vector<int32_t>& vnMatches12,
                                   int windowSize = 10, float nnRatio = 0.9f, bool checkOrientation = true, int areaMode = 0) {
  vnMatches12.assign(kps1.size(), -1);
  if (kps1.empty()) return 0;
  vbPrevMatched.resize(2 * kps1.size());
  int32_t n = 0;
  check(ydorb_match_windowed_top2(_frame2.handle(), kps1.data(), desc1.data(), (int)kps1.size(), vbPrevMatched.data(), windowSize,
                                  nnRatio, OrbMatcher::m_int_lowThd, checkOrientation ? 1 : 0, areaMode, vnMatches12.data(), &n));
  return n;
}

// Frame::computeStereoMatches (src/frame.cpp:362-472) on the pyramids the two extractors hold from their last call.
// referenceQuirks = true reproduces the reference (leftIdx lag etc.).
inline void computeStereoMatches(OrbMatcher& m, const OrbExtractor& left, const OrbExtractor& right, const std::vector<KeyPoint>& kpsLeft,
                                 const std::vector<uint8_t>& descLeft, const std::vector<KeyPoint>& kpsRight,
                                 const std::vector<uint8_t>& descRight, float baseLineTimesFx, float baseLine,
                                 std::vector<float>& rightXcords, std::vector<float>& depth, bool referenceQuirks = true) {
  rightXcords.assign(kpsLeft.size(), -1.f);  // :363-364
  depth.assign(kpsLeft.size(), -1.f);
  if (kpsLeft.empty()) return;
  const float maxD = baseLine > 0.f ? baseLineTimesFx / baseLine : std::numeric_limits<float>::infinity();  // :382; +inf on the first frame
  check(ydorb_stereo_match(m.handle(), left.handle(), 0, right.handle(), 0, kpsLeft.data(), descLeft.data(), (int)kpsLeft.size(),
                           kpsRight.data(), descRight.data(), (int)kpsRight.size(), baseLineTimesFx, maxD, referenceQuirks ? 0 : 1,
                           rightXcords.data(), depth.data()));
}

// Frame::computeStereoFromRGBD (:214-223) reading the raw depth map (CV_16U or CV_32F) with the depth-map factor of
// Tracking::grabImageRGBD (src/tracking.cpp:109-111) applied on the fly
inline void computeStereoFromRGBD(OrbMatcher& m, const std::vector<KeyPoint>& kps, const void* depthMap, int depthType, int width,
                                  int height, int strideBytes, float depthMapFactor, float baseLineTimesFx,
                                  std::vector<float>& rightXcords, std::vector<float>& depth) {
  rightXcords.assign(kps.size(), -1.f);
  depth.assign(kps.size(), -1.f);
  if (kps.empty()) return;
  check(ydorb_stereo_from_rgbd(m.handle(), kps.data(), (int)kps.size(), depthMap, depthType, width, height, strideBytes, depthMapFactor,
                               baseLineTimesFx, rightXcords.data(), depth.data()));
}

// Frame::undistortKeyPoints (:193-213), in place, with the reference's dist[0] != 0 guard
inline void undistortKeyPoints(OrbMatcher& m, std::vector<KeyPoint>& kps, const float K[9], const std::vector<float>& dist, const float P[9]) {
  if (kps.empty() || dist.empty() || dist[0] == 0.0f) return;
  std::vector<float> xy(2 * kps.size());
  for (size_t i = 0; i < kps.size(); i++) { xy[2 * i] = kps[i].x; xy[2 * i + 1] = kps[i].y; }
  check(ydorb_undistort_points(m.handle(), xy.data(), (int)kps.size(), K, dist.data(), (int)dist.size(), P, xy.data()));
  for (size_t i = 0; i < kps.size(); i++) { kps[i].x = xy[2 * i]; kps[i].y = xy[2 * i + 1]; }
}

// ------------------------------------------------------------------------------------------------ Vocabulary
// DBoW3::Vocabulary for the one call Frame::computeBoW makes (src/frame.cpp:265-272)
struct BowVector {  // std::map<WordId, WordValue> in key order
  std::vector<int32_t> wordIds;
  std::vector<double> values;
};

class Vocabulary {
 public:
  Vocabulary(OrbMatcher& m, const std::string& path) : m_h(nullptr) { check(ydorb_vocabulary_load(m.handle(), path.c_str(), &m_h)); }
  Vocabulary(OrbMatcher& m, const YdVocabularyDesc& d) : m_h(nullptr) { check(ydorb_vocabulary_create(m.handle(), &d, &m_h)); }
  ~Vocabulary() { ydorb_vocabulary_destroy(m_h); }
  Vocabulary(const Vocabulary&) = delete;
  Vocabulary& operator=(const Vocabulary&) = delete;

  // Vocabulary::transform(features, v, fv, levelsup)  (thirdParty/DBow3/src/Vocabulary.cpp:752-816)
  void transform(const std::vector<uint8_t>& features, BowVector& v, FeatureVector& fv, int levelsup) const {
    const int n = (int)(features.size() / 32);
    v.wordIds.assign(n, 0);
    v.values.assign(n, 0.0);
    fv.nodeIds.assign(n, 0);
    fv.nodeStart.assign(n + 1, 0);
    fv.items.assign(n, 0);
    YdBowResult r;
    std::memset(&r, 0, sizeof r);
    r.bow_ids = v.wordIds.data();
    r.bow_values = v.values.data();
    r.fv_node_ids = fv.nodeIds.data();
    r.fv_node_start = fv.nodeStart.data();
    r.fv_items = fv.items.data();
    check(ydorb_vocabulary_transform(m_h, features.data(), n, levelsup, &r));
    v.wordIds.resize(r.n_bow);
    v.values.resize(r.n_bow);
    fv.nodeIds.resize(r.n_fv_nodes);
    fv.nodeStart.resize(r.n_fv_nodes + 1);
    fv.items.resize(fv.nodeStart[r.n_fv_nodes]);
  }
  YdVocabulary* handle() const { return m_h; }

 private:
  YdVocabulary* m_h;
};

}  // namespace YDORB_NAMESPACE
#endif  // YDORB_HPP
