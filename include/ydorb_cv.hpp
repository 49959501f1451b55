// ydorb_cv.hpp -- the OpenCV-typed drop-in above the C ABI of libydorb.so: the reference's own signatures.
//
// Guarded: it needs <opencv2/core.hpp> (use ydorb.hpp where OpenCV is absent).  What a reference maintainer includes
// instead of calling the CPU bodies:
//   ydorb_cv::OrbExtractor            ctor and extractAndCompute exactly as src/orbExtractor.hpp:33-40, the getters of
//                                     :42-49 and m_v_imagePyramid (:50)
//   ydorb_cv::computeDescriptorsDistance(const cv::Mat&, const cv::Mat&)              src/orbMatcher.hpp:28
//   ydorb_cv::searchByProjectionInFrameAndMapPoint(matcher, Frame&, mapPoints, thd)   src/orbMatcher.hpp:32, on the
//                                     reference's own Frame / MapPoint objects (templates: this header does not include
//                                     the reference's headers, the caller's translation unit does)
//   ydorb_cv::getKeyPointsInArea(matcher, Frame&, x, y, r, minLevel, maxLevel)        src/frame.hpp:83
//   ydorb_cv::computeStereoMatches(matcher, Frame&, leftExtractor, rightExtractor)    src/frame.hpp:96
// Compiled and run against the reference's headers in this repo by oracle/refshim/cv_shim_check.cpp (with the OpenCV
// stand-in of oracle/refshim/) and tests/test_gpu_cv_shim.py; INTEGRATION.md shows where each call goes.
// Error behaviour: a failing C-ABI call throws YDORBSLAM::Error (ydorb.hpp); there is no CPU fallback.
#ifndef YDORB_CV_HPP
#define YDORB_CV_HPP

#if !defined(__has_include)
#error "ydorb_cv.hpp needs __has_include to find OpenCV; use ydorb.hpp"
#elif !__has_include(<opencv2/core.hpp>)
#error "ydorb_cv.hpp needs OpenCV's core headers; without OpenCV use ydorb.hpp"
#endif
#include <opencv2/core.hpp>

#include <memory>
#include <vector>

#include "ydorb.hpp"

namespace ydorb_cv {

static_assert(sizeof(cv::KeyPoint) == sizeof(YdKeyPoint), "cv::KeyPoint and YdKeyPoint share one 28-byte layout");

inline std::vector<uint8_t> descriptor_rows(const cv::Mat& d) {  // N x 32 CV_8U, any row step -> continuous bytes
  std::vector<uint8_t> out((size_t)d.rows * 32);
  for (int i = 0; i < d.rows; i++) std::memcpy(out.data() + (size_t)32 * i, d.ptr(i), 32);
  return out;
}

// ------------------------------------------------------------------------------------------------ OrbExtractor
class OrbExtractor {
 public:
  // src/orbExtractor.hpp:33-34; maxWidth / maxHeight size the device buffers (the reference allocates per call)
  OrbExtractor(int _keyPointsNum, float _scaleFactor, int _levelsNum, int _initFastThd, int _minFastThd, int maxWidth = 1920,
               int maxHeight = 1080, int device = 0)
      : m_impl(_keyPointsNum, _scaleFactor, _levelsNum, _initFastThd, _minFastThd, maxWidth, maxHeight, device),
        m_keepPyramid(false) {}
  ~OrbExtractor() {}

  // src/orbExtractor.hpp:39-40 / src/orbExtractor.cpp:355-399
  void extractAndCompute(const cv::InputArray& _image, std::vector<cv::KeyPoint>& _keyPoints, cv::OutputArray& _descriptors) {
    _keyPoints.clear();
    if (_image.empty()) return;  // :356-358
    const cv::Mat image = _image.getMat();
    std::vector<YdKeyPoint> kps;
    std::vector<uint8_t> desc;
    m_impl.extractAndCompute(YDORB_NAMESPACE::ImageView(image.data, image.cols, image.rows, (int)image.step), kps, desc);
    if (kps.empty()) {
      _descriptors.release();  // :373-375
    } else {
      _descriptors.create((int)kps.size(), 32, CV_8UC1);
      cv::Mat out = _descriptors.getMat();
      for (size_t i = 0; i < kps.size(); i++) std::memcpy(out.ptr((int)i), desc.data() + 32 * i, 32);
    }
    _keyPoints.resize(kps.size());
    if (!kps.empty()) std::memcpy(static_cast<void*>(_keyPoints.data()), kps.data(), sizeof(YdKeyPoint) * kps.size());
    if (m_keepPyramid) m_v_imagePyramid = getImagePyramid();
  }

  int inline getKeyPointsNum() { return m_impl.getLevelsNum(); }  // sic: the reference returns the level count (:42)
  float inline getScaleFactor() { return m_impl.getScaleFactor(); }
  int inline getLevelsNum() { return m_impl.getLevelsNum(); }
  std::vector<float> inline getScaleFactors() { return m_impl.getScaleFactors(); }
  std::vector<float> inline getInvScaleFactors() { return m_impl.getInvScaleFactors(); }
  std::vector<float> inline getScaleFactorSquares() { return m_impl.getScaleFactorSquares(); }
  std::vector<float> inline getInvScaleFactorSquares() { return m_impl.getInvScaleFactorSquares(); }

  // :49-50.  The levels live on the device; they are copied to the host on request only (the stereo matcher reads them
  // on the device).  keepPyramid(true) refreshes m_v_imagePyramid after every extractAndCompute, as the reference does.
  std::vector<cv::Mat> getImagePyramid() {
    std::vector<cv::Mat> levels(m_impl.getLevelsNum());
    for (int l = 0; l < m_impl.getLevelsNum(); l++) {
      int w = 0, h = 0;
      const std::vector<uint8_t> px = m_impl.getImagePyramidLevel(l, &w, &h, false);
      levels[l].create(h, w, CV_8UC1);
      for (int y = 0; y < h; y++) std::memcpy(levels[l].ptr(y), px.data() + (size_t)y * w, (size_t)w);
    }
    return levels;
  }
  void keepPyramid(bool on) { m_keepPyramid = on; }
  std::vector<cv::Mat> m_v_imagePyramid;

  YDORB_NAMESPACE::OrbExtractor& impl() { return m_impl; }

 private:
  YDORB_NAMESPACE::OrbExtractor m_impl;
  bool m_keepPyramid;
};

// ------------------------------------------------------------------------------------------------ OrbMatcher pieces
// src/orbMatcher.hpp:28 / src/orbMatcher.cpp:11-23
inline int computeDescriptorsDistance(YDORB_NAMESPACE::OrbMatcher& m, const cv::Mat& _firstDescriptor, const cv::Mat& _secondDescriptor) {
  return m.computeDescriptorsDistance(_firstDescriptor.ptr(0), _secondDescriptor.ptr(0));
}

// the device-side feature state of a reference Frame: key points, descriptors, rightX, image bounds, 64 x 48 grid
template <class FrameT>
std::unique_ptr<YDORB_NAMESPACE::FrameFeatures> frameFeatures(YDORB_NAMESPACE::OrbMatcher& m, const FrameT& f) {
  std::vector<YdKeyPoint> kps(f.m_v_keyPoints.size());
  if (!kps.empty()) std::memcpy(static_cast<void*>(kps.data()), f.m_v_keyPoints.data(), sizeof(YdKeyPoint) * kps.size());
  return std::unique_ptr<YDORB_NAMESPACE::FrameFeatures>(new YDORB_NAMESPACE::FrameFeatures(
      m, kps, descriptor_rows(f.m_cvMat_descriptors), f.m_v_rightXcords, FrameT::m_flt_minX, FrameT::m_flt_minY, FrameT::m_flt_maxX,
      FrameT::m_flt_maxY, f.m_v_scaleFactors));
}

// src/frame.hpp:83 / src/frame.cpp:337-361
template <class FrameT>
std::vector<int> getKeyPointsInArea(YDORB_NAMESPACE::OrbMatcher& m, const FrameT& f, const float& _posX, const float& _posY,
                                    const float& _radius, const int& _minScaleLevel = -1, const int& _maxScaleLevel = -1) {
  return frameFeatures(m, f)->getKeyPointsInArea(_posX, _posY, _radius, _minScaleLevel, _maxScaleLevel);
}

// src/orbMatcher.hpp:32 / src/orbMatcher.cpp:24-64: fills _frame.m_v_sptrMapPoints like the reference, returns its count
template <class FrameT, class MapPointPtr>
int searchByProjectionInFrameAndMapPoint(YDORB_NAMESPACE::OrbMatcher& m, FrameT& _frame, const std::vector<MapPointPtr>& _vSptrMapPoints,
                                         const float _thd = 3) {
  const size_t n = _frame.m_v_keyPoints.size(), nmp = _vSptrMapPoints.size();
  std::vector<YdMapPointProj> mps(nmp);
  std::vector<uint8_t> mpDesc(nmp * 32), obs(nmp), occupied(n);
  for (size_t i = 0; i < nmp; i++) {
    const MapPointPtr& p = _vSptrMapPoints[i];
    mps[i].proj_x = p->m_flt_trackProjX;
    mps[i].proj_y = p->m_flt_trackProjY;
    mps[i].proj_right_x = p->m_flt_trackProjRightX;
    mps[i].view_cos = p->m_flt_trackViewCos;
    mps[i].level = p->m_int_trackScaleLevel;
    mps[i].active = (p->m_b_isTrackInView && !p->isBad()) ? 1 : 0;  // :27
    const cv::Mat d = p->getDescriptor();
    if (!d.empty()) std::memcpy(mpDesc.data() + 32 * i, d.ptr(0), 32);
    obs[i] = p->getObservationsNum() > 0 ? 1 : 0;
  }
  for (size_t i = 0; i < n; i++)
    occupied[i] = (_frame.m_v_sptrMapPoints[i] && _frame.m_v_sptrMapPoints[i]->getObservationsNum() > 0) ? 1 : 0;  // :40
  std::unique_ptr<YDORB_NAMESPACE::FrameFeatures> ff = frameFeatures(m, _frame);
  std::vector<int32_t> match;
  const int ret = m.searchByProjectionInFrameAndMapPoint(*ff, mps, mpDesc.data(), obs.data(), occupied.data(), match, _thd);
  for (size_t i = 0; i < n; i++)
    if (match[i] >= 0) _frame.m_v_sptrMapPoints[i] = _vSptrMapPoints[match[i]];  // :57
  return ret;
}

// src/frame.hpp:96 / src/frame.cpp:362-472 on the pyramids the two extractors hold from their last extractAndCompute:
// fills _frame.m_v_rightXcords and _frame.m_v_depth
template <class FrameT>
void computeStereoMatches(YDORB_NAMESPACE::OrbMatcher& m, FrameT& _frame, OrbExtractor& left, OrbExtractor& right) {
  std::vector<YdKeyPoint> kl(_frame.m_v_keyPoints.size()), kr(_frame.m_v_rightKeyPoints.size());
  if (!kl.empty()) std::memcpy(static_cast<void*>(kl.data()), _frame.m_v_keyPoints.data(), sizeof(YdKeyPoint) * kl.size());
  if (!kr.empty()) std::memcpy(static_cast<void*>(kr.data()), _frame.m_v_rightKeyPoints.data(), sizeof(YdKeyPoint) * kr.size());
  YDORB_NAMESPACE::computeStereoMatches(m, left.impl(), right.impl(), kl, descriptor_rows(_frame.m_cvMat_descriptors), kr,
                                  descriptor_rows(_frame.m_cvMat_rightDescriptors), FrameT::m_flt_baseLineTimesFx,
                                  FrameT::m_flt_baseLine, _frame.m_v_rightXcords, _frame.m_v_depth, true);
}

}  // namespace ydorb_cv

#endif  // YDORB_CV_HPP
