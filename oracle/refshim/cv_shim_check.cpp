/*
 * cv_shim_check.cpp -- drives include/ydorb_cv.hpp (the OpenCV-typed drop-in over libydorb.so) next to the REFERENCE's
 * own classes, compiled unmodified, on the same inputs.  TEST INFRASTRUCTURE (tests/test_gpu_cv_shim.py runs it on the
 * GPU box; it is built here, where /root/reference exists, by oracle/refshim/Makefile).
 *
 *   usage: cv_shim_check left.raw right.raw width height n_features
 *   E1  ydorb_cv::OrbExtractor::extractAndCompute            == YDORBSLAM::OrbExtractor (inside the stereo Frame ctor)
 *   S1  ydorb_cv::computeStereoMatches(Frame&)               == Frame::computeStereoMatches (same ctor)
 *   G2  ydorb_cv::getKeyPointsInArea(Frame&)                 == Frame::getKeyPointsInArea
 *   M0  ydorb_cv::computeDescriptorsDistance                 == OrbMatcher::computeDescriptorsDistance
 *   M1  ydorb_cv::searchByProjectionInFrameAndMapPoint       == OrbMatcher::searchByProjectionInFrameAndMapPoint
 * Prints one line per check and exits non-zero on the first difference.
 */
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <vector>

#include "frame.hpp"
#include "mapPoint.hpp"
#include "orbMatcher.hpp"
#include "ydorb_cv.hpp"

namespace YDORBSLAM {
std::vector<cv::Mat> Converter::descriptors_cvMat_cvMatVector(const cv::Mat& d) {
  std::vector<cv::Mat> v;
  for (int i = 0; i < d.rows; i++) v.push_back(d.row(i));
  return v;
}
}  // namespace YDORBSLAM

namespace {

struct ProbePoint : YDORBSLAM::MapPoint {
  ProbePoint(const cv::Mat& pos, YDORBSLAM::Frame& f, int idx) : YDORBSLAM::MapPoint(pos, std::shared_ptr<YDORBSLAM::Map>(), f, idx) {}
  void observations(int n) { m_int_observationsNum = n; }
};

std::vector<uint8_t> read_raw(const char* path, size_t n) {
  std::vector<uint8_t> v(n);
  FILE* f = fopen(path, "rb");
  if (!f || fread(v.data(), 1, n, f) != n) { fprintf(stderr, "cannot read %s\n", path); exit(2); }
  fclose(f);
  return v;
}

int fail(const char* what) {
  printf("FAIL %s\n", what);
  return 1;
}

bool same_kps(const std::vector<cv::KeyPoint>& a, const std::vector<cv::KeyPoint>& b) {
  return a.size() == b.size() && (a.empty() || memcmp(a.data(), b.data(), a.size() * sizeof(cv::KeyPoint)) == 0);
}
bool same_desc(const cv::Mat& a, const cv::Mat& b) {
  if (a.rows != b.rows) return false;
  for (int i = 0; i < a.rows; i++)
    if (memcmp(a.ptr(i), b.ptr(i), 32) != 0) return false;
  return true;
}
bool same_bits(const std::vector<float>& a, const std::vector<float>& b) {
  return a.size() == b.size() && (a.empty() || memcmp(a.data(), b.data(), a.size() * sizeof(float)) == 0);
}

}  // namespace

int main(int argc, char** argv) {
  using namespace YDORBSLAM;
  if (argc < 6) { fprintf(stderr, "usage: %s left.raw right.raw width height n_features\n", argv[0]); return 2; }
  const int w = atoi(argv[3]), h = atoi(argv[4]), nf = atoi(argv[5]);
  std::vector<uint8_t> lbuf = read_raw(argv[1], (size_t)w * h), rbuf = read_raw(argv[2], (size_t)w * h);
  cv::Mat L(h, w, CV_8UC1, lbuf.data(), (size_t)w), R(h, w, CV_8UC1, rbuf.data(), (size_t)w);
  const float fx = 718.856f, fy = 718.856f, cx = 607.1928f, cy = 185.2157f, bf = 386.1448f;
  cv::Mat K = cv::Mat::eye(3, 3, CV_32F);
  K.at<float>(0, 0) = fx; K.at<float>(1, 1) = fy; K.at<float>(0, 2) = cx; K.at<float>(1, 2) = cy;
  cv::Mat dist = cv::Mat::zeros(4, 1, CV_32F);

  // ---- the reference: stereo Frame constructor (two CPU extractors on two threads, computeStereoMatches, grid)
  Frame::m_b_isComputeInit = true;
  Frame::m_flt_baseLine = 0.f;  // first frame of a run: maxD = bf / 0 (src/frame.cpp:101 runs before :108)
  Frame ref(L, R, 0.0, K, dist, dist, bf, 35.0f, std::make_shared<OrbExtractor>(nf, 1.2f, 8, 20, 7),
            std::make_shared<OrbExtractor>(nf, 1.2f, 8, 20, 7), std::shared_ptr<DBoW3::Vocabulary>());
  printf("reference frame: %zu left, %zu right key points\n", ref.m_v_keyPoints.size(), ref.m_v_rightKeyPoints.size());

  // ---- E1 through the OpenCV-typed drop-in
  ydorb_cv::OrbExtractor exl(nf, 1.2f, 8, 20, 7, w, h), exr(nf, 1.2f, 8, 20, 7, w, h);
  std::vector<cv::KeyPoint> kl, kr;
  cv::Mat dl, dr;
  exl.extractAndCompute(L, kl, dl);
  exr.extractAndCompute(R, kr, dr);
  if (!same_kps(kl, ref.m_v_keyPoints) || !same_kps(kr, ref.m_v_rightKeyPoints)) return fail("E1 key points");
  if (!same_desc(dl, ref.m_cvMat_descriptors) || !same_desc(dr, ref.m_cvMat_rightDescriptors)) return fail("E1 descriptors");
  printf("OK E1 extractAndCompute: key points and descriptors bit-identical\n");
  {
    const std::vector<cv::Mat> pyr = exl.getImagePyramid();
    for (int l = 0; l < 8; l++) {
      const cv::Mat& a = pyr[l];
      const cv::Mat& b = ref.m_sptr_leftOrbExtractor->m_v_imagePyramid[l];
      if (a.rows != b.rows || a.cols != b.cols) return fail("E2 pyramid level size");
      for (int y = 0; y < a.rows; y++)
        if (memcmp(a.ptr(y), b.ptr(y), (size_t)a.cols) != 0) return fail("E2 pyramid level bytes");
    }
    printf("OK E2 getImagePyramid: 8 levels bit-identical\n");
  }

  // ---- S1 on a copy of the reference frame whose stereo results are wiped
  ydorb_abi::OrbMatcher gm(0.8f, true);  // the C-ABI matcher handle (ydorb.hpp, compiled with -DYDORB_NAMESPACE=ydorb_abi)
  {
    Frame g(ref);
    g.m_v_rightXcords.clear();
    g.m_v_depth.clear();
    const float keep = Frame::m_flt_baseLine;
    Frame::m_flt_baseLine = 0.f;  // what the reference constructor saw
    ydorb_cv::computeStereoMatches(gm, g, exl, exr);
    Frame::m_flt_baseLine = keep;
    if (!same_bits(g.m_v_rightXcords, ref.m_v_rightXcords) || !same_bits(g.m_v_depth, ref.m_v_depth)) return fail("S1 rightX / depth");
    int matched = 0;
    for (float v : g.m_v_rightXcords) matched += v >= 0;
    printf("OK S1 computeStereoMatches: rightX and depth bit-identical (%d stereo matches)\n", matched);
  }

  // ---- G2
  {
    unsigned s = 12345;
    int nonempty = 0;
    for (int t = 0; t < 200; t++) {
      s = s * 1664525u + 1013904223u; const float x = (float)(s >> 8) / 16777216.f * (w + 40) - 20;
      s = s * 1664525u + 1013904223u; const float y = (float)(s >> 8) / 16777216.f * (h + 40) - 20;
      s = s * 1664525u + 1013904223u; const float r = 1.f + (float)(s >> 8) / 16777216.f * 80.f;
      s = s * 1664525u + 1013904223u; const int mn = (int)((s >> 8) % 9) - 1;
      s = s * 1664525u + 1013904223u; const int mx = (int)((s >> 8) % 9) - 1;
      const std::vector<int> a = ydorb_cv::getKeyPointsInArea(gm, ref, x, y, r, mn, mx), b = ref.getKeyPointsInArea(x, y, r, mn, mx);
      if (a != b) return fail("G2 getKeyPointsInArea");
      nonempty += !b.empty();
    }
    printf("OK G2 getKeyPointsInArea: 200 windows identical, %d non-empty\n", nonempty);
  }

  // ---- M0
  for (int i = 0; i + 1 < ref.m_cvMat_descriptors.rows && i < 300; i++) {
    const cv::Mat a = ref.m_cvMat_descriptors.row(i), b = ref.m_cvMat_descriptors.row(i + 1);
    if (ydorb_cv::computeDescriptorsDistance(gm, a, b) != YDORBSLAM::OrbMatcher::computeDescriptorsDistance(a, b)) return fail("M0");
  }
  printf("OK M0 computeDescriptorsDistance\n");

  // ---- M1: map points built by the reference's own constructor, projected just outside the window radius in x
  {
    const int n = (int)ref.m_v_keyPoints.size(), nmp = 3 * n;
    const float thd = 3.f;
    Frame src;
    src.m_v_keyPoints.assign(nmp + 1, cv::KeyPoint());
    src.m_v_scaleFactors = ref.m_v_scaleFactors;
    src.m_cvMat_descriptors.create(nmp + 1, 32, CV_8UC1);
    for (int i = 0; i <= nmp; i++) {
      memcpy(src.m_cvMat_descriptors.ptr(i), ref.m_cvMat_descriptors.ptr((i * 7919) % n), 32);
      src.m_cvMat_descriptors.ptr(i)[i % 32] ^= (uint8_t)(1u << (i % 8));
      src.m_cvMat_descriptors.ptr(i)[(i * 5) % 32] ^= (uint8_t)(1u << ((i / 3) % 8));
    }
    src.setCameraPoseByTransform_c2w(cv::Mat::eye(4, 4, CV_32F));
    cv::Mat pos = (cv::Mat_<float>(3, 1) << 0.f, 0.f, 1.f);
    std::vector<std::shared_ptr<MapPoint>> points(nmp);
    for (int i = 0; i < nmp; i++) {
      const cv::KeyPoint& k = ref.m_v_keyPoints[(i * 7919) % n];
      std::shared_ptr<ProbePoint> mp = std::make_shared<ProbePoint>(pos, src, i);
      mp->observations(i % 8 != 0 ? 1 : 0);
      mp->m_b_isTrackInView = i % 16 != 0;
      mp->m_int_trackScaleLevel = k.octave > 0 && i % 3 == 0 ? k.octave - 1 : k.octave;
      const float r = thd * 4.0f * ref.m_v_scaleFactors[mp->m_int_trackScaleLevel];
      mp->m_flt_trackProjX = k.pt.x + ((i & 1) ? 1.f : -1.f) * (r + 0.5f + (float)(i % 5));
      mp->m_flt_trackProjY = k.pt.y + (float)((i % 9) - 4);
      mp->m_flt_trackProjRightX = mp->m_flt_trackProjX - 10.f;
      mp->m_flt_trackViewCos = 0.9f;
      points[i] = mp;
    }
    std::shared_ptr<ProbePoint> held = std::make_shared<ProbePoint>(pos, src, nmp);
    held->observations(1);
    Frame a(ref), b(ref);
    a.m_v_sptrMapPoints.assign(n, std::shared_ptr<MapPoint>());
    for (int i = 0; i < n; i += 11) a.m_v_sptrMapPoints[i] = held;
    b.m_v_sptrMapPoints = a.m_v_sptrMapPoints;
    YDORBSLAM::OrbMatcher refm(0.8f, true);  // the reference's own class
    const int want = refm.searchByProjectionInFrameAndMapPoint(a, points, thd);
    const int got = ydorb_cv::searchByProjectionInFrameAndMapPoint(gm, b, points, thd);
    if (got != want) return fail("M1 return value");
    for (int i = 0; i < n; i++)
      if (a.m_v_sptrMapPoints[i].get() != b.m_v_sptrMapPoints[i].get()) return fail("M1 assignments");
    printf("OK M1 searchByProjectionInFrameAndMapPoint: %d matches, every slot identical\n", got);
  }
  printf("ALL OK\n");
  return 0;
}
