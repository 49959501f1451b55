/*
 * ref_frame_driver.cpp -- extern "C" entry points around the REFERENCE's own Frame and OrbMatcher.  TEST INFRASTRUCTURE.
 *
 * Linked with /root/reference/src/{frame,orbMatcher,mapPoint,keyFrame,map,keyFrameDatabase,orbExtractor}.cpp and
 * thirdParty/DBow3/src/{BowVector,FeatureVector,ScoringObject,QueryResults}.cpp, all compiled UNMODIFIED
 * (oracle/refshim/Makefile) against the OpenCV stand-in.  Nothing here restates the algorithm:
 *   M0  OrbMatcher::computeDescriptorsDistance          src/orbMatcher.cpp:11-23
 *   G1  Frame::assignKeyPointsToGrid                     src/frame.cpp:249-264  (through the stereo constructor)
 *   G2  Frame::getKeyPointsInArea                        src/frame.cpp:337-361
 *   S1  Frame::computeStereoMatches                      src/frame.cpp:362-472  (through the stereo constructor, which
 *       also runs the two extractors on two threads, :84-87)
 *   M1  OrbMatcher::searchByProjectionInFrameAndMapPoint src/orbMatcher.cpp:24-64, on MapPoint objects made by the
 *       reference's own MapPoint(pos, map, frame, idx) constructor (src/mapPoint.cpp:26-51)
 * What could NOT be built from reference sources and is stood in for here:
 *   src/converter.hpp        needs Eigen + g2o's cmake-generated config.h; its include guard is pre-defined and
 *                            stubs/converter_stub.hpp declares the one member frame.cpp names (defined below)
 *   DBoW3 Vocabulary.cpp / Database.cpp   need cv::FileStorage (YAML persistence); the members the linked objects
 *                            reference are defined below and abort if ever called (Frame::computeBoW is not on this path)
 */
#include <stdint.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <memory>
#include <set>
#include <vector>

#include "frame.hpp"
#include "keyFrame.hpp"
#include "mapPoint.hpp"
#include "orbMatcher.hpp"

namespace YDORBSLAM {
std::vector<cv::Mat> Converter::descriptors_cvMat_cvMatVector(const cv::Mat& d) {
  std::vector<cv::Mat> v;
  v.reserve(d.rows);
  for (int i = 0; i < d.rows; i++) v.push_back(d.row(i));
  return v;
}
}  // namespace YDORBSLAM

namespace {

struct RefKeyPoint {  // = cv::KeyPoint = OrcKeyPoint = YdKeyPoint (28 B)
  float x, y, size, angle, response;
  int32_t octave, class_id;
};

void put(const std::vector<cv::KeyPoint>& in, RefKeyPoint* out) {
  if (!out) return;
  for (size_t i = 0; i < in.size(); i++) {
    out[i].x = in[i].pt.x;
    out[i].y = in[i].pt.y;
    out[i].size = in[i].size;
    out[i].angle = in[i].angle;
    out[i].response = in[i].response;
    out[i].octave = in[i].octave;
    out[i].class_id = in[i].class_id;
  }
}

void put(const cv::Mat& d, uint8_t* out) {
  if (!out) return;
  for (int i = 0; i < d.rows; i++) memcpy(out + (size_t)32 * i, d.ptr(i), 32);
}

/* protected members are reachable from a derived class; nothing is overridden */
struct GridFrame : YDORBSLAM::Frame {
  void assign() { assignKeyPointsToGrid(); }
};

/* protected state of a MapPoint is reachable from a derived class; nothing is overridden */
struct ProbePoint : YDORBSLAM::MapPoint {
  ProbePoint(const cv::Mat& pos, YDORBSLAM::Frame& f, int idx) : YDORBSLAM::MapPoint(pos, std::shared_ptr<YDORBSLAM::Map>(), f, idx) {}
  ProbePoint(const cv::Mat& pos, std::shared_ptr<YDORBSLAM::Map> map, YDORBSLAM::Frame& f, int idx)
      : YDORBSLAM::MapPoint(pos, map, f, idx) {}
  void observations(int n) { m_int_observationsNum = n; }
  void distances(float mn, float mx) { m_flt_minDistance = mn; m_flt_maxDistance = mx; }
  void bad(bool b) { m_b_bad = b; }
  void normal(float x, float y, float z) { m_cvMat_normalVector = (cv::Mat_<float>(3, 1) << x, y, z); }
};

struct RefMapPoint3d {  // = OrcMapPoint3d = YdMapPoint3d
  float pos[3], normal[3], max_distance, min_distance;
};

struct RefMapPointProj {  // = OrcMapPointProj = YdMapPointProj
  float proj_x, proj_y, proj_right_x, view_cos;
  int32_t level;
  int32_t active;
};

void fill_features(YDORBSLAM::Frame* f, const uint8_t* desc, const float* right_x, const float* sf, int n_levels) {
  const int n = (int)f->m_v_keyPoints.size();
  f->m_cvMat_descriptors.create(n, 32, CV_8UC1);
  for (int i = 0; i < n; i++) memcpy(f->m_cvMat_descriptors.ptr(i), desc + (size_t)32 * i, 32);
  f->m_v_rightXcords.assign(n, -1.0f);
  if (right_x) f->m_v_rightXcords.assign(right_x, right_x + n);
  f->m_v_scaleFactors.assign(sf, sf + n_levels);
  f->m_int_scaleLevelsNum = n_levels;
  f->m_flt_scaleFactor = n_levels > 1 ? sf[1] : 1.2f;
  f->m_flt_logScaleFactor = log(f->m_flt_scaleFactor);  // as the Frame constructors do (src/frame.cpp:78)
  f->m_v_sptrMapPoints.assign(n, std::shared_ptr<YDORBSLAM::MapPoint>());
}

}  // namespace

extern "C" {

/* descriptors, rightX and scale factors of a frame made by ref_frame_from_keypoints */
void ref_frame_set_features(void* p, const uint8_t* desc, const float* right_x, const float* sf, int n_levels) {
  fill_features(static_cast<YDORBSLAM::Frame*>(p), desc, right_x, sf, n_levels);
}

/* M1.  kp_occupied[i] != 0: slot i holds a map point with observations > 0 before the call.  match_out[i] = index of the
 * map point the call assigned to key point i, or -1.  Returns the function's return value. */
int ref_search_by_projection(void* p, const RefMapPointProj* mps, const uint8_t* mp_desc, const uint8_t* mp_obs_positive,
                             int nmp, const uint8_t* kp_occupied, float thd, float ratio, int32_t* match_out) {
  using namespace YDORBSLAM;
  Frame* f = static_cast<Frame*>(p);
  const int n = (int)f->m_v_keyPoints.size();
  /* a source frame at the identity pose whose "key points" carry the map points' descriptors: the reference's own
   * constructor then builds each MapPoint (descriptor copy, normal, distances) */
  Frame src;
  src.m_v_keyPoints.assign(nmp + 1, cv::KeyPoint());
  src.m_v_scaleFactors = f->m_v_scaleFactors;
  src.m_cvMat_descriptors.create(nmp + 1, 32, CV_8UC1);
  for (int i = 0; i < nmp; i++) memcpy(src.m_cvMat_descriptors.ptr(i), mp_desc + (size_t)32 * i, 32);
  memset(src.m_cvMat_descriptors.ptr(nmp), 0, 32);
  src.setCameraPoseByTransform_c2w(cv::Mat::eye(4, 4, CV_32F));
  cv::Mat pos = (cv::Mat_<float>(3, 1) << 0.f, 0.f, 1.f);
  std::vector<std::shared_ptr<MapPoint>> points(nmp);
  for (int i = 0; i < nmp; i++) {
    std::shared_ptr<ProbePoint> mp = std::make_shared<ProbePoint>(pos, src, i);
    mp->observations(mp_obs_positive && mp_obs_positive[i] ? 1 : 0);
    mp->m_b_isTrackInView = mps[i].active != 0;
    mp->m_int_trackScaleLevel = mps[i].level;
    mp->m_flt_trackProjX = mps[i].proj_x;
    mp->m_flt_trackProjY = mps[i].proj_y;
    mp->m_flt_trackProjRightX = mps[i].proj_right_x;
    mp->m_flt_trackViewCos = mps[i].view_cos;
    points[i] = mp;
  }
  std::shared_ptr<ProbePoint> held = std::make_shared<ProbePoint>(pos, src, nmp);
  held->observations(1);
  for (int i = 0; i < n; i++)
    f->m_v_sptrMapPoints[i] = (kp_occupied && kp_occupied[i]) ? std::static_pointer_cast<MapPoint>(held) : std::shared_ptr<MapPoint>();
  OrbMatcher matcher(ratio, true);
  const int ret = matcher.searchByProjectionInFrameAndMapPoint(*f, points, thd);
  std::map<const MapPoint*, int> index;
  for (int i = 0; i < nmp; i++) index[points[i].get()] = i;
  for (int i = 0; i < n; i++) {
    const MapPoint* q = f->m_v_sptrMapPoints[i].get();
    match_out[i] = (q && q != held.get()) ? index[q] : -1;
  }
  return ret;
}

/* M2  OrbMatcher::searchByProjectionInLastAndCurrentFrame (src/orbMatcher.cpp:65-155).  `p` is the CURRENT frame (made by
 * ref_frame_from_keypoints + ref_frame_set_features), its pose is the identity.  The LAST frame is built here: key points
 * last_kps, one MapPoint per key point at world position mp_pos[i] (built by the reference's own constructor; a NaN x
 * means "no map point"), outlier flags, pose [I | last_t] (its z against the base line decides forward / backward,
 * :79-80).  cam = {fx, fy, cx, cy, bf, baseline}.  match_out[i] = index in the last frame of the map point key point i
 * of the current frame holds after the call, -1 = none.  Returns the function's return value. */
int ref_search_last_current(void* p, const RefKeyPoint* last_kps, int n_last, const float* mp_pos, const uint8_t* mp_desc,
                            const uint8_t* mp_obs_positive, const uint8_t* outlier, const float* last_t, const float* cam,
                            const uint8_t* kp_occupied, float thd, float ratio, int check_orientation, int32_t* match_out) {
  using namespace YDORBSLAM;
  Frame* cur = static_cast<Frame*>(p);
  const int n = (int)cur->m_v_keyPoints.size();
  Frame::m_flt_fx = cam[0]; Frame::m_flt_fy = cam[1]; Frame::m_flt_cx = cam[2]; Frame::m_flt_cy = cam[3];
  Frame::m_flt_baseLineTimesFx = cam[4];
  Frame::m_flt_baseLine = cam[5];
  cur->setCameraPoseByTransform_c2w(cv::Mat::eye(4, 4, CV_32F));
  Frame src;  // carries the map points' descriptors for the reference's MapPoint constructor
  src.m_v_keyPoints.assign(n_last + 1, cv::KeyPoint());
  src.m_v_scaleFactors = cur->m_v_scaleFactors;
  src.m_cvMat_descriptors.create(n_last + 1, 32, CV_8UC1);
  for (int i = 0; i < n_last; i++) memcpy(src.m_cvMat_descriptors.ptr(i), mp_desc + (size_t)32 * i, 32);
  memset(src.m_cvMat_descriptors.ptr(n_last), 0, 32);
  src.setCameraPoseByTransform_c2w(cv::Mat::eye(4, 4, CV_32F));
  Frame last;
  last.m_v_keyPoints.resize(n_last);
  for (int i = 0; i < n_last; i++) {
    cv::KeyPoint k;
    k.pt.x = last_kps[i].x; k.pt.y = last_kps[i].y; k.size = last_kps[i].size; k.angle = last_kps[i].angle;
    k.response = last_kps[i].response; k.octave = last_kps[i].octave; k.class_id = last_kps[i].class_id;
    last.m_v_keyPoints[i] = k;
  }
  last.m_v_isOutliers.assign(n_last, false);
  last.m_v_sptrMapPoints.assign(n_last, std::shared_ptr<MapPoint>());
  std::map<const MapPoint*, int> index;
  for (int i = 0; i < n_last; i++) {
    last.m_v_isOutliers[i] = outlier && outlier[i];
    if (mp_pos[3 * i] != mp_pos[3 * i]) continue;  // NaN: this key point has no map point
    cv::Mat pos = (cv::Mat_<float>(3, 1) << mp_pos[3 * i], mp_pos[3 * i + 1], mp_pos[3 * i + 2]);
    std::shared_ptr<ProbePoint> mp = std::make_shared<ProbePoint>(pos, src, i);
    mp->observations(mp_obs_positive && mp_obs_positive[i] ? 1 : 0);
    last.m_v_sptrMapPoints[i] = mp;
    index[mp.get()] = i;
  }
  cv::Mat T = cv::Mat::eye(4, 4, CV_32F);
  for (int k = 0; k < 3; k++) T.at<float>(k, 3) = last_t[k];
  last.setCameraPoseByTransform_c2w(T);
  cv::Mat pos0 = (cv::Mat_<float>(3, 1) << 0.f, 0.f, 1.f);
  std::shared_ptr<ProbePoint> held = std::make_shared<ProbePoint>(pos0, src, n_last);
  held->observations(1);
  for (int i = 0; i < n; i++)
    cur->m_v_sptrMapPoints[i] = (kp_occupied && kp_occupied[i]) ? std::static_pointer_cast<MapPoint>(held) : std::shared_ptr<MapPoint>();
  OrbMatcher matcher(ratio, check_orientation != 0);
  const int ret = matcher.searchByProjectionInLastAndCurrentFrame(*cur, last, thd);
  for (int i = 0; i < n; i++) {
    const MapPoint* q = cur->m_v_sptrMapPoints[i].get();
    match_out[i] = (q && q != held.get()) ? index[q] : -1;
  }
  return ret;
}

/* M4a OrbMatcher::searchByProjectionInSim (src/orbMatcher.cpp:240-302) with the identity similarity.  `p` (made by
 * ref_frame_from_keypoints + ref_frame_set_features) becomes the KeyFrame through the reference's constructor; the map
 * points are built like in M3.  kp_occupied[i]: _vSptrMatchedMapPoints[i] is non-null on entry.  match_out[i] = index of
 * the map point the call put into slot i, -1 = none / was occupied.  dbg as in ref_search_keyframe_current. */
int ref_search_in_sim(void* p, int n_mp, const float* mp_pos, const uint8_t* mp_desc, const float* min_dist, const float* max_dist,
                      const uint8_t* is_bad, const float* cam, const uint8_t* kp_occupied, int thd, float ratio,
                      int32_t* match_out, float* dbg) {
  using namespace YDORBSLAM;
  Frame* fr = static_cast<Frame*>(p);
  const int n = (int)fr->m_v_keyPoints.size();
  Frame::m_flt_fx = cam[0]; Frame::m_flt_fy = cam[1]; Frame::m_flt_cx = cam[2]; Frame::m_flt_cy = cam[3];
  Frame::m_flt_baseLineTimesFx = cam[4];
  Frame::m_flt_baseLine = cam[5];
  fr->setCameraPoseByTransform_c2w(cv::Mat::eye(4, 4, CV_32F));
  fr->m_int_keyPointsNum = n;
  std::shared_ptr<KeyFrame> kf = std::make_shared<KeyFrame>(*fr, std::shared_ptr<Map>(), std::shared_ptr<KeyFrameDatabase>());
  Frame src;
  src.m_v_keyPoints.assign(n_mp + 1, cv::KeyPoint());
  src.m_v_scaleFactors = fr->m_v_scaleFactors;
  src.m_cvMat_descriptors.create(n_mp + 1, 32, CV_8UC1);
  for (int i = 0; i < n_mp; i++) memcpy(src.m_cvMat_descriptors.ptr(i), mp_desc + (size_t)32 * i, 32);
  memset(src.m_cvMat_descriptors.ptr(n_mp), 0, 32);
  src.setCameraPoseByTransform_c2w(cv::Mat::eye(4, 4, CV_32F));
  std::vector<std::shared_ptr<MapPoint>> points(n_mp);
  std::map<const MapPoint*, int> index;
  for (int i = 0; i < n_mp; i++) {
    if (mp_pos[3 * i] != mp_pos[3 * i]) continue;  // NaN: a null entry of _vSptrMapPoints
    cv::Mat pos = (cv::Mat_<float>(3, 1) << mp_pos[3 * i], mp_pos[3 * i + 1], mp_pos[3 * i + 2]);
    std::shared_ptr<ProbePoint> mp = std::make_shared<ProbePoint>(pos, src, i);
    mp->distances(min_dist[i], max_dist[i]);
    mp->bad(is_bad && is_bad[i]);
    points[i] = mp;
    index[mp.get()] = i;
  }
  cv::Mat pos0 = (cv::Mat_<float>(3, 1) << 0.f, 0.f, 1.f);
  std::shared_ptr<MapPoint> held = std::make_shared<ProbePoint>(pos0, src, n_mp);
  std::vector<std::shared_ptr<MapPoint>> matched(n);
  for (int i = 0; i < n; i++)
    if (kp_occupied && kp_occupied[i]) matched[i] = held;
  cv::Mat sim = cv::Mat::eye(4, 4, CV_32F);
  if (dbg) {  // the expressions of :242-268 on the same objects
    cv::Mat sR = sim.rowRange(0, 3).colRange(0, 3);
    const float scale = sqrt(sR.row(0).dot(sR.row(0)));
    cv::Mat R = sR / scale;
    cv::Mat t = sim.rowRange(0, 3).col(3);
    cv::Mat origin = -R.t() * t;
    for (int i = 0; i < n_mp; i++) {
      float* d = dbg + 4 * i;
      d[0] = d[1] = d[2] = d[3] = 0.f;
      std::shared_ptr<MapPoint> mp = points[i];
      if (!mp || mp->isBad()) continue;
      cv::Mat pw = mp->getPosInWorld();
      cv::Mat pc = R * pw + t;
      const float x = pc.at<float>(0), y = pc.at<float>(1), z = pc.at<float>(2);
      const float u = Frame::m_flt_fx * x / z + Frame::m_flt_cx;
      const float v = Frame::m_flt_fy * y / z + Frame::m_flt_cy;
      const float maxD = mp->getMaxDistanceInvariance(), minD = mp->getMinDistanceInvariance();
      cv::Mat vec = pw - origin;
      const float distance = cv::norm(vec);
      d[0] = u;
      d[1] = v;
      d[2] = (float)mp->predictScaleLevel(distance, kf);
      d[3] = (z >= 0.0 && kf->isInImage(u, v) && distance >= minD && distance <= maxD && vec.dot(mp->getNormal()) >= 0.5 * distance) ? 1.f : 0.f;
    }
  }
  OrbMatcher matcher(ratio, true);
  const int ret = matcher.searchByProjectionInSim(kf, sim, points, matched, thd);
  for (int i = 0; i < n; i++) {
    const MapPoint* q = matched[i].get();
    match_out[i] = (q && q != held.get()) ? index[q] : -1;
  }
  return ret;
}

/* M4b OrbMatcher::searchBySim3 (src/orbMatcher.cpp:566-681).  Two KeyFrames at the identity pose built by the reference's
 * constructor from (key points, descriptors, one MapPoint per key point at mp_pos (NaN x: none) with its own descriptor,
 * invariance distances, bad flag); the grids by its assignKeyPointsToGrid.  pre1[i] >= 0: _vSptrMatchedMapPoints[i] is
 * non-null on entry and that map point is observed by key frame 2 at key point pre1[i] (:580-588).
 * match_out[i] = key point of key frame 2 whose map point the call put into slot i, -1 = none, -2 = pre-matched.
 * dbg1 / dbg2: per map point of side 1 / 2 {projX, projY, predicted level, gates passed} in the OTHER key frame. */
int ref_search_by_sim3(const RefKeyPoint* kps1, const uint8_t* desc1, int n1, const float* pos1, const uint8_t* mpd1,
                       const float* mind1, const float* maxd1, const uint8_t* bad1, const int32_t* pre1, const RefKeyPoint* kps2,
                       const uint8_t* desc2, int n2, const float* pos2, const uint8_t* mpd2, const float* mind2,
                       const float* maxd2, const uint8_t* bad2, const float* cam, const float* bounds, const float* sf,
                       int n_levels, float thd, int32_t* match_out, float* dbg1, float* dbg2) {
  using namespace YDORBSLAM;
  Frame::m_flt_fx = cam[0]; Frame::m_flt_fy = cam[1]; Frame::m_flt_cx = cam[2]; Frame::m_flt_cy = cam[3];
  Frame::m_flt_baseLineTimesFx = cam[4];
  Frame::m_flt_baseLine = cam[5];
  Frame::m_flt_minX = bounds[0]; Frame::m_flt_minY = bounds[1]; Frame::m_flt_maxX = bounds[2]; Frame::m_flt_maxY = bounds[3];
  Frame::m_flt_gridCellWidthInv = static_cast<float>(Frame::m_int_gridColsNum) / (bounds[2] - bounds[0]);
  Frame::m_flt_gridCellHeightInv = static_cast<float>(Frame::m_int_gridRowsNum) / (bounds[3] - bounds[1]);
  const int nmax = std::max(n1, n2);
  Frame src;
  src.m_v_keyPoints.assign(nmax + 1, cv::KeyPoint());
  src.m_v_scaleFactors.assign(sf, sf + n_levels);
  src.setCameraPoseByTransform_c2w(cv::Mat::eye(4, 4, CV_32F));
  auto make = [&](const RefKeyPoint* kps, const uint8_t* desc, int n, const float* pos, const uint8_t* mpd, const float* mind,
                  const float* maxd, const uint8_t* bad, std::map<const MapPoint*, int>& index) {
    GridFrame f;
    f.m_v_keyPoints.resize(n);
    for (int i = 0; i < n; i++) {
      cv::KeyPoint k;
      k.pt.x = kps[i].x; k.pt.y = kps[i].y; k.size = kps[i].size; k.angle = kps[i].angle;
      k.response = kps[i].response; k.octave = kps[i].octave; k.class_id = kps[i].class_id;
      f.m_v_keyPoints[i] = k;
    }
    f.m_int_keyPointsNum = n;
    f.m_cvMat_descriptors.create(std::max(n, 1), 32, CV_8UC1);
    for (int i = 0; i < n; i++) memcpy(f.m_cvMat_descriptors.ptr(i), desc + (size_t)32 * i, 32);
    f.m_v_rightXcords.assign(n, -1.0f);
    f.m_v_scaleFactors.assign(sf, sf + n_levels);
    f.m_int_scaleLevelsNum = n_levels;
    f.m_flt_scaleFactor = n_levels > 1 ? sf[1] : 1.2f;
    f.m_flt_logScaleFactor = log(f.m_flt_scaleFactor);
    f.setCameraPoseByTransform_c2w(cv::Mat::eye(4, 4, CV_32F));
    f.assign();
    src.m_cvMat_descriptors.create(nmax + 1, 32, CV_8UC1);
    for (int i = 0; i < n; i++) memcpy(src.m_cvMat_descriptors.ptr(i), mpd + (size_t)32 * i, 32);
    f.m_v_sptrMapPoints.assign(n, std::shared_ptr<MapPoint>());
    for (int i = 0; i < n; i++) {
      if (pos[3 * i] != pos[3 * i]) continue;
      cv::Mat p3 = (cv::Mat_<float>(3, 1) << pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]);
      std::shared_ptr<ProbePoint> mp = std::make_shared<ProbePoint>(p3, src, i);
      mp->distances(mind[i], maxd[i]);
      mp->bad(bad && bad[i]);
      f.m_v_sptrMapPoints[i] = mp;
      index[mp.get()] = i;
    }
    return std::make_shared<KeyFrame>(f, std::shared_ptr<Map>(), std::shared_ptr<KeyFrameDatabase>());
  };
  std::map<const MapPoint*, int> index1, index2;
  std::shared_ptr<KeyFrame> kf1 = make(kps1, desc1, n1, pos1, mpd1, mind1, maxd1, bad1, index1);
  std::shared_ptr<KeyFrame> kf2 = make(kps2, desc2, n2, pos2, mpd2, mind2, maxd2, bad2, index2);
  // pre-matched slots: a map point observed by key frame 2 at pre1[i]
  std::vector<std::shared_ptr<MapPoint>> matched(n1);
  std::vector<std::shared_ptr<MapPoint>> held;
  cv::Mat pos0 = (cv::Mat_<float>(3, 1) << 0.f, 0.f, 1.f);
  src.m_cvMat_descriptors.create(nmax + 1, 32, CV_8UC1);
  for (int i = 0; i < n1; i++)
    if (pre1 && pre1[i] >= 0) {
      std::shared_ptr<ProbePoint> h = std::make_shared<ProbePoint>(pos0, src, nmax);
      h->addObservation(kf2, pre1[i]);
      matched[i] = h;
      held.push_back(h);
    }
  auto debug = [&](const std::shared_ptr<KeyFrame>& from, const std::shared_ptr<KeyFrame>& into, int n, float* dbg) {
    if (!dbg) return;
    cv::Mat R = into->getRotation_c2w(), t = into->getTranslation_c2w();
    const std::vector<std::shared_ptr<MapPoint>> pts = from->getMatchedMapPointsVec();
    for (int i = 0; i < n; i++) {
      float* d = dbg + 4 * i;
      d[0] = d[1] = d[2] = d[3] = 0.f;
      std::shared_ptr<MapPoint> mp = pts[i];
      if (!mp || mp->isBad()) continue;
      cv::Mat pw = mp->getPosInWorld();
      cv::Mat pc = R * pw + t;
      const float x = pc.at<float>(0), y = pc.at<float>(1), z = pc.at<float>(2);
      const float u = Frame::m_flt_fx * x / z + Frame::m_flt_cx;
      const float v = Frame::m_flt_fy * y / z + Frame::m_flt_cy;
      const float dist3D = cv::norm(pc);
      const float maxD = mp->getMaxDistanceInvariance(), minD = mp->getMinDistanceInvariance();
      d[0] = u;
      d[1] = v;
      d[2] = (float)mp->predictScaleLevel(dist3D, into);
      d[3] = (z >= 0.0 && into->isInImage(u, v) && dist3D >= minD && dist3D <= maxD) ? 1.f : 0.f;
    }
  };
  debug(kf1, kf2, n1, dbg1);
  debug(kf2, kf1, n2, dbg2);
  OrbMatcher matcher(0.6f, true);
  cv::Mat R12 = cv::Mat::eye(3, 3, CV_32F), t12 = cv::Mat::zeros(3, 1, CV_32F);
  const int ret = matcher.searchBySim3(kf1, kf2, matched, 1.0f, R12, t12, thd);
  for (int i = 0; i < n1; i++) {
    const MapPoint* q = matched[i].get();
    if (!q) match_out[i] = -1;
    else if (index2.count(q)) match_out[i] = index2[q];
    else match_out[i] = -2;
  }
  return ret;
}

/* M4  OrbMatcher::fuseByProjection (src/orbMatcher.cpp:682-745; mode 0) and OrbMatcher::fuseBySim3 (:746-807; mode 1,
 * similarity = scale * I | sim_t) on a KeyFrame at the identity pose built by the reference's constructor.
 * Key point i holds: nothing (slot_kind 0), a good map point (1) or a bad one (2), observing the key frame, with
 * slot_obs[i] observations counted before that one.  The nmp candidate map points carry mp_obs[j] observations.
 * After the call, entities are named by id: slot point i -> i, candidate j -> n + j, nothing -> -1.
 *   slot_out[i]   occupant of key point i           slot_repl[i] / mp_repl[j]   replacement of that entity
 *   slot_obs_out / mp_obs_out   observation counts  mp_bad_out[j]               isBad() of candidate j
 *   dbg[j] = {projX, projY, projRightX, predicted level, gates (:703-707 / :771-775 without the empty-window test)}  */
int ref_fuse(int mode, const RefKeyPoint* kps, const uint8_t* desc, const float* right_x, int n, const uint8_t* slot_kind,
             const int32_t* slot_obs, const float* pos, const uint8_t* mpd, const float* mind, const float* maxd,
             const float* normal, const uint8_t* bad, const int32_t* mp_obs, int nmp, const float* cam, const float* bounds,
             const float* sf, int n_levels, float thd, float sim_scale, const float* sim_t, int32_t* slot_out,
             int32_t* slot_repl, int32_t* slot_obs_out, int32_t* mp_repl, int32_t* mp_obs_out, uint8_t* mp_bad_out, float* dbg) {
  using namespace YDORBSLAM;
  Frame::m_flt_fx = cam[0]; Frame::m_flt_fy = cam[1]; Frame::m_flt_cx = cam[2]; Frame::m_flt_cy = cam[3];
  Frame::m_flt_baseLineTimesFx = cam[4];
  Frame::m_flt_baseLine = cam[5];
  Frame::m_flt_minX = bounds[0]; Frame::m_flt_minY = bounds[1]; Frame::m_flt_maxX = bounds[2]; Frame::m_flt_maxY = bounds[3];
  Frame::m_flt_gridCellWidthInv = static_cast<float>(Frame::m_int_gridColsNum) / (bounds[2] - bounds[0]);
  Frame::m_flt_gridCellHeightInv = static_cast<float>(Frame::m_int_gridRowsNum) / (bounds[3] - bounds[1]);
  std::shared_ptr<Map> map = std::make_shared<Map>();
  const int nmax = std::max(n, nmp);
  Frame src;
  src.m_v_keyPoints.assign(nmax + 1, cv::KeyPoint());
  src.m_v_scaleFactors.assign(sf, sf + n_levels);
  src.setCameraPoseByTransform_c2w(cv::Mat::eye(4, 4, CV_32F));
  src.m_cvMat_descriptors.create(nmax + 1, 32, CV_8UC1);
  GridFrame f;
  f.m_v_keyPoints.resize(n);
  for (int i = 0; i < n; i++) {
    cv::KeyPoint k;
    k.pt.x = kps[i].x; k.pt.y = kps[i].y; k.size = kps[i].size; k.angle = kps[i].angle;
    k.response = kps[i].response; k.octave = kps[i].octave; k.class_id = kps[i].class_id;
    f.m_v_keyPoints[i] = k;
  }
  f.m_int_keyPointsNum = n;
  fill_features(&f, desc, right_x, sf, n_levels);
  f.m_v_invScaleFactorSquares.resize(n_levels);
  for (int l = 0; l < n_levels; l++) f.m_v_invScaleFactorSquares[l] = 1.0f / (sf[l] * sf[l]);
  f.setCameraPoseByTransform_c2w(cv::Mat::eye(4, 4, CV_32F));
  f.assign();
  std::map<const MapPoint*, int> id;
  std::vector<std::shared_ptr<ProbePoint>> slot(n), cand(nmp);
  cv::Mat pos0 = (cv::Mat_<float>(3, 1) << 0.f, 0.f, 1.f);
  for (int i = 0; i < n; i++) memcpy(src.m_cvMat_descriptors.ptr(i), desc + (size_t)32 * i, 32);
  for (int i = 0; i < n; i++)
    if (slot_kind[i]) {
      slot[i] = std::make_shared<ProbePoint>(pos0, map, src, i);
      slot[i]->observations(slot_obs[i]);
      f.m_v_sptrMapPoints[i] = slot[i];
      id[slot[i].get()] = i;
    }
  std::shared_ptr<KeyFrame> kf = std::make_shared<KeyFrame>(f, map, std::shared_ptr<KeyFrameDatabase>());
  for (int i = 0; i < n; i++)
    if (slot[i]) {
      slot[i]->addObservation(kf, i);
      slot[i]->bad(slot_kind[i] == 2);
    }
  for (int j = 0; j < nmp; j++) memcpy(src.m_cvMat_descriptors.ptr(j), mpd + (size_t)32 * j, 32);
  std::vector<std::shared_ptr<MapPoint>> points(nmp);
  for (int j = 0; j < nmp; j++) {
    if (pos[3 * j] != pos[3 * j]) continue;  // null entry of the list
    cv::Mat p3 = (cv::Mat_<float>(3, 1) << pos[3 * j], pos[3 * j + 1], pos[3 * j + 2]);
    cand[j] = std::make_shared<ProbePoint>(p3, map, src, j);
    cand[j]->distances(mind[j], maxd[j]);
    cand[j]->normal(normal[3 * j], normal[3 * j + 1], normal[3 * j + 2]);
    cand[j]->observations(mp_obs[j]);
    cand[j]->bad(bad[j]);
    points[j] = cand[j];
    id[cand[j].get()] = n + j;
  }
  cv::Mat sim = cv::Mat::eye(4, 4, CV_32F);
  for (int r = 0; r < 3; r++) {
    sim.at<float>(r, r) = sim_scale;
    sim.at<float>(r, 3) = sim_t[r];
  }
  // the projections the calls are about to make, by the same cv::Mat operations
  cv::Mat R, t, origin;
  if (mode == 0) {
    R = kf->getRotation_c2w(); t = kf->getTranslation_c2w(); origin = kf->getCameraOriginInWorld();
  } else {
    cv::Mat sR = sim.rowRange(0, 3).colRange(0, 3);
    const float sc = sqrt(sR.row(0).dot(sR.row(0)));
    R = sR / sc;
    t = sim.rowRange(0, 3).col(3);
    origin = -R.t() * t;
  }
  for (int j = 0; j < nmp; j++) {
    float* d = dbg + 5 * j;
    d[0] = d[1] = d[2] = d[3] = d[4] = 0.f;
    if (!cand[j] || cand[j]->isBad()) continue;
    cv::Mat pw = cand[j]->getPosInWorld();
    cv::Mat pc = R * pw + t;
    const float x = pc.at<float>(0), y = pc.at<float>(1), z = pc.at<float>(2);
    const float u = Frame::m_flt_fx * x / z + Frame::m_flt_cx;
    const float v = Frame::m_flt_fy * y / z + Frame::m_flt_cy;
    cv::Mat ray = pw - origin;
    const float dist3D = cv::norm(ray);
    d[0] = u;
    d[1] = v;
    d[2] = u - Frame::m_flt_baseLineTimesFx / z;
    d[3] = (float)cand[j]->predictScaleLevel(dist3D, kf);
    d[4] = (z >= 0.0 && kf->isInImage(u, v) && dist3D >= cand[j]->getMinDistanceInvariance() &&
            dist3D <= cand[j]->getMaxDistanceInvariance() && ray.dot(cand[j]->getNormal()) >= 0.5 * dist3D)
               ? 1.f : 0.f;
  }
  OrbMatcher matcher(0.6f, true);
  int ret;
  if (mode == 0) {
    ret = matcher.fuseByProjection(kf, points, thd);
  } else {
    std::vector<std::shared_ptr<MapPoint>> replacement(nmp);
    ret = matcher.fuseBySim3(kf, sim, points, thd, replacement);
  }
  auto name = [&](const std::shared_ptr<MapPoint>& q) { return q ? id[q.get()] : -1; };
  for (int i = 0; i < n; i++) {
    slot_out[i] = name(kf->getMapPoint(i));
    slot_repl[i] = slot[i] ? name(slot[i]->getReplacement()) : -1;
    slot_obs_out[i] = slot[i] ? slot[i]->getObservationsNum() : 0;
  }
  for (int j = 0; j < nmp; j++) {
    mp_repl[j] = cand[j] ? name(cand[j]->getReplacement()) : -1;
    mp_obs_out[j] = cand[j] ? cand[j]->getObservationsNum() : 0;
    mp_bad_out[j] = cand[j] ? (cand[j]->isBad() ? 1 : 0) : 0;
  }
  return ret;
}

/* M5  OrbMatcher::searchByBowInKeyFrameAndFrame (src/orbMatcher.cpp:303-379) on a KeyFrame built by the reference's own
 * constructor from side 1 (key points, descriptors, FeatureVector in CSR form, a MapPoint where valid1[i]) and a Frame
 * from side 2.  match_out[i2] = key point of the key frame whose map point the call gave to key point i2, or -1. */
int ref_search_by_bow_kf_frame(const RefKeyPoint* kps1, const uint8_t* desc1, const uint8_t* valid1, int n1,
                               const int32_t* node_ids1, const int32_t* node_start1, const int32_t* items1, int n_nodes1,
                               const RefKeyPoint* kps2, const uint8_t* desc2, int n2, const int32_t* node_ids2,
                               const int32_t* node_start2, const int32_t* items2, int n_nodes2, float ratio,
                               int check_orientation, int32_t* match_out) {
  using namespace YDORBSLAM;
  auto fill = [](Frame& f, const RefKeyPoint* kps, const uint8_t* desc, int n, const int32_t* ids, const int32_t* start,
                 const int32_t* items, int nn) {
    f.m_v_keyPoints.resize(n);
    for (int i = 0; i < n; i++) {
      cv::KeyPoint k;
      k.pt.x = kps[i].x; k.pt.y = kps[i].y; k.size = kps[i].size; k.angle = kps[i].angle;
      k.response = kps[i].response; k.octave = kps[i].octave; k.class_id = kps[i].class_id;
      f.m_v_keyPoints[i] = k;
    }
    f.m_int_keyPointsNum = n;
    f.m_cvMat_descriptors.create(std::max(n, 1), 32, CV_8UC1);
    for (int i = 0; i < n; i++) memcpy(f.m_cvMat_descriptors.ptr(i), desc + (size_t)32 * i, 32);
    for (int k = 0; k < nn; k++)
      for (int t = start[k]; t < start[k + 1]; t++) f.m_bow_keyPointsVec.addFeature((DBoW3::NodeId)ids[k], (unsigned)items[t]);
    f.m_v_sptrMapPoints.assign(n, std::shared_ptr<MapPoint>());
  };
  Frame side1, side2, src;
  fill(side1, kps1, desc1, n1, node_ids1, node_start1, items1, n_nodes1);
  fill(side2, kps2, desc2, n2, node_ids2, node_start2, items2, n_nodes2);
  src.m_v_keyPoints.assign(std::max(n1, 1), cv::KeyPoint());
  src.m_v_scaleFactors.assign(8, 1.0f);
  src.m_cvMat_descriptors = side1.m_cvMat_descriptors.clone();
  src.setCameraPoseByTransform_c2w(cv::Mat::eye(4, 4, CV_32F));
  cv::Mat pos = (cv::Mat_<float>(3, 1) << 0.f, 0.f, 1.f);
  std::map<const MapPoint*, int> index;
  for (int i = 0; i < n1; i++)
    if (valid1[i]) {
      side1.m_v_sptrMapPoints[i] = std::make_shared<ProbePoint>(pos, src, i);
      index[side1.m_v_sptrMapPoints[i].get()] = i;
    }
  std::shared_ptr<KeyFrame> kf = std::make_shared<KeyFrame>(side1, std::shared_ptr<Map>(), std::shared_ptr<KeyFrameDatabase>());
  OrbMatcher matcher(ratio, check_orientation != 0);
  std::vector<std::shared_ptr<MapPoint>> matched;
  const int ret = matcher.searchByBowInKeyFrameAndFrame(kf, side2, matched);
  for (int i = 0; i < n2; i++) match_out[i] = matched[i] ? index[matched[i].get()] : -1;
  return ret;
}

/* M3  OrbMatcher::searchByProjectionInKeyFrameAndCurrentFrame (src/orbMatcher.cpp:156-239).  `p` = current frame at the
 * identity pose; the key frame holds n_kf key points (angles) with one MapPoint each at mp_pos[i] (NaN x: none), built by
 * the reference's constructor, with the given invariance distances / bad flag / membership in _setFoundMapPoints.
 * dbg (may be null) receives per map point {projX, projY, predicted level, passes the gates of :187-189} evaluated with
 * the reference's own methods and the same float expressions as :171-183 -- the inputs a caller of the C ABI supplies. */
int ref_search_keyframe_current(void* p, const RefKeyPoint* kf_kps, int n_kf, const float* mp_pos, const uint8_t* mp_desc,
                                const float* min_dist, const float* max_dist, const uint8_t* is_bad, const uint8_t* in_found,
                                const float* cam, const uint8_t* kp_occupied, float thd, int orb_dist, float ratio,
                                int check_orientation, int32_t* match_out, float* dbg) {
  using namespace YDORBSLAM;
  Frame* cur = static_cast<Frame*>(p);
  const int n = (int)cur->m_v_keyPoints.size();
  Frame::m_flt_fx = cam[0]; Frame::m_flt_fy = cam[1]; Frame::m_flt_cx = cam[2]; Frame::m_flt_cy = cam[3];
  Frame::m_flt_baseLineTimesFx = cam[4];
  Frame::m_flt_baseLine = cam[5];
  cur->setCameraPoseByTransform_c2w(cv::Mat::eye(4, 4, CV_32F));
  Frame src;
  src.m_v_keyPoints.assign(n_kf + 1, cv::KeyPoint());
  src.m_v_scaleFactors = cur->m_v_scaleFactors;
  src.m_cvMat_descriptors.create(n_kf + 1, 32, CV_8UC1);
  for (int i = 0; i < n_kf; i++) memcpy(src.m_cvMat_descriptors.ptr(i), mp_desc + (size_t)32 * i, 32);
  memset(src.m_cvMat_descriptors.ptr(n_kf), 0, 32);
  src.setCameraPoseByTransform_c2w(cv::Mat::eye(4, 4, CV_32F));
  Frame kff;
  kff.m_v_keyPoints.resize(n_kf);
  for (int i = 0; i < n_kf; i++) {
    cv::KeyPoint k;
    k.pt.x = kf_kps[i].x; k.pt.y = kf_kps[i].y; k.size = kf_kps[i].size; k.angle = kf_kps[i].angle;
    k.response = kf_kps[i].response; k.octave = kf_kps[i].octave; k.class_id = kf_kps[i].class_id;
    kff.m_v_keyPoints[i] = k;
  }
  kff.m_int_keyPointsNum = n_kf;
  kff.m_v_sptrMapPoints.assign(n_kf, std::shared_ptr<MapPoint>());
  std::map<const MapPoint*, int> index;
  std::set<std::shared_ptr<MapPoint>> found;
  for (int i = 0; i < n_kf; i++) {
    if (mp_pos[3 * i] != mp_pos[3 * i]) continue;
    cv::Mat pos = (cv::Mat_<float>(3, 1) << mp_pos[3 * i], mp_pos[3 * i + 1], mp_pos[3 * i + 2]);
    std::shared_ptr<ProbePoint> mp = std::make_shared<ProbePoint>(pos, src, i);
    mp->distances(min_dist[i], max_dist[i]);
    mp->bad(is_bad && is_bad[i]);
    kff.m_v_sptrMapPoints[i] = mp;
    index[mp.get()] = i;
    if (in_found && in_found[i]) found.insert(mp);
  }
  std::shared_ptr<KeyFrame> kf = std::make_shared<KeyFrame>(kff, std::shared_ptr<Map>(), std::shared_ptr<KeyFrameDatabase>());
  if (dbg) {
    const cv::Mat R = cur->m_cvMat_T_c2w.rowRange(0, 3).colRange(0, 3), t = cur->m_cvMat_T_c2w.rowRange(0, 3).col(3);
    const cv::Mat origin = -R * t;
    for (int i = 0; i < n_kf; i++) {
      float* d = dbg + 4 * i;
      d[0] = d[1] = d[2] = d[3] = 0.f;
      std::shared_ptr<MapPoint> mp = kff.m_v_sptrMapPoints[i];
      if (!mp || mp->isBad() || found.count(mp)) continue;
      cv::Mat pw = mp->getPosInWorld();
      cv::Mat pc = R * pw + t;
      const float x = pc.at<float>(0), y = pc.at<float>(1), z = pc.at<float>(2);
      const float u = Frame::m_flt_fx * x / z + Frame::m_flt_cx;
      const float v = Frame::m_flt_fy * y / z + Frame::m_flt_cy;
      const float dist3D = cv::norm(pw - origin);
      const float maxD = mp->getMaxDistanceInvariance(), minD = mp->getMinDistanceInvariance();
      d[0] = u;
      d[1] = v;
      d[2] = (float)mp->predictScaleLevel(dist3D, *cur);
      d[3] = (z >= 0.0 && cur->isInImage(u, v) && dist3D >= minD && dist3D <= maxD) ? 1.f : 0.f;
    }
  }
  cv::Mat pos0 = (cv::Mat_<float>(3, 1) << 0.f, 0.f, 1.f);
  std::shared_ptr<ProbePoint> held = std::make_shared<ProbePoint>(pos0, src, n_kf);
  for (int i = 0; i < n; i++)
    cur->m_v_sptrMapPoints[i] = (kp_occupied && kp_occupied[i]) ? std::static_pointer_cast<MapPoint>(held) : std::shared_ptr<MapPoint>();
  OrbMatcher matcher(ratio, check_orientation != 0);
  const int ret = matcher.searchByProjectionInKeyFrameAndCurrentFrame(*cur, kf, found, thd, orb_dist);
  for (int i = 0; i < n; i++) {
    const MapPoint* q = cur->m_v_sptrMapPoints[i].get();
    match_out[i] = (q && q != held.get()) ? index[q] : -1;
  }
  return ret;
}

/* M5b OrbMatcher::searchByBowInTwoKeyFrames (src/orbMatcher.cpp:380-462): both sides are KeyFrames with a MapPoint where
 * valid[i].  match_out[i1] = key point of key frame 2 whose map point the call gave to key point i1, or -1. */
int ref_search_by_bow_two_keyframes(const RefKeyPoint* kps1, const uint8_t* desc1, const uint8_t* valid1, int n1,
                                    const int32_t* ids1, const int32_t* start1, const int32_t* items1, int nn1,
                                    const RefKeyPoint* kps2, const uint8_t* desc2, const uint8_t* valid2, int n2,
                                    const int32_t* ids2, const int32_t* start2, const int32_t* items2, int nn2, float ratio,
                                    int check_orientation, int32_t* match_out) {
  using namespace YDORBSLAM;
  Frame src;
  src.m_v_keyPoints.assign(std::max(n2, 1), cv::KeyPoint());
  src.m_v_scaleFactors.assign(8, 1.0f);
  src.m_cvMat_descriptors = cv::Mat::zeros(std::max(n2, 1), 32, CV_8UC1);
  src.setCameraPoseByTransform_c2w(cv::Mat::eye(4, 4, CV_32F));
  cv::Mat pos = (cv::Mat_<float>(3, 1) << 0.f, 0.f, 1.f);
  std::map<const MapPoint*, int> index2;
  auto fill = [&](Frame& f, const RefKeyPoint* kps, const uint8_t* desc, const uint8_t* valid, int n, const int32_t* ids,
                  const int32_t* start, const int32_t* items, int nn, bool second) {
    f.m_v_keyPoints.resize(n);
    for (int i = 0; i < n; i++) {
      cv::KeyPoint k;
      k.pt.x = kps[i].x; k.pt.y = kps[i].y; k.size = kps[i].size; k.angle = kps[i].angle;
      k.response = kps[i].response; k.octave = kps[i].octave; k.class_id = kps[i].class_id;
      f.m_v_keyPoints[i] = k;
    }
    f.m_int_keyPointsNum = n;
    f.m_cvMat_descriptors.create(std::max(n, 1), 32, CV_8UC1);
    for (int i = 0; i < n; i++) memcpy(f.m_cvMat_descriptors.ptr(i), desc + (size_t)32 * i, 32);
    for (int k = 0; k < nn; k++)
      for (int t = start[k]; t < start[k + 1]; t++) f.m_bow_keyPointsVec.addFeature((DBoW3::NodeId)ids[k], (unsigned)items[t]);
    f.m_v_sptrMapPoints.assign(n, std::shared_ptr<MapPoint>());
    for (int i = 0; i < n; i++)
      if (valid[i]) {
        f.m_v_sptrMapPoints[i] = std::make_shared<ProbePoint>(pos, src, 0);
        if (second) index2[f.m_v_sptrMapPoints[i].get()] = i;
      }
  };
  Frame a, b;
  fill(a, kps1, desc1, valid1, n1, ids1, start1, items1, nn1, false);
  fill(b, kps2, desc2, valid2, n2, ids2, start2, items2, nn2, true);
  std::shared_ptr<KeyFrame> kf1 = std::make_shared<KeyFrame>(a, std::shared_ptr<Map>(), std::shared_ptr<KeyFrameDatabase>());
  std::shared_ptr<KeyFrame> kf2 = std::make_shared<KeyFrame>(b, std::shared_ptr<Map>(), std::shared_ptr<KeyFrameDatabase>());
  OrbMatcher matcher(ratio, check_orientation != 0);
  std::vector<std::shared_ptr<MapPoint>> matched;
  const int ret = matcher.searchByBowInTwoKeyFrames(kf1, kf2, matched);
  for (int i = 0; i < n1; i++) match_out[i] = matched[i] ? index2[matched[i].get()] : -1;
  return ret;
}

/* M6  OrbMatcher::searchForTriangulation (src/orbMatcher.cpp:463-565, epipolar gate :808-819) on two KeyFrames built by the
 * reference's constructor: side i = key points, descriptors, rightX, FeatureVector (CSR), a MapPoint where has_mp[i].
 * Key frame 2 sits at the identity pose, key frame 1 at [I | -origin1], so that the reference's own cv::Mat arithmetic
 * (:465-470) puts the epipole where the caller wants it; epipole_out receives the value it computed.
 * pairs_out: [2 * n] (first, second) indices; returns the function's return value. */
int ref_search_for_triangulation(const RefKeyPoint* kps1, const uint8_t* desc1, const float* rx1, const uint8_t* has_mp1, int n1,
                                 const int32_t* ids1, const int32_t* start1, const int32_t* items1, int nn1,
                                 const RefKeyPoint* kps2, const uint8_t* desc2, const float* rx2, const uint8_t* has_mp2, int n2,
                                 const int32_t* ids2, const int32_t* start2, const int32_t* items2, int nn2, const float* F,
                                 const float* origin1, const float* cam, const float* sf, const float* sf2, int n_levels,
                                 int stereo_only, float ratio, int check_orientation, int32_t* pairs_out, float* epipole_out) {
  using namespace YDORBSLAM;
  Frame::m_flt_fx = cam[0]; Frame::m_flt_fy = cam[1]; Frame::m_flt_cx = cam[2]; Frame::m_flt_cy = cam[3];
  Frame::m_flt_baseLineTimesFx = cam[4];
  Frame::m_flt_baseLine = cam[5];
  Frame src;
  src.m_v_keyPoints.assign(1, cv::KeyPoint());
  src.m_v_scaleFactors.assign(sf, sf + n_levels);
  src.m_cvMat_descriptors = cv::Mat::zeros(1, 32, CV_8UC1);
  src.setCameraPoseByTransform_c2w(cv::Mat::eye(4, 4, CV_32F));
  cv::Mat pos = (cv::Mat_<float>(3, 1) << 0.f, 0.f, 1.f);
  std::shared_ptr<MapPoint> some = std::make_shared<ProbePoint>(pos, src, 0);
  auto fill = [&](Frame& f, const RefKeyPoint* kps, const uint8_t* desc, const float* rx, const uint8_t* has, int n,
                  const int32_t* ids, const int32_t* start, const int32_t* items, int nn) {
    f.m_v_keyPoints.resize(n);
    for (int i = 0; i < n; i++) {
      cv::KeyPoint k;
      k.pt.x = kps[i].x; k.pt.y = kps[i].y; k.size = kps[i].size; k.angle = kps[i].angle;
      k.response = kps[i].response; k.octave = kps[i].octave; k.class_id = kps[i].class_id;
      f.m_v_keyPoints[i] = k;
    }
    f.m_int_keyPointsNum = n;
    f.m_cvMat_descriptors.create(std::max(n, 1), 32, CV_8UC1);
    for (int i = 0; i < n; i++) memcpy(f.m_cvMat_descriptors.ptr(i), desc + (size_t)32 * i, 32);
    f.m_v_rightXcords.assign(rx, rx + n);
    f.m_v_scaleFactors.assign(sf, sf + n_levels);
    f.m_v_scaleFactorSquares.assign(sf2, sf2 + n_levels);
    f.m_int_scaleLevelsNum = n_levels;
    for (int k = 0; k < nn; k++)
      for (int t = start[k]; t < start[k + 1]; t++) f.m_bow_keyPointsVec.addFeature((DBoW3::NodeId)ids[k], (unsigned)items[t]);
    f.m_v_sptrMapPoints.assign(n, std::shared_ptr<MapPoint>());
    for (int i = 0; i < n; i++)
      if (has[i]) f.m_v_sptrMapPoints[i] = some;
  };
  Frame a, b;
  fill(a, kps1, desc1, rx1, has_mp1, n1, ids1, start1, items1, nn1);
  fill(b, kps2, desc2, rx2, has_mp2, n2, ids2, start2, items2, nn2);
  cv::Mat T1 = cv::Mat::eye(4, 4, CV_32F);
  for (int k = 0; k < 3; k++) T1.at<float>(k, 3) = -origin1[k];
  a.setCameraPoseByTransform_c2w(T1);
  b.setCameraPoseByTransform_c2w(cv::Mat::eye(4, 4, CV_32F));
  std::shared_ptr<KeyFrame> kf1 = std::make_shared<KeyFrame>(a, std::shared_ptr<Map>(), std::shared_ptr<KeyFrameDatabase>());
  std::shared_ptr<KeyFrame> kf2 = std::make_shared<KeyFrame>(b, std::shared_ptr<Map>(), std::shared_ptr<KeyFrameDatabase>());
  {
    cv::Mat o1 = kf1->getCameraOriginInWorld(), R2 = kf2->getRotation_c2w(), t2 = kf2->getTranslation_c2w();
    cv::Mat o = R2 * o1 + t2;
    epipole_out[0] = Frame::m_flt_fx * o.at<float>(0) / o.at<float>(2) + Frame::m_flt_cx;
    epipole_out[1] = Frame::m_flt_fy * o.at<float>(1) / o.at<float>(2) + Frame::m_flt_cy;
  }
  cv::Mat Fm(3, 3, CV_32F);
  for (int i = 0; i < 9; i++) Fm.at<float>(i / 3, i % 3) = F[i];
  OrbMatcher matcher(ratio, check_orientation != 0);
  std::vector<std::pair<int, int>> pairs;
  const int ret = matcher.searchForTriangulation(kf1, kf2, Fm, pairs, stereo_only != 0);
  for (size_t i = 0; i < pairs.size(); i++) {
    pairs_out[2 * i] = pairs[i].first;
    pairs_out[2 * i + 1] = pairs[i].second;
  }
  return ret;
}

/* N2  Frame::isInCameraFrustum + MapPoint::predictScaleLevel (src/frame.cpp:295-326, src/mapPoint.cpp:265-278) for n map
 * points seen from a Frame with pose [R | t].  cam = {fx, fy, cx, cy, bf, baseline}, bounds = {minX, minY, maxX, maxY}.
 * out[i] = the track fields the reference wrote into the MapPoint (active = its return value). */
void ref_frustum(const float* R, const float* t, const float* cam, const float* bounds, float scale_factor, int n_levels,
                 const RefMapPoint3d* mps, int n, float view_cos_limit, RefMapPointProj* out) {
  using namespace YDORBSLAM;
  Frame::m_flt_fx = cam[0]; Frame::m_flt_fy = cam[1]; Frame::m_flt_cx = cam[2]; Frame::m_flt_cy = cam[3];
  Frame::m_flt_baseLineTimesFx = cam[4];
  Frame::m_flt_baseLine = cam[5];
  Frame::m_flt_minX = bounds[0]; Frame::m_flt_minY = bounds[1]; Frame::m_flt_maxX = bounds[2]; Frame::m_flt_maxY = bounds[3];
  Frame f, src;
  f.m_int_scaleLevelsNum = n_levels;
  f.m_flt_scaleFactor = scale_factor;
  f.m_flt_logScaleFactor = log(f.m_flt_scaleFactor);  // src/frame.cpp:78
  cv::Mat T = cv::Mat::eye(4, 4, CV_32F);
  for (int i = 0; i < 3; i++) {
    for (int j = 0; j < 3; j++) T.at<float>(i, j) = R[3 * i + j];
    T.at<float>(i, 3) = t[i];
  }
  f.setCameraPoseByTransform_c2w(T);
  src.m_v_keyPoints.assign(1, cv::KeyPoint());
  src.m_v_scaleFactors.assign(n_levels, 1.0f);
  src.m_cvMat_descriptors = cv::Mat::zeros(1, 32, CV_8UC1);
  src.setCameraPoseByTransform_c2w(cv::Mat::eye(4, 4, CV_32F));
  for (int i = 0; i < n; i++) {
    cv::Mat pos = (cv::Mat_<float>(3, 1) << mps[i].pos[0], mps[i].pos[1], mps[i].pos[2]);
    std::shared_ptr<ProbePoint> mp = std::make_shared<ProbePoint>(pos, src, 0);
    mp->distances(mps[i].min_distance, mps[i].max_distance);
    mp->normal(mps[i].normal[0], mps[i].normal[1], mps[i].normal[2]);
    const bool in = f.isInCameraFrustum(mp, view_cos_limit);
    out[i].proj_x = mp->m_flt_trackProjX;
    out[i].proj_y = mp->m_flt_trackProjY;
    out[i].proj_right_x = mp->m_flt_trackProjRightX;
    out[i].view_cos = mp->m_flt_trackViewCos;
    out[i].level = mp->m_int_trackScaleLevel;
    out[i].active = (in && mp->m_b_isTrackInView) ? 1 : 0;
  }
}

/* N3  MapPoint::computeDistinctiveDescriptors (src/mapPoint.cpp:169-218).  Point p has the observations
 * desc[start[p] .. start[p+1]); observation k of every point lives in key frame k of a pool whose members are ordered by
 * address, so that the std::map<shared_ptr<KeyFrame>, int> the reference iterates (:181) visits them in the given
 * order.  out receives the 32 bytes the reference left in the MapPoint (unchanged = all zero when it has no observations). */
void ref_distinctive_descriptors(const uint8_t* desc, const int32_t* start, int n_points, uint8_t* out) {
  using namespace YDORBSLAM;
  int kmax = 0;
  for (int p = 0; p < n_points; p++) kmax = std::max(kmax, start[p + 1] - start[p]);
  Frame src;
  src.m_v_keyPoints.assign(1, cv::KeyPoint());
  src.m_v_scaleFactors.assign(8, 1.0f);
  src.m_cvMat_descriptors = cv::Mat::zeros(1, 32, CV_8UC1);
  src.setCameraPoseByTransform_c2w(cv::Mat::eye(4, 4, CV_32F));
  std::vector<std::shared_ptr<KeyFrame>> pool(kmax);
  {
    Frame proto;
    proto.m_v_rightXcords.assign(std::max(n_points, 1), -1.0f);
    proto.m_cvMat_descriptors = cv::Mat::zeros(std::max(n_points, 1), 32, CV_8UC1);
    for (int k = 0; k < kmax; k++) pool[k] = std::make_shared<KeyFrame>(proto, std::shared_ptr<Map>(), std::shared_ptr<KeyFrameDatabase>());
  }
  std::sort(pool.begin(), pool.end(), [](const std::shared_ptr<KeyFrame>& a, const std::shared_ptr<KeyFrame>& b) { return a.get() < b.get(); });
  for (int p = 0; p < n_points; p++)
    for (int k = 0; k < start[p + 1] - start[p]; k++) memcpy(pool[k]->m_cvMat_descriptors.ptr(p), desc + (size_t)32 * (start[p] + k), 32);
  cv::Mat pos = (cv::Mat_<float>(3, 1) << 0.f, 0.f, 1.f);
  for (int p = 0; p < n_points; p++) {
    std::shared_ptr<ProbePoint> mp = std::make_shared<ProbePoint>(pos, src, 0);
    for (int k = 0; k < start[p + 1] - start[p]; k++) mp->addObservation(pool[k], p);
    mp->computeDistinctiveDescriptors();
    const cv::Mat d = mp->getDescriptor();
    memcpy(out + (size_t)32 * p, d.ptr(0), 32);
  }
}

/* M0 */
int ref_descriptor_distance(const uint8_t* a, const uint8_t* b) {
  cv::Mat ma(1, 32, CV_8UC1, const_cast<uint8_t*>(a)), mb(1, 32, CV_8UC1, const_cast<uint8_t*>(b));
  return YDORBSLAM::OrbMatcher::computeDescriptorsDistance(ma, mb);
}

/* E1 x 2 + S1 + G1: the reference's stereo Frame constructor (src/frame.cpp:57-110) on a rectified pair, zero distortion.
 * Fresh extractors per call; Frame::m_b_isComputeInit is re-armed so the image bounds follow THIS image. */
void* ref_frame_stereo_create(int n_features, float scale_factor, int n_levels, int ini_th, int min_th, const uint8_t* left,
                              const uint8_t* right, int w, int h, int stride, float fx, float fy, float cx, float cy, float bf,
                              float depth_thd, float baseline_before) {
  using namespace YDORBSLAM;
  cv::Mat L(h, w, CV_8UC1, const_cast<uint8_t*>(left), (size_t)stride), R(h, w, CV_8UC1, const_cast<uint8_t*>(right), (size_t)stride);
  cv::Mat K = cv::Mat::eye(3, 3, CV_32F);
  K.at<float>(0, 0) = fx;
  K.at<float>(1, 1) = fy;
  K.at<float>(0, 2) = cx;
  K.at<float>(1, 2) = cy;
  cv::Mat dist = cv::Mat::zeros(4, 1, CV_32F);
  std::shared_ptr<OrbExtractor> el = std::make_shared<OrbExtractor>(n_features, scale_factor, n_levels, ini_th, min_th);
  std::shared_ptr<OrbExtractor> er = std::make_shared<OrbExtractor>(n_features, scale_factor, n_levels, ini_th, min_th);
  Frame::m_b_isComputeInit = true;
  /* the static base line is assigned AFTER computeStereoMatches (src/frame.cpp:101 vs :108): the first frame of a run
   * searches with maxD = bf / 0 = +inf, later frames with bf / (bf / fx).  The caller says which frame this is. */
  Frame::m_flt_baseLine = baseline_before;
  return new Frame(L, R, 0.0, K, dist, dist, bf, depth_thd, el, er, std::shared_ptr<DBoW3::Vocabulary>());
}

/* G1 alone: a Frame holding the caller's key points, grid built by the reference's assignKeyPointsToGrid */
void* ref_frame_from_keypoints(const RefKeyPoint* kps, int n, float min_x, float min_y, float max_x, float max_y) {
  using namespace YDORBSLAM;
  GridFrame* f = new GridFrame();
  f->m_v_keyPoints.resize(n);
  for (int i = 0; i < n; i++) {
    cv::KeyPoint k;
    k.pt.x = kps[i].x;
    k.pt.y = kps[i].y;
    k.size = kps[i].size;
    k.angle = kps[i].angle;
    k.response = kps[i].response;
    k.octave = kps[i].octave;
    k.class_id = kps[i].class_id;
    f->m_v_keyPoints[i] = k;
  }
  f->m_int_keyPointsNum = n;
  Frame::m_flt_minX = min_x;
  Frame::m_flt_minY = min_y;
  Frame::m_flt_maxX = max_x;
  Frame::m_flt_maxY = max_y;
  Frame::m_flt_gridCellWidthInv = static_cast<float>(Frame::m_int_gridColsNum) / (max_x - min_x);   // src/frame.cpp:99
  Frame::m_flt_gridCellHeightInv = static_cast<float>(Frame::m_int_gridRowsNum) / (max_y - min_y);  // src/frame.cpp:100
  f->assign();
  return static_cast<Frame*>(f);
}

void ref_frame_destroy(void* p) { delete static_cast<YDORBSLAM::Frame*>(p); }

/* sizes: n[0] = left key points, n[1] = right key points */
void ref_frame_sizes(void* p, int32_t* n) {
  const YDORBSLAM::Frame* f = static_cast<YDORBSLAM::Frame*>(p);
  n[0] = (int32_t)f->m_v_keyPoints.size();
  n[1] = (int32_t)f->m_v_rightKeyPoints.size();
}

/* bounds[4] = minX, minY, maxX, maxY as computeImageBounds left them */
void ref_frame_get(void* p, RefKeyPoint* kpl, uint8_t* dl, RefKeyPoint* kpr, uint8_t* dr, float* right_x, float* depth,
                   float* bounds) {
  const YDORBSLAM::Frame* f = static_cast<YDORBSLAM::Frame*>(p);
  put(f->m_v_keyPoints, kpl);
  put(f->m_v_rightKeyPoints, kpr);
  put(f->m_cvMat_descriptors, dl);
  put(f->m_cvMat_rightDescriptors, dr);
  for (size_t i = 0; i < f->m_v_rightXcords.size(); i++) {
    if (right_x) right_x[i] = f->m_v_rightXcords[i];
    if (depth) depth[i] = f->m_v_depth[i];
  }
  if (bounds) {
    bounds[0] = YDORBSLAM::Frame::m_flt_minX;
    bounds[1] = YDORBSLAM::Frame::m_flt_minY;
    bounds[2] = YDORBSLAM::Frame::m_flt_maxX;
    bounds[3] = YDORBSLAM::Frame::m_flt_maxY;
  }
}

/* G2 */
int ref_frame_keypoints_in_area(void* p, float x, float y, float r, int min_level, int max_level, int32_t* out, int cap) {
  const YDORBSLAM::Frame* f = static_cast<YDORBSLAM::Frame*>(p);
  const std::vector<int> v = f->getKeyPointsInArea(x, y, r, min_level, max_level);
  for (size_t i = 0; i < v.size() && (int)i < cap; i++) out[i] = v[i];
  return (int)v.size();
}

}  // extern "C"
