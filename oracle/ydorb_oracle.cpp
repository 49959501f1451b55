/*
 * ydorb_oracle.cpp -- CPU oracle (TEST INFRASTRUCTURE, see ydorb_oracle.h).
 *
 * Every function cites the reference lines it restates (paths relative to the
 * YDORBSLAM checkout).  OpenCV primitives are restated from their published
 * algorithms (OpenCV 4.x imgproc/features2d) and pinned against cv2 4.13.0.
 */
#include "ydorb_oracle.h"

#include <algorithm>
#include <cassert>
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <list>
#include <thread>
#include <utility>
#include <vector>

namespace {

const int kPad = 19;        // m_int_maxPadSize   src/orbExtractor.hpp:72
const int kHalfPatch = 15;  // m_int_halfPatchSize src/orbExtractor.hpp:71
const int kPatch = 31;      // m_int_patchSize     src/orbExtractor.hpp:70
const int kBorder = kPad - 3;  // src/orbExtractor.cpp:550

const int8_t kPattern[1024] = {
#include "../ydorbslam_b200/csrc/brief_pattern.inc"
};

inline int cv_round(float v) { return (int)lrintf(v); }  // cvRound(float): round-half-even

}  // namespace

extern "C" {

int orc_cvround(float v) { return cv_round(v); }

/* ------------------------------------------------------------------------------------------------
 * E0  OrbExtractor::OrbExtractor  src/orbExtractor.cpp:315-354
 * ---------------------------------------------------------------------------------------------- */
int orc_params_init(OrcParams* p, int n_features, float scale_factor, int n_levels, int ini_th, int min_th) {
  if (n_levels < 1 || n_levels > ORC_MAX_LEVELS) return -1;
  memset(p, 0, sizeof(*p));
  p->n_features = n_features;
  p->scale_factor = scale_factor;
  p->n_levels = n_levels;
  p->ini_th = ini_th;
  p->min_th = min_th;
  const float sf = scale_factor;
  // :325-326  (float promoted to double inside the expression)
  int per_level = (int)round(n_features * (1 - 1.0 / sf) / (1.0 - pow(1.0 / sf, n_levels)));
  int sum = 0;
  for (int i = 0; i < n_levels; i++) {
    p->sf[i] = (float)pow((double)sf, (double)i);            // :329
    p->sf2[i] = (float)pow((double)sf, (double)(2 * i));     // :330
    p->inv_sf[i] = (float)pow((double)sf, (double)(-i));     // :331
    p->inv_sf2[i] = (float)pow((double)sf, (double)(-2 * i));  // :332
    if (i < n_levels - 1) {
      p->quota[i] = per_level;  // :334
      sum += per_level;
      per_level = (int)round((double)((float)per_level / sf));  // :336
    } else {
      p->quota[i] = std::max(n_features - sum, 0);  // :338
    }
  }
  // :341-353 -- literal, including resize()+push_back
  std::vector<int> umax;
  umax.resize(kHalfPatch + 1);
  int vmax = (int)floor(kHalfPatch * sqrt(2.0) / 2.0 + 1.0);
  int vmin = (int)ceil(kHalfPatch * sqrt(2.0) / 2.0);
  for (int v = 0; v <= vmax; v++) umax.push_back((int)round(sqrt(pow(kHalfPatch, 2) + pow(v, 2))));
  for (int v = kHalfPatch, i = 0; v >= vmin; v--) {
    while (umax[i] == umax[i + 1]) i++;
    umax[v] = i;
    i++;
  }
  for (int v = 0; v < 16; v++) p->umax[v] = umax[v];
  return 0;
}

/* E2 sizes  src/orbExtractor.cpp:608-609 */
void orc_level_size(const OrcParams* p, int w, int h, int level, int* lw, int* lh) {
  float scale = p->inv_sf[level];
  *lw = cv_round((float)w * scale);
  *lh = cv_round((float)h * scale);
}

/* ------------------------------------------------------------------------------------------------
 * cv::resize INTER_LINEAR 8UC1  (OpenCV imgproc/resize.cpp: resizeGeneric_ + HResizeLinear + VResizeLinear,
 * INTER_RESIZE_COEF_BITS = 11).  Call site src/orbExtractor.cpp:614.
 * ---------------------------------------------------------------------------------------------- */
static void linear_tab(int sn, int dn, std::vector<int>& ofs, std::vector<short>& a0, std::vector<short>& a1) {
  double scale = 1.0 / ((double)dn / (double)sn);
  ofs.resize(dn);
  a0.resize(dn);
  a1.resize(dn);
  for (int d = 0; d < dn; d++) {
    float f = (float)((d + 0.5) * scale - 0.5);
    int s = (int)floorf(f);
    f -= (float)s;
    if (s < 0) { s = 0; f = 0.f; }
    if (s >= sn - 1) { s = sn - 1; f = 0.f; }
    ofs[d] = s;
    a0[d] = (short)cv_round((1.f - f) * 2048.f);
    a1[d] = (short)cv_round(f * 2048.f);
  }
}

void orc_resize_linear(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw, int dh, int dstride) {
  std::vector<int> xo, yo;
  std::vector<short> xa0, xa1, ya0, ya1;
  linear_tab(sw, dw, xo, xa0, xa1);
  linear_tab(sh, dh, yo, ya0, ya1);
  std::vector<int> r0(dw), r1(dw);
  for (int y = 0; y < dh; y++) {
    const uint8_t* s0 = src + (size_t)yo[y] * sstride;
    const uint8_t* s1 = src + (size_t)std::min(yo[y] + 1, sh - 1) * sstride;
    for (int x = 0; x < dw; x++) {
      int xs = xo[x], xs1 = std::min(xs + 1, sw - 1);
      r0[x] = s0[xs] * xa0[x] + s0[xs1] * xa1[x];
      r1[x] = s1[xs] * xa0[x] + s1[xs1] * xa1[x];
    }
    const int b0 = ya0[y], b1 = ya1[y];
    uint8_t* d = dst + (size_t)y * dstride;
    for (int x = 0; x < dw; x++) {
      int v = (((b0 * (r0[x] >> 4)) >> 16) + ((b1 * (r1[x] >> 4)) >> 16) + 2) >> 2;
      d[x] = (uint8_t)v;
    }
  }
}

/* cv::copyMakeBorder(BORDER_REFLECT_101)  call sites src/orbExtractor.cpp:615-618 */
static inline int reflect101(int i, int n) {
  if (n == 1) return 0;
  while (i < 0 || i >= n) {
    if (i < 0) i = -i;
    else i = 2 * (n - 1) - i;
  }
  return i;
}

void orc_fill_border101(uint8_t* buf, int w, int h, int pad, int stride) {
  // centre at (pad,pad); rows first on the centre rows, then whole rows copied
  for (int y = 0; y < h; y++) {
    uint8_t* row = buf + (size_t)(y + pad) * stride;
    for (int x = -pad; x < 0; x++) row[x + pad] = row[reflect101(x, w) + pad];
    for (int x = w; x < w + pad; x++) row[x + pad] = row[reflect101(x, w) + pad];
  }
  for (int y = -pad; y < h + pad; y++) {
    if (y >= 0 && y < h) continue;
    int sy = reflect101(y, h);
    memcpy(buf + (size_t)(y + pad) * stride, buf + (size_t)(sy + pad) * stride, (size_t)(w + 2 * pad));
  }
}

/* ------------------------------------------------------------------------------------------------
 * cv::FAST(img, kps, th, nonmaxSuppression=true), TYPE_9_16 (OpenCV features2d/fast.cpp, fast_score.cpp).
 * Call sites src/orbExtractor.cpp:581,583.  SURVEY A3.
 * ---------------------------------------------------------------------------------------------- */
static const int kCircle[16][2] = {{0, 3},  {1, 3},   {2, 2},   {3, 1},   {3, 0},  {3, -1}, {2, -2}, {1, -3},
                                   {0, -3}, {-1, -3}, {-2, -2}, {-3, -1}, {-3, 0}, {-3, 1}, {-2, 2}, {-1, 3}};

static inline int fast_m(const uint8_t* p, int stride) {
  int d[25];
  const int c = p[0];
  for (int k = 0; k < 16; k++) d[k] = c - p[kCircle[k][1] * stride + kCircle[k][0]];
  for (int k = 16; k < 25; k++) d[k] = d[k - 16];
  int m = -255;
  for (int k = 0; k < 16; k++) {
    int lo = d[k], hi = d[k];
    for (int j = 1; j < 9; j++) {
      lo = std::min(lo, d[k + j]);
      hi = std::max(hi, d[k + j]);
    }
    m = std::max(m, std::max(lo, -hi));
  }
  return m;
}

int orc_fast(const uint8_t* img, int w, int h, int stride, int th, int* xs, int* ys, int* resp, int cap) {
  th = std::min(std::max(th, 0), 255);
  if (w < 7 || h < 7) return 0;
  // score map: 0 = not a corner, else m (corner iff m > th; cornerScore returns m-1)
  std::vector<int> score((size_t)w * h, 0);
  const int o0 = 3 * stride, o4 = 3, o8 = -3 * stride, o12 = -3;
  for (int y = 3; y < h - 3; y++) {
    const uint8_t* row = img + (size_t)y * stride;
    for (int x = 3; x < w - 3; x++) {
      // any 9 contiguous circle pixels contain at least two of the four compass pixels (same pruning idea as
      // OpenCV's pre-test; does not change the result)
      const uint8_t* p = row + x;
      const int c = p[0], lo = c - th, hi = c + th;
      const int a = p[o0], b = p[o4], e = p[o8], f = p[o12];
      const int dark = (a < lo) + (b < lo) + (e < lo) + (f < lo);
      const int bright = (a > hi) + (b > hi) + (e > hi) + (f > hi);
      if (dark < 2 && bright < 2) continue;
      const int m = fast_m(p, stride);
      if (m > th) score[(size_t)y * w + x] = m;
    }
  }
  int n = 0;
  for (int y = 3; y < h - 3; y++)
    for (int x = 3; x < w - 3; x++) {
      const int* sc = &score[(size_t)y * w + x];
      const int s = sc[0];
      if (s == 0) continue;
      // strictly greater than all 8 neighbours, non-corners count as 0 (comparing m is the same as comparing m-1)
      if (s > sc[-1] && s > sc[1] && s > sc[-w - 1] && s > sc[-w] && s > sc[-w + 1] && s > sc[w - 1] && s > sc[w] &&
          s > sc[w + 1]) {
        if (n < cap) {
          xs[n] = x;
          ys[n] = y;
          resp[n] = s - 1;
        }
        n++;
      }
    }
  return n;
}

/* ------------------------------------------------------------------------------------------------
 * cv::GaussianBlur(7x7, sigma=2, BORDER_REFLECT_101) 8UC1: fixed-point 8.8 taps.
 * Call site src/orbExtractor.cpp:386.
 * ---------------------------------------------------------------------------------------------- */
static const int kGauss[7] = {18, 34, 48, 56, 48, 34, 18};

void orc_gauss7(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride) {
  std::vector<uint16_t> hb((size_t)w * h);
  for (int y = 0; y < h; y++) {
    const uint8_t* s = src + (size_t)y * sstride;
    for (int x = 0; x < w; x++) {
      int acc = 0;
      for (int k = -3; k <= 3; k++) acc += kGauss[k + 3] * s[reflect101(x + k, w)];
      hb[(size_t)y * w + x] = (uint16_t)acc;
    }
  }
  for (int y = 0; y < h; y++) {
    uint8_t* d = dst + (size_t)y * dstride;
    for (int x = 0; x < w; x++) {
      uint32_t acc = 0;
      for (int k = -3; k <= 3; k++) acc += (uint32_t)kGauss[k + 3] * hb[(size_t)reflect101(y + k, h) * w + x];
      d[x] = (uint8_t)((acc + 32768u) >> 16);
    }
  }
}

/* cv::fastAtan2 scalar (OpenCV core/mathfuncs_core: atan_f32).  Call site src/orbExtractor.cpp:419.  SURVEY A4. */
float orc_fast_atan2(float y, float x) {
  const float s = (float)(180.0 / 3.14159265358979323846);
  const float p1 = 0.9997878412794807f * s, p3 = -0.3258083974640975f * s, p5 = 0.1555786518463281f * s,
              p7 = -0.04432655554792128f * s;
  volatile float ax = fabsf(x), ay = fabsf(y);  // volatile: forbid contraction/reassociation
  volatile float a, c, c2, t;
  if (ax >= ay) {
    c = ay / (ax + (float)DBL_EPSILON);
    c2 = c * c;
    t = p7 * c2; t = t + p5; t = t * c2; t = t + p3; t = t * c2; t = t + p1;
    a = t * c;
  } else {
    c = ax / (ay + (float)DBL_EPSILON);
    c2 = c * c;
    t = p7 * c2; t = t + p5; t = t * c2; t = t + p3; t = t * c2; t = t + p1;
    t = t * c;
    a = 90.f - t;
  }
  if (x < 0) a = 180.f - a;
  if (y < 0) a = 360.f - a;
  return a;
}

/* ------------------------------------------------------------------------------------------------
 * E2  computePyramid  src/orbExtractor.cpp:605-621 (fresh-extractor semantics)
 * ---------------------------------------------------------------------------------------------- */
int64_t orc_pyramid_bytes(const OrcParams* p, int w, int h, int64_t* off) {
  int64_t tot = 0;
  for (int l = 0; l < p->n_levels; l++) {
    int lw, lh;
    orc_level_size(p, w, h, l, &lw, &lh);
    if (off) off[l] = tot;
    tot += (int64_t)(lw + 2 * kPad) * (lh + 2 * kPad);
  }
  return tot;
}

void orc_pyramid(const OrcParams* p, const uint8_t* img, int w, int h, int stride, uint8_t* out) {
  int64_t off[ORC_MAX_LEVELS];
  orc_pyramid_bytes(p, w, h, off);
  int pw = w, ph = h;
  const uint8_t* prev = nullptr;
  int pstride = 0;
  for (int l = 0; l < p->n_levels; l++) {
    int lw, lh;
    orc_level_size(p, w, h, l, &lw, &lh);
    int st = lw + 2 * kPad;
    uint8_t* buf = out + off[l];
    uint8_t* centre = buf + (size_t)kPad * st + kPad;
    if (l == 0) {
      for (int y = 0; y < h; y++) memcpy(centre + (size_t)y * st, img + (size_t)y * stride, (size_t)w);
    } else {
      orc_resize_linear(prev, pw, ph, pstride, centre, lw, lh, st);  // :614 (level l from level l-1)
    }
    orc_fill_border101(buf, lw, lh, kPad, st);  // :615-618
    prev = centre;
    pstride = st;
    pw = lw;
    ph = lh;
  }
}

/* ------------------------------------------------------------------------------------------------
 * E3 (cells)  computeKeyPointsPyramid  src/orbExtractor.cpp:545-593
 * ---------------------------------------------------------------------------------------------- */
int orc_level_candidates(const uint8_t* lvl, int lw, int lh, int stride, int ini_th, int min_th, OrcCandidate* out,
                         int cap) {
  const float widthHeight = 30.0f;                                  // :548
  const int minBorderX = kBorder, minBorderY = kBorder;             // :550-551
  const int maxBorderX = lw - kBorder, maxBorderY = lh - kBorder;   // :552-553
  const float width = (float)(maxBorderX - minBorderX);             // :556
  const float height = (float)(maxBorderY - minBorderY);
  const int colsNum = (int)(width / widthHeight);                   // :558
  const int rowsNum = (int)(height / widthHeight);
  if (colsNum <= 0 || rowsNum <= 0) return 0;  // reference would divide by zero; loops do not run
  const int cellWidth = (int)ceilf(width / colsNum);                // :560
  const int cellHeight = (int)ceilf(height / rowsNum);
  std::vector<int> xs(4096), ys(4096), rs(4096);
  int n = 0;
  for (int i = 0; i < rowsNum; i++) {
    const float initY = (float)(minBorderY + i * cellHeight);  // :563
    float maxY = initY + cellHeight + 6;
    if (initY >= maxBorderY - 3) continue;  // :565
    if (maxY > maxBorderY) maxY = (float)maxBorderY;
    for (int j = 0; j < colsNum; j++) {
      const float initX = (float)(minBorderX + j * cellWidth);  // :572
      float maxX = initX + cellWidth + 6;
      if (initX >= maxBorderX - 6) continue;  // :574
      if (maxX > maxBorderX) maxX = (float)maxBorderX;
      const int x0 = (int)initX, x1 = (int)maxX, y0 = (int)initY, y1 = (int)maxY;
      const uint8_t* roi = lvl + (size_t)y0 * stride + x0;
      int cw = x1 - x0, ch = y1 - y0;
      if ((size_t)cw * ch > xs.size()) { xs.resize((size_t)cw * ch); ys.resize((size_t)cw * ch); rs.resize((size_t)cw * ch); }
      int k = orc_fast(roi, cw, ch, stride, ini_th, xs.data(), ys.data(), rs.data(), (int)xs.size());  // :581
      if (k == 0) k = orc_fast(roi, cw, ch, stride, min_th, xs.data(), ys.data(), rs.data(), (int)xs.size());  // :583
      for (int t = 0; t < k; t++) {  // :586-590
        if (n < cap) {
          out[n].x = xs[t] + j * cellWidth;
          out[n].y = ys[t] + i * cellHeight;
          out[n].response = (float)rs[t];
        }
        n++;
      }
    }
  }
  return n;
}

}  // extern "C"

/* ------------------------------------------------------------------------------------------------
 * E4  QuadTreeNode::divideNode src/orbExtractor.cpp:4-54, distributeQuadTree :455-544
 * Literal std::list transcription (keeps SURVEY F5/F6/F7 behaviour by construction).
 * ---------------------------------------------------------------------------------------------- */
namespace {

struct P2 { int x = 0, y = 0; };
struct QKp { float x, y, response; int idx; };

struct QNode {
  std::vector<QKp> kps;
  P2 tl, tr, bl, br;
  bool indivisible = false;
  void divide(QNode& n1, QNode& n2, QNode& n3, QNode& n4) const {
    const int cx = (int)ceil(static_cast<float>(tr.x + tl.x) / 2.0);  // :7
    const int cy = (int)ceil(static_cast<float>(bl.y + tl.y) / 2.0);  // :8
    const P2 center{cx, cy};
    n1.tl = tl; n1.tr = P2{cx, tl.y}; n1.bl = P2{tl.x, cy}; n1.br = center;         // :12-15
    n2.tl = n1.tr; n2.tr = tr; n2.bl = center; n2.br = P2{tr.x, cy};                // :18-21
    n3.tl = n1.bl; n3.tr = center; n3.bl = bl; n3.br = P2{cx, tl.y};                // :24-27 (F6: br.y = tl.y)
    n4.tl = center; n4.tr = n2.br; n4.bl = n3.br; n4.br = br;                       // :30-33
    for (const QKp& kp : kps) {  // :37-47
      if (kp.x < cx && kp.y < cy) n1.kps.push_back(kp);
      else if (kp.x < cx && kp.y >= cy) n3.kps.push_back(kp);
      else if (kp.x >= cx && kp.y < cy) n2.kps.push_back(kp);
      else n4.kps.push_back(kp);
    }
    n1.indivisible = n1.kps.size() == 1;  // :50-53
    n2.indivisible = n2.kps.size() == 1;
    n3.indivisible = n3.kps.size() == 1;
    n4.indivisible = n4.kps.size() == 1;
  }
};

}  // namespace

extern "C" int orc_quadtree_list(const OrcCandidate* cand, int n, int rootw, int rooth, int want, int32_t* sel) {
  std::list<QNode> nodes;
  QNode root;  // :457-461
  root.tl = P2{0, 0};
  root.tr = P2{rootw, 0};
  root.bl = P2{0, rooth};
  root.br = P2{rootw, rooth};
  for (int i = 0; i < n; i++) root.kps.push_back(QKp{(float)cand[i].x, (float)cand[i].y, cand[i].response, i});
  nodes.push_back(root);
  for (auto it = nodes.begin(); it != nodes.end();) {  // :466-475
    if (it->kps.size() == 1) { it->indivisible = true; ++it; }
    else if (it->kps.empty()) it = nodes.erase(it);
    else ++it;
  }
  int sizeInLastIter = 0;
  while ((int)nodes.size() > sizeInLastIter && (int)nodes.size() < want) {  // :479
    sizeInLastIter = (int)nodes.size();
    // :483 is always true (expandableNodesNum was just reset to 0); the else branch :507-530 is dead.
    for (auto it = nodes.begin(); it != nodes.end();) {
      if (it->indivisible) { ++it; continue; }
      QNode c[4];
      it->divide(c[0], c[1], c[2], c[3]);
      for (QNode& ch : c) {
        if (!ch.kps.empty()) {
          if (ch.kps.size() == 1) ch.indivisible = true;
          nodes.push_front(ch);  // :496
        }
      }
      it = nodes.erase(it);  // :504
    }
  }
  int m = 0;
  for (const QNode& nd : nodes) {  // :535-540
    std::vector<QKp> v = nd.kps;
    std::sort(v.begin(), v.end(), [](QKp& a, QKp& b) { return a.response > b.response; });
    if (m < want) sel[m] = v.front().idx;  // :541-543 resize(want)
    m++;
  }
  return std::min(m, std::max(want, 0));
}

/* ------------------------------------------------------------------------------------------------
 * E4 restated as a sort over per-point path codes.  This is the formulation the
 * CUDA kernel uses; tests assert it equals orc_quadtree_list.
 * ---------------------------------------------------------------------------------------------- */
namespace {
const int kQtDepth = 16;  // digits per code (2 bits each)

inline uint32_t qt_code(int x, int y, int rootw, int rooth) {
  int tlx = 0, trx = rootw, tly = 0, bly = rooth;
  uint32_t code = 0;
  for (int d = 0; d < kQtDepth; d++) {
    int cx = (trx + tlx + 1) >> 1;  // ceil((a+b)/2) for a+b >= 0
    int cy = (bly + tly + 1) >> 1;
    bool right = x >= cx, low = y >= cy;
    uint32_t q;  // child-1 => 0, child-2 => 1, child-3 => 2, child-4 => 3
    if (!right && !low) { q = 0; trx = cx; bly = cy; }
    else if (right && !low) { q = 1; tlx = cx; bly = cy; }
    else if (!right && low) { q = 2; trx = cx; tly = cy; }
    else { q = 3; int t = tly; tlx = cx; tly = cy; bly = t; }  // F6: child-4 bl.y = parent tl.y
    code = (code << 2) | q;
  }
  return code;
}
inline int lcp_digits(uint32_t a, uint32_t b) {
  uint32_t x = a ^ b;
  if (x == 0) return kQtDepth;
  return __builtin_clz(x) >> 1;
}
}  // namespace

extern "C" int orc_quadtree_sortkey(const OrcCandidate* cand, int n, int rootw, int rooth, int want, int32_t* sel) {
  if (n <= 0) return 0;
  std::vector<std::pair<uint32_t, int>> v(n);
  for (int i = 0; i < n; i++) v[i] = {qt_code(cand[i].x, cand[i].y, rootw, rooth), i};
  std::stable_sort(v.begin(), v.end(), [](const std::pair<uint32_t, int>& a, const std::pair<uint32_t, int>& b) {
    return a.first < b.first;
  });
  // cnt(d) = number of distinct d-digit prefixes
  std::vector<int> lcp(n, 0);  // lcp[i] = common digits of v[i-1], v[i]
  int cnt[kQtDepth + 2];
  for (int d = 0; d <= kQtDepth; d++) cnt[d] = 1;
  for (int i = 1; i < n; i++) {
    lcp[i] = lcp_digits(v[i - 1].first, v[i].first);
    for (int d = lcp[i] + 1; d <= kQtDepth; d++) cnt[d]++;
  }
  int D = 0;
  {
    int last = 0, size = 1;
    while (size > last && size < want && D < kQtDepth) {
      last = size;
      D++;
      size = cnt[D];
    }
  }
  struct Node { uint64_t key; int best; };
  std::vector<Node> nodes;
  const int shift = 2 * (kQtDepth - D);
  for (int i = 0; i < n;) {
    int j = i + 1;
    while (j < n && lcp[j] >= D) j++;
    // node = v[i..j)
    int s;
    if (j - i == 1) {
      int l = 0;
      if (i > 0) l = std::max(l, lcp[i]);
      if (j < n) l = std::max(l, lcp[j]);
      s = std::min(D, (n == 1) ? 0 : l + 1);
    } else {
      s = D;
    }
    // best response; std::sort semantics
    int best;
    {
      std::vector<QKp> t;
      for (int k = i; k < j; k++) t.push_back(QKp{0, 0, cand[v[k].second].response, v[k].second});
      // members of a node keep candidate order (stable partition, :37-47); the full 16-digit code order is finer
      std::sort(t.begin(), t.end(), [](const QKp& a, const QKp& b) { return a.idx < b.idx; });
      std::sort(t.begin(), t.end(), [](QKp& a, QKp& b) { return a.response > b.response; });
      best = t.front().idx;
    }
    // order key: (D-s) ascending, then e ascending, e_k = 3-c_k if (s-k) even else c_k   (c = q-1)
    uint32_t pre = s ? (v[i].first >> (2 * (kQtDepth - s))) : 0;
    uint32_t mask = 0;
    for (int k = s; k >= 1; k -= 2) mask |= 3u << (2 * (s - k));
    uint64_t key = ((uint64_t)(D - s) << 32) | (pre ^ mask);
    nodes.push_back(Node{key, best});
    (void)shift;
    i = j;
  }
  std::sort(nodes.begin(), nodes.end(), [](const Node& a, const Node& b) { return a.key < b.key; });
  int m = std::min((int)nodes.size(), std::max(want, 0));
  for (int i = 0; i < m; i++) sel[i] = nodes[i].best;
  return m;
}

extern "C" {

/* E5  computeOrientation  src/orbExtractor.cpp:400-421 */
float orc_ic_angle(const OrcParams* p, const uint8_t* lvl, int stride, int x, int y) {
  int m01 = 0, m10 = 0;
  const uint8_t* c = lvl + (ptrdiff_t)y * stride + x;
  for (int u = -kHalfPatch; u <= kHalfPatch; u++) m10 += u * c[u];  // :404-406
  for (int v = 1; v <= kHalfPatch; v++) {                          // :407-418
    int vSum = 0;
    int d = p->umax[v];
    for (int u = -d; u <= d; u++) {
      int pos = c[(ptrdiff_t)v * stride + u];
      int neg = c[-(ptrdiff_t)v * stride + u];
      vSum += pos - neg;
      m10 += u * (pos + neg);
    }
    m01 += v * vSum;
  }
  return orc_fast_atan2((float)m01, (float)m10);  // :419
}

/* glibc 2.39 sincosf, x86-64 FMA variant (sysdeps/ieee754/flt-32/s_sincosf.c + sysdeps/x86_64/fpu/sincosf_poly.h,
 * selected by ifunc on every CPU with FMA+AVX2), restated for the device kernel to follow: the published algorithm of
 * the ARM optimized-routines sincosf -- argument reduction and two short polynomials evaluated in double, every
 * `a + b*c` a fused multiply-add, result rounded once to float.  Only the branches an angle in [0, 120) can take.
 * tests/test_oracle_ref.py checks this model against libm's sincosf over the whole input domain of
 * src/orbExtractor.cpp:425 (every float in [0, 2*pi]).  Constants: __sincosf_table of the same glibc. */
__attribute__((target("fma"))) void orc_sincosf_model(float y, float* sinp, float* cosp) {
  static const double kHpiInv = 0x1.45F306DC9C883p+23, kHpi = 0x1.921FB54442D18p0;
  static const double C0 = 0x1p0, C1 = -0x1.ffffffd0c621cp-2, C2 = 0x1.55553e1068f19p-5, C3 = -0x1.6c087e89a359dp-10,
                      C4 = 0x1.99343027bf8c3p-16;
  static const double S1 = -0x1.555545995a603p-3, S2 = 0x1.1107605230bc4p-7, S3 = -0x1.994eb3774cf24p-13;
  uint32_t bits;
  memcpy(&bits, &y, 4);
  const uint32_t top12 = (bits >> 20) & 0x7ff;
  double x = (double)y, xs = x, neg = 1.0;
  int n = 0;
  if (top12 < 0x3f4) {       // |y| < 0.75 (abstop12(y) < abstop12(pi/4))
    if (top12 < 0x398) {     // |y| < 2^-12
      *sinp = y;
      *cosp = 1.0f;
      return;
    }
  } else {                   // reduce_fast, valid below 120.0f
    const double r = x * kHpiInv;
    n = ((int32_t)r + 0x800000) >> 24;
    x = __builtin_fma(-(double)n, kHpi, x);
    xs = x * ((n + 1) & 2 ? -1.0 : 1.0);  // sign[n & 3] = {1, -1, -1, 1}
    if (n & 2) neg = -1.0;                 // __sincosf_table[1]: cosine coefficients negated
  }
  const double x2 = x * x;
  const double x3 = x2 * xs, x4 = x2 * x2;
  const double s1 = __builtin_fma(x2, S3, S2), c2 = __builtin_fma(x2, neg * C4, neg * C3);
  const double c1 = __builtin_fma(x2, neg * C1, neg * C0);
  const double x5 = x2 * x3, x6 = x2 * x4;
  const double s = __builtin_fma(x3, S1, xs), c = __builtin_fma(x4, neg * C2, c1);
  const float sv = (float)__builtin_fma(s1, x5, s), cv = (float)__builtin_fma(c2, x6, c);
  if (n & 1) {
    *sinp = cv;
    *cosp = sv;
  } else {
    *sinp = sv;
    *cosp = cv;
  }
}

/* count of floats in [lo_bits, hi_bits] (step `stride`) where orc_sincosf_model differs from libm sincosf */
int64_t orc_sincosf_model_mismatches(uint32_t lo_bits, uint32_t hi_bits, uint32_t stride, int threads) {
  if (threads < 1) threads = 1;
  std::vector<int64_t> bad((size_t)threads, 0);
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; t++) {
    pool.emplace_back([&, t]() {
      for (uint64_t u = (uint64_t)lo_bits + (uint64_t)t * stride; u <= hi_bits; u += (uint64_t)stride * threads) {
        const uint32_t b = (uint32_t)u;
        float y, s0, c0, s1, c1;
        memcpy(&y, &b, 4);
        sincosf(y, &s0, &c0);
        orc_sincosf_model(y, &s1, &c1);
        if (memcmp(&s0, &s1, 4) || memcmp(&c0, &c1, 4)) bad[t]++;
      }
    });
  }
  for (auto& th : pool) th.join();
  int64_t n = 0;
  for (int64_t b : bad) n += b;
  return n;
}

/* libm sincosf (what src/orbExtractor.cpp:426 calls) for `count` consecutive float bit patterns; out[2i] = sin, [2i+1] = cos */
void orc_sincosf_range(uint32_t first_bits, uint32_t count, float* out, int threads) {
  if (threads < 1) threads = 1;
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; t++) {
    pool.emplace_back([=]() {
      const uint64_t lo = (uint64_t)count * t / threads, hi = (uint64_t)count * (t + 1) / threads;
      for (uint64_t i = lo; i < hi; i++) {
        const uint32_t b = first_bits + (uint32_t)i;
        float y;
        memcpy(&y, &b, 4);
        sincosf(y, &out[2 * i], &out[2 * i + 1]);
      }
    });
  }
  for (auto& th : pool) th.join();
}

/* E7  computeDescriptors  src/orbExtractor.cpp:422-454 */
void orc_descriptor(const uint8_t* img, int stride, int x, int y, float angle_deg, uint8_t* desc32) {
  const float deg2rad = (float)(M_PI / 180.0);  // src/orbExtractor.hpp:57
  volatile float angle = angle_deg * deg2rad;   // :425
  // :426  `(float)cos(angle), (float)sin(angle)` with a float argument: <math.h>/<cmath> of libstdc++ put the float
  // overloads into the global namespace, so the reference calls cosf/sinf (g++ -O3 fuses the pair into one sincosf
  // call -- checked with nm on the compiled reference object, oracle/_ref/orbExtractor.o).  NOT double cos/sin: those
  // differ from cosf/sinf by one float ulp on 0.04 % / 0.09 % of the floats in [0, 2*pi].
  float cosA, sinB;
  sincosf(angle, &sinB, &cosA);
  auto get = [&](int idx) -> int {  // GET_VALUE :427-429
    const int px = kPattern[2 * idx], py = kPattern[2 * idx + 1];
    volatile float a = px * sinB, b = py * cosA, r = a + b;
    volatile float c = px * cosA, d = py * sinB, q = c - d;
    return img[(ptrdiff_t)(y + cv_round(r)) * stride + (x + cv_round(q))];
  };
  for (int j = 0; j < 32; j++) {
    int off = j * 16;
    unsigned val = 0;
    for (int b = 0; b < 8; b++) {
      int t0 = get(off + 2 * b), t1 = get(off + 2 * b + 1);
      val |= (unsigned)(t0 < t1) << b;
    }
    desc32[j] = (uint8_t)val;
  }
}

/* E1  extractAndCompute  src/orbExtractor.cpp:355-399 (+ :545-604 per level) */
int orc_extract(const OrcParams* p, const uint8_t* img, int w, int h, int stride, OrcKeyPoint* kps, uint8_t* desc,
                int cap, int32_t* per_level_n) {
  if (!img || w <= 0 || h <= 0) return 0;  // :357
  int64_t off[ORC_MAX_LEVELS];
  int64_t bytes = orc_pyramid_bytes(p, w, h, off);
  std::vector<uint8_t> pyr((size_t)bytes);
  orc_pyramid(p, img, w, h, stride, pyr.data());
  int total = 0;
  std::vector<OrcCandidate> cand;
  std::vector<int32_t> sel;
  std::vector<uint8_t> blurred;
  for (int l = 0; l < p->n_levels; l++) {
    int lw, lh;
    orc_level_size(p, w, h, l, &lw, &lh);
    const int st = lw + 2 * kPad;
    const uint8_t* lvl = pyr.data() + off[l] + (size_t)kPad * st + kPad;
    int n = orc_level_candidates(lvl, lw, lh, st, p->ini_th, p->min_th, nullptr, 0);
    cand.resize((size_t)std::max(n, 1));
    orc_level_candidates(lvl, lw, lh, st, p->ini_th, p->min_th, cand.data(), n);
    sel.resize((size_t)std::max(p->quota[l], 1));
    int m = orc_quadtree_list(cand.data(), n, (lw - kBorder) - kBorder, (lh - kBorder) - kBorder, p->quota[l],
                              sel.data());  // :594
    if (per_level_n) per_level_n[l] = m;
    if (m == 0) continue;  // :381-383
    blurred.resize((size_t)lw * lh);
    orc_gauss7(lvl, lw, lh, st, blurred.data(), lw);  // :385-386 (clone drops the pad, blur reflects on own edge)
    const int scaledPatch = (int)(kPatch * p->sf[l]);  // :595
    for (int i = 0; i < m; i++) {
      const OrcCandidate& c = cand[sel[i]];
      const int x = c.x + kBorder, y = c.y + kBorder;  // :597-598
      if (total < cap) {
        OrcKeyPoint& k = kps[total];
        k.angle = orc_ic_angle(p, lvl, st, x, y);  // :603
        orc_descriptor(blurred.data(), lw, x, y, k.angle, desc + (size_t)total * 32);  // :388
        k.x = (float)x;
        k.y = (float)y;
        if (l != 0) {  // :391-396
          k.x *= p->sf[l];
          k.y *= p->sf[l];
        }
        k.size = (float)scaledPatch;  // :600
        k.response = c.response;
        k.octave = l;  // :599
        k.class_id = -1;
      }
      total++;
    }
  }
  return total;
}

int orc_extract_batch(const OrcParams* p, const uint8_t* imgs, int w, int h, int stride, int64_t frame_stride,
                      int batch, OrcKeyPoint* kps, uint8_t* desc, int cap_per_frame, int32_t* n_out, int threads) {
  if (threads < 1) threads = 1;
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; t++) {
    pool.emplace_back([=]() {
      for (int b = t; b < batch; b += threads) {
        int n = orc_extract(p, imgs + (size_t)b * frame_stride, w, h, stride, kps + (size_t)b * cap_per_frame,
                            desc + (size_t)b * cap_per_frame * 32, cap_per_frame, nullptr);
        n_out[b] = std::min(n, cap_per_frame);
      }
    });
  }
  for (auto& th : pool) th.join();
  return 0;
}

/* ------------------------------------------------------------------------------------------------
 * M0  OrbMatcher::computeDescriptorsDistance  src/orbMatcher.cpp:11-23
 * ---------------------------------------------------------------------------------------------- */
int orc_hamming(const uint8_t* a, const uint8_t* b) {
  int32_t pa[8], pb[8];
  memcpy(pa, a, 32);
  memcpy(pb, b, 32);
  int dist = 0;
  for (int i = 0; i < 8; i++) {
    unsigned int v = (unsigned)(pa[i] ^ pb[i]);
    v = v - ((v >> 1) & 0x55555555);
    v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
    dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
  }
  return dist;
}

/* top-2 scan: rule of src/orbMatcher.cpp:319-334 (strict <, else-if strict <) */
void orc_top2(const uint8_t* q, int nq, const uint8_t* db, int ndb, int idx_base, OrcTop2* out, int threads) {
  if (threads < 1) threads = 1;
  auto work = [=](int t) {
    for (int i = t; i < nq; i += threads) {
      int best = 256, second = 256, bi = -1, si = -1;
      for (int j = 0; j < ndb; j++) {
        int d = orc_hamming(q + (size_t)i * 32, db + (size_t)j * 32);
        if (d < best) { second = best; si = bi; best = d; bi = j + idx_base; }
        else if (d < second) { second = d; si = j + idx_base; }
      }
      out[i] = OrcTop2{bi, best, second, si};
    }
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < threads; t++) pool.emplace_back(work, t);
  work(0);
  for (auto& th : pool) th.join();
}

void orc_top2_merge(const OrcTop2* in, int nshards, int nq, OrcTop2* out) {
  // Equivalent to one sequential scan over the concatenated shards (ascending global index).
  for (int i = 0; i < nq; i++) {
    int best = 256, second = 256, bi = -1, si = -1;
    for (int s = 0; s < nshards; s++) {
      const OrcTop2& t = in[(size_t)s * nq + i];
      const int ds[2] = {t.best_dist, t.second_dist};
      const int is[2] = {t.best_idx, t.second_idx};
      for (int k = 0; k < 2; k++) {
        if (is[k] < 0) continue;
        int d = ds[k];
        // candidates arrive in (distance, index) order inside a shard, not index order: use the order-free rule
        if (d < best || (d == best && is[k] < bi)) { second = best; si = bi; best = d; bi = is[k]; }
        else if (d < second || (d == second && (si < 0 || is[k] < si))) { second = d; si = is[k]; }
      }
    }
    out[i] = OrcTop2{bi, best, second, si};
  }
}

void orc_ratio_accept(const OrcTop2* t, int nq, float ratio, int th, uint8_t* accept) {
  for (int i = 0; i < nq; i++)
    accept[i] = (t[i].best_idx >= 0 && t[i].best_dist <= th &&
                 static_cast<float>(t[i].best_dist) < ratio * static_cast<float>(t[i].second_dist))
                    ? 1 : 0;  // src/orbMatcher.cpp:336
}

/* ------------------------------------------------------------------------------------------------
 * G1  Frame::assignKeyPointsToGrid src/frame.cpp:249-264, computeLocationInGrid :327-336
 * G2  Frame::getKeyPointsInArea    src/frame.cpp:337-361
 * ---------------------------------------------------------------------------------------------- */
}  // extern "C"

namespace {
const int kGridCols = 64, kGridRows = 48;  // src/frame.hpp:137-138

struct Grid {
  std::vector<int> cell[kGridCols][kGridRows];
  float inv_w, inv_h;
};

void build_grid(const OrcFrameGeom* g, const OrcKeyPoint* kps, int nkp, Grid& grid) {
  grid.inv_w = static_cast<float>(kGridCols) / (g->max_x - g->min_x);  // src/frame.cpp:99
  grid.inv_h = static_cast<float>(kGridRows) / (g->max_y - g->min_y);  // src/frame.cpp:100
  for (int i = 0; i < nkp; i++) {
    int lx = (int)round((double)((kps[i].x - g->min_x) * grid.inv_w));  // :328
    int ly = (int)round((double)((kps[i].y - g->min_x) * grid.inv_h));  // :329 (uses minX for y, as the reference does)
    if (lx < 0 || lx >= kGridCols || ly < 0 || ly >= kGridRows) continue;
    grid.cell[lx][ly].push_back(i);
  }
}

void keypoints_in_area(const OrcFrameGeom* g, const Grid& grid, const OrcKeyPoint* kps, float x, float y, float r,
                       int minL, int maxL, std::vector<int>& out) {
  out.clear();
  const int minCellX = std::max(0, (int)floor((double)((x - g->min_x - r) * grid.inv_w)));
  const int maxCellX = std::min(kGridCols - 1, (int)ceil((double)((x - g->min_x + r) * grid.inv_w)));
  const int minCellY = std::max(0, (int)floor((double)((y - g->min_y - r) * grid.inv_h)));
  const int maxCellY = std::min(kGridRows - 1, (int)ceil((double)((y - g->min_y + r) * grid.inv_h)));
  if (minCellX < kGridCols && maxCellX >= 0 && minCellY < kGridRows && maxCellY >= 0) {
    for (int ix = minCellX; ix <= maxCellX; ix++)
      for (int iy = minCellY; iy <= maxCellY; iy++)
        for (int idx : grid.cell[ix][iy]) {
          if (minL > 0 || maxL >= 0) {  // :348-352
            if (kps[idx].octave < minL || (maxL >= 0 && kps[idx].octave < maxL)) continue;
          }
          if (fabs(kps[idx].x - x) > r && fabs(kps[idx].y - y) < r) out.push_back(idx);  // :353 (F9)
        }
  }
}
}  // namespace

extern "C" {

int orc_keypoints_in_area(const OrcFrameGeom* g, const OrcKeyPoint* kps, int nkp, float x, float y, float r,
                          int min_level, int max_level, int32_t* out, int cap) {
  Grid* grid = new Grid();
  build_grid(g, kps, nkp, *grid);
  std::vector<int> v;
  keypoints_in_area(g, *grid, kps, x, y, r, min_level, max_level, v);
  delete grid;
  for (size_t i = 0; i < v.size() && (int)i < cap; i++) out[i] = v[i];
  return (int)v.size();
}

/* M1  searchByProjectionInFrameAndMapPoint  src/orbMatcher.cpp:24-64 */
int orc_search_by_projection(const OrcFrameGeom* g, const OrcKeyPoint* kps, const uint8_t* desc, const float* right_x,
                             int nkp, const uint8_t* kp_occupied, const OrcMapPointProj* mps, const uint8_t* mp_desc,
                             const uint8_t* mp_obs_positive, int nmp, float th, float ratio, int32_t* match_out) {
  Grid* grid = new Grid();
  build_grid(g, kps, nkp, *grid);
  // slot state: -2 = holds a pre-existing point with obs>0, -1 = free, >=0 = map point index assigned in this call
  std::vector<int> slot(nkp);
  for (int i = 0; i < nkp; i++) slot[i] = (kp_occupied && kp_occupied[i]) ? -2 : -1;
  int matches = 0;
  std::vector<int> cands;
  for (int m = 0; m < nmp; m++) {
    if (!mps[m].active) continue;  // :27
    const int L = mps[m].level;
    float radius = th * ((mps[m].view_cos > 0.998) ? 2.5f : 4.0f);  // :30, :820-826
    const float r = radius * g->sf[L];
    keypoints_in_area(g, *grid, kps, mps[m].proj_x, mps[m].proj_y, r, L - 1, L, cands);  // :31
    if (cands.empty()) continue;
    int bestDist = 256, bestLevel = -1, secondDist = 256, secondLevel = -1, bestIdx = -1;
    for (int idx : cands) {
      bool free_slot = (slot[idx] == -1) || (slot[idx] >= 0 && !(mp_obs_positive && mp_obs_positive[slot[idx]]));
      if (free_slot && (right_x[idx] <= 0 || fabs(mps[m].proj_right_x - right_x[idx]) <= radius * g->sf[L])) {  // :40-41
        const int d = orc_hamming(mp_desc + (size_t)m * 32, desc + (size_t)idx * 32);
        if (d < bestDist) {
          secondDist = bestDist; bestDist = d; secondLevel = bestLevel; bestLevel = kps[idx].octave; bestIdx = idx;
        } else if (d < secondDist) {
          secondLevel = kps[idx].octave; secondDist = d;
        }
      }
    }
    if (bestDist <= 100 && (bestLevel != secondLevel || bestDist <= ratio * secondDist)) {  // :56
      if (slot[bestIdx] < 0) {
        // nothing
      }
      slot[bestIdx] = m;  // :57 (overwrites; a previous holder with obs<=0 loses the slot)
      matches++;
    }
  }
  for (int i = 0; i < nkp; i++) match_out[i] = slot[i] >= 0 ? slot[i] : -1;
  delete grid;
  return matches;
}

/* ------------------------------------------------------------------------------------------------
 * S1  Frame::computeStereoMatches  src/frame.cpp:362-472
 * ---------------------------------------------------------------------------------------------- */
void orc_stereo_match(const OrcParams* p, int w, int h, const uint8_t* pyr_l, const uint8_t* pyr_r,
                      const OrcKeyPoint* kpl, const uint8_t* dl, int nl, const OrcKeyPoint* kpr, const uint8_t* dr,
                      int nr, float bf, float max_d, int mode, float* right_x, float* depth) {
  for (int i = 0; i < nl; i++) { right_x[i] = -1.0f; depth[i] = -1.0f; }  // :363-364
  const int orbDistThd = (100 + 50) / 2;                                   // :365
  int64_t off[ORC_MAX_LEVELS];
  orc_pyramid_bytes(p, w, h, off);
  int lw[ORC_MAX_LEVELS], lh[ORC_MAX_LEVELS];
  for (int l = 0; l < p->n_levels; l++) orc_level_size(p, w, h, l, &lw[l], &lh[l]);
  const int rowsNum = lh[0];  // :366
  std::vector<std::vector<int>> rows((size_t)rowsNum);
  for (int i = 0; i < nr; i++) {  // :373-379
    const OrcKeyPoint& k = kpr[i];
    const float lo = std::max((float)floor((double)(k.y - 2.0f * p->sf[k.octave])), 0.0f);
    const float hi = std::min((float)ceil((double)(k.y + 2.0f * p->sf[k.octave])), (float)rowsNum - 1.0f);
    for (int iy = (int)lo; iy <= hi; iy++) rows[(size_t)iy].push_back(i);
  }
  const float minD = 0.0f, maxD = max_d;  // :381-382
  std::vector<std::pair<int, int>> distIdx;
  int leftIdx = 0;
  for (int k = 0; k < nl; k++) {  // :387
    const OrcKeyPoint& kp = kpl[k];
    const int slot = (mode == 0) ? leftIdx : k;
    const std::vector<int>& row = rows[(size_t)kp.y];
    if (!row.empty() && kp.x >= minD) {  // :388
      int bestDist = 256, bestRightIdx = 0;
      if (kp.y < (float)rows.size()) {  // :392
        for (int ri : row) {
          const OrcKeyPoint& kr = kpr[ri];
          if (kr.octave >= kp.octave - 1 && kr.octave <= kp.octave + 1 && kr.x >= (kp.x - maxD) && kr.x <= (kp.x - minD)) {
            const int d = orc_hamming(dl + (size_t)slot * 32, dr + (size_t)ri * 32);  // :396 (row leftIdx, F10)
            if (d < bestDist) { bestDist = d; bestRightIdx = ri; }
          }
        }
      }
      if (bestDist < orbDistThd) {  // :405
        const int o = kp.octave;
        const float lsx = (float)round((double)(kp.x * p->inv_sf[o]));  // :407-409
        const float lsy = (float)round((double)(kp.y * p->inv_sf[o]));
        const float rsx = (float)round((double)(kpr[bestRightIdx].x * p->inv_sf[o]));
        const int W = 5;  // :411
        if ((lsy - W) < 0 || (lsy + W + 1) >= lh[o] || (lsx - W) < 0 || (lsx + W + 1) >= lw[o]) continue;  // :412-414
        const int st = lw[o] + 2 * kPad;
        const uint8_t* L = pyr_l + off[o] + (size_t)kPad * st + kPad;
        const uint8_t* R = pyr_r + off[o] + (size_t)kPad * st + kPad;
        const int ly0 = (int)(lsy - W), lx0 = (int)(lsx - W);
        float lp[11][11];
        for (int yy = 0; yy < 11; yy++)
          for (int xx = 0; xx < 11; xx++) lp[yy][xx] = (float)L[(ptrdiff_t)(ly0 + yy) * st + lx0 + xx];
        const float lc = lp[W][W];
        for (int yy = 0; yy < 11; yy++)
          for (int xx = 0; xx < 11; xx++) lp[yy][xx] -= lc;  // :417
        int bestSad = 256;  // :418 shadows the outer bestDist (F10)
        int bestCol = 0;
        const int S = 5;
        float dists[11];
        if ((rsx + S - W) < 0 || (rsx + S + W + 1) >= lw[o]) continue;  // :423-425
        for (int i = -S; i <= S; i++) {
          const int rx0 = (int)(rsx + i - W);
          assert(rx0 >= 0 && rx0 + 11 <= lw[o]);
          double acc = 0;
          const float rc = (float)R[(ptrdiff_t)(ly0 + W) * st + rx0 + W];
          for (int yy = 0; yy < 11; yy++)
            for (int xx = 0; xx < 11; xx++) {
              float rv = (float)R[(ptrdiff_t)(ly0 + yy) * st + rx0 + xx] - rc;
              acc += fabs((double)(lp[yy][xx] - rv));
            }
          const float dist = (float)acc;  // :430
          if (dist < bestSad) { bestSad = (int)dist; bestCol = i; }  // :431-434
          dists[i + S] = dist;
        }
        if (bestCol == -S || bestCol == S) continue;  // :437-439
        const float d1 = dists[S + bestCol - 1], d2 = dists[S + bestCol], d3 = dists[S + bestCol + 1];
        const float delta = (float)((d1 - d3) / (2.0 * ((d1 + d3) - 2.0 * d2)));  // :441-443
        if (delta < -1 || delta > 1) continue;  // :444-446
        float bestRightX = p->sf[o] * ((float)rsx + (float)delta + (float)bestCol);  // :448
        float disparity = kp.x - bestRightX;
        if (disparity >= minD && disparity < maxD) {  // :450
          if (disparity <= 0) {
            disparity = 0.01f;
            bestRightX = (float)(kp.x - 0.01);
          }
          depth[slot] = bf / disparity;  // :455-456
          right_x[slot] = bestRightX;
          distIdx.push_back(std::pair<int, int>(bestSad, slot));  // :457
        }
      }
      leftIdx++;  // :460
    }
  }
  std::sort(distIdx.begin(), distIdx.end());  // :463
  for (const auto& di : distIdx) {            // :464-471
    if (di.first >= 1.5 * 1.4 * distIdx[distIdx.size() / 2].first) {
      right_x[di.second] = -2;
      depth[di.second] = -2;
    } else {
      break;
    }
  }
}


/* ------------------------------------------------------------------------------------------------
 * M7  computeThreeMaxima  src/orbMatcher.cpp:827-854
 * ---------------------------------------------------------------------------------------------- */
void orc_three_maxima(const int32_t* sizes, int len, int32_t* idx1, int32_t* idx2, int32_t* idx3) {
  int max1 = 0, max2 = 0, max3 = 0;
  int i1 = -1, i2 = -1, i3 = -1;  // callers initialise to -1 (:139-141)
  for (int i = 0; i < len; i++) {
    const int n = sizes[i];
    if (n > max1) {
      max3 = max2; max2 = max1; max1 = n; i3 = i2; i2 = i1; i1 = i;
    } else if (n > max2) {
      max3 = max2; max2 = n; i3 = i2; i2 = i;
    } else if (n > max3) {
      max3 = n; i3 = i;
    }
  }
  if (max2 < max1 / 10) {  // integer division, as written (:848)
    i3 = -1; i2 = -1;
  } else if (max3 < max1 / 10) {
    i3 = -1;
  }
  *idx1 = i1; *idx2 = i2; *idx3 = i3;
}

}  // extern "C"

namespace {
const int kHistLen = 30;  // OrbMatcher::m_int_histLen src/orbMatcher.cpp:9

// bin of the rotation histogram (:121-129).  add_is_float: searchForTriangulation adds 360.0f (:521), the others 360.0
int rotation_bin(float angle_a, float angle_b, bool add_is_float) {
  const float factor = 1.0 / kHistLen;  // :78 (double 1/30 rounded to float)
  float rotation = angle_a - angle_b;
  if (rotation < 0.0) {
    if (add_is_float) rotation += 360.0f; else rotation += 360.0;
  }
  int bin = (int)round(rotation * factor);
  if (bin == kHistLen) bin = 0;
  return bin;
}

struct FvIter {  // DBoW3::FeatureVector::const_iterator over the CSR form
  const OrcFeatureVector* fv;
  int pos;
  bool end() const { return pos >= fv->n_nodes; }
  int id() const { return fv->node_ids[pos]; }
  void lower_bound(int id) { pos = (int)(std::lower_bound(fv->node_ids, fv->node_ids + fv->n_nodes, id) - fv->node_ids); }
};
}  // namespace

extern "C" {

/* ------------------------------------------------------------------------------------------------
 * M2 / M3 / M4  window searches  src/orbMatcher.cpp:65-302, :566-807 (see ydorb_oracle.h)
 * ---------------------------------------------------------------------------------------------- */
int orc_search_windows(const OrcFrameGeom* g, const OrcKeyPoint* kps, const uint8_t* desc, const float* right_x, int nkp,
                       const uint8_t* kp_occupied, const OrcWindowQuery* queries, const uint8_t* q_desc, int nq,
                       const OrcWindowParams* prm, int32_t* best_of, int32_t* match_out) {
  Grid* grid = new Grid();
  build_grid(g, kps, nkp, *grid);
  // slot state: -3 cleared by the orientation check, -2 blocked before the call, -1 untouched, >= 0 last claimant
  std::vector<int> slot(nkp);
  for (int i = 0; i < nkp; i++) slot[i] = (kp_occupied && kp_occupied[i]) ? -2 : -1;
  std::vector<std::vector<int>> hist(kHistLen);
  int matches = 0;
  std::vector<int> cands;
  for (int j = 0; j < nq; j++) {
    const OrcWindowQuery& q = queries[j];
    best_of[j] = -1;
    if (!(q.flags & 1)) continue;
    keypoints_in_area(g, *grid, kps, q.x, q.y, q.r, q.min_level, q.max_level, cands);
    if (cands.empty()) continue;  // :102, :190
    int bestDist = 256, bestIdx = -1;
    for (int idx : cands) {
      if (prm->exclusive) {  // :106-107, :194, :284
        const bool blocked = slot[idx] == -2 || (slot[idx] >= 0 && (queries[slot[idx]].flags & 2));
        if (blocked) continue;
      }
      if (q.right_r >= 0 && !(right_x[idx] <= 0.0 || fabs(q.right_x - right_x[idx]) <= q.right_r)) continue;  // :108-109
      const int oct = kps[idx].octave;
      if (oct < q.oct_lo || oct > q.oct_hi) continue;  // :285-286, :616, :716-717, :782
      if (prm->chi2_gate) {                            // :713-719
        const float mono = pow(kps[idx].x - q.x, 2.0) + pow(kps[idx].y - q.y, 2.0);
        const float stereo = mono + pow(right_x[idx] - q.right_x, 2.0);
        const bool ok = (right_x[idx] >= 0 && stereo * prm->inv_sf2[oct] <= 7.81) ||
                        (right_x[idx] < 0 && mono * prm->inv_sf2[oct] <= 5.99);
        if (!ok) continue;
      }
      const int d = orc_hamming(q_desc + (size_t)j * 32, desc + (size_t)idx * 32);
      if (d < bestDist) { bestDist = d; bestIdx = idx; }  // :111-114
    }
    const bool accept = prm->strict ? (bestDist < prm->max_dist) : (bestDist <= prm->max_dist);
    if (!accept || bestIdx < 0) continue;
    best_of[j] = bestIdx;
    slot[bestIdx] = j;  // :118 (overwrites a non-blocking earlier claimant)
    matches++;
    if (prm->check_orientation) hist[rotation_bin(q.angle, kps[bestIdx].angle, false)].push_back(bestIdx);  // :120-131
  }
  if (prm->check_orientation) {  // :138-153
    int32_t sizes[kHistLen], i1, i2, i3;
    for (int b = 0; b < kHistLen; b++) sizes[b] = (int32_t)hist[b].size();
    orc_three_maxima(sizes, kHistLen, &i1, &i2, &i3);
    for (int b = 0; b < kHistLen; b++)
      if (b != i1 && b != i2 && b != i3)
        for (int idx : hist[b]) { slot[idx] = -3; matches--; }
  }
  for (int i = 0; i < nkp; i++) match_out[i] = slot[i] >= 0 ? slot[i] : (slot[i] == -3 ? -2 : -1);
  delete grid;
  return matches;
}

/* ------------------------------------------------------------------------------------------------
 * EXTRA  searchForInitialization.  `north_star` names it, the reference does not have it (absent from
 * src/orbMatcher.hpp:24-66).  This restates the published algorithm the reference descends from -- ORB-SLAM2,
 * ORBmatcher::SearchForInitialization (Mur-Artal & Tardos, ORB_SLAM2 src/ORBmatcher.cc) -- on this reference's own
 * pieces: Frame::getKeyPointsInArea (G2, area_mode 0: the reference's predicate; area_mode 1: ORB-SLAM2's
 * |dx| < r && |dy| < r with minLevel <= octave <= maxLevel), computeDescriptorsDistance (M0), the rotation histogram
 * convention of src/orbMatcher.cpp:120-131 and computeThreeMaxima (M7).  NO reference parity claim: it pins the CUDA
 * path to this sequential statement only.
 *   for every level-0 keypoint i1 of frame 1, in index order: window around prev_matched[i1] in frame 2 (levels 0..0);
 *   best / second distance over candidates whose recorded match distance is still larger than this one
 *   (vMatchedDistance[i2] <= dist -> skip); accept if best <= th_low and best < nn_ratio * second; a keypoint of
 *   frame 2 taken again un-matches its previous owner; orientation cull over every push (un-matched owners included);
 *   prev_matched[i1] := position of the match.
 * ---------------------------------------------------------------------------------------------- */
int orc_search_for_initialization(const OrcFrameGeom* g2, const OrcKeyPoint* kps1, const uint8_t* desc1, int n1,
                                  const OrcKeyPoint* kps2, const uint8_t* desc2, int n2, float* prev_matched,
                                  int window_size, float nn_ratio, int th_low, int check_orientation, int area_mode,
                                  int32_t* matches12) {
  Grid* grid = new Grid();
  build_grid(g2, kps2, n2, *grid);
  for (int i = 0; i < n1; i++) matches12[i] = -1;
  std::vector<int> matched_distance(n2, INT_MAX), matches21(n2, -1);
  std::vector<std::vector<int>> hist(kHistLen);
  int nmatches = 0;
  std::vector<int> cands;
  for (int i1 = 0; i1 < n1; i1++) {
    const int level1 = kps1[i1].octave;
    if (level1 > 0) continue;
    const float x = prev_matched[2 * i1], y = prev_matched[2 * i1 + 1], r = (float)window_size;
    if (area_mode == 0) {
      keypoints_in_area(g2, *grid, kps2, x, y, r, level1, level1, cands);
    } else {
      cands.clear();
      const int minCellX = std::max(0, (int)floor((double)((x - g2->min_x - r) * grid->inv_w)));
      const int maxCellX = std::min(kGridCols - 1, (int)ceil((double)((x - g2->min_x + r) * grid->inv_w)));
      const int minCellY = std::max(0, (int)floor((double)((y - g2->min_y - r) * grid->inv_h)));
      const int maxCellY = std::min(kGridRows - 1, (int)ceil((double)((y - g2->min_y + r) * grid->inv_h)));
      if (minCellX < kGridCols && maxCellX >= 0 && minCellY < kGridRows && maxCellY >= 0)
        for (int ix = minCellX; ix <= maxCellX; ix++)
          for (int iy = minCellY; iy <= maxCellY; iy++)
            for (int idx : grid->cell[ix][iy]) {
              if (kps2[idx].octave < level1 || kps2[idx].octave > level1) continue;
              if (fabsf(kps2[idx].x - x) < r && fabsf(kps2[idx].y - y) < r) cands.push_back(idx);
            }
    }
    if (cands.empty()) continue;
    int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
    for (int i2 : cands) {
      const int dist = orc_hamming(desc1 + (size_t)i1 * 32, desc2 + (size_t)i2 * 32);
      if (matched_distance[i2] <= dist) continue;
      if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
      else if (dist < bestDist2) bestDist2 = dist;
    }
    if (bestDist <= th_low && (float)bestDist < (float)bestDist2 * nn_ratio) {
      if (matches21[bestIdx2] >= 0) { matches12[matches21[bestIdx2]] = -1; nmatches--; }
      matches12[i1] = bestIdx2;
      matches21[bestIdx2] = i1;
      matched_distance[bestIdx2] = bestDist;
      nmatches++;
      if (check_orientation) hist[rotation_bin(kps1[i1].angle, kps2[bestIdx2].angle, false)].push_back(i1);
    }
  }
  if (check_orientation) {
    int32_t sizes[kHistLen], a, b, c;
    for (int k = 0; k < kHistLen; k++) sizes[k] = (int32_t)hist[k].size();
    orc_three_maxima(sizes, kHistLen, &a, &b, &c);
    for (int k = 0; k < kHistLen; k++)
      if (k != a && k != b && k != c)
        for (int i1 : hist[k])
          if (matches12[i1] >= 0) { matches12[i1] = -1; nmatches--; }
  }
  for (int i1 = 0; i1 < n1; i1++)
    if (matches12[i1] >= 0) {
      prev_matched[2 * i1] = kps2[matches12[i1]].x;
      prev_matched[2 * i1 + 1] = kps2[matches12[i1]].y;
    }
  delete grid;
  return nmatches;
}

/* ------------------------------------------------------------------------------------------------
 * M5  searchByBowInKeyFrameAndFrame :303-379 (mode 0), searchByBowInTwoKeyFrames :380-462 (mode 1)
 * ---------------------------------------------------------------------------------------------- */
int orc_search_by_bow(const OrcKeyPoint* kps1, const uint8_t* desc1, const uint8_t* valid1, int n1,
                      const OrcFeatureVector* fv1, const OrcKeyPoint* kps2, const uint8_t* desc2, const uint8_t* valid2,
                      int n2, const OrcFeatureVector* fv2, int mode, float ratio, int check_orientation,
                      int32_t* match_out) {
  const int n_out = mode == 0 ? n2 : n1;
  for (int i = 0; i < n_out; i++) match_out[i] = -1;  // :305, :385
  std::vector<char> matched2(n2, 0);                   // mode 0: _vSptrMatchedMapPoints[frameIdx] != null; mode 1: :386
  std::vector<std::vector<int>> hist(kHistLen);
  int matches = 0;
  FvIter it1{fv1, 0}, it2{fv2, 0};
  while (!it1.end() && !it2.end()) {  // :317, :397
    if (it1.id() == it2.id()) {
      for (int a = fv1->node_start[it1.pos]; a < fv1->node_start[it1.pos + 1]; a++) {
        const int i1 = fv1->items[a];
        if (!valid1[i1]) continue;  // :320, :400
        int bestDist = 256, secondDist = 256, bestIdx = -1;
        for (int b = fv2->node_start[it2.pos]; b < fv2->node_start[it2.pos + 1]; b++) {
          const int i2 = fv2->items[b];
          if (matched2[i2]) continue;                       // :325, :405
          if (mode == 1 && !valid2[i2]) continue;           // :406-407
          const int d = orc_hamming(desc1 + (size_t)i1 * 32, desc2 + (size_t)i2 * 32);
          if (d < bestDist) { secondDist = bestDist; bestDist = d; bestIdx = i2; }
          else if (d < secondDist) { secondDist = d; }
        }
        if (bestDist <= 50 && static_cast<float>(bestDist) < ratio * static_cast<float>(secondDist)) {  // :336, :418
          matched2[bestIdx] = 1;
          const int out_idx = mode == 0 ? bestIdx : i1;
          match_out[out_idx] = mode == 0 ? i1 : bestIdx;  // :337, :419
          if (check_orientation) hist[rotation_bin(kps1[i1].angle, kps2[bestIdx].angle, false)].push_back(out_idx);
          matches++;
        }
      }
      it1.pos++; it2.pos++;
    } else if (it1.id() < it2.id()) {
      it1.lower_bound(it2.id());  // :357
    } else {
      it2.lower_bound(it1.id());  // :359
    }
  }
  if (check_orientation) {  // :362-377
    int32_t sizes[kHistLen], i1, i2, i3;
    for (int b = 0; b < kHistLen; b++) sizes[b] = (int32_t)hist[b].size();
    orc_three_maxima(sizes, kHistLen, &i1, &i2, &i3);
    for (int b = 0; b < kHistLen; b++)
      if (b != i1 && b != i2 && b != i3)
        for (int idx : hist[b]) { match_out[idx] = -1; matches--; }
  }
  return matches;
}

/* ------------------------------------------------------------------------------------------------
 * M6  searchForTriangulation :463-565, isEpipolarLineDistCorrect :808-819
 * ---------------------------------------------------------------------------------------------- */
static bool epipolar_ok(const OrcKeyPoint& k1, const OrcKeyPoint& k2, const float* F, const float* sf2) {
  const float a = F[0] * k1.x + F[3] * k1.y + F[6];  // :810 (column 0 of F: at(0,0), at(1,0), at(2,0))
  const float b = F[1] * k1.x + F[4] * k1.y + F[7];  // :811
  const float c = F[2] * k1.x + F[5] * k1.y + F[8];  // :812
  if ((a * a + b * b) > 0) {
    return ((((a * k2.x + b * k2.y + c) * (a * k2.x + b * k2.y + c)) / ((a * a + b * b) * (a * a + b * b))) <
            (3.841 * sf2[k2.octave]));  // :814-815 (squared denominator, as written)
  }
  return false;
}

int orc_search_for_triangulation(const OrcKeyPoint* kps1, const uint8_t* desc1, const float* right_x1,
                                 const uint8_t* has_mp1, int n1, const OrcFeatureVector* fv1, const OrcKeyPoint* kps2,
                                 const uint8_t* desc2, const float* right_x2, const uint8_t* has_mp2, int n2,
                                 const OrcFeatureVector* fv2, const float* F, float epipole_x, float epipole_y,
                                 const float* sf, const float* sf2, int stereo_only, int check_orientation,
                                 int32_t* match_out) {
  const float xCord = epipole_x, yCord = epipole_y;  // :469-470
  for (int i = 0; i < n1; i++) match_out[i] = -1;    // :476
  std::vector<char> matched2(n2, 0);                 // :475
  std::vector<std::vector<int>> hist(kHistLen);
  int matches = 0;
  FvIter it1{fv1, 0}, it2{fv2, 0};
  while (!it1.end() && !it2.end()) {
    if (it1.id() == it2.id()) {
      for (int a = fv1->node_start[it1.pos]; a < fv1->node_start[it1.pos + 1]; a++) {
        const int i1 = fv1->items[a];
        const bool stereo1 = right_x1[i1] >= 0;  // :489
        if (has_mp1[i1] || (stereo_only && !stereo1)) continue;  // :491
        int bestDist = 50, bestIdx = -1;  // :493-494
        for (int b = fv2->node_start[it2.pos]; b < fv2->node_start[it2.pos + 1]; b++) {
          const int i2 = fv2->items[b];
          const bool stereo2 = right_x2[i2] >= 0;  // :496
          if (matched2[i2] || has_mp2[i2] || (stereo_only && !stereo2)) continue;  // :497-499
          const int d = orc_hamming(desc1 + (size_t)i1 * 32, desc2 + (size_t)i2 * 32);
          if (d <= 50 && d <= bestDist &&
              (stereo1 || stereo2 ||
               pow((float)xCord - (float)(kps2[i2].x), 2.0) + pow((float)yCord - (float)(kps2[i2].y), 2.0) >=
                   100 * sf[kps2[i2].octave])) {  // :502-505
            if (epipolar_ok(kps1[i1], kps2[i2], F, sf2)) { bestIdx = i2; bestDist = d; }  // :506-509
          }
        }
        if (bestIdx >= 0) {  // :513-527
          matched2[bestIdx] = 1;
          match_out[i1] = bestIdx;
          matches++;
          if (check_orientation) hist[rotation_bin(kps1[i1].angle, kps2[bestIdx].angle, true)].push_back(i1);
        }
      }
      it1.pos++; it2.pos++;
    } else if (it1.id() < it2.id()) {
      it1.lower_bound(it2.id());
    } else {
      it2.lower_bound(it1.id());
    }
  }
  if (check_orientation) {  // :539-554
    int32_t sizes[kHistLen], i1, i2, i3;
    for (int b = 0; b < kHistLen; b++) sizes[b] = (int32_t)hist[b].size();
    orc_three_maxima(sizes, kHistLen, &i1, &i2, &i3);
    for (int b = 0; b < kHistLen; b++)
      if (b != i1 && b != i2 && b != i3)
        for (int idx : hist[b]) { match_out[idx] = -1; matches--; }
  }
  return matches;
}

}  // extern "C"
