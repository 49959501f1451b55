/* cvshim_matops.hpp -- Mat arithmetic for the OpenCV stand-in (included from opencv2/opencv.hpp).  TEST INFRASTRUCTURE.
 *
 * Only CV_32F single-channel matrices (what src/frame.cpp and src/orbMatcher.cpp compute with).  Rounding follows what
 * was pinned against cv2 4.13.0 for the shapes the hot path uses (oracle/ydorb_oracle_map.cpp, tests/test_oracle_map.py):
 *   A * B (+ C)      gemm small-matrix path: float accumulation a0*b0 + a1*b1 + ... left to right, then
 *                    d = (float)((double)t0 * alpha + (double)c * beta)
 *   A + B, A - B     one float operation per element
 *   norm(A)          sqrt of a double sum of squares;  norm(A, B, NORM_L1)  double sum of |a - b|;  A.dot(B)  double
 * Everything is evaluated eagerly; MatExpr only remembers an unfused product so that `A * B + C` becomes one gemm as
 * in OpenCV.  inv() / cross() / mul() are declared for the translation units to compile and abort if ever called.
 */
#ifndef YDORB_REFSHIM_MATOPS_HPP
#define YDORB_REFSHIM_MATOPS_HPP
namespace cv {


static inline void cvshim_unsupported(const char* what) {
  fprintf(stderr, "opencv stand-in: %s is not implemented (not on the hot path)\n", what);
  abort();
}

static inline Mat cvshim_f32(const Mat& m) {
  assert(m.empty() || m.type() == CV_32FC1);
  return m;
}

/* d = alpha * A * B + beta * C  (C may be empty) */
static inline Mat cvshim_gemm(const Mat& A, const Mat& B, double alpha, const Mat& C, double beta) {
  cvshim_f32(A); cvshim_f32(B);
  assert(A.cols == B.rows);
  Mat D(A.rows, B.cols, CV_32FC1);
  for (int i = 0; i < A.rows; i++)
    for (int j = 0; j < B.cols; j++) {
      float t0 = 0.f;
      for (int k = 0; k < A.cols; k++) {
        const float p = A.at<float>(i, k) * B.at<float>(k, j);
        t0 = k == 0 ? p : t0 + p;
      }
      const double c = C.empty() ? 0.0 : (double)C.at<float>(i, j) * beta;
      D.at<float>(i, j) = (float)((double)t0 * alpha + c);
    }
  return D;
}

template <class F>
static inline Mat cvshim_zip(const Mat& A, const Mat& B, F f) {
  cvshim_f32(A); cvshim_f32(B);
  assert(A.rows == B.rows && A.cols == B.cols);
  Mat D(A.rows, A.cols, CV_32FC1);
  for (int i = 0; i < A.rows; i++)
    for (int j = 0; j < A.cols; j++) D.at<float>(i, j) = f(A.at<float>(i, j), B.at<float>(i, j));
  return D;
}

inline MatExpr Mat::t() const {
  cvshim_f32(*this);
  Mat D(cols, rows, CV_32FC1);
  for (int i = 0; i < rows; i++)
    for (int j = 0; j < cols; j++) D.at<float>(j, i) = at<float>(i, j);
  return MatExpr(D);
}
inline MatExpr Mat::inv() const { cvshim_unsupported("Mat::inv"); return MatExpr(); }
inline MatExpr Mat::mul(const Mat&, double) const { cvshim_unsupported("Mat::mul"); return MatExpr(); }
inline double Mat::dot(const Mat& m) const {
  cvshim_f32(*this); cvshim_f32(m);
  double s = 0;
  for (int i = 0; i < rows; i++)
    for (int j = 0; j < cols; j++) s += (double)at<float>(i, j) * (double)m.at<float>(i, j);
  return s;
}
inline Mat Mat::cross(const Mat&) const { cvshim_unsupported("Mat::cross"); return Mat(); }
inline MatExpr MatExpr::t() const { return eval().t(); }
inline MatExpr MatExpr::inv() const { cvshim_unsupported("MatExpr::inv"); return MatExpr(); }

/* ---- operators.  A product stays unfused (kind GEMM) until something needs its value. */
static inline MatExpr operator*(const Mat& a, const Mat& b) { return MatExpr::product(a, b, 1.0); }
static inline MatExpr operator*(const MatExpr& a, const Mat& b) {
  if (a.kind == MatExpr::GEMM) return MatExpr::product(a.eval(), b, 1.0);
  return MatExpr::product(a.eval(), b, 1.0);
}
static inline MatExpr operator*(const Mat& a, const MatExpr& b) { return MatExpr::product(a, b.eval(), 1.0); }
static inline MatExpr operator*(const MatExpr& a, const MatExpr& b) { return MatExpr::product(a.eval(), b.eval(), 1.0); }
/* scalar * initializer (src/frame.cpp:417, :429: value * Mat::ones) and scalar * matrix */
static inline MatExpr operator*(double s, const MatExpr& e) {
  if (e.kind == MatExpr::FILL || e.kind == MatExpr::EYE) {
    MatExpr r = e;
    r.fill = (double)((float)s * (float)e.fill);  // one float multiply per element, as MatOp_AddEx would do
    return r;
  }
  Mat v = e.eval();
  Mat D(v.rows, v.cols, CV_32FC1);
  for (int i = 0; i < v.rows; i++)
    for (int j = 0; j < v.cols; j++) D.at<float>(i, j) = (float)((double)v.at<float>(i, j) * s);
  return MatExpr(D);
}
static inline MatExpr operator*(const MatExpr& e, double s) { return s * e; }
static inline MatExpr operator*(double s, const Mat& m) { return s * MatExpr(m); }
static inline MatExpr operator*(const Mat& m, double s) { return s * MatExpr(m); }
/* matrix / scalar: OpenCV turns it into a scale by 1/s (MatOp_AddEx with alpha = 1.0 / s, double) */
static inline MatExpr operator/(const Mat& m, double s) { return (1.0 / s) * MatExpr(m); }
static inline MatExpr operator/(const MatExpr& e, double s) { return (1.0 / s) * e; }
static inline MatExpr operator-(const Mat& m) { return -1.0 * MatExpr(m); }
static inline MatExpr operator-(const MatExpr& e) { return -1.0 * e; }

static inline MatExpr operator+(const MatExpr& a, const Mat& b) {
  if (a.kind == MatExpr::GEMM && a.c.empty()) {  // A * B + C: one gemm
    MatExpr r = a;
    r.c = b;
    return r;
  }
  return MatExpr(cvshim_zip(a.eval(), b, [](float x, float y) { return x + y; }));
}
static inline MatExpr operator+(const Mat& a, const MatExpr& b) { return b + a; }
static inline MatExpr operator+(const Mat& a, const Mat& b) {
  return MatExpr(cvshim_zip(a, b, [](float x, float y) { return x + y; }));
}
static inline MatExpr operator+(const MatExpr& a, const MatExpr& b) { return a + b.eval(); }
static inline MatExpr operator-(const Mat& a, const Mat& b) {
  return MatExpr(cvshim_zip(a, b, [](float x, float y) { return x - y; }));
}
static inline MatExpr operator-(const MatExpr& a, const Mat& b) { return a.eval() - b; }
static inline MatExpr operator-(const Mat& a, const MatExpr& b) { return a - b.eval(); }
static inline MatExpr operator-(const MatExpr& a, const MatExpr& b) { return a.eval() - b.eval(); }

/* in place: every element of a (a view stays a view) */
static inline Mat& operator-=(Mat& a, const MatExpr& e) {
  const Mat b = e.eval();
  cvshim_f32(a); cvshim_f32(b);
  assert(a.rows == b.rows && a.cols == b.cols);
  for (int i = 0; i < a.rows; i++)
    for (int j = 0; j < a.cols; j++) a.at<float>(i, j) = a.at<float>(i, j) - b.at<float>(i, j);
  return a;
}
static inline Mat& operator-=(Mat& a, const Mat& b) { return a -= MatExpr(b); }
static inline Mat& operator+=(Mat& a, const MatExpr& e) {
  const Mat b = e.eval();
  cvshim_f32(a); cvshim_f32(b);
  for (int i = 0; i < a.rows; i++)
    for (int j = 0; j < a.cols; j++) a.at<float>(i, j) = a.at<float>(i, j) + b.at<float>(i, j);
  return a;
}
static inline Mat& operator+=(Mat& a, const Mat& b) { return a += MatExpr(b); }

/* cv::norm: L2 of one matrix; difference norms of two (NORM_L1 is what src/frame.cpp:430 asks for) */
static inline double norm(const Mat& a, int type = NORM_L2) {
  cvshim_f32(a);
  double s = 0;
  for (int i = 0; i < a.rows; i++)
    for (int j = 0; j < a.cols; j++) {
      const double v = a.at<float>(i, j);
      s += type == NORM_L1 ? fabs(v) : v * v;
    }
  return type == NORM_L1 ? s : sqrt(s);
}
static inline double norm(const MatExpr& e, int type = NORM_L2) { return norm(e.eval(), type); }
static inline double norm(const Mat& a, const Mat& b, int type = NORM_L2) {
  cvshim_f32(a); cvshim_f32(b);
  assert(a.rows == b.rows && a.cols == b.cols);
  double s = 0;
  for (int i = 0; i < a.rows; i++)
    for (int j = 0; j < a.cols; j++) {
      const double v = (double)a.at<float>(i, j) - (double)b.at<float>(i, j);
      s += type == NORM_L1 ? fabs(v) : v * v;
    }
  return type == NORM_L1 ? s : sqrt(s);
}

/* cv::undistortPoints(src, dst, K, dist, R = empty, P): N x 1 CV_32FC2 in place, the model pinned to cv2 in
 * tests/golden/preprocess.npz (oracle/ydorb_oracle_pre.cpp) */
static inline void undistortPoints(const Mat& src, Mat& dst, const Mat& K, const Mat& dist, const Mat& R, const Mat& P) {
  assert(R.empty() && src.type() == CV_32FC2 && K.type() == CV_32FC1 && P.type() == CV_32FC1);
  const int n = src.rows * src.cols;
  std::vector<float> xy(2 * (size_t)n), out(2 * (size_t)n), k(9), p(9), d;
  for (int i = 0; i < n; i++) {
    const float* s = src.ptr<float>(src.cols == 1 ? i : 0) + (src.cols == 1 ? 0 : 2 * i);
    xy[2 * i] = s[0];
    xy[2 * i + 1] = s[1];
  }
  for (int i = 0; i < 9; i++) { k[i] = K.at<float>(i / 3, i % 3); p[i] = P.at<float>(i / 3, i % 3); }
  for (int i = 0; i < dist.rows * dist.cols; i++) d.push_back(dist.at<float>(i));
  orc_undistort_points(xy.data(), n, k.data(), d.data(), (int)d.size(), p.data(), out.data());
  dst.create(src.rows, src.cols, CV_32FC2);
  for (int i = 0; i < n; i++) {
    float* o = dst.ptr<float>(dst.cols == 1 ? i : 0) + (dst.cols == 1 ? 0 : 2 * i);
    o[0] = out[2 * i];
    o[1] = out[2 * i + 1];
  }
}

}  // namespace cv
#endif
