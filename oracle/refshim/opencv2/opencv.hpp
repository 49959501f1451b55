/*
 * opencv2/opencv.hpp -- a STAND-IN for the OpenCV umbrella header.  TEST INFRASTRUCTURE ONLY.
 *
 * Purpose: let the reference's own translation units (/root/reference/src/orbExtractor.cpp, orbMatcher.cpp,
 * frame.cpp, mapPoint.cpp, keyFrame.cpp ...) compile UNMODIFIED in an image that has no OpenCV C++, so
 * that oracle/_ref/libydorb_ref.so runs the reference's control flow (cell loop, std::list quadtree, std::sort ties,
 * stereo lag, greedy matcher order) as the reference's compiler sees it.  The build recipe is oracle/refshim/Makefile;
 * nothing under ydorbslam_b200/ includes or links this.
 *
 * What is here: the closed set of cv:: names those files use, written from OpenCV's documented semantics.
 *  - containers (Mat, Mat_, Point_, Size_, Rect_, KeyPoint, _InputArray/_OutputArray) behave like OpenCV's where the
 *    reference depends on it: Mat is a ref-counted header over a strided buffer, ROIs alias their parent,
 *    at<T>() does no bounds check (the reference reads the 19 px pad through a ROI, src/orbExtractor.cpp:404-413),
 *    create() keeps a buffer whose size and type already match, and `m = Mat::zeros(...)` is a MatExpr assignment that
 *    fills m IN PLACE when the shape matches (src/orbExtractor.cpp:423 relies on it to write into a rowRange view).
 *  - the image primitives (resize, copyMakeBorder, FAST, GaussianBlur, fastAtan2, cvRound, norm, gemm-shaped
 *    operators, undistortPoints) delegate to the integer / float models in oracle/ydorb_oracle*.cpp, which are pinned
 *    bit-for-bit to cv2 4.13.0 outputs (tests/golden/primitives.npz, tests/golden/preprocess.npz).
 *
 * <cmath> is included the way OpenCV's own core/cvstd.hpp includes it, so unqualified cos()/sin()/ceil()/round() in the
 * reference resolve to the same libstdc++ overload set they see in a real build.
 */
#ifndef YDORB_REFSHIM_OPENCV_HPP
#define YDORB_REFSHIM_OPENCV_HPP

#include <algorithm>
#include <cassert>
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstddef>
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <list>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#define CV_CN_SHIFT 3
#define CV_8U 0
#define CV_8S 1
#define CV_16U 2
#define CV_16S 3
#define CV_32S 4
#define CV_32F 5
#define CV_64F 6
#define CV_MAT_DEPTH(t) ((t) & 7)
#define CV_MAT_CN(t) ((((t) >> CV_CN_SHIFT) & 511) + 1)
#define CV_MAKETYPE(depth, cn) (CV_MAT_DEPTH(depth) + (((cn)-1) << CV_CN_SHIFT))
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_8UC3 CV_MAKETYPE(CV_8U, 3)
#define CV_8UC4 CV_MAKETYPE(CV_8U, 4)
#define CV_16UC1 CV_MAKETYPE(CV_16U, 1)
#define CV_32SC1 CV_MAKETYPE(CV_32S, 1)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)
#define CV_32FC2 CV_MAKETYPE(CV_32F, 2)
#define CV_32FC3 CV_MAKETYPE(CV_32F, 3)
#define CV_64FC1 CV_MAKETYPE(CV_64F, 1)

typedef unsigned char uchar;
typedef unsigned short ushort;

/* the pinned primitive models (oracle/ydorb_oracle.h; linked from oracle/_build/libydorb_oracle.so) */
extern "C" {
void orc_resize_linear(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw, int dh, int dstride);
void orc_fill_border101(uint8_t* buf, int w, int h, int pad, int stride);
int orc_fast(const uint8_t* img, int w, int h, int stride, int th, int* xs, int* ys, int* resp, int cap);
void orc_gauss7(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride);
float orc_fast_atan2(float y, float x);
int orc_cvround(float v);
void orc_undistort_points(const float* xy, int n, const float* K, const float* dist, int n_dist, const float* P,
                          float* out);
}

/* cvRound: round-half-to-even in the current rounding mode (cvtss2si / cvtsd2si), core/fast_math.hpp */
static inline int cvRound(float v) { return orc_cvround(v); }
static inline int cvRound(double v) { return (int)lrint(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { return (int)floor(v); }
static inline int cvCeil(double v) { return (int)ceil(v); }

namespace cv {

template <typename T>
static inline T saturate_cast(double v) {
  return (T)v;
}
template <>
inline uchar saturate_cast<uchar>(double v) {
  int iv = (int)lrint(v);
  return (uchar)(iv < 0 ? 0 : iv > 255 ? 255 : iv);
}
template <>
inline int saturate_cast<int>(double v) {
  return (int)lrint(v);
}

enum BorderTypes {
  BORDER_CONSTANT = 0,
  BORDER_REPLICATE = 1,
  BORDER_REFLECT = 2,
  BORDER_WRAP = 3,
  BORDER_REFLECT_101 = 4,
  BORDER_REFLECT101 = 4,
  BORDER_DEFAULT = 4,
  BORDER_ISOLATED = 16
};
enum InterpolationFlags { INTER_NEAREST = 0, INTER_LINEAR = 1, INTER_CUBIC = 2, INTER_AREA = 3 };
enum NormTypes { NORM_INF = 1, NORM_L1 = 2, NORM_L2 = 4 };

/* ---- geometry PODs (core/types.hpp) ------------------------------------------------------------------------- */
template <typename T>
struct Point_ {
  T x, y;
  Point_() : x(0), y(0) {}
  Point_(T _x, T _y) : x(_x), y(_y) {}
  template <typename U>
  Point_(const Point_<U>& p) : x((T)p.x), y((T)p.y) {}
};
typedef Point_<int> Point2i;
typedef Point2i Point;
typedef Point_<float> Point2f;
typedef Point_<double> Point2d;
/* core/types.hpp: three overloads (int, float, double), each a.x = saturate_cast<T>(a.x * b) */
template <typename T>
static inline Point_<T>& operator*=(Point_<T>& a, int b) {
  a.x = (T)(a.x * b);
  a.y = (T)(a.y * b);
  return a;
}
template <typename T>
static inline Point_<T>& operator*=(Point_<T>& a, float b) {
  a.x = (T)(a.x * b);
  a.y = (T)(a.y * b);
  return a;
}
template <typename T>
static inline Point_<T>& operator*=(Point_<T>& a, double b) {
  a.x = (T)(a.x * b);
  a.y = (T)(a.y * b);
  return a;
}
template <typename T>
static inline bool operator==(const Point_<T>& a, const Point_<T>& b) {
  return a.x == b.x && a.y == b.y;
}

template <typename T>
struct Point3_ {
  T x, y, z;
  Point3_() : x(0), y(0), z(0) {}
  Point3_(T _x, T _y, T _z) : x(_x), y(_y), z(_z) {}
};
typedef Point3_<float> Point3f;
typedef Point3_<double> Point3d;

template <typename T>
struct Size_ {
  T width, height;
  Size_() : width(0), height(0) {}
  Size_(T w, T h) : width(w), height(h) {}
};
typedef Size_<int> Size2i;
typedef Size2i Size;

template <typename T>
struct Rect_ {
  T x, y, width, height;
  Rect_() : x(0), y(0), width(0), height(0) {}
  Rect_(T _x, T _y, T w, T h) : x(_x), y(_y), width(w), height(h) {}
};
typedef Rect_<int> Rect;

struct Range {
  int start, end;
  Range() : start(0), end(0) {}
  Range(int s, int e) : start(s), end(e) {}
  static Range all() { return Range(INT_MIN, INT_MAX); }
};

struct Scalar {
  double val[4];
  Scalar() { val[0] = val[1] = val[2] = val[3] = 0; }
  Scalar(double v0) {
    val[0] = v0;
    val[1] = val[2] = val[3] = 0;
  }
  double operator[](int i) const { return val[i]; }
  static Scalar all(double v) {
    Scalar s;
    s.val[0] = s.val[1] = s.val[2] = s.val[3] = v;
    return s;
  }
};

/* core/types.hpp KeyPoint: 28 bytes, same field order */
class KeyPoint {
 public:
  KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
  KeyPoint(Point2f _pt, float _size, float _angle = -1, float _response = 0, int _octave = 0, int _class_id = -1)
      : pt(_pt), size(_size), angle(_angle), response(_response), octave(_octave), class_id(_class_id) {}
  KeyPoint(float x, float y, float _size, float _angle = -1, float _response = 0, int _octave = 0, int _class_id = -1)
      : pt(x, y), size(_size), angle(_angle), response(_response), octave(_octave), class_id(_class_id) {}
  Point2f pt;
  float size;
  float angle;
  float response;
  int octave;
  int class_id;
};

/* ---- Mat ---------------------------------------------------------------------------------------------------- */
class Mat;
class MatExpr;
/* core/persistence.hpp: the YAML / XML store.  The interface is here so that translation units naming it compile;
 * the stand-in holds no parser -- opening a store reports "not opened", touching one throws.  DBoW3's binary vocabulary
 * format (Vocabulary::toStream / fromStream) does not go through it. */
class FileNode {
 public:
  FileNode() {}
  FileNode operator[](const std::string&) const { return fail(); }
  FileNode operator[](const char*) const { return fail(); }
  FileNode operator[](int) const { return fail(); }
  size_t size() const { return 0; }
  operator int() const { return fail(), 0; }
  operator float() const { return fail(), 0.f; }
  operator double() const { return fail(), 0.0; }
  operator std::string() const { return fail(), std::string(); }

 private:
  static FileNode fail() { throw std::runtime_error("cv::FileNode: YAML persistence is not part of the OpenCV stand-in"); }
};
class FileStorage {
 public:
  enum Mode { READ = 0, WRITE = 1, APPEND = 2, MEMORY = 4 };
  FileStorage() {}
  FileStorage(const std::string&, int) {}
  bool isOpened() const { return false; }
  void release() {}
  FileNode operator[](const std::string&) const { return FileNode()[0]; }
  FileNode operator[](const char*) const { return FileNode()[0]; }
  template <class T>
  FileStorage& operator<<(const T&) {
    throw std::runtime_error("cv::FileStorage: YAML persistence is not part of the OpenCV stand-in");
  }
};

static inline size_t cvshim_elem_size1(int depth) {
  static const size_t tab[8] = {1, 1, 2, 2, 4, 4, 8, 2};
  return tab[depth & 7];
}

class Mat {
 public:
  int flags;  // holds the type only
  int rows, cols;
  size_t step;  // bytes per row
  uchar* data;
  std::shared_ptr<uchar> owner;  // keeps the allocation alive for every header that views it
  uchar* datastart;              // allocation bounds, used by isSubmatrix()
  size_t datarows, datacols;

  Mat() : flags(0), rows(0), cols(0), step(0), data(nullptr), datastart(nullptr), datarows(0), datacols(0) {}
  Mat(int r, int c, int type) : Mat() { create(r, c, type); }
  Mat(Size sz, int type) : Mat() { create(sz.height, sz.width, type); }
  Mat(int r, int c, int type, const Scalar& s) : Mat() {
    create(r, c, type);
    setTo(s);
  }
  /* user-allocated data: no copy, no ownership (Mat(rows, cols, type, void* data, size_t step)) */
  Mat(int r, int c, int type, void* d, size_t st = 0)
      : flags(type), rows(r), cols(c), step(st ? st : (size_t)c * cvshim_elem_size1(CV_MAT_DEPTH(type)) * CV_MAT_CN(type)),
        data((uchar*)d), datastart((uchar*)d), datarows(r), datacols(c) {}
  Mat(const Mat&) = default;
  Mat& operator=(const Mat&) = default;
  inline Mat(const MatExpr& e);
  inline Mat& operator=(const MatExpr& e);
  Mat& operator=(const Scalar& s) { return setTo(s); }

  int type() const { return flags; }
  int depth() const { return CV_MAT_DEPTH(flags); }
  int channels() const { return CV_MAT_CN(flags); }
  size_t elemSize1() const { return cvshim_elem_size1(depth()); }
  size_t elemSize() const { return elemSize1() * channels(); }
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  size_t total() const { return (size_t)rows * cols; }
  Size size() const { return Size(cols, rows); }
  bool isContinuous() const { return rows <= 1 || step == (size_t)cols * elemSize(); }
  bool isSubmatrix() const { return data && ((size_t)rows != datarows || (size_t)cols != datacols); }

  /* Mat::create: a no-op when shape and type already match (this is what lets resize / copyMakeBorder write into
   * the ROI views of src/orbExtractor.cpp:612-617) */
  void create(int r, int c, int type) {
    if (data && rows == r && cols == c && flags == type) return;
    release();
    flags = type;
    rows = r;
    cols = c;
    step = (size_t)c * elemSize();
    size_t bytes = step * (size_t)r;
    if (bytes == 0) return;
    uchar* p = (uchar*)malloc(bytes + 64);
    owner = std::shared_ptr<uchar>(p, free);
    data = datastart = p;
    datarows = r;
    datacols = c;
  }
  void create(Size sz, int type) { create(sz.height, sz.width, type); }
  void release() {
    owner.reset();
    data = datastart = nullptr;
    rows = cols = 0;
    step = 0;
    datarows = datacols = 0;
  }

  template <typename T>
  T& at(int r, int c) {
    return ((T*)(data + step * (ptrdiff_t)r))[c];
  }
  template <typename T>
  const T& at(int r, int c) const {
    return ((const T*)(data + step * (ptrdiff_t)r))[c];
  }
  /* 1-D form (core/mat.inl.hpp): rows==1 -> column index, cols==1 -> row index, else continuous linear index */
  template <typename T>
  T& at(int i) {
    if (rows == 1) return ((T*)data)[i];
    if (cols == 1) return *(T*)(data + step * (ptrdiff_t)i);
    int r = i / cols;
    return ((T*)(data + step * (ptrdiff_t)r))[i - r * cols];
  }
  template <typename T>
  const T& at(int i) const {
    return const_cast<Mat*>(this)->at<T>(i);
  }
  template <typename T>
  T* ptr(int r = 0) {
    return (T*)(data + step * (ptrdiff_t)r);
  }
  template <typename T>
  const T* ptr(int r = 0) const {
    return (const T*)(data + step * (ptrdiff_t)r);
  }
  uchar* ptr(int r = 0) { return data + step * (ptrdiff_t)r; }
  const uchar* ptr(int r = 0) const { return data + step * (ptrdiff_t)r; }

  Mat operator()(const Rect& roi) const {
    Mat m(*this);
    m.data = data + step * (ptrdiff_t)roi.y + (ptrdiff_t)roi.x * (ptrdiff_t)elemSize();
    m.rows = roi.height;
    m.cols = roi.width;
    return m;
  }
  Mat rowRange(int a, int b) const { return (*this)(Rect(0, a, cols, b - a)); }
  Mat colRange(int a, int b) const { return (*this)(Rect(a, 0, b - a, rows)); }
  Mat row(int y) const { return rowRange(y, y + 1); }
  Mat col(int x) const { return colRange(x, x + 1); }

  Mat clone() const {
    Mat m;
    copyTo(m);
    return m;
  }
  /* copyTo: dst.create(size, type) then row-wise copy (same-buffer copies are harmless) */
  void copyTo(Mat& dst) const {
    if (empty()) {
      dst.release();
      return;
    }
    dst.create(rows, cols, flags);
    if (dst.data == data && dst.step == step) return;
    const size_t rb = (size_t)cols * elemSize();
    for (int y = 0; y < rows; y++) memmove(dst.data + dst.step * y, data + step * y, rb);
  }
  void copyTo(Mat&& dst) const {
    Mat d(dst);
    copyTo(d);
  }

  Mat& setTo(const Scalar& s) {
    for (int y = 0; y < rows; y++)
      for (int x = 0; x < cols * channels(); x++) store(y, x, s.val[x % channels() < 4 ? x % channels() : 0]);
    return *this;
  }

  /* element access in double for the generic helpers below (depth-dispatched) */
  double load(int r, int x) const {
    const uchar* p = data + step * (ptrdiff_t)r;
    switch (depth()) {
      case CV_8U: return p[x];
      case CV_8S: return ((const signed char*)p)[x];
      case CV_16U: return ((const ushort*)p)[x];
      case CV_16S: return ((const short*)p)[x];
      case CV_32S: return ((const int*)p)[x];
      case CV_32F: return ((const float*)p)[x];
      default: return ((const double*)p)[x];
    }
  }
  void store(int r, int x, double v) {
    uchar* p = data + step * (ptrdiff_t)r;
    switch (depth()) {
      case CV_8U: p[x] = saturate_cast<uchar>(v); break;
      case CV_8S: ((signed char*)p)[x] = (signed char)lrint(v); break;
      case CV_16U: ((ushort*)p)[x] = (ushort)lrint(v); break;
      case CV_16S: ((short*)p)[x] = (short)lrint(v); break;
      case CV_32S: ((int*)p)[x] = (int)lrint(v); break;
      case CV_32F: ((float*)p)[x] = (float)v; break;
      default: ((double*)p)[x] = v; break;
    }
  }

  /* convertTo(dst, rtype): plain cast (alpha=1, beta=0).  In place when dst is *this: a NEW buffer is created
   * because the type changes -- the parent of a ROI is never touched (src/frame.cpp:416, :428). */
  void convertTo(Mat& dst, int rtype) const {
    rtype = CV_MAKETYPE(CV_MAT_DEPTH(rtype), channels());
    Mat src(*this);
    Mat out;
    out.create(rows, cols, rtype);
    for (int y = 0; y < rows; y++)
      for (int x = 0; x < cols * channels(); x++) out.store(y, x, src.load(y, x));
    if (dst.data && dst.rows == rows && dst.cols == cols && dst.flags == rtype && dst.data != data)
      out.copyTo(dst);
    else
      dst = out;
  }

  /* reshape(cn): same data, new channel count (continuous or column-preserving case only) */
  Mat reshape(int cn) const {
    Mat m(*this);
    int total_ch = cols * channels();
    assert(total_ch % cn == 0);
    m.cols = total_ch / cn;
    m.flags = CV_MAKETYPE(depth(), cn);
    return m;
  }

  inline MatExpr t() const;
  inline MatExpr inv() const;
  inline MatExpr mul(const Mat& m, double scale = 1) const;
  inline double dot(const Mat& m) const;
  inline Mat cross(const Mat& m) const;

  static inline MatExpr zeros(int r, int c, int type);
  static inline MatExpr zeros(Size sz, int type);
  static inline MatExpr ones(int r, int c, int type);
  static inline MatExpr eye(int r, int c, int type);

  void push_back(const Mat& m) {
    if (empty()) {
      m.copyTo(*this);
      return;
    }
    assert(m.cols == cols && m.flags == flags);
    Mat out;
    out.create(rows + m.rows, cols, flags);
    const size_t rb = (size_t)cols * elemSize();
    for (int y = 0; y < rows; y++) memcpy(out.data + out.step * y, data + step * y, rb);
    for (int y = 0; y < m.rows; y++) memcpy(out.data + out.step * (rows + y), m.data + m.step * y, rb);
    *this = out;
  }
};

/* Mat_<T>: typed header, with the comma initialiser used by src/frame.cpp:290, :476 */
template <typename T>
struct cvshim_type;
template <>
struct cvshim_type<float> {
  enum { value = CV_32FC1 };
};
template <>
struct cvshim_type<double> {
  enum { value = CV_64FC1 };
};
template <>
struct cvshim_type<uchar> {
  enum { value = CV_8UC1 };
};
template <>
struct cvshim_type<int> {
  enum { value = CV_32SC1 };
};

template <typename T>
class MatCommaInitializer_;

template <typename T>
class Mat_ : public Mat {
 public:
  Mat_() : Mat() { flags = cvshim_type<T>::value; }
  Mat_(int r, int c) : Mat(r, c, cvshim_type<T>::value) {}
  Mat_(const Mat& m) : Mat(m) { assert(m.empty() || m.type() == cvshim_type<T>::value); }
  T& operator()(int r, int c) { return this->template at<T>(r, c); }
  const T& operator()(int r, int c) const { return this->template at<T>(r, c); }
};

template <typename T>
class MatCommaInitializer_ {
 public:
  explicit MatCommaInitializer_(Mat_<T>* m) : mat(m), idx(0) {}
  template <typename U>
  MatCommaInitializer_<T>& operator,(U v) {
    put((T)v);
    return *this;
  }
  void put(T v) {
    int r = idx / mat->cols, c = idx % mat->cols;
    mat->template at<T>(r, c) = v;
    idx++;
  }
  operator Mat_<T>() const { return *mat; }
  operator Mat() const { return *mat; }
  Mat_<T>* mat;
  int idx;
};
template <typename T, typename U>
static inline MatCommaInitializer_<T> operator<<(const Mat_<T>& m, U v) {
  MatCommaInitializer_<T> ci(const_cast<Mat_<T>*>(&m));
  ci.put((T)v);
  return ci;
}

/* ---- MatExpr: eager value + the two lazily-assigned forms the reference relies on ------------------------------- */
class MatExpr {
 public:
  enum Kind { VALUE = 0, FILL = 1, EYE = 2, GEMM = 3 };
  Kind kind;
  Mat value;            // VALUE
  Mat a, b, c;          // GEMM: alpha * a * b (+ c), kept unfused so that `A * B + C` is ONE gemm as in OpenCV
  double alpha;
  int rows, cols, type; // FILL / EYE
  double fill;
  MatExpr() : kind(VALUE), alpha(1), rows(0), cols(0), type(0), fill(0) {}
  MatExpr(const Mat& m) : kind(VALUE), value(m), alpha(1), rows(m.rows), cols(m.cols), type(m.type()), fill(0) {}
  static MatExpr product(const Mat& x, const Mat& y, double al) {
    MatExpr e;
    e.kind = GEMM;
    e.a = x;
    e.b = y;
    e.alpha = al;
    e.rows = x.rows;
    e.cols = y.cols;
    e.type = x.type();
    return e;
  }
  static MatExpr filled(Kind k, int r, int c, int t, double v) {
    MatExpr e;
    e.kind = k;
    e.rows = r;
    e.cols = c;
    e.type = t;
    e.fill = v;
    return e;
  }
  /* MatOp_Initializer::assign: m.create(size, type) [keeps a matching buffer], then fill */
  inline Mat gemm_value() const;
  void assignTo(Mat& m) const {
    if (kind == VALUE) {
      m = value;
      return;
    }
    if (kind == GEMM) {
      m = gemm_value();
      return;
    }
    m.create(rows, cols, type);
    if (kind == FILL) {
      m.setTo(Scalar(fill));
    } else {
      m.setTo(Scalar(0));
      for (int i = 0; i < std::min(rows, cols); i++) m.store(i, i * m.channels(), fill);
    }
  }
  Mat eval() const {
    Mat m;
    assignTo(m);
    return m;
  }
  operator Mat() const { return eval(); }
  template <typename T>
  operator Mat_<T>() const {
    return Mat_<T>(eval());
  }
  MatExpr t() const;
  MatExpr inv() const;
  Mat col(int x) const { return eval().col(x); }
  Mat row(int y) const { return eval().row(y); }
  double dot(const Mat& m) const { return eval().dot(m); }
  Size size() const { return Size(cols, rows); }
};

inline Mat::Mat(const MatExpr& e) : Mat() { e.assignTo(*this); }
inline Mat& Mat::operator=(const MatExpr& e) {
  e.assignTo(*this);
  return *this;
}
inline MatExpr Mat::zeros(int r, int c, int type) { return MatExpr::filled(MatExpr::FILL, r, c, type, 0); }
inline MatExpr Mat::zeros(Size sz, int type) { return MatExpr::filled(MatExpr::FILL, sz.height, sz.width, type, 0); }
inline MatExpr Mat::ones(int r, int c, int type) { return MatExpr::filled(MatExpr::FILL, r, c, type, 1); }
inline MatExpr Mat::eye(int r, int c, int type) { return MatExpr::filled(MatExpr::EYE, r, c, type, 1); }

/* ---- _InputArray / _OutputArray (core/mat.hpp): thin proxies; InputArray is `const _InputArray&` ----------------- */
class _InputArray {
 public:
  _InputArray() : m(nullptr) {}
  _InputArray(const Mat& _m) : m(const_cast<Mat*>(&_m)) {}
  _InputArray(const MatExpr& e) : held(e.eval()), m(&held) {}
  Mat getMat() const { return m ? *m : Mat(); }
  bool empty() const { return !m || m->empty(); }
  Size size() const { return m ? m->size() : Size(); }
  int type() const { return m ? m->type() : 0; }

 protected:
  Mat held;
  Mat* m;
};
class _OutputArray : public _InputArray {
 public:
  _OutputArray() {}
  _OutputArray(Mat& _m) { m = &_m; }
  void create(int r, int c, int type) const { m->create(r, c, type); }
  void create(Size sz, int type) const { m->create(sz.height, sz.width, type); }
  void release() const { m->release(); }
  Mat& getMatRef() const { return *m; }
  bool needed() const { return m != nullptr; }
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;
typedef const _OutputArray& InputOutputArray;
static inline const _InputArray& noArray() {
  static _InputArray none;
  return none;
}

/* ---- image primitives: signatures of imgproc.hpp / features2d.hpp / core.hpp ------------------------------------- */
struct cvshim_fast_call {
  const uchar* data;
  int rows, cols, threshold, count;
};
/* optional per-thread trace of cv::FAST calls (lets the driver report the reference's raw per-cell candidate lists) */
struct cvshim_trace {
  std::vector<cvshim_fast_call> fast_calls;
  std::vector<KeyPoint> fast_keypoints;
};
inline cvshim_trace*& cvshim_trace_ptr() {
  static thread_local cvshim_trace* t = nullptr;
  return t;
}

/* cv::resize: only what the reference calls -- 8UC1, dsize given, INTER_LINEAR (src/orbExtractor.cpp:614) */
static inline void resize(InputArray _src, OutputArray _dst, Size dsize, double fx = 0, double fy = 0,
                          int interpolation = INTER_LINEAR) {
  Mat src = _src.getMat();
  assert(src.type() == CV_8UC1 && interpolation == INTER_LINEAR && dsize.width > 0 && dsize.height > 0);
  (void)fx;
  (void)fy;
  _dst.create(dsize, src.type());
  Mat dst = _dst.getMat();
  assert(dst.data != src.data);
  orc_resize_linear(src.data, src.cols, src.rows, (int)src.step, dst.data, dst.cols, dst.rows, (int)dst.step);
}

/* cv::copyMakeBorder: BORDER_REFLECT_101 with equal margins, optionally | BORDER_ISOLATED
 * (src/orbExtractor.cpp:615-618).  core/copy.cpp: dst.create(...), copy the interior unless it already is the source
 * buffer, then synthesise the border from the interior.  Without BORDER_ISOLATED a sub-matrix source would take the
 * border from its parent; the reference's non-isolated call passes a whole image, asserted here. */
static inline void copyMakeBorder(InputArray _src, OutputArray _dst, int top, int bottom, int left, int right,
                                  int borderType, const Scalar& = Scalar()) {
  Mat src = _src.getMat();
  assert(src.type() == CV_8UC1);
  assert((borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101);
  assert(top == bottom && top == left && top == right);
  assert((borderType & BORDER_ISOLATED) || !src.isSubmatrix());
  _dst.create(src.rows + top + bottom, src.cols + left + right, src.type());
  Mat dst = _dst.getMat();
  uchar* inner = dst.data + dst.step * top + left;
  if (inner != src.data)
    for (int y = 0; y < src.rows; y++) memmove(inner + dst.step * y, src.data + src.step * y, (size_t)src.cols);
  orc_fill_border101(dst.data, src.cols, src.rows, top, (int)dst.step);
}

/* cv::FAST(image, keypoints, threshold, nonmaxSuppression) = TYPE_9_16 (features2d.hpp) */
static inline void FAST(InputArray _img, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression = true) {
  Mat img = _img.getMat();
  assert(img.type() == CV_8UC1 && nonmaxSuppression);
  keypoints.clear();
  const int cap = std::max(img.rows * img.cols, 1);
  std::vector<int> xs(cap), ys(cap), rs(cap);
  int n = orc_fast(img.data, img.cols, img.rows, (int)img.step, threshold, xs.data(), ys.data(), rs.data(), cap);
  keypoints.reserve(n);
  for (int i = 0; i < n; i++) keypoints.push_back(KeyPoint((float)xs[i], (float)ys[i], 7.f, -1.f, (float)rs[i]));
  if (cvshim_trace* t = cvshim_trace_ptr()) {
    cvshim_fast_call c = {img.data, img.rows, img.cols, threshold, n};
    t->fast_calls.push_back(c);
    t->fast_keypoints.insert(t->fast_keypoints.end(), keypoints.begin(), keypoints.end());
  }
}

/* cv::GaussianBlur: 7x7, sigma 2, BORDER_REFLECT_101 on 8UC1 (src/orbExtractor.cpp:386), in place allowed */
static inline void GaussianBlur(InputArray _src, OutputArray _dst, Size ksize, double sigmaX, double sigmaY = 0,
                                int borderType = BORDER_DEFAULT) {
  Mat src = _src.getMat();
  assert(src.type() == CV_8UC1 && ksize.width == 7 && ksize.height == 7 && sigmaX == 2 && sigmaY == 2);
  assert((borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101);
  assert(!src.isSubmatrix() || (borderType & BORDER_ISOLATED));
  Mat tmp;
  tmp.create(src.rows, src.cols, src.type());
  orc_gauss7(src.data, src.cols, src.rows, (int)src.step, tmp.data, (int)tmp.step);
  _dst.create(src.rows, src.cols, src.type());
  Mat dst = _dst.getMat();
  tmp.copyTo(dst);
}

static inline float fastAtan2(float y, float x) { return orc_fast_atan2(y, x); }

}  // namespace cv

#include "cvshim_matops.hpp"
namespace cv {
inline Mat MatExpr::gemm_value() const { return cvshim_gemm(a, b, alpha, c, 1.0); }
}

#endif  // YDORB_REFSHIM_OPENCV_HPP
