/* cvshim_matops.hpp -- Mat arithmetic for the OpenCV stand-in (included from opencv2/opencv.hpp). */
#ifndef YDORB_REFSHIM_MATOPS_HPP
#define YDORB_REFSHIM_MATOPS_HPP
namespace cv {
inline MatExpr Mat::t() const { return MatExpr(); }
inline MatExpr Mat::inv() const { return MatExpr(); }
inline MatExpr Mat::mul(const Mat&, double) const { return MatExpr(); }
inline double Mat::dot(const Mat&) const { return 0; }
inline Mat Mat::cross(const Mat&) const { return Mat(); }
inline MatExpr MatExpr::t() const { return MatExpr(); }
inline MatExpr MatExpr::inv() const { return MatExpr(); }
}  // namespace cv
#endif
