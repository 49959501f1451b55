/* Stand-in for the reference's src/converter.hpp, which needs Eigen and g2o's GENERATED config.h (thirdParty/g2o has
 * only config.h.in: the header cannot be included without running the reference's cmake).  The build pre-defines that
 * header's own include guard (-DYDORBSLAM_CONVERTER_HPP) and force-includes this file, so src/frame.cpp compiles
 * UNMODIFIED; the one Converter member it names (frame.cpp:267, Frame::computeBoW, not on the hot path) is declared here
 * and defined in ref_frame_driver.cpp.  TEST INFRASTRUCTURE. */
#ifndef YDORB_REFSHIM_CONVERTER_STUB_HPP
#define YDORB_REFSHIM_CONVERTER_STUB_HPP
#include <vector>
#include <opencv2/opencv.hpp>
namespace YDORBSLAM {
class Converter {
 public:
  static std::vector<cv::Mat> descriptors_cvMat_cvMatVector(const cv::Mat& _descriptors);
};
}  // namespace YDORBSLAM
#endif
