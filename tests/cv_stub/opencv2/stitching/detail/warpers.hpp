// TEST INFRASTRUCTURE ONLY -- stand-in declaring the virtual interface of
// cv::detail::RotationWarper (OpenCV 4.x stitching/detail/warpers.hpp).
#pragma once
#include "opencv2/core.hpp"

namespace cv {
namespace detail {

class RotationWarper {
 public:
  virtual ~RotationWarper() {}
  virtual Point2f warpPoint(const Point2f& pt, InputArray K, InputArray R) = 0;
  virtual Point2f warpPointBackward(const Point2f& pt, InputArray K, InputArray R) = 0;
  virtual Rect buildMaps(Size src_size, InputArray K, InputArray R, OutputArray xmap,
                         OutputArray ymap) = 0;
  virtual Point warp(InputArray src, InputArray K, InputArray R, int interp_mode,
                     int border_mode, OutputArray dst) = 0;
  virtual void warpBackward(InputArray src, InputArray K, InputArray R, int interp_mode,
                            int border_mode, Size dst_size, OutputArray dst) = 0;
  virtual Rect warpRoi(Size src_size, InputArray K, InputArray R) = 0;
  virtual float getScale() const { return 1.f; }
  virtual void setScale(float) {}
};

}  // namespace detail
}  // namespace cv
