// TEST INFRASTRUCTURE ONLY -- stand-in for the virtual interface of
// cv::detail::Blender (OpenCV 4.x stitching/detail/blenders.hpp).
#pragma once
#include "opencv2/core.hpp"

namespace cv {
namespace detail {

class Blender {
 public:
  virtual ~Blender() {}
  virtual void prepare(Rect dst_roi) { dst_roi_ = dst_roi; }
  virtual void feed(InputArray img, InputArray mask, Point tl) = 0;
  virtual void blend(InputOutputArray dst, InputOutputArray dst_mask) = 0;

 protected:
  Rect dst_roi_;
};

}  // namespace detail
}  // namespace cv
