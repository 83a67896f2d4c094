// TEST INFRASTRUCTURE ONLY -- stand-in for cv::WarperCreator (OpenCV 4.x
// stitching/warpers.hpp).
#pragma once
#include "opencv2/stitching/detail/warpers.hpp"

namespace cv {

class WarperCreator {
 public:
  virtual ~WarperCreator() {}
  virtual Ptr<detail::RotationWarper> create(float scale) const = 0;
};

}  // namespace cv
