// TEST INFRASTRUCTURE ONLY -- stand-in for the virtual interface of
// cv::detail::ExposureCompensator (OpenCV 4.x
// stitching/detail/exposure_compensate.hpp).
#pragma once
#include <utility>
#include <vector>

#include "opencv2/core.hpp"

namespace cv {
namespace detail {

class ExposureCompensator {
 public:
  ExposureCompensator() : updateGain(true) {}
  virtual ~ExposureCompensator() {}
  // the overload the stitcher calls (stitcher.cc:274): all masks are 255
  void feed(const std::vector<Point>& corners, const std::vector<UMat>& images,
            const std::vector<UMat>& masks) {
    std::vector<std::pair<UMat, uchar>> level_masks;
    for (size_t i = 0; i < masks.size(); ++i)
      level_masks.push_back(std::make_pair(masks[i], (uchar)255));
    feed(corners, images, level_masks);
  }
  virtual void feed(const std::vector<Point>& corners,
                    const std::vector<UMat>& images,
                    const std::vector<std::pair<UMat, uchar>>& masks) = 0;
  virtual void apply(int index, Point corner, InputOutputArray image,
                     InputArray mask) = 0;
  virtual void getMatGains(std::vector<Mat>&) {
    throw Exception("getMatGains: not implemented");
  }
  virtual void setMatGains(std::vector<Mat>&) {
    throw Exception("setMatGains: not implemented");
  }
  void setUpdateGain(bool b) { updateGain = b; }
  bool getUpdateGain() { return updateGain; }

 protected:
  bool updateGain;
};

}  // namespace detail
}  // namespace cv
