// xpano_b200_shim.hpp -- the reference-side binding of libxpano_b200.so: the
// classes a maintainer adds to xpano (e.g. as xpano/algorithm/b200.h) so
// that the library drops in behind the two interfaces the stitcher already
// drives through cv::Ptr,
//   cv::WarperCreator / cv::detail::RotationWarper
//       (Stitcher::SetWarper / SetPortraitWarper, xpano/algorithm/stitcher.h:174-179,
//        installed from xpano/algorithm/algorithm.cc:257-258; calls at
//        stitcher.cc:160, 265, 269, 377, 384; gui/widgets/rotate.cc:94-372)
//   cv::detail::Blender
//       (Stitcher::SetBlender, stitcher.h:205-207, installed from algorithm.cc:269-270;
//        calls at stitcher.cc:357, 405, 419; B200Blender replaces
//        blenders::Multiblend, xpano/algorithm/blenders.h:25-39, and
//        B200MultiBandBlender replaces blenders::MultiBandOpenCV,
//        blenders.h:41-50, the BlendingMethod::kOpenCV choice)
//   cv::detail::ExposureCompensator
//       (Stitcher::SetExposureCompensator, stitcher.h:188-191, installed from
//        stitcher.cc:200-201; calls at stitcher.cc:274, 276, 389; replaces
//        cv::detail::BlocksGainCompensator).
// Header-only; needs OpenCV's stitching headers and include/xpano_b200.h; link
// with -lxpano_b200. Errors follow the reference convention: a non-zero
// xp_status becomes std::runtime_error (blenders.cc:28-31, 137, 156).
//
// In this repository the header is compiled and run against a stand-in for the
// OpenCV declarations (tests/cv_stub/, OpenCV's headers are not in the image)
// by tests/shim_driver.cpp.
#pragma once

#include <stdexcept>
#include <utility>
#include <vector>

#include <opencv2/stitching/detail/blenders.hpp>
#include <opencv2/stitching/detail/exposure_compensate.hpp>
#include <opencv2/stitching/detail/warpers.hpp>
#include <opencv2/stitching/warpers.hpp>

#include "xpano_b200.h"

namespace xpano::algorithm::b200 {

inline void Check(xp_status s) {
  if (s != XP_OK) throw std::runtime_error(xp_last_error());
}

class B200Warper : public cv::detail::RotationWarper {
 public:
  // projection: xpano::algorithm::ProjectionType order (options.h:16-26);
  // a, b: ProjectionOptions::a_param / b_param (options.h:82-86)
  B200Warper(int projection, bool portrait, float scale, float a, float b,
             int device = 0) {
    Check(xp_warper_create(projection, portrait ? 1 : 0, scale, a, b, device, &w_));
  }
  ~B200Warper() override { xp_warper_destroy(w_); }
  B200Warper(const B200Warper&) = delete;
  B200Warper& operator=(const B200Warper&) = delete;

  cv::Rect warpRoi(cv::Size src, cv::InputArray K, cv::InputArray R) override {
    int r[4];
    Check(xp_warper_warp_roi(w_, src.width, src.height, F(K), F(R), r));
    return cv::Rect(r[0], r[1], r[2], r[3]);
  }

  cv::Point warp(cv::InputArray src, cv::InputArray K, cv::InputArray R, int interp,
                 int border, cv::OutputArray dst) override {
    cv::Mat s = src.getMat();
    const cv::Rect roi = warpRoi(s.size(), K, R);
    dst.create(roi.height, roi.width, s.type());
    cv::Mat d = dst.getMat();
    int corner[2];
    Check(xp_warper_warp(
        w_, s.data, s.cols, s.rows, s.step, 8 * static_cast<int>(s.elemSize1()),
        s.channels(), F(K), F(R),
        interp == cv::INTER_NEAREST ? XP_INTER_NEAREST : XP_INTER_LINEAR,
        border == cv::BORDER_CONSTANT ? XP_BORDER_CONSTANT : XP_BORDER_REFLECT, d.data,
        d.step, XP_MEM_HOST, corner));
    return cv::Point(corner[0], corner[1]);
  }

  cv::Point2f warpPoint(const cv::Point2f& pt, cv::InputArray K,
                        cv::InputArray R) override {
    const float in[2] = {pt.x, pt.y};
    float out[2];
    Check(xp_warper_warp_point(w_, in, F(K), F(R), out));
    return cv::Point2f(out[0], out[1]);
  }

  cv::Point2f warpPointBackward(const cv::Point2f& pt, cv::InputArray K,
                                cv::InputArray R) override {
    const float in[2] = {pt.x, pt.y};
    float out[2];
    Check(xp_warper_warp_point_backward(w_, in, F(K), F(R), out));
    return cv::Point2f(out[0], out[1]);
  }

  cv::Rect buildMaps(cv::Size src, cv::InputArray K, cv::InputArray R,
                     cv::OutputArray xmap, cv::OutputArray ymap) override {
    const cv::Rect roi = warpRoi(src, K, R);
    xmap.create(roi.size(), CV_32F);
    ymap.create(roi.size(), CV_32F);
    cv::Mat xm = xmap.getMat(), ym = ymap.getMat();
    int r[4];
    Check(xp_warper_build_maps(w_, src.width, src.height, F(K), F(R),
                               xm.ptr<float>(), ym.ptr<float>(), xm.step1(),
                               XP_MEM_HOST, r));
    return roi;
  }

  float getScale() const override { return xp_warper_get_scale(w_); }
  void setScale(float s) override { Check(xp_warper_set_scale(w_, s)); }

  // not on the export path; xpano never calls it
  void warpBackward(cv::InputArray, cv::InputArray, cv::InputArray, int, int, cv::Size,
                    cv::OutputArray) override {
    throw std::runtime_error("B200Warper: warpBackward is not on the export path");
  }

 private:
  // K and R reach the warper as 3x3 CV_32F (stitcher.cc:158, 372, 540)
  static const float* F(cv::InputArray m) {
    CV_Assert(m.type() == CV_32F && m.total() == 9 && m.getMat().isContinuous());
    return m.getMat().ptr<float>();
  }
  xp_warper* w_ = nullptr;
};

class B200WarperCreator : public cv::WarperCreator {
 public:
  B200WarperCreator(int projection, bool portrait, float a, float b, int device = 0)
      : projection_(projection), portrait_(portrait), a_(a), b_(b), device_(device) {}
  cv::Ptr<cv::detail::RotationWarper> create(float scale) const override {
    return cv::makePtr<B200Warper>(projection_, portrait_, scale, a_, b_, device_);
  }

 private:
  int projection_;
  bool portrait_;
  float a_, b_;
  int device_;
};

class B200Blender : public cv::detail::Blender {
 public:
  explicit B200Blender(int device = 0) {
    xp_blend_opts o;
    xp_blend_opts_default(&o);  // multiblend defaults (mb/multiblend.h:29-45)
    o.output_bpp = 8;           // blenders.cc:20, 145-149
    o.device = device;
    Check(xp_blender_create(&o, &b_));
  }
  ~B200Blender() override { xp_blender_destroy(b_); }
  B200Blender(const B200Blender&) = delete;
  B200Blender& operator=(const B200Blender&) = delete;

  void prepare(cv::Rect dst_roi) override {  // blenders.cc:116
    Check(xp_blender_prepare(b_, dst_roi.x, dst_roi.y, dst_roi.width, dst_roi.height));
  }

  void feed(cv::InputArray img, cv::InputArray mask, cv::Point tl) override {  // :120-139
    cv::Mat i = img.getMat(), m = mask.getMat();
    CV_Assert(i.type() == CV_8UC3);
    CV_Assert(m.type() == CV_8U);
    CV_Assert(m.rows == i.rows && m.cols == i.cols);
    Check(xp_blender_feed(b_, i.data, i.step, m.data, m.step, i.cols, i.rows, 8, tl.x,
                          tl.y, XP_MEM_HOST));
  }

  void blend(cv::InputOutputArray dst, cv::InputOutputArray dst_mask) override {  // :142-158
    int wh[2], min_xy[2], bpp = 0, no_mask = 0;
    Check(xp_blender_blend(b_, wh, min_xy, &bpp, &no_mask, nullptr));
    dst.create(wh[1], wh[0], bpp == 8 ? CV_8UC3 : CV_16UC3);
    dst_mask.create(wh[1], wh[0], CV_8U);
    cv::Mat d = dst.getMat(), m = dst_mask.getMat();
    Check(xp_blender_fetch(b_, d.data, d.step, m.data, m.step, XP_MEM_HOST));
  }

 private:
  xp_blender* b_ = nullptr;
};

// BlendingMethod::kOpenCV (algorithm.cc:135-136): OpenCV's MultiBandBlender
// with xpano's CV_16S feed and CV_8U result (blenders.cc:160-175)
class B200MultiBandBlender : public cv::detail::Blender {
 public:
  explicit B200MultiBandBlender(int num_bands = 5, int device = 0) {
    Check(xp_multiband_create(num_bands, device, &b_));
  }
  ~B200MultiBandBlender() override { xp_multiband_destroy(b_); }
  B200MultiBandBlender(const B200MultiBandBlender&) = delete;
  B200MultiBandBlender& operator=(const B200MultiBandBlender&) = delete;

  void prepare(cv::Rect dst_roi) override {
    Check(xp_multiband_prepare(b_, dst_roi.x, dst_roi.y, dst_roi.width, dst_roi.height));
  }

  void feed(cv::InputArray img, cv::InputArray mask, cv::Point tl) override {
    cv::Mat i = img.getMat(), m = mask.getMat();
    CV_Assert(i.type() == CV_8UC3);
    CV_Assert(m.type() == CV_8U);
    CV_Assert(m.rows == i.rows && m.cols == i.cols);
    Check(xp_multiband_feed(b_, i.data, i.step, m.data, m.step, i.cols, i.rows, tl.x,
                            tl.y, XP_MEM_HOST));
  }

  void blend(cv::InputOutputArray dst, cv::InputOutputArray dst_mask) override {
    int wh[2];
    Check(xp_multiband_blend(b_, wh, nullptr));
    dst.create(wh[1], wh[0], CV_8UC3);
    dst_mask.create(wh[1], wh[0], CV_8U);
    cv::Mat d = dst.getMat(), m = dst_mask.getMat();
    Check(xp_multiband_fetch(b_, d.data, d.step, m.data, m.step, XP_MEM_HOST));
  }

 private:
  xp_multiband* b_ = nullptr;
};

class B200GainCompensator : public cv::detail::ExposureCompensator {
 public:
  // cv::detail::BlocksGainCompensator's defaults (32x32 blocks, two gain
  // filtering iterations), as stitcher.cc:200-201 constructs it
  explicit B200GainCompensator(int bl_width = 32, int bl_height = 32,
                               int nr_gain_filtering_iterations = 2, int device = 0) {
    Check(xp_gain_compensator_create(bl_width, bl_height, nr_gain_filtering_iterations,
                                     device, &g_));
  }
  ~B200GainCompensator() override { xp_gain_compensator_destroy(g_); }
  B200GainCompensator(const B200GainCompensator&) = delete;
  B200GainCompensator& operator=(const B200GainCompensator&) = delete;
  using cv::detail::ExposureCompensator::feed;  // the all-255 overload, stitcher.cc:274

  void feed(const std::vector<cv::Point>& corners, const std::vector<cv::UMat>& images,
            const std::vector<std::pair<cv::UMat, uchar>>& masks) override {
    CV_Assert(corners.size() == images.size() && images.size() == masks.size());
    if (!getUpdateGain() && n_ == static_cast<int>(images.size())) return;
    const int n = static_cast<int>(images.size());
    std::vector<cv::Mat> im(n), mk(n);
    std::vector<int> xy(2 * n), w(n), h(n);
    std::vector<const void*> ip(n);
    std::vector<const uint8_t*> mp(n);
    std::vector<size_t> is(n), ms(n);
    for (int i = 0; i < n; ++i) {
      im[i] = images[i].getMat(cv::ACCESS_READ);
      mk[i] = masks[i].first.getMat(cv::ACCESS_READ);
      CV_Assert(im[i].type() == CV_8UC3 && mk[i].type() == CV_8U);
      CV_Assert(mk[i].rows == im[i].rows && mk[i].cols == im[i].cols);
      CV_Assert(masks[i].second == 255);  // kMaskValueOn, stitcher.cc:254
      xy[2 * i] = corners[i].x;
      xy[2 * i + 1] = corners[i].y;
      w[i] = im[i].cols;
      h[i] = im[i].rows;
      ip[i] = im[i].data;
      mp[i] = mk[i].data;
      is[i] = im[i].step;
      ms[i] = mk[i].step;
    }
    Check(xp_gain_compensator_feed(g_, n, xy.data(), ip.data(), w.data(), h.data(),
                                   is.data(), mp.data(), ms.data(), XP_MEM_HOST));
    n_ = n;
  }

  void apply(int index, cv::Point /*corner*/, cv::InputOutputArray image,
             cv::InputArray /*mask*/) override {
    cv::Mat m = image.getMat();
    CV_Assert(m.type() == CV_8UC3 || m.type() == CV_16UC3);
    Check(xp_gain_compensator_apply(g_, index, m.data, m.cols, m.rows, m.step,
                                    8 * static_cast<int>(m.elemSize1()), XP_MEM_HOST));
  }

  void getMatGains(std::vector<cv::Mat>& gains) override {
    gains.clear();
    for (int i = 0; i < n_; ++i) {
      int wh[2];
      Check(xp_gain_compensator_map_size(g_, i, wh));
      cv::Mat g(wh[1], wh[0], CV_32F);
      Check(xp_gain_compensator_get_map(g_, i, g.ptr<float>()));
      gains.push_back(g);
    }
  }

  void setMatGains(std::vector<cv::Mat>& gains) override {
    for (size_t i = 0; i < gains.size(); ++i) {
      CV_Assert(gains[i].type() == CV_32F && gains[i].isContinuous());
      Check(xp_gain_compensator_set_map(g_, static_cast<int>(i), gains[i].ptr<float>(),
                                        gains[i].cols, gains[i].rows));
    }
    n_ = static_cast<int>(gains.size());
  }

 private:
  xp_gain_compensator* g_ = nullptr;
  int n_ = 0;
};

}  // namespace xpano::algorithm::b200
