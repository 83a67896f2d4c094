// TEST INFRASTRUCTURE: compiles include/xpano_b200_shim.hpp (against the
// OpenCV stand-in in tests/cv_stub/) and drives it exactly as the reference's
// compose loop drives its warper and blender, polymorphically through cv::Ptr
// (xpano/algorithm/stitcher.cc:144-169 ComputeRoi, :357-419 the loop).
//
//   shim_driver <in.bin> <out.bin> [opencv]
//     opencv: install B200MultiBandBlender (BlendingMethod::kOpenCV) instead
//     of B200Blender
// in : int32 n, w, h, projection; float32 scale; n x (K[9], R[9]) float32;
//      n x (h*w*3) u8, rows padded to a step of w*3+12 bytes
// out: int32 W, H, corner_x, corner_y of image 0; pano (H*W*3 u8); mask (H*W);
//      int32 gw, gh; gain map of image 0 (gh*gw float32)
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "xpano_b200_shim.hpp"

namespace b200 = xpano::algorithm::b200;

static void rd(void* p, size_t n, FILE* f) {
  if (fread(p, 1, n, f) != n) {
    fprintf(stderr, "short read\n");
    exit(2);
  }
}

int main(int argc, char** argv) {
  if (argc != 3 && argc != 4) return 2;
  const bool opencv_blender = argc == 4;
  FILE* f = fopen(argv[1], "rb");
  if (!f) return 2;
  int hdr[4];
  float scale;
  rd(hdr, sizeof(hdr), f);
  rd(&scale, 4, f);
  const int n = hdr[0], w = hdr[1], h = hdr[2], projection = hdr[3];
  std::vector<cv::Mat> Ks(n), Rs(n), imgs(n);
  for (int i = 0; i < n; ++i) {
    Ks[i].create(3, 3, CV_32F);
    Rs[i].create(3, 3, CV_32F);
    rd(Ks[i].data, 36, f);
    rd(Rs[i].data, 36, f);
  }
  const size_t step = (size_t)w * 3 + 12;  // cv::Mat with a row step > width
  std::vector<std::vector<unsigned char>> store(n);
  for (int i = 0; i < n; ++i) {
    store[i].assign(step * h, 0xAB);
    for (int y = 0; y < h; ++y) rd(&store[i][step * y], (size_t)w * 3, f);
    imgs[i] = cv::Mat(h, w, CV_8UC3, store[i].data(), step);
  }
  fclose(f);
  try {
    // what algorithm.cc:257-258, 269-270 would install
    cv::Ptr<cv::WarperCreator> creator =
        cv::makePtr<b200::B200WarperCreator>(projection, false, 2.0f, 1.0f);
    cv::Ptr<cv::detail::Blender> blender;
    if (opencv_blender)  // PickBlender, algorithm.cc:132-148
      blender = cv::makePtr<b200::B200MultiBandBlender>();
    else
      blender = cv::makePtr<b200::B200Blender>();
    // what stitcher.cc:200-201 would install
    cv::Ptr<cv::detail::ExposureCompensator> exposure_comp =
        cv::makePtr<b200::B200GainCompensator>();
    cv::Ptr<cv::detail::RotationWarper> warper = creator->create(scale);
    // ComputeRoi (stitcher.cc:144-169)
    long x0 = 1L << 40, y0 = 1L << 40, x1 = -(1L << 40), y1 = -(1L << 40);
    for (int i = 0; i < n; ++i) {
      cv::Rect r = warper->warpRoi(cv::Size(w, h), Ks[i], Rs[i]);
      if (r.x < x0) x0 = r.x;
      if (r.y < y0) y0 = r.y;
      if (r.x + r.width > x1) x1 = r.x + r.width;
      if (r.y + r.height > y1) y1 = r.y + r.height;
    }
    blender->prepare(cv::Rect((int)x0, (int)y0, (int)(x1 - x0), (int)(y1 - y0)));
    // the exposure step of EstimateSeams (stitcher.cc:257-278), here at the
    // compose scale: warp images and masks, feed the compensator
    std::vector<cv::UMat> images_warped(n), masks_warped(n);
    std::vector<cv::Point> corners(n);
    for (int i = 0; i < n; ++i) {
      cv::Mat mask;
      corners[i] = warper->warp(imgs[i], Ks[i], Rs[i], cv::INTER_LINEAR,
                                cv::BORDER_REFLECT, images_warped[i]);
      mask.create(h, w, CV_8U);
      for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) mask.ptr<unsigned char>(y)[x] = 255;
      warper->warp(mask, Ks[i], Rs[i], cv::INTER_NEAREST, cv::BORDER_CONSTANT,
                   masks_warped[i]);
    }
    exposure_comp->feed(corners, images_warped, masks_warped);
    std::vector<cv::Mat> gains;
    exposure_comp->getMatGains(gains);
    cv::Point corner0 = corners[0];
    for (int i = 0; i < n; ++i) {  // stitcher.cc:358-408
      exposure_comp->apply(i, corners[i], images_warped[i], masks_warped[i]);
      blender->feed(images_warped[i], masks_warped[i], corners[i]);
    }
    cv::Mat result, result_mask;
    blender->blend(result, result_mask);  // stitcher.cc:419
    // the point maps the rotate widget uses (rotate.cc:94, 130)
    cv::Point2f p = warper->warpPoint(cv::Point2f(10.5f, 20.25f), Ks[0], Rs[0]);
    cv::Point2f q = warper->warpPointBackward(p, Ks[0], Rs[0]);
    if (!(q.x > 10.0f && q.x < 11.0f && q.y > 19.75f && q.y < 20.75f)) {
      fprintf(stderr, "warpPointBackward(warpPoint(p)) = (%f, %f)\n", q.x, q.y);
      return 3;
    }
    FILE* o = fopen(argv[2], "wb");
    if (!o) return 2;
    int out[4] = {result.cols, result.rows, corner0.x, corner0.y};
    fwrite(out, sizeof(out), 1, o);
    for (int y = 0; y < result.rows; ++y)
      fwrite(result.ptr<unsigned char>(y), 1, (size_t)result.cols * 3, o);
    for (int y = 0; y < result_mask.rows; ++y)
      fwrite(result_mask.ptr<unsigned char>(y), 1, (size_t)result_mask.cols, o);
    int gwh[2] = {gains[0].cols, gains[0].rows};
    fwrite(gwh, sizeof(gwh), 1, o);
    fwrite(gains[0].ptr<float>(), 4, gains[0].total(), o);
    fclose(o);
    // error convention: a second blend on a single-use blender, or feeding
    // after blend, throws std::runtime_error with the library's message
    bool threw = false;
    try {
      blender->feed(result, result_mask, cv::Point(0, 0));
    } catch (const std::runtime_error& e) {
      threw = true;
    }
    if (!threw) {
      fprintf(stderr, "feed after blend did not throw\n");
      return 3;
    }
  } catch (const std::exception& e) {
    fprintf(stderr, "shim_driver: %s\n", e.what());
    return 1;
  }
  printf("shim ok\n");
  return 0;
}
