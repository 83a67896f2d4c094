// TEST INFRASTRUCTURE ONLY -- not part of the product path.
//
// extern "C" driver around the UNMODIFIED reference blend library
// (external/multiblend @ 8cb2c04, sources compiled where they lie under
// /root/reference by oracle/build_ref.sh into oracle/_ref/libmb_ref.so).
// It feeds in-memory images exactly as xpano's wrapper does
// (xpano/algorithm/blenders.cc:120-158): InMemoryImage{w,h,bpp,spp,xpos,ypos,
// bytes} -> multiblend::Multiblend(images, {MB_IN_MEMORY, output_bpp}, pool),
// then hands back the three planes, the decoded full_mask RLE
// (blenders.cc:38-62) and Result.timing (mb/include/mb/multiblend.h:16-27).
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may load this library.

#include <cstdint>
#include <cstring>
#include <exception>
#include <memory>
#include <string>
#include <vector>

#include "mb/image.h"
#include "mb/multiblend.h"
#include "mb/threadpool.h"

namespace {

struct RefResult {
  int width = 0, height = 0, bpp = 0, min_x = 0, min_y = 0, no_mask = 0;
  std::vector<uint8_t> planes[3];
  std::vector<uint8_t> mask;
  double timing[10] = {0};
};

thread_local std::string g_err;

}  // namespace

// multiblend::Options::wideblend for the next mb_ref_blend call
static int g_wideblend = 0;

extern "C" {

void mb_ref_set_wideblend(int on) { g_wideblend = on; }


const char* mb_ref_last_error() { return g_err.c_str(); }

// data[i]: tightly packed rows, w*h*spp samples of bpp bits.
// Returns an opaque handle or nullptr on error.
void* mb_ref_blend(int n, const int* w, const int* h, const int* bpp,
                   const int* spp, const int* xpos, const int* ypos,
                   const uint8_t* const* data, int output_bpp,
                   int fixed_levels, int add_levels, int dither, int reverse,
                   int n_threads) {
  try {
    std::vector<multiblend::io::Image> images;
    images.reserve(n);
    for (int i = 0; i < n; ++i) {
      std::size_t bytes =
          (std::size_t)w[i] * h[i] * spp[i] * (bpp[i] == 16 ? 2 : 1);
      std::vector<uint8_t> v(data[i], data[i] + bytes);
      if (bpp[i] == 16) {
        // image.cpp:377 copies data.size() << 1 bytes for 16 bpp, i.e. the
        // library counts `data` in samples; shrink the logical size while the
        // allocation keeps all bytes so that copy reads valid memory.
        v.resize(bytes >> 1);
      }
      images.emplace_back(multiblend::io::InMemoryImage{
          .tiff_width = w[i],
          .tiff_height = h[i],
          .bpp = (uint16_t)bpp[i],
          .spp = (uint16_t)spp[i],
          .xpos_add = xpos[i],
          .ypos_add = ypos[i],
          .data = std::move(v)});
    }
    multiblend::mt::Threadpool pool(n_threads < 1 ? 1 : n_threads);
    multiblend::Options opts;
    opts.output_type = multiblend::io::ImageType::MB_IN_MEMORY;
    opts.output_bpp = output_bpp;
    opts.fixed_levels = fixed_levels;
    opts.add_levels = add_levels;
    opts.dither = dither != 0;
    opts.reverse = reverse != 0;
    opts.wideblend = g_wideblend != 0;
    auto result = multiblend::Multiblend(images, opts,
                                         multiblend::mt::ThreadpoolPtr{&pool});

    auto out = std::make_unique<RefResult>();
    out->width = result.width;
    out->height = result.height;
    out->bpp = result.output_bpp;
    out->min_x = result.min_xpos;
    out->min_y = result.min_ypos;
    out->no_mask = result.no_mask ? 1 : 0;
    std::size_t plane_bytes =
        (std::size_t)result.width * result.height * (result.output_bpp >> 3);
    for (int c = 0; c < 3; ++c) {
      const auto* p = (const uint8_t*)result.output_channels[c].get();
      out->planes[c].assign(p, p + plane_bytes);
    }
    // full_mask RLE -> dense 0/255 (same decoding as blenders.cc:38-62)
    out->mask.assign((std::size_t)result.width * result.height, 0);
    result.full_mask.Start();
    for (int y = 0; y < result.height; ++y) {
      int x = 0;
      while (x < result.width) {
        uint32_t v = result.full_mask.SafeReadForwards32();
        uint32_t len = v & 0x7fffffffu;
        if (len == 0 || x + (int)len > result.width) {
          throw std::runtime_error("bad full_mask RLE");
        }
        if (v & 0x80000000u) {
          memset(&out->mask[(std::size_t)y * result.width + x], 255, len);
        }
        x += (int)len;
      }
    }
    const auto& t = result.timing;
    double tt[10] = {t.images_time,  t.copy_time,     t.seam_time,
                     t.shrink_mask_time, t.shrink_time, t.laplace_time,
                     t.blend_time,   t.collapse_time, t.wrap_time,
                     t.out_time};
    memcpy(out->timing, tt, sizeof(tt));
    return out.release();
  } catch (const std::exception& e) {
    g_err = e.what();
    return nullptr;
  }
}

void mb_ref_info(void* hnd, int* out8) {
  auto* r = (RefResult*)hnd;
  out8[0] = r->width;
  out8[1] = r->height;
  out8[2] = r->bpp;
  out8[3] = r->min_x;
  out8[4] = r->min_y;
  out8[5] = r->no_mask;
}

void mb_ref_copy(void* hnd, uint8_t* p0, uint8_t* p1, uint8_t* p2,
                 uint8_t* mask, double* timing10) {
  auto* r = (RefResult*)hnd;
  uint8_t* dst[3] = {p0, p1, p2};
  for (int c = 0; c < 3; ++c) {
    if (dst[c]) memcpy(dst[c], r->planes[c].data(), r->planes[c].size());
  }
  if (mask) memcpy(mask, r->mask.data(), r->mask.size());
  if (timing10) memcpy(timing10, r->timing, sizeof(r->timing));
}

void mb_ref_free(void* hnd) { delete (RefResult*)hnd; }

}  // extern "C"
