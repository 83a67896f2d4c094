// Internal declarations shared by the blend translation units.
#pragma once

#include <cstdint>
#include <vector>

#include "common.h"
#include "scan_problems.h"

namespace xp {

constexpr int kMaxLevels = 24;
constexpr int kMaxImages = 63;

// One fed image after trim, as the kernels see it.
struct ImgDev {
  const void* data;  // interleaved 4-sample pixels, untrimmed
  int tiff_w, tiff_h;
  int left, top, w, h;  // trim bbox
  int xpos, ypos;       // canvas position of the bbox after tighten
  // opacity bit plane of the untrimmed image (bit x & 31 of word x >> 5 of a
  // row of abits_pitch words), or null: what the composer's warps leave
  // beside the pixels, so that trim and the cover map need not read them
  const uint32_t* abits;
  int abits_pitch;
};

struct CanvasStats {
  unsigned long long totals[4];  // 3 channel sums + pixel count over xor px
  int uncovered;                 // some canvas pixel has no image
  int pad;
  int box[64][4];                // owner-region bbox per image (x0,y0,x1,y1)
};

// Pyramid level geometry (pyramid.cpp:21-98)
struct LevelGeom {
  int w, h, pitch, rows;
  int x, y, xs, ys;
  size_t offset;  // float offset of this level inside one channel
};

struct PyrGeom {
  int n = 0;
  LevelGeom lv[kMaxLevels];
  size_t channel_floats = 0;  // floats per channel (all levels)
};

PyrGeom make_pyramid(int width, int height, int levels, int x, int y);

// Dense mask pyramid level of one image, stored over a sub-rectangle of the
// canvas level grid (zero outside).
struct MaskLevel {
  int x0, y0, w, h;  // rect on the canvas level grid; w == 0: empty
  int cw, ch;        // canvas level size (functions.cpp:188-191)
  size_t offset;     // float offset in the image's mask buffer
};

// seams.cu
// box = (INT_MAX, INT_MAX, -1, -1), count = 0 for n images
void launch_trim_init(int* d_bbox, unsigned long long* d_opaque, int n,
                      cudaStream_t s);
// pitch: pixels per row of each image
void launch_trim_bits(const uint32_t* const* bits, const int* pitch,
                      const int* h, int n, int* d_bbox,
                      unsigned long long* d_opaque_count, cudaStream_t s);
void launch_trim(const void* const* data, const int* w, const int* h,
                 const int* pitch, int n, int bpp, int* d_bbox,
                 unsigned long long* d_opaque_count, cudaStream_t s);
// cover word of every canvas pixel in one pass (format chosen by n, see
// scan_problems.h Cov<>)
void launch_cover(void* cover, int cover_bytes, int W, int H,
                  const ImgDev* d_imgs, int n, int bpp, int reverse,
                  cudaStream_t s, bool with_bits = false);
void launch_seams(const void* cover, int cover_bytes, int n_images, int W,
                  int H, int reverse, uint32_t* bwd, uint32_t* fwd, uint8_t* labels,
                  int* counter, uint8_t* flags, cudaStream_t s,
                  cudaStream_t beside = nullptr, cudaEvent_t ev = nullptr);
// blocks (= SMs) the wavefront launches of the calling thread may use; 0: all
void set_wavefront_blocks(int n);
// bytes of tile flags launch_seams / launch_inpaint need
size_t wavefront_flag_bytes(int n_bands, int max_w);
void launch_seam_unpack(const uint32_t* in, size_t n, uint64_t* out,
                        cudaStream_t s);
void launch_inpaint(const InpaintJob* d_jobs, const InpaintBandRef* d_bands,
                    int n_bands, int max_w, int bpp, int* counter,
                    uint8_t* flags, cudaStream_t s);
void launch_canvas_stats(const void* cover, int cover_bytes,
                         const uint8_t* labels, int W, int H,
                         const ImgDev* d_imgs, int n, int bpp,
                         uint8_t* full_mask, uint8_t* xor_mask,
                         CanvasStats* d_stats, int mode, cudaStream_t s,
                         int sum_x0 = 0, int sum_x1 = 0x7fffffff);
// canvas-tiled runs: the window [wx0, wx0 + ww) of the label map with labels
// renumbered through `relabel` (64 entries), of the xor mask (zero outside the
// owned columns [ox0, ox1)) and of the full mask, as compact ww x H arrays
void launch_window_copy(const uint8_t* labels, const uint8_t* xor_mask,
                        const uint8_t* full_mask, int W, int H, int wx0,
                        int ww, int ox0, int ox1, const uint8_t* d_relabel,
                        uint8_t* wlabels, uint8_t* wxor, uint8_t* wfull,
                        cudaStream_t s);
void launch_stats_init(CanvasStats* d_stats, cudaStream_t s);

// pyramid.cu
// level 0 of an image pyramid from the (inpainted) source pixels.
// src_index: h x w u64 whose low 32 bits index `im.data`, or null (identity).
void launch_copy_level0(const ImgDev& im, int bpp, const u64* src_index,
                        float mul, const PyrGeom& g, float* pyr,
                        cudaStream_t s);
void launch_shrink(const PyrGeom& g, float* pyr, int level, int channels,
                   cudaStream_t s);
// hi <- expand(lo) - hi for `level` (Laplace and Collapse are the same op)
void launch_laplace(const PyrGeom& g, float* pyr, int level, int channels,
                    cudaStream_t s);
void launch_mask_level1(const uint8_t* labels, int CW, int CH, int image,
                        const MaskLevel& out, float* mask, cudaStream_t s);
// all images, one launch per level
size_t mask_batch_bytes(int n);
void launch_mask_pyramids(const uint8_t* labels, int CW, int CH, int n,
                          int n_levels,
                          const std::vector<std::vector<MaskLevel>>& levels,
                          const std::vector<float*>& masks, void* d_jobs,
                          Mailbox& mbox, cudaStream_t s);
void launch_mask_shrink(const MaskLevel& in, const MaskLevel& out, float* mask,
                        cudaStream_t s);
// out level (+)= in level * mask (functions.cpp:319-411)
void launch_composite(const PyrGeom& in_g, const float* in_pyr,
                      const PyrGeom& out_g, float* out_pyr, int level,
                      const uint8_t* labels, int image, const MaskLevel& m,
                      const float* mask, cudaStream_t s);
void launch_pad_level(const PyrGeom& g, float* pyr, int level, int channels,
                      cudaStream_t s);
// Σ level0 over xor pixels (double), per channel -> d_sums[3]
void launch_xor_sum(const PyrGeom& g, const float* pyr,
                    const uint8_t* xor_mask, int W, int H, double* d_partials,
                    double* d_sums, cudaStream_t s);
// last collapse step fused with the xor sums; d_partials needs
// collapse0_partials(g) doubles
size_t collapse0_partials(const PyrGeom& g);
void launch_collapse0_sum(const PyrGeom& g, float* pyr,
                          const uint8_t* xor_mask, int W, double* d_partials,
                          double* d_sums, cudaStream_t s);
// offset add (from device scalars) + dither + clamp + round + planar/BGR pack
void launch_out(const PyrGeom& g, const float* pyr, int W, int H, int out_bpp,
                int dither, int apply_add, const CanvasStats* d_stats,
                const double* d_sums, int in_bpp,
                const uint8_t* full_mask, float* d_offset /* 4 floats */,
                void* planes, void* bgr, cudaStream_t s);
// pyramid_fused.cu: streaming path (level 0 and the Laplacian stay virtual)
// gather of the inpaint fill over the unflagged tiles of the backward pass
// (`scratch` as passed to launch_inpaint)
void launch_inpaint_gather_tiles(const InpaintJob* d_jobs,
                                 const InpaintBandRef* d_bands, int n_bands,
                                 int max_w, int bpp, const uint8_t* scratch,
                                 cudaStream_t s);
// d_tmaps: 128 bytes of device scratch per image for the TMA tensor maps of
// the sources (written through `mbox`); null: plain loads only
void launch_shrink0_batch(const ImgDev* im, int bpp, float mul,
                          const PyrGeom* g, float* const* pyr, int n,
                          void* d_tmaps, Mailbox* mbox, cudaStream_t s);
void launch_shrink_batch(const PyrGeom* g, float* const* pyr, int n, int level,
                         int channels, cudaStream_t s);
void launch_shrink0(const ImgDev& im, int bpp, float mul, const PyrGeom& g,
                    float* pyr, cudaStream_t s);
void launch_lapcomp(const ImgDev& im, int bpp, float mul, const PyrGeom& in_g,
                    const float* in_pyr, const PyrGeom& out_g, float* out_pyr,
                    int level, const uint8_t* labels, int image,
                    const MaskLevel& m, const float* mask, cudaStream_t s);
// debug helpers
void launch_debug_collapsed(const PyrGeom& g, const float* pyr, int apply_add,
                            const CanvasStats* d_stats, const double* d_sums,
                            int in_bpp, int out_bpp, float* out,
                            cudaStream_t s);
void launch_extract_chan(const ImgDev& im, int bpp, const u64* src_index,
                         int channel, void* out, cudaStream_t s);
void launch_extract_opaque(const ImgDev& im, int bpp, uint8_t* out,
                           cudaStream_t s);
void launch_labels_to_i32(const uint8_t* labels, size_t n, int32_t* out,
                          cudaStream_t s);
// BGR (+mask) -> 4-sample interleaved (blenders.cc:99-112)
void launch_pack_bgra(const void* bgr, size_t bgr_pitch, const uint8_t* mask,
                      size_t mask_pitch, int w, int h, int depth, int spp_in,
                      void* out, int out_pitch, cudaStream_t s);

}  // namespace xp
