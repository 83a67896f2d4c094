// Blender object behind the xp_blender_* C ABI (see blend.cu).
#pragma once

#include <memory>
#include <string>
#include <vector>

#include "blend_impl.h"

namespace xp {

struct Fed {
  int w = 0, h = 0, depth = 8, xpos = 0, ypos = 0;
  // Row pitch in pixels: w rounded up to a multiple of 4, so that rows start
  // on 16-byte boundaries (what a TMA tensor map over the pixels needs).
  int pitch = 0;
  DevBuf data;  // pitch*h pixels of 4 samples (B, G, R, alpha)
  // by-product of the composer's warp (warp.cu AlphaOut): the opacity bit
  // plane, h rows of abits_pitch words; empty for images that came through
  // feed()
  DevBuf abits;
  int abits_pitch = 0;
};

struct StageTimer {
  struct Span {
    int stage;
    cudaEvent_t a, b;
    cudaStream_t s;
  };
  std::vector<Span> spans;
  std::vector<cudaEvent_t> pool;
  size_t used = 0;
  int device = -1;  // the device the events belong to
  cudaStream_t s = nullptr;
  ~StageTimer();
  cudaEvent_t get();
  void reset(cudaStream_t stream);
  int begin(int stage, cudaStream_t on = nullptr);
  void end(int idx);
  void collect(double out[16]);
};

struct Blender {
  xp_blend_opts opts;
  cudaStream_t stream = nullptr, own_stream = nullptr;
  cudaStream_t side_stream = nullptr;  // inpaint runs beside the seam passes
  cudaStream_t aux_stream = nullptr;   // cover-only statistics
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_cover = nullptr,
              ev_aux = nullptr, ev_cls = nullptr, ev_seam = nullptr;
  // SM partition (common.h StreamSet): the seam chain on its own SMs, the work
  // beside it on the rest; null when the device is not partitioned
  cudaStream_t seam_stream = nullptr, part_side = nullptr, part_aux = nullptr;
  int seam_sms = 0, side_sms = 0;
  bool external_stream = false;
  bool device_ready = false;  // ensure_device() ran: streams may hold work
  int roi[4] = {0, 0, 0, 0};
  std::vector<std::unique_ptr<Fed>> fed;
  DevBuf staging, staging_mask;

  // ---- results of blend() ----
  bool blended = false;
  int n = 0, in_bpp = 8, out_bpp = 8;
  int W = 0, H = 0, min_x = 0, min_y = 0, no_mask = 0, levels = 0;
  std::vector<ImgDev> im;
  CanvasStats stats;
  std::vector<unsigned long long> transparent;
  PyrGeom out_g;
  DevBuf d_bbox, d_transparent, d_imgs, d_stats, d_jobs, d_bands, d_counter,
      d_progress, d_flags_inp, d_partials, d_sums, d_mask_jobs, d_tmaps;
  DevBuf cover, seam_bwd, seam_fwd, labels, full_mask, xor_mask;
  std::vector<DevBuf> inp_fwd, inp_bwd;  // per image (empty if no holes)
  DevBuf img_pyr, mask_buf, out_pyr, planes, bgr;
  std::vector<DevBuf> mask_bufs;     // streaming path: mask pyramid per image
  std::vector<DevBuf> img_pyrs;      // streaming path: levels >= 1 per image
  std::vector<PyrGeom> img_geoms;    // ... their geometry (compact stride)
  std::vector<float*> img_bases;     // ... base pointers for level offsets
  int cover_bytes = 4;
  int apply_add = 0;
  double stage_bytes[16] = {0};
  int stage_launches[16] = {0};
  StageTimer timer;
  Mailbox mbox;
  // workload sizes (pixels)
  double s_t = 0, c_px = 0;

  // ---- debug copies (keep_debug) ----
  std::vector<DevBuf> dbg_lap, dbg_mask;
  std::vector<PyrGeom> dbg_lap_g;
  std::vector<std::vector<MaskLevel>> dbg_mask_lv;
  DevBuf dbg_out, dbg_collapsed;

  // ---- canvas tiling (blend_tile.inc): this blender computes one strip ----
  struct Tile {
    int index = 0, count = 1;
    int abs_x0 = 0, abs_x1 = 0;  // owned columns in feed (ROI) coordinates
  } tile;
  std::vector<char> fed_local;   // per fed image: its colours are present
  bool tile_pending = false;     // between blend_tile_begin and _end
  int own_x0 = 0, own_x1 = 0;    // owned canvas columns
  int win_x0 = 0, win_w = 0;     // window (strip + halo) on the canvas
  int out_w = 0;                 // row length of planes / bgr (W, or win_w)
  std::vector<int> win_img;      // window image -> fed index
  DevBuf wlabels, wxor, wfull, d_relabel;

  // run-length form of the mask (crop.cu), built on first request
  bool rle_ready = false;
  size_t rle_words = 0;
  DevBuf rle_runs, rle_offs;

  explicit Blender(const xp_blend_opts& o) : opts(o) {}
  ~Blender();
  void ensure_device();
  Fed* new_fed(int w, int h, int depth, int tl_x, int tl_y);
  void feed(const void* data, size_t stride, const uint8_t* mask,
            size_t mask_stride, int w, int h, int depth, int spp, int tl_x,
            int tl_y, int mem);
  void blend(xp_timing* timing);        // blend_launch + blend_finish
  void blend_launch();
  void blend_finish(xp_timing* timing);
  bool launch_pending = false;
  void blend_tile_begin(double partial[8]);
  void blend_tile_end(const double reduced[8], xp_timing* timing);
  void fetch(void* dst_bgr, size_t dst_stride, uint8_t* dst_mask,
             size_t mask_stride, int mem);
  void fetch_planes(void* planes_out, int mem);
  void debug_fetch(const char* what, int image, int channel, int level,
                   void* dst, size_t cap, int dims[3], size_t* bytes);
  const u64* src_index(int i) const {
    return inp_bwd.size() > (size_t)i && inp_bwd[i].p ? inp_bwd[i].as<u64>()
                                                      : nullptr;
  }
};

int blend_levels(int n, const int* widths, const int* heights,
                 int fixed_levels, int add_levels);
// halo of a canvas-tiled run, in canvas pixels, on either side of a strip
int tile_halo(int levels);

}  // namespace xp
