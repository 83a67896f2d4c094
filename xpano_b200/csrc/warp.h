// Warper object behind the xp_warper_* C ABI (see warp.cu).
#pragma once

#include "common.h"
#include "projector.h"

namespace xp {

struct MapSrc;

struct Warper {
  int type = XP_PROJ_SPHERICAL, portrait = 0, device = 0;
  float scale = 1.f, a = 1.f, b = 1.f;
  cudaStream_t stream = nullptr, own_stream = nullptr;
  bool external_stream = false;
  bool device_ready = false;  // ensure_device() ran: the stream may hold work
  bool detached = false;      // the owner (composer) synchronised already
  DevBuf scratch, tables, aux_tables, staging_src, staging_dst, dilated,
      aux_gain,
      aux_seam;
  // Host-side trig tables of the separable projections (prepare_map): pinned,
  // mapped memory the table kernel reads; a slot per call until tab_reset().
  Mailbox tab_box;
  // the stream is idle: every slot of tab_box may be reused (with room for
  // `bytes` more)
  void tab_reset(size_t bytes);

  ~Warper();
  void ensure_device();
  Projector projector(const float* K, const float* R) const;
  // {tl.x, tl.y, br.x, br.y}, inclusive
  // border_ext: extent of the border walk if the caller already has it
  void detect_roi(const Projector& p, int w, int h, int out[4],
                  const float* border_ext = nullptr);
  static void border_extent(const Projector& p, int w, int h, int part,
                            int parts, float ext[4]);
  MapSrc prepare_map(const Projector& p, const int roi[4]);
  const void* stage_src(const void* src, int sw, int sh, size_t stride,
                        size_t row_bytes, int mem, size_t* dev_pitch);

  void warp_roi(int sw, int sh, const float* K, const float* R,
                int out_xywh[4]);
  void build_maps(int sw, int sh, const float* K, const float* R, float* xmap,
                  float* ymap, size_t stride, int mem, int out_xywh[4]);
  void warp(const void* src, int sw, int sh, size_t src_stride, int depth,
            int channels, const float* K, const float* R, int interp,
            int border, void* dst, size_t dst_stride, int mem,
            int out_corner[2]);
  void warp_bgra(const void* src, int sw, int sh, size_t src_stride, int depth,
                 const float* K, const float* R, const xp_compose_aux* aux,
                 int aux_mem, void* dst, size_t dst_stride, int mem,
                 int out_xywh[4]);
  // all-device form used by the composer; no synchronisation. dsrc == null:
  // only the alpha channel is produced (colours zero), see warp_alpha_kernel
  void warp_bgra_device(const void* dsrc, int sw, int sh, size_t spitch,
                        int depth, const Projector& p, const int roi[4],
                        const float* d_gain, int gw, int gh,
                        const uint8_t* d_seam, int mw, int mh, void* ddst,
                        size_t dpitch, bool skip_transparent = false,
                        uint32_t* abits = nullptr, int abits_pitch = 0);
  // by-product for the blender (see AlphaOut in warp.cu): words per row of
  // the opacity bit plane
  static int alpha_bits_pitch(int W) { return (W + 31) / 32; }
};

}  // namespace xp
