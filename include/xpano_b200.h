/* xpano_b200 -- C ABI of the B200-native full-resolution export hot path.
 *
 * Drop-in boundary: the two OpenCV interfaces xpano's
 * stitcher drives polymorphically,
 *   cv::detail::RotationWarper   (xpano/algorithm/stitcher.cc:160,377,384;
 *                                 creators xpano/algorithm/algorithm.cc:43-105,
 *                                 xpano/algorithm/warpers.h:9-31)
 *   cv::detail::Blender          (xpano/algorithm/blenders.h:25-39,
 *                                 xpano/algorithm/blenders.cc:116-158)
 * plus the compose loop that sits between them
 *   Stitcher::ComposePanorama    (xpano/algorithm/stitcher.cc:299-432).
 *
 * Plain pointers and sizes only; every call returns an xp_status, never
 * throws; xp_last_error() gives the message of the last failure on the
 * calling thread. All calls are synchronous at return (the reference API is),
 * except the explicitly split ones (xp_composer_run_begin / run_tile_begin,
 * completed by run_end / run_tile_end).
 * Pointers are host pointers unless the `mem` argument says XP_MEM_DEVICE.
 * There is no CPU fallback: without a CUDA device every compute entry point
 * fails with XP_ERR_CUDA.
 */
#ifndef XPANO_B200_H_
#define XPANO_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef int xp_status;
enum {
  XP_OK = 0,
  XP_ERR_INVALID = 1,     /* bad argument (CV_Assert in the reference)      */
  XP_ERR_CUDA = 2,        /* CUDA runtime failure / no device               */
  XP_ERR_UNSUPPORTED = 3, /* combination the reference path never uses      */
  XP_ERR_EMPTY = 4,       /* image without an opaque pixel / nothing fed    */
  XP_ERR_NOMEM = 5,
  XP_CANCELLED = 6        /* Status::kCancelled (stitcher.cc:408-412)       */
};

enum { XP_MEM_HOST = 0, XP_MEM_DEVICE = 1 };

/* xpano::algorithm::ProjectionType order (xpano/algorithm/options.h:16-26) */
enum {
  XP_PROJ_PERSPECTIVE = 0,
  XP_PROJ_CYLINDRICAL = 1,
  XP_PROJ_SPHERICAL = 2,
  XP_PROJ_FISHEYE = 3,
  XP_PROJ_STEREOGRAPHIC = 4,
  XP_PROJ_COMPRESSED_RECTILINEAR = 5,
  XP_PROJ_PANINI = 6,
  XP_PROJ_MERCATOR = 7,
  XP_PROJ_TRANSVERSE_MERCATOR = 8
};

/* cv::InterpolationFlags / cv::BorderTypes values used on the path */
enum { XP_INTER_NEAREST = 0, XP_INTER_LINEAR = 1 };
enum { XP_BORDER_CONSTANT = 0, XP_BORDER_REFLECT = 2 };

const char* xp_last_error(void);
const char* xp_version(void);
/* number of CUDA devices visible; 0 when there is none */
int xp_device_count(void);

/* ------------------------------------------------------------------------
 * Warper: replaces cv::WarperCreator::create(scale) ->
 * cv::Ptr<cv::detail::RotationWarper> (algorithm.cc:43-105). `portrait`
 * selects the *PortraitWarper of warpers.h:9-31 / PickWarperPortrait;
 * a, b are ProjectionOptions::a_param / b_param (options.h:82-86).
 * ---------------------------------------------------------------------- */
typedef struct xp_warper xp_warper;

xp_status xp_warper_create(int projection, int portrait, float scale, float a,
                           float b, int device, xp_warper** out);
void xp_warper_destroy(xp_warper* w);
/* RotationWarper::getScale / setScale */
float xp_warper_get_scale(const xp_warper* w);
xp_status xp_warper_set_scale(xp_warper* w, float scale);
/* Use the caller's CUDA stream (a cudaStream_t) instead of the warper's own */
xp_status xp_warper_set_stream(xp_warper* w, void* cuda_stream);

/* Rect warpRoi(Size src, K, R) (stitcher.cc:160): out = {x, y, w, h};
 * K, R are 3x3 row-major float32 (CV_32F, stitcher.cc:158,372,540). */
xp_status xp_warper_warp_roi(xp_warper* w, int src_w, int src_h,
                             const float K[9], const float R[9],
                             int out_xywh[4]);
/* Point2f warpPoint / warpPointBackward (gui/widgets/rotate.cc:94,130,196,372);
 * scalar host arithmetic, no device needed. */
xp_status xp_warper_warp_point(xp_warper* w, const float xy[2],
                               const float K[9], const float R[9],
                               float uv_out[2]);
xp_status xp_warper_warp_point_backward(xp_warper* w, const float uv[2],
                                        const float K[9], const float R[9],
                                        float xy_out[2]);
/* Rect buildMaps(Size, K, R, xmap, ymap): maps are roi_h x roi_w float32 with
 * the given strides in floats (0 = tightly packed). */
xp_status xp_warper_build_maps(xp_warper* w, int src_w, int src_h,
                               const float K[9], const float R[9], float* xmap,
                               float* ymap, size_t stride, int mem,
                               int out_xywh[4]);
/* Point warp(src, K, R, interp, border, dst) (stitcher.cc:377-385).
 * depth 8|16, channels 1|3, strides in bytes (0 = packed). dst must hold the
 * warp_roi rectangle. Supported (= what the path calls): INTER_LINEAR with
 * BORDER_REFLECT, INTER_NEAREST with BORDER_CONSTANT. */
xp_status xp_warper_warp(xp_warper* w, const void* src, int src_w, int src_h,
                         size_t src_stride, int depth, int channels,
                         const float K[9], const float R[9], int interp,
                         int border, void* dst, size_t dst_stride, int mem,
                         int out_corner[2]);

/* Allocation-free variant of xp_warper_warp. It has NO counterpart in the
 * reference: it was added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it). Same arithmetic and result as
 * xp_warper_warp, with these differences:
 *  - src and dst are DEVICE pointers (the caller's buffers; dst must hold the
 *    warp_roi rectangle); the work is queued on `cuda_stream` (a cudaStream_t,
 *    NULL = the default stream) and the call returns without synchronising;
 *  - no cudaMalloc / cudaFree / stream creation / pool allocation: the only
 *    scratch of the warp, the per-column and per-row trig table of a separable
 *    projection, lives in a static device array of
 *    XP_NOALLOC_MAX_ROI_PERIMETER float2 entries (512 KiB, held by every GPU
 *    that loads the library until exit). Its values -- width + height sines
 *    and cosines -- are formed on the host with the host's libm, as in
 *    xp_warper_warp (the maps are OpenCV's bit for bit that way), in a static
 *    host array, and reach the device array as one pageable
 *    cudaMemcpyAsync of at most 512 KiB on `cuda_stream` (no allocation; not
 *    capturable in a CUDA graph -- XP_HOST_TABLES=0 in the environment makes
 *    the device evaluate the table instead, last-bit differences in ~7 % of
 *    the map values);
 *  - limits: spherical and cylindrical projections only (their ROI is host
 *    arithmetic); result ROI width + height <= XP_NOALLOC_MAX_ROI_PERIMETER,
 *    larger ones fail with XP_ERR_INVALID and a message naming the limit;
 *  - every call on a device shares that array: calls that overlap on the GPU
 *    (different streams, or graph branches) race. Nothing in it needs to
 *    start at zero -- each call rewrites every entry it reads. */
#define XP_NOALLOC_MAX_ROI_PERIMETER 65536
xp_status xp_warper_warp_noalloc(xp_warper* w, const void* d_src, int src_w,
                                 int src_h, size_t src_stride, int depth,
                                 int channels, const float K[9],
                                 const float R[9], int interp, int border,
                                 void* d_dst, size_t dst_stride,
                                 void* cuda_stream, int out_corner[2]);

/* Fused form of stitcher.cc:377-405 + blenders.cc:99-112 for one image: warp
 * image (LINEAR/REFLECT) and all-255 mask (NEAREST/CONSTANT) from one map
 * evaluation, apply the exposure gain map (BlocksCompensator::apply: bilinear
 * upsample of gain, multiply, saturate), upsample the dilated seam mask
 * (dilate 3x3 + resize INTER_LINEAR_EXACT), AND, binarise, and pack
 * B,G,R,alpha. gain / seam_mask may be NULL (no compensation / all-255).
 * dst is roi_h x roi_w x 4 samples of `depth` bits. */
typedef struct {
  const float* gain;        /* gain_h x gain_w float32, or NULL             */
  int gain_w, gain_h;
  const uint8_t* seam_mask; /* seam_h x seam_w u8 (undilated), or NULL      */
  int seam_w, seam_h;
} xp_compose_aux;

xp_status xp_warper_warp_bgra(xp_warper* w, const void* src, int src_w,
                              int src_h, size_t src_stride, int depth,
                              const float K[9], const float R[9],
                              const xp_compose_aux* aux, int aux_mem, void* dst,
                              size_t dst_stride, int mem, int out_xywh[4]);

/* ------------------------------------------------------------------------
 * Blender: replaces xpano::algorithm::blenders::Multiblend
 * (blenders.h:25-39) and, underneath, multiblend::Multiblend()
 * (external/multiblend/src/multiblend.cpp:66-1307) with its Options
 * (external/multiblend/include/mb/multiblend.h:29-45).
 * ---------------------------------------------------------------------- */
typedef struct xp_blender xp_blender;

typedef struct {
  int output_bpp;   /* 0 = auto (8 unless inputs are 16), 8, 16             */
  int fixed_levels; /* 0 = auto                                             */
  int add_levels;
  int dither;       /* reference default: 1                                 */
  int reverse;
  int device;       /* CUDA device ordinal                                  */
  int keep_debug;   /* keep intermediates for xp_blender_debug_fetch        */
  int wideblend;    /* multiblend::Options::wideblend (mb/multiblend.h:33):
                       levels from the canvas size, plus one
                       (multiblend.cpp:164-193); xpano leaves it off          */
} xp_blend_opts;

/* multiblend::TimingResult (mb/multiblend.h:16-27), milliseconds on device.
 * Stages may overlap (side stream), so they need not add up to `total`;
 * `pyramid_tail` is the wall time from the end of the seam stage to the end
 * of `out` (mask pyramids, composite, collapse, out). */
typedef struct {
  double images, seam, shrink_mask, copy, shrink, laplace, blend, collapse,
      out, total, pyramid_tail;
} xp_timing;

void xp_blend_opts_default(xp_blend_opts* o);
xp_status xp_blender_create(const xp_blend_opts* opts, xp_blender** out);
void xp_blender_destroy(xp_blender* b);
xp_status xp_blender_set_stream(xp_blender* b, void* cuda_stream);
/* Blender::prepare(Rect dst_roi) (blenders.cc:116) */
xp_status xp_blender_prepare(xp_blender* b, int x, int y, int w, int h);
/* Blender::feed(img CV_8UC3|CV_16UC3, mask CV_8U, tl) (blenders.cc:120-139):
 * copies (uploads) the data; the caller keeps ownership. */
xp_status xp_blender_feed(xp_blender* b, const void* bgr, size_t stride,
                          const uint8_t* mask, size_t mask_stride, int w,
                          int h, int depth, int tl_x, int tl_y, int mem);
/* multiblend::io::InMemoryImage with spp 3 or 4 (mb/image.h:28-37): already
 * interleaved samples, alpha last when spp == 4. */
xp_status xp_blender_feed_interleaved(xp_blender* b, const void* data,
                                      size_t stride, int w, int h, int depth,
                                      int spp, int tl_x, int tl_y, int mem);
/* Blender::blend(dst, dst_mask) (blenders.cc:142-158). Two-step form: the
 * first call runs the blend on the device and reports the size (dst NULL is
 * allowed: size query); xp_blender_fetch copies the result out.
 * out_wh = {width, height}; out_min_xy = Result.min_xpos/min_ypos. */
xp_status xp_blender_blend(xp_blender* b, int out_wh[2], int out_min_xy[2],
                           int* out_bpp, int* out_no_mask, xp_timing* timing);
/* interleaved B,G,R (samples of output bpp) with pixels outside the mask set
 * to 0 (blenders.cc:151-154) and the 0/255 mask; either pointer may be NULL.
 * Strides in bytes (0 = packed). */
xp_status xp_blender_fetch(xp_blender* b, void* dst_bgr, size_t dst_stride,
                           uint8_t* dst_mask, size_t mask_stride, int mem);
/* Result.full_mask in the reference's own wire format (utils::Flex,
 * mb/include/mb/flex.h:22-75; decoded by xpano's ToMat, blenders.cc:38-62):
 * per row the maximal runs as 32-bit words (white << 31) | length, rows back
 * to back; row_offsets (H ints, may be NULL, host memory) = Flex::rows_, the
 * byte offset of every row. runs == NULL: size query (*n_words). The dense
 * dst_mask of xp_blender_fetch stays available; this form is ~100x smaller
 * on the way to the host. */
xp_status xp_blender_fetch_mask_rle(xp_blender* b, uint32_t* runs,
                                    size_t cap_words, size_t* n_words,
                                    int32_t* row_offsets, int mem);
/* Result.output_channels as the library returns them: 3 planes, not zeroed
 * outside the mask. */
xp_status xp_blender_fetch_planes(xp_blender* b, void* planes, int mem);

/* Algorithmic bytes (inputs once + outputs once) and kernel launches per
 * stage of the last blend, in xp_timing field order (9 entries). */
xp_status xp_blender_stage_bytes(xp_blender* b, double bytes[9],
                                 int launches[9]);

/* Intermediates for stage-level parity tests (needs keep_debug). `what`:
 *  "labels" (i32 HxW), "seam_bwd"/"seam_fwd" (u64 HxW), "xor_mask" (u8),
 *  "trim" (i32[6]: left, top, w, h, xpos, ypos; image), "opaque" (u8; image),
 *  "chan" (u8/u16; image, channel), "collapsed" (f32, pitch in dims[2];
 *  channel), "totals" (u64[4]: 3 channel totals + total_pixels),
 *  "mask" (f32; image, level), "lap" (f32; image, channel, level),
 *  "out" (f32; channel, level). dims = {rows, cols, pitch_elems}.
 * dst NULL = size query (dims and *bytes filled). */
xp_status xp_blender_debug_fetch(xp_blender* b, const char* what, int image,
                                 int channel, int level, void* dst,
                                 size_t cap, int dims[3], size_t* bytes);

/* Stage entry points (device-run) mirroring the oracle's, used by the
 * parity tests: pyramid geometry (pyramid.cpp:21-98) and level count
 * (multiblend.cpp:161-206) are host arithmetic. */
int xp_pyramid_geometry(int width, int height, int levels, int x, int y,
                        int* lw, int* lh, int* lx, int* ly, int* xshift,
                        int* yshift, int* pitch);
int xp_blend_levels(int n, const int* widths, const int* heights,
                    int fixed_levels, int add_levels);

/* ------------------------------------------------------------------------
 * Compose driver: Stitcher::ComposePanorama (stitcher.cc:299-432) with
 * ComputeRoi (:144-169) and the size cap (:318-335), cameras given
 * (registration is out of scope). Everything between the uploaded sources
 * and the downloaded panorama stays on the device.
 * ---------------------------------------------------------------------- */
typedef struct {
  const void* data; /* 8UC3 / 16UC3 interleaved                            */
  int w, h;
  size_t stride;    /* bytes, 0 = packed                                    */
  float K[9];       /* already scaled by 1/work_scale (utils/opencv.cc:14-26)*/
  float R[9];
  xp_compose_aux aux; /* host pointers                                      */
} xp_compose_image;

typedef struct {
  int projection, portrait;
  float a, b;
  float warp_scale;     /* warped_image_scale * compose_work_aspect, used
                           when warped_image_scale == 0                     */
  double max_pano_mpx;  /* StitchUserOptions::max_pano_mpx (constants.h:106);
                           0 = no cap                                       */
  /* the two doubles of stitcher.cc:306-313 (warped_image_scale_,
   * 1.0 / work_scale_): given separately, the capped scale is formed exactly
   * as the reference forms it (xp_compose_capped_scale)                    */
  double warped_image_scale, compose_work_aspect;
  int depth;            /* 8 or 16                                          */
  xp_blend_opts blend;
} xp_compose_opts;

typedef struct xp_composer xp_composer;
typedef struct {
  int width, height, min_x, min_y, bpp, levels, resolution_capped;
  int roi[4];           /* ComputeRoi rect                                  */
  float warp_scale;     /* after the cap                                    */
  double warp_ms, blend_ms, h2d_ms, d2h_ms;
  xp_timing timing;
  /* workload sizes for the roofline */
  double s_src, s_w, s_t, c;
  /* algorithmic bytes and kernel launches per stage, in xp_timing order
   * (images, seam, shrink_mask, copy, shrink, laplace, blend, collapse, out);
   * warp stage separately */
  double stage_bytes[9];
  int stage_launches[9];
  double warp_bytes;
  int warp_launches;
  /* canvas tiling: the whole canvas, the first column this rank owns
   * (width / min_x above then describe the owned strip) and the window
   * (strip + halo) it computed */
  int canvas_width, canvas_height, tile_x0, window_x0, window_width;
} xp_compose_info;

void xp_compose_opts_default(xp_compose_opts* o);
/* The size cap of stitcher.cc:318-335 as host arithmetic: pano_mpx in float
 * (utils/opencv.cc:35-38), compared with float(max_pano_mpx), float sqrt,
 * warped_image_scale (double) *= ratio, then *warp_scale =
 * float(warped_image_scale * compose_work_aspect). Returns 1 when capped. */
int xp_compose_capped_scale(double warped_image_scale,
                            double compose_work_aspect, int roi_w, int roi_h,
                            double max_pano_mpx, float* warp_scale);
xp_status xp_composer_create(const xp_compose_opts* opts, xp_composer** out);
void xp_composer_destroy(xp_composer* c);
xp_status xp_composer_set_stream(xp_composer* c, void* cuda_stream);
/* Upload (or adopt, mem == XP_MEM_DEVICE) the source images. */
xp_status xp_composer_set_images(xp_composer* c, int n,
                                 const xp_compose_image* images, int mem);
/* Cooperative cancellation, as Stitcher::Cancelled() is polled between images
 * (stitcher.cc:408-412, :345-347): `flag` is read before every image's warp
 * and once more before the blend; a non-zero value makes xp_composer_run (or
 * run_tile_begin) return XP_CANCELLED with nothing left queued. The blend
 * itself is not cancellable, as in the reference. NULL removes the flag. */
xp_status xp_composer_set_cancel_flag(xp_composer* c, const volatile int* flag);
/* Run warp -> feed -> blend on the device; result stays resident. */
xp_status xp_composer_run(xp_composer* c, xp_compose_info* info);
/* The same in two halves, for a batch export that keeps several panoramas
 * in flight on one GPU (stitcher_pipeline.h:48-52 hands the pipeline one
 * group after another): run_begin returns once the last kernel is queued --
 * after the blend's last host round trip, while its pyramid tail is still
 * running -- and run_end (or xp_composer_fetch) waits for it. Between the
 * two, another composer's run_begin may start on the same GPU: its warps and
 * seam passes overlap this panorama's bandwidth-bound tail. info may be
 * NULL. */
xp_status xp_composer_run_begin(xp_composer* c);
xp_status xp_composer_run_end(xp_composer* c, xp_compose_info* info);
/* Canvas tiling of ONE panorama over `count` GPUs (SURVEY.md section 8e;
 * no counterpart in the reference, whose ComposePanorama runs on one thread,
 * stitcher_pipeline.cc:340-344): every rank creates the same composer, calls
 * xp_composer_set_tile(index = rank, count = ranks) BEFORE set_images, and
 * then, instead of xp_composer_run,
 *   xp_composer_run_tile_begin(c, partial)   warp, seams (whole canvas, every
 *        rank) and the pyramids / composite / collapse of this rank's strip;
 *   all-reduce (sum) the 8 doubles of `partial` over the ranks -- the only
 *        exchange: 3 channel sums of the collapsed level, 3 input channel
 *        totals and the pixel count of the offset correction
 *        (multiblend.cpp:1232-1267), and a flag;
 *   xp_composer_run_tile_end(c, reduced, info)   offset + output pass.
 * The strips partition the canvas into vertical slices; xp_composer_fetch
 * returns this rank's slice (info.width x info.height at canvas column
 * info.tile_x0), bit-identical to the same columns of the single-GPU result.
 * xp_composer_tile_needs tells which sources the rank uploads (the others
 * only contribute their opacity). xp_tile_halo(levels): halo in pixels. */
xp_status xp_composer_set_tile(xp_composer* c, int index, int count);
xp_status xp_composer_tile_needs(xp_composer* c, int n, int* needed);
xp_status xp_composer_run_tile_begin(xp_composer* c, double partial[8]);
xp_status xp_composer_run_tile_end(xp_composer* c, const double reduced[8],
                                   xp_compose_info* info);
int xp_tile_halo(int levels);
/* Copy the panorama (interleaved BGR, zero outside mask) and mask out. */
xp_status xp_composer_fetch(xp_composer* c, void* dst_bgr, size_t dst_stride,
                            uint8_t* dst_mask, size_t mask_stride, int mem);
/* the blender a composer drives (for debug_fetch) */
xp_blender* xp_composer_blender(xp_composer* c);
/* ------------------------------------------------------------------------
 * Auto-crop of the panorama mask: crop::FindLargestCrop
 * (xpano/algorithm/auto_crop.cc:118-143), which consumes Blender::blend's
 * dst_mask (stitcher_pipeline.cc:374-403). out_rect = {start.x, start.y,
 * end.x, end.y} (utils::RectPPi); *found = 0 where the reference returns an
 * empty optional (empty matrix, no column with a set pixel at the seeds).
 * The per-column longest runs are found on the device; the seeded rectangle
 * growth (auto_crop.cc:65-113) is host arithmetic, exposed on its own as
 * xp_largest_crop_from_lines (lines = n_columns x {start, end}, {rows, 0}
 * for a column without a set pixel).
 * ---------------------------------------------------------------------- */
xp_status xp_find_largest_crop(const uint8_t* mask, int w, int h,
                               size_t stride, int mem, int device,
                               int out_rect[4], int* found);
/* the same on the mask of a blended blender, without a download */
xp_status xp_blender_auto_crop(xp_blender* b, int out_rect[4], int* found);
xp_status xp_largest_crop_from_lines(const int* lines_start_end, int n_columns,
                                     int rows, int out_rect[4], int* found);
/* The inputs of the auto-fill step (StitcherPipeline::RunInpainting,
 * stitcher_pipeline.cc:374-403): inpaint_mask = bitwise_not(pano_mask) and
 * pixels_filled = countNonZero(inpaint_mask), from the mask on the device.
 * dst_mask may be NULL (count only). The fill itself, cv::inpaint (Telea's
 * fast marching / Navier-Stokes, algorithm.cc:344-353), advances a
 * heap-ordered front pixel by pixel and stays the caller's CPU code. */
xp_status xp_blender_inpaint_mask(xp_blender* b, uint8_t* dst_mask,
                                  size_t mask_stride, int mem,
                                  long long* pixels_to_fill);

/* ------------------------------------------------------------------------
 * Exposure step of Stitcher::EstimateSeams (xpano/algorithm/stitcher.cc:
 * 273-278): cv::detail::BlocksGainCompensator as xpano constructs it
 * (stitcher.cc:200-201; OpenCV defaults: 32x32 blocks, one feed, two gain
 * filtering iterations, similarity threshold 1). SURVEY.md 8(f1).
 *
 *   feed   : BlocksCompensator::feed(corners, images_warped, masks_warped)
 *            -- seam-scale warped images (8-bit BGR interleaved) and masks
 *            (255 = inside), as xp_warper_warp produces them. Block-pair
 *            statistics on the device, the per-block system in double on
 *            the host; one float gain map per image.
 *   get_map: ExposureCompensator::getMatGains (the arrays
 *            xp_composer_set_images / xp_warp_aux.gain take).
 *   set_map: ExposureCompensator::setMatGains.
 *   apply  : BlocksCompensator::apply(index, corner, image, mask) in place
 *            on an interleaved BGR image (8 or 16 bit): gain map resized to
 *            the image (INTER_LINEAR), multiply, saturate.
 * The graph-cut seam finder that follows (stitcher.cc:290) is sequential
 * max-flow and stays with the caller's CPU code.
 * ---------------------------------------------------------------------- */
typedef struct xp_gain_compensator xp_gain_compensator;
xp_status xp_gain_compensator_create(int block_w, int block_h,
                                     int filter_iterations, int device,
                                     xp_gain_compensator** out);
void xp_gain_compensator_destroy(xp_gain_compensator* g);
/* corners_xy: n x {x, y}; strides in bytes (NULL or 0 = packed) */
xp_status xp_gain_compensator_feed(xp_gain_compensator* g, int n,
                                   const int* corners_xy,
                                   const void* const* bgr, const int* widths,
                                   const int* heights, const size_t* strides,
                                   const uint8_t* const* masks,
                                   const size_t* mask_strides, int mem);
/* feed in three steps, the middle one being the device kernel; host-only,
 * for callers that hold the block-pair statistics already and for the CPU
 * tests of the host arithmetic: plan (block grids, overlapping block pairs),
 * pairs (n_pairs x {block i, block j, x, y, w, h}), solve (gains and maps
 * from per-pair {count, sum |a|, sum |b|}). */
xp_status xp_gain_compensator_plan(xp_gain_compensator* g, int n,
                                   const int* corners_xy, const int* widths,
                                   const int* heights, int* n_pairs);
xp_status xp_gain_compensator_pairs(const xp_gain_compensator* g,
                                    int* pairs6);
xp_status xp_gain_compensator_solve(xp_gain_compensator* g, int n_pairs,
                                    const long long* counts,
                                    const double* sums_first,
                                    const double* sums_second);
xp_status xp_gain_compensator_map_size(const xp_gain_compensator* g,
                                       int index, int wh[2]);
/* dst: host, wh[1] x wh[0] float32 */
xp_status xp_gain_compensator_get_map(const xp_gain_compensator* g, int index,
                                      float* dst);
xp_status xp_gain_compensator_set_map(xp_gain_compensator* g, int index,
                                      const float* src, int w, int h);
xp_status xp_gain_compensator_apply(const xp_gain_compensator* g, int index,
                                    void* bgr, int w, int h, size_t stride,
                                    int depth, int mem);
/* out = {block pairs, unknowns of the system, solver iterations} of the
 * last feed */
xp_status xp_gain_compensator_stats(const xp_gain_compensator* g, int out[3]);

/* ------------------------------------------------------------------------
 * xpano's second blender, blenders::MultiBandOpenCV (xpano/algorithm/
 * blenders.h:41-50, blenders.cc:160-175; BlendingMethod::kOpenCV,
 * algorithm.cc:135-136): cv::detail::MultiBandBlender(false, num_bands = 5,
 * CV_32F) fed with CV_16SC3 images, result converted to CV_8U.
 * SURVEY.md 8(f4).
 *   prepare: MultiBandBlender::prepare(dst_roi)      (stitcher.cc:357)
 *   feed   : MultiBandOpenCV::feed(img CV_8UC3, mask CV_8U, tl) -- the
 *            conversion to CV_16S of blenders.cc:164 included; mask / 255 is
 *            the weight                                (stitcher.cc:405)
 *   blend  : MultiBandOpenCV::blend                  (stitcher.cc:419)
 *   fetch  : dst (h x w x 3 u8, zero outside the mask) and dst_mask
 *            (255 where the accumulated weight exceeds 1e-5)
 * Integer pyramids as OpenCV's, see csrc/mbcv.cu. strides in bytes, 0 =
 * packed.
 * ---------------------------------------------------------------------- */
typedef struct xp_multiband xp_multiband;
xp_status xp_multiband_create(int num_bands, int device, xp_multiband** out);
void xp_multiband_destroy(xp_multiband* b);
xp_status xp_multiband_prepare(xp_multiband* b, int x, int y, int w, int h);
xp_status xp_multiband_feed(xp_multiband* b, const void* bgr, size_t stride,
                            const uint8_t* mask, size_t mask_stride, int w,
                            int h, int tl_x, int tl_y, int mem);
/* num_bands: the band count prepare() settled on (may be NULL) */
xp_status xp_multiband_blend(xp_multiband* b, int out_wh[2], int* num_bands);
xp_status xp_multiband_fetch(xp_multiband* b, void* dst_bgr, size_t dst_stride,
                             uint8_t* dst_mask, size_t mask_stride, int mem);

/* ------------------------------------------------------------------------
 * JPEG either side of the path: cv::imread of the sources
 * (xpano/algorithm/image.cc:83) and cv::imwrite of the panorama
 * (xpano/pipeline/stitcher_pipeline.cc:94-116; parameters as
 * CompressionParameters, :40-65: quality 95, baseline, 4:2:2 by default).
 * Done on the device, so that only compressed bytes cross PCIe. Decoding:
 * baseline / sequential Huffman streams (8-bit, one interleaved scan, grey or
 * YCbCr 4:4:4 / 4:2:2 / 4:2:0, with or without restart intervals -- what
 * cameras write) are decoded by this library's own kernels (csrc/jpeg_sm.cu:
 * parallel Huffman decoding, libjpeg-turbo's inverse DCT, fancy upsampling
 * and colour conversion restated), and the pixels equal cv::imread's bit for
 * bit. Encoding: baseline streams (the standard Huffman tables that
 * CompressionParameters' defaults ask for, or tables optimised for the image)
 * are written by this library's own kernels too (csrc/jpeg_enc.cu:
 * libjpeg-turbo's colour conversion, downsampling, forward DCT, quantisation
 * and Huffman coding restated), and the stream equals cv::imwrite's byte for
 * byte. Everything else (progressive, ...) goes to nvJPEG (library code,
 * linked statically),
 * whose decoded pixels / streams are close to libjpeg-turbo's, not identical.
 * XP_JPEG_SM=0 in the environment sends everything to nvJPEG, =2 makes a
 * stream or setting the kernels do not cover an error.
 * Pixels are interleaved 8-bit BGR, oriented by EXIF, as cv::imread returns.
 *   info   : width, height, channel count of a JPEG stream (host bytes)
 *   decode : -> dst_bgr (mem = XP_MEM_DEVICE: e.g. straight into the buffer
 *            xp_composer_set_images adopts; XP_MEM_HOST: downloaded)
 *   encode : bgr (host or device) -> JPEG bytes on the host. out == NULL
 *            encodes, keeps the stream and returns its size in *len; the next
 *            call with a buffer retrieves it. subsampling: 444, 422 or 420
 *   xp_blender_encode_jpeg: the same from the blended panorama on the device,
 *            optionally cropped (utils::GetCvRect, stitcher_pipeline.cc:
 *            99-102); crop_xywh may be NULL
 * ---------------------------------------------------------------------- */
typedef struct xp_codec xp_codec;
xp_status xp_codec_create(int device, xp_codec** out);
void xp_codec_destroy(xp_codec* c);
xp_status xp_codec_jpeg_info(xp_codec* c, const uint8_t* data, size_t len,
                             int wh[2], int* channels);
/* out = {images decoded by the library's own kernels, images decoded by
 * nvJPEG, images encoded by the library's own kernels, images encoded by
 * nvJPEG} since the codec was created */
xp_status xp_codec_stats(xp_codec* c, int out[4]);
xp_status xp_codec_jpeg_decode(xp_codec* c, const uint8_t* data, size_t len,
                               void* dst_bgr, size_t stride, int mem);
/* n streams into n device buffers at once, several host threads each with
 * its own stream and decoder state (nvJPEG, where it is used, decodes the
 * Huffman stage on the calling host thread): the sources of a panorama */
xp_status xp_codec_jpeg_decode_batch(xp_codec* c, int n,
                                     const uint8_t* const* data,
                                     const size_t* lens,
                                     void* const* dst_bgr_device,
                                     const size_t* strides);
xp_status xp_codec_jpeg_encode(xp_codec* c, const void* bgr, int w, int h,
                               size_t stride, int mem, int quality,
                               int subsampling, int progressive, int optimize,
                               uint8_t* out, size_t cap, size_t* len);
xp_status xp_blender_encode_jpeg(xp_blender* b, xp_codec* c,
                                 const int crop_xywh[4], int quality,
                                 int subsampling, int progressive,
                                 int optimize, uint8_t* out, size_t cap,
                                 size_t* len);

/* Device buffers are cached across objects (a blender is single-use, like the
 * reference's); this returns the cached blocks to the driver. */
void xp_trim_memory(void);
/* number of kernels this library launched since process start */
unsigned long long xp_kernel_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* XPANO_B200_H_ */
