/* TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference blend algorithm.
 *
 * Plain-C restatement (dense arrays instead of the reference's RLE codecs) of
 * multiblend::Multiblend() as xpano calls it (in-memory images, spp 3 or 4,
 * 8/16 bit, no gamma, no wrap). Every function cites the reference file:line
 * it follows; paths are relative to /root/reference/external/multiblend.
 *
 * Parity pinning: tests/test_oracle_blend.py checks this restatement
 *  - bit-for-bit against oracle/_ref/libmb_ref.so (the unmodified reference
 *    compiled from its own sources) on the reference's fixtures and on seeded
 *    random inputs, and
 *  - against the reference's golden tests/data/expected.png.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
 * call into this file. The product path (xpano_b200) never links it.
 */
#ifndef MB_ORACLE_H_
#define MB_ORACLE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MBO_MAX_LEVELS 24

typedef struct {
  int width, height; /* untrimmed size                                  */
  int bpp;           /* 8 or 16                                         */
  int spp;           /* 3 (no alpha) or 4 (alpha last)                  */
  int xpos, ypos;    /* top-left on the canvas (blenders.cc:128-135)    */
  const void* data;  /* tightly packed interleaved samples              */
} mbo_image;

typedef struct {
  int output_bpp;   /* 0 = auto, 8, 16   (multiblend.cpp:100-117)       */
  int fixed_levels; /* 0 = auto          (multiblend.cpp:161-198)       */
  int add_levels;
  int dither;       /* default 1         (mb/multiblend.h:38)           */
  int reverse;
  int wideblend;    /* levels from the canvas size, plus one
                       (multiblend.cpp:164-193)                           */
} mbo_options;

/* Per-image state after trim + inpaint (image.cpp:396-868). */
typedef struct {
  int left, top;     /* trim offset inside the untrimmed image           */
  int width, height; /* trimmed size                                     */
  int xpos, ypos;    /* canvas position after trim and tighten           */
  uint8_t* opaque;   /* width*height, 1 where alpha == max before inpaint */
  void* chan[3];     /* planar channels after inpaint, u8 or u16          */
} mbo_trimmed;

typedef struct {
  int width, height, bpp;
  int min_xpos, min_ypos;
  int no_mask;
  int levels;
  void* planes[3];     /* width*height u8/u16                             */
  uint8_t* full_mask;  /* width*height 0/255 (>=1 image opaque)           */
  uint8_t* xor_mask;   /* width*height 0/1   (exactly 1 image opaque)     */
  int32_t* labels;     /* width*height seam owner                          */
  uint64_t* seam_bwd;  /* width*height backward DT values                  */
  uint64_t* seam_fwd;  /* width*height forward DT values                   */
  uint64_t channel_totals[3];
  uint64_t total_pixels;
  float* collapsed[3]; /* level-0 floats before Out, pitch = out_pitch     */
  int out_pitch;
  int n_images;
  mbo_trimmed* images;
} mbo_result;

/* Full pipeline. Returns 0 on success; on failure returns non-zero and
 * mbo_last_error() describes it. `keep_debug` keeps the intermediate arrays
 * (labels, seam DTs, collapsed floats, trimmed images) in the result. */
int mbo_blend(int n, const mbo_image* images, const mbo_options* opts,
              int keep_debug, mbo_result* out);
void mbo_free_result(mbo_result* r);
const char* mbo_last_error(void);

/* Stage-level entry points used by kernel parity tests. */

/* image.cpp:396-512: bbox of alpha==max; out = {left, top, width, height}.
 * Returns non-zero if no opaque pixel exists. */
int mbo_trim(const void* data, int width, int height, int bpp, int out[4]);

/* image.cpp:514-812 on the bbox given by trim; modifies `data` in place and
 * writes opaque[w*h] (1 where alpha==max before the fill). */
void mbo_inpaint(void* data, int tiff_width, int bpp, const int bbox[4],
                 uint8_t* opaque);

/* multiblend.cpp:254-823: both seam passes on dense opacity maps.
 * opaque[i] is w[i]*h[i]; labels/bwd/fwd are width*height (bwd/fwd may be
 * NULL). */
void mbo_seams(int n, const int* xpos, const int* ypos, const int* w,
               const int* h, const uint8_t* const* opaque, int width,
               int height, int reverse, int32_t* labels, uint64_t* bwd,
               uint64_t* fwd);

/* pyramid.cpp:21-98 level geometry. Arrays sized MBO_MAX_LEVELS.
 * Returns the number of levels. */
int mbo_pyramid_geometry(int width, int height, int levels, int x, int y,
                         int* lw, int* lh, int* lx, int* ly, int* xshift,
                         int* yshift, int* pitch);

/* multiblend.cpp:161-206 */
int mbo_blend_levels(int n, const int* widths, const int* heights,
                     int fixed_levels, int add_levels);

#ifdef __cplusplus
}
#endif
#endif /* MB_ORACLE_H_ */
