/* TEST INFRASTRUCTURE ONLY -- see mb_oracle.h. CPU restatement of the
 * reference blend algorithm (external/multiblend @ 8cb2c04) on dense arrays.
 * Build: gcc -O2 -std=c99 -ffp-contract=off -fPIC -shared (oracle/Makefile).
 * -ffp-contract=off matters: the reference is SSE4.1 without FMA, so every
 * multiply and add below rounds separately, in the order written.
 */
#include "mb_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static char g_err[256];
const char* mbo_last_error(void) { return g_err; }

#define FAIL(...)                                 \
  do {                                            \
    snprintf(g_err, sizeof(g_err), __VA_ARGS__);  \
    return 1;                                     \
  } while (0)

/* ------------------------------------------------------------------------
 * Trim: image.cpp:396-512
 * ---------------------------------------------------------------------- */
static inline int is_opaque(const void* data, size_t idx, int bpp) {
  if (bpp == 8) return ((const uint8_t*)data)[idx * 4 + 3] == 0xff;
  return ((const uint16_t*)data)[idx * 4 + 3] == 0xffff;
}

int mbo_trim(const void* data, int width, int height, int bpp, int out[4]) {
  int left = -1, top = -1, right = -1, bottom = -1;
  size_t n = (size_t)width * height;
  size_t i;
  /* first opaque pixel in raster order gives (left, top): image.cpp:402-411 */
  for (i = 0; i < n; ++i) {
    if (is_opaque(data, i, bpp)) {
      left = (int)(i % width);
      top = (int)(i / width);
      break;
    }
  }
  if (left < 0) return 1;
  /* last opaque pixel gives (right, bottom): image.cpp:413-424 */
  for (i = n; i-- > 0;) {
    if (is_opaque(data, i, bpp)) {
      right = (int)(i % width);
      bottom = (int)(i / width);
      break;
    }
  }
  if (right < left) { /* image.cpp:426-428 */
    int t = left;
    left = right;
    right = t;
  }
  /* widen left/right over rows top..bottom: image.cpp:430-447 */
  for (int y = top; y <= bottom; ++y) {
    size_t row = (size_t)y * width;
    for (int x = 0; x < left; ++x) {
      if (is_opaque(data, row + x, bpp)) {
        left = x;
        break;
      }
    }
    for (int x = width - 1; x > right; --x) {
      if (is_opaque(data, row + x, bpp)) {
        right = x;
        break;
      }
    }
  }
  out[0] = left;
  out[1] = top;
  out[2] = right + 1 - left;
  out[3] = bottom + 1 - top;
  return 0;
}

/* ------------------------------------------------------------------------
 * Inpaint: image.cpp:514-812. Two raster passes of a 3-4 chamfer distance
 * (u32) whose winner's whole pixel is copied in place. Tie-breaking follows
 * the order of the strict '<' tests in the reference.
 * ---------------------------------------------------------------------- */
static inline void copy_px(void* data, int bpp, ptrdiff_t dst, ptrdiff_t src) {
  if (bpp == 8)
    ((uint32_t*)data)[dst] = ((uint32_t*)data)[src];
  else
    ((uint64_t*)data)[dst] = ((uint64_t*)data)[src];
}

void mbo_inpaint(void* data, int tiff_width, int bpp, const int bbox[4],
                 uint8_t* opaque) {
  const int left = bbox[0], top = bbox[1], w = bbox[2], h = bbox[3];
  const ptrdiff_t tw = tiff_width;
  uint32_t* dist = (uint32_t*)malloc((size_t)w * h * sizeof(uint32_t));
  /* record opacity before anything is overwritten (tiff_mask_,
   * image.cpp:652-676) */
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x)
      opaque[(size_t)y * w + x] =
          (uint8_t)is_opaque(data, (size_t)(top + y) * tw + left + x, bpp);

  /* forward pass: image.cpp:550-694 */
  for (int y = 0; y < h; ++y) {
    const ptrdiff_t row = (ptrdiff_t)(top + y) * tw + left;
    uint32_t* cur = dist + (size_t)y * w;
    const uint32_t* prev = cur - w;
    const uint8_t* op = opaque + (size_t)y * w;
    for (int x = 0; x < w; ++x) {
      if (op[x]) {
        cur[x] = 0; /* :659-662 */
        continue;
      }
      if (y == 0) { /* :573-583 */
        if (x == 0) {
          cur[0] = 0x80000000u;
        } else {
          cur[x] = cur[x - 1] + 3;
          copy_px(data, bpp, row + x, row + x - 1);
        }
      } else if (x == 0) { /* :585-603, tie goes to up-right */
        uint32_t b = prev[0] + 3;
        uint32_t c = prev[1] + 4;
        if (b < c) {
          cur[0] = b;
          copy_px(data, bpp, row, row - tw);
        } else {
          cur[0] = c;
          copy_px(data, bpp, row, row - tw + 1);
        }
      } else { /* :609-641 */
        uint32_t d = cur[x - 1] + 3;
        uint32_t a = prev[x - 1] + 4;
        uint32_t b = prev[x] + 3;
        uint32_t c = (x < w - 1) ? prev[x + 1] + 4 : 0xffffffffu;
        ptrdiff_t src = row + x - 1;
        if (a < d) {
          d = a;
          src = row + x - tw - 1;
        }
        if (b < d) {
          d = b;
          src = row + x - tw;
        }
        if (c < d) {
          d = c;
          src = row + x - tw + 1;
        }
        cur[x] = d;
        copy_px(data, bpp, row + x, src);
      }
    }
  }

  /* backward pass, last line: image.cpp:711-747 */
  {
    const int y = h - 1;
    const ptrdiff_t row = (ptrdiff_t)(top + y) * tw + left;
    uint32_t* cur = dist + (size_t)y * w;
    const uint8_t* op = opaque + (size_t)y * w;
    uint32_t d = 0;
    for (int x = w - 1; x >= 0; --x) {
      if (op[x]) {
        d = 3;
      } else if (x == w - 1) {
        d = cur[x] + 3;
      } else {
        uint32_t best = cur[x];
        if (d < best) {
          cur[x] = d;
          copy_px(data, bpp, row + x, row + x + 1);
          d += 3;
        } else {
          d = best + 3;
        }
      }
    }
  }
  /* backward pass, other lines: image.cpp:752-812 */
  for (int y = h - 2; y >= 0; --y) {
    const ptrdiff_t row = (ptrdiff_t)(top + y) * tw + left;
    uint32_t* cur = dist + (size_t)y * w;
    const uint32_t* prev = cur + w; /* row y+1, already final */
    const uint8_t* op = opaque + (size_t)y * w;
    uint32_t d = 0x80000000u;
    for (int x = w - 1; x >= 0; --x) {
      if (op[x]) {
        cur[x] = 0;
        d = 3;
        continue;
      }
      uint32_t a = (x > 0) ? prev[x - 1] + 4 : 0x80000000u;
      uint32_t b = prev[x] + 3;
      uint32_t c = (x < w - 1) ? prev[x + 1] + 4 : 0x80000000u;
      uint32_t best = cur[x]; /* forward value (ReadInpaintDT) */
      ptrdiff_t src = -1;
      if (a < best) {
        best = a;
        src = row + tw + x - 1;
      }
      if (b < best) {
        best = b;
        src = row + tw + x;
      }
      if (c < best) {
        best = c;
        src = row + tw + x + 1;
      }
      if (d < best) {
        best = d;
        src = row + x + 1;
      }
      if (src >= 0) copy_px(data, bpp, row + x, src);
      cur[x] = best;
      d = best + 3;
    }
  }
  free(dist);
}

/* ------------------------------------------------------------------------
 * Seams: multiblend.cpp:254-823. State per canvas pixel is
 * u64 = dist << 32 | image id, top bit = "that image does not cover this
 * pixel" (maskval, multiblend.cpp:74-76). The reference walks runs of
 * constant mask state and updates neighbours incrementally inside a run;
 * because the state is constant inside a run that is identical to
 * evaluating maskval() on the raw neighbours per pixel, which is what this
 * dense restatement does. The one place where run structure is visible is
 * the first run of the bottom row (multiblend.cpp:333-337).
 * ---------------------------------------------------------------------- */
#define TOPBIT 0x8000000000000000ull
#define LOW63 0x7fffffffffffffffull
#define DT_MAX 0x9000000000000000ull
#define U32(x) ((uint64_t)(x) << 32)

typedef struct {
  int n;
  const int *xpos, *ypos, *w, *h;
  const uint8_t* const* opaque;
} seam_ctx;

static inline int covers(const seam_ctx* s, int i, int x, int y) {
  int lx = x - s->xpos[i], ly = y - s->ypos[i];
  if (lx < 0 || ly < 0 || lx >= s->w[i] || ly >= s->h[i]) return 0;
  return s->opaque[i][(size_t)ly * s->w[i] + lx];
}

/* cover bitmask for one canvas row; images >= 64 use extra words */
/* Also marks run starts (rs[x] = 1 when a new run of the reference's RLE
 * walk begins at x, multiblend.cpp:509-534): for every image on this row a
 * run starts at its bbox edges and wherever its opacity changes. */
static void cover_row(const seam_ctx* s, int y, int width, uint64_t* bits,
                      int words, uint8_t* rs) {
  memset(bits, 0, (size_t)width * words * sizeof(uint64_t));
  if (rs) {
    memset(rs, 0, (size_t)width);
    rs[0] = 1;
  }
  for (int i = 0; i < s->n; ++i) {
    int ly = y - s->ypos[i];
    if (ly < 0 || ly >= s->h[i]) continue;
    const uint8_t* op = s->opaque[i] + (size_t)ly * s->w[i];
    for (int lx = 0; lx < s->w[i]; ++lx) {
      if (op[lx])
        bits[(size_t)(s->xpos[i] + lx) * words + (i >> 6)] |= 1ull << (i & 63);
      if (rs && (lx == 0 || op[lx] != op[lx - 1])) rs[s->xpos[i] + lx] = 1;
    }
    if (rs && s->xpos[i] + s->w[i] < width) rs[s->xpos[i] + s->w[i]] = 1;
  }
}

static inline int cover_count(const uint64_t* b, int words, int* which) {
  int cnt = 0;
  for (int k = 0; k < words; ++k) {
    uint64_t v = b[k];
    if (v) {
      if (cnt == 0) *which = k * 64 + __builtin_ctzll(v);
      cnt += __builtin_popcountll(v);
    }
  }
  return cnt;
}

static inline uint64_t maskval(uint64_t v, const uint64_t* b) {
  uint32_t lab = (uint32_t)v;
  uint64_t covered = (b[lab >> 6] >> (lab & 63)) & 1;
  return (v & LOW63) | (covered ? 0 : TOPBIT);
}

/* length of the first run (from the right) of the bottom row,
 * multiblend.cpp:276-325 evaluated at x = width-1 */
static int bottom_first_run(const seam_ctx* s, int width, int height) {
  int y = height - 1;
  int min_count = width;
  for (int i = 0; i < s->n; ++i) {
    int ly = y - s->ypos[i];
    int cnt;
    if (ly < 0 || ly >= s->h[i]) {
      cnt = width;
    } else {
      cnt = width - (s->xpos[i] + s->w[i]);
      if (cnt == 0) {
        /* last RLE run of this image row (tiff_mask_, image.cpp:652-676) */
        const uint8_t* op = s->opaque[i] + (size_t)ly * s->w[i];
        int x = s->w[i] - 1;
        uint8_t v = op[x];
        cnt = 0;
        while (x >= 0 && op[x] == v) {
          --x;
          ++cnt;
        }
      }
    }
    if (cnt < min_count) min_count = cnt;
  }
  return min_count;
}

void mbo_seams(int n, const int* xpos, const int* ypos, const int* w,
               const int* h, const uint8_t* const* opaque, int width,
               int height, int reverse, int32_t* labels, uint64_t* bwd_out,
               uint64_t* fwd_out) {
  seam_ctx s = {n, xpos, ypos, w, h, opaque};
  const int words = (n + 63) / 64;
  uint64_t* bits = (uint64_t*)malloc((size_t)width * words * sizeof(uint64_t));
  uint64_t* bwd = (uint64_t*)malloc((size_t)width * height * sizeof(uint64_t));
  uint64_t* fprev = (uint64_t*)malloc((size_t)width * sizeof(uint64_t));
  uint64_t* fcur = (uint64_t*)malloc((size_t)width * sizeof(uint64_t));
  uint8_t* rs = (uint8_t*)malloc((size_t)width);
  int which = 0;

  /* backward pass: multiblend.cpp:271-438 */
  const int first_run = bottom_first_run(&s, width, height);
  for (int y = height - 1; y >= 0; --y) {
    uint64_t* cur = bwd + (size_t)y * width;
    const uint64_t* prev = cur + width;
    cover_row(&s, y, width, bits, words, NULL);
    for (int x = width - 1; x >= 0; --x) {
      const uint64_t* b = bits + (size_t)x * words;
      if (cover_count(b, words, &which) == 1) {
        cur[x] = (uint64_t)which; /* :327-331 */
        continue;
      }
      if (y == height - 1) { /* bottom row :333-346 */
        if (x > width - 1 - first_run) {
          cur[x] = DT_MAX;
        } else {
          cur[x] = maskval(cur[x + 1], b) + U32(3);
        }
        continue;
      }
      uint64_t d = ~0ull, t;
      if (x < width - 1) {
        d = maskval(cur[x + 1] + U32(3), b);
        t = maskval(prev[x + 1] + U32(4), b);
        if (t < d) d = t;
      }
      if (x > 0) {
        t = maskval(prev[x - 1] + U32(4), b);
        if (t < d) d = t;
      }
      t = maskval(prev[x] + U32(3), b);
      if (t < d) d = t;
      cur[x] = d;
    }
  }

  /* forward pass: multiblend.cpp:493-815 */
  for (int y = 0; y < height; ++y) {
    const uint64_t* bw = bwd + (size_t)y * width;
    cover_row(&s, y, width, bits, words, rs);
    for (int x = 0; x < width; ++x) {
      const uint64_t* b = bits + (size_t)x * words;
      int cnt = cover_count(b, words, &which);
      uint64_t best;
      if (cnt == 1) {
        best = (uint64_t)which; /* :607-616 */
      } else {
        uint64_t t;
        best = bw[x]; /* ReadSeamDT / this_line of row 0, unmasked */
        if (y > 0) {
          if (x > 0) {
            t = maskval(fprev[x - 1] + U32(4), b);
            if (t < best) best = t;
          }
          t = maskval(fprev[x] + U32(3), b);
          if (t < best) best = t;
          if (x < width - 1) {
            t = maskval(fprev[x + 1] + U32(4), b);
            if (t < best) best = t;
          }
        }
        if (x > 0) {
          /* d is re-masked only on the top row, at x == 1 and at a run
           * start (:629-630, :704-705); inside a run it is best + 3 raw
           * (:750), which keeps the top bit of an "arbitrary seam" value. */
          t = fcur[x - 1] + U32(3);
          if (y == 0 || x == 1 || rs[x]) t = maskval(t, b);
          if (t < best) best = t;
        }
        if ((best & TOPBIT) && cnt != 0) { /* arbitrary seam :732-742 */
          for (int i = 0; i < n; ++i) {
            if ((b[i >> 6] >> (i & 63)) & 1) {
              best = TOPBIT | (uint64_t)i;
              if (!reverse) break;
            }
          }
        }
      }
      fcur[x] = best;
      labels[(size_t)y * width + x] = (int32_t)(best & 0xffffffffu);
      if (fwd_out) fwd_out[(size_t)y * width + x] = best;
    }
    uint64_t* tmp = fprev;
    fprev = fcur;
    fcur = tmp;
  }
  if (bwd_out) memcpy(bwd_out, bwd, (size_t)width * height * sizeof(uint64_t));
  free(rs);
  free(bits);
  free(bwd);
  free(fprev);
  free(fcur);
}

/* ------------------------------------------------------------------------
 * Level count: multiblend.cpp:161-206
 * ---------------------------------------------------------------------- */
static int cmp_int(const void* a, const void* b) {
  return *(const int*)a - *(const int*)b;
}

int mbo_blend_levels(int n, const int* widths, const int* heights,
                     int fixed_levels, int add_levels) {
  int levels;
  if (fixed_levels == 0) {
    int* ws = (int*)malloc(sizeof(int) * n);
    int* hs = (int*)malloc(sizeof(int) * n);
    memcpy(ws, widths, sizeof(int) * n);
    memcpy(hs, heights, sizeof(int) * n);
    qsort(ws, n, sizeof(int), cmp_int);
    qsort(hs, n, sizeof(int), cmp_int);
    int half = (n - 1) >> 1;
    int mw = (n & 1) ? ws[half] : (ws[half] + ws[half + 1] + 1) >> 1;
    int mh = (n & 1) ? hs[half] : (hs[half] + hs[half + 1] + 1) >> 1;
    int blend_wh = mw > mh ? mw : mh;
    free(ws);
    free(hs);
    /* std::log2(blend_wh + 4.0f) is evaluated in float */
    levels = (int)floorf(log2f((float)blend_wh + 4.0f) - 1);
  } else {
    levels = fixed_levels;
  }
  levels += add_levels;
  if (n == 1) levels = 0;
  return levels;
}

/* ------------------------------------------------------------------------
 * Pyramid geometry: pyramid.cpp:21-98 (bands are a threading detail and do
 * not change results).
 * ---------------------------------------------------------------------- */
typedef struct {
  int w, h, pitch, rows; /* rows = allocated rows */
  int x, y;
  int xs, ys; /* x_shift, y_shift */
  float* data;
} level_t;

typedef struct {
  int n;
  level_t lv[MBO_MAX_LEVELS];
} pyr_t;

int mbo_pyramid_geometry(int width, int height, int levels, int x, int y,
                         int* lw, int* lh, int* lx, int* ly, int* xshift,
                         int* yshift, int* pitch) {
  int b = 0, req = 2;
  for (int n = 0; n < levels; ++n) {
    int xs = ((x - b) & (req - 1)) != 0;
    int ys = ((y - b) & (req - 1)) != 0;
    lw[n] = width;
    lh[n] = height;
    lx[n] = x;
    ly[n] = y;
    xshift[n] = xs;
    yshift[n] = ys;
    pitch[n] = (width + xs + 7) & ~7;
    x -= xs << n;
    y -= ys << n;
    x -= req;
    y -= req;
    b -= req;
    req <<= 1;
    width = (width + xs + 6) >> 1;
    height = (height + ys + 6) >> 1;
  }
  return levels;
}

static void pyr_init(pyr_t* p, int width, int height, int levels, int x,
                     int y) {
  int lw[MBO_MAX_LEVELS], lh[MBO_MAX_LEVELS], lx[MBO_MAX_LEVELS],
      ly[MBO_MAX_LEVELS], xs[MBO_MAX_LEVELS], ys[MBO_MAX_LEVELS],
      pitch[MBO_MAX_LEVELS];
  p->n = mbo_pyramid_geometry(width, height, levels, x, y, lw, lh, lx, ly, xs,
                              ys, pitch);
  for (int l = 0; l < p->n; ++l) {
    level_t* L = &p->lv[l];
    L->w = lw[l];
    L->h = lh[l];
    L->x = lx[l];
    L->y = ly[l];
    L->xs = xs[l];
    L->ys = ys[l];
    L->pitch = pitch[l];
    L->rows = (lh[l] + ys[l] + 3) & ~3;
    L->data = (float*)calloc((size_t)L->pitch * L->rows, sizeof(float));
  }
}

static void dbg_dump(const char* tag, int c, int i, const pyr_t* p) {
  const char* dir = getenv("MBO_DUMP");
  if (!dir) return;
  for (int l = 0; l < p->n; ++l) {
    char fn[256];
    snprintf(fn, sizeof fn, "%s/or_%s_c%d_i%d_l%d.bin", dir, tag, c, i, l);
    FILE* f = fopen(fn, "wb");
    int hdr[3] = {p->lv[l].w, p->lv[l].h, p->lv[l].pitch};
    fwrite(hdr, 4, 3, f);
    fwrite(p->lv[l].data, 4, (size_t)p->lv[l].pitch * p->lv[l].h, f);
    fclose(f);
  }
}

static void pyr_free(pyr_t* p) {
  for (int l = 0; l < p->n; ++l) free(p->lv[l].data);
  p->n = 0;
}

/* Copy: pyramid.cpp:233-408 (planar, no gamma): u8/u16 -> float, pad the
 * pitch by replicating the last pixel. */
static void pyr_copy(pyr_t* p, const void* src, int src_pitch, int bpp) {
  level_t* L = &p->lv[0];
  for (int y = 0; y < L->h; ++y) {
    float* row = L->data + (size_t)y * L->pitch;
    if (bpp == 8) {
      const uint8_t* s = (const uint8_t*)src + (size_t)y * src_pitch;
      for (int x = 0; x < L->w; ++x) row[x] = (float)s[x];
    } else {
      const uint16_t* s = (const uint16_t*)src + (size_t)y * src_pitch;
      for (int x = 0; x < L->w; ++x) row[x] = (float)s[x];
    }
    for (int x = L->w; x < L->pitch; ++x) row[x] = row[L->w - 1];
  }
}

static void pyr_multiply(pyr_t* p, float mul) { /* pyramid.cpp:1168-1193 */
  level_t* L = &p->lv[0];
  size_t n = (size_t)L->pitch * L->h;
  for (size_t i = 0; i < n; ++i) L->data[i] = L->data[i] * mul;
}

/* ------------------------------------------------------------------------
 * Shrink: pyramid.cpp:584-823. Vertical [1 4 6 4 1] with the reference's
 * exact edge-row forms into a line buffer, then Squeeze horizontally.
 * ---------------------------------------------------------------------- */
static void squeeze(const float* line, int hi_pitch, float* lo, int lo_pitch,
                    float final_mul, int x_shift) {
  /* pyramid.cpp:756-823. In shifted coordinates L[k] = line[k - x_shift]
   * with replicate on both sides (left: a = splat(b[0]) and the memmove
   * leaving line[0] in place; right: splat of the last float of the line,
   * which is index hi_pitch-1 of the (shifted) line buffer). */
  for (int j = 0; j < lo_pitch; ++j) {
    float t[5];
    for (int k = 0; k < 5; ++k) {
      int idx = 2 * j - 4 + k; /* shifted-line coordinate */
      if (idx < 0) idx = 0;
      if (idx > hi_pitch - 1) idx = hi_pitch - 1;
      idx -= x_shift; /* memmove by one float */
      if (idx < 0) idx = 0;
      t[k] = line[idx];
    }
    float d = t[0] + t[4];             /* hadd of outer pairs      */
    float hh = 4.0f * (t[1] + t[3]);   /* four * hadd(inner pairs) */
    float jj = 6.0f * t[2];            /* six * centre             */
    lo[j] = (d + (hh + jj)) * final_mul;
  }
}

static void pyr_shrink(pyr_t* p) {
  const float _16th = 1.0f / 16, _256th = 1.0f / 256;
  for (int l = 0; l < p->n - 1; ++l) {
    const level_t* H = &p->lv[l];
    level_t* Lo = &p->lv[l + 1];
    const int hp = H->pitch, lp = Lo->pitch;
    const float* hi = H->data;
    float* line = (float*)malloc(sizeof(float) * hp);
    const int ys = H->ys;
    const int height_odd = (H->h & 1) ^ ys;
    const int first_bad = Lo->h - (3 - height_odd);
    const int last = H->h - 1;
#define HI(r) (hi + (size_t)(r) * hp)
    for (int y = 0; y < Lo->h; ++y) {
      float* lo = Lo->data + (size_t)y * lp;
      float mul = _256th;
      int c = 2 * y - 2 - ys; /* centre row in hi */
      if (y == 0) {           /* :632-635 */
        memcpy(line, HI(0), sizeof(float) * hp);
        mul = _16th;
      } else if (y == 1 && !ys) { /* :639-644 */
        for (int x = 0; x < hp; ++x)
          line[x] = (HI(0)[x] * 11.0f + HI(1)[x] * 4.0f) + HI(2)[x];
      } else if (y == 1 && ys) { /* :651-654 */
        for (int x = 0; x < hp; ++x) line[x] = HI(0)[x] * 15.0f + HI(1)[x];
      } else if (y < first_bad || (y == first_bad && height_odd == 0)) {
        /* good lines :690-701, line 2 of the y_shift case :659-670 and the
         * prepenultimate line :710-719 share one form with clamped taps */
        int r0 = c - 2 < 0 ? 0 : c - 2, r1 = c - 1 < 0 ? 0 : c - 1;
        int r3 = c + 1 > last ? last : c + 1, r4 = c + 2 > last ? last : c + 2;
        const float *t0 = HI(r0), *t1 = HI(r1), *t2 = HI(c), *t3 = HI(r3),
                    *t4 = HI(r4);
        for (int x = 0; x < hp; ++x)
          line[x] = ((t0[x] + t4[x]) + (t1[x] + t3[x]) * 4.0f) + t2[x] * 6.0f;
      } else if (y == Lo->h - 2 && height_odd == 0) { /* :727-731 */
        for (int x = 0; x < hp; ++x)
          line[x] = HI(last - 1)[x] + HI(last)[x] * 15.0f;
      } else if (y == Lo->h - 2) { /* penultimate, height_odd :734-739 */
        for (int x = 0; x < hp; ++x)
          line[x] =
              (HI(last)[x] * 11.0f + HI(last - 1)[x] * 4.0f) + HI(last - 2)[x];
      } else { /* last line :749-751 */
        memcpy(line, HI(last), sizeof(float) * hp);
        mul = _16th;
      }
      squeeze(line, hp, lo, lp, mul, H->xs);
    }
#undef HI
    free(line);
  }
}

/* ------------------------------------------------------------------------
 * Laplace / Collapse: pyramid.cpp:828-1038. hi <- expand(lo) - hi.
 * Horizontal expand into a temp line (rounded to float), then vertical.
 * The grouping of the three-term sums depends on x mod 4 because the
 * reference evaluates them with two levels of _mm_hadd_ps.
 * ---------------------------------------------------------------------- */
static inline float lo_at(const float* lo, int lo_pitch, int k) {
  /* beyond the pitch the reference splats the last loaded float
   * (pyramid.cpp:877-881) */
  return lo[k < lo_pitch ? k : lo_pitch - 1];
}

static void expand_line(const float* lo, int lo_pitch, float* out,
                        int hi_pitch, int shifted) {
  for (int x = 0; x < hi_pitch; ++x) {
    if (!shifted) { /* LaplaceExpand :851-896 */
      int m = x >> 1;
      if ((x & 1) == 0) {
        float a = 0.125f * lo_at(lo, lo_pitch, m);
        float b = 0.75f * lo_at(lo, lo_pitch, m + 1);
        float c = 0.125f * lo_at(lo, lo_pitch, m + 2);
        out[x] = ((x & 3) == 0) ? (a + b) + c : a + (b + c);
      } else {
        out[x] = 0.5f * lo_at(lo, lo_pitch, m + 1) +
                 0.5f * lo_at(lo, lo_pitch, m + 2);
      }
    } else { /* LaplaceExpandShifted :898-946 */
      int m = x >> 1;
      if ((x & 1) == 0) {
        out[x] = 0.5f * lo_at(lo, lo_pitch, m + 1) +
                 0.5f * lo_at(lo, lo_pitch, m + 2);
      } else {
        float a = 0.125f * lo_at(lo, lo_pitch, m + 1);
        float b = 0.75f * lo_at(lo, lo_pitch, m + 2);
        float c = 0.125f * lo_at(lo, lo_pitch, m + 3);
        out[x] = ((x & 3) == 1) ? a + (b + c) : (a + b) + c;
      }
    }
  }
}

static void laplace_level(level_t* H, const level_t* Lo) {
  const int hp = H->pitch;
  float* e0 = (float*)malloc(sizeof(float) * hp);
  float* e1 = (float*)malloc(sizeof(float) * hp);
  float* e2 = (float*)malloc(sizeof(float) * hp);
  for (int y = 0; y < H->h; ++y) {
    float* hi = H->data + (size_t)y * hp;
    /* LaplaceThread :1004-1038: with j = (y + 2 + ys) / 2 the rows of lo
     * used are j-1, j, j+1 (Line3) or j, j+1 (Line2). */
    int t = y + H->ys;
    int j = (t >> 1) + 1;
#define LO(r) (Lo->data + (size_t)(r) * Lo->pitch)
    if ((t & 1) == 0) { /* LaplaceLine3 :960-975 */
      expand_line(LO(j - 1), Lo->pitch, e0, hp, H->xs);
      expand_line(LO(j), Lo->pitch, e1, hp, H->xs);
      expand_line(LO(j + 1), Lo->pitch, e2, hp, H->xs);
      for (int x = 0; x < hp; ++x)
        hi[x] = ((e0[x] + e2[x]) * 0.125f + e1[x] * 0.75f) - hi[x];
    } else { /* LaplaceLine2 :948-958 */
      expand_line(LO(j), Lo->pitch, e1, hp, H->xs);
      expand_line(LO(j + 1), Lo->pitch, e2, hp, H->xs);
      for (int x = 0; x < hp; ++x) hi[x] = ((e1[x] + e2[x]) * 0.5f) - hi[x];
    }
#undef LO
  }
  free(e0);
  free(e1);
  free(e2);
}

static void pyr_laplace_collapse(pyr_t* p, int n_levels, int collapse) {
  for (int j = 0; j < n_levels - 1; ++j) {
    int l = collapse ? (n_levels - 2) - j : j;
    laplace_level(&p->lv[l], &p->lv[l + 1]);
  }
}

/* ------------------------------------------------------------------------
 * Mask pyramid: functions.cpp:24-314, kept in the run-length domain because
 * the token structure is observable: (1) inside a run Squish copies the value
 * instead of filtering it, (2) CompositeLine overwrites for an int run of 1s
 * but multiplies and accumulates for a bare float that happens to be 1.0f.
 * Token kinds: 0 = bare float (count 1), 1 = int run (value 0 or 1),
 * 2 = run of an arbitrary float.
 * ---------------------------------------------------------------------- */
typedef struct {
  float v;
  int n;
  int kind;
} tok_t;

typedef struct {
  tok_t* t;
  int n, cap;
} tokrow_t;

typedef struct {
  int w, h;
  tokrow_t* rows;
} rlemask_t;

static void tok_push(tokrow_t* r, float v, int n, int kind) {
  if (r->n == r->cap) {
    r->cap = r->cap ? r->cap * 2 : 16;
    r->t = (tok_t*)realloc(r->t, sizeof(tok_t) * r->cap);
  }
  r->t[r->n].v = v;
  r->t[r->n].n = n;
  r->t[r->n].kind = kind;
  r->n++;
}

static void rle_free(rlemask_t* m) {
  if (!m->rows) return;
  for (int y = 0; y < m->h; ++y) free(m->rows[y].t);
  free(m->rows);
  m->rows = NULL;
}

static void rle_alloc(rlemask_t* m, int w, int h) {
  m->w = w;
  m->h = h;
  m->rows = (tokrow_t*)calloc(h, sizeof(tokrow_t));
}

/* Record(): multiblend.cpp:29-57 -> maximal runs of owner / not owner */
static void rle_level0(rlemask_t* m, const int32_t* labels, int i, int w,
                       int h) {
  rle_alloc(m, w, h);
  for (int y = 0; y < h; ++y) {
    const int32_t* l = labels + (size_t)y * w;
    int x = 0;
    while (x < w) {
      int own = l[x] == i, s = x;
      while (x < w && (l[x] == i) == own) ++x;
      tok_push(&m->rows[y], (float)own, x - s, 1);
    }
  }
}

/* Squish: functions.cpp:24-170 */
static void squish(const tokrow_t* in, int in_width, int out_width,
                   tokrow_t* out) {
  int ip = 0, read = 0, wrote = 0, in_count, out_count, last_int = -1;
  float current_val, a, b, c, d, e;
  int c_read, e_read = 0;
  out->n = 0;
#define READ_TOKEN()                                          \
  do {                                                        \
    const tok_t* tk = &in->t[ip++];                           \
    in_count = tk->n;                                         \
    current_val = tk->v;                                      \
    last_int = tk->kind == 1 ? (int)tk->v : -1;               \
  } while (0)
  READ_TOKEN();
  if (in_count == in_width) { /* :58-66 */
    tok_push(out, current_val, out_width, last_int >= 0 ? 1 : 2);
    return;
  }
  out_count = (in_count + 1) >> 1; /* :68-78 */
  if (last_int >= 0)
    tok_push(out, current_val, out_count, 1);
  else
    tok_push(out, current_val, out_count, out_count > 1 ? 2 : 0);
  wrote = out_count;
  a = b = c = current_val;
  read = in_count;
  in_count -= (in_count - 1) | 1;
  c_read = ip;
  while (wrote < out_width) { /* :87-167 */
    if (in_count == 0 && read < in_width) {
      READ_TOKEN();
      read += in_count;
    }
    if (read == in_width) in_count = 0x7fffffff;
    if (in_count >= 2) {
      d = e = current_val;
      e_read = ip;
      in_count -= 2;
    } else {
      d = current_val;
      READ_TOKEN();
      read += in_count;
      e = current_val;
      e_read = ip;
      --in_count;
    }
    tok_push(out, (b + d) * 0.25f + (a + e + 6 * c) * 0.0625f, 1, 0);
    wrote++;
    if (c_read == ip && in_count >= 4) {
      out_count = in_count >> 1;
      if (out_count > out_width - wrote) out_count = out_width - wrote;
      if (last_int >= 0)
        tok_push(out, (float)last_int, out_count, 1);
      else
        tok_push(out, e, out_count, 2);
      wrote += out_count;
      in_count -= out_count << 1;
    }
    a = c;
    b = d;
    c = e;
    c_read = e_read;
  }
#undef READ_TOKEN
}

/* ShrinkMasks, one level: functions.cpp:187-313 */
static void rle_shrink(const rlemask_t* in, rlemask_t* out) {
  const int ow = (in->w + 6) >> 1, oh = (in->h + 6) >> 1;
  rle_alloc(out, ow, oh);
  tokrow_t* sq = (tokrow_t*)calloc(in->h, sizeof(tokrow_t));
  for (int y = 0; y < in->h; ++y) squish(&in->rows[y], in->w, ow, &sq[y]);
  /* first line: the squished first input line verbatim (:195-214) */
  for (int k = 0; k < sq[0].n; ++k)
    tok_push(&out->rows[0], sq[0].t[k].v, sq[0].t[k].n, sq[0].t[k].kind);
  for (int y = 1; y < oh; ++y) {
    const tokrow_t* ln[5];
    int ptr[5] = {0, 0, 0, 0, 0}, cnt[5] = {0, 0, 0, 0, 0};
    float vals[5];
    for (int k = 0; k < 5; ++k) {
      int r = 2 * y - 4 + k;
      r = r < 0 ? 0 : (r > in->h - 1 ? in->h - 1 : r);
      ln[k] = &sq[r];
    }
    int wrote = 0;
    while (wrote < ow) { /* :237-281 */
      int min_count = 0x7fffffff;
      for (int k = 0; k < 5; ++k) {
        if (cnt[k] == 0) {
          const tok_t* tk = &ln[k]->t[ptr[k]++];
          vals[k] = tk->v;
          cnt[k] = tk->n;
        }
        if (cnt[k] < min_count) min_count = cnt[k];
      }
      float val = (vals[0] + vals[4]) * 0.0625f + (vals[1] + vals[3]) * 0.25f +
                  vals[2] * 0.375f;
      if (val == 0 || val == 1)
        tok_push(&out->rows[y], val, min_count, min_count == 1 ? 0 : 1);
      else
        tok_push(&out->rows[y], val, min_count, min_count > 1 ? 2 : 0);
      wrote += min_count;
      for (int k = 0; k < 5; ++k) cnt[k] -= min_count;
    }
  }
  for (int y = 0; y < in->h; ++y) free(sq[y].t);
  free(sq);
}

static void rle_dense(const rlemask_t* m, float* out) {
  for (int y = 0; y < m->h; ++y) {
    int x = 0;
    for (int k = 0; k < m->rows[y].n; ++k)
      for (int j = 0; j < m->rows[y].t[k].n && x < m->w; ++j)
        out[(size_t)y * m->w + x++] = m->rows[y].t[k].v;
  }
}

/* ------------------------------------------------------------------------
 * Composite: multiblend.cpp:1087-1123 + functions.cpp:319-411.
 * ---------------------------------------------------------------------- */
static void composite_level(const level_t* in, level_t* out, int level,
                            int first, const rlemask_t* mask) {
  const int x_offset = (in->x - out->x) >> level;
  const int y_offset = (in->y - out->y) >> level;
  for (int y = 0; y < out->h; ++y) {
    int in_line = y - y_offset;
    if (in_line < 0) in_line = 0;
    if (in_line > in->h - 1) in_line = in->h - 1;
    const float* ip = in->data + (size_t)in_line * in->pitch;
    float* op = out->data + (size_t)y * out->pitch;
    const tokrow_t* row = &mask->rows[y];
    int x = 0;
    for (int k = 0; k < row->n && x < out->w; ++k) {
      const tok_t* tk = &row->t[k];
      int lim = x + tk->n;
      if (lim > out->w) lim = out->w;
      for (; x < lim; ++x) {
        int ix = x - x_offset;
        if (ix < 0) ix = 0;
        if (ix > in->w - 1) ix = in->w - 1;
        if (tk->kind == 1 && tk->v == 0.0f) { /* :347-351 */
          if (first) op[x] = 0.0f;
        } else if (tk->kind == 1) { /* :352-370 */
          op[x] = ip[ix];
        } else { /* :371-401 */
          float f = ip[ix] * tk->v;
          if (first)
            op[x] = f;
          else
            op[x] += f;
        }
      }
    }
    for (x = out->w; x < out->pitch; ++x) op[x] = op[out->w - 1];
  }
}

/* Out: pyramid.cpp:1563-1794 with Loader<kClampDither> :2054-2066
 * (or kClamp :2023-2030 without dither), round-half-even. */
static const float kDither[4][4] = {
    {-0.125f, 0.375f, 0.0f, 0.4999f},
    {0.125f, -0.375f, 0.25f, -0.25f},
    {-0.0625f, 0.4375f, -0.1875f, 0.3125f},
    {0.1875f, -0.3125f, 0.0625f, -0.4375f}};

static void pyr_out(const pyr_t* p, void* dst, int bpp, int dither) {
  const level_t* L = &p->lv[0];
  const float maxv = bpp == 8 ? 255.0f : 65535.0f;
  for (int y = 0; y < L->h; ++y) {
    const float* row = L->data + (size_t)y * L->pitch;
    for (int x = 0; x < L->w; ++x) {
      float v = row[x];
      if (dither) v = v + kDither[y & 3][x & 3];
      /* _mm_max_ps(v, 0) then _mm_min_ps(., max) */
      v = v > 0.0f ? v : 0.0f;
      v = v < maxv ? v : maxv;
      long q = lrintf(v); /* cvtps_epi32: round to nearest even */
      if (bpp == 8)
        ((uint8_t*)dst)[(size_t)y * L->w + x] = (uint8_t)q;
      else
        ((uint16_t*)dst)[(size_t)y * L->w + x] = (uint16_t)q;
    }
  }
}

/* ------------------------------------------------------------------------
 * Driver: multiblend.cpp:66-1307 for in-memory inputs.
 * ---------------------------------------------------------------------- */
void mbo_free_result(mbo_result* r) {
  if (!r) return;
  for (int c = 0; c < 3; ++c) {
    free(r->planes[c]);
    free(r->collapsed[c]);
  }
  free(r->full_mask);
  free(r->xor_mask);
  free(r->labels);
  free(r->seam_bwd);
  free(r->seam_fwd);
  if (r->images) {
    for (int i = 0; i < r->n_images; ++i) {
      free(r->images[i].opaque);
      for (int c = 0; c < 3; ++c) free(r->images[i].chan[c]);
    }
    free(r->images);
  }
  memset(r, 0, sizeof(*r));
}

int mbo_blend(int n, const mbo_image* images, const mbo_options* opts,
              int keep_debug, mbo_result* out) {
  memset(out, 0, sizeof(*out));
  if (n < 1) FAIL("no images");
  int output_bpp = opts->output_bpp;
  for (int i = 0; i < n; ++i) { /* multiblend.cpp:100-117 */
    if (output_bpp == 0 && images[i].bpp == 16) output_bpp = 16;
    if (images[i].bpp != images[0].bpp)
      FAIL("mixture of 8bpp and 16bpp images");
    if (images[i].spp != 3 && images[i].spp != 4) FAIL("spp must be 3 or 4");
  }
  if (output_bpp == 0) output_bpp = 8;
  const int bpp = images[0].bpp;
  const int bs = bpp >> 3; /* bytes per sample */

  mbo_trimmed* T = (mbo_trimmed*)calloc(n, sizeof(mbo_trimmed));
  out->images = T;
  out->n_images = n;

  /* Read / trim / inpaint / extract: image.cpp:336-919 */
  for (int i = 0; i < n; ++i) {
    const mbo_image* im = &images[i];
    size_t px = (size_t)im->width * im->height;
    int bbox[4] = {0, 0, im->width, im->height};
    void* scratch = malloc(px * im->spp * bs);
    memcpy(scratch, im->data, px * im->spp * bs);
    if (im->spp == 4) {
      if (mbo_trim(scratch, im->width, im->height, bpp, bbox)) {
        free(scratch);
        mbo_free_result(out);
        FAIL("image %d has no opaque pixel", i);
      }
    }
    T[i].left = bbox[0];
    T[i].top = bbox[1];
    T[i].width = bbox[2];
    T[i].height = bbox[3];
    T[i].xpos = im->xpos + bbox[0];
    T[i].ypos = im->ypos + bbox[1];
    size_t tp = (size_t)bbox[2] * bbox[3];
    T[i].opaque = (uint8_t*)malloc(tp);
    if (im->spp == 4)
      mbo_inpaint(scratch, im->width, bpp, bbox, T[i].opaque);
    else
      memset(T[i].opaque, 1, tp);
    for (int c = 0; c < 3; ++c) T[i].chan[c] = malloc(tp * bs);
    for (int y = 0; y < bbox[3]; ++y) { /* image.cpp:826-915 */
      for (int x = 0; x < bbox[2]; ++x) {
        size_t s = ((size_t)(bbox[1] + y) * im->width + bbox[0] + x) * im->spp;
        size_t d = (size_t)y * bbox[2] + x;
        for (int c = 0; c < 3; ++c) {
          if (bpp == 8)
            ((uint8_t*)T[i].chan[c])[d] = ((uint8_t*)scratch)[s + c];
          else
            ((uint16_t*)T[i].chan[c])[d] = ((uint16_t*)scratch)[s + c];
        }
      }
    }
    free(scratch);
  }

  /* Tighten: multiblend.cpp:139-154 */
  int min_x = 0x7fffffff, min_y = 0x7fffffff, width = 0, height = 0;
  for (int i = 0; i < n; ++i) {
    if (T[i].xpos < min_x) min_x = T[i].xpos;
    if (T[i].ypos < min_y) min_y = T[i].ypos;
  }
  for (int i = 0; i < n; ++i) {
    T[i].xpos -= min_x;
    T[i].ypos -= min_y;
    if (T[i].xpos + T[i].width > width) width = T[i].xpos + T[i].width;
    if (T[i].ypos + T[i].height > height) height = T[i].ypos + T[i].height;
  }
  int* tmp = (int*)malloc(sizeof(int) * n * 4);
  int *xs = tmp, *ys = tmp + n, *ws = tmp + 2 * n, *hs = tmp + 3 * n;
  const uint8_t** ops = (const uint8_t**)malloc(sizeof(void*) * n);
  for (int i = 0; i < n; ++i) {
    xs[i] = T[i].xpos;
    ys[i] = T[i].ypos;
    ws[i] = T[i].width;
    hs[i] = T[i].height;
    ops[i] = T[i].opaque;
  }
  int blend_levels =
      mbo_blend_levels(n, ws, hs, opts->fixed_levels, opts->add_levels);
  if (opts->wideblend && opts->fixed_levels == 0 && n > 1) {
    /* multiblend.cpp:189-193: blend_wh = max(width, height); one more */
    const int blend_wh = width > height ? width : height;
    blend_levels = (int)floorf(log2f((float)blend_wh + 4.0f) - 1) + 1 +
                   opts->add_levels;
  }
  const size_t C = (size_t)width * height;
  out->width = width;
  out->height = height;
  out->bpp = output_bpp;
  out->min_xpos = min_x;
  out->min_ypos = min_y;
  out->levels = blend_levels;

  /* Seams, full/xor masks, channel sums: multiblend.cpp:254-823 */
  out->labels = (int32_t*)malloc(C * sizeof(int32_t));
  if (keep_debug) {
    out->seam_bwd = (uint64_t*)malloc(C * sizeof(uint64_t));
    out->seam_fwd = (uint64_t*)malloc(C * sizeof(uint64_t));
  }
  mbo_seams(n, xs, ys, ws, hs, ops, width, height, opts->reverse, out->labels,
            out->seam_bwd, out->seam_fwd);
  out->full_mask = (uint8_t*)calloc(C, 1);
  out->xor_mask = (uint8_t*)calloc(C, 1);
  {
    uint8_t* cnt = (uint8_t*)calloc(C, 1);
    int32_t* who = (int32_t*)malloc(C * sizeof(int32_t));
    for (int i = 0; i < n; ++i)
      for (int y = 0; y < hs[i]; ++y)
        for (int x = 0; x < ws[i]; ++x)
          if (ops[i][(size_t)y * ws[i] + x]) {
            size_t p = (size_t)(ys[i] + y) * width + xs[i] + x;
            if (cnt[p] < 255) cnt[p]++;
            who[p] = i;
          }
    int alpha = 0;
    for (int y = 0; y < height; ++y)
      for (int x = 0; x < width; ++x) {
        size_t p = (size_t)y * width + x;
        if (cnt[p] == 0) alpha = 1;
        out->full_mask[p] = cnt[p] ? 255 : 0;
        if (cnt[p] == 1) { /* :542-605 */
          int i = who[p];
          size_t q = (size_t)(y - ys[i]) * ws[i] + (x - xs[i]);
          out->xor_mask[p] = 1;
          out->total_pixels++;
          for (int c = 0; c < 3; ++c)
            out->channel_totals[c] +=
                bpp == 8 ? ((uint8_t*)T[i].chan[c])[q]
                         : ((uint16_t*)T[i].chan[c])[q];
        }
      }
    out->no_mask = !alpha; /* :821-823 */
    free(cnt);
    free(who);
  }

  /* Output pyramid and blend loop: multiblend.cpp:988-1131 */
  const int total_levels = blend_levels > 1 ? blend_levels : 1;
  pyr_t outp[3];
  for (int c = 0; c < 3; ++c) pyr_init(&outp[c], width, height, total_levels, 0, 0);
  out->out_pitch = outp[0].lv[0].pitch;
  const float conv = output_bpp == 8 ? 1.0f / 257 : 257.0f;

  if (n > 1) {
    rlemask_t* mask = (rlemask_t*)calloc(blend_levels > 0 ? blend_levels : 1,
                                         sizeof(rlemask_t));
    for (int i = 0; i < n; ++i) {
      if (blend_levels > 0) rle_level0(&mask[0], out->labels, i, width, height);
      for (int l = 1; l < blend_levels; ++l) rle_shrink(&mask[l - 1], &mask[l]);
      if (getenv("MBO_DUMP")) {
        for (int l = 0; l < blend_levels; ++l) {
          char fn[256];
          snprintf(fn, sizeof fn, "%s/or_mask_i%d_l%d.bin", getenv("MBO_DUMP"),
                   i, l);
          FILE* f = fopen(fn, "wb");
          int hdr[3] = {mask[l].w, mask[l].h, mask[l].w};
          float* dm = (float*)malloc(sizeof(float) * (size_t)hdr[0] * hdr[1]);
          rle_dense(&mask[l], dm);
          fwrite(hdr, 4, 3, f);
          fwrite(dm, 4, (size_t)hdr[0] * hdr[1], f);
          fclose(f);
          free(dm);
        }
      }
      for (int c = 0; c < 3; ++c) {
        pyr_t ip;
        pyr_init(&ip, ws[i], hs[i], blend_levels, xs[i], ys[i]);
        if (ip.n > 0) {
          pyr_copy(&ip, T[i].chan[c], ws[i], bpp);
          if (output_bpp != bpp) pyr_multiply(&ip, conv);
          pyr_shrink(&ip);
          pyr_laplace_collapse(&ip, ip.n, 0);
          dbg_dump("lap", c, i, &ip);
        }
        for (int l = 0; l < blend_levels; ++l)
          composite_level(&ip.lv[l], &outp[c].lv[l], l, i == 0, &mask[l]);
        pyr_free(&ip);
      }
      for (int l = 0; l < blend_levels; ++l) rle_free(&mask[l]);
    }
    free(mask);
    for (int c = 0; c < 3; ++c) dbg_dump("out", c, 0, &outp[c]);
    for (int c = 0; c < 3; ++c) pyr_laplace_collapse(&outp[c], blend_levels, 1);
  } else { /* multiblend.cpp:1131-1145 */
    for (int c = 0; c < 3; ++c) {
      pyr_copy(&outp[c], T[0].chan[c], ws[0], bpp);
      if (output_bpp != bpp) pyr_multiply(&outp[c], conv);
    }
  }

  /* Offset correction + Out: multiblend.cpp:1232-1286 */
  for (int c = 0; c < 3; ++c) {
    level_t* L0 = &outp[c].lv[0];
    if (out->total_pixels != 0) {
      double channel_total = 0;
      for (int y = 0; y < height; ++y) {
        const float* row = L0->data + (size_t)y * L0->pitch;
        const uint8_t* xm = out->xor_mask + (size_t)y * width;
        for (int x = 0; x < width; ++x)
          if (xm[x]) channel_total += row[x];
      }
      float avg = (float)out->channel_totals[c] / (float)out->total_pixels;
      if (output_bpp != bpp) {
        if (output_bpp == 8)
          avg /= 256;
        else
          avg *= 256;
      }
      float output_avg = (float)channel_total / (float)out->total_pixels;
      float add = avg - output_avg;
      if (outp[c].n - 1 >= 1) { /* Pyramid::Add lim, pyramid.cpp:1087 */
        size_t cnt = (size_t)L0->pitch * L0->h;
        for (size_t k = 0; k < cnt; ++k) L0->data[k] = add + L0->data[k];
      }
    }
    out->planes[c] = malloc(C * (output_bpp >> 3));
    pyr_out(&outp[c], out->planes[c], output_bpp, opts->dither);
    if (keep_debug) {
      size_t cnt = (size_t)L0->pitch * L0->h;
      out->collapsed[c] = (float*)malloc(cnt * sizeof(float));
      memcpy(out->collapsed[c], L0->data, cnt * sizeof(float));
    }
    pyr_free(&outp[c]);
  }

  free(tmp);
  free(ops);
  if (!keep_debug) {
    for (int i = 0; i < n; ++i) {
      free(T[i].opaque);
      T[i].opaque = NULL;
      for (int c = 0; c < 3; ++c) {
        free(T[i].chan[c]);
        T[i].chan[c] = NULL;
      }
    }
  }
  return 0;
}
