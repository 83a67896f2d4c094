/* TEST INFRASTRUCTURE: plain C (gcc, no ctypes) user of include/xpano_b200.h:
 * the header must be valid C, and the sequence warp_roi -> warp x2 -> prepare
 * -> feed -> blend -> fetch must run through the ABI from C. Prints a checksum
 * the Python test compares with the ctypes-driven result. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "xpano_b200.h"

#define CHECK(call)                                                   \
  do {                                                                \
    xp_status s_ = (call);                                            \
    if (s_ != XP_OK) {                                                \
      fprintf(stderr, "%s -> %d: %s\n", #call, s_, xp_last_error());  \
      return 1;                                                       \
    }                                                                 \
  } while (0)

int main(void) {
  enum { W = 320, H = 200, N = 3 };
  const float K[9] = {260.f, 0.f, W / 2.f, 0.f, 260.f, H / 2.f, 0.f, 0.f, 1.f};
  static unsigned char img[N][H][W][3], mask[H][W];
  int i, x, y, c;
  if (xp_device_count() < 1) {
    fprintf(stderr, "no CUDA device\n");
    return 77;
  }
  for (i = 0; i < N; ++i)
    for (y = 0; y < H; ++y)
      for (x = 0; x < W; ++x)
        for (c = 0; c < 3; ++c)
          img[i][y][x][c] = (unsigned char)(40 + 60 * i + ((x * 3 + y * 5 + c * 17) % 90));
  memset(mask, 255, sizeof(mask));

  xp_warper* w = NULL;
  xp_blender* b = NULL;
  xp_blend_opts o;
  CHECK(xp_warper_create(XP_PROJ_SPHERICAL, 0, 250.f, 2.f, 1.f, 0, &w));
  xp_blend_opts_default(&o);
  o.fixed_levels = 4;
  CHECK(xp_blender_create(&o, &b));
  CHECK(xp_blender_prepare(b, 0, 0, 1, 1));
  for (i = 0; i < N; ++i) {
    const float a = 0.45f * (float)(i - 1);
    /* R = Ry(a) */
    float R[9] = {0};
    float ca = 1.f - a * a / 2.f + a * a * a * a / 24.f, sa = a - a * a * a / 6.f;
    int roi[4], corner[2], corner2[2];
    R[0] = ca, R[2] = sa, R[4] = 1.f, R[6] = -sa, R[8] = ca;
    CHECK(xp_warper_warp_roi(w, W, H, K, R, roi));
    {
      unsigned char* wi = (unsigned char*)malloc((size_t)roi[2] * roi[3] * 3);
      unsigned char* wm = (unsigned char*)malloc((size_t)roi[2] * roi[3]);
      CHECK(xp_warper_warp(w, img[i], W, H, 0, 8, 3, K, R, XP_INTER_LINEAR,
                           XP_BORDER_REFLECT, wi, 0, XP_MEM_HOST, corner));
      CHECK(xp_warper_warp(w, mask, W, H, 0, 8, 1, K, R, XP_INTER_NEAREST,
                           XP_BORDER_CONSTANT, wm, 0, XP_MEM_HOST, corner2));
      if (corner[0] != roi[0] || corner[1] != roi[1] || corner2[0] != roi[0]) return 1;
      CHECK(xp_blender_feed(b, wi, 0, wm, 0, roi[2], roi[3], 8, corner[0], corner[1],
                            XP_MEM_HOST));
      free(wi);
      free(wm);
    }
  }
  {
    int wh[2], mn[2], bpp, no_mask;
    xp_timing t;
    unsigned long long sum = 0, msum = 0;
    size_t k, npx;
    unsigned char *pano, *pm;
    CHECK(xp_blender_blend(b, wh, mn, &bpp, &no_mask, &t));
    npx = (size_t)wh[0] * wh[1];
    pano = (unsigned char*)malloc(npx * 3);
    pm = (unsigned char*)malloc(npx);
    CHECK(xp_blender_fetch(b, pano, 0, pm, 0, XP_MEM_HOST));
    for (k = 0; k < npx * 3; ++k) sum = sum * 31 + pano[k];
    for (k = 0; k < npx; ++k) msum += pm[k] != 0;
    printf("c_abi ok %d %d %d %d %d %llu %llu\n", wh[0], wh[1], mn[0], mn[1], bpp, sum,
           msum);
    free(pano);
    free(pm);
  }
  xp_blender_destroy(b);
  xp_warper_destroy(w);
  return 0;
}
