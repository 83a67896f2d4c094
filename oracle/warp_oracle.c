/* TEST INFRASTRUCTURE ONLY -- CPU restatement of the OpenCV half of the hot
 * path: cv::detail::RotationWarper (stitching/detail/warpers_inl.hpp,
 * warpers.cpp) + imgproc remap / resize(INTER_LINEAR_EXACT) / dilate as called
 * from xpano/algorithm/stitcher.cc:144-169, 377-401.
 *
 * OpenCV is a third-party dependency that is NOT vendored under
 * /root/reference (CI pins 4.11.0, .github/workflows/test.yml:16); this file
 * restates its published algorithm in float32 with glibc's libm, which is
 * what OpenCV itself calls. Parity pinning: tests/test_oracle_warp.py checks
 * it against Python cv2 (PyRotationWarper.buildMaps / warpRoi / warpPoint,
 * cv2.remap, cv2.resize, cv2.dilate) where cv2 is importable, and against
 * vectors recorded from cv2 4.13.0 in tests/golden/warp_cv2_goldens.npz.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
 * call into this file.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define WO_PI 3.14159265358979323846 /* CV_PI */

/* projection ids = xpano ProjectionType order (algorithm/options.h:16-26) */
enum {
  WO_PLANE = 0,
  WO_CYLINDRICAL,
  WO_SPHERICAL,
  WO_FISHEYE,
  WO_STEREOGRAPHIC,
  WO_COMPRESSED_RECTILINEAR,
  WO_PANINI,
  WO_MERCATOR,
  WO_TRANSVERSE_MERCATOR
};

typedef struct {
  int type, portrait;
  float scale, a, b;
  float k[9], rinv[9], r_kinv[9], k_rinv[9], t[3];
} wo_projector;

/* ProjectorBase::setCameraParams (warpers_inl.hpp): K, R are 3x3 float32.
 * K.inv(): 3x3 closed form evaluated in double, rounded to float;
 * R*Kinv and K*Rinv: float gemm with float accumulation. */
static void mat3_mul(const float* a, const float* b, float* out) {
  /* cv::gemm on 3x3 CV_32F accumulates in float, k ascending (checked
   * bit-exact against cv2 buildMaps) */
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      float s = 0;
      for (int k = 0; k < 3; ++k) s += a[i * 3 + k] * b[k * 3 + j];
      out[i * 3 + j] = s;
    }
}

static void mat3_inv(const float* m, float* out) {
  double d = (double)m[0] * ((double)m[4] * m[8] - (double)m[5] * m[7]) -
             (double)m[1] * ((double)m[3] * m[8] - (double)m[5] * m[6]) +
             (double)m[2] * ((double)m[3] * m[7] - (double)m[4] * m[6]);
  if (d == 0) {
    memset(out, 0, sizeof(float) * 9);
    return;
  }
  d = 1. / d;
  out[0] = (float)(((double)m[4] * m[8] - (double)m[5] * m[7]) * d);
  out[1] = (float)(((double)m[2] * m[7] - (double)m[1] * m[8]) * d);
  out[2] = (float)(((double)m[1] * m[5] - (double)m[2] * m[4]) * d);
  out[3] = (float)(((double)m[5] * m[6] - (double)m[3] * m[8]) * d);
  out[4] = (float)(((double)m[0] * m[8] - (double)m[2] * m[6]) * d);
  out[5] = (float)(((double)m[2] * m[3] - (double)m[0] * m[5]) * d);
  out[6] = (float)(((double)m[3] * m[7] - (double)m[4] * m[6]) * d);
  out[7] = (float)(((double)m[1] * m[6] - (double)m[0] * m[7]) * d);
  out[8] = (float)(((double)m[0] * m[4] - (double)m[1] * m[3]) * d);
}

void wo_init(wo_projector* p, int type, int portrait, float scale, float a,
             float b, const float* K, const float* R) {
  float kinv[9];
  p->type = type;
  p->portrait = portrait;
  p->scale = scale;
  p->a = a;
  p->b = b;
  memcpy(p->k, K, sizeof(float) * 9);
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) p->rinv[i * 3 + j] = R[j * 3 + i];
  mat3_inv(K, kinv);
  mat3_mul(R, kinv, p->r_kinv);
  mat3_mul(K, p->rinv, p->k_rinv);
  p->t[0] = p->t[1] = p->t[2] = 0.f;
}

static inline void rk(const wo_projector* p, float x, float y, float* x_,
                      float* y_, float* z_) {
  const float* m = p->r_kinv;
  *x_ = m[0] * x + m[1] * y + m[2];
  *y_ = m[3] * x + m[4] * y + m[5];
  *z_ = m[6] * x + m[7] * y + m[8];
}

static inline void project(const wo_projector* p, float x_, float y_, float z_,
                           float* x, float* y) {
  const float* m = p->k_rinv;
  float z;
  *x = m[0] * x_ + m[1] * y_ + m[2] * z_;
  *y = m[3] * x_ + m[4] * y_ + m[5] * z_;
  z = m[6] * x_ + m[7] * y_ + m[8] * z_;
  if (z > 0) {
    *x /= z;
    *y /= z;
  } else {
    *x = *y = -1;
  }
}

#define PIf ((float)WO_PI)

/* mapForward for every projector (warpers_inl.hpp) */
void wo_map_forward(const wo_projector* p, float x, float y, float* u,
                    float* v) {
  float x_, y_, z_;
  rk(p, x, y, &x_, &y_, &z_);
  const float scale = p->scale, a = p->a, b = p->b;
  if (p->portrait) {
    float x0 = x_, y0 = y_;
    switch (p->type) {
      case WO_PLANE: { /* PlanePortraitProjector */
        x_ = y0;
        y_ = x0;
        x_ = p->t[0] + x_ / z_ * (1 - p->t[2]);
        y_ = p->t[1] + y_ / z_ * (1 - p->t[2]);
        *u = -(scale * x_);
        *v = scale * y_;
        return;
      }
      case WO_CYLINDRICAL: { /* CylindricalPortraitProjector */
        x_ = y0;
        y_ = x0;
        float uu = scale * atan2f(x_, z_);
        *v = scale * y_ / sqrtf(x_ * x_ + z_ * z_);
        *u = -uu;
        return;
      }
      case WO_SPHERICAL: { /* SphericalPortraitProjector */
        x_ = y0;
        y_ = x0;
        float uu = scale * atan2f(x_, z_);
        *v = scale *
             (PIf - acosf(y_ / sqrtf(x_ * x_ + y_ * y_ + z_ * z_)));
        *u = -uu;
        return;
      }
      case WO_COMPRESSED_RECTILINEAR: {
        float yy = x0, xx = y0; /* y_ = row 0, x_ = row 1 */
        float u_ = atan2f(xx, z_);
        float v_ = asinf(yy / sqrtf(xx * xx + yy * yy + z_ * z_));
        *u = -scale * a * tanf(u_ / a);
        *v = scale * b * tanf(v_) / cosf(u_);
        return;
      }
      case WO_PANINI: {
        float yy = x0, xx = y0;
        float u_ = atan2f(xx, z_);
        float v_ = asinf(yy / sqrtf(xx * xx + yy * yy + z_ * z_));
        float tg = a * tanf(u_ / a);
        *u = -scale * tg;
        float sinu = sinf(u_);
        if (fabsf(sinu) < 1E-7)
          *v = scale * b * tanf(v_);
        else
          *v = scale * b * tg * tanf(v_) / sinu;
        return;
      }
      default:
        break; /* no portrait variant: fall through to the landscape one */
    }
  }
  switch (p->type) {
    case WO_PLANE:
      x_ = p->t[0] + x_ / z_ * (1 - p->t[2]);
      y_ = p->t[1] + y_ / z_ * (1 - p->t[2]);
      *u = scale * x_;
      *v = scale * y_;
      return;
    case WO_CYLINDRICAL:
      *u = scale * atan2f(x_, z_);
      *v = scale * y_ / sqrtf(x_ * x_ + z_ * z_);
      return;
    case WO_SPHERICAL: {
      *u = scale * atan2f(x_, z_);
      float w = y_ / sqrtf(x_ * x_ + y_ * y_ + z_ * z_);
      *v = scale * (PIf - acosf(w == w ? w : 0));
      return;
    }
    case WO_FISHEYE: {
      float u_ = atan2f(x_, z_);
      float v_ = PIf - acosf(y_ / sqrtf(x_ * x_ + y_ * y_ + z_ * z_));
      *u = scale * v_ * cosf(u_);
      *v = scale * v_ * sinf(u_);
      return;
    }
    case WO_STEREOGRAPHIC: {
      float u_ = atan2f(x_, z_);
      float v_ = PIf - acosf(y_ / sqrtf(x_ * x_ + y_ * y_ + z_ * z_));
      float r = sinf(v_) / (1 - cosf(v_));
      *u = scale * r * cosf(u_);
      *v = scale * r * sinf(u_);
      return;
    }
    case WO_COMPRESSED_RECTILINEAR: {
      float u_ = atan2f(x_, z_);
      float v_ = asinf(y_ / sqrtf(x_ * x_ + y_ * y_ + z_ * z_));
      *u = scale * a * tanf(u_ / a);
      *v = scale * b * tanf(v_) / cosf(u_);
      return;
    }
    case WO_PANINI: {
      float u_ = atan2f(x_, z_);
      float v_ = asinf(y_ / sqrtf(x_ * x_ + y_ * y_ + z_ * z_));
      float tg = a * tanf(u_ / a);
      *u = scale * tg;
      float sinu = sinf(u_);
      if (fabsf(sinu) < 1E-7)
        *v = scale * b * tanf(v_);
      else
        *v = scale * b * tg * tanf(v_) / sinu;
      return;
    }
    case WO_MERCATOR: {
      float u_ = atan2f(x_, z_);
      float v_ = asinf(y_ / sqrtf(x_ * x_ + y_ * y_ + z_ * z_));
      *u = scale * u_;
      *v = scale * logf(tanf((float)(WO_PI / 4) + v_ / 2));
      return;
    }
    case WO_TRANSVERSE_MERCATOR: {
      float u_ = atan2f(x_, z_);
      float v_ = asinf(y_ / sqrtf(x_ * x_ + y_ * y_ + z_ * z_));
      float B = cosf(v_) * sinf(u_);
      *u = scale / 2 * logf((1 + B) / (1 - B));
      *v = scale * atan2f(tanf(v_), cosf(u_));
      return;
    }
  }
}

/* mapBackward for every projector (warpers_inl.hpp) */
void wo_map_backward(const wo_projector* p, float u, float v, float* x,
                     float* y) {
  const float scale = p->scale, a = p->a, b = p->b;
  float x_, y_, z_;
  if (p->portrait) {
    switch (p->type) {
      case WO_PLANE: {
        float uu = -u, vv = v, z;
        const float* m = p->k_rinv;
        uu = uu / scale - p->t[0];
        vv = vv / scale - p->t[1];
        *x = m[0] * vv + m[1] * uu + m[2] * (1 - p->t[2]);
        *y = m[3] * vv + m[4] * uu + m[5] * (1 - p->t[2]);
        z = m[6] * vv + m[7] * uu + m[8] * (1 - p->t[2]);
        *x /= z;
        *y /= z;
        return;
      }
      case WO_CYLINDRICAL: {
        float uu = -u / scale, vv = v / scale;
        float x0 = sinf(uu), y0 = vv;
        z_ = cosf(uu);
        project(p, y0, x0, z_, x, y);
        return;
      }
      case WO_SPHERICAL: {
        float uu = -u / scale, vv = v / scale;
        float sinv = sinf(PIf - vv);
        float x0 = sinv * sinf(uu);
        float y0 = cosf(PIf - vv);
        z_ = sinv * cosf(uu);
        project(p, y0, x0, z_, x, y);
        return;
      }
      case WO_COMPRESSED_RECTILINEAR: {
        u /= -scale;
        v /= scale;
        float aatg = a * atanf(u / a);
        float u_ = aatg;
        float v_ = atanf(v * cosf(aatg) / b);
        float cosv = cosf(v_);
        y_ = cosv * sinf(u_);
        x_ = sinf(v_);
        z_ = cosv * cosf(u_);
        project(p, x_, y_, z_, x, y);
        return;
      }
      case WO_PANINI: {
        u /= -scale;
        v /= scale;
        float lamda = a * atanf(u / a);
        float u_ = lamda, v_;
        if (fabsf(lamda) > 1E-7)
          v_ = atanf(v * sinf(lamda) / (b * a * tanf(lamda / a)));
        else
          v_ = atanf(v / b);
        float cosv = cosf(v_);
        y_ = cosv * sinf(u_);
        x_ = sinf(v_);
        z_ = cosv * cosf(u_);
        project(p, x_, y_, z_, x, y);
        return;
      }
      default:
        break;
    }
  }
  switch (p->type) {
    case WO_PLANE: {
      const float* m = p->k_rinv;
      float z;
      u = u / scale - p->t[0];
      v = v / scale - p->t[1];
      *x = m[0] * u + m[1] * v + m[2] * (1 - p->t[2]);
      *y = m[3] * u + m[4] * v + m[5] * (1 - p->t[2]);
      z = m[6] * u + m[7] * v + m[8] * (1 - p->t[2]);
      *x /= z;
      *y /= z;
      return;
    }
    case WO_CYLINDRICAL:
      u /= scale;
      v /= scale;
      x_ = sinf(u);
      y_ = v;
      z_ = cosf(u);
      break;
    case WO_SPHERICAL: {
      u /= scale;
      v /= scale;
      float sinv = sinf(PIf - v);
      x_ = sinv * sinf(u);
      y_ = cosf(PIf - v);
      z_ = sinv * cosf(u);
      break;
    }
    case WO_FISHEYE: {
      u /= scale;
      v /= scale;
      float u_ = atan2f(v, u);
      float v_ = sqrtf(u * u + v * v);
      float sinv = sinf(PIf - v_);
      x_ = sinv * sinf(u_);
      y_ = cosf(PIf - v_);
      z_ = sinv * cosf(u_);
      break;
    }
    case WO_STEREOGRAPHIC: {
      u /= scale;
      v /= scale;
      float u_ = atan2f(v, u);
      float r = sqrtf(u * u + v * v);
      float v_ = 2 * atanf(1.f / r);
      float sinv = sinf(PIf - v_);
      x_ = sinv * sinf(u_);
      y_ = cosf(PIf - v_);
      z_ = sinv * cosf(u_);
      break;
    }
    case WO_COMPRESSED_RECTILINEAR: {
      u /= scale;
      v /= scale;
      float aatg = a * atanf(u / a);
      float u_ = aatg;
      float v_ = atanf(v * cosf(aatg) / b);
      float cosv = cosf(v_);
      x_ = cosv * sinf(u_);
      y_ = sinf(v_);
      z_ = cosv * cosf(u_);
      break;
    }
    case WO_PANINI: {
      u /= scale;
      v /= scale;
      float lamda = a * atanf(u / a);
      float u_ = lamda, v_;
      if (fabsf(lamda) > 1E-7)
        v_ = atanf(v * sinf(lamda) / (b * a * tanf(lamda / a)));
      else
        v_ = atanf(v / b);
      float cosv = cosf(v_);
      x_ = cosv * sinf(u_);
      y_ = sinf(v_);
      z_ = cosv * cosf(u_);
      break;
    }
    case WO_MERCATOR: {
      u /= scale;
      v /= scale;
      float v_ = atanf(sinhf(v));
      float u_ = u;
      float cosv = cosf(v_);
      x_ = cosv * sinf(u_);
      y_ = sinf(v_);
      z_ = cosv * cosf(u_);
      break;
    }
    default: { /* WO_TRANSVERSE_MERCATOR */
      u /= scale;
      v /= scale;
      float v_ = asinf(sinf(v) / coshf(u));
      float u_ = atan2f(sinhf(u), cosf(v));
      float cosv = cosf(v_);
      x_ = cosv * sinf(u_);
      y_ = sinf(v_);
      z_ = cosv * cosf(u_);
      break;
    }
  }
  project(p, x_, y_, z_, x, y);
}

/* ------------------------------------------------------------------------
 * Result ROI: RotationWarperBase::detectResultRoi / detectResultRoiByBorder,
 * PlaneWarper / SphericalWarper / SphericalPortraitWarper::detectResultRoi
 * (stitching/src/warpers.cpp). out = {tl.x, tl.y, br.x, br.y} (inclusive).
 * ---------------------------------------------------------------------- */
typedef struct {
  float tl_u, tl_v, br_u, br_v;
} wo_ext;

static inline void ext_init(wo_ext* e) {
  e->tl_u = e->tl_v = 3.402823466e+38f;
  e->br_u = e->br_v = -3.402823466e+38f;
}
static inline void ext_add(wo_ext* e, float u, float v) {
  /* std::min(a, b) = b < a ? b : a, std::max(a, b) = a < b ? b : a */
  e->tl_u = u < e->tl_u ? u : e->tl_u;
  e->tl_v = v < e->tl_v ? v : e->tl_v;
  e->br_u = e->br_u < u ? u : e->br_u;
  e->br_v = e->br_v < v ? v : e->br_v;
}

static void roi_by_border(const wo_projector* p, int w, int h, wo_ext* e) {
  float u, v;
  for (int i = 0; i < w; ++i) {
    wo_map_forward(p, (float)i, 0, &u, &v);
    ext_add(e, u, v);
    wo_map_forward(p, (float)i, (float)(h - 1), &u, &v);
    ext_add(e, u, v);
  }
  for (int i = 0; i < h; ++i) {
    wo_map_forward(p, 0, (float)i, &u, &v);
    ext_add(e, u, v);
    wo_map_forward(p, (float)(w - 1), (float)i, &u, &v);
    ext_add(e, u, v);
  }
}

void wo_detect_roi(const wo_projector* p, int w, int h, int out[4]) {
  wo_ext e;
  float u, v;
  ext_init(&e);
  int by_border = (p->type == WO_CYLINDRICAL || p->type == WO_SPHERICAL);
  if (p->type == WO_PLANE && !p->portrait) {
    wo_map_forward(p, 0, 0, &u, &v);
    ext_add(&e, u, v);
    wo_map_forward(p, 0, (float)(h - 1), &u, &v);
    ext_add(&e, u, v);
    wo_map_forward(p, (float)(w - 1), 0, &u, &v);
    ext_add(&e, u, v);
    wo_map_forward(p, (float)(w - 1), (float)(h - 1), &u, &v);
    ext_add(&e, u, v);
  } else if (by_border) {
    roi_by_border(p, w, h, &e);
  } else {
    for (int y = 0; y < h; ++y)
      for (int x = 0; x < w; ++x) {
        wo_map_forward(p, (float)x, (float)y, &u, &v);
        ext_add(&e, u, v);
      }
  }
  int tlx = (int)e.tl_u, tly = (int)e.tl_v, brx = (int)e.br_u,
      bry = (int)e.br_v;
  if (p->type == WO_SPHERICAL) {
    /* pole handling, SphericalWarper::detectResultRoi */
    float tl_uf = (float)tlx, tl_vf = (float)tly, br_uf = (float)brx,
          br_vf = (float)bry;
    const int c = p->portrait ? 0 : 1; /* rinv column used for the pole */
    for (int pole = 0; pole < 2; ++pole) {
      float x = p->rinv[c];
      float y = pole == 0 ? p->rinv[3 + c] : -p->rinv[3 + c];
      float z = p->rinv[6 + c];
      if (y > 0.f) {
        float x_ = (p->k[0] * x + p->k[1] * y) / z + p->k[2];
        float y_ = p->k[4] * y / z + p->k[5];
        if (x_ > 0.f && x_ < w && y_ > 0.f && y_ < h) {
          float pv = pole == 0 ? (float)(WO_PI * p->scale) : 0.f;
          tl_uf = tl_uf < 0.f ? tl_uf : 0.f;
          tl_vf = tl_vf < pv ? tl_vf : pv;
          br_uf = br_uf > 0.f ? br_uf : 0.f;
          br_vf = br_vf > pv ? br_vf : pv;
        }
      }
    }
    tlx = (int)tl_uf;
    tly = (int)tl_vf;
    brx = (int)br_uf;
    bry = (int)br_vf;
  }
  out[0] = tlx;
  out[1] = tly;
  out[2] = brx;
  out[3] = bry;
}

/* RotationWarperBase::buildMaps: maps are (br-tl+1) rows/cols */
void wo_build_maps(const wo_projector* p, const int roi[4], float* xmap,
                   float* ymap) {
  const int W = roi[2] - roi[0] + 1, H = roi[3] - roi[1] + 1;
  for (int v = 0; v < H; ++v)
    for (int u = 0; u < W; ++u)
      wo_map_backward(p, (float)(u + roi[0]), (float)(v + roi[1]),
                      &xmap[(size_t)v * W + u], &ymap[(size_t)v * W + u]);
}

/* ------------------------------------------------------------------------
 * remap (imgproc/src/imgwarp.cpp). Float maps are converted per pixel to
 * 1/32-px fixed point with round-half-even (cvRound / cvtps_epi32):
 *   sx = rint(x*32), ix = sx >> 5, fx = sx & 31.
 * 8-bit INTER_LINEAR: 4 integer weights at 15 bits, (sum + 2^14) >> 15.
 * 16-bit INTER_LINEAR: float weights from the 1-D float tabs, rint.
 * BORDER_REFLECT = fedcba|abcdefgh|hgfedcb (borderInterpolate).
 * ---------------------------------------------------------------------- */
static inline int reflect(int p, int len) {
  if ((unsigned)p < (unsigned)len) return p;
  if (len == 1) return 0;
  do {
    if (p < 0)
      p = -p - 1;
    else
      p = len - 1 - (p - len);
  } while ((unsigned)p >= (unsigned)len);
  return p;
}

static inline int sat_short(long v) {
  return v < -32768 ? -32768 : (v > 32767 ? 32767 : (int)v);
}

static inline long round_coord(float x) {
  float t = x * 32.0f;
  /* cvtps_epi32 / lrintf: out-of-range and NaN give INT_MIN */
  if (!(t > -2147483648.0f && t < 2147483648.0f)) return -2147483648L;
  return lrintf(t);
}

void wo_remap_linear_reflect(const void* src, int sw, int sh, int cn, int bpp,
                             const float* xmap, const float* ymap, int dw,
                             int dh, void* dst) {
  for (size_t i = 0; i < (size_t)dw * dh; ++i) {
    long sx = round_coord(xmap[i]), sy = round_coord(ymap[i]);
    int fx = (int)(sx & 31), fy = (int)(sy & 31);
    int ix = sat_short(sx >> 5), iy = sat_short(sy >> 5);
    int x0 = reflect(ix, sw), x1 = reflect(ix + 1, sw);
    int y0 = reflect(iy, sh), y1 = reflect(iy + 1, sh);
    for (int c = 0; c < cn; ++c) {
      if (bpp == 8) {
        const uint8_t* s = (const uint8_t*)src;
        int w00 = (32 - fy) * (32 - fx) * 32, w01 = (32 - fy) * fx * 32,
            w10 = fy * (32 - fx) * 32, w11 = fy * fx * 32;
        int p00 = s[((size_t)y0 * sw + x0) * cn + c],
            p01 = s[((size_t)y0 * sw + x1) * cn + c],
            p10 = s[((size_t)y1 * sw + x0) * cn + c],
            p11 = s[((size_t)y1 * sw + x1) * cn + c];
        int v = (p00 * w00 + p01 * w01 + p10 * w10 + p11 * w11 + (1 << 14)) >>
                15;
        ((uint8_t*)dst)[i * cn + c] = (uint8_t)(v > 255 ? 255 : v);
      } else {
        const uint16_t* s = (const uint16_t*)src;
        float ax1 = (float)fx * (1.f / 32), ax0 = 1.f - ax1;
        float ay1 = (float)fy * (1.f / 32), ay0 = 1.f - ay1;
        float w00 = ay0 * ax0, w01 = ay0 * ax1, w10 = ay1 * ax0,
              w11 = ay1 * ax1;
        float p00 = s[((size_t)y0 * sw + x0) * cn + c],
              p01 = s[((size_t)y0 * sw + x1) * cn + c],
              p10 = s[((size_t)y1 * sw + x0) * cn + c],
              p11 = s[((size_t)y1 * sw + x1) * cn + c];
        float v = p00 * w00 + p01 * w01 + p10 * w10 + p11 * w11;
        long q = lrintf(v);
        ((uint16_t*)dst)[i * cn + c] =
            (uint16_t)(q < 0 ? 0 : (q > 65535 ? 65535 : q));
      }
    }
  }
}

/* INTER_NEAREST + BORDER_CONSTANT(0) of an all-255 mask
 * (stitcher.cc:382-385): 255 where rint(x), rint(y) is inside the source */
void wo_remap_nearest_mask(int sw, int sh, const float* xmap,
                           const float* ymap, int dw, int dh, uint8_t* dst) {
  for (size_t i = 0; i < (size_t)dw * dh; ++i) {
    float x = xmap[i], y = ymap[i];
    long sx = (x > -2147483648.0f && x < 2147483648.0f) ? lrintf(x)
                                                        : -2147483648L;
    long sy = (y > -2147483648.0f && y < 2147483648.0f) ? lrintf(y)
                                                        : -2147483648L;
    int ix = sat_short(sx), iy = sat_short(sy);
    dst[i] = (ix >= 0 && ix < sw && iy >= 0 && iy < sh) ? 255 : 0;
  }
}

/* ------------------------------------------------------------------------
 * resize(INTER_LINEAR_EXACT), 8-bit single channel (imgproc/src/resize.cpp,
 * resize_bitExact with ufixedpoint16 = 8.8 coefficients):
 *   fval = scale*(d+0.5)-0.5 (double), ofs = floor, c1 = rint(frac*256),
 *   c0 = 256-c1; out of range on the low side or a 1-px source -> (0, 256,0),
 *   on the high side -> (size-1, 256, 0);
 *   dst = (cy0*(cx0*p00+cx1*p01) + cy1*(cx0*p10+cx1*p11) + 2^15) >> 16.
 * Used for the seam-mask upsample (stitcher.cc:398-399).
 * ---------------------------------------------------------------------- */
static void exact_coeffs(int srcsize, int dstsize, int* ofs, int* c0,
                         int* c1) {
  double inv_scale = (double)dstsize / srcsize;
  double scale = 1.0 / inv_scale;
  for (int d = 0; d < dstsize; ++d) {
    double fval = scale * ((double)d + 0.5) - 0.5;
    int ival = (int)floor(fval);
    if (ival >= 0 && srcsize > 1) {
      if (ival < srcsize - 1) {
        ofs[d] = ival;
        c1[d] = (int)lrint((fval - (double)ival) * 256.0);
        c0[d] = 256 - c1[d];
      } else {
        ofs[d] = srcsize - 1;
        c0[d] = 256;
        c1[d] = 0;
      }
    } else {
      ofs[d] = 0;
      c0[d] = 256;
      c1[d] = 0;
    }
  }
}

void wo_resize_linear_exact_u8(const uint8_t* src, int sw, int sh,
                               uint8_t* dst, int dw, int dh) {
  int* xo = (int*)malloc(sizeof(int) * dw * 3);
  int* yo = (int*)malloc(sizeof(int) * dh * 3);
  exact_coeffs(sw, dw, xo, xo + dw, xo + 2 * dw);
  exact_coeffs(sh, dh, yo, yo + dh, yo + 2 * dh);
  for (int y = 0; y < dh; ++y) {
    int oy = yo[y], cy0 = yo[dh + y], cy1 = yo[2 * dh + y];
    const uint8_t* r0 = src + (size_t)oy * sw;
    const uint8_t* r1 = src + (size_t)(oy + 1 < sh ? oy + 1 : oy) * sw;
    for (int x = 0; x < dw; ++x) {
      int ox = xo[x], cx0 = xo[dw + x], cx1 = xo[2 * dw + x];
      int ox1 = ox + 1 < sw ? ox + 1 : ox;
      uint32_t h0 = (uint32_t)(cx0 * r0[ox] + cx1 * r0[ox1]);
      uint32_t h1 = (uint32_t)(cx0 * r1[ox] + cx1 * r1[ox1]);
      uint32_t v = (uint32_t)cy0 * h0 + (uint32_t)cy1 * h1;
      v = (v + (1u << 15)) >> 16;
      dst[(size_t)y * dw + x] = (uint8_t)(v > 255 ? 255 : v);
    }
  }
  free(xo);
  free(yo);
}

/* dilate(src, dst, Mat()): 3x3 rectangle, pixels outside the image ignored
 * (stitcher.cc:397) */
void wo_dilate3x3_u8(const uint8_t* src, int w, int h, uint8_t* dst) {
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      uint8_t m = 0;
      for (int dy = -1; dy <= 1; ++dy)
        for (int dx = -1; dx <= 1; ++dx) {
          int yy = y + dy, xx = x + dx;
          if (yy < 0 || yy >= h || xx < 0 || xx >= w) continue;
          uint8_t v = src[(size_t)yy * w + xx];
          if (v > m) m = v;
        }
      dst[(size_t)y * w + x] = m;
    }
}

/* resize(INTER_LINEAR) for float32, 1 channel (resize.cpp resizeGeneric_,
 * HResizeLinear / VResizeLinear): used by BlocksCompensator::apply to bring
 * the gain map to the warped image size (exposure_compensate.cpp). */
static void linear_coeffs_f32(int srcsize, int dstsize, int* ofs, float* a1) {
  double scale = 1.0 / ((double)dstsize / srcsize);
  for (int d = 0; d < dstsize; ++d) {
    float f = (float)(((double)d + 0.5) * scale - 0.5);
    int s = (int)floorf(f);
    f -= (float)s;
    if (s < 0) {
      f = 0;
      s = 0;
    }
    if (s >= srcsize - 1) {
      f = 0;
      s = srcsize - 1;
    }
    ofs[d] = s;
    a1[d] = f;
  }
}

void wo_resize_linear_f32(const float* src, int sw, int sh, float* dst,
                          int dw, int dh) {
  int* xo = (int*)malloc(sizeof(int) * dw);
  int* yo = (int*)malloc(sizeof(int) * dh);
  float* xa = (float*)malloc(sizeof(float) * dw);
  float* ya = (float*)malloc(sizeof(float) * dh);
  linear_coeffs_f32(sw, dw, xo, xa);
  linear_coeffs_f32(sh, dh, yo, ya);
  for (int y = 0; y < dh; ++y) {
    int oy = yo[y];
    int oy1 = oy + 1 < sh ? oy + 1 : oy;
    float b1 = ya[y], b0 = 1.f - b1;
    const float* r0 = src + (size_t)oy * sw;
    const float* r1 = src + (size_t)oy1 * sw;
    for (int x = 0; x < dw; ++x) {
      int ox = xo[x];
      int ox1 = ox + 1 < sw ? ox + 1 : ox;
      float al1 = xa[x], al0 = 1.f - al1;
      float h0 = r0[ox] * al0 + r0[ox1] * al1;
      float h1 = r1[ox] * al0 + r1[ox1] * al1;
      dst[(size_t)y * dw + x] = h0 * b0 + h1 * b1;
    }
  }
  free(xo);
  free(yo);
  free(xa);
  free(ya);
}

/* multiply(CV_8UC3 image, CV_32FC3 gain, image): work type float,
 * saturate_cast<uchar>(rint(a*g)) */
void wo_apply_gain_u8(uint8_t* img, int w, int h, const float* gain) {
  for (size_t i = 0; i < (size_t)w * h; ++i) {
    float g = gain[i];
    for (int c = 0; c < 3; ++c) {
      float v = (float)img[i * 3 + c] * g;
      long q = lrintf(v);
      img[i * 3 + c] = (uint8_t)(q < 0 ? 0 : (q > 255 ? 255 : q));
    }
  }
}
