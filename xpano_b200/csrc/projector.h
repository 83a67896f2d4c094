// Projection surfaces of the warp: forward / backward point maps for the nine
// xpano projections and their portrait variants.
//
// Behavioural spec: cv::detail::ProjectorBase::setCameraParams and the
// *Projector::mapForward / mapBackward of OpenCV's stitching module, as used
// through cv::detail::RotationWarper by xpano/algorithm/stitcher.cc:160,377,384
// with the creators of xpano/algorithm/algorithm.cc:43-105 and
// xpano/algorithm/warpers.h:9-31. All arithmetic is
// float32 with every multiply and add rounded separately (the reference is
// built without FMA contraction).
//
// The same source serves the host (warpPoint, border ROI detection: scalar
// work the reference also does on the host, with libm floats) and the device
// (per-pixel maps). On the device transcendental functions are evaluated in
// double and rounded once to float, which reproduces libm's float result
// except for rare last-bit differences ("warp parity").
#pragma once

#include <cmath>
#include <cstdint>

#include "hd.h"

namespace xp {

struct Projector {
  int type, portrait;
  float scale, a, b;
  float k[9], rinv[9], r_kinv[9], k_rinv[9], t[3];
};

#define XP_PI_D 3.14159265358979323846 /* CV_PI */
#define XP_PI_F ((float)XP_PI_D)

#if defined(__CUDA_ARCH__)
// separately rounded float ops (nvcc would otherwise contract to FMA)
#define FMUL(a, b) __fmul_rn((a), (b))
#define FADD(a, b) __fadd_rn((a), (b))
#define FSUB(a, b) __fsub_rn((a), (b))
#define FDIV(a, b) __fdiv_rn((a), (b))
#define FSQRT(a) __fsqrt_rn((a))
#define TR1(fn, x) __double2float_rn(fn((double)(x)))
#define TR2(fn, x, y) __double2float_rn(fn((double)(x), (double)(y)))
#define XSIN(x) TR1(sin, x)
#define XCOS(x) TR1(cos, x)
#define XTAN(x) TR1(tan, x)
#define XATAN(x) TR1(atan, x)
#define XASIN(x) TR1(asin, x)
#define XACOS(x) TR1(acos, x)
#define XSINH(x) TR1(sinh, x)
#define XCOSH(x) TR1(cosh, x)
#define XLOG(x) TR1(log, x)
#define XATAN2(y, x) TR2(atan2, y, x)
#else
#define FMUL(a, b) ((float)((float)(a) * (float)(b)))
#define FADD(a, b) ((float)((float)(a) + (float)(b)))
#define FSUB(a, b) ((float)((float)(a) - (float)(b)))
#define FDIV(a, b) ((float)((float)(a) / (float)(b)))
#define FSQRT(a) sqrtf(a)
#define XSIN(x) sinf(x)
#define XCOS(x) cosf(x)
#define XTAN(x) tanf(x)
#define XATAN(x) atanf(x)
#define XASIN(x) asinf(x)
#define XACOS(x) acosf(x)
#define XSINH(x) sinhf(x)
#define XCOSH(x) coshf(x)
#define XLOG(x) logf(x)
#define XATAN2(y, x) atan2f(y, x)
#endif

// (m0*x + m1*y) + m2*z, each op rounded
XP_HD float dot3(const float* m, float x, float y, float z) {
  return FADD(FADD(FMUL(m[0], x), FMUL(m[1], y)), FMUL(m[2], z));
}

// r_kinv applied to (x, y, 1): m0*x + m1*y + m2
XP_HD void apply_rkinv(const Projector& p, float x, float y, float& x_,
                       float& y_, float& z_) {
  const float* m = p.r_kinv;
  x_ = FADD(FADD(FMUL(m[0], x), FMUL(m[1], y)), m[2]);
  y_ = FADD(FADD(FMUL(m[3], x), FMUL(m[4], y)), m[5]);
  z_ = FADD(FADD(FMUL(m[6], x), FMUL(m[7], y)), m[8]);
}

// k_rinv applied to a direction, perspective divide; (-1,-1) behind the camera
XP_HD void project_dir(const Projector& p, float x_, float y_, float z_,
                       float& x, float& y) {
  const float* m = p.k_rinv;
  x = dot3(m, x_, y_, z_);
  y = dot3(m + 3, x_, y_, z_);
  float z = dot3(m + 6, x_, y_, z_);
  if (z > 0) {
    x = FDIV(x, z);
    y = FDIV(y, z);
  } else {
    x = y = -1.f;
  }
}

XP_HD float norm3(float x, float y, float z) {
  return FSQRT(FADD(FADD(FMUL(x, x), FMUL(y, y)), FMUL(z, z)));
}

// mapForward: source pixel (x, y) -> surface (u, v)
XP_HD void map_forward(const Projector& p, float x, float y, float& u,
                       float& v) {
  float x_, y_, z_;
  apply_rkinv(p, x, y, x_, y_, z_);
  const float scale = p.scale, a = p.a, b = p.b;
  if (p.portrait) {
    const float x0 = x_, y0 = y_;
    switch (p.type) {
      case XP_PROJ_PERSPECTIVE: {
        float xx = FADD(p.t[0], FMUL(FDIV(y0, z_), FSUB(1.f, p.t[2])));
        float yy = FADD(p.t[1], FMUL(FDIV(x0, z_), FSUB(1.f, p.t[2])));
        u = -FMUL(scale, xx);
        v = FMUL(scale, yy);
        return;
      }
      case XP_PROJ_CYLINDRICAL: {
        float xx = y0, yy = x0;
        float uu = FMUL(scale, XATAN2(xx, z_));
        v = FDIV(FMUL(scale, yy), FSQRT(FADD(FMUL(xx, xx), FMUL(z_, z_))));
        u = -uu;
        return;
      }
      case XP_PROJ_SPHERICAL: {
        float xx = y0, yy = x0;
        float uu = FMUL(scale, XATAN2(xx, z_));
        v = FMUL(scale, FSUB(XP_PI_F, XACOS(FDIV(yy, norm3(xx, yy, z_)))));
        u = -uu;
        return;
      }
      case XP_PROJ_COMPRESSED_RECTILINEAR: {
        float yy = x0, xx = y0;
        float u_ = XATAN2(xx, z_);
        float v_ = XASIN(FDIV(yy, norm3(xx, yy, z_)));
        u = FMUL(FMUL(-scale, a), XTAN(FDIV(u_, a)));
        v = FDIV(FMUL(FMUL(scale, b), XTAN(v_)), XCOS(u_));
        return;
      }
      case XP_PROJ_PANINI: {
        float yy = x0, xx = y0;
        float u_ = XATAN2(xx, z_);
        float v_ = XASIN(FDIV(yy, norm3(xx, yy, z_)));
        float tg = FMUL(a, XTAN(FDIV(u_, a)));
        u = FMUL(-scale, tg);
        float sinu = XSIN(u_);
        if (fabsf(sinu) < 1E-7)
          v = FMUL(FMUL(scale, b), XTAN(v_));
        else
          v = FDIV(FMUL(FMUL(FMUL(scale, b), tg), XTAN(v_)), sinu);
        return;
      }
      default:
        break;  // no portrait variant: landscape projector
    }
  }
  switch (p.type) {
    case XP_PROJ_PERSPECTIVE: {
      float xx = FADD(p.t[0], FMUL(FDIV(x_, z_), FSUB(1.f, p.t[2])));
      float yy = FADD(p.t[1], FMUL(FDIV(y_, z_), FSUB(1.f, p.t[2])));
      u = FMUL(scale, xx);
      v = FMUL(scale, yy);
      return;
    }
    case XP_PROJ_CYLINDRICAL:
      u = FMUL(scale, XATAN2(x_, z_));
      v = FDIV(FMUL(scale, y_), FSQRT(FADD(FMUL(x_, x_), FMUL(z_, z_))));
      return;
    case XP_PROJ_SPHERICAL: {
      u = FMUL(scale, XATAN2(x_, z_));
      float w = FDIV(y_, norm3(x_, y_, z_));
      v = FMUL(scale, FSUB(XP_PI_F, XACOS(w == w ? w : 0.f)));
      return;
    }
    case XP_PROJ_FISHEYE: {
      float u_ = XATAN2(x_, z_);
      float v_ = FSUB(XP_PI_F, XACOS(FDIV(y_, norm3(x_, y_, z_))));
      u = FMUL(FMUL(scale, v_), XCOS(u_));
      v = FMUL(FMUL(scale, v_), XSIN(u_));
      return;
    }
    case XP_PROJ_STEREOGRAPHIC: {
      float u_ = XATAN2(x_, z_);
      float v_ = FSUB(XP_PI_F, XACOS(FDIV(y_, norm3(x_, y_, z_))));
      float r = FDIV(XSIN(v_), FSUB(1.f, XCOS(v_)));
      u = FMUL(FMUL(scale, r), XCOS(u_));
      v = FMUL(FMUL(scale, r), XSIN(u_));
      return;
    }
    case XP_PROJ_COMPRESSED_RECTILINEAR: {
      float u_ = XATAN2(x_, z_);
      float v_ = XASIN(FDIV(y_, norm3(x_, y_, z_)));
      u = FMUL(FMUL(scale, a), XTAN(FDIV(u_, a)));
      v = FDIV(FMUL(FMUL(scale, b), XTAN(v_)), XCOS(u_));
      return;
    }
    case XP_PROJ_PANINI: {
      float u_ = XATAN2(x_, z_);
      float v_ = XASIN(FDIV(y_, norm3(x_, y_, z_)));
      float tg = FMUL(a, XTAN(FDIV(u_, a)));
      u = FMUL(scale, tg);
      float sinu = XSIN(u_);
      if (fabsf(sinu) < 1E-7)
        v = FMUL(FMUL(scale, b), XTAN(v_));
      else
        v = FDIV(FMUL(FMUL(FMUL(scale, b), tg), XTAN(v_)), sinu);
      return;
    }
    case XP_PROJ_MERCATOR: {
      float u_ = XATAN2(x_, z_);
      float v_ = XASIN(FDIV(y_, norm3(x_, y_, z_)));
      u = FMUL(scale, u_);
      v = FMUL(scale, XLOG(XTAN(FADD((float)(XP_PI_D / 4), FDIV(v_, 2.f)))));
      return;
    }
    default: {  // XP_PROJ_TRANSVERSE_MERCATOR
      float u_ = XATAN2(x_, z_);
      float v_ = XASIN(FDIV(y_, norm3(x_, y_, z_)));
      float B = FMUL(XCOS(v_), XSIN(u_));
      u = FMUL(FDIV(scale, 2.f), XLOG(FDIV(FADD(1.f, B), FSUB(1.f, B))));
      v = FMUL(scale, XATAN2(XTAN(v_), XCOS(u_)));
      return;
    }
  }
}

// mapBackward: surface (u, v) -> source pixel (x, y)
XP_HD void map_backward(const Projector& p, float u, float v, float& x,
                        float& y) {
  const float scale = p.scale, a = p.a, b = p.b;
  float x_, y_, z_;
  if (p.portrait) {
    switch (p.type) {
      case XP_PROJ_PERSPECTIVE: {
        const float* m = p.k_rinv;
        float uu = FSUB(FDIV(-u, scale), p.t[0]);
        float vv = FSUB(FDIV(v, scale), p.t[1]);
        float w = FSUB(1.f, p.t[2]);
        x = dot3(m, vv, uu, w);
        y = dot3(m + 3, vv, uu, w);
        float z = dot3(m + 6, vv, uu, w);
        x = FDIV(x, z);
        y = FDIV(y, z);
        return;
      }
      case XP_PROJ_CYLINDRICAL: {
        float uu = FDIV(-u, scale), vv = FDIV(v, scale);
        project_dir(p, vv, XSIN(uu), XCOS(uu), x, y);
        return;
      }
      case XP_PROJ_SPHERICAL: {
        float uu = FDIV(-u, scale), vv = FDIV(v, scale);
        float sinv = XSIN(FSUB(XP_PI_F, vv));
        float x0 = FMUL(sinv, XSIN(uu));
        float y0 = XCOS(FSUB(XP_PI_F, vv));
        z_ = FMUL(sinv, XCOS(uu));
        project_dir(p, y0, x0, z_, x, y);
        return;
      }
      case XP_PROJ_COMPRESSED_RECTILINEAR: {
        u = FDIV(u, -scale);
        v = FDIV(v, scale);
        float u_ = FMUL(a, XATAN(FDIV(u, a)));
        float v_ = XATAN(FDIV(FMUL(v, XCOS(u_)), b));
        float cosv = XCOS(v_);
        y_ = FMUL(cosv, XSIN(u_));
        x_ = XSIN(v_);
        z_ = FMUL(cosv, XCOS(u_));
        project_dir(p, x_, y_, z_, x, y);
        return;
      }
      case XP_PROJ_PANINI: {
        u = FDIV(u, -scale);
        v = FDIV(v, scale);
        float lamda = FMUL(a, XATAN(FDIV(u, a)));
        float u_ = lamda, v_;
        if (fabsf(lamda) > 1E-7)
          v_ = XATAN(FDIV(FMUL(v, XSIN(lamda)),
                          FMUL(FMUL(b, a), XTAN(FDIV(lamda, a)))));
        else
          v_ = XATAN(FDIV(v, b));
        float cosv = XCOS(v_);
        y_ = FMUL(cosv, XSIN(u_));
        x_ = XSIN(v_);
        z_ = FMUL(cosv, XCOS(u_));
        project_dir(p, x_, y_, z_, x, y);
        return;
      }
      default:
        break;
    }
  }
  switch (p.type) {
    case XP_PROJ_PERSPECTIVE: {
      const float* m = p.k_rinv;
      u = FSUB(FDIV(u, scale), p.t[0]);
      v = FSUB(FDIV(v, scale), p.t[1]);
      float w = FSUB(1.f, p.t[2]);
      x = dot3(m, u, v, w);
      y = dot3(m + 3, u, v, w);
      float z = dot3(m + 6, u, v, w);
      x = FDIV(x, z);
      y = FDIV(y, z);
      return;
    }
    case XP_PROJ_CYLINDRICAL:
      u = FDIV(u, scale);
      v = FDIV(v, scale);
      x_ = XSIN(u);
      y_ = v;
      z_ = XCOS(u);
      break;
    case XP_PROJ_SPHERICAL: {
      u = FDIV(u, scale);
      v = FDIV(v, scale);
      float sinv = XSIN(FSUB(XP_PI_F, v));
      x_ = FMUL(sinv, XSIN(u));
      y_ = XCOS(FSUB(XP_PI_F, v));
      z_ = FMUL(sinv, XCOS(u));
      break;
    }
    case XP_PROJ_FISHEYE: {
      u = FDIV(u, scale);
      v = FDIV(v, scale);
      float u_ = XATAN2(v, u);
      float v_ = FSQRT(FADD(FMUL(u, u), FMUL(v, v)));
      float sinv = XSIN(FSUB(XP_PI_F, v_));
      x_ = FMUL(sinv, XSIN(u_));
      y_ = XCOS(FSUB(XP_PI_F, v_));
      z_ = FMUL(sinv, XCOS(u_));
      break;
    }
    case XP_PROJ_STEREOGRAPHIC: {
      u = FDIV(u, scale);
      v = FDIV(v, scale);
      float u_ = XATAN2(v, u);
      float r = FSQRT(FADD(FMUL(u, u), FMUL(v, v)));
      float v_ = FMUL(2.f, XATAN(FDIV(1.f, r)));
      float sinv = XSIN(FSUB(XP_PI_F, v_));
      x_ = FMUL(sinv, XSIN(u_));
      y_ = XCOS(FSUB(XP_PI_F, v_));
      z_ = FMUL(sinv, XCOS(u_));
      break;
    }
    case XP_PROJ_COMPRESSED_RECTILINEAR: {
      u = FDIV(u, scale);
      v = FDIV(v, scale);
      float u_ = FMUL(a, XATAN(FDIV(u, a)));
      float v_ = XATAN(FDIV(FMUL(v, XCOS(u_)), b));
      float cosv = XCOS(v_);
      x_ = FMUL(cosv, XSIN(u_));
      y_ = XSIN(v_);
      z_ = FMUL(cosv, XCOS(u_));
      break;
    }
    case XP_PROJ_PANINI: {
      u = FDIV(u, scale);
      v = FDIV(v, scale);
      float lamda = FMUL(a, XATAN(FDIV(u, a)));
      float u_ = lamda, v_;
      if (fabsf(lamda) > 1E-7)
        v_ = XATAN(FDIV(FMUL(v, XSIN(lamda)),
                        FMUL(FMUL(b, a), XTAN(FDIV(lamda, a)))));
      else
        v_ = XATAN(FDIV(v, b));
      float cosv = XCOS(v_);
      x_ = FMUL(cosv, XSIN(u_));
      y_ = XSIN(v_);
      z_ = FMUL(cosv, XCOS(u_));
      break;
    }
    case XP_PROJ_MERCATOR: {
      u = FDIV(u, scale);
      v = FDIV(v, scale);
      float v_ = XATAN(XSINH(v));
      float cosv = XCOS(v_);
      x_ = FMUL(cosv, XSIN(u));
      y_ = XSIN(v_);
      z_ = FMUL(cosv, XCOS(u));
      break;
    }
    default: {  // XP_PROJ_TRANSVERSE_MERCATOR
      u = FDIV(u, scale);
      v = FDIV(v, scale);
      float v_ = XASIN(FDIV(XSIN(v), XCOSH(u)));
      float u_ = XATAN2(XSINH(u), XCOS(v));
      float cosv = XCOS(v_);
      x_ = FMUL(cosv, XSIN(u_));
      y_ = XSIN(v_);
      z_ = FMUL(cosv, XCOS(u_));
      break;
    }
  }
  project_dir(p, x_, y_, z_, x, y);
}

// Separable projections: direction = (A_row * S_col, B_row, A_row * C_col)
// (or with x and y swapped for portrait). Lets the warp kernel read two small
// tables instead of evaluating trig per pixel.
XP_HD bool is_separable(int type) {
  return type == XP_PROJ_SPHERICAL || type == XP_PROJ_CYLINDRICAL ||
         type == XP_PROJ_MERCATOR;
}

// column table entry for surface abscissa u (already in pixels): (S, C)
XP_HD void sep_col(const Projector& p, float u, float& S, float& C) {
  float uu = (p.portrait && p.type != XP_PROJ_MERCATOR) ? FDIV(-u, p.scale)
                                                        : FDIV(u, p.scale);
  S = XSIN(uu);
  C = XCOS(uu);
}

// row table entry for surface ordinate v: (A, B)
XP_HD void sep_row(const Projector& p, float v, float& A, float& B) {
  float vv = FDIV(v, p.scale);
  if (p.type == XP_PROJ_SPHERICAL) {
    A = XSIN(FSUB(XP_PI_F, vv));
    B = XCOS(FSUB(XP_PI_F, vv));
  } else if (p.type == XP_PROJ_CYLINDRICAL) {
    A = 1.f;
    B = vv;
  } else {  // mercator
    float v_ = XATAN(XSINH(vv));
    A = XCOS(v_);
    B = XSIN(v_);
  }
}

XP_HD void sep_combine(const Projector& p, float A, float B, float S, float C,
                       float& x, float& y) {
  float xs = FMUL(A, S), zs = FMUL(A, C);
  if (p.portrait && p.type != XP_PROJ_MERCATOR)
    project_dir(p, B, xs, zs, x, y);
  else
    project_dir(p, xs, B, zs, x, y);
}

}  // namespace xp
