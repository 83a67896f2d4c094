// Compose driver: the full-resolution loop of Stitcher::ComposePanorama
// (xpano/algorithm/stitcher.cc:299-432) with ComputeRoi (:144-169) and the
// panorama size cap (:318-335). Cameras, gain maps and low-resolution seam
// masks are inputs (registration and seam estimation are outside the path).
// Sources are uploaded once; warp -> feed -> blend stays on the device.
#include <chrono>
#include <climits>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <condition_variable>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <thread>

#include "blender.h"
#include "common.h"
#include "warp.h"

namespace xp {

__global__ void noop_kernel() {}

struct SrcImage {
  int w = 0, h = 0;
  size_t pitch = 0;
  const void* dev = nullptr;  // device pointer (owned in `buf` unless adopted)
  DevBuf buf, gain, seam;
  int gw = 0, gh = 0, mw = 0, mh = 0;
  bool skip = false;  // countNonZero(seam mask) == 0 (stitcher.cc:360-364)
  float K[9], R[9];
};

// A few host threads that stay parked between panoramas: the per-image ROI
// walk is ~0.25 ms of scalar trig per image and nothing else can start before
// it, so it is spread over the images without paying thread creation per run.
class HostPool {
 public:
  ~HostPool() {
    {
      std::lock_guard<std::mutex> lk(mu_);
      stop_ = true;
    }
    cv_.notify_all();
    for (std::thread& t : threads_) t.join();
  }
  // runs fn(i) for i in [0, n) on the pool and the calling thread
  void parallel_for(int n, const std::function<void(int)>& fn) {
    if (n <= 1) {
      for (int i = 0; i < n; ++i) fn(i);
      return;
    }
    ensure(host_worker_threads(std::min(n - 1, 15)));
    {
      std::lock_guard<std::mutex> lk(mu_);
      fn_ = &fn;
      next_ = 0;
      total_ = n;
      pending_ = n;
      ++generation_;
    }
    cv_.notify_all();
    work();
    std::unique_lock<std::mutex> lk(mu_);
    done_.wait(lk, [&] { return pending_ == 0; });
    fn_ = nullptr;
  }

 private:
  void ensure(int k) {
    while ((int)threads_.size() < k)
      threads_.emplace_back([this] {
        unsigned long long seen = 0;
        for (;;) {
          {
            std::unique_lock<std::mutex> lk(mu_);
            cv_.wait(lk, [&] { return stop_ || generation_ != seen; });
            if (stop_) return;
            seen = generation_;
          }
          work();
        }
      });
  }
  void work() {
    for (;;) {
      const std::function<void(int)>* fn;
      int i;
      {
        std::lock_guard<std::mutex> lk(mu_);
        if (!fn_ || next_ >= total_) return;
        i = next_++;
        fn = fn_;
      }
      (*fn)(i);
      std::lock_guard<std::mutex> lk(mu_);
      if (--pending_ == 0) done_.notify_all();
    }
  }
  std::mutex mu_;
  std::condition_variable cv_, done_;
  std::vector<std::thread> threads_;
  const std::function<void(int)>* fn_ = nullptr;
  int next_ = 0, total_ = 0, pending_ = 0;
  unsigned long long generation_ = 0;
  bool stop_ = false;
};

struct Composer {
  HostPool pool;
  xp_compose_opts opts;
  Warper warper;
  std::unique_ptr<Blender> blender;
  std::vector<std::unique_ptr<SrcImage>> src;
  cudaStream_t stream = nullptr, own_stream = nullptr;
  bool external_stream = false, device_ready = false;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  double h2d_ms = 0;

  ~Composer() {
    // the caller's stream may be the legacy default stream (a null handle)
    if (device_ready) {
      int cur = 0;
      cudaGetDevice(&cur);
      if (cur != opts.blend.device) cudaSetDevice(opts.blend.device);
      cudaStreamSynchronize(stream);
      if (cur != opts.blend.device) cudaSetDevice(cur);
    }
    // members that remember the stream must go before it is destroyed
    blender.reset();
    src.clear();
    warper.stream = nullptr;
    warper.detached = true;
    for (cudaEvent_t e : ev)
      if (e) cudaEventDestroy(e);
    if (own_stream) cudaStreamDestroy(own_stream);
  }

  void ensure_device() {
    select_device(opts.blend.device);
    device_ready = true;
    if (!stream && !external_stream) {
      own_stream = create_stream(true);
      stream = own_stream;
    }
    for (cudaEvent_t& e : ev)
      if (!e) XP_CUDA(cudaEventCreate(&e));
    warper.type = opts.projection;
    warper.portrait = opts.portrait ? 1 : 0;
    warper.a = opts.a;
    warper.b = opts.b;
    warper.device = opts.blend.device;
    warper.stream = stream;
    warper.external_stream = true;
  }

  void set_images(int n, const xp_compose_image* images, int mem) {
    XP_REQUIRE(n >= 1 && images, "set_images: no images");
    XP_REQUIRE(opts.depth == 8 || opts.depth == 16, "depth must be 8 or 16");
    XP_REQUIRE(!tile_pending, "set_images between run_tile_begin and _end");
    XP_REQUIRE(!run_pending, "set_images between run_begin and run_end");
    ensure_device();
    src.clear();
    plan.valid = false;
    const size_t bs = opts.depth / 8;
    // timed on the host clock: an event recorded here would be queued on the
    // copy engine the stream used last (the previous panorama's download)
    // and wait there behind other composers' transfers
    const auto h2d_t0 = std::chrono::steady_clock::now();
    // cameras, gain maps and low-resolution seam masks of every image
    for (int i = 0; i < n; ++i) {
      const xp_compose_image& in = images[i];
      XP_REQUIRE(in.data && in.w > 0 && in.h > 0, "image %d is empty", i);
      std::unique_ptr<SrcImage> s(new SrcImage());
      s->w = in.w;
      s->h = in.h;
      memcpy(s->K, in.K, sizeof(s->K));
      memcpy(s->R, in.R, sizeof(s->R));
      if (in.aux.gain) {
        s->gw = in.aux.gain_w;
        s->gh = in.aux.gain_h;
        XP_REQUIRE(s->gw > 0 && s->gh > 0, "image %d: empty gain map", i);
        const size_t b = sizeof(float) * (size_t)s->gw * s->gh;
        s->gain.reserve(b);
        XP_CUDA(cudaMemcpyAsync(s->gain.p, in.aux.gain, b,
                                cudaMemcpyHostToDevice, stream));
      }
      if (in.aux.seam_mask) {
        s->mw = in.aux.seam_w;
        s->mh = in.aux.seam_h;
        XP_REQUIRE(s->mw > 0 && s->mh > 0, "image %d: empty seam mask", i);
        const size_t b = (size_t)s->mw * s->mh;
        bool any = false;
        for (size_t k = 0; k < b && !any; ++k) any = in.aux.seam_mask[k] != 0;
        s->skip = !any;
        s->seam.reserve(b);
        XP_CUDA(cudaMemcpyAsync(s->seam.p, in.aux.seam_mask, b,
                                cudaMemcpyHostToDevice, stream));
      }
      src.push_back(std::move(s));
    }
    // a canvas tile uploads only the sources whose colours reach its strip
    if (tile_count > 1) make_plan();
    for (int i = 0; i < n; ++i) {
      const xp_compose_image& in = images[i];
      SrcImage* s = src[i].get();
      if (tile_count > 1 && !plan.local[i]) continue;
      const size_t row = (size_t)in.w * 3 * bs;
      const size_t stride = in.stride ? in.stride : row;
      if (mem == XP_MEM_DEVICE) {
        s->dev = in.data;
        s->pitch = stride;
      } else {
        s->buf.reserve(row * (size_t)in.h);
        copy_in_2d(s->buf.p, row, in.data, stride, row, in.h, mem, stream);
        s->dev = s->buf.p;
        s->pitch = row;
      }
    }
    // ROIs and the size cap are host arithmetic on the cameras: done here,
    // under the uploads, instead of at the head of run()
    if (!plan.valid) make_plan();
    sync_blocking(stream);
    h2d_ms = std::chrono::duration<double, std::milli>(
                 std::chrono::steady_clock::now() - h2d_t0)
                 .count();
  }

  // ComputeRoi (stitcher.cc:144-169)
  void compute_roi(float scale, std::vector<int>& rois, int rect[4]) {
    warper.scale = scale;
    const int n = (int)src.size();
    rois.resize(4 * n);
    long x0 = LONG_MAX, y0 = LONG_MAX, x1 = LONG_MIN, y1 = LONG_MIN;
    // border-walk projections are pure host math (one forward map per
    // perimeter pixel): one thread per image
    const bool host_roi = warper.type == XP_PROJ_SPHERICAL ||
                          warper.type == XP_PROJ_CYLINDRICAL;
    if (host_roi) {
      // the walk of every image is cut in kParts pieces; merging the
      // pieces' min/max is exact
      constexpr int kParts = 4;
      std::vector<Projector> proj(n);
      std::vector<float> ext(4 * kParts * (size_t)n);
      for (int i = 0; i < n; ++i) proj[i] = warper.projector(src[i]->K, src[i]->R);
      pool.parallel_for(n * kParts, [&](int t) {
        const int i = t / kParts, part = t % kParts;
        Warper::border_extent(proj[i], src[i]->w, src[i]->h, part, kParts,
                              &ext[4 * (size_t)t]);
      });
      for (int i = 0; i < n; ++i) {
        float m[4] = {3.402823466e+38f, 3.402823466e+38f, -3.402823466e+38f,
                      -3.402823466e+38f};
        for (int k = 0; k < kParts; ++k) {
          const float* e = &ext[4 * ((size_t)i * kParts + k)];
          m[0] = e[0] < m[0] ? e[0] : m[0];
          m[1] = e[1] < m[1] ? e[1] : m[1];
          m[2] = m[2] < e[2] ? e[2] : m[2];
          m[3] = m[3] < e[3] ? e[3] : m[3];
        }
        int roi[4];
        warper.detect_roi(proj[i], src[i]->w, src[i]->h, roi, m);
        int* r = &rois[4 * i];
        r[0] = roi[0];
        r[1] = roi[1];
        r[2] = roi[2] - roi[0] + 1;
        r[3] = roi[3] - roi[1] + 1;
      }
    }
    for (int i = 0; i < n; ++i) {
      int* r = &rois[4 * i];
      if (!host_roi)
        warper.warp_roi(src[i]->w, src[i]->h, src[i]->K, src[i]->R, r);
      x0 = std::min<long>(x0, r[0]);
      y0 = std::min<long>(y0, r[1]);
      x1 = std::max<long>(x1, (long)r[0] + r[2]);
      y1 = std::max<long>(y1, (long)r[1] + r[3]);
    }
    rect[0] = (int)x0;
    rect[1] = (int)y0;
    rect[2] = (int)(x1 - x0);
    rect[3] = (int)(y1 - y0);
  }

  // ---- plan: ROIs, size cap, and (tiled) the strip of this rank ----------
  struct Plan {
    bool valid = false;
    float scale = 0;
    bool capped = false;
    std::vector<int> rois;
    int rect[4] = {0, 0, 0, 0};
    int abs_x0 = 0, abs_x1 = 0;   // owned columns (feed coordinates)
    std::vector<char> local;      // per source image: colours needed
  } plan;
  int tile_index = 0, tile_count = 1;
  bool tile_pending = false;
  const volatile int* cancel_flag = nullptr;
  void check_cancelled() {
    if (cancel_flag && *cancel_flag) {
      // nothing of this run may stay queued behind the caller's back
      if (blender) {
        cudaStreamSynchronize(stream);
        blender.reset();
      }
      fail(XP_CANCELLED, "cancelled (stitcher.cc:408-412)");
    }
  }
  double t_s_src = 0, t_s_w = 0;
  int t_warp_launches = 0;

  void make_plan() {
    // stitcher.cc:306-313: warp_scale = float(warped_image_scale_ *
    // compose_work_aspect), both doubles there
    double wis = opts.warped_image_scale, cwa = opts.compose_work_aspect;
    if (!(wis > 0)) {
      wis = (double)opts.warp_scale;
      cwa = 1.0;
    }
    plan.scale = (float)(wis * cwa);
    compute_roi(plan.scale, plan.rois, plan.rect);
    plan.capped = xp_compose_capped_scale(wis, cwa, plan.rect[2], plan.rect[3],
                                          opts.max_pano_mpx, &plan.scale) != 0;
    if (plan.capped) compute_roi(plan.scale, plan.rois, plan.rect);
    const int n = (int)src.size();
    plan.local.assign(n, 1);
    plan.abs_x0 = plan.rect[0];
    plan.abs_x1 = plan.rect[0] + plan.rect[2];
    if (tile_count > 1) {
      // Strips partition the ROI rectangle (known before any pixel is
      // warped; the trimmed canvas lies inside it). An image's colours are
      // needed when its ROI reaches into the strip widened by the halo, one
      // alignment unit and the slack of the level count (levels are fixed
      // after trim; from the ROI sizes they can only come out larger).
      plan.abs_x0 = plan.rect[0] + (int)((long)plan.rect[2] * tile_index / tile_count);
      plan.abs_x1 = plan.rect[0] + (int)((long)plan.rect[2] * (tile_index + 1) / tile_count);
      std::vector<int> ws, hs;
      for (int i = 0; i < n; ++i)
        if (!src[i]->skip) {
          ws.push_back(plan.rois[4 * i + 2]);
          hs.push_back(plan.rois[4 * i + 3]);
        }
      XP_REQUIRE(ws.size() >= 2, "canvas tiling needs at least two images");
      const int lv = blend_levels((int)ws.size(), ws.data(), hs.data(),
                                  opts.blend.fixed_levels,
                                  opts.blend.add_levels);
      XP_REQUIRE(lv >= 2 && lv <= kMaxLevels,
                 "canvas tiling needs 2..%d pyramid levels", kMaxLevels);
      const long margin = (long)tile_halo(lv) + (2L << lv);
      for (int i = 0; i < n; ++i) {
        const long x0 = plan.rois[4 * i], x1 = x0 + plan.rois[4 * i + 2];
        plan.local[i] =
            x1 > plan.abs_x0 - margin && x0 < plan.abs_x1 + margin ? 1 : 0;
      }
    }
    plan.valid = true;
  }

  // warp every image into a new blender (tiled: alpha only for the images
  // whose colours cannot reach this rank's strip)
  void warp_all() {
    xp_blend_opts bo = opts.blend;
    {  // the stream is idle here (the previous run ended, set_images waited):
       // the warper's host-side table slots of the previous run are free
      size_t tab = 0;
      for (size_t i = 0; i < src.size(); ++i)
        tab += 16 + sizeof(float2) * ((size_t)plan.rois[4 * i + 2] +
                                      (size_t)plan.rois[4 * i + 3]);
      warper.tab_reset(tab);
    }
    blender.reset(new Blender(bo));
    blender->stream = stream;
    blender->external_stream = true;
    for (int k = 0; k < 4; ++k) blender->roi[k] = plan.rect[k];
    warper.scale = plan.scale;
    t_s_src = t_s_w = 0;
    t_warp_launches = 0;
    const size_t bs = opts.depth / 8;
    std::vector<char> fed_local;
    for (size_t i = 0; i < src.size(); ++i) {
      SrcImage& s = *src[i];
      if (s.skip) continue;
      check_cancelled();
      const int* r = &plan.rois[4 * i];
      const int roi_incl[4] = {r[0], r[1], r[0] + r[2] - 1, r[1] + r[3] - 1};
      Fed* f = blender->new_fed(r[2], r[3], opts.depth, r[0], r[1]);
      const Projector p = warper.projector(s.K, s.R);
      const bool colours = plan.local[i] != 0;
      XP_REQUIRE(!colours || s.dev, "image %zu was not uploaded", i);
      f->abits_pitch = Warper::alpha_bits_pitch(r[2]);
      f->abits.reserve(sizeof(uint32_t) * (size_t)f->abits_pitch * r[3]);
      warper.warp_bgra_device(colours ? s.dev : nullptr, s.w, s.h, s.pitch,
                              opts.depth, p, roi_incl, s.gain.as<float>(),
                              s.gw, s.gh, s.seam.as<uint8_t>(), s.mw, s.mh,
                              f->data.p, (size_t)f->pitch * 4 * bs, true,
                              f->abits.as<uint32_t>(), f->abits_pitch);
      fed_local.push_back(colours ? 1 : 0);
      t_warp_launches += 1 + (s.seam.p ? 1 : 0) +
                         (is_separable(p.type) ? 1 : 0) +
                         (s.seam.p || (colours && s.gain.p) ? 1 : 0);
      if (colours) t_s_src += (double)s.w * s.h;
      t_s_w += (double)r[2] * r[3];
    }
    check_cancelled();
    if (tile_count > 1) {
      blender->tile.index = tile_index;
      blender->tile.count = tile_count;
      blender->tile.abs_x0 = plan.abs_x0;
      blender->tile.abs_x1 = plan.abs_x1;
      blender->fed_local = fed_local;
    }
  }

  void fill_info(xp_compose_info* info, const xp_timing& tm, float warp_ms,
                 float blend_ms) {
    if (!info) return;
    memset(info, 0, sizeof(*info));
    const bool tiled = tile_count > 1;
    info->width = tiled ? blender->own_x1 - blender->own_x0 : blender->W;
    info->height = blender->H;
    info->min_x = blender->min_x + (tiled ? blender->own_x0 : 0);
    info->min_y = blender->min_y;
    info->bpp = blender->out_bpp;
    info->levels = blender->levels;
    info->resolution_capped = plan.capped ? 1 : 0;
    memcpy(info->roi, plan.rect, sizeof(plan.rect));
    info->warp_scale = plan.scale;
    info->warp_ms = warp_ms;
    info->blend_ms = blend_ms;
    info->h2d_ms = h2d_ms;
    info->timing = tm;
    info->s_src = t_s_src;
    info->s_w = t_s_w;
    info->s_t = blender->s_t;
    info->c = tiled ? (double)info->width * info->height : blender->c_px;
    for (int k = 0; k < 9; ++k) {
      info->stage_bytes[k] = blender->stage_bytes[k];
      info->stage_launches[k] = blender->stage_launches[k];
    }
    const size_t bs = opts.depth / 8;
    info->warp_bytes = (3.0 * t_s_src + 4.0 * t_s_w) * bs;
    info->warp_launches = t_warp_launches;
    info->canvas_width = blender->W;
    info->canvas_height = blender->H;
    info->tile_x0 = tiled ? blender->own_x0 : 0;
    info->window_x0 = tiled ? blender->win_x0 : 0;
    info->window_width = tiled ? blender->win_w : blender->W;
  }

  // The stream's last operation before a run is a host-to-device copy
  // (set_images). An event recorded right after it is queued on that copy
  // engine -- in order behind the uploads and downloads of every other
  // composer of the process -- and the first kernel waits for it: measured
  // 2-5 ms of dead time per panorama with three panoramas in flight. One
  // empty kernel moves the stream to the compute engine first.
  void begin_on_compute() {
    noop_kernel<<<1, 32, 0, stream>>>();
    XP_LAUNCHED();
  }

  void run(xp_compose_info* info) {
    run_begin();
    run_end(info);
  }

  // run() in two halves for callers that keep several panoramas in flight:
  // run_begin returns when the last kernel of the panorama is queued (after
  // the blend's last host round trip, before its pyramid tail has run);
  // run_end waits for it.
  bool run_pending = false;
  double host_ms[3] = {0, 0, 0};
  void run_begin() {
    XP_REQUIRE(!src.empty(), "run: set_images first");
    XP_REQUIRE(tile_count == 1,
               "run: this composer holds a canvas tile, use run_tile_begin / "
               "run_tile_end");
    XP_REQUIRE(!run_pending, "run_begin: the previous run was not ended");
    auto now = [] {
      return std::chrono::duration<double, std::milli>(
                 std::chrono::steady_clock::now().time_since_epoch())
          .count();
    };
    const double h0 = now();
    ensure_device();
    if (!plan.valid) make_plan();
    const double h1 = now();
    begin_on_compute();
    XP_CUDA(cudaEventRecord(ev[0], stream));
    warp_all();
    XP_CUDA(cudaEventRecord(ev[1], stream));
    const double h2 = now();
    blender->blend_launch();
    XP_CUDA(cudaEventRecord(ev[2], stream));
    host_ms[0] = h1 - h0, host_ms[1] = h2 - h1, host_ms[2] = now() - h2;
    run_pending = true;
  }

  void run_end(xp_compose_info* info) {
    XP_REQUIRE(run_pending && blender, "run_end before run_begin");
    xp_timing tm;
    memset(&tm, 0, sizeof(tm));
    blender->blend_finish(&tm);
    run_pending = false;
    if (getenv("XP_TRACE"))
      fprintf(stderr,
              "trace host: roi %.3f ms, new blender + warp launches %.3f ms, "
              "blend launches and round trips %.3f ms\n",
              host_ms[0], host_ms[1], host_ms[2]);
    float a = 0, b = 0;
    cudaEventElapsedTime(&a, ev[0], ev[1]);
    cudaEventElapsedTime(&b, ev[1], ev[2]);
    fill_info(info, tm, a, b);
  }

  // Canvas-tiled run: everything up to the collapsed strip;
  // `partial` = this rank's share of the sums the offset correction needs.
  void run_tile_begin(double partial[8]) {
    XP_REQUIRE(!src.empty(), "run_tile_begin: set_images first");
    XP_REQUIRE(tile_count > 1, "run_tile_begin: set_tile first");
    ensure_device();
    if (!plan.valid) make_plan();
    begin_on_compute();
    XP_CUDA(cudaEventRecord(ev[0], stream));
    warp_all();
    XP_CUDA(cudaEventRecord(ev[1], stream));
    blender->blend_tile_begin(partial);
    tile_pending = true;
  }

  void run_tile_end(const double reduced[8], xp_compose_info* info) {
    XP_REQUIRE(tile_pending && blender, "run_tile_end before run_tile_begin");
    xp_timing tm;
    memset(&tm, 0, sizeof(tm));
    blender->blend_tile_end(reduced, &tm);
    tile_pending = false;
    float a = 0;
    cudaEventElapsedTime(&a, ev[0], ev[1]);
    fill_info(info, tm, a, (float)tm.total);
  }
};

}  // namespace xp

struct xp_composer : xp::Composer {};

extern "C" {

int xp_compose_capped_scale(double warped_image_scale,
                            double compose_work_aspect, int roi_w, int roi_h,
                            double max_pano_mpx, float* warp_scale) {
  // utils::opencv::MPx (utils/opencv.cc:35-38): float product over 1e6f;
  // max_pano_mpx_ is a float (stitcher.h:141-143, 282); std::sqrt of a float
  // is the float overload; warped_image_scale_ (double) *= ratio, and only
  // then is the product with the aspect rounded to float (stitcher.cc:318-329)
  const float pano_mpx = (float)roi_w * (float)roi_h / 1e6f;
  const float max_mpx = (float)max_pano_mpx;
  int capped = 0;
  if (max_pano_mpx > 0 && pano_mpx > max_mpx) {
    const float downscale_ratio = sqrtf(max_mpx / pano_mpx);
    warped_image_scale *= downscale_ratio;
    capped = 1;
  }
  if (warp_scale) *warp_scale = (float)(warped_image_scale * compose_work_aspect);
  return capped;
}

void xp_compose_opts_default(xp_compose_opts* o) {
  if (!o) return;
  memset(o, 0, sizeof(*o));
  o->projection = XP_PROJ_SPHERICAL;  // options.h:83
  o->a = 2.0f;
  o->b = 1.0f;
  o->warp_scale = 1.0f;
  o->warped_image_scale = 0.0;  // 0: use warp_scale as given
  o->compose_work_aspect = 1.0;
  o->max_pano_mpx = 100.0;  // constants.h:106
  o->depth = 8;
  xp_blend_opts_default(&o->blend);
}

xp_status xp_composer_create(const xp_compose_opts* opts, xp_composer** out) {
  return xp::guard([&] {
    XP_REQUIRE(opts && out, "xp_composer_create: null argument");
    XP_REQUIRE(opts->projection >= 0 && opts->projection <= 8,
               "unknown projection %d", opts->projection);
    auto* c = new xp_composer();
    c->opts = *opts;
    *out = c;
  });
}

void xp_composer_destroy(xp_composer* c) { delete c; }

xp_status xp_composer_set_stream(xp_composer* c, void* cuda_stream) {
  return xp::guard([&] {
    XP_REQUIRE(c, "null composer");
    c->stream = (cudaStream_t)cuda_stream;
    c->external_stream = true;
  });
}

xp_status xp_composer_set_images(xp_composer* c, int n,
                                 const xp_compose_image* images, int mem) {
  return xp::guard([&] {
    XP_REQUIRE(c, "null composer");
    c->set_images(n, images, mem);
  });
}

xp_status xp_composer_run(xp_composer* c, xp_compose_info* info) {
  return xp::guard([&] {
    XP_REQUIRE(c, "null composer");
    c->run(info);
  });
}

xp_status xp_composer_run_begin(xp_composer* c) {
  return xp::guard([&] {
    XP_REQUIRE(c, "null composer");
    c->run_begin();
  });
}

xp_status xp_composer_run_end(xp_composer* c, xp_compose_info* info) {
  return xp::guard([&] {
    XP_REQUIRE(c, "null composer");
    c->run_end(info);
  });
}

xp_status xp_composer_set_cancel_flag(xp_composer* c,
                                      const volatile int* flag) {
  return xp::guard([&] {
    XP_REQUIRE(c, "null composer");
    c->cancel_flag = flag;
  });
}

xp_status xp_composer_set_tile(xp_composer* c, int index, int count) {
  return xp::guard([&] {
    XP_REQUIRE(c, "null composer");
    XP_REQUIRE(count >= 1 && index >= 0 && index < count,
               "set_tile: index %d of %d", index, count);
    XP_REQUIRE(!c->tile_pending, "set_tile between run_tile_begin and _end");
    c->tile_index = index;
    c->tile_count = count;
    c->plan.valid = false;
  });
}

xp_status xp_composer_tile_needs(xp_composer* c, int n, int* needed) {
  return xp::guard([&] {
    XP_REQUIRE(c && needed, "tile_needs: null argument");
    XP_REQUIRE(c->plan.valid && (int)c->plan.local.size() == n,
               "tile_needs: set_images first");
    for (int i = 0; i < n; ++i) needed[i] = c->plan.local[i];
  });
}

xp_status xp_composer_run_tile_begin(xp_composer* c, double partial[8]) {
  return xp::guard([&] {
    XP_REQUIRE(c && partial, "run_tile_begin: null argument");
    c->run_tile_begin(partial);
  });
}

xp_status xp_composer_run_tile_end(xp_composer* c, const double reduced[8],
                                   xp_compose_info* info) {
  return xp::guard([&] {
    XP_REQUIRE(c && reduced, "run_tile_end: null argument");
    c->run_tile_end(reduced, info);
  });
}

int xp_tile_halo(int levels) {
  return levels >= 0 && levels <= xp::kMaxLevels ? xp::tile_halo(levels) : -1;
}

xp_status xp_composer_fetch(xp_composer* c, void* dst_bgr, size_t dst_stride,
                            uint8_t* dst_mask, size_t mask_stride, int mem) {
  return xp::guard([&] {
    XP_REQUIRE(c && c->blender, "fetch before run");
    if (c->run_pending) c->run_end(nullptr);  // fetch completes the run
    c->blender->fetch(dst_bgr, dst_stride, dst_mask, mask_stride, mem);
  });
}

xp_blender* xp_composer_blender(xp_composer* c) {
  return c ? reinterpret_cast<xp_blender*>(c->blender.get()) : nullptr;
}

}  // extern "C"
