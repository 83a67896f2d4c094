// Blender: device-resident restatement of multiblend::Multiblend() as xpano
// calls it (in-memory images, no gamma, no wrap), behind the
// cv::detail::Blender-shaped prepare / feed / blend surface of
// xpano/algorithm/blenders.cc:116-158.
//
// Stage order follows external/multiblend/src/multiblend.cpp:66-1307:
// read/trim/inpaint (image.cpp:336-919) -> tighten (:139-154) -> level count
// (:161-206) -> seams (:254-823) -> mask pyramids (:919-930) -> per image
// copy/shrink/laplace/composite (:1057-1126) -> collapse (:1129) -> offset
// correction (:1232-1267) -> out (:1274-1286) -> blenders.cc:151-154.
#include <algorithm>
#include <climits>
#include <cmath>
#include <mutex>

#include "blender.h"
#include "common.h"

namespace xp {

// ---------------------------------------------------------------------------
// level count: multiblend.cpp:161-206
// ---------------------------------------------------------------------------
int blend_levels(int n, const int* widths, const int* heights,
                 int fixed_levels, int add_levels) {
  int levels;
  if (fixed_levels == 0) {
    std::vector<int> ws(widths, widths + n), hs(heights, heights + n);
    std::sort(ws.begin(), ws.end());
    std::sort(hs.begin(), hs.end());
    const int half = (n - 1) >> 1;
    const int mw = (n & 1) ? ws[half] : (ws[half] + ws[half + 1] + 1) >> 1;
    const int mh = (n & 1) ? hs[half] : (hs[half] + hs[half + 1] + 1) >> 1;
    const int blend_wh = std::max(mw, mh);
    // std::log2(blend_wh + 4.0f) is evaluated in float
    levels = (int)floorf(log2f((float)blend_wh + 4.0f) - 1);
  } else {
    levels = fixed_levels;
  }
  levels += add_levels;
  if (n == 1) levels = 0;
  return levels;
}

// ---------------------------------------------------------------------------
// stage timing with CUDA events on the blender's stream
// ---------------------------------------------------------------------------
// timing events are kept across blenders (a blender is single-use and
// records ~40 of them; cudaEventCreate is host time on every panorama)
namespace {
struct EventCache {
  std::mutex mu;
  std::vector<cudaEvent_t> free_events[16];
};
EventCache& event_cache() {
  static EventCache* c = new EventCache();
  return *c;
}
}  // namespace

StageTimer::~StageTimer() {
  if (pool.empty()) return;
  EventCache& c = event_cache();
  std::lock_guard<std::mutex> lk(c.mu);
  auto& f = c.free_events[device & 15];
  f.insert(f.end(), pool.begin(), pool.end());
}
cudaEvent_t StageTimer::get() {
  if (used == pool.size()) {
    if (device < 0) XP_CUDA(cudaGetDevice(&device));
    cudaEvent_t e = nullptr;
    {
      EventCache& c = event_cache();
      std::lock_guard<std::mutex> lk(c.mu);
      auto& f = c.free_events[device & 15];
      if (!f.empty()) {
        e = f.back();
        f.pop_back();
      }
    }
    if (!e) XP_CUDA(cudaEventCreate(&e));
    pool.push_back(e);
  }
  return pool[used++];
}
void StageTimer::reset(cudaStream_t stream) {
  spans.clear();
  used = 0;
  s = stream;
}
int StageTimer::begin(int stage, cudaStream_t on) {
  Span sp{stage, get(), get(), on ? on : s};
  XP_CUDA(cudaEventRecord(sp.a, sp.s));
  spans.push_back(sp);
  return (int)spans.size() - 1;
}
void StageTimer::end(int idx) {
  XP_CUDA(cudaEventRecord(spans[idx].b, spans[idx].s));
}
void StageTimer::collect(double out[16]) {
  for (int i = 0; i < 16; ++i) out[i] = 0;
  for (const Span& sp : spans) {
    float ms = 0;
    if (cudaEventElapsedTime(&ms, sp.a, sp.b) == cudaSuccess)
      out[sp.stage] += ms;
  }
}

enum {
  ST_IMAGES = 0, ST_SEAM, ST_SHRINK_MASK, ST_COPY, ST_SHRINK, ST_LAPLACE,
  ST_BLEND, ST_COLLAPSE, ST_OUT, ST_TOTAL, ST_TAIL
};

struct Scope {
  StageTimer& t;
  int idx;
  Scope(StageTimer& tt, int stage, cudaStream_t on = nullptr)
      : t(tt), idx(tt.begin(stage, on)) {}
  ~Scope() {
    if (cudaEventRecord(t.spans[idx].b, t.spans[idx].s) != cudaSuccess)
      cudaGetLastError();
  }
};

// ---------------------------------------------------------------------------
// Blender basics
// ---------------------------------------------------------------------------
Blender::~Blender() {
  // buffers go back to the caching pool: nothing may still be using them.
  // The main stream may be the caller's legacy default stream (a null
  // handle), which is synchronised like any other -- on the blender's device.
  if (device_ready) {
    int cur = 0;
    cudaGetDevice(&cur);
    if (cur != opts.device) cudaSetDevice(opts.device);
    if (side_stream) cudaStreamSynchronize(side_stream);
    if (aux_stream) cudaStreamSynchronize(aux_stream);
    if (seam_stream) cudaStreamSynchronize(seam_stream);
    if (part_side) cudaStreamSynchronize(part_side);
    if (part_aux) cudaStreamSynchronize(part_aux);
    cudaStreamSynchronize(stream);
    if (cur != opts.device) cudaSetDevice(cur);
  }
  // the side streams and events are idle now: back to the per-device cache
  if (side_stream) {
    StreamSet set;
    set.device = opts.device;
    set.side = side_stream;
    set.aux = aux_stream;
    set.ev[0] = ev_fork, set.ev[1] = ev_join, set.ev[2] = ev_cover;
    set.ev[3] = ev_aux, set.ev[4] = ev_cls, set.ev[5] = ev_seam;
    set.seam = seam_stream, set.part_side = part_side, set.part_aux = part_aux;
    set.seam_sms = seam_sms, set.side_sms = side_sms;
    release_stream_set(set);
  }
  if (own_stream) cudaStreamDestroy(own_stream);
}

void Blender::ensure_device() {
  select_device(opts.device);
  device_ready = true;
  if (!stream && !external_stream) {
    own_stream = create_stream(true);
    stream = own_stream;
  }
  if (!side_stream) {
    const StreamSet set = acquire_stream_set(opts.device);
    side_stream = set.side;
    aux_stream = set.aux;
    ev_fork = set.ev[0], ev_join = set.ev[1], ev_cover = set.ev[2];
    ev_aux = set.ev[3], ev_cls = set.ev[4], ev_seam = set.ev[5];
    seam_stream = set.seam, part_side = set.part_side, part_aux = set.part_aux;
    seam_sms = set.seam_sms, side_sms = set.side_sms;
  }
}

Fed* Blender::new_fed(int w, int h, int depth, int tl_x, int tl_y) {
  XP_REQUIRE(w > 0 && h > 0, "feed: empty image");
  XP_REQUIRE(depth == 8 || depth == 16, "feed: depth must be 8 or 16");
  XP_REQUIRE(((size_t)w + 3) * h < (1ull << 32), "feed: image too large");
  XP_REQUIRE(!blended, "blender instances are single-use (blenders.cc)");
  if (!fed.empty())
    XP_REQUIRE(fed[0]->depth == depth,
               "mixture of 8bpp and 16bpp images (multiblend.cpp:100-117)");
  XP_REQUIRE((int)fed.size() < kMaxImages, "more than %d images", kMaxImages);
  ensure_device();
  std::unique_ptr<Fed> f(new Fed());
  f->w = w;
  f->pitch = (w + 3) & ~3;
  f->h = h;
  f->depth = depth;
  f->xpos = tl_x;
  f->ypos = tl_y;
  f->data.reserve((size_t)f->pitch * h * 4 * (depth / 8));
  fed.push_back(std::move(f));
  return fed.back().get();
}

void Blender::feed(const void* data, size_t stride, const uint8_t* mask,
                   size_t mask_stride, int w, int h, int depth, int spp,
                   int tl_x, int tl_y, int mem) {
  XP_REQUIRE(data, "feed: null image");
  XP_REQUIRE(spp == 3 || spp == 4, "feed: spp must be 3 or 4");
  const size_t bs = depth / 8;
  const size_t row = (size_t)w * spp * bs;
  if (stride == 0) stride = row;
  if (mask_stride == 0) mask_stride = w;
  Fed* f = new_fed(w, h, depth, tl_x, tl_y);
  const void* dsrc = data;
  size_t dpitch = stride;
  const uint8_t* dmask = mask;
  size_t mpitch = mask_stride;
  if (mem == XP_MEM_HOST) {
    staging.reserve(row * (size_t)h);
    copy_in_2d(staging.p, row, data, stride, row, h, mem, stream);
    dsrc = staging.p;
    dpitch = row;
    if (mask) {
      staging_mask.reserve((size_t)w * h);
      copy_in_2d(staging_mask.p, w, mask, mask_stride, w, h, mem, stream);
      dmask = staging_mask.as<uint8_t>();
      mpitch = w;
    }
  }
  launch_pack_bgra(dsrc, dpitch, dmask, mpitch, w, h, depth, spp, f->data.p,
                   f->pitch, stream);
  XP_CUDA(cudaStreamSynchronize(stream));
}

// mask pyramid rectangle of the next level (functions.cpp:188-191 geometry:
// output j reads inputs 2j-4 .. 2j, clamped into the level)
static void next_mask_rect(const MaskLevel& in, MaskLevel& out) {
  out.cw = (in.cw + 6) >> 1;
  out.ch = (in.ch + 6) >> 1;
  if (in.w <= 0 || in.h <= 0) {
    out.x0 = out.y0 = out.w = out.h = 0;
    return;
  }
  auto range = [](int s0, int s1, int in_size, int out_size, int& j0, int& j1) {
    j0 = s0 == 0 ? 0 : (s0 + 1) >> 1;
    j1 = s1 == in_size - 1 ? out_size - 1 : (s1 + 4) >> 1;
    if (j0 < 0) j0 = 0;
    if (j1 > out_size - 1) j1 = out_size - 1;
  };
  int j0, j1, k0, k1;
  range(in.x0, in.x0 + in.w - 1, in.cw, out.cw, j0, j1);
  range(in.y0, in.y0 + in.h - 1, in.ch, out.ch, k0, k1);
  out.x0 = j0;
  out.y0 = k0;
  out.w = j1 - j0 + 1;
  out.h = k1 - k0 + 1;
}

// SM partition policy (blend_run.inc): XP_PARTITION_HOLES overrides the ratio
static double partition_max_hole_ratio() {
  static const double v = [] {
    const char* e = getenv("XP_PARTITION_HOLES");
    return e ? atof(e) : 0.5;
  }();
  return v;
}

}  // namespace xp

#include "blend_run.inc"
#include "blend_tile.inc"
#include "blend_api.inc"
