#!/usr/bin/env python
"""Benchmark of the full-resolution export hot path (warp + seams/masks +
multiband blend), BASELINE.json metric "output megapixels/sec".

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A step is one pass of the hot path over one synthetic image set: N=1 runs
BASELINE.json configs[1] (8 x 12-MP inputs -> ~60-MP spherical panorama,
5 pyramid levels). With N > 1 (torchrun, one rank per GPU) every rank
composes its own panorama group -- the path shards by group with no data-path
collective -- and `value` is the total output MP of all
ranks over the slowest rank's time ("weak" scaling).

`value`   : sources resident in HBM at t0, panorama resident at t1.
`e2e`     : the same through the public call with HOST buffers: upload of the
            sources from pinned memory and download of panorama + mask inside
            the timed region.
`roofline`: dominant HBM-bound stage, algorithmic bytes / CUDA-event time
            against MEASURED_PEAKS.json.
`--impl reference` times the reference's own CPU implementation (cv2 warp
exactly as stitcher.cc:377-385 calls it + the unmodified multiblend library,
oracle/_ref) on a bounded sample of the same workload.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "output megapixels/sec (warp+multiband blend)"
UNIT = "MP/s"

CONFIGS = {
    # BASELINE.json configs[1]
    "cfg2": dict(n_cols=8, n_rows=1, w=4000, h=3000, focal=3600.0,
                 step=0.66, pitches=[0.0], projection=2, fixed_levels=5,
                 bpp=8, max_pano_mpx=100.0,
                 name="8x12MP u8 -> ~60MP spherical, 5 levels"),
    # BASELINE.json configs[2] (one GPU holds it; canvas tiling is future work)
    "cfg3": dict(n_cols=8, n_rows=3, w=6000, h=4000, focal=4500.0,
                 step=0.55, pitches=[-math.radians(25), 0.0,
                                     math.radians(25)],
                 projection=2, fixed_levels=7, bpp=8, max_pano_mpx=100.0,
                 name="24x24MP u8 -> 100MP spherical (capped), 7 levels"),
    # BASELINE.json configs[3]: 16-bit, cylindrical / Mercator, with the
    # pitch/yaw/roll adjustment R <- M R (algorithm.cc:362-369), auto levels
    "cfg4c": dict(n_cols=8, n_rows=2, w=5472, h=3648, focal=4000.0, step=0.6,
                  pitches=[-math.radians(12), math.radians(12)], projection=1,
                  fixed_levels=0, bpp=16, max_pano_mpx=1000.0,
                  extra_rot=(10.0, 5.0, 3.0),
                  name="16x20MP u16 cylindrical + rotation, auto levels"),
    "cfg4m": dict(n_cols=8, n_rows=2, w=5472, h=3648, focal=4000.0, step=0.6,
                  pitches=[-math.radians(12), math.radians(12)], projection=7,
                  fixed_levels=0, bpp=16, max_pano_mpx=1000.0,
                  extra_rot=(10.0, 5.0, 3.0),
                  name="16x20MP u16 Mercator + rotation, auto levels"),
    # BASELINE.json configs[4]: one group of the 64-group batch export
    "cfg5": dict(n_cols=6, n_rows=1, w=4000, h=3000, focal=3600.0, step=0.66,
                 pitches=[0.0], projection=2, fixed_levels=0, bpp=8,
                 max_pano_mpx=100.0,
                 name="6x12MP u8 spherical group (batch export), auto levels"),
    # reduced shapes for smoke runs
    "small": dict(n_cols=4, n_rows=1, w=800, h=600, focal=700.0, step=0.62,
                  pitches=[0.0], projection=2, fixed_levels=5, bpp=8,
                  max_pano_mpx=100.0, name="4x0.5MP u8 smoke shape"),
}


def make_image(w, h, seed, bpp=8):
    """Band-limited synthetic texture of SURVEY.md section 8d (sinusoids with
    periods 97/131/211 px, +-4 hash noise, per-image gain), vectorised."""
    rng = np.random.default_rng(seed)
    x = np.arange(w, dtype=np.float32)
    y = np.arange(h, dtype=np.float32)
    cy = np.cos(2 * np.pi * y / 131.0)[:, None]
    diag = 40 * np.sin(2 * np.pi * np.arange(w + h, dtype=np.float32) / 211.0)
    dsum = diag[np.add.outer(np.arange(h), np.arange(w))]
    gain = rng.uniform(0.9, 1.1)
    out = np.empty((h, w, 3), np.uint8 if bpp == 8 else np.uint16)
    for c in range(3):
        sx = 60 * np.sin(2 * np.pi * (x / 97.0 + rng.uniform(0, 1)))[None, :]
        img = (128 + sx * cy + dsum) * gain
        img += rng.integers(-4, 5, (h, w)).astype(np.float32)
        np.clip(img, 0, 255, out=img)
        out[..., c] = img if bpp == 8 else img * 257
    return out


PROJ_NAMES = {2: "spherical", 1: "cylindrical", 7: "mercator", 0: "plane"}


def estimate_seams(ims, Ks, Rs, cfg, masks):
    """The path's two small inputs, produced as the reference produces them:
    Stitcher::EstimateSeams (stitcher.cc:241-296) run once on the CPU with
    cv2 -- images resized to the seam scale (0.1 MPx, stitcher.cc:182),
    warped at warped_image_scale * seam_work_aspect, BlocksGainCompensator
    fed and applied, GraphCutSeamFinder(COST_COLOR) (stitcher.cc:184-185).
    This step is outside the path and outside every timed
    region; both arms get the same arrays.

    masks = "graphcut": the seam finder's masks (the reference's default);
            "full": the warped all-255 masks untouched by a seam finder
            (isolates multiblend's own seaming).
    -> (gains | None, seam_masks): per image, low resolution."""
    import cv2
    w, h = cfg["w"], cfg["h"]
    seam_scale = min(1.0, math.sqrt(0.1 * 1e6 / (w * h)))
    warper = cv2.PyRotationWarper(PROJ_NAMES[cfg["projection"]],
                                  float(np.float32(cfg["focal"] * seam_scale)))
    corners, warped, wmasks = [], [], []
    for im, K, R in zip(ims, Ks, Rs):
        small = cv2.resize(im, None, fx=seam_scale, fy=seam_scale,
                           interpolation=cv2.INTER_LINEAR_EXACT)
        if small.dtype != np.uint8:  # the compensator and finder are 8-bit
            small = (small >> 8).astype(np.uint8)
        Ks_ = np.array(K, np.float32)
        Ks_[0, 0] *= seam_scale
        Ks_[1, 1] *= seam_scale
        Ks_[0, 2] *= seam_scale
        Ks_[1, 2] *= seam_scale
        c, iw = warper.warp(small, Ks_, R, cv2.INTER_LINEAR,
                            cv2.BORDER_REFLECT)
        _, mw = warper.warp(np.full(small.shape[:2], 255, np.uint8), Ks_, R,
                            cv2.INTER_NEAREST, cv2.BORDER_CONSTANT)
        corners.append(c)
        warped.append(iw)
        wmasks.append(mw)
    gains = None
    if cfg["bpp"] == 8:  # BlocksCompensator::apply asserts CV_8UC3
        comp = cv2.detail_BlocksGainCompensator()
        comp.feed(corners, warped, wmasks)
        warped = [comp.apply(i, corners[i], warped[i], wmasks[i])
                  for i in range(len(ims))]
        gains = [np.ascontiguousarray(g, np.float32)
                 for g in comp.getMatGains()]
    if masks == "graphcut":
        finder = cv2.detail_GraphCutSeamFinder("COST_COLOR")
        out = finder.find([iw.astype(np.float32) for iw in warped], corners,
                          wmasks)
        wmasks = [m.get() if hasattr(m, "get") else np.asarray(m)
                  for m in out]
    return gains, [np.ascontiguousarray(m, np.uint8) for m in wmasks]


def make_workload(cfg, group, masks="graphcut"):
    """Synthetic registered image set + the EstimateSeams
    inputs. With XP_BENCH_CACHE=<dir> the arrays are kept on disk, so that
    successive bench runs on one box (N = 1, 2, 4, 8) do not regenerate them;
    nothing computed by the path itself is cached."""
    cache = os.environ.get("XP_BENCH_CACHE")
    if cache:
        import pickle
        os.makedirs(cache, exist_ok=True)
        key = "".join(c if c.isalnum() else "_" for c in cfg["name"])
        path = os.path.join(cache, f"{key}_g{group}_{masks}.pkl")
        if os.path.exists(path):
            with open(path, "rb") as f:
                return pickle.load(f)
        wl = _make_workload(cfg, group, masks)
        tmp = f"{path}.{os.getpid()}.tmp"
        with open(tmp, "wb") as f:
            pickle.dump(wl, f, protocol=4)
        os.replace(tmp, path)
        return wl
    return _make_workload(cfg, group, masks)


def _make_workload(cfg, group, masks="graphcut"):
    from xpano_b200 import synthetic as co
    hfov = 2 * math.atan(cfg["w"] / 2 / cfg["focal"])
    extra = None
    if cfg.get("extra_rot"):
        yaw, pitch, roll = (math.radians(v) for v in cfg["extra_rot"])
        extra = co.rot(yaw, pitch, roll)
    Ks, Rs = co.make_cameras(cfg["n_cols"], cfg["n_rows"], cfg["w"], cfg["h"],
                             cfg["focal"], cfg["step"] * hfov, cfg["pitches"],
                             2000 + 10 * group, extra=extra)
    ims = [make_image(cfg["w"], cfg["h"], 2000 + 10 * group + i, cfg["bpp"])
           for i in range(len(Ks))]
    gains = seams = None
    if masks != "none":
        gains, seams = estimate_seams(ims, Ks, Rs, cfg, masks)
    return {"ims": ims, "Ks": Ks, "Rs": Rs, "gains": gains, "seams": seams,
            "masks": masks}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,"
         "clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index, interval_ms=20):
        # every query takes driver locks that kernel launches also need: with
        # several ranks on one box only a slow poll is harmless
        self.p = None
        if index is None:
            return
        try:
            self.p = subprocess.Popen(
                ["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", str(interval_ms)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            pass

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        self.p.terminate()
        try:
            out, _ = self.p.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
            out, _ = self.p.communicate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                 "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [v.strip() for v in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------
# reference arm: the reference's own CPU implementation of the path
# ---------------------------------------------------------------------------
def host_cores():
    return len(os.sched_getaffinity(0))


def reference_step(wl, cfg, cores, keep=False):
    """One pass of the CPU path over the workload. Returns (canvas MP,
    seconds, kind, result | None). kind "reference": the loop of
    stitcher.cc:358-419 with cv2 -- RotationWarper.warp twice per image
    (:377-385), ExposureCompensator::apply (:389-390), dilate + resize
    INTER_LINEAR_EXACT + bitwise_and (:397-401), Merge (blenders.cc:99-112)
    -- then the unmodified multiblend (oracle/_ref) with max(2, cores-1)
    threads (stitcher_pipeline.h:148-149); kind "port": the oracle's C
    restatement (single thread) when either is unavailable."""
    from oracle import mb_ref
    try:
        import cv2
    except Exception:  # pragma: no cover
        cv2 = None
    ims, Ks, Rs = wl["ims"], wl["Ks"], wl["Rs"]
    gains, seams = wl["gains"], wl["seams"]
    use_ref = cv2 is not None and mb_ref.available() \
        and cfg["projection"] in PROJ_NAMES
    t0 = time.perf_counter()
    result = None
    if use_ref:
        cv2.setNumThreads(cores)
        scale, capped = reference_warp_scale(wl, cfg, cv2)
        warper = cv2.PyRotationWarper(PROJ_NAMES[cfg["projection"]], scale)
        comp = None
        if gains is not None:
            comp = cv2.detail_BlocksGainCompensator()
            comp.setMatGains(gains)
        fed, pos = [], []
        for i, (im, K, R) in enumerate(zip(ims, Ks, Rs)):
            if seams is not None and cv2.countNonZero(seams[i]) == 0:
                continue
            corner, warped = warper.warp(im, K, R, cv2.INTER_LINEAR,
                                         cv2.BORDER_REFLECT)
            mask = np.full(im.shape[:2], 255, np.uint8)
            _, mw = warper.warp(mask, K, R, cv2.INTER_NEAREST,
                                cv2.BORDER_CONSTANT)
            if comp is not None:
                warped = comp.apply(i, corner, warped, mw)
            if seams is not None:
                dil = cv2.dilate(seams[i], None)
                sm = cv2.resize(dil, (mw.shape[1], mw.shape[0]),
                                interpolation=cv2.INTER_LINEAR_EXACT)
                mw = cv2.bitwise_and(sm, mw)
            alpha = cv2.compare(mw, 0, cv2.CMP_NE)          # blenders.cc:106
            if warped.dtype == np.uint16:
                alpha = alpha.astype(np.uint16) * 257
            fed.append(cv2.merge(list(cv2.split(warped)) + [alpha]))
            pos.append(corner)
        res = mb_ref.blend(fed, pos, fixed_levels=cfg["fixed_levels"],
                           n_threads=max(2, cores - 1))
        mp = res.width * res.height / 1e6
        kind = "reference"
        if keep:
            result = {"pano": res.interleaved(True), "mask": res.mask,
                      "min": (res.min_x, res.min_y), "capped": capped,
                      "warp_scale": scale}
    else:
        import compose_oracle as co
        res = co.compose(ims, Ks, Rs, projection=cfg["projection"],
                         warp_scale=cfg["focal"],
                         max_pano_mpx=cfg["max_pano_mpx"], gains=gains,
                         seam_masks=seams, fixed_levels=cfg["fixed_levels"])
        mp = res["pano"].shape[0] * res["pano"].shape[1] / 1e6
        kind = "port"
        if keep:
            result = {"pano": res["pano"], "mask": res["mask"],
                      "min": (res["min_x"], res["min_y"]),
                      "capped": res["capped"],
                      "warp_scale": res["warp_scale"]}
    return mp, time.perf_counter() - t0, kind, result


def reference_warp_scale(wl, cfg, cv2):
    """ComputeRoi + the size cap (stitcher.cc:144-169, 312-335) with cv2's
    warpRoi; float / double mix as in the reference."""
    f32 = np.float32
    wis = float(f32(cfg["focal"]))
    scale = float(f32(wis))
    cap = cfg["max_pano_mpx"]
    warper = cv2.PyRotationWarper(PROJ_NAMES[cfg["projection"]], scale)
    rois = [warper.warpRoi((im.shape[1], im.shape[0]), K, R)
            for im, K, R in zip(wl["ims"], wl["Ks"], wl["Rs"])]
    x0 = min(r[0] for r in rois)
    y0 = min(r[1] for r in rois)
    x1 = max(r[0] + r[2] for r in rois)
    y1 = max(r[1] + r[3] for r in rois)
    pano_mpx = f32(f32(f32(x1 - x0) * f32(y1 - y0)) / f32(1e6))
    if cap > 0 and pano_mpx > f32(cap):
        ratio = f32(np.sqrt(f32(f32(cap) / pano_mpx)))
        return float(f32(wis * float(ratio))), True
    return scale, False


def parity_check(pano, mask, min_xy, ref, bpp):
    """North-star bars at full size: masks bit-exact, panorama within 1 LSB
    (8-bit) / 1e-3 relative (16-bit) at >= 99.9 % of pixels, PSNR >= 50 dB."""
    out = {"against": "reference CPU arm, all images, same inputs"}
    if pano.shape != ref["pano"].shape or tuple(min_xy) != tuple(ref["min"]):
        out.update(ok=False, size=[list(pano.shape), list(ref["pano"].shape)],
                   min_xy=[list(min_xy), list(ref["min"])])
        return out
    nm = int(np.count_nonzero(mask != ref["mask"]))
    both = (mask > 0) & (ref["mask"] > 0)
    a = pano[both].astype(np.int32)
    b = ref["pano"][both].astype(np.int32)
    d = np.abs(a - b)
    tol = 1 if bpp == 8 else np.maximum(1, (1e-3 * b).astype(np.int32))
    within = float((d <= tol).mean())
    mse = float(np.mean(d.astype(np.float64) ** 2))
    maxv = 255.0 if bpp == 8 else 65535.0
    psnr = 200.0 if mse == 0 else 10 * math.log10(maxv * maxv / mse)
    out.update(mask_equal=nm == 0, mask_diff_pixels=nm,
               within_tol=within, identical=float((d == 0).mean()),
               max_diff=int(d.max()), psnr=round(psnr, 2),
               pixels=int(mask.size),
               ok=bool(nm <= 1e-6 * mask.size and within >= 0.999
                       and psnr >= 50.0))
    return out


def sample_of(cfg, wl, n_sample):
    """The bounded CPU sample: the first n images of the workload (0 = all)."""
    n_all = len(wl["ims"])
    n = n_all if n_sample <= 0 else min(n_sample, n_all)
    if n == n_all:
        return wl, (f"the whole workload ({n} x {cfg['w']}x{cfg['h']} images,"
                    f" ratio 1.0)")
    sub = dict(wl)
    for k in ("ims", "Ks", "Rs", "gains", "seams"):
        sub[k] = wl[k][:n] if wl[k] is not None else None
    return sub, (f"first {n} of {n_all} images of the workload "
                 f"({n} x {cfg['w']}x{cfg['h']}, ratio {n / n_all:.3f})")


def config_of(cfg, wl):
    """`config` of the JSON line: identical on both arms."""
    return {"workload": cfg["name"], "images": len(wl["ims"]),
            "masks": {"graphcut": "cv2 GraphCutSeamFinder(COST_COLOR) at the "
                      "seam scale, computed once on the CPU",
                      "full": "warped all-255 masks (no seam finder)",
                      "none": "no seam masks, no gains"}[wl["masks"]],
            "gains": "BlocksGainCompensator maps from the seam-scale images"
            if wl["gains"] is not None else "none",
            "l2": "inputs and intermediates (>2 GB per step) exceed the "
            "126 MB L2"}


def run_reference(args, cfg):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = host_cores()
    full = make_workload(cfg, 0, args.masks)
    wl, sample = sample_of(cfg, full, args.ref_images)
    kind = "port"
    for _ in range(args.warmup):
        reference_step(wl, cfg, cores)
    total_mp, total_s = 0.0, 0.0
    for _ in range(args.steps):
        mp, s, kind, _ = reference_step(wl, cfg, cores)
        total_mp += mp
        total_s += s
    value = total_mp / total_s
    used = cores if kind == "reference" else 1
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * total_s / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None,
        "dtype": "u8" if cfg["bpp"] == 8 else "u16", "data": "synthetic",
        "config": config_of(cfg, full),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": used,
                         "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


# ---------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------
HBM_STAGES = ("warp", "shrink_mask", "copy", "shrink", "laplace", "blend",
              "collapse", "out")


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def ncu_traffic(workload):
    """dram bytes per step of the pyramid/blend group from the committed ncu
    capture (profiles/traffic.json, written by tools/ncu_traffic.py), or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
        return t.get(workload, {}).get("blend_group_dram_bytes_per_step")
    except Exception:
        return None


def operator_surface_leg(xb, wl, cfg, steps):
    """The path as xpano would drive it through the shim
    (include/xpano_b200_shim.hpp), one image at a time with HOST buffers:
    ComputeRoi, then per image warp(image) + warp(mask) on the GPU,
    ExposureCompensator::apply and dilate / resize / bitwise_and by OpenCV on
    the CPU exactly where stitcher.cc:389-401 has them, feed(), and finally
    blend() + download. Returns wall ms per panorama and the split between
    library calls and the CPU steps in between."""
    try:
        import cv2
    except Exception:
        return None
    ims, Ks, Rs = wl["ims"], wl["Ks"], wl["Rs"]
    gains, seams = wl["gains"], wl["seams"]
    if cfg["max_pano_mpx"] > 0:
        pass  # the cap changes the warp scale below, as in the composer
    creator = xb.PickWarper(cfg["projection"])
    comp = None
    if gains is not None:
        comp = cv2.detail_BlocksGainCompensator()
        comp.setMatGains(gains)
    ones = [np.full(im.shape[:2], 255, np.uint8) for im in ims[:1]]
    lib = xb._lib.load()
    import ctypes

    def one():
        t_lib = t_cv = 0.0
        t0 = time.perf_counter()
        warper = creator.create(cfg["focal"])
        rois = [warper.warpRoi((im.shape[1], im.shape[0]), K, R)
                for im, K, R in zip(ims, Ks, Rs)]
        x0 = min(r[0] for r in rois)
        y0 = min(r[1] for r in rois)
        x1 = max(r[0] + r[2] for r in rois)
        y1 = max(r[1] + r[3] for r in rois)
        sc = ctypes.c_float()
        if lib.xp_compose_capped_scale(float(np.float32(cfg["focal"])), 1.0,
                                       x1 - x0, y1 - y0, cfg["max_pano_mpx"],
                                       ctypes.byref(sc)):
            warper = creator.create(sc.value)
            rois = [warper.warpRoi((im.shape[1], im.shape[0]), K, R)
                    for im, K, R in zip(ims, Ks, Rs)]
            x0 = min(r[0] for r in rois)
            y0 = min(r[1] for r in rois)
            x1 = max(r[0] + r[2] for r in rois)
            y1 = max(r[1] + r[3] for r in rois)
        blender = xb.Multiblend(output_bpp=0,
                                fixed_levels=cfg["fixed_levels"])
        blender.prepare((x0, y0, x1 - x0, y1 - y0))
        t_lib += time.perf_counter() - t0
        for i, (im, K, R) in enumerate(zip(ims, Ks, Rs)):
            if seams is not None and not seams[i].any():
                continue
            t0 = time.perf_counter()
            corner, warped = warper.warp(im, K, R, xb.INTER_LINEAR,
                                         xb.BORDER_REFLECT)
            _, mw = warper.warp(ones[0], K, R, xb.INTER_NEAREST,
                                xb.BORDER_CONSTANT)
            t1 = time.perf_counter()
            if comp is not None:
                warped = comp.apply(i, corner, warped, mw)
            if seams is not None:
                dil = cv2.dilate(seams[i], None)
                sm = cv2.resize(dil, (mw.shape[1], mw.shape[0]),
                                interpolation=cv2.INTER_LINEAR_EXACT)
                mw = cv2.bitwise_and(sm, mw)
            t2 = time.perf_counter()
            blender.feed(warped, mw, corner)
            t3 = time.perf_counter()
            t_lib += (t1 - t0) + (t3 - t2)
            t_cv += t2 - t1
        t0 = time.perf_counter()
        pano, mask = blender.blend()
        t_lib += time.perf_counter() - t0
        return pano, mask, t_lib, t_cv

    one()  # warm-up
    lib_s = cv_s = 0.0
    t0 = time.perf_counter()
    for _ in range(steps):
        pano, mask, a, b = one()
        lib_s += a
        cv_s += b
    wall = time.perf_counter() - t0
    return {"ms_per_step": 1e3 * wall / steps,
            "ms_in_library_calls": 1e3 * lib_s / steps,
            "ms_in_opencv_cpu_steps_between": 1e3 * cv_s / steps,
            "steps": steps, "pano": pano, "mask": mask,
            "what": "warp x2 + feed per image and blend through host "
            "buffers (shim call sequence, stitcher.cc:358-419); gain apply "
            "and seam-mask dilate/resize/and on the CPU by cv2 as in the "
            "reference"}


def jpeg_leg(xb, torch, cfg, wl, local, steps, slots=3):
    """Compressed in -> compressed out, pipelined: the sources
    arrive as JPEG bytes in ordinary host memory (cv::imread's input,
    image.cc:83) and are decoded on the device by the library's kernels, the
    panorama is composed there and leaves as JPEG bytes (cv::imwrite's
    output, stitcher_pipeline.cc:94-116): no raw pixel crosses PCIe. Three
    host threads -- decode, compose, encode -- hand `slots` panoramas round.
    Returns wall ms per panorama and the bytes moved, or None."""
    import queue
    import threading
    if cfg["bpp"] != 8:
        return None
    ims, Ks, Rs = wl["ims"], wl["Ks"], wl["Rs"]
    h, w = ims[0].shape[:2]
    enc = xb.JpegCodec(device=local)
    dec = xb.JpegCodec(device=local)
    # camera-like sources: baseline, 4:2:2, quality 95
    jpegs = [np.frombuffer(enc.encode(im, quality=95, subsampling=422),
                           np.uint8) for im in ims]
    kw = dict(projection=cfg["projection"], warp_scale=cfg["focal"],
              max_pano_mpx=cfg["max_pano_mpx"], depth=8,
              fixed_levels=cfg["fixed_levels"], device=local)
    comps = [xb.Composer(**kw) for _ in range(slots)]
    bufs = [[torch.empty((h, w, 3), dtype=torch.uint8, device="cuda")
             for _ in ims] for _ in range(slots)]
    free, decoded, composed = queue.Queue(), queue.Queue(), queue.Queue()
    errs, sizes = [], []

    def guard(fn):
        def run(n):
            try:
                torch.cuda.set_device(local)
                fn(n)
            except Exception as e:
                errs.append(e)
        return run

    # The decoder's kernels and a panorama's warps and seam passes take turns
    # (as two panoramas do in the raw end-to-end leg): the seam wavefronts are
    # latency chains that every co-resident warp stretches. The pyramid tail
    # and the encoder run beside whatever has the turn.
    turn = threading.Lock()

    @guard
    def decode(n):
        for _ in range(n):
            k = free.get(timeout=120)
            with turn:
                dec.decode_batch_device(jpegs,
                                        [t.data_ptr() for t in bufs[k]])
            decoded.put(k)

    @guard
    def compose(n):
        for _ in range(n):
            k = decoded.get(timeout=120)
            comps[k].set_images_device([t.data_ptr() for t in bufs[k]],
                                       [(w, h)] * len(ims), Ks, Rs,
                                       wl["gains"], wl["seams"])
            with turn:
                comps[k].run_begin()
            comps[k].run_end()
            composed.put(k)

    @guard
    def encode(n):
        for _ in range(n):
            k = composed.get(timeout=120)
            sizes.append(len(comps[k].encode_jpeg(enc)))
            free.put(k)

    def one_pass(n):
        for k in range(slots):
            free.put(k)
        ts = [threading.Thread(target=f, args=(n,), daemon=True)
              for f in (decode, compose, encode)]
        for t in ts:
            t.start()
        for t in ts:
            t.join(timeout=300)
        while not free.empty():
            free.get()
        if errs:
            raise errs[0]
        if any(t.is_alive() for t in ts):
            raise RuntimeError("jpeg leg: a stage did not finish")

    # warm-up until the device pool holds every buffer size the three stages
    # ask for at once (a cudaMalloc in the timed pass stalls the whole device)
    one_pass(4 * slots)
    torch.cuda.synchronize()
    # three timed passes, the median reported: the pipeline settles into one
    # of two interleavings of its stages (~10 and ~15 ms per panorama on cfg2)
    walls = []
    for _ in range(3):
        t0 = time.perf_counter()
        one_pass(steps)
        torch.cuda.synchronize()
        walls.append(time.perf_counter() - t0)
    wall = sorted(walls)[1]
    aux = sum(a.nbytes for k in ("gains", "seams") if wl[k] is not None
              for a in wl[k])
    own, lib = dec.stats()
    return {"ms_per_step": 1e3 * wall / steps, "steps": steps,
            "passes_ms_per_step": [round(1e3 * x / steps, 3) for x in walls],
            "h2d_bytes_per_step": int(sum(j.size for j in jpegs) + aux),
            "d2h_bytes_per_step": int(sizes[-1]),
            "sources_decoded_by": {"own_kernels": own, "nvjpeg": lib},
            "panoramas_in_flight": slots,
            "what": "JPEG bytes in (baseline 4:2:2 q95, pageable host memory) "
                    "-> decode on the device -> compose -> JPEG bytes out "
                    "(4:2:2 q95); decode / compose / encode of successive "
                    "panoramas overlap"}



def sharding_group(world, rank):
    from xpano_b200 import sharding
    return sharding.assign_groups(world, world, rank)[0]


def init_ranks():
    """-> (torch, dist, rank, world, local): one process per GPU, NCCL."""
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; xpano_b200 has no CPU "
                         "fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        # NCCL announces its version on stdout at communicator creation;
        # stdout carries exactly one JSON line, so park fd 1 on stderr until
        # the first collective has run
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl",
                                    device_id=torch.device("cuda", local))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    return torch, dist, rank, world, local


def pin_images(torch, ims):
    pinned = []
    for im in ims:
        t = torch.empty(im.shape, dtype=torch.uint8 if im.dtype == np.uint8
                        else torch.uint16, pin_memory=True)
        t.numpy()[...] = im
        pinned.append(t)
    return pinned


def run_tiled(args, cfg):
    """ONE panorama over all ranks: canvas strips with recomputed halos
    (BASELINE.json configs[2]); strong scaling. Every rank
    computes the seams of the whole canvas, the pyramids / composite /
    collapse of its strip; the exchange is one all-reduce of 8 doubles."""
    torch, dist, rank, world, local = init_ranks()
    import xpano_b200 as xb
    from xpano_b200 import sharding
    lib = xb._lib.load()
    wl = make_workload(cfg, 0, args.masks)     # the same panorama everywhere
    ims, Ks, Rs = wl["ims"], wl["Ks"], wl["Rs"]
    pinned = pin_images(torch, ims)
    host = [t.numpy() for t in pinned]
    stream = torch.cuda.current_stream().cuda_stream
    kw = dict(projection=cfg["projection"], warp_scale=cfg["focal"],
              max_pano_mpx=cfg["max_pano_mpx"], depth=cfg["bpp"],
              fixed_levels=cfg["fixed_levels"], device=local)
    comp = xb.Composer(**kw)
    comp.set_stream(stream)
    if world > 1:
        comp.set_tile(rank, world)

    def step():
        return comp.run_tile() if world > 1 else comp.run()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    comp.set_images(host, Ks, Rs, wl["gains"], wl["seams"])
    needs = comp.tile_needs(len(ims)) if world > 1 else [True] * len(ims)
    info = None
    for _ in range(args.warmup):
        info = step()
    barrier()
    clocks = ClockSampler(local if rank == 0 else None,
                          20 if world == 1 else 250)
    launches0 = lib.xp_kernel_launch_count()
    ev0, ev1 = torch.cuda.Event(True), torch.cuda.Event(True)
    stage_ms = {}
    ev0.record()
    for _ in range(args.steps):
        info = step()
        for k, v in info["timing"].items():
            stage_ms[k] = stage_ms.get(k, 0.0) + v
        stage_ms["warp"] = stage_ms.get("warp", 0.0) + info["warp_ms"]
    ev1.record()
    barrier()
    launches = lib.xp_kernel_launch_count() - launches0
    clk = clocks.stop()
    ms_rank = ev0.elapsed_time(ev1)
    print(f"[rank {rank}] {ms_rank / args.steps:.3f} ms/step, strip "
          f"{info['width']}x{info['height']} at x={info['tile_x0']}, window "
          f"{info['window']}, uploads {sum(needs)}/{len(ims)} sources",
          file=sys.stderr)
    ms = sharding.reduce_max(ms_rank)
    CW, CH = info["canvas"]
    canvas_mp = CW * CH / 1e6
    value = canvas_mp * args.steps / (ms / 1e3)

    # e2e: upload of this rank's sources, tiled run, download of its strip
    dt = torch.uint8 if info["bpp"] == 8 else torch.uint16
    pano = torch.empty((info["height"], max(info["width"], 1), 3), dtype=dt,
                       pin_memory=True)
    mask = torch.empty((info["height"], max(info["width"], 1)),
                       dtype=torch.uint8, pin_memory=True)
    h2d = sum(im.nbytes for im, nd in zip(host, needs) if nd)
    d2h = pano.numpy().nbytes + mask.numpy().nbytes
    e2e_steps = min(max(args.steps, 3), 8)

    def e2e_step():
        comp.set_images(host, Ks, Rs, wl["gains"], wl["seams"])
        step()
        if info["width"] > 0:
            comp.fetch(pano.numpy(), mask.numpy())

    e2e_step()
    barrier()
    ev0.record()
    for _ in range(e2e_steps):
        e2e_step()
    torch.cuda.synchronize()
    ev1.record()
    barrier()
    e2e_ms = sharding.reduce_max(ev0.elapsed_time(ev1))
    h2d_all = sharding.reduce_sum(h2d)
    d2h_all = sharding.reduce_sum(d2h)

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None,
        "dtype": "u8" if cfg["bpp"] == 8 else "u16", "data": "synthetic",
        "config": config_of(cfg, wl), "canvas": [CW, CH],
        "levels": info["levels"],
        "resolution_capped": info["resolution_capped"],
        "sharding": (f"one panorama, {world} canvas strips + recomputed halo "
                     f"of {sharding.tile_halo(info['levels'])} px, seams "
                     "replicated, all-reduce of 8 doubles per step")
        if world > 1 else "1 GPU",
        "e2e": {"value": canvas_mp * e2e_steps / (e2e_ms / 1e3), "unit": UNIT,
                "h2d_bytes_per_step": h2d_all, "d2h_bytes_per_step": d2h_all,
                "ms_per_step": e2e_ms / e2e_steps, "steps": e2e_steps},
        "gpu_launches": int(launches), "clocks": clk,
        "rank0": {"ms_per_step": ms_rank / args.steps,
                  "strip": [info["width"], info["height"]],
                  "window": list(info["window"]),
                  "sources_uploaded": int(sum(needs)),
                  "stages_ms": {k: round(v / args.steps, 3)
                                for k, v in stage_ms.items()}},
    }
    if args.check and world > 1:
        # the strips of all ranks against the single-GPU panorama (rank 0)
        sp, sm = comp.fetch() if info["width"] > 0 else (
            np.zeros((CH, 0, 3), pano.numpy().dtype),
            np.zeros((CH, 0), np.uint8))
        gp, gm = sharding.gather_strips(sp, sm, info["tile_x0"], CW)
        if rank == 0:
            one = xb.Composer(**kw)
            one.set_images(host, Ks, Rs, wl["gains"], wl["seams"])
            one.run()
            p1, m1 = one.fetch()
            same = bool(np.array_equal(gp, p1) and np.array_equal(gm, m1))
            line["tiled_equals_single_gpu"] = same
            if not same:
                d = np.abs(gp.astype(np.int32) - p1)
                line["tiled_diff"] = {
                    "max": int(d.max()),
                    "columns": [int(v) for v in
                                np.nonzero(d.max(axis=(0, 2)))[0][[0, -1]]]}
    if rank == 0:
        print(json.dumps(line))
        if line.get("tiled_equals_single_gpu") is False:
            raise SystemExit("bench.py: tiled panorama differs from the "
                             "single-GPU one")
    if world > 1:
        dist.destroy_process_group()


def run_ours(args, cfg):
    torch, dist, rank, world, local = init_ranks()
    import xpano_b200 as xb
    lib = xb._lib.load()

    group = sharding_group(world, rank)
    wl = make_workload(cfg, group, args.masks)
    ims, Ks, Rs = wl["ims"], wl["Ks"], wl["Rs"]
    # host sources in pinned memory (e2e leg copies from here every step)
    pinned = pin_images(torch, ims)
    host = [t.numpy() for t in pinned]

    stream = torch.cuda.current_stream().cuda_stream
    comp = xb.Composer(projection=cfg["projection"], warp_scale=cfg["focal"],
                       max_pano_mpx=cfg["max_pano_mpx"], depth=cfg["bpp"],
                       fixed_levels=cfg["fixed_levels"], device=local)
    comp.set_stream(stream)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # N > 1: one panorama group per rank and step (weak scaling), no
    # data-path collective; only the counters below are reduced
    from xpano_b200 import sharding
    max_over_ranks = sharding.reduce_max
    sum_over_ranks = sharding.reduce_sum

    # ---- resident leg ------------------------------------------------------
    comp.set_images(host, Ks, Rs, wl["gains"], wl["seams"])
    info = None
    for _ in range(args.warmup):
        info = comp.run()
    barrier()
    # one nvidia-smi poller per job: every query takes driver locks on all
    # GPUs of the box, and kernel launches of every rank wait for them
    clocks = ClockSampler(local if rank == 0 else None,
                          20 if world == 1 else 250)
    launches0 = lib.xp_kernel_launch_count()
    ev0, ev1 = torch.cuda.Event(True), torch.cuda.Event(True)
    stage_ms = {}
    ev0.record()
    for _ in range(args.steps):
        info = comp.run()
        for k, v in info["timing"].items():
            stage_ms[k] = stage_ms.get(k, 0.0) + v
        stage_ms["warp"] = stage_ms.get("warp", 0.0) + info["warp_ms"]
    ev1.record()
    barrier()
    launches = lib.xp_kernel_launch_count() - launches0
    clk = clocks.stop()
    ms_rank = ev0.elapsed_time(ev1)
    if world > 1:  # per-rank record: the job's time is the slowest rank's
        print(f"[rank {rank}] {ms_rank / args.steps:.3f} ms/step, "
              f"{host_cores()} host cores visible", file=sys.stderr)
    ms = max_over_ranks(ms_rank)
    mp_rank = info["width"] * info["height"] / 1e6
    total_mp = sum_over_ranks(mp_rank)
    value = total_mp * args.steps / (ms / 1e3)

    # ---- e2e leg: host buffers in, host panorama out, every step -----------
    # Batch export keeps several panoramas in flight (one composer, stream and
    # pinned output buffer each, driven by one host thread each through the
    # public API), so the upload of one overlaps the kernels of another and
    # the download of a third. Every step still uploads its own inputs and
    # downloads its own result inside the timed region.
    import threading
    W, H = info["width"], info["height"]
    dt = torch.uint8 if info["bpp"] == 8 else torch.uint16
    in_flight = max(1, args.in_flight)
    comps = [comp]
    for _ in range(in_flight - 1):
        c2 = xb.Composer(projection=cfg["projection"],
                         warp_scale=cfg["focal"],
                         max_pano_mpx=cfg["max_pano_mpx"], depth=cfg["bpp"],
                         fixed_levels=cfg["fixed_levels"], device=local)
        comps.append(c2)
    outs = []
    for _ in comps:
        pano = torch.empty((H, W, 3), dtype=dt, pin_memory=True)
        mask = torch.empty((H, W), dtype=torch.uint8, pin_memory=True)
        outs.append((pano.numpy(), mask.numpy(), pano, mask))
    h2d = sum(im.nbytes for im in host) + sum(
        a.nbytes for k in ("gains", "seams") if wl[k] is not None
        for a in wl[k])
    d2h = outs[0][0].nbytes + outs[0][1].nbytes
    e2e_steps = max(args.steps, 2 * in_flight)

    # one panorama at a time runs its warps and seam passes, while the others
    # upload / download. Letting two do that at once is slower (measured:
    # 11.3 -> 13.0 ms per panorama): the seam wavefronts are latency chains
    # and every co-resident warp on their SMs stretches each step. The turn is
    # handed over at run_begin's return, when only the bandwidth-bound pyramid
    # tail of the panorama is still queued: it overlaps the next one's warps.
    compute_turn = threading.Lock()
    # each direction of the bus is one resource: two uploads at once only
    # finish later, both of them, so transfers take turns as well
    upload_turn, fetch_turn = threading.Lock(), threading.Lock()

    def e2e_worker(j, n_steps, errs):
        try:
            torch.cuda.set_device(local)
            for _ in range(j, n_steps, in_flight):
                with upload_turn:
                    comps[j].set_images(host, Ks, Rs, wl["gains"],
                                        wl["seams"])
                with compute_turn:
                    comps[j].run_begin()
                comps[j].run_end()
                with fetch_turn:
                    comps[j].fetch(outs[j][0], outs[j][1])
        except Exception as e:  # surfaced after join
            errs.append(e)

    def e2e_pass(n_steps):
        errs = []
        ts = [threading.Thread(target=e2e_worker, args=(j, n_steps, errs))
              for j in range(in_flight)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        if errs:
            raise errs[0]

    e2e_pass(in_flight * min(args.warmup, 2))
    barrier()
    t0 = time.perf_counter()
    ev0.record()
    e2e_pass(e2e_steps)
    torch.cuda.synchronize()
    ev1.record()
    barrier()
    wall = time.perf_counter() - t0
    for o in outs[1:]:  # same inputs: the in-flight panoramas must agree
        if not (np.array_equal(o[0], outs[0][0]) and
                np.array_equal(o[1], outs[0][1])):
            raise SystemExit("bench.py: concurrent panoramas differ")
    e2e_ms = max_over_ranks(max(ev0.elapsed_time(ev1), 0.0))
    e2e_value = total_mp * e2e_steps / (e2e_ms / 1e3)

    # ---- resident, two panoramas overlapped (informational) ----------------
    # The same K steps with the next panorama's run_begin issued while the
    # previous one's pyramid tail is still running (two composers, sources
    # resident in both, one host thread). Not the headline `value`, whose
    # steps are strictly one after another; wall clock between two device
    # synchronisations.
    pipelined = None
    if len(comps) >= 3:
        a, b = comps[1], comps[2]
        for c in (a, b):
            c.run()
        torch.cuda.synchronize()
        tp0 = time.perf_counter()
        a.run_begin()
        for _ in range(args.steps - 1):
            b.run_begin()
            a.run_end()
            a, b = b, a
        a.run_end()
        torch.cuda.synchronize()
        p_ms = max_over_ranks(1e3 * (time.perf_counter() - tp0))
        pipelined = {"ms_per_step": p_ms / args.steps,
                     "value": total_mp * args.steps / (p_ms / 1e3),
                     "unit": UNIT, "steps": args.steps,
                     "what": "run_begin of panorama k+1 issued while the "
                             "pyramid tail of panorama k runs (two composers)"}

    # ---- compressed in -> compressed out (informational) -------------------
    e2e_jpeg = None
    if not args.no_jpeg:
        try:
            e2e_jpeg = jpeg_leg(xb, torch, cfg, wl, local, e2e_steps)
        except Exception as e:  # never costs the headline numbers
            e2e_jpeg = {"error": str(e)[:200]}
        if e2e_jpeg and "ms_per_step" in e2e_jpeg:
            j_ms = max_over_ranks(e2e_jpeg["ms_per_step"])
            e2e_jpeg["ms_per_step"] = j_ms
            e2e_jpeg["value"] = total_mp / (j_ms / 1e3)
            e2e_jpeg["unit"] = UNIT
        elif world > 1:
            max_over_ranks(0.0)  # keep the ranks' collectives in step

    # ---- e2e through the drop-in operator surface -------------------------
    op_surface = None
    if world == 1 and not args.no_surface:
        op_surface = operator_surface_leg(xb, wl, cfg, min(args.steps, 3))
        if op_surface is not None:
            op_surface["value"] = mp_rank / (op_surface["ms_per_step"] / 1e3)
            op_surface["unit"] = UNIT
            po, mo = op_surface.pop("pano"), op_surface.pop("mask")
            same = po.shape == outs[0][0].shape
            op_surface["vs_composer"] = {
                "mask_equal": bool(same and np.array_equal(mo, outs[0][1])),
                "max_diff": int(np.abs(po.astype(np.int32) -
                                       outs[0][0]).max()) if same else None}

    # ---- roofline of the dominant HBM-bound stage --------------------------
    peak, peak_src = measured_peaks()
    sbytes = dict(info["stage_bytes"])
    sbytes["warp"] = info["warp_bytes"]
    slaunch = dict(info["stage_launches"])
    slaunch["warp"] = info["warp_launches"]
    per_stage = {}
    for k in HBM_STAGES:
        t = stage_ms.get(k, 0.0) / args.steps
        if t > 0 and sbytes.get(k, 0) > 0:
            gbs = sbytes[k] / (t / 1e3) / 1e9
            per_stage[k] = {"ms": round(t, 4), "launches": slaunch.get(k, 0),
                            "alg_GB": round(sbytes[k] / 1e9, 4),
                            "GBps": round(gbs, 1),
                            "frac": round(gbs / peak, 4)}
    # streaming-minimum algorithmic bytes of the pyramid/blend
    # group (K9-K16) and of warp+pack (K1-K6); seams/inpaint are latency-bound
    # recurrences and are reported as time, not against the roofline
    b = cfg["bpp"] // 8
    s_src, s_w, s_t, c_px = (info["s_src"], info["s_w"], info["s_t"],
                             info["c"])
    blend_stages = [k for k in per_stage if k != "warp"]
    # Stages overlap (Gaussian pyramids run on a side stream under the seam
    # passes), so the group's time is not the sum of its spans: it is the
    # wall time of the tail after the seam stage plus the full span of the
    # overlapped shrink stage (conservative: as if it were serialised).
    tail_ms = stage_ms.get("pyramid_tail", 0.0) / args.steps
    shrink_ms = per_stage.get("shrink", {}).get("ms", 0.0)
    blend_ms = tail_ms + shrink_ms
    blend_launches = sum(per_stage[k]["launches"] for k in blend_stages)
    b_blend = (6 * b + 61) * s_t + (3 * b + 32) * c_px
    b_warp = 3 * b * s_src + 4 * b * s_w
    blend_gbs = b_blend / (blend_ms / 1e3) / 1e9
    roof = {"bound": "hbm", "kernel": "pyramid+blend group (" +
            ",".join(blend_stages) + ")",
            "achieved": round(blend_gbs, 1), "peak": peak,
            "peak_source": peak_src, "unit": "GB/s",
            "frac": round(blend_gbs / peak, 4),
            "traffic": ncu_traffic(cfg["name"]),
            "model": "SURVEY 8(d): (6b+61)*S_t + (3b+32)*C bytes per step",
            "bytes_per_step": b_blend, "ms_per_step": round(blend_ms, 4),
            "ms_tail_after_seams": round(tail_ms, 4),
            "ms_shrink_under_seams": round(shrink_ms, 4),
            "launches_per_step": blend_launches,
            "warp": {"bytes_per_step": b_warp,
                     "ms_per_step": per_stage.get("warp", {}).get("ms"),
                     "GBps": round(b_warp / max(per_stage.get("warp", {})
                                                .get("ms", 0), 1e-9) / 1e6, 1)
                     if "warp" in per_stage else None}}
    roof["warp"]["frac"] = (round(roof["warp"]["GBps"] / peak, 4)
                            if roof["warp"]["GBps"] else None)
    serial = {k: round(stage_ms.get(k, 0.0) / args.steps, 3)
              for k in ("images", "seam")}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8" if cfg["bpp"] == 8
        else "u16", "data": "synthetic",
        "config": config_of(cfg, wl), "canvas": [W, H],
        "levels": info["levels"], "resolution_capped":
        info["resolution_capped"], "sharding": "panorama group per GPU, no "
        "data-path collective" if world > 1 else "1 GPU",
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms / e2e_steps,
                "steps": e2e_steps, "wall_ms_per_step": 1e3 * wall / e2e_steps,
                "panoramas_in_flight": in_flight},
        "e2e_operator_surface": op_surface,
        "resident_pipelined": pipelined, "e2e_jpeg": e2e_jpeg,
        "gpu_launches": int(launches), "clocks": clk, "roofline": roof,
        "stages": per_stage, "latency_bound_stages_ms": serial,
        "sizes_mp": {"s_src": info["s_src"] / 1e6, "s_w": info["s_w"] / 1e6,
                     "s_t": info["s_t"] / 1e6, "c": info["c"] / 1e6},
    }
    if rank == 0 and world == 1 and not args.no_cpu:
        # CPU leg beside it, on the box's host cores: one warm-up pass and the
        # mean of two timed ones over the bounded sample (default: all images)
        cores = host_cores()
        swl, sample = sample_of(cfg, wl, args.ref_images)
        passes = max(1, args.cpu_passes)
        if passes > 1:
            reference_step(swl, cfg, cores)
        tot_mp = tot_s = 0.0
        ref = None
        for k in range(passes):
            mp, sec, kind, r = reference_step(
                swl, cfg, cores,
                keep=(k == passes - 1 and swl is wl and args.check))
            tot_mp += mp
            tot_s += sec
            ref = r or ref
        line["cpu_baseline"] = {"value": tot_mp / tot_s, "unit": UNIT,
                                "cores": cores if kind == "reference" else 1,
                                "kind": kind, "sample": sample,
                                "seconds": round(tot_s, 2), "passes": passes}
        if ref is not None:
            # full-size parity of the GPU panorama against the reference arm
            pano, mask = comp.fetch()
            line["parity"] = parity_check(
                pano, mask, (info["min_x"], info["min_y"]), ref, cfg["bpp"])
            line["parity"]["capped"] = [bool(info["resolution_capped"]),
                                        bool(ref["capped"])]
            if not line["parity"]["ok"]:
                print(json.dumps(line))
                raise SystemExit("bench.py: full-size parity check failed: "
                                 + json.dumps(line["parity"]))
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="cfg2", choices=sorted(CONFIGS))
    ap.add_argument("--ref-images", type=int, default=0,
                    help="images in the bounded CPU sample (0 = all)")
    ap.add_argument("--masks", default="graphcut",
                    choices=["graphcut", "full", "none"],
                    help="seam masks fed to the path (SURVEY.md 8d)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--in-flight", type=int, default=4,
                    help="panoramas in flight in the end-to-end leg (one "
                    "composer, host thread and pinned output buffer each)")
    ap.add_argument("--tiled", action="store_true",
                    help="N > 1: ONE panorama over all GPUs as canvas strips "
                    "(strong scaling; use with --config cfg3) instead of one "
                    "panorama group per GPU")
    ap.add_argument("--cpu-passes", type=int, default=2,
                    help="timed passes of the inline CPU leg (one warm-up "
                    "pass before them when > 1)")
    ap.add_argument("--no-jpeg", action="store_true",
                    help="skip the compressed-in / compressed-out leg")
    ap.add_argument("--no-surface", action="store_true",
                    help="skip the operator-surface e2e leg")
    ap.add_argument("--no-check", dest="check", action="store_false",
                    help="skip the full-size parity check against the CPU arm")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    cfg = CONFIGS[args.config]
    if args.impl == "reference":
        run_reference(args, cfg)
    elif args.tiled:
        run_tiled(args, cfg)
    else:
        run_ours(args, cfg)


if __name__ == "__main__":
    main()
