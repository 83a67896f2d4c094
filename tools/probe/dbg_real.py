import sys, os
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import numpy as np, cv2
import test_real_photos as t
import xpano_b200 as xb
from oracle import warp_oracle as wo
fxf = t.fx.__wrapped__ if hasattr(t.fx, "__wrapped__") else None
g = np.load("tests/golden/real_photos.npz")
# rebuild fixture by hand
import hashlib
imgs = []
for i in range(3):
    im = cv2.imdecode(g[f"jpeg_{i}"], cv2.IMREAD_COLOR); h, w = im.shape[:2]; s = 1024 / max(h, w)
    imgs.append(cv2.resize(im, (int(round(w*s)), int(round(h*s))), interpolation=cv2.INTER_AREA))
fx = dict(g=g, imgs=imgs, work_scale=float(g["work_scale"]), seam_scale=float(g["seam_scale"]), wis=float(g["warped_image_scale"]),
          K=[g["K"][i] for i in range(3)], R=[g["R"][i] for i in range(3)])
gains = [g[f"gain_{i}"] for i in range(3)]
seams = [(np.unpackbits(g[f"seam_{i}"])[:int(np.prod(g[f"seam_shape_{i}"]))].reshape(g[f"seam_shape_{i}"]) * 255).astype(np.uint8) for i in range(3)]
cwa = 1.0 / fx["work_scale"]
Ks = [t.scaled_K(k, cwa) for k in fx["K"]]
for variant, (gg, ss) in {"gains+seams": (gains, seams), "seams only": (None, seams), "none": (None, None)}.items():
    c = xb.Composer(projection=wo.SPHERICAL, warped_image_scale=fx["wis"], compose_work_aspect=cwa, depth=8, keep_debug=True)
    c.set_images(imgs, Ks, fx["R"], gg, ss)
    info = c.run(); pano, mask = c.fetch()
    want = t.co.compose(imgs, Ks, fx["R"], projection=wo.SPHERICAL, warped_image_scale=fx["wis"], compose_work_aspect=cwa, gains=gg, seam_masks=ss, keep_fed=True)
    dm = mask != want["mask"]
    print(variant, "mask diff", int(dm.sum()), "of", mask.size, "levels", info["levels"], want["levels"])
    if dm.any():
        ys, xs = np.nonzero(dm); print("  at", list(zip(xs[:8], ys[:8])), "dev", mask[ys[:8], xs[:8]], "want", want["mask"][ys[:8], xs[:8]])
    # compare fed alpha per image via warp_bgra
    w_ = xb.RotationWarper(xb.ProjectionType.kSpherical, float(np.float32(want["warp_scale"])))
    for i in range(3):
        roi, bgra = w_.warp_bgra(imgs[i], Ks[i], fx["R"][i], gain=None if gg is None else gg[i], seam_mask=None if ss is None else ss[i])
        fo, corner = want["fed"][i]
        da = (bgra[..., 3] != 0) != (fo[..., 3] != 0)
        dc = np.abs(bgra[..., :3].astype(int) - fo[..., :3].astype(int))[(bgra[...,3]!=0)&(fo[...,3]!=0)]
        print("  image", i, "alpha diff", int(da.sum()), "colour max diff", int(dc.max()), "frac", float((dc>0).mean()))
