"""Regenerates tests/golden/*.npz. Run in the build container only (it reads
/root/reference, which does not exist on the GPU box):

    python tests/golden/make_golden.py

blend_reference_goldens.npz
    The reference's own golden vectors for the blend path
    (external/multiblend/tests/data, used by tests/multiblend_test.cpp:39-44
    and tests/test_multiblend.py:47-185), decoded to arrays: inputs, positions
    and expected outputs for expected.png, expected_alpha.png,
    expected_seams.png, expected.tif and expected_16bit.tif.

blend_ref_outputs.npz
    Outputs of the UNMODIFIED reference library (oracle/_ref/libmb_ref.so) on
    the seeded synthetic cases of tests/synth.py (inputs are regenerated from
    the seed by the tests, only outputs are stored).
"""
import os
import sys

import cv2
import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import mb_ref  # noqa: E402
import synth  # noqa: E402

D = "/root/reference/external/multiblend/tests/data/"

# (seed, kwargs for synth.blend_case, kwargs for blend)
REF_CASES = [
    (1, dict(n=2, mode="ellipse", bpp=8), dict()),
    (2, dict(n=3, mode="full", bpp=8), dict(fixed_levels=3)),
    (3, dict(n=5, mode="blobs", bpp=8), dict(fixed_levels=5)),
    (4, dict(n=4, mode="ellipse", bpp=16), dict()),
    (5, dict(n=3, mode="blobs", bpp=16), dict(fixed_levels=4, dither=False)),
    (6, dict(n=6, mode="ellipse", bpp=8, w=320, h=240), dict(fixed_levels=7)),
    (7, dict(n=3, mode="full", bpp=8, spp=3), dict()),
    (8, dict(n=4, mode="ellipse", bpp=8), dict(reverse=True, fixed_levels=2)),
    (9, dict(n=1, mode="ellipse", bpp=8), dict()),
    (10, dict(n=3, mode="ellipse", bpp=8), dict(fixed_levels=1)),
]


def main():
    g = {}
    pos_png = [(-91, 204), (-167, 202), (-41, 193)]
    pos_tif = [(76, 11), (0, 9), (126, 0)]
    for i in range(3):
        raw, p = mb_ref.read_cereal_raw(D + f"img_{i}.png.raw")
        assert p == pos_png[i]
        g[f"raw_{i}"] = raw                                   # RGB order
        g[f"alpha_{i}"] = cv2.imread(D + f"img_{i}_alpha.png",
                                     cv2.IMREAD_UNCHANGED)    # BGRA
        g[f"tif16_{i}"] = cv2.imread(D + f"img_{i}_16bit.tif",
                                     cv2.IMREAD_UNCHANGED)    # BGR u16
    g["pos_png"] = np.array(pos_png)
    g["pos_tif"] = np.array(pos_tif)
    g["expected_png"] = cv2.imread(D + "expected.png", cv2.IMREAD_UNCHANGED)
    g["expected_alpha_png"] = cv2.imread(D + "expected_alpha.png",
                                         cv2.IMREAD_UNCHANGED)
    g["expected_16bit_tif"] = cv2.imread(D + "expected_16bit.tif",
                                         cv2.IMREAD_UNCHANGED)
    g["expected_seams_png"] = np.array(Image.open(D + "expected_seams.png"))
    # multiblend_test.cpp:39-44
    g["kat_size_minpos"] = np.array([331, 244, -167, 193])
    np.savez_compressed(os.path.join(HERE, "blend_reference_goldens.npz"), **g)

    r = {}
    for seed, ckw, bkw in REF_CASES:
        ims, pos = synth.blend_case(seed, **ckw)
        res = mb_ref.blend(ims, pos, n_threads=4, **bkw)
        r[f"planes_{seed}"] = res.planes
        r[f"mask_{seed}"] = res.mask
        r[f"meta_{seed}"] = np.array([res.min_x, res.min_y, int(res.no_mask)])
    np.savez_compressed(os.path.join(HERE, "blend_ref_outputs.npz"), **r)
    for f in ("blend_reference_goldens.npz", "blend_ref_outputs.npz"):
        print(f, os.path.getsize(os.path.join(HERE, f)) // 1024, "KiB")


if __name__ == "__main__":
    main()
