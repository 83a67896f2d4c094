"""Writes xpano_b200/csrc/jpeg_std_tables.h: the Annex K quantisation and
Huffman tables, read from the DQT / DHT segments of a stream that cv2.imencode
wrote at quality 50 (where libjpeg leaves the standard tables unscaled)."""
import os
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def arr(name, vals, per=16):
    out = f"static const uint8_t {name}[{len(vals)}] = {{\n"
    for k in range(0, len(vals), per):
        out += "    " + ", ".join(str(v) for v in vals[k:k + per]) + ",\n"
    return out + "};\n"


def main():
    img = np.random.default_rng(0).integers(0, 256, (32, 48, 3)).astype(np.uint8)
    d = cv2.imencode(".jpg", img, [cv2.IMWRITE_JPEG_QUALITY, 50])[1].tobytes()
    i, dqt, dht = 2, [], []
    while i < len(d):
        m, n = d[i + 1], (d[i + 2] << 8) | d[i + 3]
        body = d[i + 4:i + 2 + n]
        if m == 0xDB:
            dqt.append(body)
        if m == 0xC4:
            dht.append(body)
        if m == 0xDA:
            break
        i += 2 + n
    path = os.path.join(ROOT, "xpano_b200", "csrc", "jpeg_std_tables.h")
    head = open(path).read().split("namespace xp {")[0]
    h = head + "namespace xp {\nnamespace jpg {\n\n"
    h += arr("kStdQuantLumaZz", list(dqt[0][1:65]))
    h += arr("kStdQuantChromaZz", list(dqt[1][1:65]))
    names = {0x00: "DcLuma", 0x10: "AcLuma", 0x01: "DcChroma", 0x11: "AcChroma"}
    for b in dht:
        h += arr(f"kStd{names[b[0]]}Bits", list(b[1:17]))
        h += arr(f"kStd{names[b[0]]}Vals", list(b[17:]))
    h += "\n}  // namespace jpg\n}  // namespace xp\n"
    if "--check" in sys.argv:
        assert open(path).read() == h, "jpeg_std_tables.h is out of date"
    else:
        open(path, "w").write(h)


if __name__ == "__main__":
    main()
