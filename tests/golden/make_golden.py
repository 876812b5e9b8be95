"""Generates tests/golden/ref_png_crops.npz from the reference's own rendered images (/root/reference/image/*.png).

The PNGs are the outputs of the reference's example binaries at their native resolutions ("Golden
images").  They cannot travel to the GPU box, so small windows of them are committed as fixtures: each entry is the
8-bit content of one window (Color::clamp units, src/color.rs:42-46) plus the scene name, the native resolution and
the window, so that a test can render exactly that window with the oracle or the CUDA path and compare.

Run here (where /root/reference exists):  python tests/golden/make_golden.py
"""
import json
import os

import numpy as np
from PIL import Image

REF = "/root/reference/image"
HERE = os.path.dirname(os.path.abspath(__file__))

# (fixture key, scene name in scenes/scenes.hpp, png file, windows [(x0, y0, w, h), ...])
CROPS = [
    ("glass_air_bubble", "glass_air_bubble", "chapter11_glass_air_bubble.png",
     [(2016, 2016, 64, 64), (1180, 2040, 48, 16), (3024, 2016, 48, 48), (2600, 1500, 48, 48)]),
    ("hexagon", "hexagon", "chapter14_hexagon.png", [(2688, 800, 64, 32), (1300, 1150, 64, 32), (1728, 1440, 64, 32)]),
    ("benchmark", "benchmark", "chapter14_benchmark.png", [(2000, 1000, 96, 48), (1984, 1280, 64, 32), (100, 1900, 48, 24)]),
    ("teapot", "teapot", "chapter15_teapot.png", [(2040, 1100, 24, 12), (1560, 1560, 24, 12), (888, 576, 24, 12), (2304, 792, 24, 12)]),
    # the remaining bins (glass spheres / cube / cylinder, cones, a mirror, 4 lights, nested groups)
    ("chapter11_title", "chapter11_title", "chapter11_title.png", [(2300, 1300, 64, 32), (1500, 1450, 64, 32), (1900, 700, 48, 24), (3000, 1100, 48, 24)]),
    ("chapter12_title", "chapter12_title", "chapter12_title.png", [(2030, 1000, 64, 32), (1200, 900, 48, 24), (2600, 1200, 48, 24), (700, 700, 48, 24)]),
    ("chapter13_title", "chapter13_title", "chapter13_title.png", [(1500, 1100, 64, 32), (1050, 1250, 48, 24), (2700, 1300, 48, 24), (2000, 600, 48, 24)]),
    ("chapter14_title", "chapter14_title", "chapter14_title.png", [(2040, 1060, 64, 32), (900, 800, 48, 24), (3100, 1300, 48, 24), (2040, 850, 48, 24)]),
    ("cover", "cover", "cover.png", [(1700, 1700, 64, 32), (2600, 2000, 48, 24), (1200, 2800, 48, 24), (1500, 2700, 48, 24)]),
]


def main():
    arrays, manifest = {}, []
    for key, scene, png, windows in CROPS:
        im = Image.open(os.path.join(REF, png))
        a = np.asarray(im)
        if a.ndim == 2:  # pnmtopng wrote a greyscale file because r == g == b everywhere
            a = np.repeat(a[:, :, None], 3, axis=2)
        a = a[:, :, :3]
        for i, (x0, y0, w, h) in enumerate(windows):
            name = f"{key}_{i}"
            arrays[name] = np.ascontiguousarray(a[y0:y0 + h, x0:x0 + w]).astype(np.uint8)
            manifest.append({"key": name, "scene": scene, "png": png, "hsize": im.size[0], "vsize": im.size[1], "window": [x0, y0, w, h],
                             "distinct_values": int(len(np.unique(arrays[name])))})
    np.savez_compressed(os.path.join(HERE, "ref_png_crops.npz"), **arrays)
    with open(os.path.join(HERE, "ref_png_crops.json"), "w") as f:
        json.dump(manifest, f, indent=1)
    for m in manifest:
        print(m)


if __name__ == "__main__":
    main()
