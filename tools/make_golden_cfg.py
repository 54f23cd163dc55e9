"""Goldens for the FULL-SIZE parity tests of BASELINE.json configs 3 and 4' (SURVEY.md 8(d)):
    spheres1024_4k   1,024 mixed-material spheres over the textured plane at 3840x2160  (wide oracle build, X2)
    spheres255_8k    255 spheres (the most the reference's uchar counts allow) at 7680x4320  (narrow build)
Per-row CRC32 of the colour words and of the primary hit ids, produced HERE from oracle/_ref -- the reference's own
.cl text compiled as C++ (needs /root/reference for the build) -- in row chunks, plus the work counters of the counted
port on a row sample (every `--count-stride`-th row; the port equals the reference text word for word, asserted on
those rows).
    python tools/make_golden_cfg.py spheres1024_4k [--count-stride 16]  ->  tests/golden/spheres1024_4k_rows.npz
"""
import argparse
import os
import sys
import time
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("workload", choices=["spheres1024_4k", "spheres255_8k"])
ap.add_argument("--count-stride", type=int, default=16)
ap.add_argument("--chunk", type=int, default=120)
args = ap.parse_args()

ref = g.load_oracle()
pkg = g.load_package()
from example_gui_opencl_raytracer_b200 import host  # noqa: E402  (scene generators only: no rendering code)

assert ref.have_ref(), "build oracle/_ref first (needs /root/reference)"
if args.workload == "spheres1024_4k":
    s, p, l, camp = host.sphere_grid_scene(32)
    W, H = 3840, 2160
else:
    s, p, l, camp = host.sphere_grid_scene(16)
    s = s[:255 * 96]
    W, H = 7680, 4320
tex, sky = ref.default_textures(ROOT)
scene = ref.Scene(s, p, l, tex, sky)
wide = max(scene.S, scene.P, scene.L) > 255
cam = host.camera(W=W, H=H, **camp)

rows_rgb = np.zeros(H, np.uint32)
rows_hit = np.zeros(H, np.uint32)
t0 = time.time()
for y0 in range(0, H, args.chunk):
    y1 = min(H, y0 + args.chunk)
    rgb, hit = ref.ref().render(scene, cam, W, H, y0=y0, y1=y1, wide=wide)
    for y in range(y0, y1):
        rows_rgb[y] = zlib.crc32(rgb[y * W:(y + 1) * W].tobytes())
        rows_hit[y] = zlib.crc32(hit[y * W:(y + 1) * W].tobytes())
    print("rows %d..%d done, %.0f s" % (y0, y1, time.time() - t0), flush=True)

# work counters of every count_stride-th row (the port's; its words must equal the reference text's on those rows)
names = ("segments", "shadow_rays", "sphere_tests", "sphere_dpos", "plane_tests", "texels")
count_rows = np.arange(0, H, args.count_stride, dtype=np.uint32)
counters = np.zeros((len(count_rows), len(names)), np.uint64)
for i, y in enumerate(count_rows):
    y = int(y)
    rgb, hit, cnt = ref.port().render_counted(scene, cam, W, H, y0=y, y1=y + 1, wide=wide)
    assert zlib.crc32(rgb[y * W:(y + 1) * W].tobytes()) == int(rows_rgb[y]), "port and reference text disagree on row %d" % y
    assert zlib.crc32(hit[y * W:(y + 1) * W].tobytes()) == int(rows_hit[y]), "port and reference text disagree on row %d" % y
    counters[i] = [cnt[k] for k in names]
out = os.path.join(ROOT, "tests", "golden", args.workload + "_rows.npz")
np.savez_compressed(out, rows_rgb=rows_rgb, rows_hit=rows_hit, count_rows=count_rows, counters=counters,
                    counter_names=np.array(names), width=np.uint32(W), height=np.uint32(H))
print("ok", out, "%.0f s" % (time.time() - t0), dict(zip(names, counters.sum(axis=0).tolist())))
