"""Sweep of split_min_cost: kernel ms at 4K for the whole frame and for one rank's share at N = 2, 4, 8."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry
pkg = entry.load_package()
from example_gui_opencl_raytracer_b200 import host
W, H = 3840, 2160
tex, sky = host.default_images(ROOT)
s, p, l = host.load_map(os.path.join(ROOT, "scenes", "render.map"))
eng = pkg.Engine(0)
eng.upload_scene(s, p, l); eng.upload_images(0, tex); eng.upload_images(1, sky)
eng.set_camera(host.camera(W=W, H=H, **host.RAYPNG_CAMERA))
ref_frame = None
BAND = int(sys.argv[1]) if len(sys.argv) > 1 else 8
for v, v32 in ((0, 0),):
    eng.set_option("split_min_cost", v); eng.set_option("split32_min_cost", v32)
    row = []
    allparts = {}
    for parts in (1, 2, 4, 8):
        worst = 0.0
        for part in range(parts):
            ms = []
            for i in range(4):
                eng.render(W, H, part=pkg.Partition(BAND, parts, part) if parts > 1 else None)
                ms.append(eng.last_kernel_ms())
            worst = max(worst, min(ms[1:]))
            allparts.setdefault(parts, []).append(round(min(ms[1:]), 3))
        row.append(worst)
    rgb, _ = eng.render_to_host(W, H, want_hit=False)
    if ref_frame is None: ref_frame = rgb
    print("split4>=%2d split32>=%2d  kernel ms (slowest part): N=1 %.3f  N=2 %.3f  N=4 %.3f  N=8 %.3f  same_pixels=%s" %
          (v, v32, row[0], row[1], row[2], row[3], bool(np.array_equal(rgb, ref_frame))), flush=True)
    print('  band', BAND, 'per-part kernel ms at N=8:', allparts[8], ' N=4:', allparts[4])
