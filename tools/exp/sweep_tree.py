"""Thresholds of the CTA-wide pixel-tree phase: slowest part's render_frame ms at N = 1, 4, 8 (one GPU standing in for
each rank of a cyclic 8-row band split), per (tree, tree_min_cost, split4, split32) on the relative cost scale
(24 = mean tile, +4 per octave; 0 = the engine's automatic value).  Six frames per part: the sequence shows whether the
measured-time schedule is stable.
    python tools/exp/sweep_tree.py [workload [combo,combo,...]]   combo = tree:tmin:s4:s32[:group_max_cost]"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry
import bench
pkg = entry.load_package()
from example_gui_opencl_raytracer_b200 import host
wl = bench.make_workload(sys.argv[1] if len(sys.argv) > 1 else "render_map_4k", host)
W, H = wl["W"], wl["H"]
tex, sky = host.default_images(ROOT)
eng = pkg.Engine(0)
eng.upload_scene(wl["spheres"], wl["planes"], wl["lights"]); eng.upload_images(0, tex); eng.upload_images(1, sky)
eng.set_camera(wl["cam"])
for kv in filter(None, os.environ.get("SWEEP_OPTS", "").split(",")):   # e.g. SWEEP_OPTS=accel=1
    k, v = kv.split("=")
    eng.set_option(k, int(v))
if len(sys.argv) > 2:
    combos = [tuple(int(v) for v in c.split(":")) for c in sys.argv[2].split(",")]
else:
    combos = [(0, 0, 0, 0), (1, 0, 0, 0)] + [(1, t, s4, s32) for t, s4, s32 in
              ((28, 64, 64), (32, 64, 64), (36, 64, 64), (40, 64, 64), (44, 64, 64), (32, 40, 64), (32, 44, 64), (36, 44, 64), (36, 44, 52), (32, 40, 48), (28, 36, 64), (28, 40, 48))]
ref_frame = None
for combo in combos:
    tree, tmin, s4, s32 = combo[:4]
    gmax = combo[4] if len(combo) > 4 else 0
    eng.set_option("tree_ctas_per_sm", combo[5] if len(combo) > 5 else 3)
    eng.set_option("tree", tree); eng.set_option("tree_min_cost", tmin); eng.set_option("tree_group_max_cost", gmax)
    eng.set_option("split_min_cost", s4); eng.set_option("split32_min_cost", s32)
    out = []
    for parts in (1, 4, 8):
        worst, seq = 0.0, None
        for part in range(parts):
            ms = []
            for i in range(6):
                eng.render(W, H, part=pkg.Partition(8, parts, part) if parts > 1 else None)
                ms.append(eng.last_kernel_ms())
            m = float(np.median(ms[2:]))
            if m > worst: worst, seq = m, ms
        out.append("N=%d %.3f [%s]" % (parts, worst, " ".join("%.2f" % v for v in seq)))
    rgb, _ = eng.render_to_host(W, H, want_hit=False)
    if ref_frame is None: ref_frame = rgb
    print("tree=%d tmin=%2d split4>=%2d split32>=%2d gmax=%2d ctas/sm=%d same=%s :: %s" % (tree, tmin, s4, s32, gmax, combo[5] if len(combo) > 5 else 3, bool(np.array_equal(rgb, ref_frame)), " | ".join(out)), flush=True)
