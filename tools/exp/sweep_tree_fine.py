"""Fine sweep of tree_min_cost / split_min_cost per part count: median and spread of 10 history frames of part 0 and of
the slowest part (one GPU standing in for each rank).  python tools/exp/sweep_tree_fine.py"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry
import bench
pkg = entry.load_package()
from example_gui_opencl_raytracer_b200 import host
wl = bench.make_workload("render_map_4k", host)
W, H = wl["W"], wl["H"]
tex, sky = host.default_images(ROOT)
eng = pkg.Engine(0)
eng.upload_scene(wl["spheres"], wl["planes"], wl["lights"]); eng.upload_images(0, tex); eng.upload_images(1, sky)
eng.set_camera(wl["cam"])
for parts in (2, 4, 8):
    for tmin, s4 in ((0, 0), (34, 64), (36, 64), (37, 64), (38, 64), (39, 64), (40, 64), (42, 64), (38, 46), (38, 48), (40, 48), (36, 46)):
        eng.set_option("tree_min_cost", tmin); eng.set_option("split_min_cost", s4)
        worst_med, worst_max = 0.0, 0.0
        for part in range(parts):
            ms = []
            for i in range(12):
                eng.render(W, H, part=pkg.Partition(8, parts, part))
                ms.append(eng.last_kernel_ms())
            worst_med = max(worst_med, float(np.median(ms[2:])))
            worst_max = max(worst_max, max(ms[2:]))
        print("N=%d tmin=%2d split4=%2d: slowest part median %.3f, worst frame %.3f" % (parts, tmin, s4, worst_med, worst_max), flush=True)
