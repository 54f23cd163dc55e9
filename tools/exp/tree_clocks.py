"""Where a pixel-tree entry's time goes (diagnostic build: make variant NAME=clocks DEFS=-DRT_TREE_CLOCKS):
SM clocks per phase summed over the entries of one N=8 part, with entry / node / level counts.
    RTB_LIB=.../librtb200_clocks.so python tools/exp/tree_clocks.py"""
import os, sys
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
for parts in (8, 1):
    for tmin, s4 in ((0, 0), (36, 64), (32, 44)):
        eng.set_option("tree_min_cost", tmin); eng.set_option("split_min_cost", s4)
        part = pkg.Partition(8, parts, 1) if parts > 1 else None
        for i in range(4):
            eng.render(W, H, part=part)
            ms = eng.last_kernel_ms()
        c = eng.debug_counters()
        n = max(c[5], 1)
        print("N=%d tmin=%d split4=%d: kernel %.3f ms; tree entries %d, nodes/entry %.0f, levels/entry %.1f; clocks/entry: expand %.0f rank-back %.0f "
              "rank-fwd %.0f shade %.0f replay %.0f  total %.0f (%.1f us)" % (parts, tmin, s4, ms, c[5], c[6] / n, c[7] / n, c[0] / n, c[1] / n, c[2] / n, c[3] / n, c[4] / n,
                                                                               sum(c[:5]) / n, sum(c[:5]) / n / 1965.0))
