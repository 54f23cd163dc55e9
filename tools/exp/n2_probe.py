"""Why is a history frame with the pixel-tree kernel slower inside bench.py's N=2 loop (1.56 ms) than alone (1.36 ms)?
One GPU standing in for rank 0 of 2: kernel ms with / without the frame mirror and the completion word, tree on / off."""
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
frame = eng.dev_alloc(W * H * 4 + 8192)
flag = frame.value + W * H * 4 + 256
for parts in (2, 4, 8):
    part = pkg.Partition(8, parts, 0)
    for tree in (0, 1):
        eng.set_option("tree", tree)
        for mirror, done in ((0, 0), (1, 0), (0, 1), (1, 1)):
            ms = []
            for i in range(8):
                if mirror: eng.set_mirror_frame(frame.value)
                if done: eng.set_completion(flag, i + 1)
                eng.render(W, H, part=part)
                ms.append(eng.last_kernel_ms())
            print("N=%d tree=%d mirror=%d completion=%d: %s  launch %s" % (parts, tree, mirror, done, " ".join("%.3f" % v for v in ms), eng.last_launch()["tree_grid"]), flush=True)
