"""Sweep of the coop_min_cost threshold (and trace variants): kernel ms at 4K."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry
pkg = entry.load_package()
from example_gui_opencl_raytracer_b200 import host
W, H = 3840, 2160
tex, sky = host.default_images(ROOT)
def run(name, s, p, l, camp, reps, values):
    eng = pkg.Engine(0)
    eng.upload_scene(s, p, l); eng.upload_images(0, tex); eng.upload_images(1, sky)
    eng.set_camera(host.camera(W=W, H=H, **camp))
    for trace, v in values:
        eng.set_option("trace", trace); eng.set_option("coop_min_cost", v)
        ms = []
        for i in range(reps + 1):
            eng.render(W, H); ms.append(eng.last_kernel_ms())
        print("%s trace=%d coop_min_cost=%2d kernel ms min %.3f mean %.3f" % (name, trace, v, min(ms[1:]), np.mean(ms[1:])), flush=True)
    eng.close()
s, p, l = host.load_map(os.path.join(ROOT, "scenes", "render.map"))
run("render.map", s, p, l, host.RAYPNG_CAMERA, 8, [(2, 0)] + [(0, v) for v in (0, 4, 8, 12, 16, 20, 24, 32, 40, 48, 64)])
if len(sys.argv) > 1:
    s, p, l, camp = host.sphere_grid_scene(32)
    run("grid1024", s, p, l, camp, 1, [(2, 0), (0, 0), (0, 16), (0, 32)])
