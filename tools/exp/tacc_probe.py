"""N4 over triangles: mesh50k at 3840x2160 through the hierarchy vs brute force (words must be equal); ms per frame.
    python tools/exp/tacc_probe.py [W H]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402

pkg = g.load_package()
from example_gui_opencl_raytracer_b200 import host  # noqa: E402

W, H = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (3840, 2160)
tex, sky = host.default_images(ROOT)
eng = pkg.Engine(0)
eng.upload_images(0, tex)
eng.upload_images(1, sky)
s, p, l, tris, camp = host.heightfield_mesh_scene()
eng.upload_scene(s, p, l)
eng.upload_triangles(tris)
eng.set_camera(host.camera(W=W, H=H, **camp))
for kv in filter(None, os.environ.get("ENGINE_OPTS", "").split(",")):
    k, v = kv.split("=")
    eng.set_option(k, int(v))
brute = eng.render_to_host(W, H, want_hit=True)
print("brute force: %.1f ms" % eng.last_render_ms(), flush=True)
eng.set_option("accel", 1)
t0 = time.time()
acc = eng.render_to_host(W, H, want_hit=True)
print("first accel frame incl. host build: %.1f ms wall, render %.2f ms" % ((time.time() - t0) * 1e3, eng.last_render_ms()))
for i in range(6):
    acc = eng.render_to_host(W, H, want_hit=True)
    print("accel frame %d: render %.3f ms kernel %.3f ms launch %s" % (i + 1, eng.last_render_ms(), eng.last_kernel_ms(), eng.last_launch()))
print("equal words:", np.array_equal(acc[0], brute[0]), "equal hit ids:", np.array_equal(acc[1], brute[1]),
      "differing words:", int((acc[0] != brute[0]).sum()), "hit ids:", int((acc[1] != brute[1]).sum()))
eng.close()
