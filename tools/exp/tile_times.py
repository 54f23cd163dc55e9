"""Diagnostic: per-tile wall time of render_frame (library built with -DRT_TILE_TIMING, RTB_LIB=...)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry
pkg = entry.load_package()
from example_gui_opencl_raytracer_b200 import host
W, H = 3840, 2160
wlname = sys.argv[1] if len(sys.argv) > 1 else "render"
if wlname == "render":
    s, p, l = host.load_map(os.path.join(ROOT, "scenes", "render.map")); camp = host.RAYPNG_CAMERA
else:
    s, p, l, camp = host.sphere_grid_scene(32)
tex, sky = host.default_images(ROOT)
eng = pkg.Engine(0)
eng.upload_scene(s, p, l); eng.upload_images(0, tex); eng.upload_images(1, sky)
eng.set_camera(host.camera(W=W, H=H, **camp))
for _ in range(3):
    rgb, t = eng.render_to_host(W, H, want_hit=True)
print("kernel ms", eng.last_kernel_ms())
t = t.reshape(H // 4, 4, W // 8, 8)[:, 0, :, 0].astype(np.float64) / 1e3   # us per tile
print("tiles", t.size, "sum of tile times %.1f ms-warp" % (t.sum() / 1e3), "-> / (148*20 warps) = %.3f ms" % (t.sum() / 1e3 / 2960))
for q in (50, 90, 99, 99.9, 99.99, 100):
    print("p%-6s %.1f us" % (q, np.percentile(t, q)))
idx = np.argsort(t.ravel())[-5:]
print("slowest tiles (ty,tx,us):", [(int(i // t.shape[1]), int(i % t.shape[1]), float(t.ravel()[i])) for i in idx])
