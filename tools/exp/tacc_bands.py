"""Where does the time of the accel mesh frame go?  Renders row bands one at a time (tree=0) and prints the slowest; then
whole frames for scene variants (no spheres / no mirrors)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402

pkg = g.load_package()
from example_gui_opencl_raytracer_b200 import host  # noqa: E402

W, H = 1920, 1080
tex, sky = host.default_images(ROOT)
eng = pkg.Engine(0)
eng.upload_images(0, tex)
eng.upload_images(1, sky)
s, p, l, tris, camp = host.heightfield_mesh_scene()
eng.set_camera(host.camera(W=W, H=H, **camp))


def frames(tag, n=3):
    out = []
    for _ in range(n):
        eng.render(W, H)
        eng.sync()
        out.append(eng.last_render_ms())
    print(tag, " ".join("%.2f" % v for v in out), flush=True)


for variant in ("full", "no_spheres", "glass_only", "mirror_only"):
    ss = {"full": s, "no_spheres": b"", "glass_only": s[:96], "mirror_only": s[96:]}[variant]
    eng.upload_scene(ss, p, l)
    eng.upload_triangles(tris)
    eng.set_option("accel", 1)
    for tree in (0, 1):
        eng.set_option("tree", tree)
        frames("%s tree=%d:" % (variant, tree))
eng.upload_scene(s, p, l)
eng.upload_triangles(tris)
eng.set_option("tree", 0)
parts = H // 8
times = []
for part in range(parts):
    pp = pkg.Partition(8, parts, part)
    eng.render(W, H, part=pp); eng.sync()
    eng.render(W, H, part=pp); eng.sync()
    times.append(eng.last_render_ms())
times = np.array(times)
top = np.argsort(-times)[:12]
print("bands of 8 rows, slowest:", [(int(b) * 8, round(float(times[b]), 2)) for b in top], "sum %.1f median %.2f" % (times.sum(), np.median(times)))
eng.close()
