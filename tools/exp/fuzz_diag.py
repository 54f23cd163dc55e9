"""Diagnose one seed of tests/test_fuzz_gpu.py: which pixels differ from the oracle, per trace variant."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as entry
pkg = entry.load_package()
from example_gui_opencl_raytracer_b200 import host
from oracle import pyref as ref
import test_fuzz_gpu as t
seed = int(sys.argv[1])
rng = np.random.default_rng(1000 + seed)
s, p, l = t.random_scene(host, rng)
W, H = 96, 64
origin = tuple(rng.normal(size=3) * 2 + np.array([0, 1.5, -4]))
if seed % 5 == 0 and len(s):
    origin = tuple(np.frombuffer(s[:12], np.float32).astype(float))
cam = host.camera(origin=origin, look=tuple(rng.normal(size=3) * 0.3 + np.array([0, -0.1, 1])),
                  fov=float(rng.choice([60.0, 90.0, 120.0])), focal=1.0, W=W, H=H)
tex, sky = host.default_images(ROOT)
sc = ref.Scene(s, p, l, tex, sky)
want_ref, hit_ref = ref.ref().render(sc, cam, W, H) if ref.have_ref() else (None, None)
want_port, hit_port = ref.port().render(sc, cam, W, H)
if want_ref is not None:
    print("ref vs port differing words:", int((np.asarray(want_ref) != np.asarray(want_port)).sum()))
want = np.asarray(want_ref if want_ref is not None else want_port).reshape(H, W)
print("S P L", len(s) // 96, len(p) // 96, len(l) // 48)
eng = pkg.Engine(0)
eng.upload_images(0, tex); eng.upload_images(1, sky); eng.upload_scene(s, p, l); eng.set_camera(cam)
for trace in (0, 1, 2, 3):
    eng.set_option("trace", trace)
    rgb, hit = eng.render_to_host(W, H, want_hit=True)
    rgb = np.asarray(rgb).reshape(H, W)
    bad = np.argwhere(rgb != want)
    print("trace", trace, "differing", len(bad), "hit equal", bool((np.asarray(hit).reshape(-1) == np.asarray(hit_ref if hit_ref is not None else hit_port).reshape(-1)).all()))
    for (y, x) in bad[:12]:
        print("   y=%d x=%d got %08x want %08x hit %d" % (y, x, rgb[y, x], want[y, x], np.asarray(hit).reshape(H, W)[y, x]))
    if len(bad):
        ys, xs = bad[:, 0], bad[:, 1]
        print("   rows %d..%d cols %d..%d; hit ids of bad:" % (ys.min(), ys.max(), xs.min(), xs.max()),
              np.unique(np.asarray(hit).reshape(H, W)[ys, xs]))
