"""Small renders for compute-sanitizer (memcheck / racecheck): history frames, forced 32-way and 4-way splits, AOV, triangles."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry
pkg = entry.load_package()
from example_gui_opencl_raytracer_b200 import host
W, H = int(sys.argv[1]), int(sys.argv[2])
s, p, l = host.load_map(os.path.join(ROOT, "scenes", "render.map"))
tex, sky = host.default_images(ROOT)
eng = pkg.Engine(0)
eng.upload_scene(s, p, l); eng.upload_images(0, tex); eng.upload_images(1, sky)
eng.set_camera(host.camera(W=W, H=H, **host.RAYPNG_CAMERA))
a, ha = eng.render_to_host(W, H, want_hit=True)
for s4, s32 in ((0, 0), (1, 64), (1, 1)):
    eng.set_option("split_min_cost", s4); eng.set_option("split32_min_cost", s32)
    for _ in range(2):
        b, hb = eng.render_to_host(W, H, want_hit=True)
        assert np.array_equal(a, b) and np.array_equal(ha, hb)
_, _, _, tris, _ = host.heightfield_mesh_scene(n=6)
eng.upload_triangles(tris)
for _ in range(2):
    eng.render_to_host(W, H, want_hit=True)
eng.render_counted(W, H)
# WIDE variant (>= 64 light+sphere records, no triangles): a fresh engine on a 9x9 sphere grid
gs, gp, gl, cam = host.sphere_grid_scene(9)
eng2 = pkg.Engine(0)
eng2.upload_scene(gs, gp, gl); eng2.upload_images(0, tex); eng2.upload_images(1, sky)
eng2.set_camera(host.camera(W=W, H=H, **cam))
w0, hw0 = eng2.render_to_host(W, H, want_hit=True)
eng2.set_option("wide_min_records", -1)
w1, hw1 = eng2.render_to_host(W, H, want_hit=True)
assert np.array_equal(w0, w1) and np.array_equal(hw0, hw1)
print("ok")
