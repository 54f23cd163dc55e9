"""Golden for the FULL-SIZE parity test: per-row CRC32 of the reference's 3840x2160 render.map frame (colour words and
hit ids) and its work counters, produced here from oracle/_ref (the reference's .cl text) and the counted port.
    python tools/make_golden_4k.py   ->  tests/golden/render_map_4k_rows.npz  (~35 KB)"""
import os, sys, zlib
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g
ref = g.load_oracle()
W, H = 3840, 2160
scene = ref.render_map_scene(ROOT)
cam = ref.camera(W=W, H=H, **ref.RAYPNG_CAMERA)
assert ref.have_ref(), "build oracle/_ref first (needs /root/reference)"
rgb, hit = ref.ref().render(scene, cam, W, H)
rgb2, hit2, cnt = ref.port().render_counted(scene, cam, W, H)
assert np.array_equal(rgb, rgb2) and np.array_equal(hit, hit2), "port and reference text disagree at 4K"
rows_rgb = np.array([zlib.crc32(rgb[y * W:(y + 1) * W].tobytes()) for y in range(H)], np.uint32)
rows_hit = np.array([zlib.crc32(hit[y * W:(y + 1) * W].tobytes()) for y in range(H)], np.uint32)
names = ("segments", "shadow_rays", "sphere_tests", "sphere_dpos", "plane_tests", "texels")
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "render_map_4k_rows.npz"), rows_rgb=rows_rgb, rows_hit=rows_hit,
                    counters=np.array([cnt[k] for k in names], np.uint64), counter_names=np.array(names))
print("ok", {k: cnt[k] for k in names})
