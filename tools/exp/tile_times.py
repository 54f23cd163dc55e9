"""Diagnostic: per-tile start time and duration of render_frame (library built with `make timing`, RTB_LIB=...)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry
pkg = entry.load_package()
from example_gui_opencl_raytracer_b200 import host
W, H = 3840, 2160
parts = int(sys.argv[1]) if len(sys.argv) > 1 else 1
split = int(sys.argv[2]) if len(sys.argv) > 2 else 64
s, p, l = host.load_map(os.path.join(ROOT, "scenes", "render.map")); camp = host.RAYPNG_CAMERA
tex, sky = host.default_images(ROOT)
eng = pkg.Engine(0)
eng.upload_scene(s, p, l); eng.upload_images(0, tex); eng.upload_images(1, sky)
eng.set_camera(host.camera(W=W, H=H, **camp))
eng.set_option("split_min_cost", split)
part = pkg.Partition(8, parts, 0) if parts > 1 else None
rows = eng.partition_rows(H, part)
for _ in range(3):
    rgb, t = eng.render_to_host(W, H, part=part, want_hit=True)
print("parts", parts, "split", split, "kernel ms", eng.last_kernel_ms())
t = t.reshape(rows // 4, 4, W // 8, 8)
start = ((t >> 16) & 0xffff).astype(np.int64)
dur = (t & 0xffff).astype(np.float64) * 1.024
start = (start - start.min()) % 65536 * 1.024          # us since the first tile began
per_row = start[:, :, :, 0], dur[:, :, :, 0]            # one sample per tile row (sub-tiles differ per row)
st, du = per_row[0].reshape(-1), per_row[1].reshape(-1)
end = st + du
print("last end %.1f us; throughput bound (sum dur / 2960 / 4 rows) %.1f us" % (end.max(), du.sum() / 4 / 2960))
order = np.argsort(end)[-8:]
for i in order:
    print("  tile-row idx %7d  start %8.1f  dur %8.1f  end %8.1f" % (i, st[i], du[i], end[i]))
late = st > 0.5 * end.max()
print("tiles starting in the second half: %d, of which dur > 100 us: %d" % (late.sum(), (late & (du > 100)).sum()))
for q in (50, 90, 99, 99.9, 100):
    print("dur p%-5s %.1f us" % (q, np.percentile(du, q)))
