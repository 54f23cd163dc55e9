"""A/B of the batched single-pixel walk (rt_trace_solo.cuh, option "solo"): kernel ms at 4K on render.map for the
whole frame and for the slowest rank's share at N = 2, 4, 8 (partitions rendered one after the other on one GPU);
then the 15x15 sphere grid (narrow variant forced)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry
pkg = entry.load_package()
from example_gui_opencl_raytracer_b200 import host
W, H = 3840, 2160
tex, sky = host.default_images(ROOT)


def sweep(eng, label, parts_list=(1, 2, 4, 8), reps=5):
    frames = {}
    for solo in (0, 1, 0, 1):
        eng.set_option("solo", solo)
        row = []
        for parts in parts_list:
            worst = 0.0
            for part in range(parts):
                ms = []
                for i in range(reps):
                    eng.render(W, H, part=pkg.Partition(8, parts, part) if parts > 1 else None)
                    ms.append(eng.last_kernel_ms())
                worst = max(worst, min(ms[2:]))
            row.append(worst)
        rgb, _ = eng.render_to_host(W, H, want_hit=False)
        frames[solo] = rgb
        print("%s solo=%d kernel ms (slowest part): %s" % (label, solo, "  ".join("N=%d %.3f" % (n, t) for n, t in zip(parts_list, row))), flush=True)
    print("%s same_pixels=%s" % (label, bool(np.array_equal(frames[0], frames[1]))), flush=True)


eng = pkg.Engine(0)
eng.upload_images(0, tex); eng.upload_images(1, sky)
s, p, l = host.load_map(os.path.join(ROOT, "scenes", "render.map"))
eng.upload_scene(s, p, l)
eng.set_camera(host.camera(W=W, H=H, **host.RAYPNG_CAMERA))
sweep(eng, "render.map 4K")
if len(sys.argv) > 1:
    s, p, l, camp = host.sphere_grid_scene(int(sys.argv[1]))
    eng.upload_scene(s, p, l)
    eng.set_camera(host.camera(W=W, H=H, **camp))
    eng.set_option("wide_min_records", -1)
    sweep(eng, "grid %sx%s 4K (narrow)" % (sys.argv[1], sys.argv[1]), parts_list=(1, 8), reps=4)
