"""A/B of a contraction-enabled build (-fmad=true) against the bit-exact default build.

    make variant NAME=fma KFLAGS="-fmad=true -prec-div=true -prec-sqrt=true -ftz=false"
    python tools/exp/contract_ab.py example-gui-opencl-raytracer_b200/librtb200_fma.so

Renders, with each library in its own process: scenes/render.map at 3840x2160, the 1,024-sphere grid at 3840x2160 and the
48 fuzz scenes of tests/test_fuzz_gpu.py at 96x64; then reports for the variant against the default build (which is
word-identical to the reference text on all of these: tests/test_fullsize_gpu.py, test_configs_fullsize_gpu.py,
test_fuzz_gpu.py): kernel ms/frame of both, hit-id mismatches (count + where), % of pixels within 1 LSB, PSNR --
BASELINE.json's tolerance is hit ids exact bar documented grazing ties, <= 1 LSB on >= 99.9 %, PSNR >= 50 dB.
Writes gpurun_out/contract_ab.json."""
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def scenes(host):
    import test_fuzz_gpu as fz
    s, p, l = host.load_map(os.path.join(ROOT, "scenes", "render.map"))
    yield "render_map_4k", s, p, l, host.camera(W=3840, H=2160, **host.RAYPNG_CAMERA), 3840, 2160
    s, p, l, camp = host.sphere_grid_scene(32)
    yield "spheres1024_4k", s, p, l, host.camera(W=3840, H=2160, **camp), 3840, 2160
    for seed in range(48):
        rng = np.random.default_rng(1000 + seed)
        s, p, l = fz.random_scene(host, rng)
        W, H = 96, 64
        origin = tuple(rng.normal(size=3) * 2 + np.array([0, 1.5, -4]))
        if seed % 5 == 0 and len(s):
            origin = tuple(np.frombuffer(s[:12], np.float32).astype(float))
        cam = host.camera(origin=origin, look=tuple(rng.normal(size=3) * 0.3 + np.array([0, -0.1, 1])),
                          fov=float(rng.choice([60.0, 90.0, 120.0])), focal=1.0, W=W, H=H)
        yield "fuzz%02d" % seed, s, p, l, cam, W, H


def child(tag):
    import __graft_entry__ as g
    pkg = g.load_package()
    from example_gui_opencl_raytracer_b200 import host
    tex, sky = host.default_images(ROOT)
    eng = pkg.Engine(0)
    eng.upload_images(0, tex)
    eng.upload_images(1, sky)
    times = {}
    for name, s, p, l, cam, W, H in scenes(host):
        eng.upload_scene(s, p, l)
        eng.set_camera(cam)
        ms = []
        for _ in range(4 if W > 1000 else 1):
            rgb, hit = eng.render_to_host(W, H, want_hit=True)
            ms.append(eng.last_kernel_ms())
        times[name] = min(ms)
        np.save("/tmp/contract_%s_%s_rgb.npy" % (tag, name), rgb)
        np.save("/tmp/contract_%s_%s_hit.npy" % (tag, name), hit)
    json.dump(times, open("/tmp/contract_%s_times.json" % tag, "w"))
    eng.close()


def main():
    if len(sys.argv) > 2 and sys.argv[1] == "--child":
        return child(sys.argv[2])
    variant = os.path.abspath(sys.argv[1])
    for tag, lib in (("base", None), ("var", variant)):
        env = dict(os.environ)
        if lib:
            env["RTB_LIB"] = lib
        subprocess.run([sys.executable, os.path.abspath(__file__), "--child", tag], check=True, env=env)
    import __graft_entry__ as g
    ref = g.load_oracle()
    g.load_package()
    from example_gui_opencl_raytracer_b200 import host
    tb, tv = json.load(open("/tmp/contract_base_times.json")), json.load(open("/tmp/contract_var_times.json"))
    out = {"variant": os.path.basename(variant), "scenes": {}}
    fz = dict(pixels=0, hit_mismatch=0, beyond1=0, sq=0.0, scenes_with_hit_mismatch=[])
    for name, s, p, l, cam, W, H in scenes(host):
        a = np.load("/tmp/contract_base_%s_rgb.npy" % name)
        b = np.load("/tmp/contract_var_%s_rgb.npy" % name)
        ha = np.load("/tmp/contract_base_%s_hit.npy" % name)
        hb = np.load("/tmp/contract_var_%s_hit.npy" % name)
        A = ref.unpack_rgb(a, W, H).astype(np.int32)
        B = ref.unpack_rgb(b, W, H).astype(np.int32)
        d = np.abs(A - B)
        bad = np.flatnonzero(ha != hb)
        if name.startswith("fuzz"):
            fz["pixels"] += W * H
            fz["hit_mismatch"] += int(bad.size)
            fz["beyond1"] += int((d.max(axis=2) > 1).sum())
            fz["sq"] += float((d.astype(np.float64) ** 2).sum())
            if bad.size:
                fz["scenes_with_hit_mismatch"].append(name)
            continue
        c = ref.compare_frames(b, a, W, H)
        out["scenes"][name] = {"kernel_ms_default": tb[name], "kernel_ms_variant": tv[name], "speedup": tb[name] / tv[name],
                               "hit_id_mismatches": int(bad.size),
                               "hit_id_mismatch_pixels": [[int(i % W), int(i // W), int(ha[i]), int(hb[i])] for i in bad[:16]],
                               "differing_pixels": int((d.max(axis=2) > 0).sum()),
                               "within1": c["within1"], "psnr": c["psnr"], "max_abs": c["max_abs"]}
    mse = fz["sq"] / (3.0 * fz["pixels"])
    out["fuzz48"] = {"pixels": fz["pixels"], "hit_id_mismatches": fz["hit_mismatch"],
                     "scenes_with_hit_mismatch": fz["scenes_with_hit_mismatch"],
                     "within1": 1.0 - fz["beyond1"] / fz["pixels"], "psnr": float("inf") if mse == 0 else 10 * np.log10(255.0 ** 2 / mse),
                     "kernel_ms_default_sum": sum(v for k, v in tb.items() if k.startswith("fuzz")),
                     "kernel_ms_variant_sum": sum(v for k, v in tv.items() if k.startswith("fuzz"))}
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "contract_ab.json"), "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
