"""The reference's own drivers (raypng.c, rayinteractive.c; compiled UNCHANGED by drivers/Makefile against
include/cuda_wrap.h) run on the GPU and their output is checked against the oracle."""
import json
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "drivers", "_bin")


def need(name):
    p = os.path.join(BIN, name)
    if not os.path.exists(p):
        pytest.fail("%s missing: run `make -C drivers` where /root/reference exists (build() does)" % p)
    return p


def read_png_words(pkg, path):
    from example_gui_opencl_raytracer_b200 import host
    a = host.load_png_layers([path])[0]
    return (a[..., 0].astype(np.uint32) << 16) | (a[..., 1].astype(np.uint32) << 8) | a[..., 2].astype(np.uint32)


def test_raypng_unchanged_renders_scene_png(pkg, ref, scene, tmp_path):
    """Config 1: raypng at its compiled-in 800x600 -> out/scene.png, compared with the oracle's frame."""
    out = os.path.join(ROOT, "out", "scene.png")
    if os.path.exists(out):
        os.remove(out)
    r = subprocess.run([need("raypng")], cwd=ROOT, capture_output=True, text=True, timeout=120)
    print(r.stdout, r.stderr)
    assert r.returncode == 0 and "Done, took:" in r.stdout
    got = read_png_words(pkg, out).reshape(-1)
    W, H = 800, 600
    cam = ref.camera(W=W, H=H, **ref.RAYPNG_CAMERA)
    want, _ = ref.judge().render(scene, cam, W, H, want_hit=False)
    c = ref.compare_frames(got, want, W, H)
    print(c)
    assert c["within1"] >= 0.999 and c["psnr"] >= 50.0


def test_rayinteractive_unchanged_headless(pkg, ref, scene, tmp_path):
    """The interactive loop against the headless MiniFB stub: 24 frames, RIGHT + D per frame; the last frame
    must equal a direct render with the camera the callback arithmetic arrives at (recomputed here in float32
    through the same libm calls via a 1-frame run of orbit_bench at 800x600)."""
    last = str(tmp_path / "last.png")
    env = dict(os.environ, RTB_FRAMES="24", RTB_KEYS="RIGHT,D", RTB_LAST_FRAME=last)
    r = subprocess.run([need("rayinteractive")], cwd=ROOT, env=env, capture_output=True, text=True, timeout=120)
    print(r.stdout, r.stderr)
    assert r.returncode == 0 and "headless MiniFB: 24 frames" in r.stdout
    a = read_png_words(pkg, last)
    last2 = str(tmp_path / "orbit.png")
    r2 = subprocess.run([need("orbit_bench"), "800", "600", "24", last2], cwd=ROOT, capture_output=True, text=True, timeout=120)
    print(r2.stdout, r2.stderr)
    assert r2.returncode == 0
    line = json.loads(r2.stdout.strip().splitlines()[-1])
    assert line["frames"] == 24 and line["frame_ms_p50"] > 0
    b = read_png_words(pkg, last2)
    assert np.array_equal(a, b)   # same key script, same arithmetic, same backend -> same words


def test_wrapper_error_behaviour(pkg):
    """opencl_wrap.h:9: failures print 'ERROR:\\t<msg>' on stdout and exit(1) -- checked in a child process."""
    code = r'''
import sys, ctypes as C
sys.path.insert(0, %r)
import __graft_entry__ as g
pkg = g.load_package()
L = pkg.wrap_lib()
w = pkg.ClWrap()
L.cl_wrap_init(C.byref(w), C.c_uint64(pkg.CL_DEVICE_TYPE_GPU), C.c_char_p(b"a.cl"), C.c_char_p(b"raygen"), C.c_char_p(b"b.cl"), C.c_char_p(b"raytracer"), C.c_char_p(None))
L.cl_wrap_load_global_data(C.byref(w), 1, 1, None, 96, pkg.CL_MEM_READ_ONLY)
sys.stdout.flush()
L.cl_wrap_load_global_data(C.byref(w), 1, 1, None, 96, pkg.CL_MEM_READ_ONLY)   # slot already in use
print("not reached")
''' % ROOT
    r = subprocess.run(["python", "-c", code], capture_output=True, text=True, timeout=120)
    print(r.stdout, r.stderr)
    assert r.returncode == 1
    assert "ERROR:\t" in r.stdout and "not reached" not in r.stdout


@pytest.mark.parametrize("mode", ["auto", "copy", "mirror"])
def test_readback_mirror_equals_copy(pkg, ref, scene, mode, monkeypatch):
    """from the second use of a host buffer the kernel stores the frame straight into it;
    the words must be those of the plain D2H copy, and a new host buffer must start with a copy again."""
    from example_gui_opencl_raytracer_b200.wrap_session import WrapSession
    monkeypatch.setenv("RTB_HOST_OUTPUT", mode)
    W, H = 640, 360
    cam = ref.camera(W=W, H=H, **ref.RAYPNG_CAMERA)
    want, _ = ref.judge().render(scene, cam, W, H, want_hit=False)
    ses = WrapSession(W, H, scene.spheres, scene.planes, scene.lights, cam)
    bufs = [np.zeros(W * H + 8, np.uint32) for _ in range(2)]
    for b in bufs:
        view = b[(-(b.ctypes.data // 4)) % 4:][:W * H]   # 16-byte aligned window
        assert view.ctypes.data % 16 == 0
        for _ in range(3):
            view[:] = 0xFFFFFFFF
            ses.frame(view.ctypes.data)
            assert np.array_equal(view, want.astype(np.uint32))
    stats = ses.readback_stats()
    print(mode, stats)
    assert stats == {"auto": {"mirrored": 4, "copied": 2}, "copy": {"mirrored": 0, "copied": 6},
                     "mirror": {"mirrored": 6, "copied": 0}}[mode]
    # an unaligned host pointer always takes the copy path
    odd = np.zeros(W * H + 8, np.uint32)
    v = odd[(-(odd.ctypes.data // 4)) % 4 + 1:][:W * H]
    ses.frame(v.ctypes.data)
    ses.frame(v.ctypes.data)
    assert np.array_equal(v, want.astype(np.uint32))
    assert ses.readback_stats()["copied"] == stats["copied"] + 2
    ses.release()
