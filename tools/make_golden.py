#!/usr/bin/env python
"""Generate tests/golden/*.npz from the reference's own .cl text (oracle/_ref/libref_cl.so).

Run in the build container (needs /root/reference to have built oracle/_ref).  The fixtures are what
pins the plain-C oracle (oracle/rt_oracle.c) and, through it, the CUDA path on machines where the
reference source is absent.  Inputs are seeded; edge cases the reference is sensitive to are added by
hand (cube-face ties, zero components, grazing / inside-sphere rays, parallel planes, TIR, INT_MIN texel
coordinates).
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import pyref  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
f32p = C.POINTER(C.c_float)


def fp(a):
    return a.ctypes.data_as(f32p)


def kat_inputs(rng):
    n = 256
    d = {}
    dirs = rng.normal(size=(n, 3)).astype(np.float32)
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True).astype(np.float32)
    ties = np.array([[1, 1, 0], [1, -1, 0], [0, 1, 1], [1, 0, 1], [-1, 0, -1], [1, 1, 1], [-1, -1, -1], [0, 0, 1],
                     [0, 0, -1], [0, 1, 0], [0, -1, 0], [1, 0, 0], [-1, 0, 0], [0.5, 0.5, -0.5]], np.float32)
    d["dirs"] = np.concatenate([dirs, ties]).astype(np.float32)
    d["origins"] = (rng.uniform(-6, 6, size=(d["dirs"].shape[0], 3))).astype(np.float32)
    d["centres"] = (rng.uniform(-4, 4, size=(d["dirs"].shape[0], 3))).astype(np.float32)
    d["radii"] = rng.uniform(0.2, 3.0, size=d["dirs"].shape[0]).astype(np.float32)
    d["origins"][:8] = d["centres"][:8]                      # ray starts at the centre (inside)
    d["origins"][8:16] = d["centres"][8:16] + d["dirs"][8:16] * d["radii"][8:16, None]  # on the surface
    nrm = rng.normal(size=(n + 14, 3)).astype(np.float32)
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True).astype(np.float32)
    nrm[:6] = np.array([[0, 1, 0], [0, 0, -1], [1, 0, 0], [0, -1, 0], [0, 1, 0], [1, 1, 1]], np.float32)
    d["normals"] = nrm
    d["points"] = rng.uniform(-5, 5, size=(n + 14, 3)).astype(np.float32)
    d["n1"] = rng.choice(np.array([1.0, 1.52, 1.57, 1.33, 1.0, 2.4], np.float32), size=n + 14)
    d["n2"] = rng.choice(np.array([1.0, 1.52, 1.57, 1.33, 1.52, 1.0], np.float32), size=n + 14)
    d["emod_a"] = np.concatenate([rng.integers(-2 ** 31, 2 ** 31 - 1, size=200), [-(2 ** 31), -1, 0, 1, 255, 256, -256, -257]]).astype(np.int32)
    d["emod_b"] = np.concatenate([rng.choice([256, 4096, 3072, 7, 1], size=200), [256] * 8]).astype(np.int32)
    d["seeds"] = np.array([0, 1, 479999, 8294399, 0xDEADBEEF, 12345], np.uint32)
    d["texel_pts"] = np.concatenate([rng.uniform(-50, 50, size=(120, 3)), [[1e12, 0, 0], [-1e12, 0, 3e12], [0, 0, 0], [-0.001, 0, -0.001]]]).astype(np.float32)
    return d


def main():
    os.makedirs(OUT, exist_ok=True)
    lib = pyref.ref().lib
    lib.ref_xorshift32.restype = C.c_float
    lib.ref_schlick.restype = C.c_float
    lib.ref_shadow.restype = C.c_float
    lib.ref_schlick.argtypes = [C.c_float, C.c_float, f32p, f32p]
    lib.ref_refract.argtypes = [C.c_float, C.c_float, f32p, f32p, f32p]
    lib.ref_isect_sphere.argtypes = [f32p, f32p, f32p, C.c_float, f32p]
    scene = pyref.render_map_scene(ROOT)
    rng = np.random.default_rng(20261104)
    k = kat_inputs(rng)
    out = dict(k)
    n = k["dirs"].shape[0]

    # xorshift32: 8 draws per seed
    xs = np.zeros((len(k["seeds"]), 8), np.float32)
    xstate = np.zeros((len(k["seeds"]), 8), np.uint32)
    for i, s in enumerate(k["seeds"]):
        st = C.c_uint32(int(s))
        for j in range(8):
            xs[i, j] = lib.ref_xorshift32(C.byref(st))
            xstate[i, j] = st.value
    out["xorshift_vals"], out["xorshift_states"] = xs, xstate

    uv = np.zeros((n, 2), np.int32)
    refl = np.zeros((n, 3), np.float32)
    refr = np.zeros((n, 3), np.float32)
    sch = np.zeros(n, np.float32)
    sph_hit = np.zeros(n, np.int32)
    sph_t = np.zeros(n, np.float32)
    pl_hit = np.zeros(n, np.int32)
    pl_t = np.zeros(n, np.float32)
    for i in range(n):
        d_, o_, c_, nn, pp = (np.ascontiguousarray(k[x][i]) for x in ("dirs", "origins", "centres", "normals", "points"))
        lib.ref_map_to_cube(fp(d_), 1024, uv[i].ctypes.data_as(C.POINTER(C.c_int)))
        lib.ref_reflect(fp(d_), fp(nn), fp(refl[i]))
        lib.ref_refract(float(k["n1"][i]), float(k["n2"][i]), fp(d_), fp(nn), fp(refr[i]))
        sch[i] = lib.ref_schlick(float(k["n1"][i]), float(k["n2"][i]), fp(d_), fp(nn))
        t = C.c_float(-1.0)
        sph_hit[i] = lib.ref_isect_sphere(fp(o_), fp(d_), fp(c_), float(k["radii"][i]), C.byref(t))
        sph_t[i] = t.value if sph_hit[i] else -1.0
        t = C.c_float(-1.0)
        pl_hit[i] = lib.ref_isect_plane(fp(o_), fp(d_), fp(nn), fp(pp), C.byref(t))
        pl_t[i] = t.value if pl_hit[i] else -1.0
    out.update(cube_uv=uv, reflect=refl, refract=refr, schlick=sch, sph_hit=sph_hit, sph_t=sph_t, pl_hit=pl_hit, pl_t=pl_t)
    out["emod"] = np.array([lib.ref_emod(int(a), int(b)) for a, b in zip(k["emod_a"], k["emod_b"])], np.int32)

    # scene-level functions on render.map: rays from the raypng camera at 64x48 plus random rays
    W, H = 64, 48
    cam = pyref.camera(W=W, H=H, **pyref.RAYPNG_CAMERA)
    ro = np.zeros((W * H, 3), np.float32)
    rd = np.zeros((W * H, 3), np.float32)
    for i in range(W * H):
        lib.ref_raygen(fp(cam), W, H, i, fp(ro[i]), fp(rd[i]))
    out["cam_64x48"], out["raygen_o"], out["raygen_d"] = cam, ro, rd
    sp, pl, li = scene.spheres, scene.planes, scene.lights
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    tex = scene.tex
    fl_hit = np.zeros(W * H, np.int32); fl_col = np.zeros((W * H, 3), np.float32)
    fs_hit = np.zeros(W * H, np.int32); fs_ip = np.zeros((W * H, 3), np.float32); fs_n = np.zeros((W * H, 3), np.float32)
    fs_m = np.zeros((W * H, 64), np.uint8)
    shadow = np.zeros(W * H, np.float32)
    light0 = np.frombuffer(li.tobytes()[48:60], np.float32).copy()  # light 1 origin
    for i in range(W * H):
        fl_hit[i] = lib.ref_find_light(fp(ro[i]), fp(rd[i]), vp(sp), scene.S, vp(pl), scene.P, vp(li), scene.L, fp(fl_col[i]))
        fs_hit[i] = lib.ref_find_solid(fp(ro[i]), fp(rd[i]), vp(sp), scene.S, vp(pl), scene.P, vp(tex), 256, 256, 4,
                                       fp(fs_ip[i]), fp(fs_n[i]), vp(fs_m[i]))
        if fs_hit[i]:
            shadow[i] = lib.ref_shadow(fp(light0), fp(fs_ip[i]), vp(sp), scene.S, vp(pl), scene.P)
    out.update(fl_hit=fl_hit, fl_col=fl_col, fs_hit=fs_hit, fs_ip=fs_ip, fs_n=fs_n, fs_m=fs_m, shadow=shadow)
    tx = np.zeros((len(k["texel_pts"]), 3), np.float32)
    plane0 = np.ascontiguousarray(np.frombuffer(pl.tobytes()[:96], np.uint8))
    for i, p in enumerate(k["texel_pts"]):
        lib.ref_plane_texel(vp(plane0), fp(np.ascontiguousarray(p)), vp(tex), 256, 256, 4, fp(tx[i]))
    out["plane_texel"] = tx
    np.savez_compressed(os.path.join(OUT, "kat.npz"), **out)

    # full frames: render.map at 160x120 (raypng camera) and at 96x54 (16:9), colour + hit id
    frames = {}
    for (w, h) in ((160, 120), (96, 54)):
        c = pyref.camera(W=w, H=h, **pyref.RAYPNG_CAMERA)
        rgb, hit = pyref.ref().render(scene, c, w, h)
        frames["cam_%dx%d" % (w, h)] = c
        frames["rgb_%dx%d" % (w, h)] = rgb
        frames["hit_%dx%d" % (w, h)] = hit
    np.savez_compressed(os.path.join(OUT, "render_map_frames.npz"), **frames)
    print("wrote", os.listdir(OUT))


if __name__ == "__main__":
    main()
