"""Randomised parity: seeded random scenes (tie-prone duplicates, degenerate radii, non-unit plane normals, odd
materials, cameras inside objects) rendered by the CUDA path and the oracle.  Same tolerance as everywhere; the
point is the rare branches: total internal reflection, zero denominators, NaN directions, equal-t ties."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def host(pkg):
    from example_gui_opencl_raytracer_b200 import host as h
    return h


def random_scene(host, rng):
    def mat():
        base = [host.STONE, host.PLASTIC, host.MIRROR, host.GLASS][rng.integers(4)]
        m = dict(base, rgb=tuple(rng.random(3)))
        if rng.random() < 0.3: m["reflectivity"] = float(rng.choice([0.0, 0.5, 1.0]))
        if rng.random() < 0.2: m["n"] = float(rng.choice([1.0, 1.33, 2.4, 0.7]))
        if rng.random() < 0.2: m["shininess"] = int(rng.choice([0, 1, 7, 300]))
        if rng.random() < 0.2: m["transperent"] = int(rng.integers(2)); 
        if rng.random() < 0.2: m["dielectric"] = int(rng.integers(2))
        if rng.random() < 0.15: m["diffuse"] = 0.0
        if rng.random() < 0.15: m["specular"] = 0.0
        return m
    S, P, L = int(rng.integers(0, 9)), int(rng.integers(0, 4)), int(rng.integers(0, 5))
    sp = []
    for i in range(S):
        c = rng.normal(size=3) * 2.0 + np.array([0, 1, 3])
        r = float(rng.choice([0.0, 0.3, 0.8, 1.5])) if rng.random() < 0.3 else float(rng.random() * 1.2 + 0.1)
        sp.append(host.sphere(tuple(c), r, mat()))
        if rng.random() < 0.2:                       # exact duplicate: first of equal t must win
            sp.append(host.sphere(tuple(c), r, mat()))
    pl = []
    for j in range(P):
        n = rng.normal(size=3)
        if rng.random() < 0.7: n = n / np.linalg.norm(n)
        if rng.random() < 0.3: n = np.array([[0, 1, 0], [0, 0, -1], [1, 0, 0]][rng.integers(3)], float)
        m = mat()
        m["texture_id"] = int(rng.choice([-1, 0, 1, 2, 3, 7]))
        m["texture_scale"] = float(rng.choice([100.0, 1.0, -13.0, 1e9, 0.0]))
        pl.append(host.plane(tuple(n), tuple(rng.normal(size=3) * 2 + np.array([0, -1, 4])), m))
        if rng.random() < 0.15:
            pl.append(pl[-1])
    li = [host.light(tuple(rng.normal(size=3) * 3 + np.array([0, 3, 2])), float(rng.choice([0.1, 0.5, 0.0])),
                     float(rng.random() * 60), tuple(rng.random(3))) for _ in range(L)]
    return b"".join(sp), b"".join(pl), b"".join(li)


@pytest.mark.parametrize("seed", range(48))
def test_random_scene(pkg, ref, scene, host, seed):
    rng = np.random.default_rng(1000 + seed)
    s, p, l = random_scene(host, rng)
    W, H = 96, 64
    origin = tuple(rng.normal(size=3) * 2 + np.array([0, 1.5, -4]))
    if seed % 5 == 0 and len(s):                       # camera at a sphere centre
        origin = tuple(np.frombuffer(s[:12], np.float32).astype(float))
    cam = host.camera(origin=origin, look=tuple(rng.normal(size=3) * 0.3 + np.array([0, -0.1, 1])),
                      fov=float(rng.choice([60.0, 90.0, 120.0])), focal=1.0, W=W, H=H)
    sc = ref.Scene(s, p, l, scene.tex, scene.sky)
    want_rgb, want_hit = ref.judge().render(sc, cam, W, H)
    eng = pkg.Engine(0)
    try:
        eng.upload_images(0, scene.tex); eng.upload_images(1, scene.sky)
        eng.upload_scene(s, p, l)
        eng.set_camera(cam)
        for trace in (0, 1, 2):
            eng.set_option("trace", trace)
            rgb, hit = eng.render_to_host(W, H, want_hit=True)
            c = ref.compare_frames(rgb, want_rgb, W, H)
            same = float((hit == want_hit).mean())
            assert same == 1.0, (seed, trace, same)
            assert c["within1"] >= 0.999 and c["psnr"] >= 50.0, (seed, trace, c)
            if trace == 0:
                # the WIDE variant (one ray over 32 lanes), forced onto these small scenes: every round, then a mix
                for max_live in (32, 3):
                    eng.set_option("wide_min_records", 0); eng.set_option("wide_max_live", max_live)
                    rgb_w, hit_w = eng.render_to_host(W, H, want_hit=True)
                    eng.set_option("wide_min_records", 64); eng.set_option("wide_max_live", -1)
                    assert np.array_equal(rgb_w, rgb) and np.array_equal(hit_w, hit), (seed, "wide", max_live)
                # the level-wise pixel-tree trace: every tile as one pooled tree (2), every pixel on its own (3) -- what
                # the split queue entries of heavy tiles run; must be the stack machine's words
                for tree in (2, 3):
                    eng.set_option("tree", tree)
                    rgb_t, hit_t = eng.render_to_host(W, H, want_hit=True)
                    assert np.array_equal(rgb_t, rgb) and np.array_equal(hit_t, hit), (seed, "tree", tree)
                eng.set_option("tree", 1)
    finally:
        eng.close()
