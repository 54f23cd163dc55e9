"""the cluster acceleration structure (csrc/rt_trace_accel.cuh, engine option "accel" = 1) must return
the brute-force walk's words -- the reference's tie rules (first of equal t, spheres before planes), the same shadow
values -- on every pixel.  Judges: the brute-force CUDA kernel (bit for bit) and the reference text (oracle/_ref)."""
import os
import zlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def host(pkg):
    from example_gui_opencl_raytracer_b200 import host as h
    return h


@pytest.fixture()
def fresh(pkg, scene):
    e = pkg.Engine(0)
    e.upload_images(0, scene.tex)
    e.upload_images(1, scene.sky)
    yield e
    e.close()


def both(eng, W, H):
    eng.set_option("accel", 0)
    brute = eng.render_to_host(W, H, want_hit=True)
    eng.set_option("accel", 1)
    acc = eng.render_to_host(W, H, want_hit=True)
    eng.set_option("accel", 0)
    return brute, acc


@pytest.mark.parametrize("n_side,W,H", [(8, 320, 180), (15, 256, 144), (17, 192, 108), (32, 256, 144), (81, 96, 54)])
def test_sphere_grids_accel_equals_brute_force_and_reference(fresh, ref, scene, host, n_side, W, H):
    s, p, l, camp = host.sphere_grid_scene(n_side)
    cam = host.camera(W=W, H=H, **camp)
    sc = ref.Scene(s, p, l, scene.tex, scene.sky)
    want_rgb, want_hit = ref.judge().render(sc, cam, W, H, wide=max(sc.S, sc.P, sc.L) > 255)
    fresh.upload_scene(s, p, l)
    fresh.set_camera(cam)
    (b_rgb, b_hit), (a_rgb, a_hit) = both(fresh, W, H)
    assert np.array_equal(a_hit, b_hit) and np.array_equal(a_rgb, b_rgb)
    assert np.array_equal(a_hit, want_hit)
    c = ref.compare_frames(a_rgb, want_rgb, W, H)
    print("n_side=%d: exact %.6f within1 %.6f psnr %.2f" % (n_side, c["exact"], c["within1"], c["psnr"]))
    assert c["within1"] >= 0.999 and c["psnr"] >= 50.0
    # the pixel-tree kernel on top of the cluster walk
    fresh.set_option("accel", 1)
    for tree in (2, 3):
        fresh.set_option("tree", tree)
        rgb, hit = fresh.render_to_host(W, H, want_hit=True)
        assert np.array_equal(rgb, b_rgb) and np.array_equal(hit, b_hit), tree
    fresh.set_option("tree", 1)
    fresh.set_option("accel", 0)


def crowd(host, rng, n):
    """n spheres in a small volume: overlaps, exact duplicates (equal-t ties), zero and huge radii, glass, mirrors."""
    out = []
    while len(out) < n:
        base = [host.STONE, host.PLASTIC, host.MIRROR, host.GLASS][rng.integers(4)]
        m = dict(base, rgb=tuple(rng.random(3)))
        if rng.random() < 0.2: m["reflectivity"] = float(rng.choice([0.0, 0.5, 1.0]))
        if rng.random() < 0.2: m["transperent"] = int(rng.integers(2))
        c = rng.normal(size=3) * np.array([3.0, 1.0, 3.0]) + np.array([0, 1.2, 4])
        r = float(rng.choice([0.0, 0.25, 0.6, 2.5])) if rng.random() < 0.25 else float(rng.random() * 0.7 + 0.05)
        out.append(host.sphere(tuple(c), r, m))
        if rng.random() < 0.15:
            out.append(host.sphere(tuple(c), r, dict(m, rgb=tuple(rng.random(3)))))   # duplicate geometry, other colour
    return b"".join(out[:n])


@pytest.mark.parametrize("seed", range(24))
def test_random_crowds_accel_equals_brute_force_and_reference(pkg, ref, scene, host, seed):
    rng = np.random.default_rng(7000 + seed)
    n = int(rng.choice([16, 17, 40, 97, 200]))
    s = crowd(host, rng, n)
    p = host.plane((0, 1, 0), (0, 0, 0), dict(host.STONE, rgb=(0, 0, 0), texture_id=2, texture_scale=100.0))
    if seed % 3 == 0:
        p += host.plane((0, 0, -1), (0, 0, 9), dict(host.MIRROR, reflectivity=1.0))
    l = b"".join(host.light(tuple(rng.normal(size=3) * 3 + np.array([0, 5, 2])), float(rng.choice([0.1, 0.4])), float(20 + rng.random() * 60),
                            tuple(rng.random(3))) for _ in range(int(rng.integers(1, 4))))
    W, H = 96, 64
    origin = tuple(rng.normal(size=3) * 2 + np.array([0, 1.5, -3]))
    if seed % 4 == 0:                               # camera at a sphere centre
        origin = tuple(np.frombuffer(s[:12], np.float32).astype(float))
    cam = host.camera(origin=origin, look=tuple(rng.normal(size=3) * 0.3 + np.array([0, -0.1, 1])), fov=float(rng.choice([60.0, 90.0, 120.0])),
                      focal=1.0, W=W, H=H)
    sc = ref.Scene(s, p, l, scene.tex, scene.sky)
    want_rgb, want_hit = ref.judge().render(sc, cam, W, H)
    eng = pkg.Engine(0)
    try:
        eng.upload_images(0, scene.tex); eng.upload_images(1, scene.sky)
        eng.upload_scene(s, p, l)
        eng.set_camera(cam)
        (b_rgb, b_hit), (a_rgb, a_hit) = both(eng, W, H)
        assert np.array_equal(a_hit, b_hit) and np.array_equal(a_rgb, b_rgb), seed
        assert np.array_equal(a_hit, want_hit), seed
        c = ref.compare_frames(a_rgb, want_rgb, W, H)
        assert c["within1"] >= 0.999 and c["psnr"] >= 50.0, (seed, c)
    finally:
        eng.close()


def test_accel_with_non_finite_sphere_falls_back(fresh, ref, scene, host):
    """A NaN centre cannot be clustered: the scene is rendered by brute force and still matches."""
    s, p, l, camp = host.sphere_grid_scene(8)
    bad = bytearray(s)
    bad[96 * 5: 96 * 5 + 4] = np.float32(np.nan).tobytes()
    W, H = 160, 90
    cam = host.camera(W=W, H=H, **camp)
    fresh.upload_scene(bytes(bad), p, l)
    fresh.set_camera(cam)
    (b_rgb, b_hit), (a_rgb, a_hit) = both(fresh, W, H)
    assert np.array_equal(a_hit, b_hit) and np.array_equal(a_rgb, b_rgb)


def test_fullsize_1024_spheres_accel_matches_reference_rows(pkg, ref, scene, host):
    """BASELINE.json configs[2] at 3840x2160 through the cluster walk: the same golden rows as the brute-force kernel."""
    golden = np.load(os.path.join(ROOT, "tests", "golden", "spheres1024_4k_rows.npz"))
    s, p, l, camp = host.sphere_grid_scene(32)
    W, H = 3840, 2160
    eng = pkg.Engine(0)
    try:
        eng.upload_images(0, scene.tex); eng.upload_images(1, scene.sky)
        eng.upload_scene(s, p, l)
        eng.set_camera(host.camera(W=W, H=H, **camp))
        eng.set_option("accel", 1)
        for frame_no in range(3):
            rgb, hit = eng.render_to_host(W, H, want_hit=True)
            bad_rgb = [y for y in range(H) if zlib.crc32(rgb[y * W:(y + 1) * W].tobytes()) != int(golden["rows_rgb"][y])]
            bad_hit = [y for y in range(H) if zlib.crc32(hit[y * W:(y + 1) * W].tobytes()) != int(golden["rows_hit"][y])]
            print("accel frame %d: %.1f ms; rows with different colour words %d, hit ids %d" % (frame_no, eng.last_render_ms(), len(bad_rgb), len(bad_hit)))
            assert not bad_hit and not bad_rgb
    finally:
        eng.close()


# ---- N4 over the triangles of extension X1 (rt_kernels_tri_accel.cu) -------------------------------------------------
# Judge = the brute-force CUDA walk of the same build family (itself compared with the port and the numpy pin in
# tests/test_triangles_gpu.py).  X1 has no reference counterpart: parity unpinned either way.
def both_tri(eng, W, H, frames=1):
    eng.set_option("accel", 0)
    brute = eng.render_to_host(W, H, want_hit=True)
    eng.set_option("accel", 1)
    acc = None
    for _ in range(frames):                      # frame 0: skeleton schedule; later frames: measured-time schedule, tree kernel
        acc = eng.render_to_host(W, H, want_hit=True)
        assert np.array_equal(acc[1], brute[1]) and np.array_equal(acc[0], brute[0])
    eng.set_option("accel", 0)
    return brute, acc


@pytest.mark.parametrize("n,W,H", [(4, 160, 90), (12, 192, 108), (23, 256, 144), (40, 320, 180)])
def test_heightfield_meshes_tri_accel_equals_brute_force(fresh, ref, scene, host, n, W, H):
    """32 / 288 / 1,058 / 3,200 triangles (2 .. 4 levels of the hierarchy, partial last cluster), glass and mirror spheres
    above the mesh, three soft lights: closest hits, occlusion of the lights and every shadow ray go through the walk."""
    s, p, l, tris, camp = host.heightfield_mesh_scene(n=n)
    fresh.upload_scene(s, p, l)
    fresh.upload_triangles(tris)
    fresh.set_camera(host.camera(W=W, H=H, **camp))
    (b_rgb, b_hit), _ = both_tri(fresh, W, H, frames=3)
    assert (b_hit > 1 + 2 + 1).sum() > W * H // 10                 # the mesh is in view
    if n == 12:                                                     # and the port agrees (as test_triangles_gpu.py checks for the brute-force walk)
        port = ref.port()
        port.set_triangles(tris)
        try:
            want_rgb, want_hit = port.render(ref.Scene(s, p, l, scene.tex, scene.sky), host.camera(W=W, H=H, **camp), W, H, wide=True)
        finally:
            port.set_triangles(b"")
        assert np.array_equal(b_hit, want_hit)
    fresh.set_option("accel", 1)
    for tree in (0, 2, 3):                                          # cooperative walk only; every tile / every pixel through the tree kernel
        fresh.set_option("tree", tree)
        rgb, hit = fresh.render_to_host(W, H, want_hit=True)
        assert np.array_equal(rgb, b_rgb) and np.array_equal(hit, b_hit), tree
    fresh.set_option("tree", 1)
    fresh.set_option("accel", 0)


def soup(host, rng, n):
    """n triangles in a small volume: overlaps, exact duplicates with another colour (equal-t ties: the smaller index must
    win), zero-area and needle triangles, tiny and huge ones, mirrors and glass-like materials."""
    out = []
    while len(out) < n:
        base = [host.STONE, host.PLASTIC, host.MIRROR][rng.integers(3)]
        m = dict(base, rgb=tuple(rng.random(3)))
        if rng.random() < 0.2: m["reflectivity"] = float(rng.choice([0.0, 0.5, 1.0]))
        c = rng.normal(size=3) * np.array([3.0, 1.0, 3.0]) + np.array([0, 1.5, 4])
        size = float(rng.choice([0.01, 0.2, 0.8, 6.0])) if rng.random() < 0.3 else float(rng.random() * 0.9 + 0.1)
        a, b = rng.normal(size=3) * size, rng.normal(size=3) * size
        kind = rng.random()
        if kind < 0.05: b = a * float(rng.random() * 2)             # zero area
        elif kind < 0.10: b = a + rng.normal(size=3) * 1e-4 * size   # needle
        v0, v1, v2 = c, c + a, c + b
        out.append(host.triangle(tuple(v0), tuple(v1), tuple(v2), m))
        if rng.random() < 0.15:
            out.append(host.triangle(tuple(v0), tuple(v1), tuple(v2), dict(m, rgb=tuple(rng.random(3)))))
        if rng.random() < 0.05:                                      # same plane, same vertices, other winding
            out.append(host.triangle(tuple(v0), tuple(v2), tuple(v1), dict(m, rgb=tuple(rng.random(3)))))
    return b"".join(out[:n])


@pytest.mark.parametrize("seed", range(24))
def test_triangle_soups_tri_accel_equals_brute_force(pkg, scene, host, seed):
    rng = np.random.default_rng(9000 + seed)
    n = int(rng.choice([16, 17, 63, 130, 520, 1100]))
    tris = soup(host, rng, n)
    s = crowd(host, rng, int(rng.choice([0, 3, 20]))) if seed % 2 else b""      # 20: the sphere clusters are walked too
    p = host.plane((0, 1, 0), (0, -0.5, 0), dict(host.STONE, rgb=(0, 0, 0), texture_id=2, texture_scale=100.0))
    if seed % 3 == 0:
        p += host.plane((0, 0, -1), (0, 0, 9), dict(host.MIRROR, reflectivity=1.0))
    l = b"".join(host.light(tuple(rng.normal(size=3) * 3 + np.array([0, 5, 2])), float(rng.choice([0.1, 0.4])), float(20 + rng.random() * 60),
                            tuple(rng.random(3))) for _ in range(int(rng.integers(1, 4))))
    W, H = 96, 64
    origin = tuple(rng.normal(size=3) * 2 + np.array([0, 1.5, -3]))
    if seed % 4 == 0:                                                # camera on a triangle vertex
        origin = tuple(np.frombuffer(tris[:12], np.float32).astype(float))
    look = np.array([0, 1.5, 4]) - np.array(origin)                 # towards the soup
    look = look / max(np.linalg.norm(look), 1e-3) + rng.normal(size=3) * 0.2
    cam = host.camera(origin=origin, look=tuple(look), fov=float(rng.choice([60.0, 90.0, 120.0])), focal=1.0, W=W, H=H)
    eng = pkg.Engine(0)
    try:
        eng.upload_images(0, scene.tex); eng.upload_images(1, scene.sky)
        eng.upload_scene(s, p, l)
        eng.upload_triangles(tris)
        eng.set_camera(cam)
        (b_rgb, b_hit), _ = both_tri(eng, W, H, frames=2)
        eng.set_option("accel", 1)
        eng.set_option("tree", 2)
        rgb, hit = eng.render_to_host(W, H, want_hit=True)
        assert np.array_equal(rgb, b_rgb) and np.array_equal(hit, b_hit), seed
    finally:
        eng.close()


def test_tri_accel_with_non_finite_triangle_falls_back(fresh, scene, host):
    """A NaN vertex cannot be clustered: the mesh is rendered by the brute-force build and still matches."""
    s, p, l, tris, camp = host.heightfield_mesh_scene(n=8)
    bad = bytearray(tris)
    bad[112 * 5: 112 * 5 + 4] = np.float32(np.nan).tobytes()
    W, H = 160, 90
    fresh.upload_scene(s, p, l)
    fresh.upload_triangles(bytes(bad))
    fresh.set_camera(host.camera(W=W, H=H, **camp))
    both_tri(fresh, W, H)


def test_mesh50k_tri_accel_equals_brute_force_at_1080p(pkg, scene, host):
    """BASELINE.json configs[3]'s mesh (49,928 triangles, 5 levels) at 1920x1080: every word of the frame and of the hit
    ids equal to the brute-force walk's (which renders every tile in the tree kernel from TMA-staged chunks)."""
    s, p, l, tris, camp = host.heightfield_mesh_scene()
    W, H = 1920, 1080
    eng = pkg.Engine(0)
    try:
        eng.upload_images(0, scene.tex); eng.upload_images(1, scene.sky)
        eng.upload_scene(s, p, l)
        eng.upload_triangles(tris)
        eng.set_camera(host.camera(W=W, H=H, **camp))
        brute = eng.render_to_host(W, H, want_hit=True)
        t_brute = eng.last_render_ms()
        eng.set_option("accel", 1)
        for frame_no in range(3):
            acc = eng.render_to_host(W, H, want_hit=True)
            print("mesh50k 1080p frame %d: hierarchy %.2f ms, brute force %.1f ms" % (frame_no, eng.last_render_ms(), t_brute))
            assert np.array_equal(acc[1], brute[1]) and np.array_equal(acc[0], brute[0]), frame_no
    finally:
        eng.close()
