"""EXTENSION X1 (SURVEY.md 8(a')): triangles.  The reference has no triangle type, so these tests pin the CUDA
path to oracle/rt_oracle.c's OWN definition (self-consistency; "parity unpinned" w.r.t. the reference)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


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


def both(ref, scene0, eng, s, p, l, tris, cam, W, H):
    sc = ref.Scene(s, p, l, scene0.tex, scene0.sky)
    eng.upload_scene(s, p, l)
    eng.upload_triangles(tris)
    eng.set_camera(cam)
    port = ref.port()
    port.set_triangles(tris)
    try:
        want_rgb, want_hit, want_cnt = port.render_counted(sc, cam, W, H, wide=True)
    finally:
        port.set_triangles(b"")
    out = {}
    for trace in (0, 1, 2):
        eng.set_option("trace", trace)
        rgb, hit = eng.render_to_host(W, H, want_hit=True)
        c = ref.compare_frames(rgb, want_rgb, W, H)
        same = float((hit == want_hit).mean())
        print("trace %d T=%d: hit %.7f exact %.6f within1 %.6f psnr %.2f" % (trace, len(tris) // 112, same, c["exact"],
                                                                          c["within1"], c["psnr"]))
        assert same == 1.0 and c["within1"] >= 0.999 and c["psnr"] >= 50.0
        out[trace] = rgb
    eng.set_option("trace", 0)
    assert np.array_equal(out[0], out[2])
    got_cnt = eng.render_counted(W, H)
    for k in ("segments", "shadow_rays", "sphere_tests", "plane_tests", "tri_tests"):
        assert abs(got_cnt[k] - want_cnt[k]) <= 1e-4 * want_cnt[k] + 2, (k, got_cnt[k], want_cnt[k])
    return want_hit


def test_single_triangles_with_render_map(fresh, ref, scene, host):
    tris = (host.triangle((-3, 0.2, 2), (-1, 2.5, 3), (-2, 0.2, 4), dict(host.PLASTIC, rgb=(.9, .6, .1))) +
            host.triangle((2, 0.1, -2), (4, 0.1, -1), (3, 2.0, -1.5), dict(host.MIRROR, reflectivity=1.0)) +
            host.triangle((0, 3.0, 6.5), (1.5, 4.5, 6.5), (-1.5, 4.5, 6.5), dict(host.STONE, rgb=(.2, .7, .9))))
    W, H = 320, 200
    hit = both(ref, scene, fresh, bytes(scene.spheres), bytes(scene.planes), bytes(scene.lights), tris,
               ref.camera(W=W, H=H, **ref.RAYPNG_CAMERA), W, H)
    ids = set(np.unique(hit).tolist())
    assert {1 + 4 + 2 + 0, 1 + 4 + 2 + 1} <= ids     # triangles are visible and carry ids after the planes


def test_heightfield_mesh(fresh, ref, scene, host):
    s, p, l, tris, camp = host.heightfield_mesh_scene(n=12)      # 288 triangles
    W, H = 192, 108
    both(ref, scene, fresh, s, p, l, tris, host.camera(W=W, H=H, **camp), W, H)


def test_triangles_only_and_removal(fresh, ref, scene, host):
    tris = host.triangle((-2, 0, 3), (2, 0, 3), (0, 3, 3), dict(host.PLASTIC, rgb=(.1, .8, .3)))
    W, H = 128, 96
    cam = host.camera(origin=(0, 1, -3), look=(0, 0, 1), fov=90.0, focal=1.0, W=W, H=H)
    both(ref, scene, fresh, b"", b"", bytes(scene.lights), tris, cam, W, H)
    fresh.upload_triangles(b"")                       # T = 0 removes them
    rgb, hit = fresh.render_to_host(W, H, want_hit=True)
    want_rgb, want_hit = ref.port().render(ref.Scene(b"", b"", bytes(scene.lights), scene.tex, scene.sky), cam, W, H)
    assert np.array_equal(hit, want_hit) and np.array_equal(rgb, want_rgb)


def test_triangle_primary_hits_equal_the_numpy_pin(fresh, ref, scene, host):
    """Second, independent judge for X1 (still not reference parity -- the reference has no triangles): the CUDA kernel's
    primary hit ids on a triangles-only frame against vectorised float32 numpy Moeller-Trumbore over rays from the
    reference's raygen.cl text (tests/test_extensions_cpu.py: numpy_triangle_hit_ids)."""
    import test_extensions_cpu as ext
    tris, camp = ext.triangle_pin_scene(host)
    W, H = 80, 45
    cam = host.camera(W=W, H=H, **camp)
    want = ext.numpy_triangle_hit_ids(ref, tris, cam, W, H)
    fresh.upload_scene(b"", b"", b"")
    fresh.upload_triangles(tris)
    fresh.set_camera(cam)
    for tree in (1, 2):
        fresh.set_option("tree", tree)
        _, hit = fresh.render_to_host(W, H, want_hit=True)
        assert np.array_equal(hit, want), tree
    fresh.set_option("tree", 1)


@pytest.mark.parametrize("n,W,H", [(12, 192, 108), (20, 160, 90), (23, 96, 54)])
def test_staged_triangle_walk_in_the_tree_kernel(fresh, ref, scene, host, n, W, H):
    """Large meshes render every tile in the CTA-wide tree kernel with the triangle records staged into shared memory in
    256-record chunks by TMA (rt_trace_tree.cuh: tri_walk_staged): 288 / 800 / 1,058 triangles = 2 / 4 / 5 chunks, the
    last one partial.  Forced here with "tri_tree_min" = 1 and with "tree" = 2 / 3; judge = the port (X1: parity unpinned),
    and the un-staged cooperative walk must give the same words."""
    s, p, l, tris, camp = host.heightfield_mesh_scene(n=n)
    cam = host.camera(W=W, H=H, **camp)
    sc = ref.Scene(s, p, l, scene.tex, scene.sky)
    port = ref.port()
    port.set_triangles(tris)
    try:
        want_rgb, want_hit = port.render(sc, cam, W, H, wide=True)
    finally:
        port.set_triangles(b"")
    fresh.upload_scene(s, p, l)
    fresh.upload_triangles(tris)
    fresh.set_camera(cam)
    fresh.set_option("tri_tree_min", 0)
    fresh.set_option("tree", 0)
    base_rgb, base_hit = fresh.render_to_host(W, H, want_hit=True)
    assert np.array_equal(base_hit, want_hit)
    c = ref.compare_frames(base_rgb, want_rgb, W, H)
    assert c["within1"] >= 0.999 and c["psnr"] >= 50.0
    for opts in ({"tree": 1, "tri_tree_min": 1}, {"tree": 2, "tri_tree_min": 0}, {"tree": 3, "tri_tree_min": 0}):
        for k, v in opts.items():
            fresh.set_option(k, v)
        for rep in range(2):               # second frame: history-ordered queue
            rgb, hit = fresh.render_to_host(W, H, want_hit=True)
            assert np.array_equal(hit, base_hit), (opts, rep)
            assert np.array_equal(rgb, base_rgb), (opts, rep)
        if opts["tri_tree_min"]:
            assert fresh.last_launch()["tree_grid"] > 0
    fresh.set_option("tree", 1)
    fresh.set_option("tri_tree_min", 1024)
