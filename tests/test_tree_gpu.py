"""The level-wise pixel-tree trace (csrc/rt_trace_tree.cuh; CPU proof: oracle/rt_split.c, tests/test_split_cpu.py) against the
reference text and against the cooperative stack-machine walk: "tree" = 2 runs EVERY tile as one pooled tree pass,
"tree" = 3 every pixel on its own (what a single-pixel queue entry of a heavy tile does), "tree" = 0 never.
All three must give the same words; the oracle is oracle/_ref (the reference's .cl text)."""
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


def all_modes(eng, W, H, **kw):
    out = {}
    for tree in (0, 2, 3):
        eng.set_option("tree", tree)
        out[tree] = eng.render_to_host(W, H, want_hit=True, **kw)
    eng.set_option("tree", 1)
    return out


@pytest.mark.parametrize("W,H", [(800, 600), (320, 240), (101, 67), (8, 4), (1, 1)])
def test_render_map_tree_modes_match_reference(fresh, ref, scene, W, H):
    cam = ref.camera(W=W, H=H, **ref.RAYPNG_CAMERA)
    fresh.upload_scene(scene.spheres, scene.planes, scene.lights)
    fresh.set_camera(cam)
    want_rgb, want_hit = ref.judge().render(scene, cam, W, H)
    got = all_modes(fresh, W, H)
    for tree, (rgb, hit) in got.items():
        assert np.array_equal(hit, want_hit), tree
        c = ref.compare_frames(rgb, want_rgb, W, H)
        print("tree=%d: exact %.6f within1 %.6f psnr %.2f" % (tree, c["exact"], c["within1"], c["psnr"]))
        assert c["within1"] >= 0.999 and c["psnr"] >= 50.0
        assert np.array_equal(rgb, got[0][0]), tree          # and the same words as the stack-machine walk


def test_tree_with_row_bands_and_total_guard(fresh, ref, scene, pkg):
    """Global pixel ids seed the jump-ahead: a row-band part and a clipped total give the 1-part frame's words."""
    W, H = 328, 250
    fresh.upload_scene(scene.spheres, scene.planes, scene.lights)
    fresh.set_camera(ref.camera(W=W, H=H, **ref.RAYPNG_CAMERA))
    fresh.set_option("tree", 0)
    whole, whole_hit = fresh.render_to_host(W, H, want_hit=True)
    for tree in (2, 3):
        fresh.set_option("tree", tree)
        bufs = [fresh.render_to_host(W, H, part=pkg.Partition(8, 3, r), want_hit=True) for r in range(3)]
        assert np.array_equal(pkg.assemble_bands([b[0] for b in bufs], W, H, 8), whole)
        assert np.array_equal(pkg.assemble_bands([b[1] for b in bufs], W, H, 8), whole_hit)
        total = W * 100 + 17
        fresh.render(W, H, want_hit=False)
        fresh.render(W, H, total=total, want_hit=False)
        assert np.array_equal(fresh.read_frame(W * H), whole)


@pytest.mark.parametrize("n_side,W,H", [(8, 320, 180), (15, 256, 144), (32, 160, 90)])
def test_sphere_grids_tree_modes(fresh, ref, scene, host, n_side, W, H):
    """Many glass spheres: deep trees, several passes per level, the WIDE kernel variant around it (>= 64 records)."""
    s, p, l, camp = host.sphere_grid_scene(n_side)
    cam = host.camera(W=W, H=H, **camp)
    sc = ref.Scene(s, p, l, scene.tex, scene.sky)
    want_rgb, want_hit = ref.judge().render(sc, cam, W, H, wide=max(sc.S, sc.P, sc.L) > 255)
    fresh.upload_scene(s, p, l)
    fresh.set_camera(cam)
    for tree, (rgb, hit) in all_modes(fresh, W, H).items():
        assert np.array_equal(hit, want_hit), tree
        c = ref.compare_frames(rgb, want_rgb, W, H)
        print("n_side=%d tree=%d: exact %.6f within1 %.6f psnr %.2f" % (n_side, tree, c["exact"], c["within1"], c["psnr"]))
        assert c["within1"] >= 0.999 and c["psnr"] >= 50.0


def test_tree_overflow_falls_back_to_the_stack_machine(fresh, ref, scene, host):
    """Nested glass + facing mirrors grow trees of hundreds of nodes per pixel; with a 64-node slab most tiles
    overflow and are re-run by the cooperative walk -- the words cannot change."""
    sp = (host.sphere((0, 1.2, 2), 1.2, dict(host.GLASS)) + host.sphere((0, 1.2, 2), 0.6, dict(host.GLASS, n=1.33)) +
          host.sphere((2.0, 0.7, 0.5), 0.7, dict(host.MIRROR, reflectivity=1.0)) +
          host.sphere((0.5, 2.6, -2.0), 0.4, dict(host.GLASS)))
    pl = (host.plane((0, 1, 0), (0, 0, 0), dict(host.STONE, texture_id=2, texture_scale=100.0)) +
          host.plane((1, 0, 0), (-4, 0, 0), dict(host.MIRROR, reflectivity=1.0)) +
          host.plane((-1, 0, 0), (4, 0, 0), dict(host.MIRROR, reflectivity=1.0)))
    li = host.light((0.5, 3.2, -1.0), .25, 40.0, (1, 1, 1)) + host.light((-1, 2, 6), .1, 25.0, (.2, .4, 1))
    W, H = 320, 200
    cam = host.camera(origin=(0.3, 2.0, -6.0), look=(0.0, -0.1, 1.0), fov=90.0, focal=1.0, W=W, H=H)
    sc = ref.Scene(sp, pl, li, scene.tex, scene.sky)
    want_rgb, want_hit = ref.judge().render(sc, cam, W, H)
    fresh.upload_scene(sp, pl, li)
    fresh.set_camera(cam)
    for cap in (2048, 64, 256):
        fresh.set_option("tree_cap", cap)
        for tree, (rgb, hit) in all_modes(fresh, W, H).items():
            assert np.array_equal(hit, want_hit), (cap, tree)
            c = ref.compare_frames(rgb, want_rgb, W, H)
            assert c["within1"] >= 0.999 and c["psnr"] >= 50.0, (cap, tree, c)
    fresh.set_option("tree_cap", 2048)


def test_tree_many_lights_and_none(fresh, ref, scene, host):
    """The jump-ahead matrices depend on the light count (4L draws per shaded segment): L = 0, 1, 5, 16 use the tree,
    L = 17 exceeds the slab's per-node addition slots and stays on the cooperative walk."""
    W, H = 160, 100
    cam = ref.camera(W=W, H=H, **ref.RAYPNG_CAMERA)
    for L in (0, 1, 5, 16, 17):
        li = b"".join(host.light((-4 + 0.5 * k, 3 + 0.1 * k, -2 + 0.4 * k), .1 + .01 * k, 20.0 + k, (1, 1 - .05 * k, .2 + .05 * k))
                      for k in range(L))
        sc = ref.Scene(bytes(scene.spheres), bytes(scene.planes), li, scene.tex, scene.sky)
        want_rgb, want_hit = ref.judge().render(sc, cam, W, H)
        fresh.upload_scene(bytes(scene.spheres), bytes(scene.planes), li)
        fresh.set_camera(cam)
        for tree, (rgb, hit) in all_modes(fresh, W, H).items():
            assert np.array_equal(hit, want_hit), (L, tree)
            c = ref.compare_frames(rgb, want_rgb, W, H)
            assert c["within1"] >= 0.999 and c["psnr"] >= 50.0, (L, tree, c)


def test_tree_entries_at_4k_every_tile_split(engine, ref):
    """Every tile of a 4K frame forced into 32 single-pixel queue entries, each one a pixel tree (and, with tree = 0,
    each one a lone stack-machine walk): the frame's words do not change."""
    W, H = 3840, 2160
    engine.set_camera(ref.camera(W=W, H=H, **ref.RAYPNG_CAMERA))
    engine.set_option("history", 0)
    want = engine.render_to_host(W, H, want_hit=True)
    engine.set_option("history", 1)
    engine.set_option("split_min_cost", 1)
    engine.set_option("split32_min_cost", 1)
    try:
        for tree in (1, 0):
            engine.set_option("tree", tree)
            for rep in range(2):
                rgb, hit = engine.render_to_host(W, H, want_hit=True)
                print("tree=%d frame %d: %.2f ms" % (tree, rep, engine.last_render_ms()))
                assert np.array_equal(rgb, want[0]) and np.array_equal(hit, want[1]), (tree, rep)
    finally:
        engine.set_option("tree", 1)
        engine.set_option("split_min_cost", 0)
        engine.set_option("split32_min_cost", 0)


def test_tree_with_triangles_extension(fresh, ref, scene, host):
    """The triangle build of the kernels (rt_kernels_tri.cu) carries the same tree trace.  Judge: the port (X1 has no
    reference code: parity unpinned)."""
    tris = b"".join(host.triangle((-3 + k, 0.0, 2 + k), (-2 + k, 2.5, 3 + k), (-1 + k, 0.0, 2.5 + k),
                                  dict(host.MIRROR if k % 2 else host.PLASTIC, rgb=(.2 + .1 * k, .8 - .1 * k, .4), reflectivity=.5))
                    for k in range(5))
    W, H = 200, 120
    cam = ref.camera(W=W, H=H, **ref.RAYPNG_CAMERA)
    port = ref.port()
    port.set_triangles(tris)
    try:
        want_rgb, want_hit = port.render(scene, cam, W, H)
    finally:
        port.set_triangles(b"")
    fresh.upload_scene(scene.spheres, scene.planes, scene.lights)
    fresh.upload_triangles(tris)
    fresh.set_camera(cam)
    for tree, (rgb, hit) in all_modes(fresh, W, H).items():
        assert np.array_equal(hit, want_hit), tree
        c = ref.compare_frames(rgb, want_rgb, W, H)
        assert c["within1"] >= 0.999 and c["psnr"] >= 50.0, (tree, c)
