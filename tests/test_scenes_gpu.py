"""GPU parity on scenes other than render.map: the synthetic sphere grids of BASELINE.json config 3 (narrow
uchar counts and the widened >255 counts), and the edge cases of SURVEY.md 8(a): empty
arrays, a plane whose tangent basis is undefined (Q16), glass inside glass (Q4), a light inside the view,
mirror corridors that run the ray stack to MAX_DEPTH.  Same tolerances as test_parity_gpu.py."""
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


def run_both(ref, scene0, eng, s, p, l, cam, W, H):
    sc = ref.Scene(s, p, l, scene0.tex, scene0.sky)
    wide = max(sc.S, sc.P, sc.L) > 255
    eng.upload_scene(s, p, l)
    eng.set_camera(cam)
    rgb, hit = eng.render_to_host(W, H, want_hit=True)
    o = ref.judge()
    want_rgb, want_hit = o.render(sc, cam, W, H, wide=wide)
    c = ref.compare_frames(rgb, want_rgb, W, H)
    same = float((hit == want_hit).mean())
    print("S=%d P=%d L=%d %dx%d: hit %.7f exact %.6f within1 %.6f psnr %.2f" % (sc.S, sc.P, sc.L, W, H, same, c["exact"],
                                                                            c["within1"], c["psnr"]))
    assert same == 1.0
    assert c["within1"] >= 0.999 and c["psnr"] >= 50.0
    return c


@pytest.mark.parametrize("n_side,W,H", [(8, 320, 180), (15, 256, 144), (17, 192, 108), (32, 160, 90)])
def test_sphere_grid_matches_oracle(fresh, ref, scene, host, n_side, W, H):
    """15x15 = 225 fits the reference's uchar counts; 17x17 = 289 and 32x32 = 1,024 need the widened build."""
    s, p, l, camp = host.sphere_grid_scene(n_side)
    run_both(ref, scene, fresh, s, p, l, host.camera(W=W, H=H, **camp), W, H)


@pytest.mark.parametrize("n_side,W,H", [(15, 384, 216), (32, 256, 144)])
def test_wide_variant_is_bit_identical(fresh, scene, host, n_side, W, H):
    """Scenes with >= 64 records run the WIDE kernel variant (rt_trace_coop.cuh); the narrow one must give the same
    words, at the automatic break-even and with every round forced ray-by-ray."""
    s, p, l, camp = host.sphere_grid_scene(n_side)
    fresh.upload_scene(s, p, l)
    fresh.set_camera(host.camera(W=W, H=H, **camp))
    fresh.set_option("wide_min_records", -1)
    narrow, nh = fresh.render_to_host(W, H, want_hit=True)
    fresh.set_option("wide_min_records", 64)
    for max_live in (-1, 32, 8):
        fresh.set_option("wide_max_live", max_live)
        for _ in range(2):      # second frame: history-ordered queue with split tiles
            wide, wh = fresh.render_to_host(W, H, want_hit=True)
            assert np.array_equal(wide, narrow) and np.array_equal(wh, nh), max_live


def test_max_uchar_counts(fresh, ref, scene, host):
    """255 spheres, 255 planes and 255 lights would take minutes on the CPU; 255 spheres + 7 planes + 5 lights."""
    s, _, _, camp = host.sphere_grid_scene(16)
    s = s[:255 * 96]
    planes = b"".join(host.plane((0, 1, 0), (0, -0.1 * k, 0), dict(host.STONE, texture_id=k % 4, texture_scale=10.0 + k))
                      for k in range(7))
    lights = b"".join(host.light((-4 + 2 * k, 6, -3 + k), .2, 60.0 + 10 * k, (1, 1 - .1 * k, .5 + .1 * k)) for k in range(5))
    run_both(ref, scene, fresh, s, planes, lights, host.camera(W=128, H=72, **camp), 128, 72)


@pytest.mark.parametrize("case", ["nothing", "no_lights", "no_spheres", "no_planes", "only_lights"])
def test_empty_arrays(fresh, ref, scene, host, case):
    s0, p0, l0 = scene.spheres, scene.planes, scene.lights
    s, p, l = {"nothing": (b"", b"", b""), "no_lights": (s0, p0, b""), "no_spheres": (b"", p0, l0),
               "no_planes": (s0, b"", l0), "only_lights": (b"", b"", l0)}[case]
    W, H = 200, 120
    run_both(ref, scene, fresh, bytes(s), bytes(p), bytes(l), ref.camera(W=W, H=H, **ref.RAYPNG_CAMERA), W, H)


def test_plane_with_undefined_tangent_basis(fresh, ref, scene, host):
    """Q16: for n proportional to (1,1,1) no axis passes the sum-of-cross test; both sides use a zero basis."""
    k = 1.0 / np.sqrt(3.0)
    planes = host.plane((k, k, k), (0, -1, 0), dict(host.STONE, texture_id=1, texture_scale=7.0)) + \
        host.plane((0, 0, -1), (0, 0, 9), dict(host.STONE, texture_id=3, texture_scale=33.0))
    W, H = 240, 160
    run_both(ref, scene, fresh, bytes(scene.spheres), planes, bytes(scene.lights),
             ref.camera(W=W, H=H, **ref.RAYPNG_CAMERA), W, H)


@pytest.mark.parametrize("trace", [0, 2, 3])
def test_infinite_specular_in_shadow_is_nan_not_zero(fresh, ref, scene, host, trace):
    """A non-unit plane normal makes max(0, n.h) exceed 1, so pow(.., 100) overflows to +inf.  In full shadow the reference computes (0 * specular) * inf = NaN and the pixel packs
    to black; where lit it saturates.  The fast traces' "exact zero" shortcuts (R5) must not apply there
    (found by tests/test_fuzz_gpu.py seed 32)."""
    shiny = dict(host.STONE, rgb=(.6, .5, .9), diffuse=0.0, specular=0.6, shininess=100)
    odd = dict(host.PLASTIC, rgb=(.2, .8, .3), shininess=0)
    pl = host.plane((0, 6.0, 0), (0, 0, 0), shiny) + host.plane((0, 0, -1), (0, 0, 8), odd)
    sp = host.sphere((0.0, 1.0, 2.0), 1.0, dict(host.PLASTIC, rgb=(1, 0, 0))) + \
        host.sphere((-2.5, 0.8, 3.0), 0.8, dict(host.PLASTIC, rgb=(0, 0, 1), shininess=0))
    li = host.light((1.5, 4.0, 0.0), .1, 30.0, (1, 1, 1)) + host.light((-3.0, 2.0, -1.0), .1, 12.0, (1, .8, .6))
    W, H = 192, 128
    fresh.set_option("trace", trace)
    run_both(ref, scene, fresh, sp, pl, li, host.camera(origin=(0.5, 2.5, -5.0), look=(0, -0.25, 1.0), fov=90.0,
                                                       focal=1.0, W=W, H=H), W, H)


def test_glass_in_glass_and_mirror_corridor(fresh, ref, scene, host):
    """Nested dielectric spheres (Q4: no nested media), two facing mirrors (stack runs to MAX_DEPTH), a light
    in view (Q20) and a glass sphere in front of it (Q19)."""
    sp = (host.sphere((0, 1.2, 2), 1.2, dict(host.GLASS)) + host.sphere((0, 1.2, 2), 0.6, dict(host.GLASS, n=1.33)) +
          host.sphere((2.0, 0.7, 0.5), 0.7, dict(host.MIRROR, reflectivity=1.0)) +
          host.sphere((-2.2, 0.6, 0.0), 0.6, dict(host.PLASTIC, rgb=(.9, .2, .1))) +
          host.sphere((0.5, 2.6, -2.0), 0.4, dict(host.GLASS)))
    pl = (host.plane((0, 1, 0), (0, 0, 0), dict(host.STONE, texture_id=2, texture_scale=100.0)) +
          host.plane((1, 0, 0), (-4, 0, 0), dict(host.MIRROR, reflectivity=1.0)) +
          host.plane((-1, 0, 0), (4, 0, 0), dict(host.MIRROR, reflectivity=1.0)))
    li = host.light((0.5, 3.2, -1.0), .25, 40.0, (1, 1, 1)) + host.light((-1, 2, 6), .1, 25.0, (.2, .4, 1))
    W, H = 320, 200
    cam = host.camera(origin=(0.3, 2.0, -6.0), look=(0.0, -0.1, 1.0), fov=90.0, focal=1.0, W=W, H=H)
    run_both(ref, scene, fresh, sp, pl, li, cam, W, H)


def test_camera_inside_glass_and_degenerate_views(fresh, ref, scene):
    """Camera origin inside the glass sphere of render.map, and a camera looking straight down the plane normal."""
    W, H = 160, 120
    for cam in (ref.camera(origin=(0.8, 0.8, 1.5), look=(0.1, 0.2, 1.0), fov=90.0, focal=1.0, W=W, H=H),
                ref.camera(origin=(0.0, 5.0, 1.0), look=(0.0, -1.0, 0.001), fov=90.0, focal=1.0, W=W, H=H),
                ref.camera(origin=(0.8, 2.5, -8.0), look=(0.2, 0.0, 1.0), fov=170.0, focal=1.0, W=W, H=H)):
        run_both(ref, scene, fresh, bytes(scene.spheres), bytes(scene.planes), bytes(scene.lights), cam, W, H)
