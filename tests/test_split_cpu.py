"""The per-pixel stack machine splits into geometry walk / independent shading / ordered replay with bit-identical
output (oracle/rt_split.c).  This is the property a kernel needs to spread one heavy pixel over many warps: the
walk never reads the RNG or the accumulated colour, each shaded segment draws exactly 4L numbers
(raytracing.cl:96-113, no early exit), and xorshift32 (primitives.cl:116-125) is linear over GF(2)."""
import ctypes as C
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SPLIT_SO = os.path.join(ROOT, "oracle", "librt_split.so")
u8p, u32p, f32p = C.POINTER(C.c_uint8), C.POINTER(C.c_uint32), C.POINTER(C.c_float)


@pytest.fixture(scope="module")
def split():
    lib = C.CDLL(SPLIT_SO)
    lib.rto_split_render.restype = C.c_int
    lib.rto_split_render.argtypes = [C.c_int, f32p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                     C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32,
                                     u8p, C.c_int, C.c_int, C.c_int, u8p, C.c_int, C.c_int, u32p, u32p]
    lib.rto_tree_render.restype = C.c_int
    lib.rto_tree_render.argtypes = lib.rto_split_render.argtypes
    lib.rto_split_jump.restype = C.c_uint32
    lib.rto_split_jump.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32]
    return lib


@pytest.fixture(scope="module")
def host(pkg):
    from example_gui_opencl_raytracer_b200 import host as h
    return h


def split_render(lib, sc, cam, W, H, wide=False, tree=False):
    if tree == "flat":                                   # ranks and replay without recursion, over the flat node array
        lib.rto_tree_mode(1)
        try:
            return split_render(lib, sc, cam, W, H, wide=wide, tree=True)
        finally:
            lib.rto_tree_mode(0)
    """tree=False: walk / shade / replay in program order of the walk; tree=True: segments expanded breadth-first
    (second output: warp passes << 24 | levels << 16 | segments instead of the shaded-segment count)."""
    cam = np.ascontiguousarray(cam, np.float32)
    out = np.zeros(W * H, np.uint32)
    shaded = np.zeros(W * H, np.uint32)
    rc = (lib.rto_tree_render if tree else lib.rto_split_render)(int(wide), cam.ctypes.data_as(f32p), W, H, 0, H, sc.spheres.ctypes.data, sc.S,
                              sc.planes.ctypes.data, sc.P, sc.lights.ctypes.data, sc.L,
                              sc.tex.ctypes.data_as(u8p), sc.tex.shape[2], sc.tex.shape[1], sc.tex.shape[0],
                              sc.sky.ctypes.data_as(u8p), sc.sky.shape[2], sc.sky.shape[1],
                              out.ctypes.data_as(u32p), shaded.ctypes.data_as(u32p))
    assert rc == 0
    return out, shaded


def test_jump_ahead_equals_stepping(ref, split):
    """xorshift32^(d*k)(seed) by bit-matrix powers == d*k calls of the port's generator; seed 0 stays 0 (Q1)."""
    lib = ref.port().lib
    lib.rto_xorshift32.restype = C.c_float
    for seed in (0, 1, 2, 479999, 8294399, 0xFFFFFFFF, 0x80000000):
        for draws in (4, 12, 20):
            st = C.c_uint32(seed)
            k_done = 0
            for k in (0, 1, 2, 3, 17, 255, 4097):
                for _ in range((k - k_done) * draws):
                    lib.rto_xorshift32(C.byref(st))
                k_done = k
                assert split.rto_split_jump(seed, draws, k) == st.value, (seed, draws, k)


@pytest.mark.parametrize("W,H", [(160, 120), (97, 61)])
def test_split_equals_stack_machine_on_render_map(ref, scene, split, W, H):
    cam = ref.camera((0.8, 2.5, -8.0), (0.2, 0.0, 1.0), 90.0, 1.0, W, H)
    want, _ = ref.port().render(scene, cam, W, H)
    got, shaded = split_render(split, scene, cam, W, H)
    assert np.array_equal(got, want)
    assert shaded.max() > 15          # glass pixels: more shaded segments than one chain can hold -> real sub-trees
    got, shape = split_render(split, scene, cam, W, H, tree=True)
    assert np.array_equal(got, want)
    got_f, shape_f = split_render(split, scene, cam, W, H, tree="flat")
    assert np.array_equal(got_f, want) and np.array_equal(shape_f, shape)
    nodes, levels, steps = shape & 0xffff, (shape >> 16) & 0xff, shape >> 24
    heavy = int(np.argmax(nodes))
    print("breadth-first: heaviest pixel %d segments in %d levels (%d passes of 32 lanes); mean %.2f segments in %.2f levels" % (
        nodes[heavy], levels[heavy], steps[heavy], nodes.mean(), levels.mean()))
    assert levels.max() <= 16 and nodes.max() > 4 * levels.max()   # the chain a warp would walk: levels, not segments


@pytest.mark.parametrize("n_side,W,H", [(8, 96, 54), (17, 64, 36)])
def test_split_equals_stack_machine_on_sphere_grids(ref, scene, split, host, n_side, W, H):
    s, p, l, camp = host.sphere_grid_scene(n_side)
    sc = ref.Scene(s, p, l, scene.tex, scene.sky)
    wide = max(sc.S, sc.P, sc.L) > 255
    cam = host.camera(W=W, H=H, **camp)
    want, _ = ref.port().render(sc, cam, W, H, wide=wide)
    got, shaded = split_render(split, sc, cam, W, H, wide=wide)
    assert np.array_equal(got, want)
    for mode in (True, "flat"):
        got_t, _ = split_render(split, sc, cam, W, H, wide=wide, tree=mode)
        assert np.array_equal(got_t, want), mode
    print("n_side %d: shaded segments per pixel max %d mean %.2f" % (n_side, shaded.max(), shaded.mean()))


@pytest.mark.parametrize("seed", range(24))
def test_split_equals_stack_machine_on_random_scenes(ref, scene, split, host, seed):
    """The fuzz generator of the GPU parity tests: TIR, NaN directions, zero radii, duplicates, 0..4 lights."""
    from test_fuzz_gpu import random_scene
    rng = np.random.default_rng(1000 + seed)
    s, p, l = random_scene(host, rng)
    W, H = 48, 32
    origin = tuple(rng.normal(size=3) * 2 + np.array([0, 1.5, -4]))
    if seed % 5 == 0 and len(s):                       # camera at a sphere centre
        origin = tuple(np.frombuffer(s[:12], np.float32).astype(float))
    cam = host.camera(origin=origin, look=tuple(rng.normal(size=3) * 0.3 + np.array([0, -0.1, 1])),
                      fov=float(rng.choice([60.0, 90.0, 120.0])), focal=1.0, W=W, H=H)
    sc = ref.Scene(s, p, l, scene.tex, scene.sky)
    want, _ = ref.port().render(sc, cam, W, H)
    got, _ = split_render(split, sc, cam, W, H)
    for mode in (True, "flat"):
        got_t, _ = split_render(split, sc, cam, W, H, tree=mode)
        assert np.array_equal(got_t, want), mode
    # NaN colours pack through an undefined float->uint conversion on the CPU; both sides run the same code, so equal
    assert np.array_equal(got, want)


def test_jump_tables_golden(split):
    """tests/golden/xorshift_jump.npz (tools/make_jump_tables.py, numpy) against the C jump-ahead: the tables the
    level-wise kernel will keep in constant memory, with known answers incl. a far jump (k = 123,457)."""
    g = np.load(os.path.join(ROOT, "tests", "golden", "xorshift_jump.npz"))
    for L, seed, k, state in g["kat"].tolist():
        assert split.rto_split_jump(int(seed), 4 * int(L), int(k)) == int(state), (L, seed, k)
    tab = g["L3"]
    assert tab.shape == (32, 32) and tab.dtype == np.uint32
    # column i of pw[0] = 12 steps applied to bit i
    for i in (0, 7, 31):
        assert int(tab[0][i]) == split.rto_split_jump(1 << i, 12, 1)
