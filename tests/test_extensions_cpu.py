"""CPU-side checks of the extensions: scene archives v1/v2 (csrc/map_io.c) and the oracle port's triangle test
(EXTENSION X1 -- defined by oracle/rt_oracle.c itself; here it is compared with an independent float32 restatement)."""
import ctypes as C
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def host(pkg):
    from example_gui_opencl_raytracer_b200 import host as h
    return h


def test_map_v1_is_read_like_the_reference_loader(host):
    s, p, l = host.load_map(os.path.join(ROOT, "scenes", "render.map"))      # python mirror of cpu_obj.c:76-101
    s2, p2, l2, t2, ver = host.load_map_any(os.path.join(ROOT, "scenes", "render.map"))
    assert ver == 1 and t2 == b"" and (s2, p2, l2) == (s, p, l)


def test_map_v2_round_trip_and_limits(host, tmp_path):
    s, p, l, camp = host.sphere_grid_scene(17)                   # 289 spheres: not expressible in v1
    _, _, _, tris, _ = host.heightfield_mesh_scene(n=4)
    f2 = str(tmp_path / "big.map")
    host.save_map(f2, s, p, l, tris, version=2)
    assert host.load_map_any(f2) == (s, p, l, tris, 2)
    with pytest.raises(IOError):
        host.save_map(str(tmp_path / "x.map"), s, p, l, b"", version=1)
    with pytest.raises(IOError):
        host.save_map(str(tmp_path / "y.map"), s[:96 * 3], p, l, tris, version=1)   # triangles do not fit v1 either
    f1 = str(tmp_path / "small.map")
    host.save_map(f1, s[:96 * 200], p, l, b"", version=1)
    assert host.load_map(f1) == (s[:96 * 200], p, l)            # what the reference's extract_robj would read
    assert host.load_map_any(f1)[4] == 1
    bad = str(tmp_path / "bad.map")
    open(bad, "wb").write(open(f2, "rb").read()[:-5])
    with pytest.raises(IOError):
        host.load_map_any(bad)
    # empty scene, both versions
    for v in (1, 2):
        fe = str(tmp_path / ("e%d.map" % v))
        host.save_map(fe, b"", b"", b"", b"", version=v)
        assert host.load_map_any(fe)[:4] == (b"", b"", b"", b"")


def mt_float32(o, d, v0, v1, v2):
    f = np.float32
    def dot(a, b): return f(f(f(a[0] * b[0]) + f(a[1] * b[1])) + f(a[2] * b[2]))
    def cross(a, b): return [f(f(a[1] * b[2]) - f(a[2] * b[1])), f(f(a[2] * b[0]) - f(a[0] * b[2])), f(f(a[0] * b[1]) - f(a[1] * b[0]))]
    def sub(a, b): return [f(a[i] - b[i]) for i in range(3)]
    e1, e2 = sub(v1, v0), sub(v2, v0)
    p = cross(d, e2)
    det = dot(e1, p)
    if det == 0: return None
    inv = f(f(1.0) / det)
    s = sub(o, v0)
    u = f(dot(s, p) * inv)
    if u < 0 or u > 1: return None
    q = cross(s, e1)
    v = f(dot(d, q) * inv)
    if v < 0 or f(u + v) > 1: return None
    t = f(dot(e2, q) * inv)
    return None if t <= 0 else t


def test_oracle_triangle_kat(ref):
    lib = ref.port().lib
    lib.rto_isect_triangle.restype = C.c_int
    rng = np.random.default_rng(7)
    fp = C.POINTER(C.c_float)
    hits = 0
    with np.errstate(all="ignore"):
        for k in range(4000):
            a = rng.normal(size=15).astype(np.float32) * np.float32(3.0)
            o, d, v0, v1, v2 = a[0:3].copy(), a[3:6].copy(), a[6:9].copy(), a[9:12].copy(), a[12:15].copy()
            if k % 7 == 0: d = (v0 + v1 + v2) / np.float32(3) - o          # aim at the centroid: certain hits
            if k % 11 == 0: v2 = v1.copy()                                  # degenerate: det == 0
            t = C.c_float(-1)
            got = lib.rto_isect_triangle(o.ctypes.data_as(fp), d.ctypes.data_as(fp), v0.ctypes.data_as(fp),
                                         v1.ctypes.data_as(fp), v2.ctypes.data_as(fp), C.byref(t))
            want = mt_float32([np.float32(x) for x in o], [np.float32(x) for x in d], [np.float32(x) for x in v0],
                              [np.float32(x) for x in v1], [np.float32(x) for x in v2])
            assert bool(got) == (want is not None)
            if got:
                hits += 1
                assert np.float32(t.value).tobytes() == np.float32(want).tobytes()
    assert hits > 400


def test_oracle_triangle_scene_is_deterministic_and_visible(ref, scene, host):
    tris = host.triangle((-2, 0, 3), (2, 0, 3), (0, 3, 3), dict(host.PLASTIC, rgb=(.1, .8, .3)))
    W, H = 64, 48
    cam = host.camera(origin=(0, 1, -3), look=(0, 0, 1), fov=90.0, focal=1.0, W=W, H=H)
    sc = ref.Scene(b"", b"", bytes(scene.lights), scene.tex, scene.sky)
    port = ref.port()
    base_rgb, base_hit = port.render(sc, cam, W, H)
    port.set_triangles(tris)
    try:
        rgb1, hit1 = port.render(sc, cam, W, H, threads=1)
        rgb2, hit2 = port.render(sc, cam, W, H, threads=4)
    finally:
        port.set_triangles(b"")
    assert np.array_equal(rgb1, rgb2) and np.array_equal(hit1, hit2)
    assert (hit1 == 1).sum() > 100 and (base_hit == 1).sum() == 0          # id 1+S+P+0 with S = P = 0
    rgb3, _ = port.render(sc, cam, W, H)
    assert np.array_equal(rgb3, base_rgb)                                   # reset really removes them


def test_triangle_prefilter_is_exact():
    """rt_device.cuh:hit_triangle decides most u-test rejects without the division (RT_TRI_PREFILTER).  Claim checked
    here in IEEE binary32 (numpy): whenever the shortcut fires, the defining sequence u = (s.p) * (1/det),
    `u < 0 || u > 1` (oracle/rt_oracle.c:hit_triangle) rejects as well -- over ordinary, near-boundary, tiny, huge,
    zero, infinite and NaN operands."""
    rng = np.random.default_rng(0x7E1)
    f = np.float32
    n = 400_000
    mag = lambda k: (rng.standard_normal(k) * np.exp2(rng.uniform(-140, 127, k))).astype(f)
    det = np.concatenate([rng.standard_normal(n).astype(f), mag(n), mag(n), rng.standard_normal(n).astype(f)])
    ratio = np.concatenate([rng.uniform(-2, 3, n), rng.uniform(-2, 3, n), np.exp2(rng.uniform(-160, 60, n)) * rng.choice([-1, 1], n),
                            1 + rng.uniform(-3e-4, 3e-4, n)])
    with np.errstate(all="ignore"):
        a = (det.astype(np.float64) * ratio).astype(f)
        special = np.array([0.0, -0.0, np.inf, -np.inf, np.nan, 1e-45, -1e-45, 3.4e38, -3.4e38, 1.17549435e-38, 1.0, -1.0], dtype=f)
        sa, sd = np.meshgrid(special, special)
        a = np.concatenate([a, sa.ravel(), rng.permutation(a)[:n]])
        det = np.concatenate([det, sd.ravel(), det[:n]])
        keep = det != 0
        a, det = a[keep], det[keep]
        ad, aa = np.abs(det), np.abs(a)
        opposite = (a < 0) != (det < 0)
        fires = np.where(opposite, aa > ad * f(1e-30), aa > ad * f(1.0001))
        inv = f(1.0) / det
        u = a * inv
        rejects = (u < 0) | (u > 1)
    assert fires.sum() > 100_000 and (~fires).sum() > 100_000
    assert not np.any(fires & ~rejects)


def test_triangle_batch_defers_the_zero_determinant_check():
    """rt_device.cuh:for_triangle_hits tests `det == 0` only for the records that survive the prefilter.  Same outcome
    as hit_triangle (which rejects det == 0 first): with det = +-0 the prefilter either rejects (a is a non-zero number)
    or lets the record through to the survivors' pass, whose first test is det == 0."""
    f = np.float32
    special = np.array([0.0, -0.0, np.inf, -np.inf, np.nan, 1e-45, -1e-45, 3.4e38, -3.4e38, 1.17549435e-38, 1.0, -1.0, 7.5, -0.3], dtype=f)
    for det in (f(0.0), f(-0.0)):
        for a in special:
            ad, aa = np.abs(det), np.abs(a)
            with np.errstate(all="ignore"):
                fires = (aa > ad * f(1e-30)) if ((a < 0) != (det < 0)) else (aa > ad * f(1.0001))
            batch_rejects = bool(fires) or det == 0      # prefilter, else the survivors' det == 0 test
            assert batch_rejects                          # hit_triangle: det == 0 -> no hit
            if not (a == 0 or np.isnan(a)):
                assert fires                              # ... and the hot loop alone already says so


def numpy_triangle_hit_ids(ref, tris, cam, W, H):
    """A SECOND, independent pin for extension X1 (the reference has no triangle type, so this is still not reference
    parity): the primary hit-id frame of a triangles-only scene computed here in vectorised IEEE binary32 numpy --
    rays from the reference's own raygen.cl text (oracle/_ref), Moeller-Trumbore in the operation order of
    mt_float32 above, closest hit = first of strictly smaller t in index order (primitives.cl:322-394's rule),
    id = 1 + S + P + k with S = P = 0."""
    f = np.float32
    lib = ref.ref().lib
    o = np.zeros((W * H, 3), f); d = np.zeros((W * H, 3), f)
    fp = C.POINTER(C.c_float)
    camf = np.ascontiguousarray(cam, f)
    oo, dd = np.zeros(3, f), np.zeros(3, f)
    for i in range(W * H):
        lib.ref_raygen(camf.ctypes.data_as(fp), C.c_uint32(W), C.c_uint32(H), C.c_uint32(i), oo.ctypes.data_as(fp), dd.ctypes.data_as(fp))
        o[i] = oo; d[i] = dd
    T = len(tris) // 112
    rec = np.frombuffer(tris, f).reshape(T, 28)
    v0, v1, v2 = rec[:, 0:3], rec[:, 4:7], rec[:, 8:11]
    dot = lambda a, b: (a[..., 0] * b[..., 0] + a[..., 1] * b[..., 1]) + a[..., 2] * b[..., 2]
    cross = lambda a, b: np.stack([a[..., 1] * b[..., 2] - a[..., 2] * b[..., 1], a[..., 2] * b[..., 0] - a[..., 0] * b[..., 2],
                                   a[..., 0] * b[..., 1] - a[..., 1] * b[..., 0]], axis=-1)
    best_t = np.full(W * H, np.inf, f)
    best_k = np.full(W * H, -1, np.int64)
    with np.errstate(all="ignore"):
        for k in range(T):                      # index order; strict improvement only
            e1, e2 = (v1[k] - v0[k]).astype(f), (v2[k] - v0[k]).astype(f)
            p = cross(d, e2[None, :])
            det = dot(e1[None, :], p)
            inv = f(1.0) / det
            s = o - v0[k][None, :]
            u = dot(s, p) * inv
            q = cross(s, e1[None, :])
            v = dot(d, q) * inv
            t = dot(e2[None, :], q) * inv
            hit = (det != 0) & ~((u < 0) | (u > 1)) & ~((v < 0) | (u + v > 1)) & ~(t <= 0)
            better = hit & ~(t >= best_t)
            best_t = np.where(better, t, best_t)
            best_k = np.where(better, k, best_k)
    assert o.dtype == f and d.dtype == f and best_t.dtype == f
    return np.where(best_k >= 0, 1 + best_k, 0).astype(np.uint32)


def triangle_pin_scene(host):
    s, p, l, tris, camp = host.heightfield_mesh_scene(n=10)      # 200 triangles, mixed materials
    return tris, camp


def test_triangle_primary_hits_numpy_pin_vs_port(ref, scene, host):
    """The port's primary hit ids on a triangles-only frame equal the numpy float32 restatement's (ties and all)."""
    tris, camp = triangle_pin_scene(host)
    W, H = 80, 45
    cam = host.camera(W=W, H=H, **camp)
    want = numpy_triangle_hit_ids(ref, tris, cam, W, H)
    sc = ref.Scene(b"", b"", b"", scene.tex, scene.sky)          # no spheres, planes or lights: ids are 1 + k
    port = ref.port()
    port.set_triangles(tris)
    try:
        _, hit = port.render(sc, cam, W, H)
    finally:
        port.set_triangles(b"")
    assert (want > 0).sum() > 500
    assert np.array_equal(hit, want)


# ---- N4 over the triangles (csrc/rt_trace_accel.cuh: tacc_cull / tacc_walk), replayed on the CPU ------------------------
def _mt_all(o, d, v0, e1, e2):
    """float32 Moeller-Trumbore of ONE ray against every record (the operation order of hit_triangle); -> (hit mask, t)."""
    f = np.float32
    dot = lambda a, b: (a[..., 0] * b[..., 0] + a[..., 1] * b[..., 1]) + a[..., 2] * b[..., 2]
    cross = lambda a, b: np.stack([a[..., 1] * b[..., 2] - a[..., 2] * b[..., 1], a[..., 2] * b[..., 0] - a[..., 0] * b[..., 2],
                                   a[..., 0] * b[..., 1] - a[..., 1] * b[..., 0]], axis=-1)
    with np.errstate(all="ignore"):
        p = cross(d[None, :], e2)
        det = dot(e1, p)
        inv = f(1.0) / det
        s = o[None, :] - v0
        u = dot(s, p) * inv
        q = cross(s, e1)
        v = dot(d[None, :], q) * inv
        t = dot(e2, q) * inv
        hit = (det != 0) & ~((u < 0) | (u > 1)) & ~((v < 0) | (u + v > 1)) & ~(t <= 0)
    return hit, t


class _Hierarchy:
    def __init__(self, pkg, rec):
        L_ = pkg.lib()
        vp = lambda a: a.ctypes.data_as(C.c_void_p)
        T = len(rec)
        nc, nl, nn = C.c_uint32(), C.c_uint32(), C.c_uint32()
        off = (C.c_uint32 * 6)()
        assert L_.rtb_debug_tri_accel_build(vp(rec), T, None, None, None, C.byref(nc), C.byref(nl), off, C.byref(nn)) == 1
        self.nodes = np.zeros(8 * nn.value, np.float32); self.srt = np.zeros(12 * 8 * nc.value, np.float32); self.idx = np.zeros(8 * nc.value, np.uint32)
        assert L_.rtb_debug_tri_accel_build(vp(rec), T, vp(self.nodes), vp(self.srt), vp(self.idx), C.byref(nc), C.byref(nl), off, C.byref(nn)) == 1
        self.nodes = self.nodes.reshape(-1, 8); self.srt = self.srt.reshape(-1, 12)
        self.nc, self.nl, self.off = nc.value, nl.value, list(off)

    def cull(self, o, inv, ainv, nd, tmax):
        """tacc_cull in float32 (the FMA of mu through float64: one rounding, as __fmaf_rn)."""
        f = np.float32
        with np.errstate(all="ignore"):
            a = nd[0:3] - o
            mu = f(np.float64((np.abs(a[0]) + np.abs(a[1])) + np.abs(a[2])) * np.float64(f(0.000244140625)) + np.float64(nd[7]))
            c = a * inv
            e = (nd[4:7] + mu) * ainv
            tn = np.fmax(np.fmax(c[0] - e[0], c[1] - e[1]), c[2] - e[2])      # fmaxf / fminf: a NaN operand drops out
            tf = np.fmin(np.fmin(c[0] + e[0], c[1] + e[1]), c[2] + e[2])
            return bool(tn > tf or tf < 0 or tn > tmax)

    def walk(self, o, d, tmax_fn, leaf):
        """tacc_walk: cluster positions in order, one node per step; leaf(k) returns True to stop.  -> (node tests, leaves)"""
        f = np.float32
        with np.errstate(all="ignore"):
            inv = f(1.0) / d
        ainv = np.abs(inv)
        top, k, l, tests, leaves = self.nl - 1, 0, self.nl - 1, 0, 0
        up = lambda kk: min(((kk & -kk).bit_length() - 1) // 3, top)
        while k < self.nc:
            tests += 1
            if self.cull(o, inv, ainv, self.nodes[self.off[l] + (k >> (3 * l))], tmax_fn()):
                k += 1 << (3 * l)
                if k < self.nc: l = up(k)
            elif l:
                l -= 1
            else:
                leaves += 1
                if leaf(k): break
                k += 1
                if k < self.nc: l = up(k)
        return tests, leaves


def test_triangle_hierarchy_walk_replay_equals_brute_force(pkg, host):
    """The device walk of rt_trace_accel.cuh restated here (same cull arithmetic in float32, same visiting order, same
    closest-hit and any-hit rules) over the hierarchy the library builds, against the brute-force loop, for rays of the
    kinds a frame produces: camera rays, shadow rays leaving the surface, rays from far away (the ground plane near the
    horizon: origins hundreds to thousands of units out, where the cull's margin is widest) and rays along an axis (a zero
    direction component: infinite slab bounds).  Closest hit = (t, original index) equal; any-hit equal; and the walk
    must actually skip most of the mesh for the near rays."""
    f = np.float32
    rng = np.random.default_rng(424242)
    _, _, _, tris, _ = host.heightfield_mesh_scene(n=24)                  # 1,152 triangles over [-8, 8] x [-4, 12]
    T = len(tris) // 112
    wire = np.frombuffer(tris, f).reshape(T, 28)
    rec = np.zeros((T, 12), f)
    rec[:, 0:3] = wire[:, 0:3]; rec[:, 4:7] = wire[:, 4:7] - wire[:, 0:3]; rec[:, 8:11] = wire[:, 8:11] - wire[:, 0:3]
    h = _Hierarchy(pkg, rec)
    v0, e1, e2 = rec[:, 0:3], rec[:, 4:7], rec[:, 8:11]
    s0, s1, s2 = h.srt[:, 0:3], h.srt[:, 4:7], h.srt[:, 8:11]
    centre = np.array([0.0, 0.6, 4.0])
    rays = []
    for _ in range(500):                                                   # camera-like
        o = np.array([rng.normal() * 3, 4 + rng.random() * 4, -9 + rng.normal()])
        rays.append((o, centre + rng.normal(size=3) * np.array([6, 0.5, 6]) - o))
    for _ in range(500):                                                   # leaving the surface towards a light
        k = rng.integers(T)
        p = v0[k] + 0.3 * e1[k] + 0.3 * e2[k] + np.array([0, 1e-3, 0])
        rays.append((p.astype(np.float64), np.array([rng.normal() * 8, 6 + rng.random() * 4, rng.normal() * 8 + 4]) - p))
    for _ in range(400):                                                   # from far away, low over the ground
        ang = rng.random() * 2 * np.pi
        dist = 10 ** (2 + 2 * rng.random())                                # 100 .. 10,000
        o = np.array([np.cos(ang) * dist, -0.6 + rng.random() * 0.01, 4 + np.sin(ang) * dist])
        aim = (np.array([rng.normal() * 8, 6 + rng.random() * 4, rng.normal() * 8 + 4]) if rng.random() < 0.5      # a light above the mesh
               else centre + rng.normal(size=3) * np.array([5, 0.3, 5]))                                             # the mesh itself
        rays.append((o, aim - o))
    for _ in range(100):                                                   # axis-parallel: a zero component
        o = np.array([rng.normal() * 6, 5.0, 4 + rng.normal() * 6])
        rays.append((o, np.array([0.0, -1.0, 0.0]) if rng.random() < 0.5 else np.array([0.0, -1.0, rng.normal() * 0.2])))
    near_tests = near_leaves = n_near = hits = blocked_n = 0
    for n, (o64, d64) in enumerate(rays):
        o = o64.astype(f)
        d = (d64 / np.linalg.norm(d64)).astype(f)
        hit, t = _mt_all(o, d, v0, e1, e2)
        assert not np.isnan(t[hit]).any()
        if hit.any():
            tb = t[hit].min()
            want = (tb, int(np.flatnonzero(hit & (t == tb))[0]))           # first of equal t in index order
        else:
            want = (f(np.inf), -1)
        best = [f(np.inf), -1]

        def leaf(k):
            hk, tk = _mt_all(o, d, s0[8 * k: 8 * k + 8], s1[8 * k: 8 * k + 8], s2[8 * k: 8 * k + 8])
            for j in np.flatnonzero(hk):
                orig = int(h.idx[8 * k + j])
                if tk[j] < best[0] or (tk[j] == best[0] and best[1] >= 0 and orig < best[1]):
                    best[0], best[1] = tk[j], orig
            return False
        tests, leaves = h.walk(o, d, lambda: f(best[0] * f(1.00001)), leaf)
        assert (best[0], best[1]) == want, (n, best, want)
        hits += want[1] >= 0
        if n < 1000:
            near_tests += tests; near_leaves += leaves; n_near += 1
        # any-hit with a random length (a shadow ray): blocked iff some member has !(t >= len)
        length = f(rng.random() * 30 if want[1] < 0 else want[0] * (0.5 + rng.random()))
        want_blocked = bool((hit & ~(t >= length)).any())
        got = [False]

        def leaf_any(k):
            hk, tk = _mt_all(o, d, s0[8 * k: 8 * k + 8], s1[8 * k: 8 * k + 8], s2[8 * k: 8 * k + 8])
            got[0] = bool((hk & ~(tk >= length)).any())
            return got[0]
        h.walk(o, d, lambda: f(length * f(1.00001)), leaf_any)
        assert got[0] == want_blocked, n
        blocked_n += want_blocked
    print("rays with a hit: %d of %d; blocked shadow rays: %d" % (hits, len(rays), blocked_n))
    assert hits > 500 and blocked_n > 200
    print("near rays: %.1f node tests and %.1f of %d clusters per closest-hit walk" % (near_tests / n_near, near_leaves / n_near, h.nc))
    assert near_leaves / n_near < 0.1 * h.nc


def test_triangle_hierarchy_walk_replay_tie_rule_on_duplicates(pkg, host):
    """Equal-t ties: a soup in which every third triangle is an exact duplicate of an earlier one (another index, the same
    geometry) and some are the same triangle with the other winding.  The brute-force loop keeps the FIRST of equal t in
    index order; the hierarchy visits in its own order and must still name that one."""
    f = np.float32
    rng = np.random.default_rng(777)
    recs = []
    while len(recs) < 400:
        c = rng.normal(size=3) * np.array([3.0, 1.0, 3.0]) + np.array([0, 1.5, 4])
        a, b = rng.normal(size=3) * 0.8, rng.normal(size=3) * 0.8
        r = np.zeros(12, f); r[0:3] = c; r[4:7] = a; r[8:11] = b
        recs.append(r)
        if rng.random() < 0.5: recs.append(r.copy())                       # exact duplicate: equal t on every ray
        if rng.random() < 0.2:
            w = r.copy(); w[4:7], w[8:11] = r[8:11], r[4:7]                # other winding: the same plane and vertices
            recs.append(w)
    order = rng.permutation(len(recs))                                      # duplicates far apart in index order
    rec = np.stack([recs[i] for i in order]).astype(f)
    T = len(rec)
    h = _Hierarchy(pkg, rec)
    v0, e1, e2 = rec[:, 0:3], rec[:, 4:7], rec[:, 8:11]
    s0, s1, s2 = h.srt[:, 0:3], h.srt[:, 4:7], h.srt[:, 8:11]
    ties = 0
    for n in range(600):
        o = (rng.normal(size=3) * 2 + np.array([0, 1.5, -3])).astype(f)
        d64 = np.array([0, 1.5, 4]) + rng.normal(size=3) * np.array([3, 1, 3]) - o
        d = (d64 / np.linalg.norm(d64)).astype(f)
        hit, t = _mt_all(o, d, v0, e1, e2)
        if not hit.any():
            continue
        tb = t[hit].min()
        at_tb = np.flatnonzero(hit & (t == tb))
        ties += len(at_tb) > 1
        want = (tb, int(at_tb[0]))
        best = [f(np.inf), -1]

        def leaf(k):
            hk, tk = _mt_all(o, d, s0[8 * k: 8 * k + 8], s1[8 * k: 8 * k + 8], s2[8 * k: 8 * k + 8])
            for j in np.flatnonzero(hk):
                orig = int(h.idx[8 * k + j])
                if tk[j] < best[0] or (tk[j] == best[0] and best[1] >= 0 and orig < best[1]):
                    best[0], best[1] = tk[j], orig
            return False
        h.walk(o, d, lambda: f(best[0] * f(1.00001)), leaf)
        assert (best[0], best[1]) == want, (n, best, want)
    assert ties > 100                                                      # the rule was actually exercised
