"""CPU replay of the N4-over-triangles walk (csrc/rt_trace_accel.cuh) on the 49,928-triangle mesh for variants the library
does not build: clusters of K = 4 instead of 8.  Counts node tests and triangle tests per ray (camera rays of the 4K
frame and the shadow rays of their hit points); float64 geometry -- counts, not parity.
    python tools/exp/tacc_counts.py [n_rays]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402

g.load_package()
from example_gui_opencl_raytracer_b200 import host  # noqa: E402

NR = int(sys.argv[1]) if len(sys.argv) > 1 else 150
s, p, l, tris, camp = host.heightfield_mesh_scene()
T = len(tris) // 112
wire = np.frombuffer(tris, np.float32).reshape(T, 28).astype(np.float64)
V = np.stack([wire[:, 0:3], wire[:, 4:7], wire[:, 8:11]], axis=1)            # [T, 3, 3]
cen = V.mean(axis=1)


def build(K):
    """the kd-style order of rt_engine.cu: build_tri_accel_host, for clusters of K"""
    nc = (T + K - 1) // K
    sizes = [nc]
    while len(sizes) < 6 and sizes[-1] > 4:
        sizes.append((sizes[-1] + 7) // 8)
    cap = K * 8 ** (len(sizes) - 1)
    order = np.arange(T)
    jobs = [(b, min(b + cap, T), cap) for b in range(0, T, cap)]
    while jobs:
        b, e, c = jobs.pop()
        if c <= K or e - b <= 1:
            continue
        half, mid = c // 2, b + c // 2
        if e > mid:
            ext = cen[order[b:e]]
            ax = int(np.argmax(ext.max(axis=0) - ext.min(axis=0)))
            seg = order[b:e]
            part = np.argpartition(cen[seg, ax], mid - b - 1) if mid - b - 1 >= 0 else np.arange(e - b)
            key = cen[seg, ax]
            idx = np.argsort(key, kind="stable")                             # a full sort is a valid nth_element
            order[b:e] = seg[idx]
            jobs.append((mid, e, half))
        jobs.append((b, min(e, mid), half))
    levels = []
    for lv, n in enumerate(sizes):
        span = K * 8 ** lv
        lo = np.zeros((n, 3)); hi = np.zeros((n, 3))
        for k in range(n):
            pts = V[order[k * span: min((k + 1) * span, T)]].reshape(-1, 3)
            lo[k], hi[k] = pts.min(axis=0), pts.max(axis=0)
        levels.append(((lo + hi) / 2, (hi - lo) / 2))
    return order, levels, nc


def cull(o, inv, c, h, tmax):
    a = c - o
    mu = (np.abs(a).sum() + h.sum()) / 4096.0
    with np.errstate(all="ignore"):
        tc = a * inv
        te = (h + mu) * np.abs(inv)
        tn = np.fmax(np.fmax(tc[0] - te[0], tc[1] - te[1]), tc[2] - te[2])
        tf = np.fmin(np.fmin(tc[0] + te[0], tc[1] + te[1]), tc[2] + te[2])
    return tn > tf or tf < 0 or tn > tmax


def mt(o, d, tri):
    e1, e2 = tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0]
    pv = np.cross(d[None, :], e2)
    det = (e1 * pv).sum(axis=1)
    with np.errstate(all="ignore"):
        sv = o[None, :] - tri[:, 0]
        u = (sv * pv).sum(axis=1) / det
        q = np.cross(sv, e1)
        v = (q * d[None, :]).sum(axis=1) / det
        t = (e2 * q).sum(axis=1) / det
    hit = (det != 0) & (u >= 0) & (u <= 1) & (v >= 0) & (u + v <= 1) & (t > 0)
    return np.where(hit, t, np.inf).min() if len(t) else np.inf


def walk(o, d, K, order, levels, nc, tmax, anyhit):
    with np.errstate(all="ignore"):
        inv = 1.0 / d
    top = len(levels) - 1
    up = lambda kk: min(((kk & -kk).bit_length() - 1) // 3, top)
    k, lv, nodes, tests, best = 0, top, 0, 0, np.inf
    while k < nc:
        nodes += 1
        c, h = levels[lv]
        i = k >> (3 * lv)
        if cull(o, inv, c[i], h[i], min(tmax, best)):
            k += 1 << (3 * lv)
            if k < nc: lv = up(k)
        elif lv:
            lv -= 1
        else:
            tri = V[order[k * K: (k + 1) * K]]
            tests += K
            t = mt(o, d, tri)
            if t < min(tmax, best):
                best = t
                if anyhit: break
            k += 1
            if k < nc: lv = up(k)
    return nodes, tests, best


W, H = 3840, 2160
cam = np.asarray(host.camera(W=W, H=H, **camp), np.float64)
corner, origin, upv, right, wf, hf = cam[0:3], cam[3:6], cam[6:9], cam[9:12], cam[12], cam[13]
lights = [(-6, 8, -4), (8, 6, 6), (0, 10, 14)]
rng = np.random.default_rng(5)
px = [(int(rng.integers(W)), int(rng.integers(H))) for _ in range(NR)]
for K in (8, 4):
    order, levels, nc = build(K)
    tot = np.zeros(4); n_p = n_s = 0
    for x, y in px:
        d = corner + right * wf * x - upv * hf * y
        d /= np.linalg.norm(d)
        nodes, tests, best = walk(origin, d, K, order, levels, nc, np.inf, False)
        tot[0] += nodes; tot[1] += tests; n_p += 1
        if np.isfinite(best):
            hp = origin + d * best
            for lp in lights:
                dd = np.array(lp, float) - hp
                ln = np.linalg.norm(dd)
                dd /= ln
                nodes, tests, _ = walk(hp + dd * 1e-3, dd, K, order, levels, nc, ln, True)
                tot[2] += nodes; tot[3] += tests; n_s += 1
    print("K=%d: %d clusters, %d levels; camera rays: %.1f node tests + %.1f triangle tests; shadow rays: %.1f + %.1f"
          % (K, nc, len(levels), tot[0] / n_p, tot[1] / n_p, tot[2] / max(n_s, 1), tot[3] / max(n_s, 1)))
