"""xorshift32 jump-ahead tables for the level-wise pixel-tree kernel: column form of the GF(2) matrices
(step^d)^(2^j), j = 0..31, d = draws per shaded segment = 4L (raytracing.cl:96-113: 2 samples x 2 draws per light).
Independent numpy restatement of oracle/rt_split.c:jump_init; writes tests/golden/xorshift_jump.npz
(tables for L = 1..4 plus known answers).  Usage: python tools/make_jump_tables.py"""
import os
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
M32 = np.uint64(0xFFFFFFFF)


def step(x):
    """primitives.cl:119-121 on uint32 arrays."""
    x = np.asarray(x, np.uint64)
    x = (x ^ (x << np.uint64(13))) & M32
    x = x ^ (x >> np.uint64(17))
    x = (x ^ (x << np.uint64(5))) & M32
    return x.astype(np.uint32)


def apply(cols, x):
    r = 0
    for i in range(32):
        if (int(x) >> i) & 1:
            r ^= int(cols[i])
    return r


def compose(a, b):
    """columns of (a after b)"""
    return np.array([apply(a, int(b[i])) for i in range(32)], np.uint32)


def tables(draws):
    one = step(np.uint32(1) << np.arange(32, dtype=np.uint32))
    acc = (np.uint32(1) << np.arange(32, dtype=np.uint32)).astype(np.uint32)
    for _ in range(draws):
        acc = compose(one, acc)
    out = [acc]
    for _ in range(31):
        out.append(compose(out[-1], out[-1]))
    return np.stack(out)          # [32 powers][32 columns]


def jump(tab, seed, k):
    j = 0
    while k:
        if k & 1:
            seed = apply(tab[j], seed)
        k >>= 1
        j += 1
    return seed


if __name__ == "__main__":
    out = {}
    kat = []
    for L in (1, 2, 3, 4):
        tab = tables(4 * L)
        out["L%d" % L] = tab
        for seed in (0, 1, 479999, 8294399, 0xFFFFFFFF):
            x = np.uint32(seed)
            for k in range(0, 40):
                if k in (0, 1, 2, 3, 17, 39):
                    kat.append((L, seed, k, int(x), jump(tab, seed, k)))
                    assert int(x) == jump(tab, seed, k)
                for _ in range(4 * L):
                    x = step(x)
            kat.append((L, seed, 123457, jump(tab, seed, 123457), jump(tab, seed, 123457)))
    out["kat"] = np.array([(a, b, c, e) for a, b, c, d, e in kat], np.uint64)   # L, seed, k, state
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "xorshift_jump.npz"), **out)
    print("wrote tests/golden/xorshift_jump.npz:", {k: v.shape for k, v in out.items()})
