"""CPU-side checks (no GPU): the C ABI exports what include/*.h declares, the host-side mirror agrees with
the oracle's helpers, the PNG codec decodes the shipped assets exactly, and the N>1 row-band gather
rebuilds the single-part frame (gloo, world_size 2)."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared(header):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return set(re.findall(r"\b((?:rtb|cl_wrap|cuda_wrap)_\w+)\s*\(", text))


def test_library_exports_every_declared_symbol(pkg):
    L = pkg.lib()   # raises if librtb200.so is missing
    names = declared("rtb200.h") | declared("cuda_wrap.h")
    assert len(names) >= 35
    for n in sorted(names):
        assert hasattr(L, n), "librtb200.so does not export " + n
    assert names == set(pkg.ENGINE_SYMBOLS) | set(pkg.WRAP_SYMBOLS)


def test_no_compute_without_a_gpu(pkg):
    """The product path must fail loudly, never fall back to a CPU implementation."""
    if pkg.lib().rtb_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(pkg.RtbError):
        pkg.Engine(0)


def test_cl_wrap_struct_layout(pkg):
    w = pkg.ClWrap
    assert w.buffers.offset == 4 * 8 + 8 + 16 * 8           # 4 handles, kernels_num (+pad), kernels[16]
    assert C.sizeof(w) == w.buffers.offset + 16 * 32 * 8 + 16 * 32 * 4 + 16 * 4
    src = subprocess.run(["gcc", "-Iinclude", "-Iinclude/shim", "-x", "c", "-", "-o", "/tmp/rtb_layout"], cwd=ROOT,
                         input=b'#include <stdio.h>\n#include <stddef.h>\n#include "cuda_wrap.h"\n'
                               b'int main(){printf("%zu %zu %zu", sizeof(cl_wrap), offsetof(cl_wrap, buffers), sizeof(cl_float3));return 0;}')
    assert src.returncode == 0
    out = subprocess.run(["/tmp/rtb_layout"], capture_output=True).stdout.decode().split()
    assert [int(v) for v in out] == [C.sizeof(w), w.buffers.offset, 16]


def test_camera_and_map_helpers_match_oracle_helpers(pkg, ref):
    from example_gui_opencl_raytracer_b200 import host
    for (W, H) in ((800, 600), (3840, 2160), (101, 67)):
        a = host.camera(W=W, H=H, **host.RAYPNG_CAMERA)
        b = ref.camera(W=W, H=H, **ref.RAYPNG_CAMERA)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    s, p, l = host.load_map(os.path.join(ROOT, "scenes", "render.map"))
    assert (len(s), len(p), len(l)) == (4 * 96, 2 * 96, 3 * 48)
    assert [s, p, l] == ref.load_map(os.path.join(ROOT, "scenes", "render.map"))
    # the packers reproduce the file (scene_dump.c:8-69 values)
    sp = (host.sphere((4.5, .5, -1), .5, dict(host.PLASTIC, rgb=(1, 0, 0))) + host.sphere((-1, 1, 4.5), .8, dict(host.PLASTIC, rgb=(0, 0, 1))) +
          host.sphere((.8, .8, 1.5), .8, dict(host.GLASS)) + host.sphere((-.6, .8, -1), .8, dict(host.GLASS, rgb=(0, 1, 0), ambient=.05)))
    assert sp == s
    assert host.render_map_lights() == l
    tmp = "/tmp/rtb_roundtrip.map"
    host.dump_map(tmp, s, p, l)
    assert open(tmp, "rb").read() == open(os.path.join(ROOT, "scenes", "render.map"), "rb").read()
    with pytest.raises(ValueError):
        host.dump_map(tmp, s * 64, p, l)      # 256 spheres do not fit the one-byte count


def test_png_codec_decodes_assets_exactly(pkg, ref):
    from example_gui_opencl_raytracer_b200 import host
    tex, sky = host.default_images(ROOT)
    rt, rs = ref.default_textures(ROOT)   # PIL
    assert tex.shape == (4, 256, 256, 4) and sky.shape == (1, 3072, 4096, 4)
    assert np.array_equal(tex, rt) and np.array_equal(sky, rs)
    # writer -> reader round trip, odd size
    L = pkg.lib()
    L.rtpng_write_rgb8.argtypes = [C.c_char_p, C.c_void_p, C.c_uint32, C.c_uint32]
    rgb = np.random.default_rng(1).integers(0, 256, size=(37, 53, 3), dtype=np.uint8)
    assert L.rtpng_write_rgb8(b"/tmp/rtb_rt.png", rgb.ctypes.data_as(C.c_void_p), 53, 37) == 0
    back = host.load_png_layers(["/tmp/rtb_rt.png"])
    assert np.array_equal(back[0, :, :, :3], rgb) and (back[..., 3] == 255).all()
    from PIL import Image
    assert np.array_equal(np.asarray(Image.open("/tmp/rtb_rt.png")), rgb)
    # the reference's format gate (opencl_wrap.c:299-307): 8-bit RGB only
    Image.fromarray(rgb[..., 0]).save("/tmp/rtb_gray.png")
    with pytest.raises(ValueError):
        host.load_png_layers(["/tmp/rtb_gray.png"])


def _png_with_filter(path, img, ftype):
    """Hand-made 8-bit RGB PNG whose every row (the first included) uses filter `ftype`."""
    import struct, zlib
    h, w, _ = img.shape
    rows = img.reshape(h, w * 3).astype(np.int32)
    raw = bytearray()
    for y in range(h):
        cur = rows[y]
        up = rows[y - 1] if y else np.zeros_like(cur)
        a = np.concatenate([np.zeros(3, np.int32), cur[:-3]])
        c = np.concatenate([np.zeros(3, np.int32), up[:-3]])
        if ftype == 0: pred = 0
        elif ftype == 1: pred = a
        elif ftype == 2: pred = up
        elif ftype == 3: pred = (a + up) >> 1
        else:
            pp = a + up - c
            pa, pb, pc = abs(pp - a), abs(pp - up), abs(pp - c)
            pred = np.where((pa <= pb) & (pa <= pc), a, np.where(pb <= pc, up, c))
        raw.append(ftype)
        raw += ((cur - pred) & 255).astype(np.uint8).tobytes()
    def chunk(t, d):
        return struct.pack(">I", len(d)) + t + d + struct.pack(">I", zlib.crc32(t + d))
    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)) +
                chunk(b"IDAT", zlib.compress(bytes(raw))) + chunk(b"IEND", b""))


def test_png_decoder_every_filter_type(pkg, tmp_path):
    from example_gui_opencl_raytracer_b200 import host
    rng = np.random.default_rng(7)
    for (h, w) in ((1, 1), (5, 2), (19, 33)):
        img = rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8)
        for ft in range(5):
            p = str(tmp_path / ("f%d_%dx%d.png" % (ft, w, h)))
            _png_with_filter(p, img, ft)
            got = host.load_png_layers([p], cache_dir="")
            assert np.array_equal(got[0, :, :, :3], img), (ft, h, w)
            assert (got[..., 3] == 255).all()


def _png_adam7(path, img, ftype):
    """Hand-made Adam7-interlaced 8-bit RGB PNG (PNG spec 8.2); every row of every pass uses filter `ftype` (0..2)."""
    import struct, zlib
    h, w, _ = img.shape
    raw = bytearray()
    for x0, y0, dx, dy in ((0, 0, 8, 8), (4, 0, 8, 8), (0, 4, 4, 8), (2, 0, 4, 4), (0, 2, 2, 4), (1, 0, 2, 2), (0, 1, 1, 2)):
        sub = img[y0::dy, x0::dx]
        if sub.shape[0] == 0 or sub.shape[1] == 0:
            continue
        rows = sub.reshape(sub.shape[0], -1).astype(np.int32)
        for y in range(rows.shape[0]):
            cur = rows[y]
            up = rows[y - 1] if y else np.zeros_like(cur)
            a = np.concatenate([np.zeros(3, np.int32), cur[:-3]])
            pred = 0 if ftype == 0 else (a if ftype == 1 else up)
            raw.append(ftype)
            raw += ((cur - pred) & 255).astype(np.uint8).tobytes()
    def chunk(t, d):
        return struct.pack(">I", len(d)) + t + d + struct.pack(">I", zlib.crc32(t + d))
    z = zlib.compress(bytes(raw))
    with open(path, "wb") as f:      # two IDAT chunks: the stream may be split anywhere
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 1)) +
                chunk(b"IDAT", z[:len(z) // 2]) + chunk(b"IDAT", z[len(z) // 2:]) + chunk(b"IEND", b""))


def test_png_decoder_adam7_interlace(pkg, tmp_path):
    """libpng's png_read_image de-interlaces (opencl_wrap.c:318), so the reference accepts Adam7 files: so do we."""
    from example_gui_opencl_raytracer_b200 import host
    from PIL import Image
    rng = np.random.default_rng(11)
    for (h, w) in ((1, 1), (2, 3), (7, 5), (8, 8), (19, 33), (64, 40)):
        img = rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8)
        for ft in range(3):
            p = str(tmp_path / ("a7_%d_%dx%d.png" % (ft, w, h)))
            _png_adam7(p, img, ft)
            assert np.array_equal(np.asarray(Image.open(p).convert("RGB")), img)      # the file is a valid PNG
            got = host.load_png_layers([p], cache_dir="")
            assert np.array_equal(got[0, :, :, :3], img), (ft, h, w)
            assert (got[..., 3] == 255).all()


def test_png_decoder_rejects_damaged_files(pkg, tmp_path):
    """A flipped payload byte (chunk CRC), a truncated zlib stream, a wrong Adler-32 trailer and trailing garbage in the
    stream are all decode errors -- never a silently half-decoded texture."""
    import struct, zlib
    from example_gui_opencl_raytracer_b200 import host
    img = np.random.default_rng(3).integers(0, 256, size=(24, 31, 3), dtype=np.uint8)
    good = str(tmp_path / "good.png")
    _png_with_filter(good, img, 4)
    assert np.array_equal(host.load_png_layers([good], cache_dir="")[0, :, :, :3], img)
    data = bytearray(open(good, "rb").read())
    at = data.index(b"IDAT") + 4

    def chunk(t, d):
        return struct.pack(">I", len(d)) + t + d + struct.pack(">I", zlib.crc32(t + d))
    raw = b"".join(b"\0" + img[y].tobytes() for y in range(img.shape[0]))
    z = zlib.compress(raw)
    head = b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", 31, 24, 8, 2, 0, 0, 0))
    cases = {}
    flipped = bytearray(data); flipped[at + 20] ^= 0x40
    cases["payload byte flipped (chunk CRC)"] = bytes(flipped)
    cases["zlib stream truncated, CRC of the chunk valid"] = head + chunk(b"IDAT", z[:-9]) + chunk(b"IEND", b"")
    cases["Adler-32 trailer wrong, CRC of the chunk valid"] = head + chunk(b"IDAT", z[:-1] + bytes([z[-1] ^ 1])) + chunk(b"IEND", b"")
    cases["stream holds one row too many"] = head + chunk(b"IDAT", zlib.compress(raw + raw[:94])) + chunk(b"IEND", b"")
    cases["stream holds one row too few"] = head + chunk(b"IDAT", zlib.compress(raw[:-94])) + chunk(b"IEND", b"")
    for what, blob in cases.items():
        p = str(tmp_path / "bad.png")
        open(p, "wb").write(blob)
        with pytest.raises(IOError):
            host.load_png_layers([p], cache_dir="")
        print("rejected:", what)


def test_decoded_texel_cache(pkg, tmp_path):
    """a hit returns the decoder's bytes; a changed source or a damaged entry is a miss."""
    from example_gui_opencl_raytracer_b200 import host
    L = pkg.lib()
    fn = L.rtpng_read_rgb8_to_rgba_cached
    fn.argtypes = [C.c_char_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_char_p, C.POINTER(C.c_int)]
    cache = tmp_path / "cache"
    cache.mkdir()
    src = str(tmp_path / "a.png")
    img = np.random.default_rng(3).integers(0, 256, size=(21, 40, 3), dtype=np.uint8)
    _png_with_filter(src, img, 4)

    def read():
        out = np.zeros((21, 40, 4), np.uint8)
        hit = C.c_int(-1)
        assert fn(src.encode(), out.ctypes.data_as(C.c_void_p), 40, 21, str(cache).encode(), C.byref(hit)) == 0
        return out, hit.value

    a, h0 = read()
    b, h1 = read()
    assert (h0, h1) == (0, 1) and np.array_equal(a, b) and np.array_equal(a[..., :3], img)
    entries = os.listdir(cache)
    assert len(entries) == 1 and entries[0].endswith(".rgba")
    # damaged entry (truncated): miss, rewritten
    e = os.path.join(cache, entries[0])
    blob = open(e, "rb").read()
    assert len(blob) == 40 + 21 * 40 * 4
    open(e, "wb").write(blob[:-5])
    c, h2 = read()
    d, h3 = read()
    assert (h2, h3) == (0, 1) and np.array_equal(c, a) and np.array_equal(d, a)
    # the source changes (other pixels, other mtime/size): miss, new texels
    img2 = 255 - img
    _png_with_filter(src, img2, 1)
    os.utime(src, ns=(1, 1))
    f, h4 = read()
    assert h4 == 0 and np.array_equal(f[..., :3], img2)
    # no cache directory -> plain decode; unwritable cache directory is not an error
    out = np.zeros((21, 40, 4), np.uint8)
    assert fn(src.encode(), out.ctypes.data_as(C.c_void_p), 40, 21, b"", None) == 0
    assert fn(src.encode(), out.ctypes.data_as(C.c_void_p), 40, 21, b"/nonexistent/dir", None) == 0
    assert np.array_equal(out[..., :3], img2)
    # the Python loader takes the same cache
    g = host.load_png_layers([src], cache_dir=str(cache))
    assert np.array_equal(g[0, :, :, :3], img2)


def test_synthetic_scene_is_deterministic(pkg):
    from example_gui_opencl_raytracer_b200 import host
    a = host.sphere_grid_scene(32)
    b = host.sphere_grid_scene(32)
    assert a[0] == b[0] and len(a[0]) == 1024 * 96 and len(a[1]) == 96 and len(a[2]) == 3 * 48


def test_partition_rows_matches_python_twin(pkg):
    from example_gui_opencl_raytracer_b200 import bands
    for H in (1, 7, 250, 2160):
        for band in (1, 4, 8, 100):
            for world in (1, 2, 3, 8):
                tot = 0
                for r in range(world):
                    n = int(pkg.lib().rtb_partition_rows(H, C.byref(pkg.Partition(band, world, r))))
                    assert n == len(bands.band_rows_of(r, world, H, band))
                    tot += n
                assert tot == H


WORKER = r'''
import os, sys
sys.path.insert(0, sys.argv[1])
import numpy as np, torch, torch.distributed as dist
import __graft_entry__ as g
pkg = g.load_package(); ref = g.load_oracle()
from example_gui_opencl_raytracer_b200 import bands
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
scene = ref.render_map_scene(sys.argv[1])
ok = True
for (W, H, band) in ((64, 48, 8), (40, 30, 4), (33, 21, 8)):
    cam = ref.camera(W=W, H=H, **ref.RAYPNG_CAMERA)
    fg = bands.FrameGather(W, H, band, world, rank, torch.device("cpu"))
    rows = bands.band_rows_of(rank, world, H, band)
    full, _ = ref.port().render(scene, cam, W, H, threads=2, want_hit=False)   # CPU stand-in for this rank's render
    mine = full.reshape(H, W)[rows].astype(np.int64).astype(np.uint32).view(np.int32).reshape(-1)
    fg.local[: mine.size] = torch.from_numpy(mine.copy())
    frame = fg.gather()
    if rank == 0:
        ok = ok and np.array_equal(frame.numpy().view(np.uint32), full)
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 3)
'''


def test_two_rank_band_gather_rebuilds_the_frame():
    """world_size 2 over gloo: the gather + interleave must reproduce the one-part frame bit for bit."""
    path = "/tmp/rtb_gloo_worker.py"
    open(path, "w").write(WORKER)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29617", path, ROOT],
                       capture_output=True, timeout=300)
    assert r.returncode == 0, r.stderr.decode()[-2000:]


def test_engine_jump_table_equals_stepping(pkg):
    """rt_engine.cu: build_jump_table -- the GF(2) matrices the pixel-tree kernel multiplies to reach the RNG state of the
    k-th shaded segment, xorshift32^(4L*k)(pixel id) (primitives.cl:111-125; 4L draws per segment, raytracing.cl:96-113):
    applying the matrices of k's set bits to a seed equals stepping the generator 4L*k times."""
    L_ = pkg.lib()
    rng = np.random.default_rng(5)

    def step(x):
        x ^= (x << 13) & 0xFFFFFFFF
        x ^= x >> 17
        x ^= (x << 5) & 0xFFFFFFFF
        return x
    for nl in (0, 1, 3, 7, 16):
        tab = (C.c_uint32 * 480)()
        assert L_.rtb_debug_jump_table(nl, tab) == 0
        tab = np.array(tab, np.uint64).reshape(15, 32)
        for seed in [0, 1, 479999, 0xFFFFFFFF] + [int(v) for v in rng.integers(0, 2**32, 6)]:
            for k in (0, 1, 2, 5, 166, 1023):
                want = seed
                for _ in range(4 * nl * k):
                    want = step(want)
                got = seed
                for j in range(15):
                    if (k >> j) & 1:
                        r = 0
                        for i in range(32):
                            if (got >> i) & 1:
                                r ^= int(tab[j][i])
                        got = r
                assert got == want, (nl, seed, k)


def test_accel_structure_contains_its_spheres(pkg):
    """rt_engine.cu: build_accel_host -- every sphere appears exactly once, with its r*r as one float32
    product and its transparency bit; it lies inside its cluster's bounding sphere and its group's (w = 1.125 R'^2, so
    dist + r <= sqrt(w / 1.125)); padding never hits; a non-finite sphere or a tiny scene is left to brute force."""
    from example_gui_opencl_raytracer_b200 import host
    L_ = pkg.lib()
    for n_side in (4, 9, 32):
        s, _, _, _ = host.sphere_grid_scene(n_side)
        S = len(s) // 96
        nc, ng = C.c_uint32(), C.c_uint32()
        cap_c = (S + 7) // 8
        cl = np.zeros(4 * cap_c, np.float32); rec = np.zeros(32 * cap_c, np.float32)
        idx = np.zeros(8 * cap_c, np.uint32); grp = np.zeros(4 * ((cap_c + 7) // 8), np.float32)
        vp = lambda a: a.ctypes.data_as(C.c_void_p)
        built = L_.rtb_debug_accel_build(s, S, vp(cl), vp(rec), vp(idx), vp(grp), C.byref(nc), C.byref(ng))
        assert built == 1 and nc.value == cap_c and ng.value == (cap_c + 7) // 8
        wire = np.frombuffer(s, np.float32).reshape(S, 24)
        transp = np.frombuffer(s, np.uint32).reshape(S, 24)[:, 16]
        real = idx != 0xFFFFFFFF
        assert sorted((idx[real] & 0x7FFFFFFF).tolist()) == list(range(S))
        rec4 = rec.reshape(-1, 4); cl4 = cl.reshape(-1, 4); grp4 = grp.reshape(-1, 4)
        for j in np.flatnonzero(real):
            i = int(idx[j] & 0x7FFFFFFF)
            assert np.array_equal(rec4[j, :3], wire[i, :3])
            assert rec4[j, 3] == np.float32(wire[i, 4] * wire[i, 4])
            assert bool(idx[j] >> 31) == bool(transp[i])
            for box in (cl4[j // 8], grp4[j // 64]):
                d = np.linalg.norm(wire[i, :3].astype(np.float64) - box[:3].astype(np.float64)) + abs(float(wire[i, 4]))
                assert d <= np.sqrt(float(box[3]) / 1.125), (n_side, i)
        assert np.all(np.isneginf(rec4[~real, 3]))
    bad = bytearray(s); bad[96 * 5: 96 * 5 + 4] = np.float32(np.inf).tobytes()
    assert L_.rtb_debug_accel_build(bytes(bad), S, None, None, None, None, C.byref(nc), C.byref(ng)) == 0
    assert L_.rtb_debug_accel_build(s, 15, None, None, None, None, C.byref(nc), C.byref(ng)) == 0


def test_triangle_hierarchy_contains_its_triangles(pkg):
    """rt_engine.cu: build_tri_accel_host (N4 over extension X1) -- every triangle appears exactly once with its record
    bytes unchanged; its three vertices v0, v0 + e1, v0 + e2 lie inside the box {centre, half extents} of EVERY ancestor
    (cluster and each level above), whose .w carries at least the sum of the half extents; level sizes shrink by 8; padding records are all zero (det == 0: never hit); a non-finite
    vertex or a tiny mesh is left to brute force."""
    from example_gui_opencl_raytracer_b200 import host
    L_ = pkg.lib()
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    for n in (3, 12, 40):
        _, _, _, tris, _ = host.heightfield_mesh_scene(n=n)
        T = len(tris) // 112
        wire = np.frombuffer(tris, np.float32).reshape(T, 28)
        rec = np.zeros((T, 12), np.float32)                      # what tri_prep makes: {v0, v1 - v0, v2 - v0}, one fp32 subtraction each
        rec[:, 0:3] = wire[:, 0:3]
        rec[:, 4:7] = wire[:, 4:7] - wire[:, 0:3]
        rec[:, 8:11] = wire[:, 8:11] - wire[:, 0:3]
        nc, nl, nn = C.c_uint32(), C.c_uint32(), C.c_uint32()
        off = (C.c_uint32 * 6)()
        built = L_.rtb_debug_tri_accel_build(vp(rec), T, None, None, None, C.byref(nc), C.byref(nl), off, C.byref(nn))
        assert built == 1 and nc.value == (T + 7) // 8
        nodes = np.zeros(8 * nn.value, np.float32); srt = np.zeros(12 * 8 * nc.value, np.float32); idx = np.zeros(8 * nc.value, np.uint32)
        assert L_.rtb_debug_tri_accel_build(vp(rec), T, vp(nodes), vp(srt), vp(idx), C.byref(nc), C.byref(nl), off, C.byref(nn)) == 1
        sizes = [nc.value]
        for l in range(1, nl.value):
            sizes.append((sizes[-1] + 7) // 8)
        ntop, top = sizes[-1], nl.value - 1
        assert ntop <= 4 or nl.value == 6
        for l in range(nl.value):                                  # top level first, level l padded to ntop * 8^(top-l) slots
            assert off[l] == ntop * ((8 ** (top - l) - 1) // 7) and sizes[l] <= ntop * 8 ** (top - l)
        assert nn.value == ntop * ((8 ** nl.value - 1) // 7)
        real = idx != 0xFFFFFFFF
        assert sorted(idx[real].tolist()) == list(range(T)) and real[:T].all() and not real[T:].any()
        srt = srt.reshape(-1, 12)
        assert np.array_equal(srt[:T].view(np.uint32), rec[idx[:T]].view(np.uint32))
        assert not srt[T:].any()
        node8 = nodes.reshape(-1, 8).astype(np.float64)
        used = np.concatenate([np.arange(off[l], off[l] + sizes[l]) for l in range(nl.value)])
        assert np.all(node8[used, 7] * 4096 >= node8[used, 4:7].sum(axis=1)) and np.all(node8[used, 4:7] > 0)
        r64 = rec.astype(np.float64)
        for j in range(T):
            i = int(idx[j])
            v0 = r64[i, 0:3]
            for l in range(nl.value):
                box = node8[off[l] + (j // 8 >> (3 * l))]
                for v in (v0, v0 + r64[i, 4:7], v0 + r64[i, 8:11]):
                    assert np.all(np.abs(v - box[0:3]) <= box[4:7]), (n, i, l)
        # the boxes are tight: a cluster's box is no larger than its members' extent (plus rounding)
        ext = np.stack([r64[idx[:T], 0:3], r64[idx[:T], 0:3] + r64[idx[:T], 4:7], r64[idx[:T], 0:3] + r64[idx[:T], 8:11]], axis=1)
        for k in range(nc.value):
            pts = ext[8 * k: 8 * k + 8].reshape(-1, 3)
            assert np.all(2 * node8[off[0] + k, 4:7] <= (pts.max(axis=0) - pts.min(axis=0)) * (1 + 1e-5) + 1e-6)
    bad = rec.copy(); bad[5, 4] = np.inf
    assert L_.rtb_debug_tri_accel_build(vp(bad), T, None, None, None, C.byref(nc), C.byref(nl), off, C.byref(nn)) == 0
    assert L_.rtb_debug_tri_accel_build(vp(rec), 15, None, None, None, C.byref(nc), C.byref(nl), off, C.byref(nn)) == 0
