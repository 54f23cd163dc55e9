"""Host-side mirror of the reference's caller code for the hot path (no rendering here).

  * load_map / dump_map  -- the .map scene archive of src/cpu_obj.c:51-101 (count byte + raw structs, x3)
  * material / sphere / plane / light packers for the wire structs of src/cpu_obj.h:10-48, and the four
    material presets of src/cpu_obj.c:6-49
  * camera()             -- rinit_camera + rgen_perspective (src/cpu_ray.c:8-35,42-106) in float32
  * load_png_layers()    -- what cl_wrap_load_images builds (src/opencl_wrap.c:189-349), decoded by the
                            library's own PNG codec (csrc/png_io.c)
  * synthetic scenes for BASELINE.json configs 3 and 4' 
"""
import ctypes as C
import os
import struct

import numpy as np

from . import MapArchive, lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SPHERE_BYTES, PLANE_BYTES, LIGHT_BYTES = 96, 96, 48


# ---- wire structs (src/cpu_obj.h:10-48; 16-byte aligned float3, cl_bool = 4 bytes) ---------------
def material(rgb=(1, 1, 1), ambient=0.3, diffuse=0.2, specular=0.6, shininess=50, transperent=0, dielectric=0, n=1.0,
             reflectivity=0.0, texture_id=-1, texture_scale=0.0):
    return struct.pack("<3f4x3fI2I2fif8x", *rgb, ambient, diffuse, specular, int(shininess), int(transperent),
                       int(dielectric), n, reflectivity, int(texture_id), texture_scale)


# src/cpu_obj.c:6-49
STONE = dict(rgb=(1, 1, 1), ambient=.4, diffuse=.2, specular=.6, shininess=50, transperent=0, dielectric=1, n=1.57, reflectivity=0.0)
PLASTIC = dict(rgb=(1, 1, 1), ambient=.3, diffuse=.2, specular=.6, shininess=50, transperent=0, dielectric=0, n=1.4, reflectivity=0.1)
MIRROR = dict(rgb=(.2, .2, .2), ambient=.3, diffuse=0., specular=.6, shininess=100, transperent=0, dielectric=1, n=1.0, reflectivity=1.0)
GLASS = dict(rgb=(0, 0, 0), ambient=.1, diffuse=0., specular=0., shininess=20, transperent=1, dielectric=1, n=1.52, reflectivity=0.04)


def sphere(origin, radius, mat):
    return struct.pack("<3f4xf12x", *origin, radius) + material(**mat)


def plane(normal, point, mat):
    return struct.pack("<3f4x3f4x", *normal, *point) + material(**mat)


def light(origin, radius, intensity, rgb):
    return struct.pack("<3f4xff8x3f4x", *origin, radius, intensity, *rgb)


TRIANGLE_BYTES = 112


def triangle(v0, v1, v2, mat):
    """EXTENSION X1 (SURVEY.md 8(a'), no reference counterpart): rtriangle = float4 v0, v1, v2 + rmaterial."""
    return struct.pack("<3f4x3f4x3f4x", *v0, *v1, *v2) + material(**mat)


def load_map(path):
    """extract_robj (src/cpu_obj.c:76-101) -> (spheres, planes, lights) as bytes."""
    d = open(path, "rb").read()
    o, out = 0, []
    for sz in (SPHERE_BYTES, PLANE_BYTES, LIGHT_BYTES):
        n = d[o]
        o += 1
        out.append(d[o:o + n * sz])
        o += n * sz
    if o != len(d):
        raise ValueError("%s: trailing or missing bytes" % path)
    return tuple(out)


def dump_map(path, spheres, planes, lights):
    """dump_robj (src/cpu_obj.c:51-74); counts are one byte each."""
    ns, np_, nl = len(spheres) // SPHERE_BYTES, len(planes) // PLANE_BYTES, len(lights) // LIGHT_BYTES
    if max(ns, np_, nl) > 255:
        raise ValueError(".map counts are a single byte (max 255 per type)")
    with open(path, "wb") as f:
        f.write(bytes([ns]) + spheres + bytes([np_]) + planes + bytes([nl]) + lights)


# ---- camera (src/cpu_ray.c) -----------------------------------------------------------------------
def camera(origin, look, fov, focal, W, H, normalize_look=True):
    """The 14 floats raygen takes (raygen.cl:5-7): corner, origin, up, right, w_factor, h_factor.
    float32 arithmetic with the C promotions of cpu_ray.c (sqrt/tan/pi steps go through double)."""
    f = np.float32
    o = [f(v) for v in origin]
    d = [f(v) for v in look]
    if normalize_look:  # cpu_ray.c:8-19
        s = f(f(f(d[0] * d[0]) + f(d[1] * d[1])) + f(d[2] * d[2]))
        num = f(1.0 / np.sqrt(np.float64(s)))
        d = [f(v * num) for v in d]
    fov, focal = f(fov), f(focal)
    half = f(np.float64(f(fov / f(360.0))) * np.pi)       # :65
    aspect = f(f(H) / f(W))                                 # :66
    fov_tan = f(np.tan(np.float64(half)))                   # :67
    iw = f(f(fov_tan * focal) * f(2))                       # :69
    ih = f(aspect * iw)                                     # :70
    wf, hf = f(iw / f(W)), f(ih / f(H))                     # :72-73
    fw = [f(-v) for v in d]                                 # :76-79
    top = [f(0), f(1), f(0)]

    def cross(a, b):
        return [f(f(a[1] * b[2]) - f(a[2] * b[1])), f(f(a[2] * b[0]) - f(a[0] * b[2])),
                f(f(a[0] * b[1]) - f(a[1] * b[0]))]
    right = cross(top, fw)                                  # :82-86
    up = cross(fw, right)                                   # :87-91
    center = [f(f(-v) * focal) for v in fw]                 # :94-98
    corner = [f(f(center[i] - f(f(right[i] * iw) / f(2))) + f(f(up[i] * ih) / f(2))) for i in range(3)]  # :101-105
    return np.array(corner + o + up + right + [wf, hf], np.float32)


RAYPNG_CAMERA = dict(origin=(0.8, 2.5, -8.0), look=(0.2, 0.0, 1.0), fov=90.0, focal=1.0)  # raypng.c:17-21


def orbit_camera(frame, frames, W, H, centre=(0.5, 1.0, 1.0), radius=9.5, height=2.5):
    """Camera on a circle around the scene, looking at `centre` (the config-5 sweep)."""
    a = 2.0 * np.pi * frame / frames - np.pi / 2
    o = (centre[0] + radius * np.cos(a), height, centre[2] + radius * np.sin(a))
    look = (centre[0] - o[0], centre[1] - o[1], centre[2] - o[2])
    return camera(o, look, 90.0, 1.0, W, H)


# ---- image arrays ---------------------------------------------------------------------------------
def load_png_layers(paths, cache_dir=None):
    """-> (layers, h, w, 4) uint8, A = 255; every file must be 8-bit RGB of equal size.
    cache_dir (default: $RTB_IMAGE_CACHE, like cl_wrap_load_images): decoded texels are kept there."""
    L = lib()
    if cache_dir is None:
        cache_dir = os.environ.get("RTB_IMAGE_CACHE", "")
    L.rtpng_read_rgb8_to_rgba_cached.argtypes = [C.c_char_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_char_p, C.POINTER(C.c_int)]
    L.rtpng_read_rgb8_to_rgba_cached.restype = C.c_int
    L.rtpng_info.argtypes = [C.c_char_p] + [C.POINTER(C.c_uint32)] * 2 + [C.POINTER(C.c_int)] * 3
    L.rtpng_info.restype = C.c_int
    L.rtpng_read_rgb8_to_rgba.argtypes = [C.c_char_p, C.c_void_p, C.c_uint32, C.c_uint32]
    L.rtpng_read_rgb8_to_rgba.restype = C.c_int
    out = None
    for i, p in enumerate(paths):
        w, h = C.c_uint32(), C.c_uint32()
        depth, ctype, il = C.c_int(), C.c_int(), C.c_int()
        rc = L.rtpng_info(p.encode(), C.byref(w), C.byref(h), C.byref(depth), C.byref(ctype), C.byref(il))
        if rc != 0:
            raise IOError("%s: cannot read PNG header (%d)" % (p, rc))
        if depth.value != 8 or ctype.value != 2:
            raise ValueError('"%s" must have a depth of 8 bits and be RGB' % p)
        if out is None:
            out = np.empty((len(paths), h.value, w.value, 4), np.uint8)
        elif (h.value, w.value) != out.shape[1:3]:
            raise ValueError("All images must have same dimensions")
        rc = L.rtpng_read_rgb8_to_rgba_cached(p.encode(), out[i].ctypes.data_as(C.c_void_p), w.value, h.value,
                                              cache_dir.encode(), None)
        if rc != 0:
            raise IOError("%s: cannot decode (%d)" % (p, rc))
    return out


TEXTURE_FILES = ("assets/cobblestone.png", "assets/sand.png", "assets/check.png", "assets/grass.png")  # raypng.c:74-78
SKYBOX_FILE = "assets/bg/stormydays.png"                                                              # raypng.c:80-81


def default_images(root=ROOT):
    return (load_png_layers([os.path.join(root, p) for p in TEXTURE_FILES]),
            load_png_layers([os.path.join(root, SKYBOX_FILE)]))


# ---- synthetic scenes ---------------------------------------------------------
def render_map_lights():
    return (light((-2, 3, 2), .1, 8.0, (0, 1, 0)) + light((2, 1.5, .2), .1, 50.3, (1, 1, 1)) +
            light((1, 4, 3), .1, 20.5, (0, 0, 1)))


def sphere_grid_scene(n_side=32, seed=0x5EED1024, spacing=1.0):
    """Config 3: n_side^2 spheres on a jittered grid over the check-textured ground plane, material by i mod 4
    (plastic / glass / mirror-like / stone), 3 lights as in render.map.  Deterministic (own LCG)."""
    state = [seed & 0xFFFFFFFF]

    def rnd():
        state[0] = (1664525 * state[0] + 1013904223) & 0xFFFFFFFF
        return state[0] / 4294967296.0
    sp = []
    half = 0.5 * (n_side - 1) * spacing
    for i in range(n_side * n_side):
        gx, gz = i % n_side, i // n_side
        r = 0.2 + 0.25 * rnd()
        x = gx * spacing - half + (rnd() - 0.5) * 0.3 * spacing
        z = gz * spacing - half + (rnd() - 0.5) * 0.3 * spacing + half * 0.5
        rgb = (rnd(), rnd(), rnd())
        k = i % 4
        if k == 0:
            m = dict(PLASTIC, rgb=rgb)
        elif k == 1:
            m = dict(GLASS)
        elif k == 2:
            m = dict(MIRROR, rgb=rgb, reflectivity=1.0)
        else:
            m = dict(STONE, rgb=rgb)
        sp.append(sphere((x, r, z), r, m))
    planes = plane((0, 1, 0), (0, 0, 0), dict(STONE, rgb=(0, 0, 0), texture_id=2, texture_scale=100.0))
    lights = (light((-6, 8, -4), .3, 260.0, (1, 1, 1)) + light((8, 6, 6), .3, 180.0, (1, .9, .8)) +
              light((0, 10, 14), .3, 200.0, (.8, .9, 1)))
    cam = dict(origin=(0.0, 9.0, -half - 6.0), look=(0.0, -0.55, 1.0), fov=90.0, focal=1.0)
    return b"".join(sp), planes, lights, cam


def heightfield_mesh_scene(n=158, seed=0x7121A461, size=16.0, amplitude=0.9):
    """Config 4 (extension X1): a seeded n x n heightfield = 2*n*n triangles (158 -> 49,928) over the textured
    ground plane, 3 lights.  Returns (spheres, planes, lights, triangles, camera kwargs)."""
    state = [seed & 0xFFFFFFFF]

    def rnd():
        state[0] = (1664525 * state[0] + 1013904223) & 0xFFFFFFFF
        return state[0] / 4294967296.0
    k1, k2, p1, p2 = 0.9 + rnd(), 0.7 + rnd(), 6.28 * rnd(), 6.28 * rnd()
    step = size / n

    def hgt(i, j):
        x, z = i * step - size / 2, j * step - size / 2
        return 0.6 + amplitude * (0.5 * np.sin(k1 * x + p1) * np.cos(k2 * z + p2) + 0.5 * np.sin(0.37 * x * z + p1))
    tris = []
    for j in range(n):
        for i in range(n):
            x0, x1 = i * step - size / 2, (i + 1) * step - size / 2
            z0, z1 = j * step - size / 2 + 4.0, (j + 1) * step - size / 2 + 4.0
            a, b, c, d = (x0, hgt(i, j), z0), (x1, hgt(i + 1, j), z0), (x0, hgt(i, j + 1), z1), (x1, hgt(i + 1, j + 1), z1)
            kind = (i + j) % 8
            if kind == 0:
                m = dict(MIRROR, rgb=(.25, .25, .3), reflectivity=0.6)
            else:
                m = dict(PLASTIC, rgb=(0.3 + 0.6 * rnd(), 0.4 + 0.5 * rnd(), 0.3 + 0.6 * rnd()))
            tris.append(triangle(a, c, b, m))   # wound so that e1 x e2 points up (+y)
            tris.append(triangle(b, c, d, m))
    planes = plane((0, 1, 0), (0, -0.6, 0), dict(STONE, rgb=(0, 0, 0), texture_id=2, texture_scale=100.0))
    lights = (light((-6, 8, -4), .3, 260.0, (1, 1, 1)) + light((8, 6, 6), .3, 180.0, (1, .9, .8)) +
              light((0, 10, 14), .3, 200.0, (.8, .9, 1)))
    spheres = sphere((0.0, 2.6, 4.0), 0.9, dict(GLASS)) + sphere((-3.0, 2.2, 6.0), 0.7, dict(MIRROR, reflectivity=1.0))
    cam = dict(origin=(0.0, 6.5, -9.0), look=(0.0, -0.45, 1.0), fov=90.0, focal=1.0)
    return spheres, planes, lights, b"".join(tris), cam


# ---- scene archives through the library's own reader/writer (csrc/map_io.c) ----
def load_map_any(path):
    """v1 (the reference's .map) or v2 ("RTM2", 32-bit counts + triangles) -> (spheres, planes, lights, triangles, version)."""
    m = MapArchive()
    rc = lib().rtb_map_load(path.encode(), C.byref(m))
    if rc != 0:
        raise IOError("%s: rtb_map_load failed (%d)" % (path, rc))
    try:
        out = [C.string_at(p, n * sz) if n else b"" for p, n, sz in
               ((m.spheres, m.S, 96), (m.planes, m.P, 96), (m.lights, m.L, 48), (m.triangles, m.T, 112))]
        return out[0], out[1], out[2], out[3], m.version
    finally:
        lib().rtb_map_free(C.byref(m))


def save_map(path, spheres, planes, lights, triangles=b"", version=2):
    bufs = [C.create_string_buffer(bytes(b), len(b)) if len(b) else None for b in (spheres, planes, lights, triangles)]
    m = MapArchive()
    m.spheres, m.planes, m.lights, m.triangles = [C.cast(b, C.c_void_p) if b is not None else None for b in bufs]
    m.S, m.P, m.L, m.T = len(spheres) // 96, len(planes) // 96, len(lights) // 48, len(triangles) // 112
    rc = lib().rtb_map_save(path.encode(), C.byref(m), version)
    if rc != 0:
        raise IOError("%s: rtb_map_save failed (%d)" % (path, rc))
