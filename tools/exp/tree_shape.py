"""Shape of the per-pixel segment trees of render.map at 3840x2160 (oracle/rt_split.c:rto_tree_render, rows of the glass
spheres): segments per pixel vs tree levels vs passes of a 32-lane warp taking one level at a time."""
import ctypes as C, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry
ref = entry.load_oracle()
lib = C.CDLL(os.path.join(ROOT, "oracle", "librt_split.so"))
u8p, u32p, f32p = C.POINTER(C.c_uint8), C.POINTER(C.c_uint32), C.POINTER(C.c_float)
lib.rto_tree_render.restype = C.c_int
lib.rto_tree_render.argtypes = [C.c_int, f32p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_uint32,
                                C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, u8p, C.c_int, C.c_int, C.c_int, u8p,
                                C.c_int, C.c_int, u32p, u32p]
W, H = 3840, 2160
sc = ref.render_map_scene(ROOT)
cam = np.ascontiguousarray(ref.camera(W=W, H=H, **ref.RAYPNG_CAMERA), np.float32)
out = np.zeros(W * H, np.uint32); shape = np.zeros(W * H, np.uint32)
y0, y1 = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1000, 1400)
assert lib.rto_tree_render(0, cam.ctypes.data_as(f32p), W, H, y0, y1, sc.spheres.ctypes.data, sc.S, sc.planes.ctypes.data, sc.P,
                           sc.lights.ctypes.data, sc.L, sc.tex.ctypes.data_as(u8p), sc.tex.shape[2], sc.tex.shape[1], sc.tex.shape[0],
                           sc.sky.ctypes.data_as(u8p), sc.sky.shape[2], sc.sky.shape[1], out.ctypes.data_as(u32p), shape.ctypes.data_as(u32p)) == 0
sh = shape[y0 * W:y1 * W]
nodes, levels, steps = sh & 0xffff, (sh >> 16) & 0xff, sh >> 24
order = np.argsort(nodes)[::-1]
print("rows %d..%d: segments per pixel max %d, mean %.2f; levels max %d; warp passes max %d" % (y0, y1, nodes.max(), nodes.mean(), levels.max(), steps.max()))
for q in (0, 10, 100, 1000, 10000):
    i = order[q]
    print("  #%d heaviest pixel: %d segments, %d levels, %d warp passes (chain %.1fx shorter)" % (q, nodes[i], levels[i], steps[i], nodes[i] / max(1, steps[i])))
