"""example-gui-opencl-raytracer_b200 -- Python view of the B200 raytracer backend.

The product is the C-ABI shared library ``librtb200.so`` in this directory (built by the top-level
Makefile from ``csrc/``): the render engine of ``include/rtb200.h`` plus the ``cl_wrap_*`` drop-in of
``include/cuda_wrap.h`` that replaces the reference's ``src/opencl_wrap.c``.  This module only binds
that library with ctypes for the tests and ``bench.py``; it contains no rendering code and there is
NO CPU fallback: importing works anywhere (so the symbol table can be checked without a GPU), but any
compute call needs the library and a B200 and raises otherwise.

The directory name is not a Python identifier; load it with ``__graft_entry__.load_package()``.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RTB_LIB") or os.path.join(HERE, "librtb200.so")  # RTB_LIB: A/B builds only

HIT_LIGHT = 0x80000000
COUNTER_NAMES = ("segments", "shadow_rays", "sphere_tests", "sphere_dpos", "plane_tests", "texels", "shadow_skipped",
                 "tri_tests")
NCOUNTERS = 8

# Every symbol include/rtb200.h and include/cuda_wrap.h declare (checked by tests/test_host_cpu.py::test_library_exports_every_declared_symbol).
ENGINE_SYMBOLS = (
    "rtb_last_error", "rtb_device_count", "rtb_create", "rtb_destroy", "rtb_upload_scene", "rtb_bind_scene_device",
    "rtb_upload_triangles", "rtb_bind_triangles_device", "rtb_upload_images", "rtb_bind_images_device", "rtb_set_camera", "rtb_partition_rows", "rtb_render",
    "rtb_render_counted", "rtb_sync", "rtb_read_frame", "rtb_read_hit", "rtb_frame_ptr", "rtb_hit_ptr",
    "rtb_last_render_ms", "rtb_last_kernel_ms", "rtb_stream", "rtb_set_option", "rtb_selftest_arith", "rtb_dev_alloc", "rtb_dev_free", "rtb_dev_write", "rtb_dev_read",
    "rtb_host_register", "rtb_host_unregister", "rtb_host_map", "rtb_set_mirror", "rtb_set_mirror_frame", "rtb_set_completion", "rtb_wait_flags", "rtb_wait_timeouts", "rtb_ipc_export", "rtb_ipc_open", "rtb_ipc_close", "rtb_assemble_bands", "rtb_last_launch", "rtb_render_noalloc", "rtb_debug_counters", "rtb_debug_jump_table", "rtb_debug_accel_build", "rtb_debug_tri_accel_build", "rtb_fp32_peak_tflops", "rtb_map_load", "rtb_map_save", "rtb_map_free",
)
WRAP_SYMBOLS = (
    "cl_wrap_init", "cl_wrap_load_global_data", "cl_wrap_load_single_data", "cl_wrap_load_images", "cl_wrap_output",
    "cl_wrap_release", "cuda_wrap_engine", "cuda_wrap_last_kernel_ms", "cuda_wrap_readback_stats",
)


class RtbError(RuntimeError):
    pass


class Partition(C.Structure):
    """rtb_partition: cyclic row bands (include/rtb200.h)."""
    _fields_ = [("band_rows", C.c_uint32), ("parts", C.c_uint32), ("part", C.c_uint32)]


class MapArchive(C.Structure):
    """rtb_map (include/rtb200.h)."""
    _fields_ = [("spheres", C.c_void_p), ("planes", C.c_void_p), ("lights", C.c_void_p), ("triangles", C.c_void_p),
                ("S", C.c_uint32), ("P", C.c_uint32), ("L", C.c_uint32), ("T", C.c_uint32), ("version", C.c_int)]


class LaunchInfo(C.Structure):
    _fields_ = [("grid", C.c_uint32), ("block", C.c_uint32), ("smem_bytes", C.c_uint32), ("tiles", C.c_uint32),
                ("resident_scene", C.c_uint32), ("tree_grid", C.c_uint32)]


_lib = None


def lib():
    """The loaded library.  Raises if it has not been built: there is nothing to fall back to."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RtbError("%s is missing: run `make` at the repo root (or __graft_entry__.build())" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        vp, u32, i, sz = C.c_void_p, C.c_uint32, C.c_int, C.c_size_t
        pp = C.POINTER(Partition)
        sig = {
            "rtb_last_error": (C.c_char_p, []),
            "rtb_device_count": (i, []),
            "rtb_create": (i, [C.POINTER(vp), i]),
            "rtb_destroy": (None, [vp]),
            "rtb_upload_scene": (i, [vp, vp, u32, vp, u32, vp, u32]),
            "rtb_bind_scene_device": (i, [vp, vp, u32, vp, u32, vp, u32]),
            "rtb_upload_triangles": (i, [vp, vp, u32]),
            "rtb_bind_triangles_device": (i, [vp, vp, u32]),
            "rtb_upload_images": (i, [vp, i, vp, u32, u32, u32]),
            "rtb_bind_images_device": (i, [vp, i, vp, u32, u32, u32]),
            "rtb_set_camera": (i, [vp, C.POINTER(C.c_float)]),
            "rtb_partition_rows": (u32, [u32, pp]),
            "rtb_render": (i, [vp, u32, u32, u32, pp, vp, vp, i]),
            "rtb_render_counted": (i, [vp, u32, u32, u32, pp, vp, vp, i, i, C.POINTER(C.c_uint64)]),
            "rtb_render_noalloc": (i, [vp, C.POINTER(C.c_float), u32, u32, u32, pp, vp, vp]),
            "rtb_set_option": (i, [vp, C.c_char_p, i]),
            "rtb_selftest_arith": (i, [vp, u32, u32, C.POINTER(u32)]),
            "rtb_last_kernel_ms": (C.c_float, [vp]),
            "rtb_sync": (i, [vp]),
            "rtb_read_frame": (i, [vp, vp, sz]),
            "rtb_read_hit": (i, [vp, vp, sz]),
            "rtb_frame_ptr": (vp, [vp]),
            "rtb_hit_ptr": (vp, [vp]),
            "rtb_last_render_ms": (C.c_float, [vp]),
            "rtb_stream": (vp, [vp]),
            "rtb_dev_alloc": (i, [vp, C.POINTER(vp), sz]),
            "rtb_dev_free": (i, [vp, vp]),
            "rtb_dev_write": (i, [vp, vp, vp, sz]),
            "rtb_dev_read": (i, [vp, vp, vp, sz]),
            "rtb_host_register": (i, [vp, sz]),
            "rtb_host_unregister": (i, [vp]),
            "rtb_host_map": (i, [vp, sz, C.POINTER(vp), C.POINTER(i)]),
            "rtb_set_mirror": (i, [vp, vp]),
            "rtb_set_mirror_frame": (i, [vp, vp]),
            "rtb_set_completion": (i, [vp, vp, u32]),
            "rtb_wait_flags": (i, [vp, vp, u32, u32, u32]),
            "rtb_wait_timeouts": (i, [vp, C.POINTER(u32)]),
            "rtb_ipc_export": (i, [vp, vp, C.c_char_p]),
            "rtb_ipc_open": (i, [vp, C.c_char_p, C.POINTER(vp)]),
            "rtb_ipc_close": (i, [vp, vp]),
            "rtb_assemble_bands": (i, [vp, vp, sz, vp, u32, u32, u32, u32]),
            "rtb_last_launch": (i, [vp, C.POINTER(LaunchInfo)]),
            "rtb_debug_counters": (i, [vp, C.POINTER(C.c_uint64)]),
            "rtb_debug_jump_table": (i, [u32, C.POINTER(u32)]),
            "rtb_debug_accel_build": (i, [vp, u32, vp, vp, vp, vp, C.POINTER(u32), C.POINTER(u32)]),
            "rtb_debug_tri_accel_build": (i, [vp, u32, vp, vp, vp, C.POINTER(u32), C.POINTER(u32), C.POINTER(u32), C.POINTER(u32)]),
            "rtb_fp32_peak_tflops": (C.c_double, [vp, i]),
            "rtb_map_load": (i, [C.c_char_p, C.POINTER(MapArchive)]),
            "rtb_map_save": (i, [C.c_char_p, C.POINTER(MapArchive), i]),
            "rtb_map_free": (None, [C.POINTER(MapArchive)]),
        }
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def _check(rc, what):
    if rc != 0:
        raise RtbError("%s failed (%d): %s" % (what, rc, lib().rtb_last_error().decode()))


def _host_ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None and a.size else None


class Engine:
    """One render engine on one GPU (rtb_engine)."""

    def __init__(self, device=0):
        self._h = C.c_void_p()
        _check(lib().rtb_create(C.byref(self._h), int(device)), "rtb_create")
        self.device = int(device)
        self.S = self.P = self.L = 0

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            lib().rtb_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- inputs -------------------------------------------------------------------------------
    def upload_scene(self, spheres, planes, lights):
        """Wire-format arrays as bytes / uint8 ndarrays (rsphere 96 B, rplane 96 B, rlight 48 B)."""
        s = np.ascontiguousarray(np.frombuffer(bytes(spheres), np.uint8)) if len(spheres) else np.zeros(0, np.uint8)
        p = np.ascontiguousarray(np.frombuffer(bytes(planes), np.uint8)) if len(planes) else np.zeros(0, np.uint8)
        l = np.ascontiguousarray(np.frombuffer(bytes(lights), np.uint8)) if len(lights) else np.zeros(0, np.uint8)
        if s.size % 96 or p.size % 96 or l.size % 48:
            raise ValueError("scene arrays are not whole wire records")
        self.S, self.P, self.L = s.size // 96, p.size // 96, l.size // 48
        _check(lib().rtb_upload_scene(self._h, _host_ptr(s), self.S, _host_ptr(p), self.P, _host_ptr(l), self.L),
               "rtb_upload_scene")

    def upload_triangles(self, tris):
        """EXTENSION X1 (no reference counterpart): rtriangle records, 112 B each; b"" removes them."""
        t = np.ascontiguousarray(np.frombuffer(bytes(tris), np.uint8)) if len(tris) else np.zeros(0, np.uint8)
        if t.size % 112:
            raise ValueError("triangle array is not whole 112-byte records")
        self.T = t.size // 112
        _check(lib().rtb_upload_triangles(self._h, _host_ptr(t), self.T), "rtb_upload_triangles")

    def bind_scene_device(self, d_spheres, S, d_planes, P, d_lights, L):
        self.S, self.P, self.L = int(S), int(P), int(L)
        _check(lib().rtb_bind_scene_device(self._h, d_spheres, S, d_planes, P, d_lights, L), "rtb_bind_scene_device")

    def upload_images(self, which, rgba):
        a = np.ascontiguousarray(rgba, np.uint8)
        if a.ndim != 4 or a.shape[3] != 4:
            raise ValueError("image array must be (layers, h, w, 4) uint8")
        _check(lib().rtb_upload_images(self._h, int(which), _host_ptr(a), a.shape[2], a.shape[1], a.shape[0]),
               "rtb_upload_images")

    def set_camera(self, cam):
        c = np.ascontiguousarray(cam, np.float32)
        if c.size != 14:
            raise ValueError("camera is 14 floats: corner, origin, up, right, w_factor, h_factor")
        _check(lib().rtb_set_camera(self._h, c.ctypes.data_as(C.POINTER(C.c_float))), "rtb_set_camera")

    # ---- launches -----------------------------------------------------------------------------
    @staticmethod
    def partition_rows(H, part):
        return H if part is None else int(lib().rtb_partition_rows(H, C.byref(part)))

    def render(self, W, H, total=None, part=None, d_out=None, d_hit=None, want_hit=False):
        """Asynchronous launch on the engine's stream."""
        total = W * H if total is None else total
        _check(lib().rtb_render(self._h, W, H, total, C.byref(part) if part is not None else None, d_out, d_hit,
                                int(bool(want_hit))), "rtb_render")

    def render_counted(self, W, H, total=None, part=None, d_out=None, d_hit=None, want_hit=False, executed=False):
        """executed=False: literal trace, the reference's own work counts; True: what the production trace does."""
        total = W * H if total is None else total
        cnt = (C.c_uint64 * NCOUNTERS)()
        _check(lib().rtb_render_counted(self._h, W, H, total, C.byref(part) if part is not None else None, d_out,
                                        d_hit, int(bool(want_hit)), int(bool(executed)), cnt), "rtb_render_counted")
        c = {k: int(cnt[i]) for i, k in enumerate(COUNTER_NAMES)}
        c["rays"] = c["segments"] + c["shadow_rays"]
        c["flop"] = algorithmic_flop(c)
        return c

    def render_noalloc(self, W, H, d_out, cam=None, total=None, part=None, d_hit=None):
        """rtb_render_noalloc: one frame into the caller's device buffers without any allocation (static device scratch;
        at most 262,144 tiles per call).  Asynchronous on the engine's stream."""
        total = W * H if total is None else total
        c = None
        if cam is not None:
            c = np.ascontiguousarray(cam, np.float32)
            if c.size != 14:
                raise ValueError("camera is 14 floats")
            c = c.ctypes.data_as(C.POINTER(C.c_float))
        _check(lib().rtb_render_noalloc(self._h, c, W, H, total, C.byref(part) if part is not None else None, d_out, d_hit),
               "rtb_render_noalloc")

    def set_mirror_frame(self, d_frame):
        """The next render also stores every word at d_frame[y*W + x] (a full frame, typically a peer GPU's)."""
        _check(lib().rtb_set_mirror_frame(self._h, C.c_void_p(d_frame)), "rtb_set_mirror_frame")

    def set_completion(self, d_flag, value):
        """The next render's last CTA release-stores `value` into the device-visible word at d_flag."""
        _check(lib().rtb_set_completion(self._h, C.c_void_p(d_flag), int(value) & 0xFFFFFFFF), "rtb_set_completion")

    def wait_flags(self, d_flags, stride_words, n, value):
        """Enqueue the wait kernel: returns (on the stream) when n flags have reached `value`."""
        _check(lib().rtb_wait_flags(self._h, C.c_void_p(d_flags), stride_words, n, int(value) & 0xFFFFFFFF), "rtb_wait_flags")

    def wait_timeouts(self):
        out = C.c_uint32()
        _check(lib().rtb_wait_timeouts(self._h, C.byref(out)), "rtb_wait_timeouts")
        return int(out.value)

    @staticmethod
    def host_map(host_ptr, nbytes):
        """Device-visible alias of a host buffer (page-locks it if needed) -> (device view, registered_here)."""
        d, r = C.c_void_p(), C.c_int()
        _check(lib().rtb_host_map(C.c_void_p(host_ptr), nbytes, C.byref(d), C.byref(r)), "rtb_host_map")
        return d.value, bool(r.value)

    @staticmethod
    def host_unregister(host_ptr):
        lib().rtb_host_unregister(C.c_void_p(host_ptr))

    def ipc_export(self, dptr):
        h = C.create_string_buffer(64)
        _check(lib().rtb_ipc_export(self._h, dptr, h), "rtb_ipc_export")
        return h.raw

    def ipc_open(self, handle):
        p = C.c_void_p()
        _check(lib().rtb_ipc_open(self._h, C.create_string_buffer(bytes(handle), 64), C.byref(p)), "rtb_ipc_open")
        return p

    def ipc_close(self, p):
        _check(lib().rtb_ipc_close(self._h, p), "rtb_ipc_close")

    def assemble_bands(self, d_compact, stride_words, d_frame, W, H, band_rows, parts):
        """Gathering rank: compact band buffers -> row-major frame (one kernel on the engine's stream)."""
        _check(lib().rtb_assemble_bands(self._h, C.c_void_p(d_compact), stride_words, C.c_void_p(d_frame), W, H, band_rows,
                                        parts), "rtb_assemble_bands")

    def sync(self):
        _check(lib().rtb_sync(self._h), "rtb_sync")

    def read_frame(self, npixels):
        out = np.empty(npixels, np.uint32)
        if npixels == 0:
            return out
        _check(lib().rtb_read_frame(self._h, _host_ptr(out), out.nbytes), "rtb_read_frame")
        return out

    def read_hit(self, npixels):
        out = np.empty(npixels, np.uint32)
        if npixels == 0:
            return out
        _check(lib().rtb_read_hit(self._h, _host_ptr(out), out.nbytes), "rtb_read_hit")
        return out

    def render_to_host(self, W, H, part=None, want_hit=True, total=None):
        """Convenience: render into the engine's buffers and read them back -> (rgb, hit or None)."""
        rows = self.partition_rows(H, part)
        self.render(W, H, total=total, part=part, want_hit=want_hit)
        rgb = self.read_frame(rows * W)
        return rgb, (self.read_hit(rows * W) if want_hit else None)

    def last_render_ms(self):
        ms = float(lib().rtb_last_render_ms(self._h))
        if ms < 0:
            raise RtbError("no timed render")
        return ms

    def last_kernel_ms(self):
        ms = float(lib().rtb_last_kernel_ms(self._h))
        if ms < 0:
            raise RtbError("no timed render")
        return ms

    def set_option(self, name, value):
        _check(lib().rtb_set_option(self._h, name.encode(), int(value)), "rtb_set_option")

    def selftest_arith(self, seed=1, per_thread=2000):
        out = (C.c_uint32 * 3)()
        _check(lib().rtb_selftest_arith(self._h, seed, per_thread, out), "rtb_selftest_arith")
        return {"div_mismatch": int(out[0]), "normalize_mismatch": int(out[1]), "sincos_max_dev_2p24": int(out[2])}

    def debug_counters(self):
        cnt = (C.c_uint64 * NCOUNTERS)()
        _check(lib().rtb_debug_counters(self._h, cnt), "rtb_debug_counters")
        return [int(v) for v in cnt]

    def last_launch(self):
        info = LaunchInfo()
        _check(lib().rtb_last_launch(self._h, C.byref(info)), "rtb_last_launch")
        return {k: int(getattr(info, k)) for k, _ in LaunchInfo._fields_}

    def stream(self):
        return lib().rtb_stream(self._h)

    def frame_ptr(self):
        return lib().rtb_frame_ptr(self._h)

    def fp32_peak_tflops(self, iters=4096):
        v = float(lib().rtb_fp32_peak_tflops(self._h, iters))
        if v <= 0:
            raise RtbError("fp32 calibration kernel failed")
        return v

    # raw device memory (for callers that keep the scene resident in HBM themselves)
    def dev_alloc(self, nbytes):
        p = C.c_void_p()
        _check(lib().rtb_dev_alloc(self._h, C.byref(p), nbytes), "rtb_dev_alloc")
        return p

    def dev_free(self, p):
        _check(lib().rtb_dev_free(self._h, p), "rtb_dev_free")

    def dev_write(self, p, host):
        a = np.ascontiguousarray(host)
        _check(lib().rtb_dev_write(self._h, p, _host_ptr(a.view(np.uint8).reshape(-1)), a.nbytes), "rtb_dev_write")

    def dev_read(self, p, nbytes):
        out = np.empty(nbytes, np.uint8)
        _check(lib().rtb_dev_read(self._h, _host_ptr(out), p, nbytes), "rtb_dev_read")
        return out


def algorithmic_flop(c):
    """21 flop per sphere test (+5 when D >= 0), 14 per plane test, 37 per triangle test
    (extension X1); nothing else counted."""
    return 21 * c["sphere_tests"] + 5 * c["sphere_dpos"] + 14 * c["plane_tests"] + 37 * c.get("tri_tests", 0)


def assemble_bands(parts_rows, W, H, band_rows):
    """Interleave the compact per-part buffers of a cyclic row-band split back into one H x W frame.
    parts_rows[r] is part r's (rows_r * W) uint32 array.  Host-side twin of the multi-GPU gather."""
    parts = len(parts_rows)
    out = np.empty((H, W), np.uint32)
    cursors = [0] * parts
    nb = (H + band_rows - 1) // band_rows
    for b in range(nb):
        r = b % parts
        y0, y1 = b * band_rows, min(H, (b + 1) * band_rows)
        n = y1 - y0
        out[y0:y1] = np.asarray(parts_rows[r]).reshape(-1, W)[cursors[r]:cursors[r] + n]
        cursors[r] += n
    return out.reshape(-1)


# ---- the drop-in boundary, bound the way a C caller sees it (include/cuda_wrap.h) -------------
MAX_KERNELS, MAX_BUFFERS = 16, 32


class ClWrap(C.Structure):
    """cl_wrap (reference src/opencl_wrap.h:11-23; include/cuda_wrap.h)."""
    _fields_ = [("context", C.c_void_p), ("device", C.c_void_p), ("program", C.c_void_p), ("queue", C.c_void_p),
                ("kernels_num", C.c_uint32), ("kernels", C.c_void_p * MAX_KERNELS),
                ("buffers", (C.c_void_p * MAX_BUFFERS) * MAX_KERNELS),
                ("buffers_ids", (C.c_uint32 * MAX_BUFFERS) * MAX_KERNELS), ("buffers_num", C.c_uint32 * MAX_KERNELS)]


CL_DEVICE_TYPE_GPU = 1 << 2
CL_MEM_READ_WRITE, CL_MEM_WRITE_ONLY, CL_MEM_READ_ONLY, CL_MEM_COPY_HOST_PTR = 1, 2, 4, 32


def wrap_lib():
    """The same library with the cl_wrap_* prototypes set (varargs are passed as c_char_p / None)."""
    L = lib()
    if not getattr(L, "_wrap_ready", False):
        wp = C.POINTER(ClWrap)
        L.cl_wrap_load_global_data.argtypes = [wp, C.c_uint32, C.c_uint32, C.c_void_p, C.c_size_t, C.c_uint64]
        L.cl_wrap_load_global_data.restype = None
        L.cl_wrap_load_single_data.argtypes = [wp, C.c_uint32, C.c_uint32, C.c_void_p, C.c_size_t]
        L.cl_wrap_load_single_data.restype = None
        L.cl_wrap_output.argtypes = [wp, C.c_size_t, C.c_size_t, C.c_uint32, C.c_uint32, C.c_int32, C.c_void_p]
        L.cl_wrap_output.restype = None
        L.cl_wrap_release.argtypes = [wp]
        L.cl_wrap_release.restype = None
        L.cl_wrap_init.restype = None      # variadic: call with explicit ctypes values
        L.cl_wrap_load_images.restype = None
        L.cuda_wrap_engine.argtypes = [wp]
        L.cuda_wrap_engine.restype = C.c_void_p
        L.cuda_wrap_readback_stats.argtypes = [wp, C.POINTER(C.c_ulong), C.POINTER(C.c_ulong)]
        L.cuda_wrap_readback_stats.restype = None
        L.cuda_wrap_last_kernel_ms.argtypes = [wp]
        L.cuda_wrap_last_kernel_ms.restype = C.c_float
        L._wrap_ready = True
    return L
