"""The reference drivers' cl_wrap_* call sequence (raypng.c:31-98, rayinteractive.c:132-196), made from Python
through ctypes against the SAME C entry points (include/cuda_wrap.h) -- used by bench.py's e2e leg and the
tests.  Host buffers in, host buffer out; nothing here touches CUDA directly."""
import ctypes as C
import os

import numpy as np

from . import (CL_DEVICE_TYPE_GPU, CL_MEM_COPY_HOST_PTR, CL_MEM_READ_ONLY, CL_MEM_READ_WRITE, CL_MEM_WRITE_ONLY, ClWrap,
               wrap_lib)
from . import host as _host


class WrapSession:
    def __init__(self, W, H, spheres, planes, lights, cam, root=_host.ROOT, texture_files=_host.TEXTURE_FILES,
                 skybox_file=_host.SKYBOX_FILE, triangles=b""):
        self.L = L = wrap_lib()
        self.W, self.H = W, H
        self.w = w = ClWrap()
        cs = C.c_char_p
        # raypng.c:31-34
        L.cl_wrap_init(C.byref(w), C.c_uint64(CL_DEVICE_TYPE_GPU), cs(b"src/cl/raygen.cl"), cs(b"raygen"),
                       cs(b"src/cl/raytracing.cl"), cs(b"raytracer"), cs(None))
        self._vec = [np.zeros(4, np.float32) for _ in range(4)]
        self._wf, self._hf = C.c_float(), C.c_float()
        self.set_camera(cam)
        pw, ph, pixels = C.c_uint32(W), C.c_uint32(H), C.c_uint32(W * H)
        L.cl_wrap_load_single_data(C.byref(w), 0, 6, C.byref(pw), 4)                      # raypng.c:56-57
        L.cl_wrap_load_single_data(C.byref(w), 0, 7, C.byref(ph), 4)
        L.cl_wrap_load_global_data(C.byref(w), 0, 8, None, 64 * W * H, CL_MEM_READ_WRITE)  # :59
        L.cl_wrap_load_single_data(C.byref(w), 1, 0, C.byref(w.buffers[0], 8 * C.sizeof(C.c_void_p)), 8)  # :61
        s, p, l = bytes(spheres), bytes(planes), bytes(lights)
        L.cl_wrap_load_global_data(C.byref(w), 1, 1, s, len(s), CL_MEM_READ_ONLY)         # :62-67
        L.cl_wrap_load_global_data(C.byref(w), 1, 2, p, len(p), CL_MEM_READ_ONLY)
        L.cl_wrap_load_global_data(C.byref(w), 1, 3, l, len(l), CL_MEM_READ_ONLY)
        ns, np_, nl = len(s) // 96, len(p) // 96, len(l) // 48
        if max(ns, np_, nl) <= 255:   # cl_uchar, as the reference passes them (:68-70)
            cnts, csz = [C.c_uint8(ns), C.c_uint8(np_), C.c_uint8(nl)], 1
        else:                         # widened counts
            cnts, csz = [C.c_uint32(ns), C.c_uint32(np_), C.c_uint32(nl)], 4
        for i in range(3):
            L.cl_wrap_load_single_data(C.byref(w), 1, 4 + i, C.byref(cnts[i]), csz)
        L.cl_wrap_load_single_data(C.byref(w), 1, 7, C.byref(pixels), 4)                  # :71
        tex = [cs(os.path.join(root, f).encode()) for f in texture_files]
        L.cl_wrap_load_images(C.byref(w), C.c_uint32(1), C.c_uint32(8), C.c_uint64(CL_MEM_COPY_HOST_PTR),
                              C.c_uint32(len(tex)), *tex)                                  # :74-78
        L.cl_wrap_load_images(C.byref(w), C.c_uint32(1), C.c_uint32(9), C.c_uint64(CL_MEM_COPY_HOST_PTR), C.c_uint32(1),
                              cs(os.path.join(root, skybox_file).encode()))               # :80-81
        L.cl_wrap_load_global_data(C.byref(w), 1, 10, None, 4 * W * H, CL_MEM_WRITE_ONLY)  # :83
        if len(triangles):  # EXTENSION X1: optional raytracer args 11 (rtriangle buffer) and 12 (cl_uint count)
            t = bytes(triangles)
            nt = C.c_uint32(len(t) // 112)
            L.cl_wrap_load_global_data(C.byref(w), 1, 11, t, len(t), CL_MEM_READ_ONLY)
            L.cl_wrap_load_single_data(C.byref(w), 1, 12, C.byref(nt), 4)

    def set_camera(self, cam):
        """The six per-frame scalars (rayinteractive.c:98-103)."""
        cam = np.ascontiguousarray(cam, np.float32)
        L, w = self.L, self.w
        for i in range(4):
            self._vec[i][:3] = cam[3 * i: 3 * i + 3]
            L.cl_wrap_load_single_data(C.byref(w), 0, i, self._vec[i].ctypes.data_as(C.c_void_p), 16)
        self._wf.value, self._hf.value = float(cam[12]), float(cam[13])
        L.cl_wrap_load_single_data(C.byref(w), 0, 4, C.byref(self._wf), 4)
        L.cl_wrap_load_single_data(C.byref(w), 0, 5, C.byref(self._hf), 4)

    def frame(self, host_ptr):
        """raypng.c:86-89: run raygen, run raytracer + blocking read-back into the HOST buffer at host_ptr."""
        n = self.W * self.H
        self.L.cl_wrap_output(C.byref(self.w), n, 0, 0, 0, 0, None)
        self.L.cl_wrap_output(C.byref(self.w), n, 4 * n, 1, 1, 10, C.c_void_p(host_ptr))

    def kernel_ms(self):
        return float(self.L.cuda_wrap_last_kernel_ms(C.byref(self.w)))

    def readback_stats(self):
        a, b = C.c_ulong(), C.c_ulong()
        self.L.cuda_wrap_readback_stats(C.byref(self.w), C.byref(a), C.byref(b))
        return {"mirrored": a.value, "copied": b.value}

    def release(self):
        if self.w is not None:
            self.L.cl_wrap_release(C.byref(self.w))
            self.w = None
