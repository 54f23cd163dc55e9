"""Experiment: frame written straight into pinned host memory by the kernel vs render + D2H copy."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry
pkg = entry.load_package()
from example_gui_opencl_raytracer_b200 import host
W, H = 3840, 2160
s, p, l = host.load_map(os.path.join(ROOT, "scenes", "render.map"))
tex, sky = host.default_images(ROOT)
eng = pkg.Engine(0)
eng.upload_scene(s, p, l); eng.upload_images(0, tex); eng.upload_images(1, sky)
eng.set_camera(host.camera(W=W, H=H, **host.RAYPNG_CAMERA))
dev = torch.device("cuda", 0)
stream = torch.cuda.ExternalStream(eng.stream(), device=dev)
pinned = torch.zeros(W * H, dtype=torch.int32, pin_memory=True)
dbuf = torch.zeros(W * H, dtype=torch.int32, device=dev)
pageable = np.zeros(W * H, np.int32)

def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) * 1e3 / n

def a():
    eng.render(W, H, d_out=dbuf.data_ptr()); eng.sync()
def b():
    eng.render(W, H, d_out=dbuf.data_ptr())
    with torch.cuda.stream(stream): pinned.copy_(dbuf, non_blocking=True)
    eng.sync()
def c():
    eng.render(W, H, d_out=pinned.data_ptr()); eng.sync()
def d():
    eng.render(W, H, d_out=dbuf.data_ptr()); eng.sync()
    import ctypes as C; pkg.lib().rtb_dev_read(eng._h, C.c_void_p(pageable.ctypes.data), C.c_void_p(dbuf.data_ptr()), C.c_size_t(W * H * 4))
print("device only          %.3f ms" % t(a))
print("device + D2H pinned  %.3f ms" % t(b))
print("zero-copy to pinned  %.3f ms" % t(c))
ref = dbuf.cpu().numpy()
print("zero-copy identical:", bool((pinned.numpy() == ref).all()))
try:
    print("device + D2H pageable %.3f ms" % t(d, 5))
except Exception as ex:
    print("pageable leg failed", ex)
t0 = time.perf_counter()
arr = np.zeros(W * H, np.int32)
import ctypes as C; rc = pkg.lib().rtb_host_register(C.c_void_p(arr.ctypes.data), C.c_size_t(arr.nbytes))
print("cudaHostRegister 33 MB: %.3f ms rc=%d" % ((time.perf_counter() - t0) * 1e3, rc))
