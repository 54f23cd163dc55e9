"""rtb_render_noalloc (include/rtb200.h): the entry point for callers that cannot call cudaMalloc -- its scheduling scratch
is a file-scope static __device__ array.  No counterpart in the reference.
Called through the C ABI and compared with the existing entry point (rtb_render) on two calls in a row: the scratch keeps
the last call's contents, so the second call (another camera, another size) proves the clearing.  A scheduling difference
cannot change a pixel, so the comparison is word for word -- tighter than BASELINE.json's tolerance, which is asserted
too.  This file is collected last on purpose."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def frames_equal(ref, got, want, W, H):
    c = ref.compare_frames(got, want, W, H)
    assert c["within1"] >= 0.999 and c["psnr"] >= 50.0, c      # BASELINE.json's tolerance
    assert np.array_equal(got, want)                           # and in fact the same words


def test_noalloc_entry_point_equals_rtb_render_twice_in_a_row(pkg, ref, scene):
    eng = pkg.Engine(0)
    try:
        eng.upload_scene(scene.spheres, scene.planes, scene.lights)
        eng.upload_images(0, scene.tex)
        eng.upload_images(1, scene.sky)
        calls = [(1024, 512, ref.camera(W=1024, H=512, **ref.RAYPNG_CAMERA)),                      # 16,384 tiles
                 (640, 360, ref.camera(origin=(1.2, 2.2, -7.5), look=(0.15, -0.05, 1.0), fov=90.0, focal=1.0, W=640, H=360)),
                 (1024, 512, ref.camera(W=1024, H=512, **ref.RAYPNG_CAMERA))]
        out = eng.dev_alloc(1024 * 512 * 4)          # the CALLER's buffers (allocated here, outside the entry point)
        hit = eng.dev_alloc(1024 * 512 * 4)
        for W, H, cam in calls:
            eng.set_camera(cam)
            want_rgb, want_hit = eng.render_to_host(W, H, want_hit=True)           # the existing entry point
            eng.render_noalloc(W, H, out, cam=cam, d_hit=hit)
            eng.sync()
            got_rgb = eng.dev_read(out, W * H * 4).view(np.uint32)
            got_hit = eng.dev_read(hit, W * H * 4).view(np.uint32)
            frames_equal(ref, got_rgb, want_rgb, W, H)
            assert np.array_equal(got_hit, want_hit)
        # a row-band part, the latched camera (cam = NULL), no hit buffer
        W, H = 328, 250
        eng.set_camera(ref.camera(W=W, H=H, **ref.RAYPNG_CAMERA))
        part = pkg.Partition(8, 3, 1)
        rows = eng.partition_rows(H, part)
        want_rgb, _ = eng.render_to_host(W, H, part=part, want_hit=False)
        eng.render_noalloc(W, H, out, part=part)
        eng.sync()
        assert np.array_equal(eng.dev_read(out, rows * W * 4).view(np.uint32), want_rgb)
        eng.dev_free(out)
        eng.dev_free(hit)
    finally:
        eng.close()


def test_noalloc_entry_point_full_4k_and_the_limit(pkg, ref, scene):
    """3840x2160 = 259,200 tiles fits the static scratch (262,144 tiles); 4096x2304 = 294,912 tiles is rejected with a clear
    error, and so is a call without an output buffer."""
    eng = pkg.Engine(0)
    try:
        eng.upload_scene(scene.spheres, scene.planes, scene.lights)
        eng.upload_images(0, scene.tex)
        eng.upload_images(1, scene.sky)
        W, H = 3840, 2160
        cam = ref.camera(W=W, H=H, **ref.RAYPNG_CAMERA)
        eng.set_camera(cam)
        want_rgb, _ = eng.render_to_host(W, H, want_hit=False)
        out = eng.dev_alloc(4096 * 2304 * 4)
        for _ in range(2):
            eng.render_noalloc(W, H, out, cam=cam)
            eng.sync()
            frames_equal(ref, eng.dev_read(out, W * H * 4).view(np.uint32), want_rgb, W, H)
        with pytest.raises(pkg.RtbError, match="exceed the static scratch"):
            eng.render_noalloc(4096, 2304, out, cam=ref.camera(W=4096, H=2304, **ref.RAYPNG_CAMERA))
        with pytest.raises(pkg.RtbError, match="caller passes the output buffer"):
            eng.render_noalloc(W, H, None, cam=cam)
        eng.render_noalloc(4096, 2304, out, cam=ref.camera(W=4096, H=2304, **ref.RAYPNG_CAMERA), part=pkg.Partition(8, 2, 0))   # half of it fits
        eng.sync()
        eng.dev_free(out)
    finally:
        eng.close()
