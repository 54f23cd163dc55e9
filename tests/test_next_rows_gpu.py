"""GPU tests for the SURVEY.md 8(f) "next" rows that had CPU tests only, and for the kernel paths no other test reaches:
  N2  a scene that went through the archive reader (rtb_map_load: RTM2, 32-bit counts) renders like the reference text,
      through the engine ABI and through cl_wrap_* with 4-byte count scalars;
  N3  frames rendered from textures that came through cl_wrap_load_images with the decoded-texel cache cold, then warm;
  non-resident geometry: a scene whose record blob exceeds the 96 KiB shared-memory budget (render_frame<false,...>),
      narrow and WIDE variants."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def host(pkg):
    from example_gui_opencl_raytracer_b200 import host as h
    return h


@pytest.fixture()
def fresh(pkg, scene):
    e = pkg.Engine(0)
    e.upload_images(0, scene.tex)
    e.upload_images(1, scene.sky)
    yield e
    e.close()


def judge_frame(ref, rgb, hit, want_rgb, want_hit, W, H, what):
    c = ref.compare_frames(rgb, want_rgb, W, H)
    same = 1.0 if hit is None else float((hit == want_hit).mean())
    print("%s: hit %.7f exact %.6f within1 %.6f psnr %.2f" % (what, same, c["exact"], c["within1"], c["psnr"]))
    assert same == 1.0
    assert c["within1"] >= 0.999 and c["psnr"] >= 50.0


def test_scene_archive_rtm2_renders_like_the_reference(pkg, ref, scene, host, fresh, tmp_path):
    """N2: 289 spheres do not fit the reference's one-byte counts; written as RTM2, read back by rtb_map_load, rendered."""
    s, p, l, camp = host.sphere_grid_scene(17)
    path = str(tmp_path / "grid289.map")
    host.save_map(path, s, p, l, version=2)
    s2, p2, l2, t2, ver = host.load_map_any(path)
    assert ver == 2 and (len(s2) // 96, len(p2) // 96, len(l2) // 48, len(t2)) == (289, 1, 3, 0)
    W, H = 192, 108
    cam = host.camera(W=W, H=H, **camp)
    sc = ref.Scene(s, p, l, scene.tex, scene.sky)         # the oracle gets the ORIGINAL arrays
    want_rgb, want_hit = ref.judge().render(sc, cam, W, H, wide=True)
    fresh.upload_scene(s2, p2, l2)
    fresh.set_camera(cam)
    rgb, hit = fresh.render_to_host(W, H, want_hit=True)
    judge_frame(ref, rgb, hit, want_rgb, want_hit, W, H, "engine ABI, scene from rtb_map_load")
    # the same archive through the drop-in boundary: counts are 4-byte scalars (slots (1,4..6), obj_size 4)
    from example_gui_opencl_raytracer_b200.wrap_session import WrapSession
    ses = WrapSession(W, H, s2, p2, l2, cam, root=ROOT)
    try:
        out = np.zeros(W * H, np.uint32)
        for _ in range(3):                                 # the third read-back goes through the mapped host buffer
            out[:] = 0
            ses.frame(out.ctypes.data)
            judge_frame(ref, out, None, want_rgb, None, W, H, "cl_wrap_*, 4-byte counts")
    finally:
        ses.release()


def test_scene_archive_v1_is_the_reference_file(pkg, ref, scene, host, fresh):
    """N2: scenes/render.map (the reference's own v1 archive) through rtb_map_load."""
    s, p, l, t, ver = host.load_map_any(os.path.join(ROOT, "scenes", "render.map"))
    assert ver == 1 and t == b"" and (s, p, l) == (bytes(scene.spheres), bytes(scene.planes), bytes(scene.lights))
    W, H = 200, 150
    cam = ref.camera(W=W, H=H, **ref.RAYPNG_CAMERA)
    want_rgb, want_hit = ref.judge().render(scene, cam, W, H)
    fresh.upload_scene(s, p, l)
    fresh.set_camera(cam)
    rgb, hit = fresh.render_to_host(W, H, want_hit=True)
    judge_frame(ref, rgb, hit, want_rgb, want_hit, W, H, "render.map via rtb_map_load")


def test_scene_archive_with_triangles_extension(pkg, ref, scene, host, fresh, tmp_path):
    """N2 + X1: an RTM2 archive with a triangle block.  The reference has no triangle type, so the judge here is the
    repo's own port (oracle/rt_oracle.c) -- PARITY UNPINNED by definition (SURVEY.md 8(a') X1)."""
    s, p, l, camp = host.sphere_grid_scene(17)
    tris = b"".join(host.triangle((-3 + k, 0.2, 2 + k), (-2 + k, 2.2, 3 + k), (-1 + k, 0.2, 2.5 + k),
                                  dict(host.PLASTIC if k % 2 else host.MIRROR, rgb=(.2 + .1 * k, .8 - .1 * k, .4)))
                    for k in range(6))
    path = str(tmp_path / "grid289_tri.map")
    host.save_map(path, s, p, l, tris, version=2)
    s2, p2, l2, t2, ver = host.load_map_any(path)
    assert ver == 2 and t2 == tris and s2 == s
    W, H = 160, 90
    cam = host.camera(W=W, H=H, **camp)
    sc = ref.Scene(s, p, l, scene.tex, scene.sky)
    port = ref.port()
    port.set_triangles(tris)
    try:
        want_rgb, want_hit = port.render(sc, cam, W, H, wide=True)
    finally:
        port.set_triangles(b"")
    assert (want_hit > 1 + 289 + 1).sum() > 50            # the triangles are in view
    fresh.upload_scene(s2, p2, l2)
    fresh.upload_triangles(t2)
    fresh.set_camera(cam)
    rgb, hit = fresh.render_to_host(W, H, want_hit=True)
    judge_frame(ref, rgb, hit, want_rgb, want_hit, W, H, "RTM2 with triangles vs the port (unpinned)")


def test_texture_ingest_cache_cold_then_warm(pkg, ref, scene, host, tmp_path, monkeypatch):
    """N3: cl_wrap_load_images with RTB_IMAGE_CACHE set -- first run decodes the PNGs and fills the cache, second run
    reads the cached texels; both frames (plane textures and skybox texels all over them) equal the oracle's."""
    from example_gui_opencl_raytracer_b200.wrap_session import WrapSession
    cache = tmp_path / "texcache"
    cache.mkdir()
    monkeypatch.setenv("RTB_IMAGE_CACHE", str(cache))
    W, H = 320, 240
    cam = ref.camera(W=W, H=H, **ref.RAYPNG_CAMERA)
    want_rgb, _ = ref.judge().render(scene, cam, W, H, want_hit=False)
    stamps = None
    for run in ("cold", "warm"):
        ses = WrapSession(W, H, bytes(scene.spheres), bytes(scene.planes), bytes(scene.lights), cam, root=ROOT)
        try:
            out = np.zeros(W * H, np.uint32)
            ses.frame(out.ctypes.data)
        finally:
            ses.release()
        judge_frame(ref, out, None, want_rgb, None, W, H, "cl_wrap_load_images, cache " + run)
        entries = sorted(os.listdir(cache))
        assert len(entries) == 5 and all(e.endswith(".rgba") for e in entries)      # 4 plane textures + the skybox
        now = [os.stat(cache / e).st_mtime_ns for e in entries]
        if stamps is not None:
            assert now == stamps                                                     # warm run: entries read, not rewritten
        stamps = now


@pytest.mark.parametrize("wide", [False, True])
def test_scene_larger_than_shared_memory(pkg, ref, scene, host, fresh, wide):
    """81 x 81 = 6,561 spheres: a 105 KB record blob, above the 96 KiB residency budget, so render_frame<false,...>
    reads the records through L1/L2 instead of shared memory.  Narrow and WIDE variants against the widened oracle."""
    s, p, l, camp = host.sphere_grid_scene(81)
    W, H = 96, 54
    cam = host.camera(W=W, H=H, **camp)
    sc = ref.Scene(s, p, l, scene.tex, scene.sky)
    want_rgb, want_hit = ref.judge().render(sc, cam, W, H, wide=True)
    fresh.upload_scene(s, p, l)
    fresh.set_camera(cam)
    fresh.set_option("wide_min_records", 64 if wide else -1)
    if wide:
        fresh.set_option("wide_max_live", 32)
    rgb, hit = fresh.render_to_host(W, H, want_hit=True)
    assert fresh.last_launch()["resident_scene"] == 0
    judge_frame(ref, rgb, hit, want_rgb, want_hit, W, H, "6,561 spheres, non-resident, wide=%s" % wide)


def test_completion_word_and_wait_kernel(pkg, ref, scene, fresh):
    """the completion signal fused into the render (rtb_set_completion: the last CTA of the frame's
    kernels release-stores the frame number) and the gathering rank's wait kernel (rtb_wait_flags), on one GPU: the flag
    word carries the value once the stream has drained, with and without the pixel-tree kernel beside render_frame; a
    flag that never arrives ends in a counted timeout (~2 s), not a hung GPU."""
    W, H = 1024, 512
    fresh.upload_scene(scene.spheres, scene.planes, scene.lights)
    fresh.set_camera(ref.camera(W=W, H=H, **ref.RAYPNG_CAMERA))
    want, _ = fresh.render_to_host(W, H, want_hit=False)
    buf = fresh.dev_alloc(W * H * 4 + 4096)
    flags = buf.value + W * H * 4
    fresh.dev_write(buf, np.zeros(W * H + 1024, np.uint32))
    part = pkg.Partition(8, 8, 3)                      # a part small enough for the tree kernel's automatic threshold
    rows = fresh.partition_rows(H, part)
    frame_no = 0
    for tree in (1, 0):
        fresh.set_option("tree", tree)
        for rep in range(3):
            frame_no += 1
            fresh.set_mirror_frame(buf.value)          # full-frame mirror, global pixel ids
            fresh.set_completion(flags + 64 * 2, frame_no)
            fresh.render(W, H, part=part)
            fresh.wait_flags(flags + 64 * 2, 16, 1, frame_no)
            fresh.sync()
            got = fresh.dev_read(buf, W * H * 4 + 4096).view(np.uint32)
            assert int(got[W * H + 32]) == frame_no
            y = np.arange(H)
            mine = (y // 8) % 8 == 3
            assert np.array_equal(got[:W * H].reshape(H, W)[mine], want.reshape(H, W)[mine]), (tree, rep)
    fresh.set_option("tree", 1)
    assert fresh.wait_timeouts() == 0
    fresh.wait_flags(flags + 64 * 5, 16, 1, 7)         # nobody ever writes this word
    assert fresh.wait_timeouts() == 1
    fresh.dev_free(buf)


def test_schedule_lookahead_never_changes_a_frame(pkg, ref, scene, host):
    """r02, rt_engine.cu: while a history-scheduled frame renders, the next frame's queue is built on a second stream
    into a second set of queue buffers.  A sequence that keeps breaking what the look-ahead assumed -- another size,
    another partition, a counted (literal) frame, an option that moves the thresholds, a new scene, a frame through
    rtb_render_noalloc, frames too small for the ordered queue -- must give, frame for frame, the words of the same
    sequence rendered with the look-ahead off; and both must be the oracle's frame where one is at hand."""
    s2, p2, l2, camp2 = host.sphere_grid_scene(5)
    part_a, part_b = pkg.Partition(8, 2, 1), pkg.Partition(8, 4, 2)
    seq = ([("render", 1024, 512, None)] * 5 + [("render", 768, 768, None)] * 2 + [("render", 1024, 512, None)] * 3 +
           [("render", 1024, 1024, part_a)] * 4 + [("render", 1024, 1024, part_b)] * 3 + [("counted", 1024, 512, None)] +
           [("render", 1024, 512, None)] * 3 + [("option", "tree", 0)] + [("render", 1024, 512, None)] * 3 + [("option", "tree", 1)] +
           [("render", 1024, 512, None)] * 3 + [("noalloc", 1024, 512, None)] + [("render", 1024, 512, None)] * 2 +
           [("render", 96, 64, None)] * 2 + [("render", 1024, 512, None)] * 3 + [("scene",)] + [("render", 1024, 512, None)] * 4)
    out = {}
    for la in (0, 1):
        eng = pkg.Engine(0)
        try:
            eng.upload_images(0, scene.tex); eng.upload_images(1, scene.sky)
            eng.upload_scene(scene.spheres, scene.planes, scene.lights)
            eng.set_option("lookahead", la)
            camp = dict(ref.RAYPNG_CAMERA)
            scratch = eng.dev_alloc(1024 * 512 * 4)
            frames = []
            for k, step in enumerate(seq):
                if step[0] == "option":
                    eng.set_option(step[1], step[2])
                    continue
                if step[0] == "scene":
                    eng.upload_scene(s2, p2, l2)
                    camp = camp2
                    continue
                W, H, part = step[1], step[2], step[3]
                o = np.array(camp["origin"], np.float64) + np.array([0.01 * k, 0.0, 0.0])      # the camera moves every frame
                eng.set_camera(host.camera(W=W, H=H, **dict(camp, origin=tuple(o))))
                if step[0] == "counted":
                    eng.render_counted(W, H)
                    frames.append(eng.read_frame(W * H))
                elif step[0] == "noalloc":
                    eng.render_noalloc(W, H, scratch.value)
                    eng.sync()
                    frames.append(eng.dev_read(scratch, W * H * 4).view(np.uint32).copy())
                else:
                    rgb, hit = eng.render_to_host(W, H, want_hit=True, part=part)
                    frames.append(np.concatenate([rgb, hit]))
            eng.dev_free(scratch)
            out[la] = frames
        finally:
            eng.close()
    assert len(out[0]) == len(out[1])
    for k, (a, b) in enumerate(zip(out[0], out[1])):
        assert np.array_equal(a, b), k
    # one frame of the sequence against the reference text: the 5th (look-ahead queue in use), 1024x512
    W, H = 1024, 512
    o = np.array(ref.RAYPNG_CAMERA["origin"], np.float64) + np.array([0.04, 0.0, 0.0])
    want_rgb, want_hit = ref.judge().render(scene, ref.camera(W=W, H=H, **dict(ref.RAYPNG_CAMERA, origin=tuple(o))), W, H)
    judge_frame(ref, out[1][4][:W * H], out[1][4][W * H:], want_rgb, want_hit, W, H, "5th frame of the look-ahead sequence")
