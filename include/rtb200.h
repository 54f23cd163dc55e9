/* include/rtb200.h -- C ABI of the B200 render engine (librtb200.so).
 *
 * This is the device half of the drop-in boundary: the `cl_wrap_*` entry points in cuda_wrap.h
 * (which replace /root/reference/src/opencl_wrap.h:29-42) are a thin C layer over these calls.
 * Plain pointers and sizes only; no C++ or torch types.  All functions return 0 on success and a
 * negative code on failure; rtb_last_error() returns a static, human-readable reason.
 *
 * What one engine holds (all in HBM of one GPU):
 *   - the raw wire arrays the host uploads, byte for byte (rsphere 96 B, rplane 96 B, rlight 48 B;
 *     reference src/cl/types.cl:4-42 == src/cpu_obj.h:10-48),
 *   - the two RGBA8 image arrays (what opencl_wrap.c:189-349 builds: layers x H x W x uchar4),
 *   - a compact copy of the geometry produced on the device by the `scene_prep` kernel,
 *   - the frame buffer (0x00RRGGBB words, raytracing.cl:193-194) and the optional primary hit-id buffer.
 * The renderer fuses the reference's two kernels, `raygen` (src/cl/raygen.cl:5-25) and
 * `raytracer` (src/cl/raytracing.cl:14-195).
 */
#ifndef RTB200_H
#define RTB200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rtb_engine rtb_engine;

/* Hit-id encoding: 0 sky, 1+i sphere i, 1+S+j plane j, 1+S+P+k triangle k (extension),
 * RTB_HIT_LIGHT|k light k. */
#define RTB_HIT_LIGHT 0x80000000u

/* Work counters filled by rtb_render_counted (same order as oracle/rt_oracle.h). */
enum {
    RTB_SEGMENTS = 0,     /* iterations of raytracing.cl:42 */
    RTB_SHADOW_RAYS = 1,  /* testShadowPath calls, raytracing.cl:108 */
    RTB_SPHERE_TESTS = 2, /* intersect_sphere calls (spheres and lights), primitives.cl:170 */
    RTB_SPHERE_DPOS = 3,  /* ... of which had D >= 0 */
    RTB_PLANE_TESTS = 4,  /* intersect_plane calls, primitives.cl:197 */
    RTB_TEXELS = 5,       /* texel fetches the reference would make (primitives.cl:250, raytracing.cl:70) */
    RTB_TRI_TESTS = 7,    /* EXTENSION X1: triangle tests */
    RTB_SHADOW_SKIPPED = 6, /* executed-work accounting only: shadow rays elided because their light term is exactly 0 */
    RTB_NCOUNTERS = 8
};

/* Row-band partition of a frame.  Bands of `band_rows` rows are dealt out
 * cyclically: band b belongs to part (b % parts).  Part `part` renders its rows into a COMPACT
 * buffer (local row lr <-> global row ((lr / band_rows) * parts + part) * band_rows + lr % band_rows)
 * while pixel ids (RNG seed, raytracing.cl:33; raygen.cl:13-14) stay global.
 * {band_rows = H, parts = 1, part = 0} is the whole frame. */
typedef struct {
    uint32_t band_rows;
    uint32_t parts;
    uint32_t part;
} rtb_partition;

const char* rtb_last_error(void);
int rtb_device_count(void);

/* Create an engine on CUDA device `device` (its own non-blocking stream). */
int rtb_create(rtb_engine** out, int device);
void rtb_destroy(rtb_engine* e);

/* Upload the scene arrays in wire format (blocking H2D), then run scene_prep on the device.
 * Pointers may be NULL when the count is 0.  Replaces any previous scene. */
int rtb_upload_scene(rtb_engine* e, const void* spheres, uint32_t S, const void* planes, uint32_t P,
                     const void* lights, uint32_t L);
/* Same, but the wire arrays already live in this device's memory (not copied, must stay valid). */
int rtb_bind_scene_device(rtb_engine* e, const void* d_spheres, uint32_t S, const void* d_planes, uint32_t P,
                          const void* d_lights, uint32_t L);

/* EXTENSION X1 -- triangles; the reference has none, so there is no reference parity for scenes
 * that use them (checked against oracle/rt_oracle.c's own definition only).  Wire record rtriangle = 112 bytes:
 * float4 v0, v1, v2 (w unused) then the 64-byte rmaterial.  Triangles are visited after the planes in all three
 * traversals, are opaque and two-sided, normal = normalize(cross(v1-v0, v2-v0)) (not flipped), never textured.
 * T = 0 removes them.  Call after rtb_upload_scene / rtb_bind_scene_device (which do not touch them). */
int rtb_upload_triangles(rtb_engine* e, const void* tris, uint32_t T);
int rtb_bind_triangles_device(rtb_engine* e, const void* d_tris, uint32_t T);

/* RGBA8 image arrays, layers*h*w*4 bytes, row 0 = top row.  which: 0 = plane textures (kernel arg 8),
 * 1 = skybox (kernel arg 9, layers must be 1). */
int rtb_upload_images(rtb_engine* e, int which, const uint8_t* rgba, uint32_t w, uint32_t h, uint32_t layers);
/* Same, but the texels already live in this device's memory (not copied, must stay valid). */
int rtb_bind_images_device(rtb_engine* e, int which, const void* d_rgba, uint32_t w, uint32_t h, uint32_t layers);

/* The raygen kernel's scalar arguments (raygen.cl:5-8): corner, origin, up, right (xyz each), w_factor,
 * h_factor.  Latched by value; takes effect at the next render. */
int rtb_set_camera(rtb_engine* e, const float cam[14]);

/* Number of rows part `part` owns in an H-row frame. */
uint32_t rtb_partition_rows(uint32_t H, const rtb_partition* p);

/* Launch one frame (asynchronously, on the engine's stream).  `total` is the raytracer kernel's
 * total_size guard (raytracing.cl:23-26); pass W*H.  d_out / d_hit are DEVICE pointers to
 * rows*W uint32 each, or NULL to use the engine's own buffers (d_hit NULL + want_hit 0 = no AOV). */
int rtb_render(rtb_engine* e, uint32_t W, uint32_t H, uint32_t total, const rtb_partition* part, uint32_t* d_out,
               uint32_t* d_hit, int want_hit);
/* Same launch with work counters (slower, for accounting only). counters[RTB_NCOUNTERS] is host memory;
 * the call synchronises.  executed = 0: the literal trace, whose counts are the reference's own control
 * flow (the definition of "rays" and algorithmic flop); executed = 1: the production
 * trace, counting the work it really does (batched tests, elided shadow rays). */
int rtb_render_counted(rtb_engine* e, uint32_t W, uint32_t H, uint32_t total, const rtb_partition* part,
                       uint32_t* d_out, uint32_t* d_hit, int want_hit, int executed, uint64_t* counters);

/* On the gathering rank: re-interleave `parts` compact band buffers (part r at d_compact + r*stride_words, as the
 * gather delivers them; stride_words >= rows of part 0 x W) into the row-major H x W frame, on the engine's stream. */
int rtb_assemble_bands(rtb_engine* e, const uint32_t* d_compact, size_t stride_words, uint32_t* d_frame, uint32_t W,
                       uint32_t H, uint32_t band_rows, uint32_t parts);

/* Engine options.  "trace": 0 = production trace (warp-cooperative shadow rays; default), 1 = literal trace
 * (reference control flow; same pixels, slower), 2 = per-lane fast trace (same pixels; the previous production path).  "tile_order": 1 = cost-ordered tile queue (default), 0 = row-major queue.
 * "history": 1 (default) = a frame of the same size/partition/scene as the previous one is scheduled from that frame's
 * MEASURED per-tile times (4*log2 cost scale) instead of the skeleton trace; 0 = always the skeleton trace.
 * "lookahead": 1 (default; environment RTB_LOOKAHEAD=0 turns it off for callers behind cl_wrap_*) = while a history
 * frame renders, the NEXT frame's queue is built on a second stream from the frame before's times, so that the two
 * scheduling kernels leave the frame's critical path; used when the next frame has the same size, partition, scene and
 * thresholds, else that frame builds its own queue.  Pixels do not change.
 * "split_min_cost" (history frames only): tiles whose cost estimate reaches this value are dealt to four warps (one 8-pixel tile row each,
 * the other 24 lanes only run shadow tasks) so that no single warp owns a whole glass/mirror tile (64 = never, 0 = automatic: tiles longer than 0.75 x the queue's expected drain time).
 * "split32_min_cost": ... and from this value on to 32 warps, one pixel each (64 = never, 0 = automatic: tiles
 * longer than 3 x the expected drain time).
 * "coop_min_cost": tiles whose cost estimate (0..63) reaches this value use the cooperative walk, cheaper tiles the
 * per-lane walk (0 = all cooperative, 64 = none; scheduling only, pixels do not change).
 * "wide_min_records": scenes without triangles whose padded light + sphere records reach this number (default 64;
 * negative = never) run the WIDE variant of the production trace: rounds with few tracing lanes walk the records one
 * ray at a time over all 32 lanes.  "wide_max_live": -1 (default) = those are the rounds below the instruction
 * break-even; 0..32 = fixed limit (32 = every round, used by the tests).  Pixels do not change.
 * "tree": level-wise pixel trees (csrc/rt_trace_tree.cuh: the segment trees of an entry's pixels expanded breadth-first
 * over the 128 threads of a CTA, RNG jump-ahead, ordered replay).  1 (default) = the heavy queue entries of
 * history-scheduled frames: tiles whose cost reaches "tree_min_cost" (0 = automatic) and every split entry; 0 = never;
 * 2 = every tile; 3 = every tile, one pixel at a time (what a single-pixel entry does; tests).
 * "tree_cap": nodes per CTA slab (default 8192; a larger tree falls back to the cooperative walk).  Pixels do not
 * change.
 * "tri_tree_min" (extension X1; default 1024, 0 = never): scenes with at least this many triangles render EVERY tile in the
 * tree kernel, where the triangle records are staged into shared memory in 12 KiB chunks by TMA bulk copies and each
 * chunk serves the 128 rays of a CTA.  Pixels do not change.
 * "accel" (EXTENSION; default 0 = brute force over every record, as the reference does): 1 = scenes
 * with >= 16 spheres walk a cluster structure built at upload (csrc/rt_trace_accel.cuh) that skips only tests whose
 * outcome is provably "no hit"; the frame is word-identical to the brute-force one.  Scenes with >= 16 triangles
 * (extension X1) walk a box hierarchy over the triangles, built on the host at the first frame after an upload; its
 * frames equal the brute-force ones wherever compared (tests/test_accel_gpu.py, bench.py), which for Moeller-Trumbore's
 * noise at grazing incidence is a measurement, not a proof (rt_trace_accel.cuh). */
int rtb_set_option(rtb_engine* e, const char* name, int value);
/* Brute-force check of the arithmetic routines against IEEE division / sqrt (trials per thread on 592x256
 * threads).  out[0] = division mismatches, out[1] = normalize mismatches, out[2] = max |sincos - libdevice|
 * in units of 2^-24. */
int rtb_selftest_arith(rtb_engine* e, uint32_t seed, uint32_t per_thread, uint32_t out[3]);

/* Wait for the engine's stream. */
int rtb_sync(rtb_engine* e);
/* Blocking D2H of the engine's own frame / hit buffers (nbytes from the start). */
int rtb_read_frame(rtb_engine* e, void* host, size_t nbytes);
int rtb_read_hit(rtb_engine* e, void* host, size_t nbytes);
/* Device pointers of the engine's own buffers (valid until the next larger render). */
uint32_t* rtb_frame_ptr(rtb_engine* e);
uint32_t* rtb_hit_ptr(rtb_engine* e);

/* Device time of the last render launch in milliseconds (CUDA events on the engine's stream;
 * synchronises).  Negative on error. */
float rtb_last_render_ms(rtb_engine* e);
/* Same, for the render_frame kernel alone (without the tile-ordering pre-pass). */
float rtb_last_kernel_ms(rtb_engine* e);
/* The engine's CUDA stream as a cudaStream_t cast to void* (for callers that enqueue their own
 * work, e.g. a collective, behind the render). */
void* rtb_stream(rtb_engine* e);

/* Raw device memory helpers for callers without a CUDA runtime of their own (the cl_wrap layer). */
int rtb_dev_alloc(rtb_engine* e, void** dptr, size_t nbytes);
int rtb_dev_free(rtb_engine* e, void* dptr);
int rtb_dev_write(rtb_engine* e, void* dptr, const void* host, size_t nbytes); /* blocking H2D */
int rtb_dev_read(rtb_engine* e, void* host, const void* dptr, size_t nbytes);  /* blocking D2H */
int rtb_host_register(void* host, size_t nbytes);   /* page-lock a caller buffer (optional fast path) */
/* Device-visible view of a HOST buffer: page-locks and maps it unless it already is (cudaHostAlloc'ed or
 * registered memory); *registered_here = 1 when this call did the registration (undo with rtb_host_unregister). */
int rtb_host_map(void* host, size_t nbytes, void** dev_view, int* registered_here);
/* One-shot: the NEXT rtb_render also writes every frame word to `mirror` (same compact layout as d_out; any
 * device-visible pointer, typically an rtb_host_map view, so the frame is in the caller's host buffer when the
 * kernel ends -- PCIe writes overlap the render, no D2H copy).  Cleared by that render.  16-byte aligned. */
int rtb_set_mirror(rtb_engine* e, uint32_t* mirror);
/* Same, but `frame` is a full H x W row-major frame and every word goes to frame[y*W + x] (global pixel id) whatever
 * the partition: with a PEER GPU's frame buffer (rtb_ipc_open) each rank's bands land in the gathering rank's frame
 * over NVLink while the kernel runs -- the band exchange fused into the render. */
int rtb_set_mirror_frame(rtb_engine* e, uint32_t* frame);
/* Completion signal fused into the render (multi-GPU, replaces a per-frame collective): one-shot -- the NEXT
 * rtb_render's last CTA stores `value` into *flag (system-scope release, after every frame store of that kernel has
 * been fenced).  `flag` is any device-visible word: a slot in the gathering rank's frame allocation (rtb_ipc_open) or
 * in mapped host memory.  The gathering rank enqueues rtb_wait_flags on its own stream: a one-warp-per-32-flags kernel
 * that returns when flags[i * stride_words] has reached `value` for all i < n (wrap-safe; bounded at ~2 s, after which
 * rtb_wait_timeouts reports a non-zero count instead of a hung GPU). */
int rtb_set_completion(rtb_engine* e, uint32_t* flag, uint32_t value);
int rtb_wait_flags(rtb_engine* e, const uint32_t* d_flags, uint32_t stride_words, uint32_t n, uint32_t value);
int rtb_wait_timeouts(rtb_engine* e, uint32_t* out);
/* CUDA IPC plumbing for that (one process per GPU): export a cudaMalloc'ed buffer of this engine's device as a 64-byte
 * handle / map a peer's handle into this process (peer access enabled lazily) / unmap. */
int rtb_ipc_export(rtb_engine* e, const void* dptr, unsigned char handle[64]);
int rtb_ipc_open(rtb_engine* e, const unsigned char handle[64], void** dptr);
int rtb_ipc_close(rtb_engine* e, void* dptr);
int rtb_host_unregister(void* host);

/* Scene archives.  v1 = the reference's .map (src/cpu_obj.c:51-101: 1-byte counts, <= 255 per
 * type); v2 = "RTM2" + four uint32 counts (spheres, planes, lights, triangles) + the raw arrays.  rtb_map_load reads
 * either (arrays malloc'ed, 16-byte aligned; release with rtb_map_free); rtb_map_save writes the version asked for
 * (-3 if the scene does not fit v1).  Returns 0, -2 on I/O errors, -3 on a malformed file. */
typedef struct {
    void *spheres, *planes, *lights, *triangles;
    uint32_t S, P, L, T;
    int version;
} rtb_map;
int rtb_map_load(const char* path, rtb_map* out);
int rtb_map_save(const char* path, const rtb_map* m, int version);
void rtb_map_free(rtb_map* m);

/* Host-only views (no GPU needed; CPU tests) of two things the engine builds at upload time:
 * the xorshift32 jump-ahead table of the pixel-tree kernel -- out[j*32 + i] = image of bit i under (step^(4L))^(2^j), j < 15
 * (primitives.cl:111-125 is linear over GF(2); raytracing.cl:96-113 draws 4L numbers per shaded segment) --
 * and the N4 cluster structure over a wire sphere array: clusters[4*nc] {C, 1.125 R'^2}, records[4*8*nc] {c, r*r} in
 * cluster order, index[8*nc] (original index | transperent << 31, 0xffffffff = padding), groups[4*ng]; nc = ceil(S/8),
 * ng = ceil(nc/8); array pointers may be NULL.  Returns 1 when built, 0 when the scene stays brute force. */
int rtb_debug_jump_table(uint32_t L, uint32_t out[480]);
int rtb_debug_accel_build(const void* spheres, uint32_t S, float* clusters, float* records, uint32_t* index, float* groups,
                          uint32_t* n_clusters, uint32_t* n_groups);
/* ... and the N4 hierarchy over the triangles of extension X1 (built on first use when option "accel" is on), from
 * {v0, e1, e2} records of 12 floats (what tri_prep makes of the wire array): nodes[8*n_nodes] = axis-aligned boxes
 * {centre.xyz, 0, half extents.xyz, 2^-12 of their sum}, level l at level_offset[l] (level 0 = clusters of 8 triangles in a
 * kd-style order -- every node hands the lower half of its capacity, along the longest axis of its centroids, to its
 * lower half --, level l+1 = 8 nodes of level l; stored top level first, each level padded to ntop * 8^(levels-1-l) slots,
 * n_nodes counts the slots); sorted_records[12*8*nc] (padding all zero); index[8*nc] original triangle index (0xffffffff =
 * padding).  Array pointers may be NULL (call once for the sizes).  Returns 1 when built, 0 when the mesh stays brute force. */
int rtb_debug_tri_accel_build(const float* records, uint32_t T, float* nodes, float* sorted_records, uint32_t* index,
                              uint32_t* n_clusters, uint32_t* n_levels, uint32_t level_offset[6], uint32_t* n_nodes);
/* Diagnostic builds only (e.g. -DRT_TREE_CLOCKS): the raw counter words the last launch left behind. */
int rtb_debug_counters(rtb_engine* e, uint64_t out[RTB_NCOUNTERS]);
/* rtb_render_noalloc -- NO COUNTERPART IN THE REFERENCE; added for callers that
 * cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Renders one frame of the
 * engine's resident scene and image arrays (uploaded / bound earlier) with camera `cam` (the 14 floats of rtb_set_camera;
 * NULL = the latched camera) into the CALLER's device buffers d_out (rows*W words, required) and d_hit (optional), on
 * the engine's stream, asynchronously.  The call itself performs no cudaMalloc / cudaFree / allocating library call: the
 * scheduling scratch of the render kernel (cursor block, tile costs, the cost-ordered tile queue) lives in a file-scope
 * `static __device__` array of 2.0 MiB that every GPU loading the library holds until exit.
 * LIMIT: at most 262,144 tiles of 8x4 pixels per call (8,388,608 pixels; 3840x2160 fits) -- larger frames are rejected
 * (-1, rtb_last_error says so): use rtb_render, or split the frame with a partition.
 * The frame is scheduled like a first frame (skeleton trace, no measured-time history, no pixel-tree kernel); the words
 * are those of rtb_render.  One-shot mirror / completion settings are neither used nor consumed.
 * Every call on a device shares the one array (it keeps the last call's contents; the words that must start at zero
 * are cleared at the head of each call): calls that OVERLAP on the GPU -- two engines on one device, two streams --
 * race on it.  Serialise them. */
int rtb_render_noalloc(rtb_engine* e, const float cam[14], uint32_t W, uint32_t H, uint32_t total, const rtb_partition* part,
                       uint32_t* d_out, uint32_t* d_hit);

/* Launch geometry the renderer chose for the last frame (for bench.py's accounting). */
typedef struct {
    uint32_t grid, block, smem_bytes, tiles, resident_scene; /* resident_scene: 1 = geometry staged in shared memory */
    uint32_t tree_grid; /* CTAs of the render_tree kernel launched beside render_frame (0 = none) */
} rtb_launch_info;
int rtb_last_launch(rtb_engine* e, rtb_launch_info* info);

/* FP32 pipe calibration: runs a dependent-FMA kernel for `iters` iterations per thread on the
 * engine's device and returns the measured TFLOP/s (FMA = 2 flop), or a negative value on error. */
double rtb_fp32_peak_tflops(rtb_engine* e, int iters);

#ifdef __cplusplus
}
#endif
#endif
