/* rt_kernels.h -- internal interface between the engine (rt_engine.cu) and the kernels (rt_kernels.cu).
 * Not part of the C ABI (that is include/rtb200.h). */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

/* Wire sizes, reference src/cl/types.cl:4-42 (== src/cpu_obj.h:10-48). */
#define RT_SPHERE_BYTES 96
#define RT_PLANE_BYTES 96
#define RT_LIGHT_BYTES 48
#define RT_MATERIAL_OFFSET 32 /* rmaterial sits at byte 32 of rsphere and rplane */
/* EXTENSION X1 (no reference counterpart): rtriangle = v0, v1, v2 as float4 (48 B) + rmaterial (64 B) */
#define RT_TRIANGLE_BYTES 112
#define RT_TRI_MATERIAL_OFFSET 48

/* Compact geometry ("scene blob"), produced on the device by scene_prep and staged into shared memory
 * by TMA bulk copies.  Units are float4 records:
 *   [geo_off  .. +Lpad)      light  {o.xyz, r*r}          \ one contiguous run of "sphere-like" records;
 *   [         .. +Spad)      sphere {c.xyz, r*r}          /  each count padded to a multiple of 4 with
 *                                                            never-hit records {0,0,0,-inf}
 *   [laux_off .. +L)         light  {rgb*intensity*M_1_PI_F, radius}
 *   [pln_off  .. +4P)        {n.xyz, -}, {p.xyz, -}, {basis0.xyz, texture_scale}, {basis1.xyz, texture_id bits}
 *   [flag_off .. +ceil(S/4)) one uint32 per sphere: material.transperent != 0
 */
struct RtBlobLayout {
    uint32_t geo_off, laux_off, pln_off, flag_off, total; /* in float4 records */
    uint32_t Lpad, Spad;
};
static inline RtBlobLayout rt_blob_layout(uint32_t S, uint32_t P, uint32_t L) {
    RtBlobLayout b;
    b.Lpad = (L + 3u) & ~3u;
    b.Spad = (S + 3u) & ~3u;
    b.geo_off = 0;
    b.laux_off = b.Lpad + b.Spad;
    b.pln_off = b.laux_off + L;
    b.flag_off = b.pln_off + 4 * P;
    b.total = b.flag_off + (S + 3) / 4;
    if (b.total == 0) b.total = 1;
    return b;
}

/* Level-wise pixel trees (rt_trace_tree.cuh): one warp's slab = RT_TREE_PLANES node planes of `cap` 16-byte records,
 * `cap` shaded-list words, then an addition stream of cap * (3 + 2L) records (1 + 2L additions per shaded node plus a
 * PUSH and a POP mark per refraction entry).  In 16-byte records. */
#define RT_TREE_PLANES 8
static inline size_t rt_tree_slab_records(uint32_t cap, uint32_t L) {
    return (size_t)RT_TREE_PLANES * cap + (cap + 3u) / 4u + (size_t)cap * (3u + 2u * L);
}

/* N4 (rt_trace_accel.cuh): spheres per cluster of the acceleration structure */
#define RT_ACCEL_K 8
#define RT_ACCEL_G 8 /* clusters per group (second level) */
#define RT_TACC_LEVELS 6 /* triangle hierarchy: clusters of RT_ACCEL_K triangles, then 8 nodes per parent, up to 8^5 clusters per top node */

struct RtFrameParams {
    /* raygen.cl:5-8 scalar arguments */
    float corner[3], origin[3], up[3], right[3], wf, hf;
    uint32_t W, H, total;
    /* row-band partition (rtb200.h) */
    uint32_t band_rows, parts, part, local_rows;
    /* scene */
    uint32_t S, P, L;
    RtBlobLayout lay;
    const float4* blob;
    const unsigned char* raw_spheres; /* wire arrays: materials are read in place */
    const unsigned char* raw_planes;
    /* extension X1: triangles (not staged: 48 B records {v0, e1, e2} read through L1, broadcast per warp) */
    uint32_t T;
    const float4* tri;
    const unsigned char* raw_tris;
    /* N4 acceleration structure over the spheres (extension; NULL / 0 = brute force as in the reference) */
    const float4* acc_cl;    /* [acc_nc] {C.xyz, 1.125 * R'^2}: inflated bounding sphere of a cluster */
    const float4* acc_rec;   /* [acc_nc * RT_ACCEL_K] {c.xyz, r*r} in cluster order, never-hit padding */
    const uint32_t* acc_idx; /* [acc_nc * RT_ACCEL_K] original sphere index | transperent << 31, 0xffffffff = padding */
    uint32_t acc_nc;
    const float4* acc_grp;   /* [ceil(acc_nc / RT_ACCEL_G)] bounding sphere of RT_ACCEL_G consecutive clusters, same encoding */
    /* image arrays */
    const uchar4* tex; int tw, th, tl;
    const uchar4* sky; int sw, sh;
    /* outputs (compact: local_rows x W) */
    uint32_t* out;
    uint32_t* hit; /* may be NULL */
    uint32_t* mirror; /* may be NULL: second copy of `out` (mapped host memory: the frame lands in the caller's buffer
                         while the kernel runs, 16 bytes per store, no D2H copy afterwards) */
    uint32_t mirror_global; /* 1: mirror is a full H x W frame indexed by global pixel id (a peer GPU's frame buffer) */
    /* scheduling */
    unsigned int* tile_counter; /* zeroed before launch */
    const uint32_t* order;      /* queue slot -> tile (heaviest first), or NULL for row-major order */
    uint32_t tiles_x, tiles_y;
    const uint32_t* cost;       /* per-tile cost estimate (tile_cost), or NULL */
    uint32_t* tile_time;        /* per-tile wall time of THIS frame in units of 1024 SM clocks (zeroed before launch), or NULL */
    unsigned long long* tile_time_sum;            /* sum of tile_time over the frame (zeroed before launch) */
    const unsigned long long* tile_time_sum_prev; /* the previous frame's */
    const uint32_t* tile_time_prev; /* the previous frame's, used as next cost estimate instead of the skeleton trace, or NULL */
    const uint32_t* queue_len;  /* number of queue entries when `order` is set (tiles + 3 per split tile) */
    const uint32_t* n_split;    /* ... of which the first *n_split are heavy (cost >= tree_min_cost): the render_tree kernel's share */
    uint32_t tree_min_cost;     /* 64 = none */
    uint32_t tree_group_max_cost; /* heavy entries below this cost are pooled four to a CTA pass; from it on they run alone */
    uint32_t cost_is_history;   /* the cost buffer still holds the previous frame's measured-time costs (hysteresis) */
    unsigned int* tree_counter; /* zeroed before launch: render_tree's queue cursor */

    uint32_t split32_min_cost;  /* ... into 32 single-pixel entries (64 = never) */
    uint32_t split_min_cost;    /* tiles whose estimate reaches this are split into four 8x1 sub-tiles (64 = never) */
    uint32_t wide_min_records;  /* light+sphere records (padded) from which the WIDE kernel variant is launched */
    int32_t wide_max_live;      /* < 0: automatic break-even; else rounds with <= this many tracing lanes go ray by ray */
    uint32_t coop_min_cost;     /* tiles whose estimate reaches this walk in cooperative mode, the rest lane per pixel */
    /* level-wise pixel trees (rt_trace_tree.cuh): per-warp slabs of tree_records 16-byte records each */
    float4* tree_slab;          /* NULL: never */
    unsigned long long tree_records;
    uint32_t tree_cap;          /* nodes per slab */
    uint32_t tree_slots;        /* slabs allocated, one per CTA */
    uint32_t tree_mode;         /* 0 off; 1 heavy queue entries (cost >= tree_min_cost); 2 every tile; 3 every tile, pixel by pixel (tests) */
    const uint32_t* jump;       /* [15][32] bit matrices (xorshift32^(4L))^(2^j), see rt_engine.cu */
    unsigned long long* counters; /* RTB_NCOUNTERS, only for the counted variant */
    /* completion signal (multi-GPU, fused into the render): when done_flag is set, the LAST CTA to finish stores
     * done_value there with a system-scope release store -- after every store of the frame (band buffer, mirror / peer
     * frame) has been fenced -- so the gathering rank can wait on a word in its own memory instead of a collective */
    unsigned int* done_counter; /* zeroed before launch */
    uint32_t done_expected;     /* CTAs of the frame's kernels (render_frame + render_tree); set by the launcher */
    uint32_t* done_flag;        /* may be NULL; device-visible: a peer GPU's memory (CUDA IPC) or mapped host memory */
    uint32_t done_value;
    /* N4 over the triangles of extension X1 (rt_trace_accel.cuh; NULL / 0 = brute force): axis-aligned boxes {centre,
     * half extents} of two float4 each; level 0 = clusters of RT_ACCEL_K triangles in a kd-style order, level l + 1 = 8
     * consecutive nodes of level l, tacc_nl levels, the top one with tacc_ntop nodes; stored top level first, every
     * level padded to tacc_ntop * 8^(tacc_nl - 1 - l) slots (rt_trace_accel.cuh: tacc_level_base) */
    const float4* tacc_node;
    const float4* tacc_rec;   /* [tacc_nc * RT_ACCEL_K * 3] {v0, e1, e2} in cluster order; padding records are all zero (det == 0: never hit) */
    const uint32_t* tacc_idx; /* [tacc_nc * RT_ACCEL_K] original triangle index */
    uint32_t tacc_nc, tacc_nl, tacc_ntop;
    /* scheduling, launcher side only: the cost buffer the hysteresis of tile_cost_history looks back at when it is not the
     * one being written (the engine keeps two sets of queue buffers, rt_engine.cu: schedule look-ahead); NULL = in place */
    const uint32_t* cost_prev;
};

/* scene_prep: wire arrays -> blob. */
cudaError_t rt_launch_scene_prep(const unsigned char* d_spheres, uint32_t S, const unsigned char* d_planes, uint32_t P,
                                 const unsigned char* d_lights, uint32_t L, float4* d_blob, RtBlobLayout lay,
                                 cudaStream_t st);

struct RtLaunchShape {
    uint32_t grid, block, smem_bytes, resident, tree_grid;
};

/* Launches the fused raygen+raytracer kernel.  mode: 0 = cooperative trace, 1 = literal trace + reference work
 * counters, 2 = cooperative trace + executed-work counters, 3 = per-lane fast trace.  Geometry is staged in shared memory when it fits. */
cudaError_t rt_launch_render(const RtFrameParams& fp, int mode, int sm_count, cudaStream_t st, RtLaunchShape* shape);

/* Cost-ordered tile queue: d_cost[ntiles], d_hist_cursor[130] (zeroed here; word 128 receives the queue length, word 129
 * the number of leading heavy entries),
 * d_order[32*ntiles]. */
cudaError_t rt_launch_tile_order(const RtFrameParams& fp, uint32_t* d_cost, uint32_t* d_hist_cursor, uint32_t* d_order,
                                 cudaStream_t st);

cudaError_t rt_launch_fma_peak(float* d_sink, int iters, int blocks, int threads, cudaStream_t st);
/* d_out[3] (zeroed by the caller): division mismatches, normalize mismatches, max sincos deviation in 2^-24 units */
cudaError_t rt_launch_arith_selftest(uint32_t seed, uint32_t per_thread, uint32_t* d_out, cudaStream_t st);

/* Gathering rank: compact per-part band buffers (parts x stride_words) -> row-major frame. */
cudaError_t rt_launch_assemble_bands(const uint32_t* d_compact, size_t stride_words, uint32_t* d_frame, uint32_t W, uint32_t H,
                                     uint32_t band_rows, uint32_t parts, cudaStream_t st);

/* Gathering rank: spin (on the stream) until flags[i * stride_words] has reached `value` for every i < n (wrap-safe
 * compare); gives up after ~2 s of SM clocks and adds 1 to *d_timeouts instead of hanging the GPU. */
cudaError_t rt_launch_wait_flags(const uint32_t* d_flags, uint32_t stride_words, uint32_t n, uint32_t value,
                                 uint32_t* d_timeouts, cudaStream_t st);

/* Extension X1 builds of the two launchers above (rt_kernels_tri.cu), used when the scene has triangles. */
cudaError_t rt_launch_render_tri(const RtFrameParams& fp, int mode, int sm_count, cudaStream_t st, RtLaunchShape* shape);
/* N4 builds (rt_kernels_accel.cu): the sphere walks go through the cluster structure. */
cudaError_t rt_launch_render_acc(const RtFrameParams& fp, int mode, int sm_count, cudaStream_t st, RtLaunchShape* shape);
cudaError_t rt_launch_tile_order_acc(const RtFrameParams& fp, uint32_t* d_cost, uint32_t* d_hist_cursor, uint32_t* d_order,
                                     cudaStream_t st);
cudaError_t rt_launch_tile_order_tri(const RtFrameParams& fp, uint32_t* d_cost, uint32_t* d_hist_cursor, uint32_t* d_order,
                                     cudaStream_t st);
/* N4 over triangles (rt_kernels_tri_accel.cu): extension X1 build with the triangle walks going through the hierarchy. */
cudaError_t rt_launch_render_tacc(const RtFrameParams& fp, int mode, int sm_count, cudaStream_t st, RtLaunchShape* shape);
cudaError_t rt_launch_tile_order_tacc(const RtFrameParams& fp, uint32_t* d_cost, uint32_t* d_hist_cursor, uint32_t* d_order,
                                      cudaStream_t st);
/* Extension X1: rtriangle wire array -> {v0, v1-v0, v2-v0} records (3 float4 each). */
cudaError_t rt_launch_tri_prep(const unsigned char* d_tris, uint32_t T, float4* d_rec, cudaStream_t st);
