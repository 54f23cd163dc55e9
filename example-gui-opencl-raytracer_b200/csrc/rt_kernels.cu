/* rt_kernels.cu -- the hot path: fused camera-ray generation + Whitted trace + shade + pack, for sm_100a.
 *
 * Behavioural contract = the reference's two OpenCL kernels and their device library
 *   /root/reference/src/cl/raygen.cl:5-25, raytracing.cl:14-195, primitives.cl:14-442, types.cl:4-57.
 * Design (nothing here is a line-by-line translation):
 *   - scene_prep compacts the wire arrays into 16-byte records once per scene (per-plane texture bases,
 *     per-light colour products and r*r hoisted out of the per-ray code -- bit-identical rewrites);
 *   - tile_cost + tile_order: a skeleton trace (no shading) of five rays per 8x4 tile estimates its cost and
 *     a counting sort orders tiles heaviest-first, so the glass/mirror hot spots do not form a tail;
 *   - render_frame: persistent CTAs; the record blob is staged into shared memory with TMA bulk copies
 *     (cp.async.bulk + mbarrier, <= 32 KiB per copy); one warp = one 8x4 tile at a time, tiles pulled from
 *     an atomic queue (prefetched one tile ahead); every intersection test is one broadcast LDS.128; the
 *     current ray lives in registers, the 14 suspended parents in a local-memory stack touched only on
 *     push/pop; the frame buffer is written with 128-bit stores after a warp shuffle transpose.
 *   - two trace variants produce the same words: rt_trace_literal.cuh (reference control flow, used for
 *     accounting) and rt_trace_fast.cuh (production).
 * Compiled with -fmad=false -prec-div=true -prec-sqrt=true -ftz=false (see rt_device.cuh).
 */
#include "rt_kernels.h"

#include "rt_device.cuh"
#include "rt_trace_fast.cuh"
#include "rt_trace_coop.cuh"
#include "rt_trace_tree.cuh"
#include "rt_trace_literal.cuh"

namespace RT_NS {

#ifndef RT_MIN_BLOCKS
#define RT_MIN_BLOCKS 5
#endif
#ifndef RT_WIDE_BLOCKS
#define RT_WIDE_BLOCKS 4 /* the WIDE variant: scenes large enough for it fill shared memory to 4 CTAs per SM anyway; 128 registers, no spills */
#endif
extern __shared__ __align__(128) unsigned char rt_smem[];

#ifndef RT_TREE_BLOCKS
#define RT_TREE_BLOCKS (RT_TRIANGLES ? 4 : 3) /* CTAs per SM the tree kernel is compiled for: 147 registers; the triangle
                                                build takes 128 (no spills), where 16 instead of 12 warps per SM pay:
                                                5k-triangle mesh at 4K 563 -> 531 ms (tools/exp/tri_stage_ab.sh) */
#endif
#ifndef RT_TREE
#define RT_TREE 1 /* 0: compile the level-wise pixel-tree phase out (A/B builds) */
#endif

/* Frame (and hit-id, and mirror) stores of one tree entry by warp 0: lanes 0 and 4 of each tile row store 4 pixels =
 * 16 bytes after a shuffle transpose; single-pixel entries and frames whose rows are not 16-byte aligned store scalars.
 * (The main loop of render_frame has the same code inline.) */
template <bool AOV>
__device__ __forceinline__ void store_tile(const RtFrameParams& fp, const TileLanes& tq, uint32_t sub, uint32_t lane, uint32_t word,
                                           uint32_t hid) {
    const bool vec_store = (fp.W % 4u) == 0u; /* rows are 16-byte aligned */
    const bool live = tq.live;
    const uint32_t id = tq.id;
    const size_t at = tq.at;
    const size_t mat = fp.mirror_global ? (size_t)id : at; /* mirror in full-frame (row-major) or compact layout */
    if (vec_store && sub <= 4u) {
        uint32_t w1 = __shfl_down_sync(0xffffffffu, word, 1);
        uint32_t w2 = __shfl_down_sync(0xffffffffu, word, 2);
        uint32_t w3 = __shfl_down_sync(0xffffffffu, word, 3);
        uint32_t live3 = __shfl_down_sync(0xffffffffu, (uint32_t)live, 3);
        if ((lane & 3u) == 0u && live) {
            if (live3) {
                *reinterpret_cast<uint4*>(fp.out + at) = make_uint4(word, w1, w2, w3);
                if (fp.mirror) __stcs(reinterpret_cast<uint4*>(fp.mirror + mat), make_uint4(word, w1, w2, w3));
            } else { /* id >= total inside this quad: scalar tail */
                fp.out[at] = word;
                if (id + 1 < fp.total) fp.out[at + 1] = w1;
                if (id + 2 < fp.total) fp.out[at + 2] = w2;
                if (fp.mirror) {
                    fp.mirror[mat] = word;
                    if (id + 1 < fp.total) fp.mirror[mat + 1] = w1;
                    if (id + 2 < fp.total) fp.mirror[mat + 2] = w2;
                }
            }
        }
        if (AOV && fp.hit) {
            uint32_t h1 = __shfl_down_sync(0xffffffffu, hid, 1);
            uint32_t h2 = __shfl_down_sync(0xffffffffu, hid, 2);
            uint32_t h3 = __shfl_down_sync(0xffffffffu, hid, 3);
            if ((lane & 3u) == 0u && live) {
                if (live3) *reinterpret_cast<uint4*>(fp.hit + at) = make_uint4(hid, h1, h2, h3);
                else {
                    fp.hit[at] = hid;
                    if (id + 1 < fp.total) fp.hit[at + 1] = h1;
                    if (id + 2 < fp.total) fp.hit[at + 2] = h2;
                }
            }
        }
    } else if (live) {
        fp.out[at] = word;
        if (fp.mirror) fp.mirror[mat] = word;
        if (AOV && fp.hit) fp.hit[at] = hid;
    }
}

/* Completion signal: every thread fences its frame stores (they may target a peer GPU or host memory), the CTA joins,
 * and the last CTA of the frame's kernels (render_frame and, beside it, render_tree: fp.done_expected CTAs in all)
 * publishes the frame number. */
__device__ __forceinline__ void signal_done(const RtFrameParams& fp) {
    if (fp.done_flag) {
        /* The CTA's frame stores are ordered before the barrier; thread 0's fence after it is cumulative (PTX memory
         * model), so one fence per CTA publishes them all -- at DEVICE scope: the counter is a device-scope atomic, and
         * the one system-scope fence the peer / host needs is the last CTA's, before its release store (cumulative again
         * over everything the counter chain made visible to it).  A system-scope fence per CTA cost 0.2 ms per frame once
         * render_tree's CTAs executed theirs in the middle of render_frame's run (tools/exp/n2_probe.py). */
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            const unsigned int prev = atomicAdd(fp.done_counter, 1u);
            if (prev + 1u == fp.done_expected) {
                __threadfence_system();
                asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(fp.done_flag), "r"(fp.done_value) : "memory");
            }
        }
    }
}

/* Per CTA: TMA-stage the record blob into shared memory (when resident), fill the /255 table, bind the geometry. */
template <bool RESIDENT>
__device__ __forceinline__ void stage_geometry(const RtFrameParams& fp, uint64_t* bar, float* lut255, Geo& g) {
    if (RESIDENT) {
        const uint32_t bytes = fp.lay.total * 16u;
        if (threadIdx.x == 0) mbar_init(bar, 1);
        __syncthreads();
        if (threadIdx.x == 0) {
            mbar_expect_tx(bar, bytes);
            const uint32_t kChunk = 32768u; /* the object array travels in chunks of <= 32 KiB */
            for (uint32_t off = 0; off < bytes; off += kChunk) {
                uint32_t nb = bytes - off < kChunk ? bytes - off : kChunk;
                tma_bulk_g2s(rt_smem + off, reinterpret_cast<const unsigned char*>(fp.blob) + off, nb, bar);
            }
        }
    }
    for (uint32_t i = threadIdx.x; i < 256u; i += kBlock) lut255[i] = (float)i / 255.0f; /* overlaps the copy */
    if (RESIDENT) {
        mbar_wait(bar, 0);
        geo_bind(g, reinterpret_cast<const float4*>(rt_smem), fp, lut255);
    } else {
        geo_bind(g, fp.blob, fp, lut255);
    }
    __syncthreads();
}

/* ---- render_tree: the heavy entries of the queue as level-wise pixel trees, one CTA per entry at a time
 * (rt_trace_tree.cuh).  A kernel of its own, launched just before render_frame in the same stream: sharing a kernel cost
 * render_frame's main loop 8 % (2.65 -> 2.87 ms per 4K frame) even with the phase switched off, and more than that as
 * an out-of-line call (1.1 KB of spills).  An entry whose
 * tree outgrows the slab is walked by warp 0 with the cooperative trace -- same words. ---- */
template <bool RESIDENT, bool AOV>
__global__ void __launch_bounds__(kBlock, RT_TREE_BLOCKS) render_tree(const __grid_constant__ RtFrameParams fp) {
    __shared__ __align__(8) uint64_t bar;
    __shared__ float lut255[256];
    __shared__ TreeShared tsh;
    /* Programmatic dependent launch: render_frame follows in the same stream and may start as soon as every CTA of this
     * grid has got here, i.e. IS RESIDENT -- the heaviest entries of the frame hold their SM slots first, and
     * render_frame's persistent CTAs fill what is left and, later, what these CTAs vacate.  (On a second stream the two
     * kernels raced for the slots: when render_frame won, the heavy entries waited for it to drain -- frames alternating
     * 0.57 / 0.87 ms at N = 8.) */
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    Geo g;
    stage_geometry<RESIDENT>(fp, &bar, lut255, g);
    const uint32_t lane = threadIdx.x & 31u;
    CoopSmem cs; /* scratch of the fallback walk (warp 0 only) */
    {
        unsigned char* base = rt_smem + (RESIDENT ? fp.lay.total * 16u : 0u);
        cs.owner = reinterpret_cast<float4*>(base);
        cs.task = cs.owner + 32;
        cs.light = cs.task + kCoopTasks;
        cs.pre = cs.light + 32 * kCoopLights;
    }
    TriStage stg; /* extension X1 build: chunked TMA staging of the triangle records (behind the fallback walk's scratch) */
    stg.buf = nullptr; stg.bar = nullptr; stg.phase = 0;
#if RT_TRIANGLES
    tri_stage_init(stg, rt_smem + (((RESIDENT ? fp.lay.total * 16u : 0u) + kCoopBytesPerWarp + 15u) & ~15u));
#endif
    const uint32_t ntiles = fp.order ? __ldg(fp.queue_len) : fp.tiles_x * fp.tiles_y;
    const uint32_t n_tree = fp.tree_mode >= 2u ? ntiles : (fp.order ? __ldg(fp.n_split) : 0u);
    const TreeSlab sl = {fp.tree_slab + (size_t)blockIdx.x * fp.tree_records, fp.tree_cap};
    unsigned long long warp_time = 0;
    const uint32_t warp = threadIdx.x >> 5;
    for (;;) {
        /* the heaviest entries one at a time (their chains bound the frame); the lighter ones four to a pass: a lone
         * tile of a few hundred segments leaves most of the 128 lanes idle at every level of its tree */
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t slot = atomicAdd(fp.tree_counter, 1u), n = 0;
            if (slot < n_tree) {
                tsh.entry[n++] = fp.order ? __ldg(fp.order + slot) : slot;
                const uint32_t e0 = tsh.entry[0];
                const bool alone = fp.tree_mode == 3u || (fp.cost && __ldg(fp.cost + (e0 & 0x00ffffffu)) >= fp.tree_group_max_cost);
                if (!alone) {
                    slot = atomicAdd(fp.tree_counter, kTreeGroup - 1u);
                    for (uint32_t k = 0; k < kTreeGroup - 1u && slot + k < n_tree; k++) tsh.entry[n++] = fp.order ? __ldg(fp.order + slot + k) : slot + k;
                }
            }
            tsh.count = n;
        }
        __syncthreads();
        const uint32_t ne = tsh.count;
        if (ne == 0u) break;
        /* one pooled pass; should it outgrow the slab, entry by entry; and an entry that does not fit alone is walked
         * by warp 0 with the cooperative trace */
        uint32_t first = 0, span = ne;
        while (first < ne) {
            const long long clk0 = clock64();
            /* "tree" = 3 (tests): every pixel of the entry on its own, as single-pixel queue entries are; else one pass */
            bool traced = true;
            const uint32_t rounds = fp.tree_mode == 3u ? 32u : 1u;
#pragma unroll 1
            for (uint32_t p = 0; p < rounds && traced; p++) {
                if (rounds > 1u && !tile_lanes(fp, tsh.entry[first], p).live) continue;
                traced = trace_entry_tree<AOV>(fp, g, sl, tsh, first, span, rounds > 1u ? p : 32u, stg);
            }
            if (!traced && span > 1u) { span = 1u; __syncthreads(); continue; } /* CTA-uniform; the barrier keeps the retry off the scratch the failed pass is still reading */
            if (warp < span) { /* warp w stores entry first + w, lane p = tile pixel p */
                const uint32_t entry = tsh.entry[first + warp];
                const TileLanes tq = tile_lanes(fp, entry, lane);
                uint32_t word = 0, hid = 0;
                if (traced) {
                    if (tq.live) { word = tsh.word[threadIdx.x]; hid = tsh.hid[threadIdx.x]; }
                } else { /* span == 1, warp 0: the tree did not fit the slab */
                    Tally<false> tl;
                    uint32_t nseg = 0;
                    word = trace_tile_coop<false, AOV, false>(fp, g, cs, tq.id, tq.live, lane, hid, nseg, tl);
                    __syncwarp();
                }
                if (fp.tile_time && lane == 0) {
                    /* warp-time units, like render_frame; a pooled pass is shared out by shaded segments */
                    uint32_t dt = (uint32_t)((clock64() - clk0) >> 10) * (traced ? kBlock / 32 : 1u);
                    if (traced && span > 1u) {
                        uint32_t tot = 0;
                        for (uint32_t k = 0; k < span; k++) tot += tsh.shaded_of[k] + 1u;
                        dt = (uint32_t)(((unsigned long long)dt * (tsh.shaded_of[warp] + 1u)) / tot);
                    }
                    dt += 1u;
                    atomicAdd(fp.tile_time + (entry & 0x00ffffffu), dt);
                    warp_time += dt;
                }
                store_tile<AOV>(fp, tq, entry >> 24, lane, word, hid);
            }
            first += span;
            __syncthreads(); /* tsh.word / shaded_of are read above and rewritten by the next pass */
        }
    }
    if (fp.tile_time && lane == 0 && warp_time) atomicAdd(fp.tile_time_sum, warp_time);
    signal_done(fp);
}

/* MODE: 0 = cooperative trace (production), 1 = literal + counters (reference accounting), 2 = cooperative +
 * counters (executed work), 3 = per-lane fast trace (the previous production path, kept for A/B),
 * 4 = cooperative trace, WIDE variant (rt_trace_coop.cuh: one ray over 32 lanes; scenes with many spheres) */
template <bool RESIDENT, int MODE, bool AOV>
__global__ void __launch_bounds__(kBlock, MODE == 4 ? RT_WIDE_BLOCKS : RT_MIN_BLOCKS) render_frame(const __grid_constant__ RtFrameParams fp) {
    constexpr bool COUNT = MODE == 1 || MODE == 2;
    constexpr bool COOP = MODE == 0 || MODE == 2 || MODE == 4;
    constexpr bool WIDE = MODE == 4;
    constexpr bool TREE = MODE == 0 || MODE == 4; /* heavy entries may run as level-wise pixel trees (phase 1) */
    __shared__ __align__(8) uint64_t bar;
    __shared__ float lut255[256];
    Geo g;
    stage_geometry<RESIDENT>(fp, &bar, lut255, g);

    const uint32_t lane = threadIdx.x & 31u;
    CoopSmem cs;
    {
        unsigned char* base = rt_smem + (RESIDENT ? fp.lay.total * 16u : 0u) + (threadIdx.x >> 5) * kCoopBytesPerWarp;
        cs.owner = reinterpret_cast<float4*>(base);
        cs.task = cs.owner + 32;
        cs.light = cs.task + kCoopTasks;
        cs.pre = cs.light + 32 * kCoopLights;
    }
    /* queue length: every tile once, plus 3 extra entries per split tile (tile_order writes the total) */
    const uint32_t ntiles = fp.order ? __ldg(fp.queue_len) : fp.tiles_x * fp.tiles_y;
    const bool vec_store = (fp.W % 4u) == 0u; /* rows are 16-byte aligned */
    Tally<COUNT> tl;
    tl.clear();

    unsigned long long warp_time = 0; /* sum of this warp's tile times: the frame total gives the mean the next frame's costs are relative to */

    /* The first n_tree entries of the cost-ordered queue (tiles at or above fp.tree_min_cost and the split entries of the
     * heaviest ones) belong to the render_tree kernel, which was launched just before this one and runs beside it. */
    const uint32_t n_tree = (RT_TREE && TREE && fp.tree_mode) ? (fp.tree_mode >= 2u ? ntiles : (fp.order ? __ldg(fp.n_split) : 0u)) : 0u;
    uint32_t next = 0;
    if (lane == 0) next = atomicAdd(fp.tile_counter, 1u) + n_tree;
    for (;;) {
        const uint32_t slot = __shfl_sync(0xffffffffu, next, 0);
        if (slot >= ntiles) break;
        if (lane == 0) next = atomicAdd(fp.tile_counter, 1u) + n_tree; /* prefetch: latency hides behind this tile */
        const uint32_t entry = fp.order ? __ldg(fp.order + slot) : slot;
        const uint32_t tile = entry & 0x00ffffffu;
        const uint32_t sub = entry >> 24; /* 0: whole 8x4 tile; 1..4: only tile row sub-1; 5..36: only pixel sub-5 (split heavy tiles) */

        const uint32_t tx = tile % fp.tiles_x, ty = tile / fp.tiles_x;
        const uint32_t x = tx * kTileW + (lane & 7u);
        const uint32_t lr = ty * kTileH + (lane >> 3);
        const uint32_t y = ((lr / fp.band_rows) * fp.parts + fp.part) * fp.band_rows + lr % fp.band_rows;
        const uint32_t id = y * fp.W + x;
        const bool live = x < fp.W && lr < fp.local_rows && y < fp.H && id < fp.total && (sub == 0u || (sub <= 4u ? (lane >> 3) + 1u == sub : lane + 5u == sub));

        const long long tile_clk0 = clock64();
        uint32_t word = 0, hid = 0, nseg = 0;
#ifdef RT_TILE_TIMING /* diagnostic build: the hit-id AOV carries each tile's wall time in ns (tools/exp/tile_times.py) */
        unsigned long long t_begin;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_begin));
#endif
        /* hybrid: the cooperative walk shortens the heavy tiles (glass, mirrors) that bound the kernel's tail; the
         * per-lane walk has less lock-step overhead on the cheap majority.  Same words either way. */
        const bool heavy = COOP && (!fp.cost || __ldg(fp.cost + tile) >= fp.coop_min_cost);
        if (heavy) {
            word = trace_tile_coop<COUNT, AOV, WIDE>(fp, g, cs, id, live, lane, hid, nseg, tl);
        } else if (live) {
            if (MODE == 1) word = trace_pixel_literal<COUNT, AOV>(fp, g, id, hid, tl);
            else word = trace_pixel_fast<COUNT, AOV>(fp, g, id, hid, tl);
        }
        __syncwarp();
        /* measured cost of this tile in units of 1024 SM clocks (sub-tiles add up): the next frame's scheduling hint.
         * (Segments traced -- nseg -- were tried as a split-invariant metric: worse order, tools/exp/sweep_split.py.) */
        if (fp.tile_time && lane == 0) {
            const uint32_t dt = (uint32_t)((clock64() - tile_clk0) >> 10) + 1u;
            atomicAdd(fp.tile_time + tile, dt);
            warp_time += dt;
        }
#ifdef RT_TILE_TIMING
        {
            unsigned long long t_end;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_end));
            /* start time and duration in units of 1.024 us: [31:16] start (mod 67 ms), [15:0] duration */
            unsigned long long dur = (t_end - t_begin) >> 10;
            hid = (uint32_t)(((t_begin >> 10) & 0xffffull) << 16) | (uint32_t)(dur > 0xffffull ? 0xffffull : dur);
        }
#endif

        const size_t at = (size_t)lr * fp.W + x;
        const size_t mat = fp.mirror_global ? (size_t)id : at; /* mirror in full-frame (row-major) or compact layout */
        if (vec_store && sub <= 4u) { /* lanes 0 and 4 of each tile row store 4 pixels = 16 bytes (single-pixel entries: scalar) */
            uint32_t w1 = __shfl_down_sync(0xffffffffu, word, 1);
            uint32_t w2 = __shfl_down_sync(0xffffffffu, word, 2);
            uint32_t w3 = __shfl_down_sync(0xffffffffu, word, 3);
            uint32_t live3 = __shfl_down_sync(0xffffffffu, (uint32_t)live, 3);
            if ((lane & 3u) == 0u && live) {
                if (live3) {
                    *reinterpret_cast<uint4*>(fp.out + at) = make_uint4(word, w1, w2, w3);
                    if (fp.mirror) __stcs(reinterpret_cast<uint4*>(fp.mirror + mat), make_uint4(word, w1, w2, w3));
                } else { /* id >= total inside this quad: scalar tail */
                    fp.out[at] = word;
                    if (id + 1 < fp.total) fp.out[at + 1] = w1;
                    if (id + 2 < fp.total) fp.out[at + 2] = w2;
                    if (fp.mirror) {
                        fp.mirror[mat] = word;
                        if (id + 1 < fp.total) fp.mirror[mat + 1] = w1;
                        if (id + 2 < fp.total) fp.mirror[mat + 2] = w2;
                    }
                }
            }
            if (AOV && fp.hit) {
                uint32_t h1 = __shfl_down_sync(0xffffffffu, hid, 1);
                uint32_t h2 = __shfl_down_sync(0xffffffffu, hid, 2);
                uint32_t h3 = __shfl_down_sync(0xffffffffu, hid, 3);
                if ((lane & 3u) == 0u && live) {
                    if (live3) *reinterpret_cast<uint4*>(fp.hit + at) = make_uint4(hid, h1, h2, h3);
                    else {
                        fp.hit[at] = hid;
                        if (id + 1 < fp.total) fp.hit[at + 1] = h1;
                        if (id + 2 < fp.total) fp.hit[at + 2] = h2;
                    }
                }
            }
        } else if (live) {
            fp.out[at] = word;
            if (fp.mirror) fp.mirror[mat] = word;
            if (AOV && fp.hit) fp.hit[at] = hid;
        }
    }

    if (fp.tile_time && lane == 0 && warp_time) atomicAdd(fp.tile_time_sum, warp_time);

    if (COUNT) {
#pragma unroll
        for (int k = 0; k < RTB_NCOUNTERS; k++) {
            unsigned long long s = tl.get(k);
            for (int off = 16; off > 0; off >>= 1) s += __shfl_down_sync(0xffffffffu, s, off);
            if (lane == 0 && s) atomicAdd(fp.counters + k, s);
        }
    }

    signal_done(fp);
}

/* ---- scene_prep: wire arrays -> compact blob (one thread per record slot) ---- */
__global__ void scene_prep(const unsigned char* __restrict__ spheres, uint32_t S, const unsigned char* __restrict__ planes,
                           uint32_t P, const unsigned char* __restrict__ lights, uint32_t L, float4* __restrict__ blob,
                           RtBlobLayout lay) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const float4 never = make_float4(0.0f, 0.0f, 0.0f, -CUDART_INF_F); /* D = b*b - 4a*(|v|^2 + inf) < 0 for every finite ray */
    float4* sph = blob + lay.geo_off + lay.Lpad;
    if (i < S) {
        const float4* q = reinterpret_cast<const float4*>(spheres + (size_t)i * RT_SPHERE_BYTES);
        float4 c = q[0];
        float r = q[1].x;
        sph[i] = make_float4(c.x, c.y, c.z, r * r);
        uint32_t transp = __float_as_uint(q[4].x); /* material @32: rgb, {amb,dif,spe,shin}, {transperent,...} */
        reinterpret_cast<uint32_t*>(blob + lay.flag_off)[i] = transp ? 1u : 0u;
    } else if (i < lay.Spad) {
        sph[i] = never;
    }
    if (i < P) {
        const float4* q = reinterpret_cast<const float4*>(planes + (size_t)i * RT_PLANE_BYTES);
        float4 n4 = q[0], p4 = q[1], m3 = q[5];
        V3 n = xyz(n4);
        /* primitives.cl:217-241: first unit axis e with sum(cross(e, n)) != 0; zero basis if none (Q16) */
        V3 b0 = mk(0.0f, 0.0f, 0.0f), b1 = mk(0.0f, 0.0f, 0.0f);
        const V3 axes[3] = {{1.0f, 0.0f, 0.0f}, {0.0f, 1.0f, 0.0f}, {0.0f, 0.0f, 1.0f}};
        for (int k = 0; k < 3; k++) {
            V3 cr = cross3(axes[k], n);
            if (dot3(mk(1.0f, 1.0f, 1.0f), cr) == 0.0f) continue;
            b0 = cr;
            b1 = cross3(n, cr);
            break;
        }
        float4* out = blob + lay.pln_off + 4 * i;
        out[0] = make_float4(n.x, n.y, n.z, 0.0f);
        out[1] = make_float4(p4.x, p4.y, p4.z, 0.0f);
        out[2] = make_float4(b0.x, b0.y, b0.z, m3.y); /* texture_scale */
        out[3] = make_float4(b1.x, b1.y, b1.z, m3.x); /* texture_id bits */
    }
    if (i < L) {
        const float4* q = reinterpret_cast<const float4*>(lights + (size_t)i * RT_LIGHT_BYTES);
        float4 o = q[0], ri = q[1], c = q[2];
        float radius = ri.x, intensity = ri.y;
        V3 pre = (xyz(c) * intensity) * kInvPi; /* rgb*intensity*M_1_PI_F, left to right */
        blob[lay.geo_off + i] = make_float4(o.x, o.y, o.z, radius * radius);
        blob[lay.laux_off + i] = make_float4(pre.x, pre.y, pre.z, radius);
    } else if (i < lay.Lpad) {
        blob[lay.geo_off + i] = never;
    }
    if (i == 0 && S + P + L == 0) blob[0] = make_float4(0, 0, 0, 0);
}

/* ---- tri_prep (extension X1): rtriangle wire records -> {v0, e1 = v1 - v0, e2 = v2 - v0} ---- */
__global__ void tri_prep(const unsigned char* __restrict__ tris, uint32_t T, float4* __restrict__ rec) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= T) return;
    const float4* q = reinterpret_cast<const float4*>(tris + (size_t)i * RT_TRIANGLE_BYTES);
    const V3 v0 = xyz(q[0]), e1 = xyz(q[1]) - v0, e2 = xyz(q[2]) - v0;
    rec[3 * i + 0] = make_float4(v0.x, v0.y, v0.z, 0.0f);
    rec[3 * i + 1] = make_float4(e1.x, e1.y, e1.z, 0.0f);
    rec[3 * i + 2] = make_float4(e2.x, e2.y, e2.z, 0.0f);
}

/* ---- tile_cost: skeleton trace (no shading, no refraction children) of five rays per 2x2 group of tiles.
 * A scheduling hint only: it cannot change any pixel. ---- */
constexpr uint32_t kCostBuckets = 64;
/* queue entries per tile: whole tile, four 8-pixel rows, or 32 single pixels (the other lanes only run shadow tasks) */
__device__ __forceinline__ uint32_t split_ways(uint32_t cost, uint32_t split4, uint32_t split32) {
    return cost >= split32 ? 32u : (cost >= split4 ? 4u : 1u);
}

__device__ uint32_t skeleton_cost(const RtFrameParams& fp, const Geo& g, uint32_t id) {
    V3 o, d;
    primary_ray(fp, id, o, d);
    Tally<false> tl;
    uint32_t cost = 0;
    for (int depth = 0; depth < kMaxDepth && cost < kCostBuckets; depth++) {
        RayF r;
        ray_setup_fast(r, o, d);
        Closest c;
        closest_pass(r, g, tl, c);
        if ((c.k >= 0 && !c.occluded) || (c.who < 0 && c.whop < 0 && c.whot < 0)) break;
        V3 ip = o + d * c.t, nrm;
        Material m;
        if (c.whot >= 0) {
            nrm = triangle_normal(g.tri + 3 * c.whot);
            m = load_material_at(fp.raw_tris + (size_t)c.whot * RT_TRIANGLE_BYTES + RT_TRI_MATERIAL_OFFSET);
        } else if (c.whop >= 0) {
            nrm = xyz(g.pln[4 * c.whop]);
            m = load_material(fp.raw_planes + (size_t)c.whop * RT_PLANE_BYTES);
        } else {
            float ql;
            nrm = normalize_len(ip - xyz(g.sph[c.who]), ql);
            m = load_material(fp.raw_spheres + (size_t)c.who * RT_SPHERE_BYTES);
        }
        const bool shaded = (m.diffuse != 0.0f) || (m.specular != 0.0f);
        cost += shaded ? 4u : 1u;
        if (m.transperent) cost += 12u; /* a refraction subtree hangs off this hit */
        ip = ip + nrm * kEps;
        float cosI = -dot3(nrm, d);
        d = d + nrm * (2.0f * cosI);
        o = ip;
    }
    return cost < kCostBuckets ? cost : kCostBuckets - 1u;
}

__global__ void tile_cost(const __grid_constant__ RtFrameParams fp, uint32_t* __restrict__ cost_out,
                          uint32_t* __restrict__ hist) {
    __shared__ float lut255[256];
    __shared__ uint32_t shist[kCostBuckets];
    for (uint32_t i = threadIdx.x; i < kCostBuckets; i += blockDim.x) shist[i] = 0;
    __syncthreads();
    Geo g;
    geo_bind(g, fp.blob, fp, lut255);
    const uint32_t gx = (fp.tiles_x + 1) / 2, gy = (fp.tiles_y + 1) / 2;
    const uint32_t grp = blockIdx.x * blockDim.x + threadIdx.x;
    if (grp < gx * gy) {
        const uint32_t bx = grp % gx, by = grp / gx;
        const uint32_t px[5] = {0, 2 * kTileW - 1, 0, 2 * kTileW - 1, kTileW};
        const uint32_t py[5] = {0, 0, 2 * kTileH - 1, 2 * kTileH - 1, kTileH};
        uint32_t cost = 0;
        for (int s = 0; s < 5; s++) {
            uint32_t x = bx * 2 * kTileW + px[s], lr = by * 2 * kTileH + py[s];
            if (x >= fp.W) x = fp.W - 1;
            if (lr >= fp.local_rows) lr = fp.local_rows - 1;
            const uint32_t y = ((lr / fp.band_rows) * fp.parts + fp.part) * fp.band_rows + lr % fp.band_rows;
            if (y >= fp.H) continue;
            uint32_t c = skeleton_cost(fp, g, y * fp.W + x);
            cost = c > cost ? c : cost;
        }
        uint32_t n = 0;
        for (uint32_t dy = 0; dy < 2; dy++)
            for (uint32_t dx = 0; dx < 2; dx++) {
                const uint32_t tx = bx * 2 + dx, ty = by * 2 + dy;
                if (tx < fp.tiles_x && ty < fp.tiles_y) { cost_out[ty * fp.tiles_x + tx] = cost; n++; }
            }
        atomicAdd(&shist[cost], split_ways(cost, fp.split_min_cost, fp.split32_min_cost) * n);
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < kCostBuckets; i += blockDim.x)
        if (shist[i]) atomicAdd(&hist[i], shist[i]);
}

/* ---- tile_cost_history: the previous frame's measured tile times as cost, RELATIVE to that frame's mean tile time:
 * cost = 24 + 4*log2(t / mean), so the mean tile sits at 24, 16x the mean at 40, 64x at 48 -- whatever the scene's
 * absolute speed (render.map: mean 29 us; the 1,024-sphere grid: 4.5 ms).  Replaces the skeleton trace whenever a frame of the same geometry has just been rendered. ---- */
__global__ void tile_cost_history(const uint32_t* __restrict__ prev, const unsigned long long* __restrict__ prev_sum,
                                  const uint32_t* cost_prev /* the costs the hysteresis looks back at: cost_out itself, or the other set's */,
                                  uint32_t* cost_out,
                                  uint32_t* __restrict__ hist, uint32_t ntiles, uint32_t split_min_cost, uint32_t split32_min_cost,
                                  uint32_t sticky_cost, uint32_t* __restrict__ time_now /* 4 header words + ntiles: zeroed here instead of by a memset */) {
    __shared__ uint32_t shist[kCostBuckets];
    for (uint32_t i = threadIdx.x; i < kCostBuckets; i += blockDim.x) shist[i] = 0;
    __syncthreads();
    const uint32_t tile = blockIdx.x * blockDim.x + threadIdx.x;
    if (time_now) {
        if (tile < 4u) time_now[tile] = 0u;
        if (tile < ntiles) time_now[4u + tile] = 0u;
    }
    if (tile < ntiles) {
        const float mean = fmaxf((float)(*prev_sum) / (float)ntiles, 1.0f);
        const float rel = 24.0f + 4.0f * __log2f(fmaxf((float)prev[tile], 0.5f) / mean);
        uint32_t c = rel <= 0.0f ? 0u : (uint32_t)rel;
        c = c < kCostBuckets ? c : kCostBuckets - 1u;
        /* hysteresis: a tile that the pixel-tree kernel took last frame (cost_out still holds that frame's costs) keeps the
         * cost it entered with until its time there has dropped two octaves below the threshold.  A tile whose tree pass
         * is cheaper than its cooperative walk would otherwise change sides every frame; and the kept cost is what
         * orders the heavy tiles. */
        if (sticky_cost < kCostBuckets && c + 8u >= sticky_cost) {
            const uint32_t old = cost_prev[tile];
            if (old >= sticky_cost && old < kCostBuckets) c = old;
        }
        cost_out[tile] = c;
        atomicAdd(&shist[c], split_ways(c, split_min_cost, split32_min_cost));
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < kCostBuckets; i += blockDim.x)
        if (shist[i]) atomicAdd(&hist[i], shist[i]);
}

/* counting sort, heaviest bucket first; order within a bucket is arbitrary (it cannot affect any pixel).
 * One global atomic per (block, bucket): ranks inside a block come from shared-memory atomics. */
__global__ void tile_order(const uint32_t* __restrict__ cost, const uint32_t* __restrict__ hist,
                           uint32_t* __restrict__ cursor, uint32_t* __restrict__ order, uint32_t ntiles,
                           uint32_t split_min_cost, uint32_t split32_min_cost, uint32_t first_heavy, uint32_t* __restrict__ queue_len,
                           uint32_t* __restrict__ sched /* 64 words: the render kernels' cursors and counters, zeroed here */) {
    __shared__ uint32_t base[kCostBuckets], local[kCostBuckets], gbase[kCostBuckets];
    __shared__ bool last_block;
    if (sched && blockIdx.x == 0 && threadIdx.x < 64u) sched[threadIdx.x] = 0u;
    for (uint32_t i = threadIdx.x; i < kCostBuckets; i += blockDim.x) local[i] = 0;
    if (threadIdx.x == 0) {
        uint32_t acc = 0, nheavy = 0;
        for (int b = (int)kCostBuckets - 1; b >= 0; b--) {
            base[b] = acc;
            acc += hist[b];
            if ((uint32_t)b >= first_heavy) nheavy = acc; /* the entries of the heavy buckets form a prefix of the queue */
        }
        if (blockIdx.x == 0) { queue_len[0] = acc; queue_len[1] = nheavy; }
    }
    __syncthreads();
    const uint32_t tile = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t c = 0, rank = 0;
    if (tile < ntiles) {
        c = cost[tile];
        rank = atomicAdd(&local[c], split_ways(c, split_min_cost, split32_min_cost));
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < kCostBuckets; i += blockDim.x)
        gbase[i] = local[i] ? atomicAdd(&cursor[i], local[i]) : 0u;
    __syncthreads();
    if (tile < ntiles) {
        uint32_t* o = order + base[c] + gbase[c] + rank;
        const uint32_t ways = split_ways(c, split_min_cost, split32_min_cost);
        if (ways == 32u) { /* the heaviest tiles: one pixel per warp, 31 helper lanes */
            for (uint32_t k = 0; k < 32u; k++) o[k] = tile | ((5u + k) << 24);
        } else if (ways == 4u) { /* a heavy tile is dealt to four warps, one tile row (8 pixels + 24 helper lanes) each */
            o[0] = tile | (1u << 24); o[1] = tile | (2u << 24); o[2] = tile | (3u << 24); o[3] = tile | (4u << 24);
        } else o[0] = tile;
    }
    /* leave the histogram and the cursors clean for the next frame (no memset per frame): the last block to get here */
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        last_block = atomicAdd(queue_len + 2, 1u) + 1u == gridDim.x;
    }
    __syncthreads();
    if (last_block) {
        uint32_t* h = const_cast<uint32_t*>(hist);
        for (uint32_t i = threadIdx.x; i < kCostBuckets; i += blockDim.x) { h[i] = 0u; cursor[i] = 0u; }
        if (threadIdx.x == 0) queue_len[2] = 0u;
    }
}

/* Row-band re-interleave on the gathering rank: `compact` holds `parts` band buffers of
 * `stride` words each (what the gather delivers); frame row y comes from part (y / band_rows) % parts.
 * One uint4 (4 pixels) per thread when W % 4 == 0. */
__global__ void assemble_bands(const uint32_t* __restrict__ compact, size_t stride, uint32_t* __restrict__ frame,
                               uint32_t W, uint32_t H, uint32_t band_rows, uint32_t parts) {
    const uint32_t wq = (W + 3u) / 4u;
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)wq * H) return;
    const uint32_t y = (uint32_t)(i / wq), xq = (uint32_t)(i % wq);
    const uint32_t b = y / band_rows, r = b % parts;
    const uint32_t lr = (b / parts) * band_rows + y % band_rows;
    const uint32_t* src = compact + (size_t)r * stride + (size_t)lr * W + 4u * xq;
    uint32_t* dst = frame + (size_t)y * W + 4u * xq;
    if ((W & 3u) == 0u && (stride & 3u) == 0u) *reinterpret_cast<uint4*>(dst) = __ldcs(reinterpret_cast<const uint4*>(src));
    else for (uint32_t k = 0; k < 4u && 4u * xq + k < W; k++) dst[k] = src[k];
}

/* The gathering rank's side of the completion signal: lane i polls flag i (written by rank i's render_frame through
 * NVLink / PCIe) with system-scope acquire loads.  Bounded: ~4e9 SM clocks, then the timeout word is bumped. */
__global__ void wait_flags(const uint32_t* flags, uint32_t stride, uint32_t n, uint32_t value, uint32_t* timeouts) {
    const uint32_t i = threadIdx.x;
    if (i >= n) return;
    const uint32_t* p = flags + (size_t)i * stride;
    const long long t0 = clock64();
    for (;;) {
        uint32_t v;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
        if ((int32_t)(v - value) >= 0) break;
        if (clock64() - t0 > 4000000000ll) { atomicAdd(timeouts, 1u); break; }
        __nanosleep(200);
    }
}

/* FP32 pipe calibration: 8 independent FMA chains per thread. */
__global__ void fma_peak(float* sink, int iters) {
    float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1.0f, a2 = a0 + 2.0f, a3 = a0 + 3.0f;
    float a4 = a0 + 4.0f, a5 = a0 + 5.0f, a6 = a0 + 6.0f, a7 = a0 + 7.0f;
    const float m = 0.999f, c = 1e-3f + (float)blockIdx.x * 1e-9f;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 8; u++) {
            a0 = __fmaf_rn(a0, m, c); a1 = __fmaf_rn(a1, m, c); a2 = __fmaf_rn(a2, m, c); a3 = __fmaf_rn(a3, m, c);
            a4 = __fmaf_rn(a4, m, c); a5 = __fmaf_rn(a5, m, c); a6 = __fmaf_rn(a6, m, c); a7 = __fmaf_rn(a7, m, c);
        }
    }
    float s = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
    if (s == 123.456f) sink[0] = s;
}

/* Arithmetic self-test: the shared-reciprocal division / fast normalize / bounded sincos against the IEEE
 * operators, on pseudo-random operands.  out[0] += mismatching divisions, out[1] += mismatching normalize
 * components or lengths, out[2] = max |sincos_bounded - sincosf| in units of 2^-24 (as uint). */
__global__ void arith_selftest(uint32_t seed, uint32_t per_thread, uint32_t* out) {
    uint32_t s = seed ^ (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u;
    if (s == 0) s = 1;
    uint32_t bad_div = 0, bad_nrm = 0, worst = 0;
    for (uint32_t it = 0; it < per_thread; it++) {
        /* operands with exponents spread over [2^-40, 2^40] and full random mantissas */
        float v[4];
        for (int k = 0; k < 4; k++) {
            s = xorshift_step(s);
            uint32_t mant = s & 0x007fffffu, sign = s & 0x80000000u;
            s = xorshift_step(s);
            uint32_t e = 127u - 40u + (s % 81u);
            if ((it & 7u) == 0u) e = 127u - 2u + (s % 5u); /* a share of near-unit operands, as in the renderer */
            v[k] = __uint_as_float(sign | (e << 23) | mant);
        }
        const float b = v[3];
        if (div_guard(b)) {
            float r = refine_rcp(b);
            for (int k = 0; k < 3; k++)
                if (div_guard(v[k]) && __float_as_uint(div_with_rcp(v[k], b, r)) != __float_as_uint(v[k] / b)) bad_div++;
        }
        float len;
        V3 q = normalize_len(mk(v[0], v[1], v[2]), len);
        float dd = dot3(mk(v[0], v[1], v[2]), mk(v[0], v[1], v[2]));
        float len_ref = sqrtf(dd);
        if (__float_as_uint(len) != __float_as_uint(len_ref)) bad_nrm++;
        if (__float_as_uint(q.x) != __float_as_uint(v[0] / len_ref)) bad_nrm++;
        if (__float_as_uint(q.y) != __float_as_uint(v[1] / len_ref)) bad_nrm++;
        if (__float_as_uint(q.z) != __float_as_uint(v[2] / len_ref)) bad_nrm++;
        /* sampling angles: u in [0, 4] exactly as the renderer forms them */
        s = xorshift_step(s);
        float u = (float)s * 9.31322574615478515625e-10f;
        float ang = (float)(kTwoPi * (double)u);
        float ss, cc, rs, rc;
        sincos_bounded(ang, ss, cc);
        sincosf(ang, &rs, &rc);
        float e1 = fabsf(ss - rs) * 16777216.0f, e2 = fabsf(cc - rc) * 16777216.0f;
        uint32_t w = (uint32_t)(e1 > e2 ? e1 : e2);
        worst = w > worst ? w : worst;
    }
    if (bad_div) atomicAdd(out + 0, bad_div);
    if (bad_nrm) atomicAdd(out + 1, bad_nrm);
    atomicMax(out + 2, worst);
}

template <bool R, int M, bool A>
cudaError_t launch_variant(const RtFrameParams& fp, uint32_t grid, uint32_t smem, cudaStream_t st) {
    render_frame<R, M, A><<<grid, kBlock, smem, st>>>(fp);
    return cudaGetLastError();
}
template <bool R, int M, bool A>
cudaError_t prep_variant(int* blocks_per_sm, uint32_t smem) {
    if (smem > 48u * 1024u) {
        cudaError_t e = cudaFuncSetAttribute(render_frame<R, M, A>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, render_frame<R, M, A>, kBlock, smem);
}

/* The tree kernel first, render_frame right behind it in the same stream with programmatic stream serialization
 * allowed: it starts once every CTA of render_tree has signalled (griddepcontrol.launch_dependents at its top). */
template <bool R, bool A>
cudaError_t go_tree(const RtFrameParams& fp, uint32_t grid, uint32_t smem, cudaStream_t st) {
    if (smem > 48u * 1024u) {
        cudaError_t e = cudaFuncSetAttribute(render_tree<R, A>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    render_tree<R, A><<<grid, kBlock, smem, st>>>(fp);
    return cudaGetLastError();
}
template <bool R, int M, bool A>
cudaError_t launch_variant_pdl(const RtFrameParams& fp, uint32_t grid, uint32_t smem, cudaStream_t st) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kBlock);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, render_frame<R, M, A>, fp);
}

template <bool R, int M, bool A>
cudaError_t go(RtFrameParams& fp, int sm_count, cudaStream_t st, uint32_t smem, RtLaunchShape* shape) {
    int bps = 0;
    cudaError_t e = prep_variant<R, M, A>(&bps, smem);
    if (e != cudaSuccess) return e;
    if (bps < 1) bps = 1;
    const uint32_t ntiles = fp.tiles_x * fp.tiles_y;
    const uint32_t warps_per_cta = kBlock / 32;
    uint32_t grid = (uint32_t)sm_count * (uint32_t)bps; /* persistent: every resident slot filled exactly once */
    const uint32_t need = (ntiles + warps_per_cta - 1) / warps_per_cta;
    if (grid > need) grid = need;
    if (grid == 0) grid = 1;
    if (shape) { shape->grid = grid; shape->block = kBlock; shape->smem_bytes = smem; shape->resident = R; shape->tree_grid = 0; }
    const bool with_tree = RT_TREE && (M == 0 || M == 4) && fp.tree_mode;
    uint32_t tgrid = 0;
    if (with_tree) {
        tgrid = fp.tree_slots;
        if (tgrid > ntiles) tgrid = ntiles;
        if (tgrid == 0) tgrid = 1;
    } else {
        fp.tree_mode = 0;
    }
    fp.done_expected = grid + tgrid;
    if (with_tree) {
        const uint32_t tsmem = (((R ? fp.lay.total * 16u : 0u) + kCoopBytesPerWarp + 15u) & ~15u) + (RT_TRIANGLES ? kTriStageBytes : 0u);
        e = go_tree<R, A>(fp, tgrid, tsmem, st);
        if (e != cudaSuccess) return e;
        if (shape) shape->tree_grid = tgrid;
        return launch_variant_pdl<R, M, A>(fp, grid, smem, st);
    }
    return launch_variant<R, M, A>(fp, grid, smem, st);
}

}  // namespace RT_NS

using namespace RT_NS;

#if RT_TRIANGLES && RT_ACCEL
#define RT_EXPORT(name) name##_tacc
#elif RT_TRIANGLES
#define RT_EXPORT(name) name##_tri
#elif RT_ACCEL
#define RT_EXPORT(name) name##_acc
#else
#define RT_EXPORT(name) name
#endif

#if !RT_TRIANGLES && !RT_ACCEL
cudaError_t rt_launch_scene_prep(const unsigned char* d_spheres, uint32_t S, const unsigned char* d_planes, uint32_t P,
                                 const unsigned char* d_lights, uint32_t L, float4* d_blob, RtBlobLayout lay,
                                 cudaStream_t st) {
    uint32_t n = lay.Spad > P ? lay.Spad : P;
    n = n > lay.Lpad ? n : lay.Lpad;
    if (n == 0) n = 1;
    scene_prep<<<(n + 127) / 128, 128, 0, st>>>(d_spheres, S, d_planes, P, d_lights, L, d_blob, lay);
    return cudaGetLastError();
}

#endif

#if RT_TRIANGLES && !RT_ACCEL
cudaError_t rt_launch_tri_prep(const unsigned char* d_tris, uint32_t T, float4* d_rec, cudaStream_t st) {
    if (T == 0) return cudaSuccess;
    tri_prep<<<(T + 127) / 128, 128, 0, st>>>(d_tris, T, d_rec);
    return cudaGetLastError();
}

#endif

cudaError_t RT_EXPORT(rt_launch_tile_order)(const RtFrameParams& fp, uint32_t* d_cost, uint32_t* d_hist_cursor, uint32_t* d_order,
                                 cudaStream_t st) {
    const uint32_t ntiles = fp.tiles_x * fp.tiles_y;
    /* d_hist_cursor: [64 histogram][64 cursors][queue length][heavy prefix][ticket]; zeroed once by the engine, left clean by
     * every tile_order.  fp.tile_time (history frames) and the 256-byte sched block are zeroed by these kernels too. */
    cudaError_t e = cudaSuccess;
    const uint32_t groups = ((fp.tiles_x + 1) / 2) * ((fp.tiles_y + 1) / 2);
    if (fp.tile_time_prev) tile_cost_history<<<(ntiles + 255) / 256, 256, 0, st>>>(fp.tile_time_prev, fp.tile_time_sum_prev, fp.cost_prev ? fp.cost_prev : d_cost, d_cost, d_hist_cursor, ntiles, fp.split_min_cost, fp.split32_min_cost,
                                                                                  fp.cost_is_history ? fp.tree_min_cost : kCostBuckets,
                                                                                  fp.tile_time ? reinterpret_cast<uint32_t*>(fp.tile_time_sum) : nullptr);
    else tile_cost<<<(groups + 127) / 128, 128, 0, st>>>(fp, d_cost, d_hist_cursor);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    tile_order<<<(ntiles + 255) / 256, 256, 0, st>>>(d_cost, d_hist_cursor, d_hist_cursor + kCostBuckets, d_order, ntiles,
                                                    fp.split_min_cost, fp.split32_min_cost, fp.tree_min_cost, d_hist_cursor + 2 * kCostBuckets,
                                                    reinterpret_cast<uint32_t*>(fp.tile_counter));
    return cudaGetLastError();
}

cudaError_t RT_EXPORT(rt_launch_render)(const RtFrameParams& fp_in, int mode, int sm_count, cudaStream_t st, RtLaunchShape* shape) {
    RtFrameParams fp = fp_in;
    const uint32_t blob_bytes = fp.lay.total * 16u;
    /* keep the geometry resident while it leaves room for >= 2 CTAs per SM */
    const bool resident = blob_bytes <= 96u * 1024u;
#if !RT_TRIANGLES && !RT_ACCEL
    if (mode == 0 && fp.T == 0u && fp.lay.Lpad + fp.lay.Spad >= fp.wide_min_records) mode = 4; /* many spheres: WIDE variant */
#endif
    const bool coop = mode == 0 || mode == 2 || mode == 4;
    const uint32_t smem = (resident ? blob_bytes : 0u) + (coop ? (kBlock / 32) * kCoopBytesPerWarp : 0u);
    const bool aov = fp.hit != nullptr;
#define RT_GO(R, M, A) return go<R, M, A>(fp, sm_count, st, smem, shape)
#define RT_GO_A(R, M) do { if (aov) RT_GO(R, M, true); else RT_GO(R, M, false); } while (0)
#if RT_TRIANGLES && RT_ACCEL
#define RT_GO_M(R) do { if (mode == 0) RT_GO_A(R, 0); } while (0) /* the engine sends mode 0 only (rt_engine.cu: use_tacc) */
#elif RT_TRIANGLES || RT_ACCEL
#define RT_GO_M(R) do { if (mode == 0) RT_GO_A(R, 0); else if (mode == 1) RT_GO_A(R, 1); else if (mode == 2) RT_GO_A(R, 2); else RT_GO_A(R, 3); } while (0)
#else
#define RT_GO_M(R) do { if (mode == 0) RT_GO_A(R, 0); else if (mode == 1) RT_GO_A(R, 1); else if (mode == 2) RT_GO_A(R, 2); else if (mode == 4) RT_GO_A(R, 4); else RT_GO_A(R, 3); } while (0)
#endif
    if (resident) RT_GO_M(true);
    else RT_GO_M(false);
#undef RT_GO
#undef RT_GO_A
#undef RT_GO_M
    return cudaErrorInvalidValue;
}

#if !RT_TRIANGLES && !RT_ACCEL
cudaError_t rt_launch_assemble_bands(const uint32_t* d_compact, size_t stride_words, uint32_t* d_frame, uint32_t W, uint32_t H,
                                     uint32_t band_rows, uint32_t parts, cudaStream_t st) {
    const size_t n = (size_t)((W + 3u) / 4u) * H;
    assemble_bands<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(d_compact, stride_words, d_frame, W, H, band_rows, parts);
    return cudaGetLastError();
}

cudaError_t rt_launch_wait_flags(const uint32_t* d_flags, uint32_t stride_words, uint32_t n, uint32_t value,
                                 uint32_t* d_timeouts, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    if (n > 1024) return cudaErrorInvalidValue;
    wait_flags<<<1, (n + 31u) & ~31u, 0, st>>>(d_flags, stride_words, n, value, d_timeouts);
    return cudaGetLastError();
}

cudaError_t rt_launch_fma_peak(float* d_sink, int iters, int blocks, int threads, cudaStream_t st) {
    fma_peak<<<blocks, threads, 0, st>>>(d_sink, iters);
    return cudaGetLastError();
}

cudaError_t rt_launch_arith_selftest(uint32_t seed, uint32_t per_thread, uint32_t* d_out, cudaStream_t st) {
    arith_selftest<<<148 * 4, 256, 0, st>>>(seed, per_thread, d_out);
    return cudaGetLastError();
}
#endif
