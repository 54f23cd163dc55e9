/* rt_trace_tree.cuh -- level-wise pixel trees: the segments of a heavy tile spread over the 128 lanes of a CTA.
 *
 * The stack machine of raytracing.cl:41-190 runs a pixel's segments one after the other; a glass pixel of render.map
 * has up to 166 of them at 3840x2160, a glass tile 2,000-3,000, and a lone warp walks them at ~7,500 clocks per step:
 * that chain (0.6 ms) was the floor of every multi-GPU frame and of the single-GPU tail.  But the segments of a pixel
 * form a binary tree -- a shaded segment continues as the reflection of its own stack entry (:150-159) and, through
 * glass, as a new entry for the refracted ray (:161-179) -- whose geometry depends on neither the RNG nor the
 * accumulated colour.  oracle/rt_split.c (rto_tree_render, tree_rank_flat, tree_replay_flat) proves on the CPU that the
 * following evaluation order gives the stack machine's words; this file is that order on one CTA (4 warps), for the
 * pixels of up to four queue entries at a time (each a whole 8x4 tile, one of its rows, or a single pixel):
 *
 *   1 EXPAND  breadth-first.  The frontier is a queue of ray states {origin, direction, f, n, depth, stack height,
 *             pixel} in the CTA's slab; each pass, thread j takes frontier node j: closest hit (one pass over lights,
 *             spheres, planes), classification, material / texel, ambient term, Fresnel, and its <= 2 continuations
 *             appended at prefix-sum positions (ballots per warp, warp totals through shared memory).  All nodes of
 *             BFS level k have depth k, so there are <= 15 levels.
 *   2 RANK    per node, last level to first: shaded segments and addition-stream slots of its sub-tree; then first
 *             level to last: the pre-order rank among shaded nodes (refraction sub-tree first -- the stack machine's
 *             order) and the position of the node's additions in its pixel's stream.  A shaded segment draws exactly
 *             4L random numbers (raytracing.cl:96-113 has no early exit), so its RNG state is
 *             xorshift32^(4L*rank)(pixel id): xorshift32 (primitives.cl:111-125) is linear over GF(2) and the state is
 *             a product of <= 15 precomputed 32x32 bit matrices (fp.jump, built per scene by the engine).
 *   3 SHADE   (shaded node, light) pairs dealt over the 128 threads: shading geometry, the four draws, two shadow
 *             rays, the light's specular and diffuse terms -- written straight into the stream slots.
 *   4 REPLAY  one thread per pixel runs down the pixel's stream: ADD v / PUSH / POP on a <= 15-deep stack of partial
 *             sums -- additions only, in exactly the order the stack machine performs them.
 *
 * Terms that the cooperative trace skips because they are exact zeros (R5) are stored as +0 and added: x + 0 == x for
 * every x but -0, and the sign of a zero cannot reach the packed word (clamp, *255, truncation).
 * A tree that outgrows the slab (fp.tree_cap nodes) makes the function return false; the caller then hands the entry
 * to the cooperative walk -- same words, the old way.
 */
#pragma once
#include "rt_trace_coop.cuh"

namespace RT_NS {

enum : uint32_t { kNdLight = 0u, kNdSky = 1u, kNdShaded = 2u };
enum : uint32_t { kOpAdd = 0u, kOpPush = 1u, kOpPop = 2u };
constexpr uint32_t kTreePlanes = RT_TREE_PLANES; /* R0 R1 M S0 S1 S2 T X, 16 bytes per node each (rt_kernels.h) */
constexpr uint32_t kWarps = kBlock / 32;

/* One queue entry -> a lane's pixel.  entry = tile | sub << 24; sub 0: whole 8x4 tile; 1..4: only tile row sub-1;
 * 5..36: only pixel sub-5 (the split entries of heavy tiles). */
struct TileLanes {
    uint32_t id;  /* global pixel id y*W + x (RNG seed, raytracing.cl:33) */
    size_t at;    /* index into the compact band buffer */
    bool live;    /* this lane owns a pixel of the frame (and of the entry) */
};
__device__ __forceinline__ TileLanes tile_lanes(const RtFrameParams& fp, uint32_t entry, uint32_t lane) {
    const uint32_t tile = entry & 0x00ffffffu, sub = entry >> 24;
    const uint32_t tx = tile % fp.tiles_x, ty = tile / fp.tiles_x;
    const uint32_t x = tx * kTileW + (lane & 7u);
    const uint32_t lr = ty * kTileH + (lane >> 3);
    const uint32_t y = ((lr / fp.band_rows) * fp.parts + fp.part) * fp.band_rows + lr % fp.band_rows;
    TileLanes t;
    t.id = y * fp.W + x;
    t.at = (size_t)lr * fp.W + x;
    t.live = x < fp.W && lr < fp.local_rows && y < fp.H && t.id < fp.total &&
             (sub == 0u || (sub <= 4u ? (lane >> 3) + 1u == sub : lane + 5u == sub));
    return t;
}

/* One CTA's slab: kTreePlanes planes of `cap` 16-byte records, cap shaded-list words, then the addition stream of
 * cap * (3 + 2L) 16-byte records.  SoA: thread j touching node base + j reads 512 contiguous bytes per warp and plane. */
struct TreeSlab {
    float4* base;
    uint32_t cap;
    __device__ __forceinline__ float4* plane(uint32_t k) const { return base + (size_t)k * cap; }
    __device__ __forceinline__ float4* R0() const { return plane(0); }   /* {o.xyz, f} */
    __device__ __forceinline__ float4* R1() const { return plane(1); }   /* {d.xyz, n} */
    __device__ __forceinline__ uint4* M() const { return reinterpret_cast<uint4*>(plane(2)); } /* {depth | sz<<8 | pix<<16 | kind<<24 | entry<<28, refr, refl, -} */
    __device__ __forceinline__ float4* S0() const { return plane(3); }   /* {ip.xyz, f*specular} */
    __device__ __forceinline__ float4* S1() const { return plane(4); }   /* {nrm.xyz, f*diffuse} */
    __device__ __forceinline__ float4* S2() const { return plane(5); }   /* {v.xyz, shininess} */
    __device__ __forceinline__ float4* T() const { return plane(6); }    /* {light / sky / ambient term, -} */
    __device__ __forceinline__ uint4* X() const { return reinterpret_cast<uint4*>(plane(7)); } /* {shaded in sub-tree, stream slots of sub-tree, rank -> RNG state, stream position} */
    __device__ __forceinline__ uint32_t* shaded() const { return reinterpret_cast<uint32_t*>(plane(kTreePlanes)); }
    __device__ __forceinline__ float4* stream() const { return plane(kTreePlanes) + (cap + 3u) / 4u; }
};

/* The CTA's shared scratch for one entry. */
constexpr uint32_t kTreeGroup = kWarps;         /* queue entries one CTA pools into one tree pass: warp w owns entry w */
constexpr uint32_t kTreePixels = 32u * kTreeGroup;
struct TreeShared {
    uint32_t lvl[kMaxDepth + 2];   /* first node of every BFS level, then the node count */
    uint32_t wcnt[kWarps][4];      /* per warp: refraction children, reflection children, shaded nodes of this pass */
    uint32_t pix_base[kTreePixels], pix_size[kTreePixels]; /* per pixel (entry slot * 32 + tile pixel): its addition stream */
    uint32_t word[kTreePixels], hid[kTreePixels];          /* per pixel: the results */
    uint32_t entry[kTreeGroup];    /* the queue entries of this pass */
    uint32_t shaded_of[kTreeGroup]; /* shaded segments per entry (cost attribution) */
    uint32_t slot, count;          /* queue cursor broadcast */
};

/* ---- triangle records staged into shared memory in chunks (extension X1 build) ----------------------------------------
 * The tree kernel runs a CTA in lock step -- every thread holds one ray of the current level (expand) or one pair of
 * shadow rays (shade) -- so a chunk of the triangle array fetched ONCE by a TMA bulk copy serves all 128 threads:
 * kTriChunk records of {v0, e1, e2} (12 KiB) per buffer, two buffers, two mbarriers; while the threads test chunk c from
 * one buffer the copy of chunk c+1 is in flight into the other.  Records are visited in index order, with the same test
 * (for_triangle_hits_ld<true>: plain shared-memory loads instead of the read-only global path): same words. */
constexpr uint32_t kTriChunk = 256;
constexpr uint32_t kTriStageBytes = 2u * kTriChunk * 48u + 16u;
struct TriStage {
    float4* buf;    /* [2][3 * kTriChunk] */
    uint64_t* bar;  /* [2] */
    uint32_t phase; /* bit b: the parity bar[b] completes next (CTA-uniform) */
};
__device__ __forceinline__ void tri_stage_init(TriStage& st, unsigned char* smem /* kTriStageBytes, 16-byte aligned */) {
    st.buf = reinterpret_cast<float4*>(smem);
    st.bar = reinterpret_cast<uint64_t*>(smem + 2u * kTriChunk * 48u);
    st.phase = 0;
    if (threadIdx.x == 0) { mbar_init(&st.bar[0], 1); mbar_init(&st.bar[1], 1); }
    __syncthreads();
}
__device__ __forceinline__ void tri_stage_issue(const TriStage& st, const float4* tri, uint32_t T, uint32_t c, uint32_t b) {
    const uint32_t first = c * kTriChunk;
    const uint32_t n = T - first < kTriChunk ? T - first : kTriChunk;
    mbar_expect_tx(&st.bar[b], n * 48u);
    tma_bulk_g2s(st.buf + (size_t)b * 3u * kTriChunk, tri + (size_t)3u * first, n * 48u, &st.bar[b]);
}
/* Every thread of the CTA calls this (uniformly).  test_chunk(records, first index, count) -> this thread still needs the
 * rest of the array; `need` = it needs anything at all.  The walk ends early when no thread does. */
template <class F>
__device__ __forceinline__ void tri_walk_staged(const float4* tri, uint32_t T, TriStage& st, bool need, F&& test_chunk) {
    const uint32_t nch = (T + kTriChunk - 1u) / kTriChunk;
    if (nch == 0u) return;
    if (!__syncthreads_or(need)) return; /* also: every thread is done with whatever used the buffers last */
    if (threadIdx.x == 0) {
        tri_stage_issue(st, tri, T, 0u, 0u);
        if (nch > 1u) tri_stage_issue(st, tri, T, 1u, 1u);
    }
#pragma unroll 1
    for (uint32_t c = 0; c < nch; c++) {
        const uint32_t b = c & 1u;
        mbar_wait(&st.bar[b], (st.phase >> b) & 1u);
        st.phase ^= 1u << b;
        if (need) {
            const uint32_t first = c * kTriChunk;
            need = test_chunk(st.buf + (size_t)b * 3u * kTriChunk, first, T - first < kTriChunk ? T - first : kTriChunk);
        }
        const int any = __syncthreads_or(need); /* everyone has finished reading buffer b */
        if (!any) {
            if (c + 1u < nch) { /* drain the copy already in flight so the barrier's parity stays in step */
                const uint32_t b1 = (c + 1u) & 1u;
                mbar_wait(&st.bar[b1], (st.phase >> b1) & 1u);
                st.phase ^= 1u << b1;
            }
            return;
        }
        if (threadIdx.x == 0 && c + 2u < nch) tri_stage_issue(st, tri, T, c + 2u, b);
    }
}

/* xorshift32^(draws*k)(seed): jump[j*32 + i] = image of bit i under (step^draws)^(2^j).  The j loop is warp-uniform
 * (runs to the highest set bit of any lane's k), so every table read is one broadcast load. */
__device__ __forceinline__ uint32_t rng_jump(const uint32_t* __restrict__ jump, uint32_t seed, uint32_t k, uint32_t kmax) {
    for (uint32_t j = 0; (kmax >> j) != 0u; j++) {
        uint32_t r = 0;
#pragma unroll 8
        for (uint32_t i = 0; i < 32u; i++) r ^= __ldg(jump + j * 32u + i) & (0u - ((seed >> i) & 1u));
        if ((k >> j) & 1u) seed = r;
    }
    return seed;
}

/* All kBlock threads of the CTA call this; the pass covers the queue entries sh.entry[first .. first + ne) (tile | sub << 24;
 * ne <= kTreeGroup: warp w makes the roots of entry first + w, and the pooled frontier fills the 128 lanes far better than
 * one tile's would) and, when only_pixel < 32, just that pixel of each (tests).  Returns false (CTA-uniformly) when the
 * tree does not fit the slab; otherwise sh.word[q] / sh.hid[q], q = w * 32 + tile pixel, hold the results. */
template <bool AOV>
__device__ __forceinline__ bool trace_entry_tree(const RtFrameParams& fp, const Geo& g, const TreeSlab sl, TreeShared& sh, uint32_t first,
                                                 uint32_t ne, uint32_t only_pixel, TriStage& stg) {
    /* extension X1 build with triangles present: their records go through the staged walk, everything else as ever */
    const bool staged = RT_TRIANGLES && g.T != 0u && !(RT_TRI_ACCEL && g.tacc_nc != 0u); /* N4 walks are per ray, not CTA-wide */
    Geo gs = g; /* the scene without its triangles */
    if (staged) gs.T = 0u;
    constexpr uint32_t kFull = 0xffffffffu;
    Tally<false> tl;
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t lt = (1u << lane) - 1u;
    const uint32_t L = g.L, cap = sl.cap;
    const uint32_t own_shaded = 1u + 2u * L;

    /* ---- roots: one node per live pixel; warp w makes those of entry first + w (lane p = tile pixel p) ---- */
    uint32_t nroots = 0;
    {
        const uint32_t entry = warp < ne ? sh.entry[first + warp] : 0u;
        const TileLanes tq = tile_lanes(fp, entry, lane);
        const bool live = warp < ne && tq.live && (only_pixel >= 32u || lane == only_pixel);
        const uint32_t live_mask = __ballot_sync(kFull, live);
        if (lane == 0u) sh.wcnt[warp][0] = (uint32_t)__popc(live_mask);
        __syncthreads();
        uint32_t before = 0;
#pragma unroll
        for (uint32_t w = 0; w < kWarps; w++) {
            const uint32_t a = sh.wcnt[w][0];
            if (w < warp) before += a;
            nroots += a;
        }
        if (live) {
            const uint32_t r = before + (uint32_t)__popc(live_mask & lt);
            V3 o, d;
            primary_ray(fp, tq.id, o, d);
            sl.R0()[r] = make_float4(o.x, o.y, o.z, 1.0f);
            sl.R1()[r] = make_float4(d.x, d.y, d.z, kAirN);
            sl.M()[r] = make_uint4(0u | (1u << 8) | ((warp * 32u + lane) << 16), 0xffffffffu, 0xffffffffu, 0u);
        }
        __syncthreads();
    }
    if (nroots == 0u) return true;

#ifdef RT_TREE_CLOCKS /* diagnostic build (tools/exp/tree_clocks.py): SM clocks per phase, summed over entries, in fp.counters */
    long long tc0 = clock64();
#define RT_TREE_TICK(k) do { if (tid == 0u) { const long long t_ = clock64(); atomicAdd(fp.counters + (k), (unsigned long long)(t_ - tc0)); tc0 = t_; } } while (0)
#else
#define RT_TREE_TICK(k) do { } while (0)
#endif
    /* ---- 1 EXPAND ---- */
    uint32_t n_nodes = nroots, n_shaded = 0, head = 0, nlev = 0;
    while (head < n_nodes) {
        const uint32_t level_end = n_nodes;
        if (tid == 0u) sh.lvl[nlev] = head;
        nlev++;
#pragma unroll 1
        for (uint32_t base = head; base < level_end; base += kBlock) {
            const uint32_t i = base + tid;
            const bool act = i < level_end;
            bool c_refr = false, c_refl = false, shaded = false;
            V3 ro = mk(0.0f, 0.0f, 0.0f), rd = ro, lo = ro, ld = ro; /* refraction / reflection continuations */
            float rf = 0.0f, rn = 0.0f, lf = 0.0f, ln = 0.0f;
            uint32_t meta = 0;
            V3 o = mk(0.0f, 0.0f, 0.0f), d = mk(0.0f, 0.0f, 1.0f);
            float f = 0.0f, n_cur = kAirN;
            Closest c;
            c.t = CUDART_INF_F; c.tL = CUDART_INF_F; c.who = -1; c.whop = -1; c.whot = -1; c.k = -1; c.occluded = false;
            if (act) {
                const float4 r0 = sl.R0()[i], r1 = sl.R1()[i];
                meta = sl.M()[i].x;
                o = xyz(r0); d = xyz(r1);
                f = r0.w; n_cur = r1.w;
                RayF r;
                ray_setup_fast(r, o, d);
                closest_pass(r, gs, tl, c);
            }
#if RT_TRIANGLES
            if (staged) /* CTA-uniform: triangles after the planes, same rules (closest_pass), from staged chunks */
                tri_walk_staged(g.tri, g.T, stg, act, [&](const float4* recs, uint32_t k0, uint32_t n) {
                    for_triangle_hits_ld<true>(o, d, recs, n, tl, [&](uint32_t k, float t1) {
                        if (!(t1 >= c.t)) { c.t = t1; c.whot = (int)(k0 + k); }
                        if (t1 <= c.tL) c.occluded = true;
                        return false;
                    });
                    return true;
                });
#endif
            if (act) {
                const uint32_t depth = meta & 0xffu, sz = (meta >> 8) & 0xffu;
                uint32_t kind, entry_child = 0u, hid = 0u;
                V3 term;
                if (c.k >= 0 && !c.occluded) { /* a visible light ends the chain, raytracing.cl:48-54 */
                    V3 lp = o + d * c.tL;
                    V3 od = o - lp;
                    float dd = sqrtf(dot3(od, od));
                    V3 col = xyz(g.laux[c.k]) * (1.0f / dd * dd); /* (1/d)*d, primitives.cl:288 */
                    term = col * f;
                    kind = kNdLight;
                    hid = RTB_HIT_LIGHT | (uint32_t)c.k;
                } else if (c.who < 0 && c.whop < 0 && c.whot < 0) { /* skybox, raytracing.cl:61-81 */
                    int u, v;
                    cube_uv(d, (uint32_t)(fp.sw / 4), u, v);
                    V3 px = texel_rgb(fp.sky, fp.sw, fp.sh, 1, u, fp.sh - v, 0, g.lut255);
                    term = px * f;
                    kind = kNdSky;
                } else {
                    kind = kNdShaded;
                    shaded = true;
                    V3 ip = o + d * c.t, nrm;
                    Material m;
                    if (c.whot >= 0) { /* extension X1 */
                        nrm = triangle_normal(g.tri + 3 * c.whot);
                        m = load_material_at(fp.raw_tris + (size_t)c.whot * RT_TRIANGLE_BYTES + RT_TRI_MATERIAL_OFFSET);
                        hid = 1u + g.S + g.P + (uint32_t)c.whot;
                    } else if (c.whop >= 0) {
                        nrm = xyz(g.pln[4 * c.whop]);
                        m = load_material(fp.raw_planes + (size_t)c.whop * RT_PLANE_BYTES);
                        if (m.texture_id >= 0) { /* primitives.cl:217-259; sampled once, for the plane that won */
                            float4 b0 = g.pln[4 * c.whop + 2], b1 = g.pln[4 * c.whop + 3];
                            float ui = dot3(xyz(b0), ip) * m.texture_scale;
                            float vi = dot3(xyz(b1), ip) * m.texture_scale;
                            m.rgb = texel_rgb(fp.tex, fp.tw, fp.th, fp.tl, emod(f2i_x86(ui), fp.tw), emod(f2i_x86(vi), fp.th),
                                              m.texture_id, g.lut255);
                        }
                        hid = 1u + g.S + (uint32_t)c.whop;
                    } else {
                        float qlen;
                        nrm = normalize_len(ip - xyz(g.sph[c.who]), qlen);
                        m = load_material(fp.raw_spheres + (size_t)c.who * RT_SPHERE_BYTES);
                        hid = 1u + (uint32_t)c.who;
                    }
                    ip = ip + nrm * kEps;
                    term = (m.rgb * f) * m.ambient; /* raytracing.cl:83-84 */
                    float vlen;
                    const V3 vv = normalize_len(o - ip, vlen);
                    sl.S0()[i] = make_float4(ip.x, ip.y, ip.z, f * m.specular);
                    sl.S1()[i] = make_float4(nrm.x, nrm.y, nrm.z, f * m.diffuse);
                    sl.S2()[i] = make_float4(vv.x, vv.y, vv.z, (float)m.shininess);
                    /* Fresnel, reflection, refraction child (raytracing.cl:138-179) */
                    const float n1 = n_cur;
                    const float n2 = (n1 == kAirN) ? m.n : kAirN;
                    const float cosI = -dot3(nrm, d);
                    const float refl = reflect_amount(m, n1, n2, cosI);
                    const uint32_t depth1 = depth + 1u;
                    if (m.transperent && sz < (uint32_t)kMaxDepth && refl < 1.0f) {
                        if (refract_child(ip, nrm, d, n1, n2, ro, rd)) {
                            entry_child = 1u; /* the entry is pushed (and popped) even when it has no segment left to run */
                            if (depth1 < (uint32_t)kMaxDepth) { c_refr = true; rf = f * (1.0f - refl); rn = n2; }
                        }
                    }
                    if (depth1 < (uint32_t)kMaxDepth) {
                        c_refl = true;
                        lo = ip;
                        ld = d + nrm * (2.0f * cosI); /* primitives.cl:127-130 */
                        lf = f * refl;
                        ln = n_cur;
                    }
                }
                sl.T()[i] = make_float4(term.x, term.y, term.z, 0.0f);
                meta |= (kind << 24) | (entry_child << 28);
                if (AOV && base == 0u) sh.hid[(meta >> 16) & 127u] = hid; /* level 0: the roots (nroots <= kBlock: one pass) */
            }
            /* positions of the continuations: ballots inside the warp, warp totals through shared memory */
            const uint32_t m1 = __ballot_sync(kFull, c_refr), m2 = __ballot_sync(kFull, c_refl), ms = __ballot_sync(kFull, shaded);
            if (lane == 0u) { sh.wcnt[warp][0] = (uint32_t)__popc(m1); sh.wcnt[warp][1] = (uint32_t)__popc(m2); sh.wcnt[warp][2] = (uint32_t)__popc(ms); }
            __syncthreads();
            uint32_t b1 = 0, b2 = 0, bs = 0, t1 = 0, t2 = 0, ts = 0;
#pragma unroll
            for (uint32_t w = 0; w < kWarps; w++) {
                const uint32_t a = sh.wcnt[w][0], b = sh.wcnt[w][1], c = sh.wcnt[w][2];
                if (w < warp) { b1 += a; b2 += b; bs += c; }
                t1 += a; t2 += b; ts += c;
            }
            if (n_nodes + t1 + t2 > cap) return false; /* CTA-uniform */
            if (act) {
                const uint32_t child_meta = ((meta & 0xffu) + 1u) | (meta & 0x00ffff00u);
                uint32_t refr = 0xffffffffu, refl_i = 0xffffffffu;
                if (c_refr) {
                    refr = n_nodes + b1 + (uint32_t)__popc(m1 & lt);
                    sl.R0()[refr] = make_float4(ro.x, ro.y, ro.z, rf);
                    sl.R1()[refr] = make_float4(rd.x, rd.y, rd.z, rn);
                    sl.M()[refr] = make_uint4((child_meta & 0xffff00ffu) | ((((meta >> 8) & 0xffu) + 1u) << 8), 0xffffffffu, 0xffffffffu, 0u);
                }
                if (c_refl) {
                    refl_i = n_nodes + t1 + b2 + (uint32_t)__popc(m2 & lt);
                    sl.R0()[refl_i] = make_float4(lo.x, lo.y, lo.z, lf);
                    sl.R1()[refl_i] = make_float4(ld.x, ld.y, ld.z, ln);
                    sl.M()[refl_i] = make_uint4(child_meta, 0xffffffffu, 0xffffffffu, 0u);
                }
                sl.M()[i] = make_uint4(meta, refr, refl_i, 0u);
                if (shaded) sl.shaded()[n_shaded + bs + (uint32_t)__popc(ms & lt)] = i;
            }
            n_nodes += t1 + t2;
            n_shaded += ts;
            __syncthreads(); /* the children are visible to whoever expands them; wcnt may be rewritten */
        }
        head = level_end;
    }
    if (tid == 0u) sh.lvl[nlev] = n_nodes;
    __syncthreads();
    RT_TREE_TICK(0);
#ifdef RT_TREE_CLOCKS
    if (tid == 0u) { atomicAdd(fp.counters + 5, 1ull); atomicAdd(fp.counters + 6, (unsigned long long)n_nodes); atomicAdd(fp.counters + 7, (unsigned long long)nlev); }
#endif

    /* ---- 2 RANK: sub-tree totals, last level first ---- */
    for (uint32_t lv = nlev; lv-- > 0u;) {
        const uint32_t b = sh.lvl[lv], e = sh.lvl[lv + 1u];
        for (uint32_t i = b + tid; i < e; i += kBlock) {
            const uint4 mm = sl.M()[i];
            const bool shd = ((mm.x >> 24) & 0xfu) == kNdShaded;
            uint32_t cnt = shd ? 1u : 0u;
            uint32_t size = (shd ? own_shaded : 1u) + (((mm.x >> 28) & 1u) ? 2u : 0u);
            if (mm.y != 0xffffffffu) { const uint4 x = sl.X()[mm.y]; cnt += x.x; size += x.y; }
            if (mm.z != 0xffffffffu) { const uint4 x = sl.X()[mm.z]; cnt += x.x; size += x.y; }
            sl.X()[i] = make_uint4(cnt, size, 0u, 0u);
        }
        __syncthreads();
    }
    /* stream base of every pixel: exclusive scan of the roots' sizes (roots are nodes 0 .. nroots-1, nroots <= kBlock);
     * shaded segments per entry for the cost attribution */
    {
        const uint4 rx = tid < nroots ? sl.X()[tid] : make_uint4(0u, 0u, 0u, 0u);
        const uint32_t my_size = rx.y;
        uint32_t scan = my_size;
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t v = __shfl_up_sync(kFull, scan, off);
            if ((int)lane >= off) scan += v;
        }
        if (lane == 31u) sh.wcnt[warp][0] = scan;
        if (tid < kTreeGroup) sh.shaded_of[tid] = 0u;
        __syncthreads();
        uint32_t before = 0;
#pragma unroll
        for (uint32_t w = 0; w < kWarps; w++) if (w < warp) before += sh.wcnt[w][0];
        if (tid < nroots) {
            uint4 x = rx;
            x.z = 0u; x.w = before + scan - my_size;
            sl.X()[tid] = x;
            const uint32_t pix = (sl.M()[tid].x >> 16) & 127u;
            sh.pix_base[pix] = x.w;
            sh.pix_size[pix] = my_size;
            atomicAdd(&sh.shaded_of[pix >> 5], rx.x);
        }
    }
    __syncthreads();
    RT_TREE_TICK(1);
    /* ranks and stream positions, first level first; the node's own first addition and the PUSH / POP marks */
    float4* const stream = sl.stream();
    for (uint32_t lv = 0; lv < nlev; lv++) {
        const uint32_t b = sh.lvl[lv], e = sh.lvl[lv + 1u];
        for (uint32_t i = b + tid; i < e; i += kBlock) {
            {
                const uint4 mm = sl.M()[i];
                const uint4 x = sl.X()[i];
                const uint32_t rank = x.z;
                const uint32_t pos = x.w;
                const bool shd = ((mm.x >> 24) & 0xfu) == kNdShaded;
                const float4 t = sl.T()[i];
                stream[pos] = make_float4(t.x, t.y, t.z, __uint_as_float(kOpAdd));
                if (shd) {
                    uint32_t p = pos + own_shaded;
                    uint32_t crank = rank + 1u;
                    if ((mm.x >> 28) & 1u) {
                        stream[p] = make_float4(0.0f, 0.0f, 0.0f, __uint_as_float(kOpPush));
                        p += 1u;
                        if (mm.y != 0xffffffffu) {
                            uint4 cx = sl.X()[mm.y];
                            const uint32_t ccnt = cx.x, csize = cx.y;
                            cx.z = crank; cx.w = p;
                            sl.X()[mm.y] = cx;
                            crank += ccnt;
                            p += csize;
                        }
                        stream[p] = make_float4(0.0f, 0.0f, 0.0f, __uint_as_float(kOpPop));
                        p += 1u;
                    }
                    if (mm.z != 0xffffffffu) {
                        uint4 cx = sl.X()[mm.z];
                        cx.z = crank; cx.w = p;
                        sl.X()[mm.z] = cx;
                    }
                }
            }
        }
        __syncthreads();
    }
    /* RNG state on entry to every shaded segment: xorshift32^(4L*rank)(pixel id), the pixel's global id y*W + x being its
     * seed (raytracing.cl:33).  One pass over the shaded list -- whole warps run it: rng_jump's table walk is
     * warp-uniform -- instead of a jump per level. */
    for (uint32_t k0 = 0; k0 < n_shaded; k0 += kBlock) {
        const uint32_t k = k0 + tid;
        const bool act = k < n_shaded;
        uint32_t node = 0, rank = 0, seed = 0;
        if (act) {
            node = sl.shaded()[k];
            const uint32_t pix = (sl.M()[node].x >> 16) & 127u;
            rank = sl.X()[node].z;
            seed = tile_lanes(fp, sh.entry[first + (pix >> 5)], pix & 31u).id;
        }
        const uint32_t kmax = __reduce_max_sync(kFull, rank);
        const uint32_t state = rng_jump(fp.jump, seed, rank, kmax);
        if (act) reinterpret_cast<uint32_t*>(sl.X() + node)[2] = state;
    }
    __syncthreads();

    RT_TREE_TICK(2);
    /* ---- 3 SHADE: (shaded node, light) pairs over the threads ---- */
    const uint32_t items = n_shaded * L;
#pragma unroll 1
    for (uint32_t it0 = 0; it0 < items; it0 += kBlock) { /* uniform trip count: the staged triangle walk is a CTA-wide step */
        const uint32_t it = it0 + tid;
        const bool act = it < items;
        if (!staged && !act) continue;
        V3 ip = mk(0.0f, 0.0f, 0.0f);
        float fs = 0.0f, fd = 0.0f, shin = 0.0f, sx = 0.0f, dif = 0.0f, dist = 1.0f;
        bool tame = true, spec_zero = true, diff_zero = true;
        float4 l1 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        float4* slot = nullptr;
        V3 smp[kShadowSamples];
        smp[0] = smp[1] = ip;
        if (act) {
            const uint32_t k = it / L, li = it - k * L;
            const uint32_t node = sl.shaded()[k];
            const float4 s0 = sl.S0()[node], s1 = sl.S1()[node], s2 = sl.S2()[node];
            const uint4 x = sl.X()[node];
            const V3 nrm = xyz(s1), vv = xyz(s2);
            ip = xyz(s0);
            fs = s0.w; fd = s1.w; shin = s2.w;
            uint32_t rng = x.z;
            for (uint32_t q = 0; q < 4u * li; q++) rng = xorshift_step(rng); /* the lights before this one drew 4 numbers each */
            const float4 l0 = g.lgeo[li];
            l1 = g.laux[li];
            const V3 sdir = normalize_len(xyz(l0) - ip, dist);
            float hlen;
            const V3 hh = normalize_len(vv + sdir, hlen);
            const float ndh = dot3(nrm, hh);
            sx = 0.0f < ndh ? ndh : 0.0f;
            const float ndl = dot3(nrm, sdir);
            dif = 0.0f < ndl ? ndl : 0.0f;
            /* R5: exact zeros -- only while pow(sx, shin) is finite (see rt_trace_fast.cuh) */
            tame = sx <= 1.0f && shin >= 0.0f;
            spec_zero = tame && ((fs == 0.0f) || (sx == 0.0f && shin > 0.0f));
            diff_zero = (fd == 0.0f) || (dif == 0.0f);
            slot = stream + x.w + 1u + 2u * li;
            if (!(spec_zero && diff_zero)) {
#pragma unroll
                for (int j = 0; j < kShadowSamples; j++) {
                    const float theta = (float)(kTwoPi * (double)rng_next(rng));
                    const float phi = (float)(kPi * (double)rng_next(rng));
                    float sp, cp, sth, cth;
                    sincos_bounded(phi, sp, cp);
                    sincos_bounded(theta, sth, cth);
                    const float rs = l1.w * sp;
                    smp[j] = xyz(l0) + mk(rs * cth, rs * sth, l1.w * cp); /* raytracing.cl:101-106 */
                }
            }
        }
        static_assert(kShadowSamples == 2, "the two samples of a light are traced together");
        const bool rays = act && !(spec_zero && diff_zero);
        float r0 = 0.0f, r1 = 0.0f;
        if (!staged) {
            if (rays) shadow_pair_fast(smp[0], smp[1], ip, g, tl, r0, r1);
        } else {
#if RT_TRIANGLES
            ShadowPair sp;
            sp.dead0 = sp.dead1 = true; sp.o0 = sp.o1 = 1.0f; sp.len0 = sp.len1 = 0.0f;
            if (rays) shadow_pair_solids(smp[0], smp[1], ip, gs, tl, sp);
            tri_walk_staged(g.tri, g.T, stg, !(sp.dead0 && sp.dead1), [&](const float4* recs, uint32_t, uint32_t n) {
                const float len0 = sp.len0, len1 = sp.len1;
                if (!sp.dead0 && for_triangle_hits_ld<true>(sp.r0.o, sp.r0.d, recs, n, tl, [&](uint32_t, float t1) { return !(t1 >= len0); })) sp.dead0 = true;
                if (!sp.dead1 && for_triangle_hits_ld<true>(sp.r1.o, sp.r1.d, recs, n, tl, [&](uint32_t, float t1) { return !(t1 >= len1); })) sp.dead1 = true;
                return !(sp.dead0 && sp.dead1);
            });
            if (rays) { r0 = sp.dead0 ? 0.0f : sp.o0; r1 = sp.dead1 ? 0.0f : sp.o1; }
#endif
        }
        if (act) {
            V3 spec = mk(0.0f, 0.0f, 0.0f), diffv = spec;
            if (rays) {
                const float soft = (0.0f + r0) + r1;
                if (!(soft == 0.0f && tame)) { /* else light_rgb *= 0: both terms are exact zeros */
                    const float ssr = soft / (float)kShadowSamples;
                    const V3 lrgb = div3_pos(xyz(l1) * 1.0f, dist * dist) * ssr;
                    if (!spec_zero) spec = (lrgb * fs) * pow_ieee(sx, shin);
                    if (!diff_zero) diffv = (lrgb * fd) * dif;
                }
            }
            slot[0] = make_float4(spec.x, spec.y, spec.z, __uint_as_float(kOpAdd));
            slot[1] = make_float4(diffv.x, diffv.y, diffv.z, __uint_as_float(kOpAdd));
        }
    }
    __syncthreads();

    RT_TREE_TICK(3);
    /* ---- 4 REPLAY: additions in the stack machine's order; thread q replays pixel q = entry slot * 32 + tile pixel ---- */
    if (warp < ne) {
        const TileLanes tq = tile_lanes(fp, sh.entry[first + warp], lane);
        if (tq.live && (only_pixel >= 32u || lane == only_pixel)) {
            /* The additions are one dependent chain per pixel (their order is the result), so the heaviest pixel of the
             * entry sets the pace: 166 segments = 1,300 records.  Records are fetched eight at a time, one batch ahead of
             * the one being added (they sit in L2: other threads wrote them), so the chain never waits for memory --
             * 169k -> clocks per entry measured before (tools/exp/tree_clocks.py), 44 % of the whole pass.  Batches are
             * padded with ADD +0 records: x + 0 == x (see the header). */
            V3 acc = mk(0.0f, 0.0f, 0.0f);
            V3 stk[kMaxDepth];
            uint32_t sp = 0;
            const float4* ev = stream + sh.pix_base[tid];
            const uint32_t n = sh.pix_size[tid];
            const float4 pad = make_float4(0.0f, 0.0f, 0.0f, __uint_as_float(kOpAdd));
            float4 cur[8], nxt[8];
#pragma unroll
            for (uint32_t k = 0; k < 8u; k++) cur[k] = k < n ? __ldcg(ev + k) : pad;
            for (uint32_t q = 0; q < n; q += 8u) {
#pragma unroll
                for (uint32_t k = 0; k < 8u; k++) nxt[k] = q + 8u + k < n ? __ldcg(ev + q + 8u + k) : pad;
#pragma unroll
                for (uint32_t k = 0; k < 8u; k++) {
                    const float4 r = cur[k];
                    const uint32_t op = __float_as_uint(r.w);
                    if (op == kOpAdd) acc = acc + xyz(r);
                    else if (op == kOpPush) { stk[sp++] = acc; acc = mk(0.0f, 0.0f, 0.0f); }
                    else { sp--; acc = stk[sp] + acc; }
                }
#pragma unroll
                for (uint32_t k = 0; k < 8u; k++) cur[k] = nxt[k];
            }
            sh.word[tid] = pack_rgb(acc);
        }
    }
    __syncthreads();
    RT_TREE_TICK(4);
    return true;
}

}  // namespace RT_NS
