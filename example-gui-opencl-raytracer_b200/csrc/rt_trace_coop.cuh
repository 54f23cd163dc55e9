/* rt_trace_coop.cuh -- the production trace: warp-cooperative shadow rays.
 *
 * Same pixel words as rt_trace_fast.cuh / rt_trace_literal.cuh (and the oracle); what changes is WHO runs the
 * shadow rays.  Measured on render.map at 4K (tools/exp/tile_times.py): with one pixel per lane from start to
 * finish, the slowest 8x4 tile (a glass-sphere pixel with >1000 segments) runs for 2.15 ms of a 2.66 ms
 * kernel, and lanes whose segment ended on the sky or a light idle while their neighbours trace 2L shadow
 * rays.  Shadow rays are 3/4 of all tests and independent of one another, so here the warp
 * walks its tile in lock step, one segment per live lane per round:
 *   A  (per lane)  closest hit, classification, material, ambient term; per light of a chunk of <= 4 lights:
 *                  shading geometry, the four RNG draws (raytracing.cl:99-100), and -- unless the light's
 *                  terms are exact zeros (R5) -- two shadow TASKS {theta, phi, light, owner lane} appended to
 *                  the warp's task list in shared memory (positions from ballots);
 *   B  (warp)      tasks are dealt round-robin to ALL 32 lanes, finished pixels included: sample point on
 *                  the light (bounded sincos), testShadowPath, opacity written back into the task slot;
 *   C  (per lane)  the owner folds its opacities into rgb in the reference's order (light by light, specular
 *                  then diffuse), then Fresnel / reflection / refraction child as before.
 * A task's value does not depend on which lane computes it (same routine, same operands), so the words are
 * identical; the per-pixel RNG stream is drawn by the owner in the reference's order.
 * Shared memory per warp: 32 owner records (hit point) + 192 task slots + 3x32 light records + 32x9 R8 records,
 * 16 bytes each = 9.5 KiB.
 */
#pragma once
#include "rt_trace_fast.cuh"

namespace RT_NS {

#ifndef RT_COOP_LIGHTS
#define RT_COOP_LIGHTS 4
#endif
#ifndef RT_PRE_S
#define RT_PRE_S 0 /* R8 measured slower on B200 (2.84 vs 2.63 ms: bank conflicts on per-owner records, smaller L1 for the ray stacks): off */
#endif
constexpr int kCoopLights = RT_COOP_LIGHTS;                      /* lights per chunk */
constexpr int kPreS = RT_PRE_S, kPreRec = kPreS ? kPreS + 1 : 0;                    /* R8: per-owner records: <= 8 padded spheres + 1 of <= 4 plane numerators */
constexpr int kCoopTasks = 32 * kCoopLights * kShadowSamples;    /* 256 */
constexpr uint32_t kCoopBytesPerWarp = 32 * 16 + kCoopTasks * 16 + 32 * kCoopLights * 16 + 32 * kPreRec * 16; /* 6656 */

struct CoopSmem {
    float4* owner; /* [32] {ip.xyz, -} */
    float4* task;  /* [kCoopTasks] {theta, phi, light (bits), owner lane (bits)}; .x becomes the opacity */
    float4* pre;   /* [32][kPreRec] R8 records of each owner's current hit point (used when Spad <= 8 and P <= 4) */
    float4* light; /* [kCoopLights][32] per owner and light of the chunk: {dist, max(0,n.h), max(0,n.l), flags | slot << 8} */
};

/* Phase B for one task. */
template <bool COUNT>
__device__ __forceinline__ float coop_run_task(const Geo& g, const CoopSmem& cs, float4 tk, Tally<COUNT>& tl) {
    const uint32_t li = __float_as_uint(tk.z), ow = __float_as_uint(tk.w);
    const float4 l0 = g.lgeo[li];
    const float radius = g.laux[li].w;
    const float4 ipo = cs.owner[ow];
    float sp, cp, sth, cth;
    sincos_bounded(tk.y, sp, cp);
    sincos_bounded(tk.x, sth, cth);
    const float rs = radius * sp;
    const V3 sample = xyz(l0) + mk(rs * cth, rs * sth, radius * cp); /* raytracing.cl:101-106 */
    if (kPreS && g.Spad <= (uint32_t)kPreS && g.P <= 4u) return shadow_path_pre(sample, xyz(ipo), cs.pre + ow * kPreRec, kPreS, g, tl);
    return shadow_path_fast(sample, xyz(ipo), g, tl);
}

/* ---- WIDE scenes: one ray, 32 lanes ------------------------------------------------------------------------
 * With hundreds of spheres a ray's pass over the records is long, and it is paid by the whole warp even when
 * only a few lanes still trace (tile tails; the split entries of the heaviest tiles).  In the WIDE variant of the
 * kernel (chosen by the launcher for scenes with >= fp.wide_min_records light+sphere records and no triangles)
 * a round with at most `max_live` tracing lanes handles their rays ONE AFTER THE OTHER, the sphere records dealt
 * over the 32 lanes interleaved (lane l tests records l, l+32, l+64, l+96 of each 128-record block: every LDS.128
 * of the warp reads 32 consecutive records, no bank conflicts); likewise the last, partial pass of shadow tasks.  The outcome per ray is
 * the serial one: the closest hit is the (t, index) minimum -- "first of equal t wins" is the lowest index --,
 * occlusion is an OR, a shadow ray's opacity is kGlassPass multiplied in once per glass hit (a count), blocked
 * is an OR.  A NaN hit distance makes the serial update order-dependent (!(t1 >= t)), so such a ray is re-run
 * serially by its owner.
 * Break-even: a serial pass costs the warp ~17 instructions per record, a cooperative ray 68 per 128 records plus
 * a fixed part (broadcast, ray set-up, lights and planes on every lane, reductions) that the sweeps put at ~900
 * instruction-equivalents -- measured optimum 12..14 lanes for 1,028 records (4K grid: 188 ms vs 193 at 24, 199 at
 * 31) and 3..5 for 260 (8K: 149 ms vs 161 at 6, 182 at 20).  max_live = 17 n / (900 + 68 n / 128). */
__device__ __forceinline__ uint32_t wide_max_live(const RtFrameParams& fp, const Geo& g) {
    if (fp.wide_max_live >= 0) return (uint32_t)fp.wide_max_live;
    const uint32_t n = g.Lpad + g.Spad;
    const uint32_t m = (17u * n) / (900u + (68u * n) / 128u);
    return m > 31u ? 31u : m;
}

/* closest_pass for the tracing lanes of a warp; every lane calls it. */
template <bool COUNT>
__device__ __forceinline__ void closest_pass_warp(const RayF& r, V3 o, V3 d, bool tracing, const Geo& g, Tally<COUNT>& tl,
                                                  Closest& c, uint32_t lane, uint32_t max_live) {
    constexpr uint32_t kFull = 0xffffffffu;
    uint32_t mask = __ballot_sync(kFull, tracing);
    if ((uint32_t)__popc(mask) > max_live) {
        if (tracing) closest_pass(r, g, tl, c);
        return;
    }
    const uint32_t L = g.L, Lpad = g.Lpad, S = g.S, end = g.Lpad + g.Spad;
    while (mask) {
        const int src = __ffs((int)mask) - 1;
        mask &= mask - 1u;
        const V3 oo = mk(__shfl_sync(kFull, o.x, src), __shfl_sync(kFull, o.y, src), __shfl_sync(kFull, o.z, src));
        const V3 dd = mk(__shfl_sync(kFull, d.x, src), __shfl_sync(kFull, d.y, src), __shfl_sync(kFull, d.z, src));
        RayF rr;
        ray_setup_fast(rr, oo, dd);
        Closest cc;
        cc.t = CUDART_INF_F; cc.tL = CUDART_INF_F;
        cc.who = -1; cc.whop = -1; cc.whot = -1; cc.k = -1;
        cc.occluded = false;
        /* lights: few, and tL must be final before the spheres -- every lane runs them */
        for_sphere_hits(rr, g.lgeo, 0, Lpad, tl, [&](uint32_t i, float t1) {
            if (i < L && !(t1 >= cc.tL)) { cc.tL = t1; cc.k = (int)i; }
            return false;
        });
        float bt = CUDART_INF_F;
        int bw = -1;
        bool occ = false, odd = false;
#pragma unroll 1
        for (uint32_t i0 = Lpad + lane; i0 < end; i0 += 128u) { /* records i0, i0+32, i0+64, i0+96: conflict-free LDS.128 */
            for_sphere_hits_strided(rr, g.lgeo, i0, end, tl, [&](uint32_t i, float t1) {
                const uint32_t s = i - Lpad;
                if (s < S) {
                    if (t1 != t1) odd = true;
                    if (!(t1 >= bt)) { bt = t1; bw = (int)s; }
                    if (t1 <= cc.tL && !g.flag[s]) occ = true;
                }
                return false;
            });
        }
        if (__any_sync(kFull, odd)) { /* NaN distance: order-dependent, the owner walks the records itself */
            if ((int)lane == src) closest_pass(r, g, tl, c);
            continue;
        }
        /* hit distances are > 0 (sphere_root), so their bit patterns order like unsigned integers; +inf = none */
        const uint32_t tb = __float_as_uint(bt);
        const uint32_t tmin = __reduce_min_sync(kFull, tb);
        const uint32_t wmin = __reduce_min_sync(kFull, (tb == tmin && bw >= 0) ? (uint32_t)bw : 0x7fffffffu);
        cc.t = __uint_as_float(tmin);
        cc.who = wmin == 0x7fffffffu ? -1 : (int)wmin;
        cc.occluded = __any_sync(kFull, occ);
        for (uint32_t j = 0; j < g.P; j++) { /* planes: few; every lane, same values */
            float t1;
            if (!hit_plane_fast(rr, xyz(g.pln[4 * j]), xyz(g.pln[4 * j + 1]), t1, tl)) continue;
            if (!(t1 >= cc.t)) { cc.t = t1; cc.whop = (int)j; }
            if (t1 <= cc.tL) cc.occluded = true;
        }
        for (uint32_t k = 0; k < g.T; k++) { /* not reached: the launcher keeps triangle scenes off this variant */
            float t1;
            if (!hit_triangle(rr.o, rr.d, g.tri + 3 * k, t1, tl)) continue;
            if (!(t1 >= cc.t)) { cc.t = t1; cc.whot = (int)k; }
            if (t1 <= cc.tL) cc.occluded = true;
        }
        if ((int)lane == src) c = cc;
    }
}

/* Phase B for one task, all 32 lanes on it; every lane returns the opacity. */
template <bool COUNT>
__device__ __forceinline__ float coop_run_task_warp(const Geo& g, const CoopSmem& cs, float4 tk, uint32_t lane, Tally<COUNT>& tl) {
    constexpr uint32_t kFull = 0xffffffffu;
    const uint32_t li = __float_as_uint(tk.z), ow = __float_as_uint(tk.w);
    const float4 l0 = g.lgeo[li];
    const float radius = g.laux[li].w;
    const V3 from = xyz(cs.owner[ow]);
    float sp, cp, sth, cth;
    sincos_bounded(tk.y, sp, cp);
    sincos_bounded(tk.x, sth, cth);
    const float rs = radius * sp;
    const V3 to = xyz(l0) + mk(rs * cth, rs * sth, radius * cp); /* raytracing.cl:101-106 */
    V3 diff = to - from;
    float len;
    V3 dir = normalize_len(diff, len);
    RayF r;
    ray_setup_fast(r, from, dir);
    const uint32_t S = g.S;
    uint32_t nglass = 0;
    bool blocked = false;
#pragma unroll 1
    for (uint32_t i0 = lane; i0 < g.Spad && !blocked; i0 += 128u) { /* records i0, i0+32, i0+64, i0+96 */
        blocked = for_sphere_hits_strided(r, g.sph, i0, g.Spad, tl, [&](uint32_t i, float t1) {
            if (i >= S || t1 >= len) return false;
            if (g.flag[i]) { nglass++; return false; }
            return true;
        });
    }
    if (__any_sync(kFull, blocked)) return 0.0f;
    const uint32_t n = __reduce_add_sync(kFull, nglass);
    float opacity = 1.0f;
    for (uint32_t k = 0; k < n; k++) opacity *= kGlassPass; /* primitives.cl:396-442: one factor per glass hit */
    for (uint32_t j = 0; j < g.P; j++) {
        float t1;
        if (!hit_plane_fast(r, xyz(g.pln[4 * j]), xyz(g.pln[4 * j + 1]), t1, tl) || t1 >= len) continue;
        return 0.0f;
    }
    for (uint32_t k = 0; k < g.T; k++) {
        float t1;
        if (!hit_triangle(r.o, r.d, g.tri + 3 * k, t1, tl) || t1 >= len) continue;
        return 0.0f;
    }
    return opacity;
}

/* One 8x4 tile, all 32 lanes in lock step.  `live` = this lane owns a pixel of the frame.  Returns the pixel word. */
template <bool COUNT, bool AOV, bool WIDE = false>
__device__ __forceinline__ uint32_t trace_tile_coop(const RtFrameParams& fp, const Geo& g, const CoopSmem& cs, uint32_t id,
                                                    bool live, uint32_t lane, uint32_t& hit_id, uint32_t& nseg, Tally<COUNT>& tl) {
    constexpr uint32_t kFull = 0xffffffffu;
    Saved st[kMaxDepth - 1];
    uint32_t rng = id; /* raytracing.cl:32-33 */
    uint32_t sz = 1;
    V3 o = mk(0.0f, 0.0f, 0.0f), d = mk(0.0f, 0.0f, 1.0f), rgb = mk(0.0f, 0.0f, 0.0f);
    if (live) primary_ray(fp, id, o, d);
    int depth = 0;
    float n_cur = kAirN, f = 1.0f;
    bool first = true, done = !live;
    if (AOV) hit_id = 0;
    const uint32_t lt = (1u << lane) - 1u;
    const uint32_t L = g.L;
    static_assert(!(WIDE && COUNT), "the wide variant does not keep executed-work counters");
    const uint32_t wide_live = WIDE ? wide_max_live(fp, g) : 0u;

    for (;;) {
        if (!done) {
            while (depth >= kMaxDepth) { /* entry finished: fold it into its parent (raytracing.cl:183-190) */
                if (sz == 1) { done = true; break; }
                const Saved& p = st[sz - 2];
                rgb = p.rgb + rgb;
                o = p.o; d = p.d; depth = p.depth; n_cur = p.n; f = p.f;
                sz--;
            }
        }
        if (__all_sync(kFull, done)) break;

        /* ---- phase A: this lane's segment up to the direct-light loop ---- */
        bool shade = false;
        V3 ip = mk(0.0f, 0.0f, 0.0f), nrm = ip, vv = ip;
        Material m;
        m.shininess = 0; m.transperent = 0; m.dielectric = 0;
        m.n = 1.0f; m.reflectivity = 0.0f; m.specular = 0.0f; m.diffuse = 0.0f;
        RayF r;
        Closest c;
        if (!done) {
            tl.add(RTB_SEGMENTS);
            nseg++;
            ray_setup_fast(r, o, d);
            if (!WIDE) closest_pass(r, g, tl, c);
        }
        if (WIDE) closest_pass_warp(r, o, d, !done, g, tl, c, lane, wide_live);
        if (!done) {
            if (c.k >= 0 && !c.occluded) { /* a visible light ends the chain, raytracing.cl:48-54 */
                V3 lp = o + d * c.tL;
                V3 od = o - lp;
                float dd = sqrtf(dot3(od, od));
                V3 col = xyz(g.laux[c.k]) * (1.0f / dd * dd); /* (1/d)*d, primitives.cl:288 */
                rgb = rgb + col * f;
                if (AOV && first) hit_id = RTB_HIT_LIGHT | (uint32_t)c.k;
                depth = kMaxDepth;
            } else if (c.who < 0 && c.whop < 0 && c.whot < 0) { /* skybox, raytracing.cl:61-81 */
                int u, v;
                cube_uv(d, (uint32_t)(fp.sw / 4), u, v);
                tl.add(RTB_TEXELS);
                V3 px = texel_rgb(fp.sky, fp.sw, fp.sh, 1, u, fp.sh - v, 0, g.lut255);
                rgb = rgb + px * f;
                depth = kMaxDepth;
            } else {
                shade = true;
                ip = o + d * c.t;
                if (c.whot >= 0) { /* extension X1 */
                    nrm = triangle_normal(g.tri + 3 * c.whot);
                    m = load_material_at(fp.raw_tris + (size_t)c.whot * RT_TRIANGLE_BYTES + RT_TRI_MATERIAL_OFFSET);
                    if (AOV && first) hit_id = 1u + g.S + g.P + (uint32_t)c.whot;
                } else if (c.whop >= 0) {
                    nrm = xyz(g.pln[4 * c.whop]);
                    m = load_material(fp.raw_planes + (size_t)c.whop * RT_PLANE_BYTES);
                    if (m.texture_id >= 0) { /* primitives.cl:217-259; sampled once, for the plane that won */
                        float4 b0 = g.pln[4 * c.whop + 2], b1 = g.pln[4 * c.whop + 3];
                        float ui = dot3(xyz(b0), ip) * m.texture_scale;
                        float vi = dot3(xyz(b1), ip) * m.texture_scale;
                        tl.add(RTB_TEXELS);
                        m.rgb = texel_rgb(fp.tex, fp.tw, fp.th, fp.tl, emod(f2i_x86(ui), fp.tw), emod(f2i_x86(vi), fp.th),
                                          m.texture_id, g.lut255);
                    }
                    if (AOV && first) hit_id = 1u + g.S + (uint32_t)c.whop;
                } else {
                    float qlen;
                    nrm = normalize_len(ip - xyz(g.sph[c.who]), qlen);
                    m = load_material(fp.raw_spheres + (size_t)c.who * RT_SPHERE_BYTES);
                    if (AOV && first) hit_id = 1u + (uint32_t)c.who;
                }
                ip = ip + nrm * kEps;
                rgb = rgb + (m.rgb * f) * m.ambient; /* raytracing.cl:83-84 */
                float vlen;
                vv = normalize_len(o - ip, vlen);
            }
            first = false;
        }

        /* ---- direct light, raytracing.cl:87-136, in chunks of kCoopLights lights (L is warp-uniform) ---- */
        if (__any_sync(kFull, shade)) {
            const float shin = (float)m.shininess;
            const float fs = f * m.specular, fd = f * m.diffuse;
            if (shade) {
                cs.owner[lane] = make_float4(ip.x, ip.y, ip.z, 0.0f);
                if (kPreS && g.Spad <= (uint32_t)kPreS && g.P <= 4u) { /* R8: origin-dependent halves of this segment's shadow tests */
                    float4* mine = cs.pre + lane * kPreRec;
                    for (uint32_t i = 0; i < g.Spad; i++) mine[i] = sphere_pre(ip, g.sph[i]);
                    float nums[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
                    for (uint32_t j = 0; j < 4u; j++)
                        if (j < g.P) nums[j] = dot3(xyz(g.pln[4 * j + 1]) - ip, xyz(g.pln[4 * j]));
                    mine[kPreS] = make_float4(nums[0], nums[1], nums[2], nums[3]);
                }
            }
            for (uint32_t c0 = 0; c0 < L; c0 += kCoopLights) {
                const uint32_t nk = L - c0 < (uint32_t)kCoopLights ? L - c0 : (uint32_t)kCoopLights; /* warp-uniform */
                uint32_t ntask = 0;
#pragma unroll 1
                for (uint32_t k = 0; k < nk; k++) {
                    const uint32_t i = c0 + k;
                    bool post = false;
                    float th0 = 0.0f, ph0 = 0.0f, th1 = 0.0f, ph1 = 0.0f;
                    float dist = 1.0f, sx = 0.0f, dif = 0.0f;
                    uint32_t bits = 0; /* bit 0: tasks posted; bit 1: spec_zero; bit 2: diff_zero; bit 3: pow may be inf; bits 8..: first task slot */
                    if (shade) {
                        const float4 l0 = g.lgeo[i];
                        V3 sdir = normalize_len(xyz(l0) - ip, dist);
                        float hlen;
                        V3 hh = normalize_len(vv + sdir, hlen);
                        float ndh = dot3(nrm, hh);
                        sx = 0.0f < ndh ? ndh : 0.0f;
                        float ndl = dot3(nrm, sdir);
                        dif = 0.0f < ndl ? ndl : 0.0f;
                        /* R5: exact zeros -- only while pow(sx, shin) is finite (see rt_trace_fast.cuh) */
                        const bool tame = sx <= 1.0f && shin >= 0.0f;
                        if (!tame) bits |= 8u;
                        const bool spec_zero = tame && ((fs == 0.0f) || (sx == 0.0f && shin > 0.0f));
                        const bool diff_zero = (fd == 0.0f) || (dif == 0.0f);
                        if (spec_zero) bits |= 2u;
                        if (diff_zero) bits |= 4u;
                        if (spec_zero && diff_zero) {
                            rng = xorshift_step(xorshift_step(xorshift_step(xorshift_step(rng)))); /* 2 samples x 2 draws */
                            tl.add(RTB_SHADOW_SKIPPED, kShadowSamples);
                        } else {
                            th0 = (float)(kTwoPi * (double)rng_next(rng));
                            ph0 = (float)(kPi * (double)rng_next(rng));
                            th1 = (float)(kTwoPi * (double)rng_next(rng));
                            ph1 = (float)(kPi * (double)rng_next(rng));
                            post = true;
                        }
                    }
                    const uint32_t mask = __ballot_sync(kFull, post);
                    if (post) {
                        const uint32_t a = ntask + 2u * (uint32_t)__popc(mask & lt);
                        bits |= 1u | (a << 8);
                        cs.task[a] = make_float4(th0, ph0, __uint_as_float(i), __uint_as_float(lane));
                        cs.task[a + 1] = make_float4(th1, ph1, __uint_as_float(i), __uint_as_float(lane));
                    }
                    if (shade) cs.light[k * 32u + lane] = make_float4(dist, sx, dif, __uint_as_float(bits));
                    ntask += 2u * (uint32_t)__popc(mask);
                }
                __syncwarp();
                /* ---- phase B: every lane, finished or not, takes tasks ---- */
                uint32_t lane_end = ntask;
                if (WIDE && (ntask & 31u) && (ntask & 31u) <= wide_live) lane_end = ntask & ~31u; /* partial pass: one task, 32 lanes */
#pragma unroll 1
                for (uint32_t t = lane; t < lane_end; t += 32u) {
                    const float4 tk = cs.task[t];
                    const float op = coop_run_task(g, cs, tk, tl);
                    reinterpret_cast<float*>(cs.task + t)[0] = op;
                }
                if (WIDE) {
#pragma unroll 1
                    for (uint32_t t = lane_end; t < ntask; t++) {
                        const float4 tk = cs.task[t];
                        const float op = coop_run_task_warp(g, cs, tk, lane, tl);
                        if (lane == 0u) reinterpret_cast<float*>(cs.task + t)[0] = op;
                    }
                }
                __syncwarp();
                /* ---- phase C: owners fold their lights in, in order ---- */
                if (shade) {
#pragma unroll 1
                    for (uint32_t k = 0; k < nk; k++) {
                        const float4 rec = cs.light[k * 32u + lane];
                        const uint32_t bits = __float_as_uint(rec.w);
                        if (!(bits & 1u)) continue;
                        const uint32_t a = bits >> 8;
                        const float r0 = reinterpret_cast<const float*>(cs.task + a)[0];
                        const float r1 = reinterpret_cast<const float*>(cs.task + a + 1)[0];
                        const float soft = (0.0f + r0) + r1;
                        if (soft == 0.0f && !(bits & 8u)) continue; /* light_rgb *= 0: both terms are exact zeros */
                        const float ssr = soft / (float)kShadowSamples;
                        const float4 l1 = g.laux[c0 + k];
                        V3 lrgb = div3_pos(xyz(l1) * 1.0f, rec.x * rec.x) * ssr;
                        if (!(bits & 2u)) rgb = rgb + (lrgb * fs) * pow_ieee(rec.y, shin);
                        if (!(bits & 4u)) rgb = rgb + (lrgb * fd) * rec.z;
                    }
                }
                __syncwarp(); /* the next chunk / round reuses the task slots */
            }
        }

        /* ---- Fresnel, reflection, refraction child (raytracing.cl:138-179) ---- */
        if (shade) {
            V3 incident = d;
            float n1 = n_cur;
            float n2 = (n1 == kAirN) ? m.n : kAirN;
            float cosI = -dot3(nrm, incident);
            float refl = reflect_amount(m, n1, n2, cosI);
            float old_f = f;
            f *= refl;
            d = incident + nrm * (2.0f * cosI); /* primitives.cl:127-130 */
            o = ip;
            depth++;
            if (m.transperent && sz < (uint32_t)kMaxDepth && refl < 1.0f) {
                V3 co, cd;
                if (refract_child(o, nrm, incident, n1, n2, co, cd)) {
                    Saved& p = st[sz - 1];
                    p.o = o; p.d = d; p.rgb = rgb; p.depth = depth; p.n = n_cur; p.f = f;
                    sz++;
                    o = co; d = cd; rgb = mk(0.0f, 0.0f, 0.0f);
                    f = old_f * (1.0f - refl);
                    n_cur = n2;
                }
            }
        }
    }
    return pack_rgb(rgb);
}

}  // namespace RT_NS
