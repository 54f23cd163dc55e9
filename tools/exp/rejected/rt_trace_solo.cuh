/* NOT BUILT.  Kept as the record of a measured experiment: bit-identical frames (17 GPU tests:
 * every tile forced into single-pixel entries on render.map at 1024x512, three cameras, and on 16 fuzz scenes at
 * 512x256) but SLOWER on render.map at 4K: slowest rank's share N=4 0.87 -> 1.00 ms, N=8 0.60 -> 0.70 ms, whole frame
 * unchanged (no single-pixel entries at the automatic thresholds); inlining it also cost the main kernel 1.2 %
 * (2.732 -> 2.769 ms per frame).  The RT_SOLO_PROFILE clocks below say why: for a 109-segment glass pixel (75 shaded,
 * 16 batches) the WALK took 816,541 clocks, the batched shading 121,138 and the replay 110,558; the segment-per-round
 * walk of rt_trace_coop.cuh needs 1,053,166 for the same pixel.  A heavy pixel's chain is ~7,500 clocks per step of the
 * walk itself (closest hit, material and texel loads, Fresnel, stack) -- batching the shading cannot shorten that.
 * What would: walking the reflection and refraction sub-trees of a pixel on different lanes (the geometry of the
 * two is independent; oracle/rt_split.c shows how the RNG ordinals and the addition order are rebuilt afterwards).
 * To build it again: include it after rt_trace_coop.cuh, add `uint32_t solo` to RtFrameParams, and in render_frame
 * call trace_pixel_solo for entries with sub >= 5 when (!COUNT && !WIDE && fp.solo && 1 <= g.L <= 16), passing the
 * live lane's id by __shfl_sync (git history: the commit after "oracle/rt_split.c"). */
/* rt_trace_solo.cuh -- one heavy pixel, a whole warp: segments shaded in batches.
 *
 * The heaviest tiles are queued one pixel per warp (rt_kernels.cu: split_ways == 32).  In trace_tile_coop such a
 * warp runs one segment per round: one closest-hit pass, then 2L shadow tasks on 2L of its 32 lanes, then the
 * fold -- and the kernel's tail is that chain.  The reference's stack machine splits exactly
 * (oracle/rt_split.c, tests/test_split_cpu.py): the walk over hits / reflections / refraction children
 * (raytracing.cl:41-81,138-190) reads neither the RNG nor the accumulated colour; every shaded segment draws
 * exactly 4L numbers (raytracing.cl:96-113, no early exit), so the RNG state at a segment is known from its
 * ordinal alone; and the colour is a sequence of additions whose ORDER is fixed by the walk.  So here
 *   walk    all 32 lanes walk the same ray (uniform control flow: nothing diverges, nothing is broadcast) for up
 *           to B = min(8, 32 / 2L) shaded segments, leaving per segment a record {hit point, normal, view vector,
 *           f*specular, f*diffuse, shininess, RNG state} and, in program order, an event list
 *           {ADD term | SHADE b | PUSH | POP} in the warp's shared memory;
 *   shade   lane j takes (segment j / 2L, light (j % 2L) / 2, sample j % 2): shading geometry, its two draws,
 *           testShadowPath; the even lane of a pair folds the two opacities into the light's specular and
 *           diffuse terms (raytracing.cl:115-135) -- B*2L shadow rays and B*L pow() per pass instead of 2L and L;
 *   replay  every lane replays the additions in event order on a stack of partial sums (raytracing.cl:188).
 * Same operations on the same operands in the same order as trace_tile_coop (incl. its exact-zero shortcuts R5):
 * identical words.  Used for single-pixel queue entries when fp.solo != 0, 1 <= L <= 16, not in the WIDE or
 * counting variants.
 */
#pragma once
#include "rt_trace_coop.cuh"

namespace RT_NS {

constexpr uint32_t kSoloMaxSeg = 8;    /* records: cs.owner[0..31] = 4 float4 per segment */
constexpr uint32_t kSoloEvents = 32 * kCoopLights; /* event slots: the cs.light area (128 float4) */
constexpr uint32_t kSoloEvAdd = 0u, kSoloEvShade = 1u, kSoloEvPush = 2u, kSoloEvPop = 3u;
static_assert(kSoloEvents >= 64, "event list too short for one walk step (<= 17 events) per check");
static_assert(kCoopTasks >= 32, "term slots");

template <bool AOV>
__device__ __forceinline__ uint32_t trace_pixel_solo(const RtFrameParams& fp, const Geo& g, const CoopSmem& cs, uint32_t id,
                                                  uint32_t lane, uint32_t& hit_id, uint32_t& nseg) {
    constexpr uint32_t kFull = 0xffffffffu;
    Tally<false> tl;
    Saved st[kMaxDepth - 1];     /* walker: o, d, depth, n, f; replay: rgb */
    float4* const ev = cs.light;
    const uint32_t L = g.L, T = 2u * L;
    const uint32_t B = (32u / T) < kSoloMaxSeg ? (32u / T) : kSoloMaxSeg;

    uint32_t rng = id;           /* state at the start of the next shaded segment, raytracing.cl:32-33 */
    uint32_t sz = 1, rsz = 1;    /* walker / replay stack heights */
    V3 o, d, rgb = mk(0.0f, 0.0f, 0.0f);
    primary_ray(fp, id, o, d);
    int depth = 0;
    float n_cur = kAirN, f = 1.0f;
    bool first = true, finished = false;
    if (AOV) hit_id = 0;

#ifdef RT_SOLO_PROFILE
    long long pc_walk = 0, pc_shade = 0, pc_replay = 0;
    uint32_t pc_batches = 0, pc_shaded = 0, pc_events = 0;
#endif
    while (!finished) {
        uint32_t nev = 0, nb = 0;
#ifdef RT_SOLO_PROFILE
        long long pc0 = clock64();
#endif
        /* ---- walk: uniform over the warp; lane 0 writes the lists ---- */
        while (nb < B && nev + 20u <= kSoloEvents) {
            while (depth >= kMaxDepth) { /* entry finished: fold it into its parent (raytracing.cl:183-190) */
                if (sz == 1) { finished = true; break; }
                const Saved& p = st[sz - 2];
                o = p.o; d = p.d; depth = p.depth; n_cur = p.n; f = p.f;
                sz--;
                if (lane == 0u) ev[nev] = make_float4(0.0f, 0.0f, 0.0f, __uint_as_float(kSoloEvPop));
                nev++;
            }
            if (finished) break;
            nseg++;
            RayF r;
            Closest c;
            ray_setup_fast(r, o, d);
            closest_pass(r, g, tl, c);
            if (c.k >= 0 && !c.occluded) { /* a visible light ends the chain, raytracing.cl:48-54 */
                V3 lp = o + d * c.tL;
                V3 od = o - lp;
                float dd = sqrtf(dot3(od, od));
                V3 col = xyz(g.laux[c.k]) * (1.0f / dd * dd); /* (1/d)*d, primitives.cl:288 */
                V3 term = col * f;
                if (lane == 0u) ev[nev] = make_float4(term.x, term.y, term.z, __uint_as_float(kSoloEvAdd));
                nev++;
                if (AOV && first) hit_id = RTB_HIT_LIGHT | (uint32_t)c.k;
                depth = kMaxDepth;
                first = false;
                continue;
            }
            if (c.who < 0 && c.whop < 0 && c.whot < 0) { /* skybox, raytracing.cl:61-81 */
                int u, v;
                cube_uv(d, (uint32_t)(fp.sw / 4), u, v);
                V3 px = texel_rgb(fp.sky, fp.sw, fp.sh, 1, u, fp.sh - v, 0, g.lut255);
                V3 term = px * f;
                if (lane == 0u) ev[nev] = make_float4(term.x, term.y, term.z, __uint_as_float(kSoloEvAdd));
                nev++;
                depth = kMaxDepth;
                first = false;
                continue;
            }
            V3 ip = o + d * c.t, nrm;
            Material m;
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
            first = false;
            ip = ip + nrm * kEps;
            {
                V3 amb = (m.rgb * f) * m.ambient; /* raytracing.cl:83-84 */
                float vlen;
                V3 vv = normalize_len(o - ip, vlen);
                if (lane == 0u) {
                    ev[nev] = make_float4(amb.x, amb.y, amb.z, __uint_as_float(kSoloEvAdd));
                    ev[nev + 1u] = make_float4(__uint_as_float(nb), 0.0f, 0.0f, __uint_as_float(kSoloEvShade));
                    cs.owner[nb] = make_float4(ip.x, ip.y, ip.z, f * m.specular);
                    cs.owner[8u + nb] = make_float4(nrm.x, nrm.y, nrm.z, f * m.diffuse);
                    cs.owner[16u + nb] = make_float4(vv.x, vv.y, vv.z, (float)m.shininess);
                    cs.owner[24u + nb] = make_float4(__uint_as_float(rng), 0.0f, 0.0f, 0.0f);
                }
                nev += 2u;
                nb++;
#pragma unroll 1
                for (uint32_t q = 0; q < T; q++) rng = xorshift_step(xorshift_step(rng)); /* 4L draws per shaded segment */
            }
            /* ---- Fresnel, reflection, refraction child (raytracing.cl:138-179) ---- */
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
                    p.o = o; p.d = d; p.depth = depth; p.n = n_cur; p.f = f;
                    sz++;
                    o = co; d = cd;
                    f = old_f * (1.0f - refl);
                    n_cur = n2;
                    if (lane == 0u) ev[nev] = make_float4(0.0f, 0.0f, 0.0f, __uint_as_float(kSoloEvPush));
                    nev++;
                }
            }
        }
        __syncwarp();
#ifdef RT_SOLO_PROFILE
        long long pc1 = clock64();
#endif

        /* ---- shade: lane j = (segment, light, sample) ---- */
        if (nb) {
            const uint32_t b = lane / T, q = lane - b * T, i = q >> 1, s = q & 1u;
            const bool mine = b < nb;
            float op = 0.0f, dist = 1.0f, sx = 0.0f, dif = 0.0f, fs = 0.0f, fd = 0.0f, shin = 0.0f;
            bool tame = true, spec_zero = true, diff_zero = true;
            if (mine) {
                const float4 r0 = cs.owner[b], r1 = cs.owner[8u + b], r2 = cs.owner[16u + b];
                uint32_t rs = __float_as_uint(cs.owner[24u + b].x);
                const V3 ip = xyz(r0), nrm = xyz(r1), vv = xyz(r2);
                fs = r0.w; fd = r1.w; shin = r2.w;
                const float4 l0 = g.lgeo[i];
                V3 sdir = normalize_len(xyz(l0) - ip, dist);
                float hlen;
                V3 hh = normalize_len(vv + sdir, hlen);
                float ndh = dot3(nrm, hh);
                sx = 0.0f < ndh ? ndh : 0.0f;
                float ndl = dot3(nrm, sdir);
                dif = 0.0f < ndl ? ndl : 0.0f;
                /* R5: exact zeros -- only while pow(sx, shin) is finite (see rt_trace_fast.cuh) */
                tame = sx <= 1.0f && shin >= 0.0f;
                spec_zero = tame && ((fs == 0.0f) || (sx == 0.0f && shin > 0.0f));
                diff_zero = (fd == 0.0f) || (dif == 0.0f);
                if (!(spec_zero && diff_zero)) {
#pragma unroll 1
                    for (uint32_t k = 0; k < q; k++) rs = xorshift_step(xorshift_step(rs)); /* draws of the samples before mine */
                    const float theta = (float)(kTwoPi * (double)rng_next(rs));
                    const float phi = (float)(kPi * (double)rng_next(rs));
                    const float radius = g.laux[i].w;
                    float sp, cp, sth, cth;
                    sincos_bounded(phi, sp, cp);
                    sincos_bounded(theta, sth, cth);
                    const float rr = radius * sp;
                    const V3 sample = xyz(l0) + mk(rr * cth, rr * sth, radius * cp); /* raytracing.cl:101-106 */
                    op = shadow_path_fast(sample, ip, g, tl);
                }
            }
            const float op1 = __shfl_down_sync(kFull, op, 1);
            if (mine && s == 0u) {
                V3 ts = mk(0.0f, 0.0f, 0.0f), td = ts;
                float as = 0.0f, ad = 0.0f; /* 1 = the term is added (the others are exact zeros the fold skips) */
                if (!(spec_zero && diff_zero)) {
                    const float soft = (0.0f + op) + op1;
                    if (!(soft == 0.0f && tame)) { /* light_rgb *= 0: both terms are exact zeros */
                        const float ssr = soft / (float)kShadowSamples;
                        const float4 l1 = g.laux[i];
                        V3 lrgb = div3_pos(xyz(l1) * 1.0f, dist * dist) * ssr;
                        if (!spec_zero) { ts = (lrgb * fs) * pow_ieee(sx, shin); as = 1.0f; }
                        if (!diff_zero) { td = (lrgb * fd) * dif; ad = 1.0f; }
                    }
                }
                cs.task[(b * L + i) * 2u] = make_float4(ts.x, ts.y, ts.z, as);
                cs.task[(b * L + i) * 2u + 1u] = make_float4(td.x, td.y, td.z, ad);
            }
        }
        __syncwarp();
#ifdef RT_SOLO_PROFILE
        long long pc2 = clock64();
#endif

        /* ---- replay: additions in the reference's order (every lane, same values) ---- */
#pragma unroll 1
        for (uint32_t e = 0; e < nev; e++) {
            const float4 v = ev[e];
            const uint32_t kind = __float_as_uint(v.w);
            if (kind == kSoloEvAdd) {
                rgb = rgb + xyz(v);
            } else if (kind == kSoloEvShade) {
                const uint32_t b = __float_as_uint(v.x);
#pragma unroll 1
                for (uint32_t t = 0; t < T; t++) {
                    const float4 term = cs.task[b * T + t];
                    if (term.w != 0.0f) rgb = rgb + xyz(term);
                }
            } else if (kind == kSoloEvPush) {
                st[rsz - 1].rgb = rgb;
                rsz++;
                rgb = mk(0.0f, 0.0f, 0.0f);
            } else {
                rgb = st[rsz - 2].rgb + rgb;
                rsz--;
            }
        }
        __syncwarp(); /* the next batch reuses the lists */
#ifdef RT_SOLO_PROFILE
        {
            long long pc3 = clock64();
            pc_walk += pc1 - pc0; pc_shade += pc2 - pc1; pc_replay += pc3 - pc2;
            pc_batches++; pc_shaded += nb; pc_events += nev;
        }
#endif
    }
#ifdef RT_SOLO_PROFILE
    if (lane == 0u && nseg > 60u && (id % 31u) == 0u)
        printf("SOLO id %u nseg %u shaded %u batches %u events %u walk %lld shade %lld replay %lld\n", id, nseg, pc_shaded,
               pc_batches, pc_events, pc_walk, pc_shade, pc_replay);
#endif
    return pack_rgb(rgb);
}

}  // namespace RT_NS
