/* NOT BUILT.  Kept as the record of a measured experiment: bit-identical frames (all GPU
 * parity tests passed) but slower -- 4K kernel 2.71 -> 3.15..3.28 ms with K forced to 1 or not, one rank's share
 * at N=8 0.683 -> 0.715 ms; ncu: +6 % warp instructions, issue-slot use 70 -> 62 %.  To build it again CoopSmem
 * needs a member `float4* own2` ([2][32], +1 KiB per warp) and render_frame must call trace_tile_deep. */
/* rt_trace_deep.cuh -- cooperative trace with DEFERRED colour: several segments of a pixel per round.
 *
 * rt_trace_coop.cuh walks a tile one segment per live lane per round.  When few lanes are still alive -- the
 * tail of every tile, and the whole life of the split entries of the heaviest tiles (one tile row or one
 * pixel per warp, rt_kernels.cu) -- a round has only 2L shadow tasks per live lane for 32 lanes, and the
 * owner's closest-hit pass and the shadow pass alternate, each waiting for the other.
 *
 * Nothing in a segment's GEOMETRY depends on its shadow rays: the next origin, direction, depth, Fresnel
 * factor and the generator state are known after the closest-hit pass (raytracing.cl:138-179 use the hit and
 * the material only; every shaded segment draws exactly 4 numbers per light, :99-100).  Only the COLOUR
 * needs the opacities, and colour is a sequence of additions.  So with nl live lanes a round is
 *   A  each live lane runs up to K = 32/nl (power of two, <= kDeepMax) consecutive segments of its pixel:
 *      pops, closest hit, material, per-light prologue, task posting, Fresnel/reflect/refract, push.  What
 *      the colour will need goes into the segment's record v = rank*K + j in shared memory:
 *      {pops before the segment, one term to add (light / sky / ambient), specular and diffuse factors,
 *      shininess, "a refraction child was pushed"}.
 *   B  all shadow tasks of all those segments, dealt over 32 lanes (same routine as rt_trace_coop.cuh).
 *   C  each lane replays its records IN ORDER on its colour register and its own colour stack: the same
 *      additions with the same operands in the same order as raytracing.cl:48-84,118-135,183-190, hence the
 *      same words.
 * K = 1 when more than 16 lanes live: that is rt_trace_coop.cuh's round with the Fresnel step moved up.
 * Scenes with more lights than one chunk (L > kCoopLights) take rt_trace_coop.cuh's chunked path.
 */
#pragma once
#include "rt_trace_coop.cuh"

namespace RT_NS {

#ifndef RT_DEEP_MAX
#define RT_DEEP_MAX 8
#endif
constexpr uint32_t kDeepMax = RT_DEEP_MAX;

struct SavedGeo {
    V3 o, d;
    int depth;
    float n, f;
};

/* record flags */
constexpr uint32_t kRecPops = 15u, kRecAdd = 16u, kRecShade = 32u, kRecPush = 64u;

template <bool COUNT, bool AOV>
__device__ __forceinline__ uint32_t trace_tile_deep(const RtFrameParams& fp, const Geo& g, const CoopSmem& cs, uint32_t id,
                                                    bool live, uint32_t lane, uint32_t& hit_id, uint32_t& nseg, Tally<COUNT>& tl) {
    constexpr uint32_t kFull = 0xffffffffu;
    const uint32_t L = g.L;
#ifndef RT_DEEP_ONLY
    if (L > (uint32_t)kCoopLights) return trace_tile_coop<COUNT, AOV>(fp, g, cs, id, live, lane, hit_id, nseg, tl);
#endif

    SavedGeo st[kMaxDepth - 1];
    V3 strgb[kMaxDepth - 1];
    uint32_t rng = id; /* raytracing.cl:32-33 */
    uint32_t sz = 1, csz = 0;
    V3 o = mk(0.0f, 0.0f, 0.0f), d = mk(0.0f, 0.0f, 1.0f), rgb = mk(0.0f, 0.0f, 0.0f);
    if (live) primary_ray(fp, id, o, d);
    int depth = 0;
    float n_cur = kAirN, f = 1.0f;
    bool first = true, done = !live;
    if (AOV) hit_id = 0;
    const uint32_t lt = (1u << lane) - 1u;

    for (;;) {
        const uint32_t livemask = __ballot_sync(kFull, !done);
        if (!livemask) break;
        const uint32_t nl = (uint32_t)__popc(livemask);
        uint32_t K = nl > 16u ? 1u : nl > 8u ? 2u : nl > 4u ? 4u : nl > 2u ? 8u : nl > 1u ? 16u : 32u;
        if (K > kDeepMax) K = kDeepMax;
        const uint32_t base = K == 1u ? lane : (uint32_t)__popc(livemask & lt) * K;

        /* ---- phase A: up to K segments of this lane's pixel ---- */
        uint32_t nrec = 0, ntask = 0;
        bool active = !done;
#pragma unroll 1
        for (uint32_t j = 0; j < K; j++) {
            const uint32_t v = base + j;
            bool shade = false;
            uint32_t flags = 0;
            V3 term = mk(0.0f, 0.0f, 0.0f);
            V3 ip = term, nrm = term, vv = term;
            Material m;
            m.shininess = 0; m.transperent = 0; m.dielectric = 0;
            m.n = 1.0f; m.reflectivity = 0.0f; m.specular = 0.0f; m.diffuse = 0.0f;
            if (active) {
                nrec = j + 1u;
                while (depth >= kMaxDepth) { /* entry finished: its parent takes over (raytracing.cl:183-190) */
                    if (sz == 1) { done = true; break; }
                    const SavedGeo& p = st[sz - 2];
                    o = p.o; d = p.d; depth = p.depth; n_cur = p.n; f = p.f;
                    sz--;
                    flags++;
                }
                if (done) active = false;
            }
            if (active) {
                tl.add(RTB_SEGMENTS);
                nseg++;
                RayF r;
                ray_setup_fast(r, o, d);
                Closest c;
                closest_pass(r, g, tl, c);
                if (c.k >= 0 && !c.occluded) { /* a visible light ends the chain, raytracing.cl:48-54 */
                    V3 lp = o + d * c.tL;
                    V3 od = o - lp;
                    float dd = sqrtf(dot3(od, od));
                    V3 col = xyz(g.laux[c.k]) * (1.0f / dd * dd); /* (1/d)*d, primitives.cl:288 */
                    term = col * f;
                    flags |= kRecAdd;
                    if (AOV && first) hit_id = RTB_HIT_LIGHT | (uint32_t)c.k;
                    depth = kMaxDepth;
                } else if (c.who < 0 && c.whop < 0 && c.whot < 0) { /* skybox, raytracing.cl:61-81 */
                    int u, w;
                    cube_uv(d, (uint32_t)(fp.sw / 4), u, w);
                    tl.add(RTB_TEXELS);
                    V3 px = texel_rgb(fp.sky, fp.sw, fp.sh, 1, u, fp.sh - w, 0, g.lut255);
                    term = px * f;
                    flags |= kRecAdd;
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
                    term = (m.rgb * f) * m.ambient; /* raytracing.cl:83-84 */
                    flags |= kRecAdd | kRecShade;
                    float vlen;
                    vv = normalize_len(o - ip, vlen);
                }
                first = false;
            }

            /* ---- direct light prologue, raytracing.cl:87-107 (L <= kCoopLights here) ---- */
            const float shin = (float)m.shininess;
            const float fs = f * m.specular, fd = f * m.diffuse;
            if (shade) cs.owner[v] = make_float4(ip.x, ip.y, ip.z, shin);
#pragma unroll 1
            for (uint32_t k = 0; k < L; k++) {
                bool post = false;
                float th0 = 0.0f, ph0 = 0.0f, th1 = 0.0f, ph1 = 0.0f;
                float dist = 1.0f, sx = 0.0f, dif = 0.0f;
                uint32_t bits = 0; /* bit 0: tasks posted; bit 1: spec_zero; bit 2: diff_zero; bit 3: pow may be inf; bits 8..: first task slot */
                if (shade) {
                    const float4 l0 = g.lgeo[k];
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
                    cs.task[a] = make_float4(th0, ph0, __uint_as_float(k), __uint_as_float(v));
                    cs.task[a + 1] = make_float4(th1, ph1, __uint_as_float(k), __uint_as_float(v));
                }
                if (shade) cs.light[k * 32u + v] = make_float4(dist, sx, dif, __uint_as_float(bits));
                ntask += 2u * (uint32_t)__popc(mask);
            }

            /* ---- Fresnel, reflection, refraction child (raytracing.cl:138-179): geometry only ---- */
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
                        SavedGeo& p = st[sz - 1];
                        p.o = o; p.d = d; p.depth = depth; p.n = n_cur; p.f = f;
                        sz++;
                        flags |= kRecPush; /* the colour so far is parked, the child starts from black */
                        o = co; d = cd;
                        f = old_f * (1.0f - refl);
                        n_cur = n2;
                    }
                }
            }
            if (nrec == j + 1u) {
                cs.own2[v] = make_float4(term.x, term.y, term.z, __uint_as_float(flags));
                cs.own2[32u + v] = make_float4(fs, fd, 0.0f, 0.0f);
            }
            if (!__any_sync(kFull, active)) break;
        }
        __syncwarp();

        /* ---- phase B: every lane, finished or not, takes tasks ---- */
#pragma unroll 1
        for (uint32_t t = lane; t < ntask; t += 32u) {
            const float4 tk = cs.task[t];
            const float op = coop_run_task(g, cs, tk, tl);
            reinterpret_cast<float*>(cs.task + t)[0] = op;
        }
        __syncwarp();

        /* ---- phase C: replay this lane's records on its colour, in order ---- */
#pragma unroll 1
        for (uint32_t j = 0; j < nrec; j++) {
            const uint32_t v = base + j;
            const float4 r1 = cs.own2[v];
            const uint32_t flags = __float_as_uint(r1.w);
            for (uint32_t p = flags & kRecPops; p; p--) { /* raytracing.cl:186-188 */
                csz--;
                rgb = strgb[csz] + rgb;
            }
            if (flags & kRecAdd) rgb = rgb + xyz(r1);
            if (flags & kRecShade) {
                const float4 r2 = cs.own2[32u + v];
                const float fs = r2.x, fd = r2.y, shin = cs.owner[v].w;
#pragma unroll 1
                for (uint32_t k = 0; k < L; k++) {
                    const float4 rec = cs.light[k * 32u + v];
                    const uint32_t bits = __float_as_uint(rec.w);
                    if (!(bits & 1u)) continue;
                    const uint32_t a = bits >> 8;
                    const float r0 = reinterpret_cast<const float*>(cs.task + a)[0];
                    const float rr = reinterpret_cast<const float*>(cs.task + a + 1)[0];
                    const float soft = (0.0f + r0) + rr;
                    if (soft == 0.0f && !(bits & 8u)) continue; /* light_rgb *= 0: both terms are exact zeros */
                    const float ssr = soft / (float)kShadowSamples;
                    const float4 l1 = g.laux[k];
                    V3 lrgb = div3_pos(xyz(l1) * 1.0f, rec.x * rec.x) * ssr;
                    if (!(bits & 2u)) rgb = rgb + (lrgb * fs) * pow_ieee(rec.y, shin);
                    if (!(bits & 4u)) rgb = rgb + (lrgb * fd) * rec.z;
                }
            }
            if (flags & kRecPush) {
                strgb[csz] = rgb;
                csz++;
                rgb = mk(0.0f, 0.0f, 0.0f);
            }
        }
        __syncwarp(); /* the next round reuses the records and the task slots */
    }
    return pack_rgb(rgb);
}

}  // namespace RT_NS
