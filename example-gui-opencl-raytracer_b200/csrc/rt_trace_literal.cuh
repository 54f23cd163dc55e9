/* rt_trace_literal.cuh -- the "literal" trace: the reference's control flow, test for test.
 *
 * Same loops, same early-outs and the same number of intersection tests as raytracing.cl:14-195 with
 * primitives.cl:262-442, so its work counters ARE the reference's counts (rays,
 * algorithmic flop).  It is the accounting variant behind rtb_render_counted and the in-library
 * cross-check of the fast trace (rt_trace_fast.cuh must produce the same words, bit for bit).
 */
#pragma once
#include "rt_device.cuh"

namespace RT_NS {

/* The ray being traced plus hoisted sphere-test invariants (bit-identical rewrites, see rt_kernels.cu). */
struct Ray {
    V3 o, d;
    V3 d2;        /* 2*d : dot(2*v, d) == dot(v, 2*d) bit for bit (scaling by 2 is exact) */
    float a4, a2; /* 4*dot(d,d), 2*dot(d,d)  (primitives.cl:173,179,189) */
};
__device__ __forceinline__ void ray_setup(Ray& r, V3 o, V3 d) {
    r.o = o;
    r.d = d;
    float a = dot3(d, d);
    r.d2 = d * 2.0f;
    r.a4 = 4.0f * a;
    r.a2 = 2.0f * a;
}

/* primitives.cl:170-195 with g = {centre, r*r}.  Comparison forms are kept (NaN behaviour). */
template <bool COUNT>
__device__ __forceinline__ bool hit_sphere(const Ray& r, float4 g, float& t, Tally<COUNT>& tl) {
    tl.add(RTB_SPHERE_TESTS);
    float vx = r.o.x - g.x, vy = r.o.y - g.y, vz = r.o.z - g.z;
    float b = (vx * r.d2.x + vy * r.d2.y) + vz * r.d2.z;
    float c = ((vx * vx + vy * vy) + vz * vz) - g.w;
    float D = b * b - r.a4 * c;
    if (D < 0.0f) return false;
    tl.add(RTB_SPHERE_DPOS);
    D = sqrtf(D);
    float nearT = (-b - D) / r.a2;
    float t2 = (nearT < 0.0f) ? (-b + D) / r.a2 : nearT;
    if (t2 <= 0.0f) return false;
    t = t2;
    return true;
}
/* primitives.cl:197-215 */
template <bool COUNT>
__device__ __forceinline__ bool hit_plane(const Ray& r, V3 n, V3 p, float& t, Tally<COUNT>& tl) {
    tl.add(RTB_PLANE_TESTS);
    float b = dot3(r.d, n);
    if (b == 0.0f) return false;
    float t2 = dot3(p - r.o, n) / b;
    if (t2 <= 0.0f) return false;
    t = t2;
    return true;
}

/* primitives.cl:396-442 */
template <bool COUNT>
__device__ __forceinline__ float shadow_path_literal(V3 to, V3 from, const Geo& g, Tally<COUNT>& tl) {
    tl.add(RTB_SHADOW_RAYS);
    V3 diff = to - from;
    float len = sqrtf(dot3(diff, diff)); /* distance(to, from) and normalize() share it */
    Ray r;
    ray_setup(r, from, diff / len);
    float opacity = 1.0f;
    for (uint32_t i = 0; i < g.S; i++) {
        float t1;
        if (!hit_sphere(r, g.sph[i], t1, tl) || t1 >= len) continue;
        if (g.flag[i]) { opacity *= kGlassPass; continue; }
        return 0.0f;
    }
    for (uint32_t j = 0; j < g.P; j++) {
        float t1;
        if (!hit_plane(r, xyz(g.pln[4 * j]), xyz(g.pln[4 * j + 1]), t1, tl) || t1 >= len) continue;
        return 0.0f;
    }
    for (uint32_t k = 0; k < g.T; k++) { /* extension X1 */
        float t1;
        if (!hit_triangle(r.o, r.d, g.tri + 3 * k, t1, tl) || t1 >= len) continue;
        return 0.0f;
    }
    return opacity;
}

template <bool COUNT, bool AOV>
__device__ __forceinline__ uint32_t trace_pixel_literal(const RtFrameParams& fp, const Geo& g, uint32_t id,
                                                        uint32_t& hit_id, Tally<COUNT>& tl) {
    Saved st[kMaxDepth - 1];
    uint32_t rng = id; /* raytracing.cl:32-33 */
    uint32_t sz = 1;
    V3 o, d, rgb = mk(0.0f, 0.0f, 0.0f);
    primary_ray(fp, id, o, d);
    int depth = 0;
    float n_cur = kAirN, f = 1.0f;
    bool first = true;
    if (AOV) hit_id = 0;

    for (;;) {
        if (depth >= kMaxDepth) { /* entry finished: fold it into its parent (raytracing.cl:183-190) */
            if (sz == 1) break;
            const Saved& p = st[sz - 2];
            rgb = p.rgb + rgb;
            o = p.o; d = p.d; depth = p.depth; n_cur = p.n; f = p.f;
            sz--;
            continue;
        }
        tl.add(RTB_SEGMENTS);
        Ray r;
        ray_setup(r, o, d);

        /* ---- findLightIntersection, primitives.cl:262-318 ---- */
        {
            float tL = CUDART_INF_F;
            int k = -1;
            for (uint32_t i = 0; i < g.L; i++) {
                float t1;
                if (!hit_sphere(r, g.lgeo[i], t1, tl) || t1 >= tL) continue;
                tL = t1;
                k = (int)i;
            }
            if (k >= 0) {
                bool blocked = false;
                for (uint32_t i = 0; i < g.S && !blocked; i++) {
                    float t1;
                    if (hit_sphere(r, g.sph[i], t1, tl) && t1 <= tL && !g.flag[i]) blocked = true;
                }
                for (uint32_t j = 0; j < g.P && !blocked; j++) {
                    float t1;
                    if (hit_plane(r, xyz(g.pln[4 * j]), xyz(g.pln[4 * j + 1]), t1, tl) && t1 <= tL) blocked = true;
                }
                for (uint32_t q = 0; q < g.T && !blocked; q++) { /* extension X1 */
                    float t1;
                    if (hit_triangle(r.o, r.d, g.tri + 3 * q, t1, tl) && t1 <= tL) blocked = true;
                }
                if (!blocked) {
                    V3 ip = o + d * tL;
                    V3 od = o - ip;
                    float dd = sqrtf(dot3(od, od));
                    V3 col = xyz(g.laux[k]) * (1.0f / dd * dd); /* (1/d)*d, primitives.cl:288 */
                    rgb = rgb + col * f;
                    if (AOV && first) hit_id = RTB_HIT_LIGHT | (uint32_t)k;
                    depth = kMaxDepth; /* `break` of raytracing.cl:53 */
                    first = false;
                    continue;
                }
            }
        }

        /* ---- findSolidIntersection, primitives.cl:322-394 (strict `_t >= t` rejection) ---- */
        float t = CUDART_INF_F;
        int who = -1;
        for (uint32_t i = 0; i < g.S; i++) {
            float t1;
            if (!hit_sphere(r, g.sph[i], t1, tl) || t1 >= t) continue;
            t = t1;
            who = (int)i;
        }
        int whop = -1;
        for (uint32_t j = 0; j < g.P; j++) {
            float t1;
            if (!hit_plane(r, xyz(g.pln[4 * j]), xyz(g.pln[4 * j + 1]), t1, tl) || t1 >= t) continue;
            t = t1;
            whop = (int)j;
            if (COUNT && __float_as_int(g.pln[4 * j + 3].w) >= 0) tl.add(RTB_TEXELS);
        }

        int whot = -1;
        for (uint32_t q = 0; q < g.T; q++) { /* extension X1 */
            float t1;
            if (!hit_triangle(r.o, r.d, g.tri + 3 * q, t1, tl) || t1 >= t) continue;
            t = t1;
            whot = (int)q;
        }

        if (who < 0 && whop < 0 && whot < 0) { /* skybox, raytracing.cl:61-81 */
            int u, v;
            cube_uv(d, (uint32_t)(fp.sw / 4), u, v);
            tl.add(RTB_TEXELS);
            V3 px = texel_rgb(fp.sky, fp.sw, fp.sh, 1, u, fp.sh - v, 0, g.lut255);
            rgb = rgb + px * f;
            depth = kMaxDepth;
            first = false;
            continue;
        }

        V3 ip = o + d * t, nrm;
        Material m;
        if (whot >= 0) { /* extension X1 */
            nrm = triangle_normal(g.tri + 3 * whot);
            m = load_material_at(fp.raw_tris + (size_t)whot * RT_TRIANGLE_BYTES + RT_TRI_MATERIAL_OFFSET);
            if (AOV && first) hit_id = 1u + g.S + g.P + (uint32_t)whot;
        } else if (whop >= 0) {
            nrm = xyz(g.pln[4 * whop]);
            m = load_material(fp.raw_planes + (size_t)whop * RT_PLANE_BYTES);
            if (m.texture_id >= 0) { /* primitives.cl:217-259 with the per-plane basis from scene_prep */
                float4 b0 = g.pln[4 * whop + 2], b1 = g.pln[4 * whop + 3];
                float ui = dot3(xyz(b0), ip) * m.texture_scale;
                float vi = dot3(xyz(b1), ip) * m.texture_scale;
                m.rgb = texel_rgb(fp.tex, fp.tw, fp.th, fp.tl, emod(f2i_x86(ui), fp.tw), emod(f2i_x86(vi), fp.th),
                                  m.texture_id, g.lut255);
            }
            if (AOV && first) hit_id = 1u + g.S + (uint32_t)whop;
        } else {
            V3 c = xyz(g.sph[who]);
            V3 q = ip - c;
            nrm = q / sqrtf(dot3(q, q));
            m = load_material(fp.raw_spheres + (size_t)who * RT_SPHERE_BYTES);
            if (AOV && first) hit_id = 1u + (uint32_t)who;
        }
        first = false;
        ip = ip + nrm * kEps;

        rgb = rgb + (m.rgb * f) * m.ambient; /* raytracing.cl:83-84 */

        /* ---- direct light, raytracing.cl:87-136 ---- */
        {
            V3 vq = o - ip;
            V3 vv = vq / sqrtf(dot3(vq, vq));
            float shin = (float)m.shininess;
            float fs = f * m.specular, fd = f * m.diffuse;
            for (uint32_t i = 0; i < g.L; i++) {
                float4 l0 = g.lgeo[i], l1 = g.laux[i];
                V3 lo = xyz(l0);
                float radius = l1.w;
                V3 ld = lo - ip;
                float dist = sqrtf(dot3(ld, ld));
                V3 sdir = ld / dist;
                float soft = 0.0f;
#pragma unroll 1
                for (int j = 0; j < kShadowSamples; j++) {
                    float theta = (float)(kTwoPi * (double)rng_next(rng));
                    float phi = (float)(kPi * (double)rng_next(rng));
                    float sp, cp, sth, cth;
                    sincosf(phi, &sp, &cp);
                    sincosf(theta, &sth, &cth);
                    float rs = radius * sp;
                    V3 sample = lo + mk(rs * cth, rs * sth, radius * cp);
                    soft += shadow_path_literal(sample, ip, g, tl);
                }
                float ssr = soft / (float)kShadowSamples;
                V3 lrgb = ((xyz(l1) * 1.0f) / (dist * dist)) * ssr;
                V3 hq = vv + sdir;
                V3 hh = hq / sqrtf(dot3(hq, hq));
                float ndh = dot3(nrm, hh);
                float spec = powf(0.0f < ndh ? ndh : 0.0f, shin);
                rgb = rgb + (lrgb * fs) * spec;
                float ndl = dot3(nrm, sdir);
                float diff = 0.0f < ndl ? ndl : 0.0f;
                rgb = rgb + (lrgb * fd) * diff;
            }
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
                p.o = o; p.d = d; p.rgb = rgb; p.depth = depth; p.n = n_cur; p.f = f;
                sz++;
                o = co; d = cd; rgb = mk(0.0f, 0.0f, 0.0f);
                f = old_f * (1.0f - refl);
                n_cur = n2;
                /* depth: the child starts at the parent's incremented depth (raytracing.cl:159,163) */
            }
        }
    }
    return pack_rgb(rgb);
}

}  // namespace RT_NS
