/* rt_trace_accel.cuh -- SURVEY.md 8(f) N4: an acceleration structure for many-sphere scenes that returns EXACTLY what
 * the brute-force walks of primitives.cl:322-394 (closest hit) and :396-442 (shadow path) return.
 *
 * Not in the reference (its loops are O(S + P) per ray by construction); reported on its own bench line, never as the
 * headline.  Structure: the spheres in Morton order of their centres, cut into clusters of kAccK; per cluster a bounding
 * sphere {C, R} (built on the host in fp64 and inflated, rt_engine.cu: prep_accel).  A walk visits every cluster, skips
 * those the ray provably cannot hit, and runs the UNCHANGED sphere test (sphere_disc / sphere_root, four records at a
 * time) on the members of the others.
 *
 * Why the result is the brute-force one, word for word:
 *  - a member test is skipped only when its computed discriminant is certainly negative, its computed root certainly
 *    <= 0, or its computed root certainly beyond the distance that matters (the closest hit so far / the shadow ray's
 *    length).  The cull compares exact-geometry distances with margins far above the rounding error of the member
 *    test: the line must miss the inflated bounding sphere by |v| / 256 (|v| = distance origin -> centre), while a
 *    computed D >= 0 needs the line within sqrt(k * 2^-24) * |v| ~ |v| / 1300 of the sphere (D = b*b - 4a*c, each term
 *    rounded to 2^-24 relative, |b*b|, |4ac| <= 4a|v|^2).  The (x+y)^2 <= 1.125 x^2 + 9 y^2 split keeps the test free
 *    of square roots; its own rounding (2^-23 of a*|v|^2) sits inside the 9/65536 * |v|^2 term.
 *  - closest hit: candidates arrive in cluster order, not index order, so "first of equal t wins" (strict `_t >= t`
 *    reject in index order) becomes "smaller t, then smaller ORIGINAL index"; a NaN distance would make the reference's
 *    `!(t1 >= t)` update order-dependent, so such a ray is re-walked by the brute-force loop (as the WIDE variant does);
 *    the light-occlusion flag is an OR.
 *  - shadow path: 0 if any opaque member with t1 < len, else kGlassPass multiplied in once per transparent member hit
 *    -- equal factors, so their order cannot change the product.
 * tests/test_accel_gpu.py: word-identical to the brute-force kernel and to the reference text on sphere grids (225 ..
 * 6,561), the full-size 1,024-sphere golden, and the 48 fuzz scenes (duplicates = equal-t ties, NaN rays, zero radii).
 */
#pragma once
#include "rt_device.cuh"

namespace RT_NS {

constexpr uint32_t kAccK = RT_ACCEL_K; /* spheres per cluster (rt_kernels.h) */
constexpr uint32_t kAccG = RT_ACCEL_G; /* clusters per group: a second level of bounding spheres over consecutive clusters */

/* {C - o, slack-free}: true when no member of the cluster can yield a hit with 0 < t <= tmax (tmax = +inf: any t). */
__device__ __forceinline__ bool accel_cull(const RayF& r, float a, float4 cl, float tmax) {
    const float vx = cl.x - r.o.x, vy = cl.y - r.o.y, vz = cl.z - r.o.z;
    const float L2 = (vx * vx + vy * vy) + vz * vz;
    const float s = (vx * r.d.x + vy * r.d.y) + vz * r.d.z;       /* |d| * (signed distance of C along the ray) */
    const float lim = a * (cl.w + 1.3733e-4f * L2);               /* a * (1.125 R'^2 + 9 |v|^2 / 65536) >= a * (R' + |v| / 256)^2 */
    const float q = L2 * a - s * s;                               /* a * (distance of C from the line)^2 */
    if (q > lim) return true;                                     /* the line misses the inflated bounding sphere */
    const float s2 = s * s;
    if (s < 0.0f && s2 > lim && L2 > cl.w) return true;           /* wholly behind the origin: every root is negative */
    const float e = s - tmax * a;                                 /* tmax = +inf gives -inf (or NaN for a = 0): no cull */
    if (e > 0.0f && e * e > lim) return true;                     /* wholly beyond tmax */
    return false;
}

/* findLightIntersection + findSolidIntersection (R7, rt_trace_fast.cuh: closest_pass) over the clusters. */
template <bool COUNT>
__device__ __forceinline__ bool closest_spheres_accel(const RayF& r, const Geo& g, Tally<COUNT>& tl, Closest& c) {
    const float a = (r.d.x * r.d.x + r.d.y * r.d.y) + r.d.z * r.d.z;
    bool odd = false;
#pragma unroll 1
    for (uint32_t k = 0; k < g.acc_nc; k++) {
        /* what still matters: (computed again per cluster: c.t shrinks as hits come in) */
        if ((k % kAccG) == 0u) { /* second level: the bounding sphere of the next kAccG clusters */
            const float tm = (c.k >= 0 && c.tL > c.t) ? c.tL : c.t;
            if (accel_cull(r, a, __ldg(g.acc_grp + k / kAccG), tm * 1.00001f)) { k += kAccG - 1u; continue; }
        }
        const float4 cl = __ldg(g.acc_cl + k);
        /* what still matters: hits up to the closest so far (one at exactly c.t with a smaller index must still be seen)
         * and, when the ray has hit a light, every opaque sphere up to that light (its occlusion test) */
        const float tmax = (c.k >= 0 && c.tL > c.t) ? c.tL : c.t;
        if (accel_cull(r, a, cl, tmax * 1.00001f)) continue;
        for_sphere_hits(r, g.acc_rec, k * kAccK, (k + 1u) * kAccK, tl, [&](uint32_t j, float t1) {
            const uint32_t w = __ldg(g.acc_idx + j);
            if (w == 0xffffffffu) return false;                   /* padding */
            const int orig = (int)(w & 0x7fffffffu);
            if (t1 != t1) odd = true;
            if (t1 < c.t || (t1 == c.t && orig < c.who)) { c.t = t1; c.who = orig; }
            if (t1 <= c.tL && !(w >> 31)) c.occluded = true;       /* bit 31: material.transperent */
            return false;
        });
    }
    return odd;
}

/* testShadowPath's sphere part (rt_trace_fast.cuh: shadow_path_fast) over the clusters: false = blocked. */
template <bool COUNT>
__device__ __forceinline__ bool shadow_spheres_accel(const RayF& r, float len, const Geo& g, Tally<COUNT>& tl, float& opacity) {
    const float a = (r.d.x * r.d.x + r.d.y * r.d.y) + r.d.z * r.d.z;
    uint32_t nglass = 0;
    bool blocked = false;
#pragma unroll 1
    for (uint32_t k = 0; k < g.acc_nc && !blocked; k++) {
        if ((k % kAccG) == 0u && accel_cull(r, a, __ldg(g.acc_grp + k / kAccG), len * 1.00001f)) { k += kAccG - 1u; continue; }
        const float4 cl = __ldg(g.acc_cl + k);
        if (accel_cull(r, a, cl, len * 1.00001f)) continue;
        blocked = for_sphere_hits(r, g.acc_rec, k * kAccK, (k + 1u) * kAccK, tl, [&](uint32_t j, float t1) {
            const uint32_t w = __ldg(g.acc_idx + j);
            if (w == 0xffffffffu || t1 >= len) return false;
            if (w >> 31) { nglass++; return false; }
            return true;
        });
    }
    if (blocked) return false;
    for (uint32_t n = 0; n < nglass; n++) opacity *= kGlassPass;   /* primitives.cl:396-442: one factor per glass hit */
    return true;
}

}  // namespace RT_NS
