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

#if RT_TRIANGLES
/* ---- N4 over the triangles of extension X1 -------------------------------------------------------------------------
 * The same idea one level deeper, with boxes: triangles in Morton order of their centroids, clusters of kAccK, then 8
 * nodes per parent up to RT_TACC_LEVELS levels (rt_engine.cu: build_tri_accel_host); every node is the axis-aligned box
 * {centre, half extents} of the vertices below it (flat patches of a mesh fill a box far better than a sphere).  A walk
 * runs over cluster positions in order; at a position that starts a level-l node it tests that node first and skips
 * its 8^l clusters when the ray provably cannot produce a hit there; clusters that survive run the UNCHANGED
 * Moeller-Trumbore walk (for_triangle_hits) on their members.
 *
 * The cull (tacc_cull): the box grown on every side by mu = (|o - c|_1 + |h|_1) / 4096 >= (distance from the origin to
 * any point of the box) * 2^-12, and the slab test against it, as a LINE first (entry after exit: miss), then "exit
 * behind the origin" and "entry beyond what matters".  (accel_cull's margin is 2^-8: a sphere test's discriminant
 * cancels, so its noise goes with sqrt(2^-24); Moeller-Trumbore's goes with 2^-24 itself, and a margin relative to
 * the origin's distance is what decides how well far-away origins -- shadow rays from the ground plane near the
 * horizon -- can still be culled: at 2^-8 those walks degenerated into the full brute-force loop.)  IEEE infinities do the parallel case
 * (1 / 0 = inf: a slab the origin is inside gives (-inf, +inf), one it is outside gives an empty interval); every
 * comparison is written so that a NaN anywhere means "do not cull".  The slab arithmetic itself rounds at 2^-22 of
 * the distances involved, 2^14 times below mu.
 *
 * Equality with the brute-force walk (closest_pass / shadow_path_fast):
 *  - closest hit: the brute-force rule over triangles in index order is `!(t1 >= c.t)` -- strictly closer than
 *    everything before, spheres and planes included -- so among triangles of equal distance the smallest index wins and
 *    a triangle never displaces an equally distant sphere or plane.  Here candidates arrive in cluster order, so the
 *    rule is spelled out: closer, or equally close as the TRIANGLE held so far with a smaller original index.  A NaN
 *    distance (in c.t on entry or from a member) makes the brute-force outcome order-dependent: such a ray is
 *    re-walked in index order.  The light-occlusion flag is an OR.
 *  - shadow ray: any member with !(t1 >= len) blocks -- an OR.
 *  - what a cull may skip: the margin argument of accel_cull carries over as long as the member test's rounding
 *    error, seen as a displacement in the triangle's plane, stays below the margin:
 *    ~ 8 * 2^-24 * |o - v0| / (sin(angle e1,e2) * |cos(angle d,normal)|) < |o - v0| / 4096, i.e. the ray is not within
 *    ~2^-9 (0.1 degree) of grazing the triangle's plane (times the triangle's shape factor).  Inside that sliver of
 *    directions the brute-force walk's own answer is rounding noise of a det near zero -- and it accepts a far-away
 *    triangle only when the ray also lies within ~2^-21 |o - v0| of that triangle's plane -- so a cull can in
 *    principle skip a triangle the noise would have accepted; the parity statement for this extension of an extension is therefore
 *    "equal wherever compared" (tests/test_accel_gpu.py: meshes, triangle soups, duplicates, the 49,928-triangle mesh
 *    at 1080p; bench.py compares the 4K frame), not a proof.  X1 itself has no reference counterpart (parity
 *    unpinned, SURVEY.md 8(a') X1). */
struct TaccRay {
    V3 o, inv, ainv; /* origin, 1 / direction per axis (IEEE: +-inf for a zero component), |1 / direction| */
};
__device__ __forceinline__ TaccRay tacc_ray(V3 o, V3 d) {
    TaccRay q;
    q.o = o;
    q.inv = mk(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    q.ainv = mk(fabsf(q.inv.x), fabsf(q.inv.y), fabsf(q.inv.z));
    return q;
}
/* false: a NaN or an infinity among the six numbers (NaN test results then depend on the visiting order) */
__device__ __forceinline__ bool tacc_finite(const RayF& r) {
    const float chk = ((fabsf(r.o.x) + fabsf(r.o.y)) + fabsf(r.o.z)) + ((fabsf(r.d.x) + fabsf(r.d.y)) + fabsf(r.d.z));
    return chk <= 3.0e38f;
}
/* true when no triangle inside the box {c, h} can yield a hit with 0 < t <= tmax (tmax = +inf: any t).  Per axis the
 * grown slab is crossed during [tc - te, tc + te], tc = (c - o) / d, te = (h + mu) / |d|.  The arithmetic here is the
 * cull's own (explicit FMAs are fine: nothing downstream depends on its bits, only on its margin). */
__device__ __forceinline__ bool tacc_cull(const TaccRay& q, float4 c, float4 h, float tmax) {
    const float ax = c.x - q.o.x, ay = c.y - q.o.y, az = c.z - q.o.z;
    const float mu = __fmaf_rn((fabsf(ax) + fabsf(ay)) + fabsf(az), 0.000244140625f, h.w); /* 2^-12 |o - c|_1 + 2^-12 |h|_1 (h.w, rounded up on the host) */
    const float cx = ax * q.inv.x, cy = ay * q.inv.y, cz = az * q.inv.z;
    const float ex = (h.x + mu) * q.ainv.x, ey = (h.y + mu) * q.ainv.y, ez = (h.z + mu) * q.ainv.z;
    const float tn = fmaxf(fmaxf(cx - ex, cy - ey), cz - ez);
    const float tf = fminf(fminf(cx + ex, cy + ey), cz + ez);
    /* a direction component of exactly 0: inv = inf, so tc = +-inf (NaN when the origin sits on the centre plane) and
     * te = inf; inf - inf = NaN drops out of fminf / fmaxf, which leaves that axis unconstrained -- right when the origin
     * is inside the slab, merely a missed cull when it is outside.  Rays with a non-finite origin or direction never get
     * here (the callers walk those by brute force); every comparison below is false on NaN: "do not cull". */
    return tn > tf || tf < 0.0f || tn > tmax;
}
/* node (level l, i): levels are stored top first, each padded to ntop * 8^(top - l) slots, so level l starts at
 * ntop * (8^(top - l) - 1) / 7 -- and (8^m - 1) / 7 is the bit pattern 001001...001 of m ones */
__device__ __forceinline__ uint32_t tacc_level_base(uint32_t ntop, uint32_t m) { return ntop * (0x09249249u & ((1u << (3u * m)) - 1u)); }
template <class TMAX, class LEAF>
__device__ __forceinline__ void tacc_walk(const TaccRay& q, const Geo& g, TMAX&& tmax, LEAF&& leaf) {
    const uint32_t top = g.tacc_nl - 1u, nc = g.tacc_nc, ntop = g.tacc_ntop;
    uint32_t k = 0, l = top;
    for (;;) {
        /* search: one node per step -- the level-l ancestor that starts at cluster position k, then down, or past it */
        bool have = false;
#pragma unroll 1
        while (k < nc) {
            const float4* nd = g.tacc_node + 2u * (tacc_level_base(ntop, top - l) + (k >> (3u * l)));
            if (tacc_cull(q, __ldg(nd), __ldg(nd + 1), tmax())) {
                k += 1u << (3u * l);
                const uint32_t up = (uint32_t)(__ffs((int)k) - 1) / 3u; /* k != 0 here */
                l = up < top ? up : top;
            } else if (l != 0u) {
                l--;
            } else {
                have = true;
                break;
            }
        }
        if (!have) return;
        if (leaf(k)) return;
        k++;
        const uint32_t up = (uint32_t)(__ffs((int)k) - 1) / 3u;
        l = up < top ? up : top;
    }
}

/* the triangle part of closest_pass (rt_trace_fast.cuh) through the hierarchy; true = re-walk this ray in index order */
template <bool COUNT>
__device__ __forceinline__ bool closest_triangles_accel(const RayF& r, const Geo& g, Tally<COUNT>& tl, Closest& c) {
    if (c.t != c.t || c.tL != c.tL || !tacc_finite(r)) return true;
    const TaccRay q = tacc_ray(r.o, r.d);
    bool odd = false;
    tacc_walk(q, g, [&]() { return ((c.k >= 0 && !c.occluded && c.tL > c.t) ? c.tL : c.t) * 1.00001f; },
              [&](uint32_t k) {
                  for_triangle_hits(r.o, r.d, g.tacc_rec + 3u * kAccK * k, kAccK, tl, [&](uint32_t j, float t1) {
                      const int orig = (int)__ldg(g.tacc_idx + kAccK * k + j);
                      if (t1 != t1) odd = true;
                      if (t1 < c.t || (t1 == c.t && c.whot >= 0 && orig < c.whot)) { c.t = t1; c.whot = orig; }
                      if (t1 <= c.tL) c.occluded = true;
                      return false;
                  });
                  return false;
              });
    return odd;
}

/* the triangle part of a shadow ray: true = blocked */
template <bool COUNT>
__device__ __forceinline__ bool shadow_triangles_accel(const RayF& r, float len, const Geo& g, Tally<COUNT>& tl) {
    if (!tacc_finite(r)) return for_triangle_hits(r.o, r.d, g.tri, g.T, tl, [&](uint32_t, float t1) { return !(t1 >= len); });
    const TaccRay q = tacc_ray(r.o, r.d);
    const float tm = len * 1.00001f;
    bool blocked = false;
    tacc_walk(q, g, [&]() { return tm; }, [&](uint32_t k) {
        blocked = for_triangle_hits(r.o, r.d, g.tacc_rec + 3u * kAccK * k, kAccK, tl, [&](uint32_t, float t1) { return !(t1 >= len); });
        return blocked;
    });
    return blocked;
}
#endif

}  // namespace RT_NS
