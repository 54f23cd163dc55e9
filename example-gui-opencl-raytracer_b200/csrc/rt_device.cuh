/* rt_device.cuh -- device-side vocabulary shared by the trace variants (included only by rt_kernels.cu).
 *
 * Arithmetic contract (must equal oracle/cl_shim.hpp): binary32, round-to-nearest-even, NO contraction.
 * The translation unit is compiled with -fmad=false -prec-div=true -prec-sqrt=true -ftz=false, so every
 * + - * / sqrtf written below is one correctly rounded operation in source order; FMAs appear only where
 * written explicitly (__fmaf_rn) inside routines whose RESULT is proven/tested to equal the IEEE operation.
 * Reference citations: /root/reference/src/cl/{raygen,raytracing,primitives,types}.cl.
 */
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

#include "../../include/rtb200.h"
#include "rt_kernels.h"

/* rt_kernels.cu is compiled twice: as is (namespace rt, no triangle code at all -- the reference's primitive set,
 * the hot path) and through rt_kernels_tri.cu (namespace rt_tri, RT_TRIANGLES=1: extension X1). */
#ifndef RT_TRIANGLES
#define RT_TRIANGLES 0
#endif
#ifndef RT_ACCEL
#define RT_ACCEL 0 /* 1: rt_kernels_accel.cu -- sphere walks through the N4 cluster structure (rt_trace_accel.cuh) */
#endif
#ifndef RT_NS
#define RT_NS rt
#endif
#ifndef RT_TRI_PREFILTER
#define RT_TRI_PREFILTER 1 /* hit_triangle: decide the common u-test rejects without the division (exact outcome) */
#endif

#define RT_TRI_ACCEL (RT_ACCEL && RT_TRIANGLES) /* rt_kernels_tri_accel.cu: triangle walks through the hierarchy of rt_trace_accel.cuh */

namespace RT_NS {

constexpr float kEps = 0.001f;                     /* primitives.cl:5  EPSILON */
constexpr float kInvPi = 0.318309886183790671538f; /* primitives.cl:6  M_1_PI_F */
constexpr float kGlassPass = 0.8f;                 /* primitives.cl:7  TRANSPERENT_THROUGH */
constexpr float kAirN = 1.0f;                      /* raytracing.cl:7  DEFAULT_N */
constexpr int kMaxDepth = 15;                      /* raytracing.cl:9  MAX_DEPTH */
constexpr int kShadowSamples = 2;                  /* raytracing.cl:10 MAX_SOFT_SHADOWS */
constexpr int kTileW = 8, kTileH = 4;              /* one warp */
constexpr int kBlock = 128;
constexpr double kTwoPi = 6.283185307179586476925286766559; /* 2*M_PI, a double in OpenCL C (raytracing.cl:99) */
constexpr double kPi = 3.14159265358979323846264338327950288;

struct V3 {
    float x, y, z;
};
__device__ __forceinline__ V3 mk(float x, float y, float z) { return V3{x, y, z}; }
__device__ __forceinline__ V3 xyz(float4 v) { return V3{v.x, v.y, v.z}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ V3 operator/(V3 a, float s) { return mk(a.x / s, a.y / s, a.z / s); }
__device__ __forceinline__ float dot3(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
__device__ __forceinline__ V3 cross3(V3 a, V3 b) {
    return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}

/* (int)f with the x86 cvttss2si result for out-of-range / NaN inputs. */
__device__ __forceinline__ int f2i_x86(float f) {
    return (f >= -2147483648.0f && f < 2147483648.0f) ? __float2int_rz(f) : (int)0x80000000;
}
/* primitives.cl:162-168 */
__device__ __forceinline__ int emod(int a, int b) {
    int m = a % b;
    if (m < 0) m = (b < 0) ? m - b : m + b;
    return m;
}
/* clamp(v, 0, 1) as min(max(v, 0), 1) with max(a,b) = a<b ? b : a, min(a,b) = b<a ? b : a (oracle/cl_shim.hpp) */
__device__ __forceinline__ float clamp01(float v) {
    float m = (v < 0.0f) ? 0.0f : v;
    return (1.0f < m) ? 1.0f : m;
}
__device__ __forceinline__ int clampi(int v, int hi) { return v < 0 ? 0 : (v >= hi ? hi - 1 : v); }
/* raytracing.cl:193-194 */
__device__ __forceinline__ uint32_t pack_rgb(V3 rgb) {
    float cr = clamp01(rgb.x) * 255.0f, cg = clamp01(rgb.y) * 255.0f, cb = clamp01(rgb.z) * 255.0f;
    return ((uint32_t)cr << 16) | ((uint32_t)cg << 8) | (uint32_t)cb;
}

/* Work counters: a no-op unless COUNT. */
template <bool COUNT>
struct Tally {
    uint32_t c[RTB_NCOUNTERS];
    __device__ __forceinline__ void clear() {
#pragma unroll
        for (int i = 0; i < RTB_NCOUNTERS; i++) c[i] = 0;
    }
    __device__ __forceinline__ void add(int k, uint32_t n = 1) { c[k] += n; }
    __device__ __forceinline__ uint32_t get(int k) const { return c[k]; }
};
template <>
struct Tally<false> {
    __device__ __forceinline__ void clear() {}
    __device__ __forceinline__ void add(int, uint32_t = 1) {}
    __device__ __forceinline__ uint32_t get(int) const { return 0; }
};

/* Geometry records (rt_kernels.h: RtBlobLayout), in shared memory when resident. */
struct Geo {
    const float4* lgeo;  /* [Lpad] light {o, r*r}; the sphere records follow directly: sph == lgeo + Lpad */
    const float4* sph;   /* [Spad] sphere {c, r*r} */
    const float4* laux;  /* [L] {premultiplied rgb, radius} */
    const float4* pln;   /* [4P] */
    const uint32_t* flag;
    const float* lut255; /* i / 255.0f for i in 0..255 (IEEE division, filled once per CTA) */
    uint32_t S, P, L, Spad, Lpad;
    const float4* tri; /* extension X1: [3T] {v0, e1, e2}, global memory */
    uint32_t T;
    const float4* acc_cl; /* N4 (rt_trace_accel.cuh), global memory */
    const float4* acc_rec;
    const uint32_t* acc_idx;
    const float4* acc_grp;
    uint32_t acc_nc;
    const float4* tacc_node; /* N4 over the triangles (rt_trace_accel.cuh), global memory */
    const float4* tacc_rec;
    const uint32_t* tacc_idx;
    uint32_t tacc_nc, tacc_nl, tacc_ntop;
};
__device__ __forceinline__ void geo_bind(Geo& g, const float4* base, const RtFrameParams& fp, const float* lut) {
    g.S = fp.S; g.P = fp.P; g.L = fp.L;
    g.Spad = fp.lay.Spad; g.Lpad = fp.lay.Lpad;
    g.lgeo = base + fp.lay.geo_off;
    g.sph = g.lgeo + fp.lay.Lpad;
    g.laux = base + fp.lay.laux_off;
    g.pln = base + fp.lay.pln_off;
    g.flag = reinterpret_cast<const uint32_t*>(base + fp.lay.flag_off);
    g.lut255 = lut;
    g.tri = fp.tri;
    g.T = RT_TRIANGLES ? fp.T : 0u; /* a literal 0 in the plain build: every triangle loop and branch folds away */
    g.acc_cl = fp.acc_cl; g.acc_rec = fp.acc_rec; g.acc_idx = fp.acc_idx; g.acc_grp = fp.acc_grp;
    g.acc_nc = RT_ACCEL ? fp.acc_nc : 0u;
    g.tacc_node = fp.tacc_node; g.tacc_rec = fp.tacc_rec; g.tacc_idx = fp.tacc_idx;
    g.tacc_nc = (RT_ACCEL && RT_TRIANGLES) ? fp.tacc_nc : 0u;
    g.tacc_nl = fp.tacc_nl; g.tacc_ntop = fp.tacc_ntop;
}

/* primitives.cl:111-125: ((float)x)/2^31*2 == (float)x * 2^-30 exactly. */
__device__ __forceinline__ uint32_t xorshift_step(uint32_t x) {
    x ^= x << 13;
    x ^= x >> 17;
    x ^= x << 5;
    return x;
}
__device__ __forceinline__ float rng_next(uint32_t& s) {
    s = xorshift_step(s);
    return (float)s * 9.31322574615478515625e-10f;
}

/* primitives.cl:14-109: six independent tests, the last match wins; zero counts as negative. */
__device__ __forceinline__ void cube_uv(V3 d, uint32_t face, int& u, int& v) {
    float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    bool px = d.x > 0.0f, py = d.y > 0.0f, pz = d.z > 0.0f;
    float m = 0.0f, uc = 0.0f, vc = 0.0f;
    uint32_t su = 0, sv = 0;
    if (px && ax >= ay && ax >= az) { m = ax; uc = -d.z; vc = d.y; su = face * 2; sv = face; }
    if (!px && ax >= ay && ax >= az) { m = ax; uc = d.z; vc = d.y; sv = face; }
    if (py && ay >= ax && ay >= az) { m = ay; uc = d.x; vc = -d.z; su = face; sv = face * 2; }
    if (!py && ay >= ax && ay >= az) { m = ay; uc = d.x; vc = d.z; su = face; }
    if (pz && az >= ax && az >= ay) { m = az; uc = d.x; vc = d.y; su = face; sv = face; }
    if (!pz && az >= ax && az >= ay) { m = az; uc = -d.x; vc = d.y; su = face * 3; sv = face; }
    float fu = 0.5f * (uc / m + 1.0f);
    float fv = 0.5f * (vc / m + 1.0f);
    u = f2i_x86((float)su + fu * (float)face);
    v = f2i_x86((float)sv + fv * (float)face);
}

/* read_imagei with clamp-to-edge (Q10) + the manual /255.0f of primitives.cl:252-256, raytracing.cl:72-76 */
__device__ __forceinline__ V3 texel_rgb(const uchar4* base, int w, int h, int layers, int x, int y, int l,
                                        const float* lut255) {
    x = clampi(x, w);
    y = clampi(y, h);
    l = clampi(l, layers);
    uchar4 p = __ldg(base + ((size_t)l * h + y) * w + x);
    return mk(lut255[p.x], lut255[p.y], lut255[p.z]);
}

/* The 64-byte wire material (types.cl:4-19), read in place with four 16-byte loads. */
struct Material {
    V3 rgb;
    float ambient, diffuse, specular;
    uint32_t shininess, transperent, dielectric;
    float n, reflectivity;
    int texture_id;
    float texture_scale;
};
__device__ __forceinline__ Material load_material(const unsigned char* rec) {
    const float4* q = reinterpret_cast<const float4*>(rec + RT_MATERIAL_OFFSET);
    float4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2), d = __ldg(q + 3);
    Material m;
    m.rgb = mk(a.x, a.y, a.z);
    m.ambient = b.x; m.diffuse = b.y; m.specular = b.z; m.shininess = __float_as_uint(b.w);
    m.transperent = __float_as_uint(c.x); m.dielectric = __float_as_uint(c.y); m.n = c.z; m.reflectivity = c.w;
    m.texture_id = __float_as_int(d.x); m.texture_scale = d.y;
    return m;
}

__device__ __forceinline__ Material load_material_at(const unsigned char* mat) {
    const float4* q = reinterpret_cast<const float4*>(mat);
    float4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2), d = __ldg(q + 3);
    Material m;
    m.rgb = mk(a.x, a.y, a.z);
    m.ambient = b.x; m.diffuse = b.y; m.specular = b.z; m.shininess = __float_as_uint(b.w);
    m.transperent = __float_as_uint(c.x); m.dielectric = __float_as_uint(c.y); m.n = c.z; m.reflectivity = c.w;
    m.texture_id = __float_as_int(d.x); m.texture_scale = d.y;
    return m;
}

/* EXTENSION X1 (no reference counterpart; oracle/rt_oracle.c:hit_triangle is the definition): Moeller-Trumbore,
 * two-sided, one rounded operation per step in exactly that order. */
/* record loads: through the read-only path from global memory, or plain loads when the records have been staged into
 * shared memory (rt_trace_tree.cuh: tri_walk_staged) */
template <bool STAGED>
__device__ __forceinline__ float4 ld_rec(const float4* p) {
    if (STAGED) return *p;
    return __ldg(p);
}
template <bool COUNT, bool STAGED = false>
__device__ __forceinline__ bool hit_triangle(V3 o, V3 d, const float4* rec, float& t, Tally<COUNT>& tl) {
    tl.add(RTB_TRI_TESTS);
    const V3 v0 = xyz(ld_rec<STAGED>(rec)), e1 = xyz(ld_rec<STAGED>(rec + 1)), e2 = xyz(ld_rec<STAGED>(rec + 2));
    V3 p = cross3(d, e2);
    float det = dot3(e1, p);
    if (det == 0.0f) return false;
    V3 s = o - v0;
    const float a = dot3(s, p);
#if RT_TRI_PREFILTER
    /* Exact-outcome shortcut that keeps the division off the common reject path (most triangles fail the u test):
     * opposite signs and |a| > |det|*1e-30  =>  u = RN(a*RN(1/det)) is a normal negative number (no underflow to -0);
     * equal signs and |a| > |det|*1.0001     =>  u > 1 whatever the two roundings (each <= 2^-22 relative).
     * NaN / inf / zero operands make the comparisons false and fall through to the defining sequence below
     * (tests/test_extensions_cpu.py::test_triangle_prefilter_is_exact checks the claim in float32). */
    {
        const float ad = fabsf(det), aa = fabsf(a);
        if (((a < 0.0f) != (det < 0.0f)) ? (aa > ad * 1e-30f) : (aa > ad * 1.0001f)) return false;
    }
#endif
    float inv = 1.0f / det;
    float u = a * inv;
    if (u < 0.0f || u > 1.0f) return false;
    V3 q = cross3(s, e1);
    float v = dot3(d, q) * inv;
    if (v < 0.0f || u + v > 1.0f) return false;
    float t2 = dot3(e2, q) * inv;
    if (t2 <= 0.0f) return false;
    t = t2;
    return true;
}
/* The triangle walk of the fast traces: RT_TRI_BATCH records per step, their 3*RT_TRI_BATCH loads issued together and
 * the cheap half of every test (det, the u numerator, the exact-outcome prefilter) run side by side, so the L1/L2
 * latency of a record is paid once per batch instead of once per triangle; survivors finish one by one IN INDEX
 * ORDER (same operations as hit_triangle, same visiting order => same closest hit under "first of equal t wins").
 * on_hit(k, t) returns true to stop the walk (shadow rays); the function then returns true. */
#ifndef RT_TRI_BATCH
#define RT_TRI_BATCH 4
#endif
template <bool STAGED, bool COUNT, class F>
__device__ __forceinline__ bool for_triangle_hits_ld(V3 o, V3 d, const float4* __restrict__ tri, uint32_t T, Tally<COUNT>& tl, F&& on_hit) {
    constexpr uint32_t kB = RT_TRI_BATCH;
    uint32_t k = 0;
#pragma unroll 1
    for (; k + kB <= T; k += kB) {
        tl.add(RTB_TRI_TESTS, kB);
        float4 rv0[kB], re1[kB], re2[kB];
#pragma unroll
        for (uint32_t j = 0; j < kB; j++) {
            rv0[j] = ld_rec<STAGED>(tri + 3u * (k + j));
            re1[j] = ld_rec<STAGED>(tri + 3u * (k + j) + 1u);
            re2[j] = ld_rec<STAGED>(tri + 3u * (k + j) + 2u);
        }
        float det[kB], a[kB];
        uint32_t cand = 0;
#pragma unroll
        for (uint32_t j = 0; j < kB; j++) {
            const V3 p = cross3(d, xyz(re2[j]));
            det[j] = dot3(xyz(re1[j]), p);
            const V3 s = o - xyz(rv0[j]);
            a[j] = dot3(s, p);
#if RT_TRI_PREFILTER
            /* det == 0 is left to the survivors' pass: with det == 0 the prefilter rejects every a != 0 by itself */
            const float ad = fabsf(det[j]), aa = fabsf(a[j]);
            const bool live = !(((a[j] < 0.0f) != (det[j] < 0.0f)) ? (aa > ad * 1e-30f) : (aa > ad * 1.0001f));
#else
            const bool live = det[j] != 0.0f;
#endif
            if (live) cand |= 1u << j;
        }
        if (cand == 0u) continue;
#pragma unroll
        for (uint32_t j = 0; j < kB; j++) {
            if (!(cand >> j & 1u) || det[j] == 0.0f) continue;
            const float inv = 1.0f / det[j];
            const float u = a[j] * inv;
            if (u < 0.0f || u > 1.0f) continue;
            const V3 s = o - xyz(rv0[j]);
            const V3 q = cross3(s, xyz(re1[j]));
            const float v = dot3(d, q) * inv;
            if (v < 0.0f || u + v > 1.0f) continue;
            const float t2 = dot3(xyz(re2[j]), q) * inv;
            if (t2 <= 0.0f) continue;
            if (on_hit(k + j, t2)) return true;
        }
    }
    for (; k < T; k++) {
        float t1;
        if (!hit_triangle<COUNT, STAGED>(o, d, tri + 3u * k, t1, tl)) continue;
        if (on_hit(k, t1)) return true;
    }
    return false;
}
template <bool COUNT, class F>
__device__ __forceinline__ bool for_triangle_hits(V3 o, V3 d, const float4* __restrict__ tri, uint32_t T, Tally<COUNT>& tl, F&& on_hit) {
    return for_triangle_hits_ld<false>(o, d, tri, T, tl, on_hit);
}
__device__ __forceinline__ V3 triangle_normal(const float4* rec) {
    V3 n = cross3(xyz(__ldg(rec + 1)), xyz(__ldg(rec + 2)));
    return n / sqrtf(dot3(n, n));
}

/* One suspended stack entry (raytracing.cl:28-30 keeps rray + n + f in three arrays). */
struct Saved {
    V3 o, d, rgb;
    int depth;
    float n, f;
};

/* raygen.cl:13-24 */
__device__ __forceinline__ void primary_ray(const RtFrameParams& fp, uint32_t id, V3& o, V3& d) {
    float w = (float)(id % fp.W);
    float h = (float)(id / fp.W);
    V3 corner = mk(fp.corner[0], fp.corner[1], fp.corner[2]);
    V3 up = mk(fp.up[0], fp.up[1], fp.up[2]);
    V3 right = mk(fp.right[0], fp.right[1], fp.right[2]);
    V3 vec = (corner + (right * fp.wf) * w) - (up * fp.hf) * h;
    d = vec / sqrtf(dot3(vec, vec));
    o = mk(fp.origin[0], fp.origin[1], fp.origin[2]);
}

/* Schlick reflectance of primitives.cl:146-160 folded into raytracing.cl:146-150; cosI = -dot(N, I). */
__device__ __forceinline__ float reflect_amount(const Material& m, float n1, float n2, float cosI) {
    float refl = m.reflectivity;
    if (m.dielectric) {
        const float rn = n1 - n2, rd = n1 + n2;
        /* 0 / (positive finite) is 0: spares the division's slow path for index-matched media (mirror, n = 1) */
        float r0 = (rn == 0.0f && rd > 1e-30f && rd < 1e30f) ? rn : rn / rd;
        r0 *= r0;
        float cosX = cosI;
        float fr;
        bool tir = false;
        if (n1 > n2) {
            float nn = n1 / n2;
            float sinT2 = nn * nn * (1.0f - cosX * cosX);
            if (sinT2 > 1.0f) tir = true;
            else cosX = sqrtf(1.0f - sinT2);
        }
        if (tir) fr = 1.0f;
        else {
            float x = 1.0f - cosX;
            fr = r0 + (1.0f - r0) * x * x * x * x * x;
        }
        refl = m.reflectivity + (1.0f - m.reflectivity) * fr;
    }
    return refl;
}

/* The refracted child of raytracing.cl:161-179 + refract() of primitives.cl:132-144.
 * Returns false when the child is dropped (total internal reflection or NaN direction). */
__device__ __forceinline__ bool refract_child(V3 o, V3 nrm, V3 incident, float n1, float n2, V3& co, V3& cd) {
    co = o;
    V3 nn = nrm;
    if (n1 < n2) co = o - nrm * (2.0f * kEps);
    else nn = mk(-nrm.x, -nrm.y, -nrm.z);
    float eta = n1 / n2;
    float cI = -dot3(nn, incident);
    float sinT2 = eta * eta * (1.0f - cI * cI);
    if (sinT2 > 1.0f) return false;
    float cosT = sqrtf(1.0f - sinT2);
    cd = incident * eta + nn * (eta * cI - cosT);
    return !(cd.x != cd.x);
}

/* ---- TMA bulk copy + mbarrier (PTX; SASS: UBLKCP / SYNCS) ---- */
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE;\n"
        "bra WAIT_LOOP;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

}  // namespace RT_NS
