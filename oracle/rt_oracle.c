/* oracle/rt_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE (see rt_oracle.h).
 *
 * Plain-C restatement of the reference hot path.  "ref:" comments cite /root/reference files.
 * Arithmetic contract (same as cl_shim.hpp): binary32, round-to-nearest-even, NO contraction
 * (build with -ffp-contract=off), every intermediate rounded in source order.
 */
#define _GNU_SOURCE
#include "rt_oracle.h"

#include <limits.h>
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ---------- wire layouts, ref: src/cl/types.cl:4-57 == src/cpu_obj.h:10-48 ---------- */
typedef struct { float x, y, z; } v3;

typedef struct {
    float rgb[4];                 /* float3 occupies 16 B */
    float ambient, diffuse, specular;
    uint32_t shininess, transperent, dielectric;
    float n, reflectivity;
    int32_t texture_id;
    float texture_scale;
    uint32_t tail_pad[2];
} w_material;                     /* 64 B */
typedef struct { float origin[4]; float radius; uint32_t pad[3]; w_material m; } w_sphere; /* 96 B */
typedef struct { float normal[4]; float point[4]; w_material m; } w_plane;                 /* 96 B */
typedef struct { float origin[4]; float radius, intensity; uint32_t pad[2]; float rgb[4]; } w_light; /* 48 B */

/* EXTENSION X1 (SURVEY.md 8(a'), no counterpart in the reference): a triangle record, 112 B. */
typedef struct { float v0[4], v1[4], v2[4]; w_material m; } w_triangle;
_Static_assert(sizeof(w_triangle) == 112, "rtriangle");

_Static_assert(sizeof(w_material) == 64, "rmaterial");
_Static_assert(sizeof(w_sphere) == 96, "rsphere");
_Static_assert(sizeof(w_plane) == 96, "rplane");
_Static_assert(sizeof(w_light) == 48, "rlight");

#define EPS_OFFSET 0.001f          /* ref: primitives.cl:5  EPSILON */
#define INV_PI_F 0.318309886183790671538f /* ref: primitives.cl:6  M_1_PI_F */
#define GLASS_PASS 0.8f            /* ref: primitives.cl:7  TRANSPERENT_THROUGH */
#define AIR_N 1.0f                 /* ref: raytracing.cl:7  DEFAULT_N */
#define STACK_MAX 15               /* ref: raytracing.cl:9  MAX_DEPTH */
#define SHADOW_SAMPLES 2           /* ref: raytracing.cl:10 MAX_SOFT_SHADOWS */
#define HIT_LIGHT_FLAG 0x80000000u

typedef struct {
    const w_sphere* sph; uint32_t S;
    const w_plane* pln; uint32_t P;
    const w_light* lit; uint32_t L;
    const uint8_t* tex; int tw, th, tl;
    const uint8_t* sky; int sw, sh;
    const w_triangle* tri; uint32_t T; /* extension X1 */
} scene_t;

/* Extension X1: triangles for the NEXT rto_render* calls (process-global so the reference-shaped signatures stay). */
static const w_triangle* g_tri = 0;
static uint32_t g_T = 0;

/* ---------- vector helpers: each component op is one rounded binary32 operation ---------- */
static inline v3 mk(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static inline v3 ld3(const float* p) { return mk(p[0], p[1], p[2]); }
static inline v3 add(v3 a, v3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 sub(v3 a, v3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 scl(v3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
static inline v3 dvs(v3 a, float s) { return mk(a.x / s, a.y / s, a.z / s); }
static inline float dot3(v3 a, v3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
static inline v3 cross3(v3 a, v3 b) {
    return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
static inline v3 unit(v3 a) { return dvs(a, sqrtf(dot3(a, a))); }            /* normalize */
static inline float dist3(v3 a, v3 b) { v3 d = sub(a, b); return sqrtf(dot3(d, d)); } /* distance */

/* (int)f as x86 cvttss2si does it: out-of-range and NaN give INT_MIN. */
static inline int f2i(float f) {
    if (!(f >= -2147483648.0f && f < 2147483648.0f)) return INT_MIN;
    return (int)f;
}

/* per-thread work counters */
typedef struct { uint64_t c[RTO_NCOUNTERS]; } counters_t;
static __thread counters_t* tl_cnt = 0;
static __thread uint32_t tl_pixel_segments = 0; /* segments of the pixel being traced (cost map) */
#define COUNT(k) do { if (tl_cnt) tl_cnt->c[k]++; } while (0)

/* clamp-to-edge texel fetch; ref: read_imagei at primitives.cl:250 / raytracing.cl:70 (Q10) */
static inline v3 texel(const uint8_t* base, int w, int h, int layers, int x, int y, int l) {
    COUNT(RTO_TEXELS);
    x = x < 0 ? 0 : (x >= w ? w - 1 : x);
    y = y < 0 ? 0 : (y >= h ? h - 1 : y);
    l = l < 0 ? 0 : (l >= layers ? layers - 1 : l);
    const uint8_t* p = base + (((size_t)l * h + y) * w + x) * 4;
    return mk((float)p[0] / 255.0f, (float)p[1] / 255.0f, (float)p[2] / 255.0f);
}

/* ref: primitives.cl:116-125 */
static inline float rng_next(uint32_t* s) {
    uint32_t x = *s;
    x ^= x << 13; x ^= x >> 17; x ^= x << 5;
    *s = x;
    return ((float)x) / 2147483648.0f * 2.0f;
}

/* ref: primitives.cl:14-109.  Six independent tests, last match wins (Q11); zero counts as negative (Q12). */
static void cube_uv(v3 d, uint32_t face, int* u_out, int* v_out) {
    float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    int px = d.x > 0, py = d.y > 0, pz = d.z > 0;
    float m = 0.0f, uc = 0.0f, vc = 0.0f;
    uint32_t su = 0, sv = 0;
    if (px && ax >= ay && ax >= az)  { m = ax; uc = -d.z; vc = d.y;  su = face * 2; sv = face; }
    if (!px && ax >= ay && ax >= az) { m = ax; uc = d.z;  vc = d.y;  sv = face; }
    if (py && ay >= ax && ay >= az)  { m = ay; uc = d.x;  vc = -d.z; su = face; sv = face * 2; }
    if (!py && ay >= ax && ay >= az) { m = ay; uc = d.x;  vc = d.z;  su = face; }
    if (pz && az >= ax && az >= ay)  { m = az; uc = d.x;  vc = d.y;  su = face; sv = face; }
    if (!pz && az >= ax && az >= ay) { m = az; uc = -d.x; vc = d.y;  su = face * 3; sv = face; }
    float fu = 0.5f * (uc / m + 1.0f);
    float fv = 0.5f * (vc / m + 1.0f);
    *u_out = f2i((float)su + fu * (float)face);
    *v_out = f2i((float)sv + fv * (float)face);
}

/* ref: primitives.cl:127-130 */
static inline v3 mirror_dir(v3 I, v3 N) {
    float cosI = -dot3(N, I);
    return add(I, scl(N, 2.0f * cosI));
}

/* ref: primitives.cl:132-144; returns 0 on total internal reflection (the NaN vector there) */
static inline int bend_dir(float n1, float n2, v3 I, v3 N, v3* out) {
    float n = n1 / n2;
    float cosI = -dot3(N, I);
    float sinT2 = n * n * (1.0f - cosI * cosI);
    if (sinT2 > 1.0f) return 0;
    float cosT = sqrtf(1.0f - sinT2);
    *out = add(scl(I, n), scl(N, n * cosI - cosT));
    return 1;
}

/* ref: primitives.cl:146-160 */
static inline float fresnel_schlick(float n1, float n2, v3 I, v3 N) {
    float r0 = (n1 - n2) / (n1 + n2);
    r0 *= r0;
    float cosX = -dot3(N, I);
    if (n1 > n2) {
        float n = n1 / n2;
        float sinT2 = n * n * (1.0f - cosX * cosX);
        if (sinT2 > 1.0f) return 1.0f;
        cosX = sqrtf(1.0f - sinT2);
    }
    float x = 1.0f - cosX;
    return r0 + (1.0f - r0) * x * x * x * x * x;
}

/* ref: primitives.cl:162-168 */
static inline int emod(int a, int b) {
    int m = a % b;
    if (m < 0) m = (b < 0) ? m - b : m + b;
    return m;
}

/* ref: primitives.cl:170-195 */
static inline int hit_sphere(v3 o, v3 d, v3 c, float r, float* t) {
    COUNT(RTO_SPHERE_TESTS);
    v3 v = sub(o, c);
    float a = dot3(d, d);
    float b = dot3(scl(v, 2.0f), d);
    float cc = dot3(v, v) - r * r;
    float D = b * b - 4.0f * a * cc;
    if (D < 0) return 0;
    COUNT(RTO_SPHERE_DPOS);
    D = sqrtf(D);
    float near = (-b - D) / (2.0f * a);
    float t2 = (near < 0) ? (-b + D) / (2.0f * a) : near;
    if (t2 <= 0) return 0;
    *t = t2;
    return 1;
}

/* ref: primitives.cl:197-215 */
static inline int hit_plane(v3 o, v3 d, v3 n, v3 p, float* t) {
    COUNT(RTO_PLANE_TESTS);
    float b = dot3(d, n);
    if (b == 0) return 0;
    float t2 = dot3(sub(p, o), n) / b;
    if (t2 <= 0) return 0;
    *t = t2;
    return 1;
}

/* EXTENSION X1 -- no reference code: Moeller-Trumbore, two-sided, every step one rounded binary32 operation in
 * this order.  The CUDA kernels restate exactly this sequence.  Parity for scenes with triangles is therefore
 * SELF-CONSISTENCY ONLY ("parity unpinned"). */
static inline int hit_triangle(v3 o, v3 d, v3 v0, v3 e1, v3 e2, float* t) {
    COUNT(RTO_TRI_TESTS);
    v3 p = cross3(d, e2);
    float det = dot3(e1, p);
    if (det == 0) return 0;
    float inv = 1.0f / det;
    v3 s = sub(o, v0);
    float u = dot3(s, p) * inv;
    if (u < 0 || u > 1) return 0;
    v3 q = cross3(s, e1);
    float v = dot3(d, q) * inv;
    if (v < 0 || u + v > 1) return 0;
    float t2 = dot3(e2, q) * inv;
    if (t2 <= 0) return 0;
    *t = t2;
    return 1;
}
#define TRI_EDGES(tr) v3 v0_ = ld3((tr)->v0), e1_ = sub(ld3((tr)->v1), v0_), e2_ = sub(ld3((tr)->v2), v0_)

/* ref: primitives.cl:217-259.  Tangent basis = first unit axis e with sum(cross(e, n)) != 0 (Q16);
 * if none qualifies the reference reads an uninitialised basis -- defined here (and in cl_shim.hpp,
 * whose float3 default-constructs to zero) as the zero vectors. */
static v3 plane_texel(const w_plane* pl, v3 P, const scene_t* sc) {
    const v3 axes[3] = {{1.0f, 0.0f, 0.0f}, {0.0f, 1.0f, 0.0f}, {0.0f, 0.0f, 1.0f}};
    v3 n = ld3(pl->normal);
    v3 b0 = {0, 0, 0}, b1 = {0, 0, 0};
    for (int i = 0; i < 3; i++) {
        v3 cr = cross3(axes[i], n);
        if (dot3(mk(1.0f, 1.0f, 1.0f), cr) == 0.0f) continue;
        b0 = cr;
        b1 = cross3(n, cr);
        break;
    }
    float ui = dot3(b0, P) * pl->m.texture_scale;
    float vi = dot3(b1, P) * pl->m.texture_scale;
    return texel(sc->tex, sc->tw, sc->th, sc->tl, emod(f2i(ui), sc->tw), emod(f2i(vi), sc->th), pl->m.texture_id);
}

/* ref: primitives.cl:262-318.  *which (optional) = index of the accepted light, for the hit-id AOV. */
static int find_light(v3 o, v3 d, const scene_t* sc, v3* color, uint32_t* which) {
    int got = 0;
    float t = INFINITY;
    v3 col = {0, 0, 0};
    uint32_t k = 0;
    for (uint32_t i = 0; i < sc->L; i++) {
        const w_light* li = &sc->lit[i];
        float t1;
        if (!hit_sphere(o, d, ld3(li->origin), li->radius, &t1) || t1 >= t) continue;
        t = t1;
        v3 ip = add(o, scl(d, t));
        float dd = dist3(o, ip);
        /* rgb*intensity*M_1_PI_F*(1/d*d): the last factor is (1/d)*d, not 1/d^2 (Q18) */
        col = scl(scl(scl(ld3(li->rgb), li->intensity), INV_PI_F), 1.0f / dd * dd);
        got = 1;
        k = i;
    }
    if (!got) return 0;
    for (uint32_t i = 0; i < sc->S; i++) {
        const w_sphere* s = &sc->sph[i];
        float t1;
        if (hit_sphere(o, d, ld3(s->origin), s->radius, &t1) && t1 <= t && !s->m.transperent) return 0; /* Q19 */
    }
    for (uint32_t i = 0; i < sc->P; i++) {
        const w_plane* p = &sc->pln[i];
        float t1;
        if (hit_plane(o, d, ld3(p->normal), ld3(p->point), &t1) && t1 <= t) return 0;
    }
    for (uint32_t i = 0; i < sc->T; i++) { /* X1: triangles occlude like planes */
        TRI_EDGES(&sc->tri[i]);
        float t1;
        if (hit_triangle(o, d, v0_, e1_, e2_, &t1) && t1 <= t) return 0;
    }
    *color = col;
    if (which) *which = k;
    return 1;
}

/* ref: primitives.cl:322-394.  Strict `_t >= t` rejection: first of equal t wins, spheres before planes.
 * *id (optional): 1+i sphere, 1+S+j plane. */
static int find_solid(v3 o, v3 d, const scene_t* sc, v3* ip_out, v3* n_out, w_material* m_out, uint32_t* id) {
    int got = 0;
    float t = INFINITY;
    v3 nrm = {0, 0, 0}, ip = {0, 0, 0};
    w_material m;
    uint32_t who = 0;
    for (uint32_t i = 0; i < sc->S; i++) {
        const w_sphere* s = &sc->sph[i];
        float t1;
        if (!hit_sphere(o, d, ld3(s->origin), s->radius, &t1) || t1 >= t) continue;
        t = t1;
        ip = add(o, scl(d, t));
        nrm = unit(sub(ip, ld3(s->origin)));
        ip = add(ip, scl(nrm, EPS_OFFSET));
        m = s->m;
        got = 1;
        who = 1 + i;
    }
    for (uint32_t j = 0; j < sc->P; j++) {
        const w_plane* p = &sc->pln[j];
        float t1;
        if (!hit_plane(o, d, ld3(p->normal), ld3(p->point), &t1) || t1 >= t) continue;
        t = t1;
        ip = add(o, scl(d, t));
        nrm = ld3(p->normal);
        m = p->m;
        if (p->m.texture_id >= 0) { /* texel sampled at the un-offset point, every time a plane takes the lead */
            v3 c = plane_texel(p, ip, sc);
            m.rgb[0] = c.x; m.rgb[1] = c.y; m.rgb[2] = c.z;
        }
        ip = add(ip, scl(nrm, EPS_OFFSET));
        got = 1;
        who = 1 + sc->S + j;
    }
    for (uint32_t k = 0; k < sc->T; k++) { /* X1: after the planes, same strict rejection; normal = unit(e1 x e2), not flipped */
        TRI_EDGES(&sc->tri[k]);
        float t1;
        if (!hit_triangle(o, d, v0_, e1_, e2_, &t1) || t1 >= t) continue;
        t = t1;
        ip = add(o, scl(d, t));
        nrm = unit(cross3(e1_, e2_));
        ip = add(ip, scl(nrm, EPS_OFFSET));
        m = sc->tri[k].m;
        got = 1;
        who = 1 + sc->S + sc->P + k;
    }
    if (!got) return 0;
    *ip_out = ip; *n_out = nrm; *m_out = m;
    if (id) *id = who;
    return 1;
}

/* ref: primitives.cl:396-442 */
static float shadow_path(v3 to, v3 from, const scene_t* sc) {
    COUNT(RTO_SHADOW_RAYS);
    v3 d = unit(sub(to, from));
    float t = dist3(to, from);
    float opacity = 1.0f;
    for (uint32_t i = 0; i < sc->S; i++) {
        const w_sphere* s = &sc->sph[i];
        float t1;
        if (!hit_sphere(from, d, ld3(s->origin), s->radius, &t1) || t1 >= t) continue;
        if (s->m.transperent) { opacity *= GLASS_PASS; continue; }
        return 0.0f;
    }
    for (uint32_t j = 0; j < sc->P; j++) {
        const w_plane* p = &sc->pln[j];
        float t1;
        if (!hit_plane(from, d, ld3(p->normal), ld3(p->point), &t1) || t1 >= t) continue;
        return 0.0f;
    }
    for (uint32_t k = 0; k < sc->T; k++) { /* X1 */
        TRI_EDGES(&sc->tri[k]);
        float t1;
        if (!hit_triangle(from, d, v0_, e1_, e2_, &t1) || t1 >= t) continue;
        return 0.0f;
    }
    return opacity;
}

/* ref: raygen.cl:5-25 */
static void primary_ray(const float* cam, uint32_t W, uint32_t id, v3* o, v3* d) {
    float w = (float)(id % W);
    float h = (float)(id / W);
    v3 corner = ld3(cam), up = ld3(cam + 6), right = ld3(cam + 9);
    float wf = cam[12], hf = cam[13];
    v3 vec = sub(add(corner, scl(scl(right, wf), w)), scl(scl(up, hf), h));
    *d = unit(vec);
    *o = ld3(cam + 3);
}

/* One stack entry.  ref: raytracing.cl:28-30 keeps rray + n + f in three parallel arrays. */
typedef struct { v3 o, d, rgb; int depth; float n, f; } entry_t;

/* ref: raytracing.cl:14-195.  Returns the packed 0x00RRGGBB word for global pixel `id`. */
static uint32_t trace_pixel(const float* cam, uint32_t W, uint32_t id, const scene_t* sc, uint32_t* hit_id) {
    entry_t st[STACK_MAX];
    uint32_t rng = id;                               /* :32-33, Q1 */
    uint32_t sz = 1;
    tl_pixel_segments = 0;
    primary_ray(cam, W, id, &st[0].o, &st[0].d);
    st[0].rgb = mk(0, 0, 0); st[0].depth = 0; st[0].n = AIR_N; st[0].f = 1.0f;

    if (hit_id) { /* primary-ray AOV; side computation, does not touch rng or colour */
        counters_t* keep = tl_cnt; tl_cnt = 0;
        v3 c, ip, nn; w_material mm; uint32_t k = 0, who = 0;
        if (find_light(st[0].o, st[0].d, sc, &c, &k)) *hit_id = HIT_LIGHT_FLAG | k;
        else if (find_solid(st[0].o, st[0].d, sc, &ip, &nn, &mm, &who)) *hit_id = who;
        else *hit_id = 0;
        tl_cnt = keep;
    }

    while (sz > 0) {
        while (st[sz - 1].depth < STACK_MAX) {
            entry_t* top = &st[sz - 1];
            COUNT(RTO_SEGMENTS);
            tl_pixel_segments++;
            v3 lcol;
            if (find_light(top->o, top->d, sc, &lcol, 0)) {          /* :48-54 */
                top->rgb = add(top->rgb, scl(lcol, top->f));
                break;
            }
            v3 ip, nrm; w_material m;
            if (!find_solid(top->o, top->d, sc, &ip, &nrm, &m, 0)) { /* :56-81 skybox */
                int u, v;
                cube_uv(top->d, (uint32_t)(sc->sw / 4), &u, &v);
                v3 px = texel(sc->sky, sc->sw, sc->sh, 1, u, sc->sh - v, 0);
                top->rgb = add(top->rgb, scl(px, top->f));
                break;
            }
            v3 mrgb = ld3(m.rgb);
            top->rgb = add(top->rgb, scl(scl(mrgb, top->f), m.ambient));   /* :83-84 (f*rgb)*ambient */

            for (uint32_t i = 0; i < sc->L; i++) {                          /* :87-136 */
                const w_light* li = &sc->lit[i];
                v3 lo = ld3(li->origin);
                float soft = 0.0f;
                v3 sdir = unit(sub(lo, ip));
                for (int j = 0; j < SHADOW_SAMPLES; j++) {
                    float theta = (float)(2 * M_PI * (double)rng_next(&rng));   /* fp64 product, narrowed */
                    float phi = (float)(M_PI * (double)rng_next(&rng));
                    float x = li->radius * sinf(phi) * cosf(theta);
                    float y = li->radius * sinf(phi) * sinf(theta);
                    float z = li->radius * cosf(phi);
                    v3 sample = add(lo, mk(x, y, z));
                    soft += shadow_path(sample, ip, sc);
                }
                float ssr = soft / (float)SHADOW_SAMPLES;
                float d = dist3(lo, ip);
                v3 lrgb = dvs(scl(scl(scl(ld3(li->rgb), li->intensity), INV_PI_F), 1.0f), d * d);  /* :118, Q7 */
                lrgb = scl(lrgb, ssr);
                v3 vv = unit(sub(top->o, ip));
                v3 hh = unit(add(vv, sdir));                                               /* Q8 */
                float ndh = dot3(nrm, hh);
                float spec = powf(0.0f < ndh ? ndh : 0.0f, (float)m.shininess);
                top->rgb = add(top->rgb, scl(scl(lrgb, top->f * m.specular), spec));      /* :129-130 */
                float ndl = dot3(nrm, sdir);
                float diff = 0.0f < ndl ? ndl : 0.0f;
                top->rgb = add(top->rgb, scl(scl(lrgb, top->f * m.diffuse), diff));       /* :134-135 */
            }

            v3 incident = top->d;
            float n1 = top->n;
            float n2 = (n1 == AIR_N) ? m.n : AIR_N;                  /* :141-144, Q4 */
            float refl = m.reflectivity;
            if (m.dielectric) {
                float fr = fresnel_schlick(n1, n2, incident, nrm);   /* un-flipped normal, Q14 */
                refl = m.reflectivity + (1.0f - m.reflectivity) * fr;
            }
            float old_f = top->f;
            top->f *= refl;
            top->d = mirror_dir(top->d, nrm);
            top->o = ip;
            top->depth++;

            if (m.transperent && sz < STACK_MAX && refl < 1.0f) {   /* :161-179 */
                entry_t* ch = &st[sz];
                *ch = *top;                                          /* after the parent's update, Q3 */
                if (n1 < n2) ch->o = sub(ch->o, scl(nrm, 2.0f * EPS_OFFSET));
                else nrm = scl(nrm, -1.0f);
                ch->f = old_f * (1.0f - refl);
                ch->rgb = mk(0, 0, 0);
                ch->n = n2;
                if (!bend_dir(n1, n2, incident, nrm, &ch->d) || ch->d.x != ch->d.x) continue;  /* TIR / isnan(dir.x): child dropped, Q5 */
                sz++;
            }
        }
        if (sz == 1) break;
        st[sz - 2].rgb = add(st[sz - 2].rgb, st[sz - 1].rgb);       /* :188, Q2 */
        sz--;
    }
    v3 c = st[0].rgb;                                                /* :193-194, Q9 */
#define CLAMP01(v) ((1.0f < (((v) < 0.0f) ? 0.0f : (v))) ? 1.0f : (((v) < 0.0f) ? 0.0f : (v)))
    float r = CLAMP01(c.x) * 255.0f;
    float g = CLAMP01(c.y) * 255.0f;
    float b = CLAMP01(c.z) * 255.0f;
    return ((uint32_t)r << 16) | ((uint32_t)g << 8) | (uint32_t)b;
}

/* ---------- frame driver: rows handed out dynamically to host threads ---------- */
typedef struct {
    const float* cam; uint32_t W, y1; const scene_t* sc;
    uint32_t *out, *hit; volatile uint32_t* next; int counted; counters_t cnt;
} job_t;

static void* row_worker(void* arg) {
    job_t* j = (job_t*)arg;
    memset(&j->cnt, 0, sizeof j->cnt);
    tl_cnt = j->counted ? &j->cnt : 0;
    for (;;) {
        uint32_t y = __atomic_fetch_add(j->next, 1u, __ATOMIC_RELAXED);
        if (y >= j->y1) break;
        for (uint32_t x = 0; x < j->W; x++) {
            uint32_t id = y * j->W + x;
            j->out[id] = trace_pixel(j->cam, j->W, id, j->sc, j->hit ? &j->hit[id] : 0);
        }
    }
    tl_cnt = 0;
    return 0;
}

int rto_render_counted(int wide, const float* cam, uint32_t W, uint32_t H, uint32_t y0, uint32_t y1,
                       const void* spheres, uint32_t S, const void* planes, uint32_t P, const void* lights,
                       uint32_t L, const uint8_t* tex, int tw, int th, int tl, const uint8_t* sky, int sw, int sh,
                       uint32_t* out, uint32_t* hit, int threads, uint64_t* counters) {
    if (!wide && (S > 255 || P > 255 || L > 255)) return -1;   /* uchar counts, ref: raytracing.cl:17 */
    if (y1 > H || y0 > y1 || W == 0) return -1;
    scene_t sc = {(const w_sphere*)spheres, S, (const w_plane*)planes, P, (const w_light*)lights, L,
                  tex, tw, th, tl, sky, sw, sh, g_tri, g_T};
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    volatile uint32_t next = y0;
    job_t* jobs = (job_t*)calloc((size_t)threads, sizeof(job_t));
    pthread_t* tid = (pthread_t*)calloc((size_t)threads, sizeof(pthread_t));
    for (int i = 0; i < threads; i++) {
        jobs[i].cam = cam; jobs[i].W = W; jobs[i].y1 = y1; jobs[i].sc = &sc;
        jobs[i].out = out; jobs[i].hit = hit; jobs[i].next = &next; jobs[i].counted = counters != 0;
        if (i > 0) pthread_create(&tid[i], 0, row_worker, &jobs[i]);
    }
    row_worker(&jobs[0]);
    for (int i = 1; i < threads; i++) pthread_join(tid[i], 0);
    if (counters) {
        memset(counters, 0, sizeof(uint64_t) * RTO_NCOUNTERS);
        for (int i = 0; i < threads; i++)
            for (int k = 0; k < RTO_NCOUNTERS; k++) counters[k] += jobs[i].cnt.c[k];
    }
    free(jobs); free(tid);
    return 0;
}

void rto_set_triangles(const void* tris, uint32_t T) { g_tri = (const w_triangle*)tris; g_T = tris ? T : 0; }
int rto_isect_triangle(const float* o, const float* d, const float* v0, const float* v1, const float* v2, float* t) {
    v3 a = ld3(v0);
    return hit_triangle(ld3(o), ld3(d), a, sub(ld3(v1), a), sub(ld3(v2), a), t);
}

int rto_render(int wide, const float* cam, uint32_t W, uint32_t H, uint32_t y0, uint32_t y1, const void* spheres,
               uint32_t S, const void* planes, uint32_t P, const void* lights, uint32_t L, const uint8_t* tex,
               int tw, int th, int tl, const uint8_t* sky, int sw, int sh, uint32_t* out, uint32_t* hit,
               int threads) {
    return rto_render_counted(wide, cam, W, H, y0, y1, spheres, S, planes, P, lights, L, tex, tw, th, tl, sky, sw,
                              sh, out, hit, threads, 0);
}

/* Cost map: segments traced per pixel (single-threaded; for load-balance analysis, not a parity surface). */
int rto_segment_map(const float* cam, uint32_t W, uint32_t H, const void* spheres, uint32_t S, const void* planes,
                    uint32_t P, const void* lights, uint32_t L, const uint8_t* tex, int tw, int th, int tl,
                    const uint8_t* sky, int sw, int sh, uint32_t* seg) {
    scene_t sc = {(const w_sphere*)spheres, S, (const w_plane*)planes, P, (const w_light*)lights, L,
                  tex, tw, th, tl, sky, sw, sh, g_tri, g_T};
    for (uint32_t id = 0; id < W * H; id++) {
        (void)trace_pixel(cam, W, id, &sc, 0);
        seg[id] = tl_pixel_segments;
    }
    return 0;
}

/* ---------- per-function known-answer entry points ---------- */
static void st3(float* p, v3 v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }
float rto_xorshift32(uint32_t* state) { return rng_next(state); }
void rto_map_to_cube(const float* dir, uint32_t face, int* uv) { cube_uv(ld3(dir), face, &uv[0], &uv[1]); }
void rto_reflect(const float* I, const float* N, float* out) { st3(out, mirror_dir(ld3(I), ld3(N))); }
void rto_refract(float n1, float n2, const float* I, const float* N, float* out) {
    v3 r;
    if (bend_dir(n1, n2, ld3(I), ld3(N), &r)) st3(out, r);
    else out[0] = out[1] = out[2] = NAN;
}
float rto_schlick(float n1, float n2, const float* I, const float* N) { return fresnel_schlick(n1, n2, ld3(I), ld3(N)); }
int rto_emod(int a, int b) { return emod(a, b); }
int rto_isect_sphere(const float* o, const float* d, const float* c, float r, float* t) {
    return hit_sphere(ld3(o), ld3(d), ld3(c), r, t);
}
int rto_isect_plane(const float* o, const float* d, const float* n, const float* p, float* t) {
    return hit_plane(ld3(o), ld3(d), ld3(n), ld3(p), t);
}
void rto_plane_texel(const void* plane, const float* P, const uint8_t* tex, int tw, int th, int tl, float* rgb) {
    scene_t sc; memset(&sc, 0, sizeof sc); sc.tex = tex; sc.tw = tw; sc.th = th; sc.tl = tl;
    st3(rgb, plane_texel((const w_plane*)plane, ld3(P), &sc));
}
void rto_raygen(const float* cam, uint32_t W, uint32_t H, uint32_t id, float* o, float* d) {
    (void)H; v3 oo, dd; primary_ray(cam, W, id, &oo, &dd); st3(o, oo); st3(d, dd);
}
int rto_find_light(const float* o, const float* d, const void* spheres, uint32_t S, const void* planes, uint32_t P,
                   const void* lights, uint32_t L, float* color) {
    scene_t sc; memset(&sc, 0, sizeof sc);
    sc.sph = (const w_sphere*)spheres; sc.S = S; sc.pln = (const w_plane*)planes; sc.P = P;
    sc.lit = (const w_light*)lights; sc.L = L;
    v3 c; int r = find_light(ld3(o), ld3(d), &sc, &c, 0);
    if (r) st3(color, c);
    return r;
}
int rto_find_solid(const float* o, const float* d, const void* spheres, uint32_t S, const void* planes, uint32_t P,
                   const uint8_t* tex, int tw, int th, int tl, float* ip, float* nrm, void* material64) {
    scene_t sc; memset(&sc, 0, sizeof sc);
    sc.sph = (const w_sphere*)spheres; sc.S = S; sc.pln = (const w_plane*)planes; sc.P = P;
    sc.tex = tex; sc.tw = tw; sc.th = th; sc.tl = tl;
    v3 a, b; w_material m;
    int r = find_solid(ld3(o), ld3(d), &sc, &a, &b, &m, 0);
    if (r) { st3(ip, a); st3(nrm, b); memcpy(material64, &m, 64); }
    return r;
}
float rto_shadow(const float* to, const float* from, const void* spheres, uint32_t S, const void* planes,
                 uint32_t P) {
    scene_t sc; memset(&sc, 0, sizeof sc);
    sc.sph = (const w_sphere*)spheres; sc.S = S; sc.pln = (const w_plane*)planes; sc.P = P;
    return shadow_path(ld3(to), ld3(from), &sc);
}
