/* oracle/ref_cl.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Builds oracle/_ref/libref_cl.so: the reference's own OpenCL C text, #included verbatim from
 * /root/reference/src/cl/ (never copied into this repo) and compiled as C++ through
 * cl_shim.hpp.  Two instantiations of the same text:
 *   narrow : uchar = 8 bit, exactly the reference (max 255 objects per type)
 *   wide   : uchar = 32 bit, types-only widening for >255 objects
 * The extern "C" functions below only marshal arguments and loop over pixels/threads.
 */
#include "cl_shim.hpp"

#include <atomic>
#include <thread>
#include <vector>

namespace clshim { thread_local uint g_global_id = 0; }

namespace ref_narrow {
CLSHIM_IMPORT
typedef unsigned char uchar;
#include "src/cl/raygen.cl"
#include "src/cl/raytracing.cl"
}  // namespace ref_narrow

#undef __TYPES_CL
#undef __PRIMITIVES_CL
#undef EPSILON
#undef INVERSE_SQUARE_LIGHT
#undef TRANSPERENT_THROUGH
#undef PRINT_VEC
#undef DEFAULT_N
#undef MAX_DEPTH
#undef MAX_SOFT_SHADOWS

namespace ref_wide {
CLSHIM_IMPORT
typedef unsigned int uchar;
#include "src/cl/raygen.cl"
#include "src/cl/raytracing.cl"
}  // namespace ref_wide

using clshim::float3;
using clshim::image2d_array_t;

static_assert(sizeof(ref_narrow::rmaterial) == 64, "rmaterial wire size");
static_assert(sizeof(ref_narrow::rsphere) == 96, "rsphere wire size");
static_assert(sizeof(ref_narrow::rplane) == 96, "rplane wire size");
static_assert(sizeof(ref_narrow::rlight) == 48, "rlight wire size");
static_assert(sizeof(ref_narrow::rray) == 64, "rray wire size");

static inline float3 f3(const float* p) { return float3(p[0], p[1], p[2]); }

/* cam = {corner xyz, origin xyz, up xyz, right xyz, w_factor, h_factor} (14 floats) */
struct RefFrame {
    const float* cam;
    uint32_t W, H;
    const void *spheres, *planes, *lights;
    uint32_t S, P, L;
    image2d_array_t tex, sky;
    uint32_t* out;
    uint32_t* hit_id;
};

#define HIT_LIGHT 0x80000000u

/* Primary-ray hit id: the reference's own closest-hit rule, evaluated with the
 * reference's intersect functions.  0 = sky, 1+i sphere, 1+S+j plane, HIT_LIGHT|k light. */
template <class NS_rray, class NS_rsphere, class NS_rplane, class NS_rlight, class FL, class IS, class IP>
static uint32_t primary_hit_id(NS_rray* ray, NS_rsphere* sp, NS_rplane* pl, NS_rlight* li, uint32_t S, uint32_t P,
                               uint32_t L, FL find_light, IS isect_sphere, IP isect_plane) {
    float3 col;
    if (find_light(ray, li, sp, pl, L, S, P, &col)) {
        float t = INFINITY; uint32_t k = 0;
        for (uint32_t i = 0; i < L; i++) {
            float _t; float3 o = li[i].origin;
            if (!isect_sphere(ray, &o, li[i].radius, &_t) || _t >= t) continue;
            t = _t; k = i;
        }
        return HIT_LIGHT | k;
    }
    float t = INFINITY; uint32_t id = 0;
    for (uint32_t i = 0; i < S; i++) {
        float _t; float3 o = sp[i].origin;
        if (!isect_sphere(ray, &o, sp[i].radius, &_t) || _t >= t) continue;
        t = _t; id = 1 + i;
    }
    for (uint32_t j = 0; j < P; j++) {
        float _t; float3 n = pl[j].normal, q = pl[j].point_in_plane;
        if (!isect_plane(ray, &n, &q, &_t) || _t >= t) continue;
        t = _t; id = 1 + S + j;
    }
    return id;
}

#define DEFINE_RENDER_ROWS(NS)                                                                          \
    static void render_rows_##NS(const RefFrame& f, uint32_t y0, uint32_t y1, NS::rray* scratch) {     \
        for (uint32_t y = y0; y < y1; y++)                                                              \
            for (uint32_t x = 0; x < f.W; x++) {                                                        \
                uint32_t id = y * f.W + x;                                                              \
                /* raygen writes rays[get_global_id]; run it with a 1-entry window onto the buffer */  \
                clshim::g_global_id = id;                                                               \
                NS::raygen(f3(f.cam), f3(f.cam + 3), f3(f.cam + 6), f3(f.cam + 9), f.cam[12], f.cam[13], \
                           f.W, f.H, scratch - id);                                                     \
                if (f.hit_id) {                                                                         \
                    NS::rray r = *scratch;                                                              \
                    f.hit_id[id] = primary_hit_id(&r, (NS::rsphere*)f.spheres, (NS::rplane*)f.planes,   \
                                                  (NS::rlight*)f.lights, f.S, f.P, f.L,                 \
                                                  NS::findLightIntersection, NS::intersect_sphere,      \
                                                  NS::intersect_plane);                                 \
                }                                                                                       \
                NS::raytracer(scratch - id, (NS::rsphere*)f.spheres, (NS::rplane*)f.planes,             \
                              (NS::rlight*)f.lights, (NS::uchar)f.S, (NS::uchar)f.P, (NS::uchar)f.L,    \
                              f.W * f.H, f.tex, f.sky, f.out - 0);                                      \
            }                                                                                           \
    }
DEFINE_RENDER_ROWS(ref_narrow)
DEFINE_RENDER_ROWS(ref_wide)

extern "C" {

/* Render rows [y0,y1) of a W x H frame into out[W*H] (global pixel ids) with `threads` host threads.
 * Returns 0 on success, -1 on bad arguments. */
int ref_render(int wide, const float* cam, uint32_t W, uint32_t H, uint32_t y0, uint32_t y1, const void* spheres,
               uint32_t S, const void* planes, uint32_t P, const void* lights, uint32_t L, const uint8_t* tex,
               int tw, int th, int tl, const uint8_t* sky, int sw, int sh, uint32_t* out, uint32_t* hit_id,
               int threads) {
    if (!wide && (S > 255 || P > 255 || L > 255)) return -1;
    if (y1 > H || y0 > y1) return -1;
    RefFrame f{cam, W, H, spheres, planes, lights, S, P, L, {tex, tw, th, tl}, {sky, sw, sh, 1}, out, hit_id};
    if (threads < 1) threads = 1;
    std::atomic<uint32_t> next(y0);
    auto worker = [&]() {
        alignas(16) unsigned char raybuf[64];
        for (;;) {
            uint32_t y = next.fetch_add(1);
            if (y >= y1) break;
            if (wide) render_rows_ref_wide(f, y, y + 1, (ref_wide::rray*)raybuf);
            else render_rows_ref_narrow(f, y, y + 1, (ref_narrow::rray*)raybuf);
        }
    };
    std::vector<std::thread> pool;
    for (int i = 1; i < threads; i++) pool.emplace_back(worker);
    worker();
    for (auto& t : pool) t.join();
    return 0;
}

int ref_hw_threads(void) { int n = (int)std::thread::hardware_concurrency(); return n > 0 ? n : 1; }

}  // extern "C"

/* ---- per-function known-answer entry points: thin marshalling onto the reference's own functions
 * (narrow instantiation).  Mirrors the rto_* set of rt_oracle.h. ---- */
namespace rn = ref_narrow;
static inline void st3(float* p, const float3& v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }
static inline rn::rray mkray(const float* o, const float* d) {
    rn::rray r; r.origin = f3(o); r.dir = f3(d); r.rgb = float3(0.0f); r.depth = 0; return r;
}
extern "C" {
float ref_xorshift32(uint32_t* state) { rn::xorshift32_state s; s.x = *state; float v = rn::xorshift32(&s); *state = s.x; return v; }
void ref_map_to_cube(const float* dir, uint32_t face, int* uv) { float3 d = f3(dir); clshim::int2 r = rn::map_to_cube(&d, face); uv[0] = r.x; uv[1] = r.y; }
void ref_reflect(const float* I, const float* N, float* out) { float3 i = f3(I), n = f3(N); st3(out, rn::reflect(&i, &n)); }
void ref_refract(float n1, float n2, const float* I, const float* N, float* out) { float3 i = f3(I), n = f3(N); st3(out, rn::refract(n1, n2, &i, &n)); }
float ref_schlick(float n1, float n2, const float* I, const float* N) { float3 i = f3(I), n = f3(N); return rn::compute_schlick(n1, n2, &i, &n); }
int ref_emod(int a, int b) { return rn::euclidean_modulo(a, b); }
int ref_isect_sphere(const float* o, const float* d, const float* c, float r, float* t) {
    rn::rray ray = mkray(o, d); float3 cc = f3(c); return rn::intersect_sphere(&ray, &cc, r, t) ? 1 : 0;
}
int ref_isect_plane(const float* o, const float* d, const float* n, const float* p, float* t) {
    rn::rray ray = mkray(o, d); float3 nn = f3(n), pp = f3(p); return rn::intersect_plane(&ray, &nn, &pp, t) ? 1 : 0;
}
void ref_plane_texel(const void* plane, const float* P, const uint8_t* tex, int tw, int th, int tl, float* rgb) {
    rn::rplane pl = *(const rn::rplane*)plane; float3 ip = f3(P); image2d_array_t im{tex, tw, th, tl};
    st3(rgb, rn::plane_texture_pixel(&pl, &ip, im));
}
void ref_raygen(const float* cam, uint32_t W, uint32_t H, uint32_t id, float* o, float* d) {
    alignas(16) unsigned char buf[64]; rn::rray* one = (rn::rray*)buf;
    clshim::g_global_id = id;
    rn::raygen(f3(cam), f3(cam + 3), f3(cam + 6), f3(cam + 9), cam[12], cam[13], W, H, one - id);
    st3(o, one->origin); st3(d, one->dir);
}
int ref_find_light(const float* o, const float* d, const void* spheres, uint32_t S, const void* planes, uint32_t P,
                   const void* lights, uint32_t L, float* color) {
    rn::rray ray = mkray(o, d); float3 c;
    bool r = rn::findLightIntersection(&ray, (rn::rlight*)lights, (rn::rsphere*)spheres, (rn::rplane*)planes, L, S, P, &c);
    if (r) st3(color, c);
    return r ? 1 : 0;
}
int ref_find_solid(const float* o, const float* d, const void* spheres, uint32_t S, const void* planes, uint32_t P,
                   const uint8_t* tex, int tw, int th, int tl, float* ip, float* nrm, void* material64) {
    rn::rray ray = mkray(o, d); float3 a, b; rn::rmaterial m; image2d_array_t im{tex, tw, th, tl};
    bool r = rn::findSolidIntersection(&ray, (rn::rsphere*)spheres, (rn::rplane*)planes, (rn::uchar)S, (rn::uchar)P, &a, &b, &m, im);
    if (r) { st3(ip, a); st3(nrm, b); memcpy(material64, &m, 64); }
    return r ? 1 : 0;
}
float ref_shadow(const float* to, const float* from, const void* spheres, uint32_t S, const void* planes, uint32_t P) {
    float3 t = f3(to), f = f3(from);
    return rn::testShadowPath(&t, &f, (rn::rsphere*)spheres, (rn::rplane*)planes, S, P);
}
}  // extern "C"
