/* oracle/rt_oracle.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C CPU restatement of the reference's per-pixel hot path
 * (/root/reference/src/cl/{raygen,raytracing,primitives,types}.cl).  It exists to CHECK the CUDA
 * path; nothing under example-gui-opencl-raytracer_b200/ may include, link or call it.
 *
 * Pinning: validated bit-for-bit (colour words and hit ids) against oracle/_ref/libref_cl.so --
 * the reference's own .cl text compiled as C++ -- by tests/test_oracle_pinning.py, and against the
 * golden buffers under tests/golden/ that were generated from that library.  The reference has no
 * tests or golden vectors of its own, and the OpenCL built-ins
 * (normalize/distance/pow/sin/cos/contraction) are defined by whichever runtime runs the .cl, so
 * relative to a *real* OpenCL device: parity unpinned at the built-in level.  The definitions
 * chosen are listed in cl_shim.hpp's header and repeated next to each helper in rt_oracle.c.
 */
#ifndef RT_ORACLE_H
#define RT_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Work counters ("ray" = segment or shadow ray). */
enum {
    RTO_SEGMENTS = 0,     /* iterations of raytracing.cl:42 */
    RTO_SHADOW_RAYS = 1,  /* testShadowPath calls, raytracing.cl:108 */
    RTO_SPHERE_TESTS = 2, /* intersect_sphere calls (spheres and lights) */
    RTO_SPHERE_DPOS = 3,  /* ... of which reached the sqrt (D >= 0) */
    RTO_PLANE_TESTS = 4,  /* intersect_plane calls */
    RTO_TEXELS = 5,       /* read_imagei calls (plane texture + skybox) */
    RTO_TRI_TESTS = 7,    /* EXTENSION X1: triangle tests (no reference counterpart) */
    RTO_NCOUNTERS = 8
};

/* cam = {corner xyz, origin xyz, up xyz, right xyz, w_factor, h_factor}.  Scene arrays are the raw
 * wire structs (rsphere 96 B, rplane 96 B, rlight 48 B).  Renders rows [y0,y1) of a W x H frame into
 * out[W*H] / hit[W*H] (hit may be NULL) at GLOBAL pixel ids.  counters (may be NULL) receives
 * RTO_NCOUNTERS totals.  Returns 0, or -1 on bad arguments (narrow mode with > 255 objects). */
int rto_render(int wide, const float* cam, uint32_t W, uint32_t H, uint32_t y0, uint32_t y1, const void* spheres,
               uint32_t S, const void* planes, uint32_t P, const void* lights, uint32_t L, const uint8_t* tex,
               int tw, int th, int tl, const uint8_t* sky, int sw, int sh, uint32_t* out, uint32_t* hit,
               int threads);
int rto_render_counted(int wide, const float* cam, uint32_t W, uint32_t H, uint32_t y0, uint32_t y1,
                       const void* spheres, uint32_t S, const void* planes, uint32_t P, const void* lights,
                       uint32_t L, const uint8_t* tex, int tw, int th, int tl, const uint8_t* sky, int sw, int sh,
                       uint32_t* out, uint32_t* hit, int threads, uint64_t* counters);

/* EXTENSION X1 (no reference counterpart; parity unpinned): triangle records (112 B: v0, v1, v2 as float4, then
 * rmaterial) used by every later rto_render* call until reset with (NULL, 0).  Slotted after the planes in the
 * three traversal loops; opaque; normal = normalize(cross(v1-v0, v2-v0)). */
void rto_set_triangles(const void* tris, uint32_t T);
int rto_isect_triangle(const float* o, const float* d, const float* v0, const float* v1, const float* v2, float* t);

/* Segments traced per pixel (cost map for load-balance analysis; single-threaded). */
int rto_segment_map(const float* cam, uint32_t W, uint32_t H, const void* spheres, uint32_t S, const void* planes,
                    uint32_t P, const void* lights, uint32_t L, const uint8_t* tex, int tw, int th, int tl,
                    const uint8_t* sky, int sw, int sh, uint32_t* seg);

/* Per-function known-answer entry points (same names with the ref_ prefix exist in _ref/libref_cl.so). */
float rto_xorshift32(uint32_t* state);
void rto_map_to_cube(const float* dir, uint32_t face, int* uv);
void rto_reflect(const float* I, const float* N, float* out);
void rto_refract(float n1, float n2, const float* I, const float* N, float* out);
float rto_schlick(float n1, float n2, const float* I, const float* N);
int rto_emod(int a, int b);
int rto_isect_sphere(const float* o, const float* d, const float* c, float r, float* t);
int rto_isect_plane(const float* o, const float* d, const float* n, const float* p, float* t);
void rto_plane_texel(const void* plane, const float* P, const uint8_t* tex, int tw, int th, int tl, float* rgb);
void rto_raygen(const float* cam, uint32_t W, uint32_t H, uint32_t id, float* o, float* d);
int rto_find_light(const float* o, const float* d, const void* spheres, uint32_t S, const void* planes, uint32_t P,
                   const void* lights, uint32_t L, float* color);
int rto_find_solid(const float* o, const float* d, const void* spheres, uint32_t S, const void* planes, uint32_t P,
                   const uint8_t* tex, int tw, int th, int tl, float* ip, float* nrm, void* material64);
float rto_shadow(const float* to, const float* from, const void* spheres, uint32_t S, const void* planes,
                 uint32_t P);

#ifdef __cplusplus
}
#endif
#endif
