/* oracle/cl_shim.hpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A C++ view of the subset of OpenCL C that the reference's device sources use, so that
 * the reference text (src/cl/types.cl, primitives.cl, raygen.cl, raytracing.cl under
 * /root/reference) can be #included verbatim and executed on the host cores.  Nothing in
 * here restates the tracer; it only supplies the language built-ins.
 *
 * Built-in definitions pinned by this oracle (the reference leaves them to whichever OpenCL
 * runtime is installed, SURVEY.md 8(c), so "parity unpinned at the built-in level"):
 *   - no floating-point contraction (compile with -ffp-contract=off), IEEE-754 binary32
 *     round-to-nearest-even for + - * / and sqrt;
 *   - dot(a,b)      = (a.x*b.x + a.y*b.y) + a.z*b.z
 *   - normalize(v)  = v / sqrtf(dot(v,v))            (three correctly rounded divisions)
 *   - distance(a,b) = sqrtf(dot(a-b, a-b))
 *   - cross         = the textbook component formula, each component a*b - c*d
 *   - sin/cos/pow   = the host libm's sinf/cosf/powf (glibc: < 1 ulp)
 *   - (int)f out of range = x86 cvttss2si result (INT_MIN)
 *   - read_imagei out of range = clamp-to-edge (undefined in OpenCL)
 */
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

namespace clshim {

typedef unsigned int uint;

struct alignas(16) float3 {
    float x, y, z, w;
    float3() : x(0), y(0), z(0), w(0) {}
    float3(float s) : x(s), y(s), z(s), w(0) {}                 /* scalar splat (raytracing.cl:128,133) */
    float3(float a, float b, float c) : x(a), y(b), z(c), w(0) {}
    float3& operator+=(const float3& o) { x += o.x; y += o.y; z += o.z; return *this; }
    float3& operator-=(const float3& o) { x -= o.x; y -= o.y; z -= o.z; return *this; }
    float3& operator*=(float s) { x *= s; y *= s; z *= s; return *this; }
    float3& operator*=(const float3& o) { x *= o.x; y *= o.y; z *= o.z; return *this; }
};
static inline float3 operator+(const float3& a, const float3& b) { return float3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline float3 operator-(const float3& a, const float3& b) { return float3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline float3 operator*(const float3& a, const float3& b) { return float3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline float3 operator*(const float3& a, float s) { return float3(a.x * s, a.y * s, a.z * s); }
static inline float3 operator*(float s, const float3& a) { return float3(s * a.x, s * a.y, s * a.z); }
static inline float3 operator*(const float3& a, int s) { return a * (float)s; }
static inline float3 operator*(int s, const float3& a) { return (float)s * a; }
static inline float3 operator/(const float3& a, float s) { return float3(a.x / s, a.y / s, a.z / s); }
static inline float3 operator-(const float3& a) { return float3(-a.x, -a.y, -a.z); }

struct int2 {
    int x, y;
    int& operator[](int i) { return i == 0 ? x : y; }
};
struct int4 {
    int x, y, z, w;
};

/* ---- geometric / math built-ins ---- */
static inline float dot(const float3& a, const float3& b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
static inline float3 cross(const float3& a, const float3& b) {
    return float3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
static inline float sqrt(float v) { return ::sqrtf(v); }
static inline float sin(float v) { return ::sinf(v); }
static inline float cos(float v) { return ::cosf(v); }
static inline float pow(float a, float b) { return ::powf(a, b); }
static inline float fabs(float v) { return ::fabsf(v); }
static inline float max(float a, float b) { return (a < b) ? b : a; }
static inline float min(float a, float b) { return (b < a) ? b : a; }
static inline bool isnan(float v) { return v != v; }
static inline float3 normalize(const float3& v) { return v / sqrt(dot(v, v)); }
static inline float distance(const float3& a, const float3& b) { float3 d = a - b; return sqrt(dot(d, d)); }
static inline float clamp(float v, float lo, float hi) { return min(max(v, lo), hi); }
static inline float3 clamp(const float3& v, float lo, float hi) {
    return float3(clamp(v.x, lo, hi), clamp(v.y, lo, hi), clamp(v.z, lo, hi));
}

/* ---- images: image2d_array_t of CL_RGBA / CL_UNSIGNED_INT8 ---- */
struct image2d_array_t {
    const uint8_t* texels; /* layers * h * w * 4 bytes, row 0 = top row */
    int w, h, layers;
};
static inline int2 get_image_dim(const image2d_array_t& im) { int2 d; d.x = im.w; d.y = im.h; return d; }
static inline int4 read_imagei(const image2d_array_t& im, const int4& c) {
    int x = c.x < 0 ? 0 : (c.x >= im.w ? im.w - 1 : c.x);
    int y = c.y < 0 ? 0 : (c.y >= im.h ? im.h - 1 : c.y);
    int l = c.z < 0 ? 0 : (c.z >= im.layers ? im.layers - 1 : c.z);
    const uint8_t* p = im.texels + (((size_t)l * im.h + y) * im.w + x) * 4;
    int4 r; r.x = p[0]; r.y = p[1]; r.z = p[2]; r.w = p[3];
    return r;
}

/* ---- work-item id ---- */
extern thread_local uint g_global_id;
static inline uint get_global_id(int) { return g_global_id; }

}  // namespace clshim

/* using-declarations (not a using-directive) so the shim's built-ins HIDE libm's global overloads */
#define CLSHIM_IMPORT                                                                             \
    using clshim::uint; using clshim::float3; using clshim::int2; using clshim::int4;             \
    using clshim::dot; using clshim::cross; using clshim::sqrt; using clshim::sin;                \
    using clshim::cos; using clshim::pow; using clshim::fabs; using clshim::max;                  \
    using clshim::min; using clshim::isnan; using clshim::normalize; using clshim::distance;      \
    using clshim::clamp; using clshim::image2d_array_t; using clshim::get_image_dim;              \
    using clshim::read_imagei; using clshim::get_global_id;

#define __kernel
#define __global
#define read_only
#define M_1_PI_F 0.318309886183790671538f
