/* rt_engine.cu -- the C ABI of include/rtb200.h: device memory, uploads, launch bookkeeping.
 * No CPU rendering path exists here: every render call ends in rt_launch_render (rt_kernels.cu). */
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <cmath>
#include <utility>
#include <vector>

#include "../../include/rtb200.h"
#include "rt_kernels.h"

namespace {

thread_local char g_err[512] = "ok";

int fail(int code, const char* what, cudaError_t ce = cudaSuccess) {
    if (ce != cudaSuccess) snprintf(g_err, sizeof g_err, "%s: %s", what, cudaGetErrorString(ce));
    else snprintf(g_err, sizeof g_err, "%s", what);
    return code;
}
#define CU(call, what)                                        \
    do {                                                      \
        cudaError_t ce_ = (call);                             \
        if (ce_ != cudaSuccess) return fail(-2, what, ce_);   \
    } while (0)

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    int ensure(size_t n) {
        if (n <= cap && p) return 0;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        size_t want = n ? n : 16;
        cudaError_t ce = cudaMalloc(&p, want);
        if (ce != cudaSuccess) return fail(-2, "cudaMalloc", ce);
        cap = want;
        return 0;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

}  // namespace

struct rtb_engine {
    int device = 0;
    int sm_count = 0;
    size_t smem_optin = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, evk = nullptr, ev1 = nullptr;
    bool timed = false;

    DevBuf spheres, planes, lights; /* owned copies of the wire arrays */
    const unsigned char *d_spheres = nullptr, *d_planes = nullptr, *d_lights = nullptr; /* owned or bound */
    uint32_t S = 0, P = 0, L = 0;
    DevBuf blob;
    DevBuf tris, tri_rec; /* extension X1 */
    const unsigned char* d_tris = nullptr;
    uint32_t T = 0;
    RtBlobLayout lay{};
    bool have_scene = false;

    DevBuf tex, sky;                             /* owned copies ... */
    const void *d_tex = nullptr, *d_sky = nullptr; /* ... or arrays bound in place */
    uint32_t tw = 0, th = 0, tl = 0, sw = 0, sh = 0;

    float cam[14];
    bool have_cam = false;

    DevBuf frame, hit, sched; /* sched: tile counter (4 B) + work counters */
    DevBuf tile_cost[2], tile_order[2], tile_hist[2]; /* two sets of queue buffers (and two 256-byte halves of `sched`): the
                                                         frame being rendered reads one while the next frame's queue is built in the other */
    int set_cur = 0;          /* the set the most recent ordered frame used */
    int cost_hist_set = -1;   /* the set whose cost buffer holds measured-time costs of a frame with the tk_* geometry (-1: none) */
    /* schedule look-ahead: right after frame n's kernels are in the stream, frame n+1's queue is built on a second
     * stream from frame n-1's tile times, so that the two small scheduling kernels leave the frame's critical path */
    cudaStream_t stream2 = nullptr;
    cudaEvent_t ev_look = nullptr;
    bool look_pending = false; /* ev_look has been recorded and the main stream has not waited for it yet */
    bool look_valid = false;   /* ... and what it describes: */
    int look_set = 0, look_time = 0, look_launcher = 0;
    uint32_t look_W = 0, look_H = 0, look_total = 0, look_band = 0, look_parts = 0, look_part = 0;
    uint32_t look_split = 0, look_split32 = 0, look_tree = 0, look_group = 0, look_cih = 0;
    unsigned long long look_scene = 0;
    int opt_lookahead = 1;
    uint32_t last_sched_off = 0; /* which half of `sched` the last frame used (rtb_debug_counters) */
    DevBuf tree_slab;         /* per-warp node storage of the level-wise pixel-tree trace (rt_trace_tree.cuh) */
    DevBuf jump;              /* [15][32] xorshift32 jump-ahead matrices for this scene's light count */
    uint32_t jump_L = 0xffffffffu;
    uint32_t* noalloc_scratch = nullptr; /* device address of g_noalloc_scratch on this engine's device (rtb_render_noalloc) */
    int opt_tree_ctas_per_sm = 3;
    int opt_tri_tree_min = 1024; /* triangle scenes from this many triangles on render every tile in the tree kernel (0 = never) */
    int opt_accel = 0;        /* N4: 1 = many-sphere scenes walk the cluster structure instead of every record */
    DevBuf accel;             /* clusters, cluster-ordered records, original indices */
    uint32_t acc_nc = 0;
    DevBuf tacc;              /* N4 over the triangles: nodes (all levels), cluster-ordered records, original indices */
    bool tacc_valid = false;  /* built (or found unbuildable: tacc_nc == 0) for the current triangle array */
    uint32_t tacc_nc = 0, tacc_nl = 0, tacc_nodes = 0, tacc_ntop = 0;
    int opt_tree = 1, opt_tree_cap = 8192, opt_tree_min_cost = 0 /* 0 = automatic */, opt_tree_group_max_cost = 0;
    DevBuf tile_time[3];      /* measured per-tile times: this frame's, the previous frame's, and the one before (which the
                                 look-ahead clears for the next frame) */
    int time_cur = 0;         /* the buffer the next history frame writes */
    bool time_valid = false;  /* tile_time[(time_cur + 2) % 3] describes a frame of the geometry below */
    uint32_t tk_W = 0, tk_H = 0, tk_total = 0, tk_band = 0, tk_parts = 0, tk_part = 0;
    unsigned long long scene_gen = 0, tk_scene_gen = 0;
    int opt_trace = 0, opt_tile_order = 1, opt_coop_min_cost = 0, opt_split_min_cost = 0 /* 0 = automatic */, opt_split32_min_cost = 0 /* 0 = automatic */, opt_history = 1, opt_wide_min_records = 64, opt_wide_max_live = -1;
    uint32_t* mirror = nullptr; /* one-shot, see rtb_set_mirror */
    bool mirror_global = false;
    uint32_t* done_flag = nullptr; /* one-shot, see rtb_set_completion */
    uint32_t done_value = 0;
    DevBuf wait_word;              /* timeout counter of rtb_wait_flags */
    rtb_launch_info last{};
};

extern "C" {

const char* rtb_last_error(void) { return g_err; }

int rtb_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

void rtb_destroy(rtb_engine* e) {
    if (!e) return;
    cudaSetDevice(e->device);
    if (e->stream) cudaStreamSynchronize(e->stream);
    if (e->stream2) cudaStreamSynchronize(e->stream2);
    e->tris.release(); e->tri_rec.release(); e->spheres.release(); e->planes.release(); e->lights.release(); e->blob.release();
    e->tex.release(); e->sky.release(); e->frame.release(); e->hit.release(); e->sched.release();
    for (int k = 0; k < 2; k++) { e->tile_cost[k].release(); e->tile_order[k].release(); e->tile_hist[k].release(); }
    for (int k = 0; k < 3; k++) e->tile_time[k].release();
    e->tree_slab.release(); e->wait_word.release(); e->jump.release(); e->accel.release(); e->tacc.release();
    if (e->ev0) cudaEventDestroy(e->ev0);
    if (e->ev1) cudaEventDestroy(e->ev1);
    if (e->evk) cudaEventDestroy(e->evk);
    if (e->ev_look) cudaEventDestroy(e->ev_look);
    if (e->stream2) cudaStreamDestroy(e->stream2);
    if (e->stream) cudaStreamDestroy(e->stream);
    delete e;
}

static int create_impl(rtb_engine* e) {
    CU(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking), "cudaStreamCreate");
    CU(cudaEventCreate(&e->ev0), "cudaEventCreate");
    CU(cudaEventCreate(&e->ev1), "cudaEventCreate");
    CU(cudaEventCreate(&e->evk), "cudaEventCreate");
    CU(cudaStreamCreateWithFlags(&e->stream2, cudaStreamNonBlocking), "cudaStreamCreate");
    if (const char* la = getenv("RTB_LOOKAHEAD")) e->opt_lookahead = atoi(la) != 0; /* A/B hook for callers behind cl_wrap_* (option "lookahead") */
    CU(cudaEventCreateWithFlags(&e->ev_look, cudaEventDisableTiming), "cudaEventCreate");
    if (e->sched.ensure(512)) return -2;
    CU(cudaMemsetAsync(e->sched.p, 0, 512, e->stream), "cudaMemsetAsync");
    CU(cudaStreamSynchronize(e->stream), "cudaStreamSynchronize");
    return 0;
}

int rtb_create(rtb_engine** out, int device) {
    if (!out) return fail(-1, "rtb_create: null out");
    int n = 0;
    CU(cudaGetDeviceCount(&n), "cudaGetDeviceCount");
    if (device < 0 || device >= n) return fail(-1, "rtb_create: no such CUDA device");
    CU(cudaSetDevice(device), "cudaSetDevice");
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device), "cudaGetDeviceProperties");
    if (prop.major < 10) return fail(-3, "rtb_create: this library is built for sm_100a only");
    rtb_engine* e = new rtb_engine();
    e->device = device;
    e->sm_count = prop.multiProcessorCount;
    e->smem_optin = prop.sharedMemPerBlockOptin;
    memset(e->cam, 0, sizeof e->cam);
    const int rc = create_impl(e);
    if (rc) { /* g_err already says why; nothing of the half-built engine survives */
        rtb_destroy(e);
        return rc;
    }
    *out = e;
    return 0;
}

/* xorshift32 (primitives.cl:111-125) is linear over GF(2).  T = step^(4L) advances the generator by one shaded segment
 * (2 samples x 2 draws per light, raytracing.cl:96-113); table[j][i] = image of bit i under T^(2^j), j < 15 (a pixel
 * has < 2^15 segments).  The k-th shaded segment of a pixel starts from prod_{j in bits(k)} T^(2^j) applied to the
 * pixel id (oracle/rt_split.c:jump_init is the CPU proof; tests/test_split_cpu.py). */
static uint32_t xs_step(uint32_t x) { x ^= x << 13; x ^= x >> 17; x ^= x << 5; return x; }
static uint32_t bm_apply(const uint32_t* col, uint32_t x) {
    uint32_t r = 0;
    for (int i = 0; i < 32; i++) if ((x >> i) & 1u) r ^= col[i];
    return r;
}
static void build_jump_table(uint32_t L, uint32_t tab[15][32]) {
    for (int i = 0; i < 32; i++) {
        uint32_t x = 1u << i;
        for (uint32_t d = 0; d < 4u * L; d++) x = xs_step(x);
        tab[0][i] = x;
    }
    for (int j = 1; j < 15; j++)
        for (int i = 0; i < 32; i++) tab[j][i] = bm_apply(tab[j - 1], tab[j - 1][i]);
}
static int prep_jump(rtb_engine* e) {
    if (e->jump_L == e->L && e->jump.p) return 0;
    uint32_t tab[15][32];
    build_jump_table(e->L, tab);
    if (e->jump.ensure(sizeof tab)) return -2;
    CU(cudaMemcpyAsync(e->jump.p, tab, sizeof tab, cudaMemcpyHostToDevice, e->stream), "H2D jump table");
    CU(cudaStreamSynchronize(e->stream), "jump table upload"); /* tab is on this stack frame */
    e->jump_L = e->L;
    return 0;
}

/* N4 (csrc/rt_trace_accel.cuh): spheres in Morton order of their centres, clusters of RT_ACCEL_K, per cluster a bounding
 * sphere computed in fp64 around the (float) centre and inflated.  `host_spheres` = the wire array (96-byte records). */
static uint32_t morton10(uint32_t v) { /* spread 10 bits to every third position */
    v &= 1023u;
    v = (v | (v << 16)) & 0x030000FFu;
    v = (v | (v << 8)) & 0x0300F00Fu;
    v = (v | (v << 4)) & 0x030C30C3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
}
struct AccelHost {
    std::vector<float> cl, rec, grp;
    std::vector<uint32_t> idx;
    uint32_t nc = 0, ng = 0;
};
/* host part: 1 = built, 0 = this scene is walked by brute force (too few spheres, a non-finite sphere, huge extents) */
static int build_accel_host(const unsigned char* host_spheres, uint32_t S, AccelHost& out) {
    const uint32_t K = RT_ACCEL_K;
    out.nc = out.ng = 0;
    if (S < 2u * K || !host_spheres) return 0;
    std::vector<double> cx(S), cy(S), cz(S), rr(S);
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
    for (uint32_t i = 0; i < S; i++) {
        float c[4], r;
        memcpy(c, host_spheres + (size_t)i * RT_SPHERE_BYTES, 16);
        memcpy(&r, host_spheres + (size_t)i * RT_SPHERE_BYTES + 16, 4);
        if (!std::isfinite(c[0]) || !std::isfinite(c[1]) || !std::isfinite(c[2]) || !std::isfinite(r)) return 0; /* brute force */
        cx[i] = c[0]; cy[i] = c[1]; cz[i] = c[2]; rr[i] = fabs((double)r);
        const double v[3] = {cx[i], cy[i], cz[i]};
        for (int k = 0; k < 3; k++) { lo[k] = v[k] < lo[k] ? v[k] : lo[k]; hi[k] = v[k] > hi[k] ? v[k] : hi[k]; }
    }
    std::vector<std::pair<uint32_t, uint32_t>> order(S);
    double span = 0;
    for (int k = 0; k < 3; k++) span = hi[k] - lo[k] > span ? hi[k] - lo[k] : span;
    for (uint32_t i = 0; i < S; i++) {
        const double v[3] = {cx[i], cy[i], cz[i]};
        uint32_t code = 0;
        for (int k = 0; k < 3; k++) { /* ONE scale for the three axes: a thin axis (spheres resting on a plane) must not scatter the order */
            const uint32_t q = span > 0 ? (uint32_t)fmin(1023.0, (v[k] - lo[k]) / span * 1024.0) : 0u;
            code |= morton10(q) << k;
        }
        order[i] = {code, i};
    }
    std::sort(order.begin(), order.end());
    const uint32_t nc = (S + K - 1) / K;
    std::vector<float>&cl = out.cl, &rec = out.rec, &grp = out.grp;
    std::vector<uint32_t>& idx = out.idx;
    cl.assign(4 * (size_t)nc, 0.0f);
    rec.assign(4 * (size_t)nc * K, 0.0f);
    idx.assign((size_t)nc * K, 0xffffffffu);
    for (uint32_t k = 0; k < nc; k++) {
        const uint32_t b = k * K, n = (b + K <= S ? K : S - b);
        double mx = 0, my = 0, mz = 0;
        for (uint32_t j = 0; j < n; j++) { const uint32_t i = order[b + j].second; mx += cx[i]; my += cy[i]; mz += cz[i]; }
        const float C[3] = {(float)(mx / n), (float)(my / n), (float)(mz / n)};
        double R = 0;
        for (uint32_t j = 0; j < n; j++) {
            const uint32_t i = order[b + j].second;
            const double dx = cx[i] - C[0], dy = cy[i] - C[1], dz = cz[i] - C[2];
            const double d = sqrt(dx * dx + dy * dy + dz * dz) + rr[i];
            R = d > R ? d : R;
        }
        R = R * (1.0 + 1e-5) + 1e-30;
        const double w = 1.125 * R * R * (1.0 + 1e-6);
        if (!(w < 1e37)) return 0;
        cl[4 * k + 0] = C[0]; cl[4 * k + 1] = C[1]; cl[4 * k + 2] = C[2]; cl[4 * k + 3] = nextafterf((float)w, INFINITY);
        for (uint32_t j = 0; j < K; j++) {
            float* o = &rec[4 * ((size_t)b + j)];
            if (j < n) {
                const uint32_t i = order[b + j].second;
                float c[4], r;
                uint32_t transp;
                memcpy(c, host_spheres + (size_t)i * RT_SPHERE_BYTES, 16);
                memcpy(&r, host_spheres + (size_t)i * RT_SPHERE_BYTES + 16, 4);
                memcpy(&transp, host_spheres + (size_t)i * RT_SPHERE_BYTES + RT_MATERIAL_OFFSET + 32, 4);
                o[0] = c[0]; o[1] = c[1]; o[2] = c[2]; o[3] = r * r; /* as scene_prep does: one fp32 product */
                idx[(size_t)b + j] = i | (transp ? 0x80000000u : 0u);
            } else {
                o[0] = o[1] = o[2] = 0.0f; o[3] = -INFINITY; /* never hit */
            }
        }
    }
    /* second level: one bounding sphere per RT_ACCEL_G consecutive clusters, around their member spheres */
    const uint32_t ng = (nc + RT_ACCEL_G - 1) / RT_ACCEL_G;
    grp.assign(4 * (size_t)ng, 0.0f);
    for (uint32_t q = 0; q < ng; q++) {
        const uint32_t b = q * RT_ACCEL_G * K, end = (b + RT_ACCEL_G * K < S ? b + RT_ACCEL_G * K : S), n = end - b;
        double mx = 0, my = 0, mz = 0;
        for (uint32_t j = b; j < end; j++) { const uint32_t i = order[j].second; mx += cx[i]; my += cy[i]; mz += cz[i]; }
        const float C[3] = {(float)(mx / n), (float)(my / n), (float)(mz / n)};
        double R = 0;
        for (uint32_t j = b; j < end; j++) {
            const uint32_t i = order[j].second;
            const double dx = cx[i] - C[0], dy = cy[i] - C[1], dz = cz[i] - C[2];
            const double d = sqrt(dx * dx + dy * dy + dz * dz) + rr[i];
            R = d > R ? d : R;
        }
        R = R * (1.0 + 1e-5) + 1e-30;
        const double w = 1.125 * R * R * (1.0 + 1e-6);
        if (!(w < 1e37)) return 0;
        grp[4 * q + 0] = C[0]; grp[4 * q + 1] = C[1]; grp[4 * q + 2] = C[2]; grp[4 * q + 3] = nextafterf((float)w, INFINITY);
    }
    out.nc = nc;
    out.ng = ng;
    return 1;
}
static int prep_accel(rtb_engine* e, const unsigned char* host_spheres) {
    e->acc_nc = 0;
    AccelHost h;
    if (build_accel_host(host_spheres, e->S, h) != 1) return 0;
    const uint32_t K = RT_ACCEL_K, nc = h.nc, ng = h.ng;
    std::vector<float>&cl = h.cl, &rec = h.rec, &grp = h.grp;
    std::vector<uint32_t>& idx = h.idx;
    const size_t b0 = (size_t)nc * 16, b1 = (size_t)nc * K * 16, b2 = ((size_t)nc * K * 4 + 15) / 16 * 16, b3 = (size_t)ng * 16;
    if (e->accel.ensure(b0 + b1 + b2 + b3)) return -2;
    CU(cudaMemcpyAsync((char*)e->accel.p + b0 + b1 + b2, grp.data(), b3, cudaMemcpyHostToDevice, e->stream), "H2D accel");
    CU(cudaMemcpyAsync(e->accel.p, cl.data(), b0, cudaMemcpyHostToDevice, e->stream), "H2D accel");
    CU(cudaMemcpyAsync((char*)e->accel.p + b0, rec.data(), b1, cudaMemcpyHostToDevice, e->stream), "H2D accel");
    CU(cudaMemcpyAsync((char*)e->accel.p + b0 + b1, idx.data(), (size_t)nc * K * 4, cudaMemcpyHostToDevice, e->stream), "H2D accel");
    CU(cudaStreamSynchronize(e->stream), "accel upload"); /* the vectors die with this frame */
    e->acc_nc = nc;
    return 0;
}

/* N4 over the triangles of extension X1 (csrc/rt_trace_accel.cuh): `rec` = the {v0, e1, e2} records tri_prep made (12
 * floats each; the kernel's test sees the triangle v0, v0 + e1, v0 + e2, so that is what the boxes bound).  Triangles in
 * a kd-style order (below), clusters of RT_ACCEL_K, then 8 nodes per parent until a level has <= 4 nodes or
 * RT_TACC_LEVELS levels exist.  Every node: the axis-aligned box of the vertices below it as {centre, half extents}, two
 * float4: the centre rounded to float, the half extents measured from that float centre in fp64 and rounded up;
 * .w of the second = 2^-12 (hx + hy + hz) rounded up (the device's margin term).  Levels are stored TOP FIRST, each padded
 * to ntop * 8^(nl - 1 - l) slots, so that a level's first node is a closed form of its number (tacc_level_base). */
struct TriAccelHost {
    std::vector<float> node, rec;
    std::vector<uint32_t> idx;
    uint32_t nc = 0, nl = 0, ntop = 0, off[RT_TACC_LEVELS] = {0};
};
static uint32_t tacc_level_base_host(uint32_t ntop, uint32_t m) { return ntop * (0x09249249u & ((1u << (3u * m)) - 1u)); }
static int build_tri_accel_host(const float* rec, uint32_t T, TriAccelHost& out) {
    const uint32_t K = RT_ACCEL_K;
    out.nc = out.nl = 0;
    if (T < 2u * K || !rec) return 0;
    std::vector<double> vx(3 * (size_t)T), vy(3 * (size_t)T), vz(3 * (size_t)T);
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
    std::vector<double> cen(3 * (size_t)T);
    for (uint32_t i = 0; i < T; i++) {
        const float* r = rec + 12 * (size_t)i;
        for (int k = 0; k < 12; k++)
            if ((k & 3) != 3 && !std::isfinite(r[k])) return 0; /* brute force */
        for (int v = 0; v < 3; v++) {
            const double ex = v == 0 ? 0.0 : r[4 * v + 0], ey = v == 0 ? 0.0 : r[4 * v + 1], ez = v == 0 ? 0.0 : r[4 * v + 2];
            vx[3 * (size_t)i + v] = (double)r[0] + ex; vy[3 * (size_t)i + v] = (double)r[1] + ey; vz[3 * (size_t)i + v] = (double)r[2] + ez;
        }
        const double c[3] = {(vx[3 * (size_t)i] + vx[3 * (size_t)i + 1] + vx[3 * (size_t)i + 2]) / 3.0,
                             (vy[3 * (size_t)i] + vy[3 * (size_t)i + 1] + vy[3 * (size_t)i + 2]) / 3.0,
                             (vz[3 * (size_t)i] + vz[3 * (size_t)i + 1] + vz[3 * (size_t)i + 2]) / 3.0};
        for (int k = 0; k < 3; k++) { cen[3 * (size_t)i + k] = c[k]; lo[k] = c[k] < lo[k] ? c[k] : lo[k]; hi[k] = c[k] > hi[k] ? c[k] : hi[k]; }
    }
    /* order: a kd-style sort by counts.  The implicit layout wants every node but the last of a level FULL, so a node of
     * capacity `cap` triangles hands its first cap/2 (along the longest axis of its centroids' bounds) to its lower
     * half and the rest to the upper half, down to clusters of K: contiguous positions, tight boxes, no space-filling-curve jumps */
    std::vector<std::pair<uint32_t, uint32_t>> order(T); /* .second = triangle (the .first is unused here) */
    for (uint32_t i = 0; i < T; i++) order[i] = {0u, i};
    uint32_t nc0 = (T + K - 1) / K, levels = 1;
    for (uint32_t n = nc0; levels < RT_TACC_LEVELS && n > 4u; n = (n + 7u) / 8u) levels++;
    size_t top_cap = K;
    for (uint32_t q = 1; q < levels; q++) top_cap *= 8u;
    struct Job { size_t b, e, cap; };
    std::vector<Job> jobs;
    for (size_t b = 0; b < T; b += top_cap) jobs.push_back({b, b + top_cap < T ? b + top_cap : (size_t)T, top_cap});
    while (!jobs.empty()) {
        const Job jb = jobs.back();
        jobs.pop_back();
        if (jb.cap <= K || jb.e - jb.b <= 1) continue;
        const size_t half = jb.cap / 2, mid = jb.b + half;
        if (jb.e > mid) {
            double blo[3] = {1e300, 1e300, 1e300}, bhi[3] = {-1e300, -1e300, -1e300};
            for (size_t q = jb.b; q < jb.e; q++)
                for (int a = 0; a < 3; a++) {
                    const double v = cen[3 * (size_t)order[q].second + a];
                    blo[a] = v < blo[a] ? v : blo[a]; bhi[a] = v > bhi[a] ? v : bhi[a];
                }
            int ax = 0;
            if (bhi[1] - blo[1] > bhi[ax] - blo[ax]) ax = 1;
            if (bhi[2] - blo[2] > bhi[ax] - blo[ax]) ax = 2;
            std::nth_element(order.begin() + jb.b, order.begin() + mid, order.begin() + jb.e,
                             [&](const std::pair<uint32_t, uint32_t>& x, const std::pair<uint32_t, uint32_t>& y) {
                                 const double cx_ = cen[3 * (size_t)x.second + ax], cy_ = cen[3 * (size_t)y.second + ax];
                                 return cx_ < cy_ || (cx_ == cy_ && x.second < y.second);
                             });
            jobs.push_back({mid, jb.e, half});
        }
        jobs.push_back({jb.b, jb.e < mid ? jb.e : mid, half});
    }
    (void)lo; (void)hi;
    const uint32_t nc = (T + K - 1) / K;
    uint32_t n_lvl[RT_TACC_LEVELS], nl = 1;
    n_lvl[0] = nc;
    while (nl < RT_TACC_LEVELS && n_lvl[nl - 1] > 4u) {
        n_lvl[nl] = (n_lvl[nl - 1] + 7u) / 8u;
        nl++;
    }
    const uint32_t ntop = n_lvl[nl - 1];
    for (uint32_t l = 0; l < nl; l++) out.off[l] = tacc_level_base_host(ntop, nl - 1u - l);
    const uint32_t total = tacc_level_base_host(ntop, nl); /* ntop * (8^nl - 1) / 7 slots; the ones past a level's last node are never read */
    out.node.assign(8 * (size_t)total, 0.0f);
    out.rec.assign(12 * (size_t)nc * K, 0.0f); /* padding: all zero -> det == 0, never hit */
    out.idx.assign((size_t)nc * K, 0xffffffffu);
    for (uint32_t j = 0; j < T; j++) {
        memcpy(&out.rec[12 * (size_t)j], rec + 12 * (size_t)order[j].second, 48);
        out.idx[j] = order[j].second;
    }
    for (uint32_t l = 0; l < nl; l++) {
        size_t span_t = K; /* triangles (sorted positions) per node of this level */
        for (uint32_t q = 0; q < l; q++) span_t *= 8u;
        for (uint32_t k = 0; k < n_lvl[l]; k++) {
            const size_t b = (size_t)k * span_t, end = b + span_t < T ? b + span_t : T;
            double blo[3] = {1e300, 1e300, 1e300}, bhi[3] = {-1e300, -1e300, -1e300};
            for (size_t j = b; j < end; j++)
                for (int v = 0; v < 3; v++) {
                    const size_t i = 3 * (size_t)order[j].second + v;
                    const double pt[3] = {vx[i], vy[i], vz[i]};
                    for (int a = 0; a < 3; a++) { blo[a] = pt[a] < blo[a] ? pt[a] : blo[a]; bhi[a] = pt[a] > bhi[a] ? pt[a] : bhi[a]; }
                }
            float* o = &out.node[8 * ((size_t)out.off[l] + k)];
            double hs = 0;
            for (int a = 0; a < 3; a++) {
                const float c = (float)(0.5 * (blo[a] + bhi[a]));
                const double up = bhi[a] - (double)c, dn = (double)c - blo[a];
                const double h = (up > dn ? up : dn) * (1.0 + 1e-6) + 1e-30;
                if (!(h < 1e37) || !std::isfinite(c)) return 0;
                o[a] = c;
                o[4 + a] = nextafterf((float)h, INFINITY);
                hs += o[4 + a];
            }
            o[3] = 0.0f;
            o[7] = nextafterf((float)(hs * (1.0 / 4096.0)), INFINITY);
        }
    }
    out.nc = nc;
    out.nl = nl;
    out.ntop = ntop;
    return 1;
}
static int prep_tri_accel(rtb_engine* e) {
    e->tacc_nc = 0;
    if (e->T < 2u * RT_ACCEL_K) return 0;
    std::vector<float> rec(12 * (size_t)e->T);
    CU(cudaMemcpyAsync(rec.data(), e->tri_rec.p, rec.size() * 4, cudaMemcpyDeviceToHost, e->stream), "D2H triangle records (accel build)");
    CU(cudaStreamSynchronize(e->stream), "triangle records");
    TriAccelHost h;
    if (build_tri_accel_host(rec.data(), e->T, h) != 1) return 0;
    const size_t b0 = h.node.size() * 4, b1 = h.rec.size() * 4, b2 = h.idx.size() * 4;
    if (e->tacc.ensure(b0 + b1 + b2)) return -2;
    CU(cudaMemcpyAsync(e->tacc.p, h.node.data(), b0, cudaMemcpyHostToDevice, e->stream), "H2D triangle accel");
    CU(cudaMemcpyAsync((char*)e->tacc.p + b0, h.rec.data(), b1, cudaMemcpyHostToDevice, e->stream), "H2D triangle accel");
    CU(cudaMemcpyAsync((char*)e->tacc.p + b0 + b1, h.idx.data(), b2, cudaMemcpyHostToDevice, e->stream), "H2D triangle accel");
    CU(cudaStreamSynchronize(e->stream), "triangle accel upload"); /* the vectors die with this frame */
    e->tacc_nc = h.nc; e->tacc_nl = h.nl; e->tacc_nodes = (uint32_t)(h.node.size() / 8);
    e->tacc_ntop = h.ntop;
    return 0;
}

static int prep_scene(rtb_engine* e) {
    if (int rc = prep_jump(e)) return rc;
    e->lay = rt_blob_layout(e->S, e->P, e->L);
    if (e->blob.ensure((size_t)e->lay.total * 16)) return -2;
    CU(rt_launch_scene_prep(e->d_spheres, e->S, e->d_planes, e->P, e->d_lights, e->L, (float4*)e->blob.p, e->lay,
                            e->stream),
       "scene_prep launch");
    e->have_scene = true;
    e->scene_gen++;
    return 0;
}

int rtb_upload_scene(rtb_engine* e, const void* spheres, uint32_t S, const void* planes, uint32_t P, const void* lights,
                     uint32_t L) {
    if (!e) return fail(-1, "null engine");
    if ((S && !spheres) || (P && !planes) || (L && !lights)) return fail(-1, "rtb_upload_scene: null array with non-zero count");
    CU(cudaSetDevice(e->device), "cudaSetDevice");
    CU(cudaStreamSynchronize(e->stream), "sync before scene upload");
    if (e->spheres.ensure((size_t)S * RT_SPHERE_BYTES) || e->planes.ensure((size_t)P * RT_PLANE_BYTES) ||
        e->lights.ensure((size_t)L * RT_LIGHT_BYTES))
        return -2;
    if (S) CU(cudaMemcpyAsync(e->spheres.p, spheres, (size_t)S * RT_SPHERE_BYTES, cudaMemcpyHostToDevice, e->stream), "H2D spheres");
    if (P) CU(cudaMemcpyAsync(e->planes.p, planes, (size_t)P * RT_PLANE_BYTES, cudaMemcpyHostToDevice, e->stream), "H2D planes");
    if (L) CU(cudaMemcpyAsync(e->lights.p, lights, (size_t)L * RT_LIGHT_BYTES, cudaMemcpyHostToDevice, e->stream), "H2D lights");
    e->d_spheres = (const unsigned char*)e->spheres.p;
    e->d_planes = (const unsigned char*)e->planes.p;
    e->d_lights = (const unsigned char*)e->lights.p;
    e->S = S; e->P = P; e->L = L;
    int rc = prep_scene(e);
    if (rc) return rc;
    CU(cudaStreamSynchronize(e->stream), "scene upload"); /* blocking, like clEnqueueWriteBuffer(CL_TRUE) */
    return prep_accel(e, (const unsigned char*)spheres);
}

int rtb_bind_scene_device(rtb_engine* e, const void* d_spheres, uint32_t S, const void* d_planes, uint32_t P,
                          const void* d_lights, uint32_t L) {
    if (!e) return fail(-1, "null engine");
    if ((S && !d_spheres) || (P && !d_planes) || (L && !d_lights)) return fail(-1, "rtb_bind_scene_device: null array");
    if (((uintptr_t)d_spheres | (uintptr_t)d_planes | (uintptr_t)d_lights) & 15) return fail(-1, "rtb_bind_scene_device: arrays must be 16-byte aligned");
    CU(cudaSetDevice(e->device), "cudaSetDevice");
    e->d_spheres = (const unsigned char*)d_spheres;
    e->d_planes = (const unsigned char*)d_planes;
    e->d_lights = (const unsigned char*)d_lights;
    e->S = S; e->P = P; e->L = L;
    if (int rc = prep_scene(e)) return rc;
    std::vector<unsigned char> host((size_t)S * RT_SPHERE_BYTES);
    if (S) CU(cudaMemcpy(host.data(), d_spheres, host.size(), cudaMemcpyDeviceToHost), "D2H spheres (accel build)");
    return prep_accel(e, S ? host.data() : nullptr);
}

static int prep_tris(rtb_engine* e) {
    if (e->T == 0) return 0;
    if (e->tri_rec.ensure((size_t)e->T * 48)) return -2;
    CU(rt_launch_tri_prep(e->d_tris, e->T, (float4*)e->tri_rec.p, e->stream), "tri_prep launch");
    return 0;
}
/* the hierarchy is built on first use (option "accel"), not at upload: 50k triangles cost tens of ms of host time */
static void tris_changed(rtb_engine* e) { e->tacc_valid = false; e->tacc_nc = 0; }
int rtb_upload_triangles(rtb_engine* e, const void* tris, uint32_t T) {
    if (!e) return fail(-1, "null engine");
    if (T && !tris) return fail(-1, "rtb_upload_triangles: null array with non-zero count");
    CU(cudaSetDevice(e->device), "cudaSetDevice");
    CU(cudaStreamSynchronize(e->stream), "sync before triangle upload");
    if (e->tris.ensure((size_t)T * RT_TRIANGLE_BYTES)) return -2;
    if (T) CU(cudaMemcpyAsync(e->tris.p, tris, (size_t)T * RT_TRIANGLE_BYTES, cudaMemcpyHostToDevice, e->stream), "H2D triangles");
    e->d_tris = (const unsigned char*)e->tris.p;
    e->T = T;
    tris_changed(e);
    int rc = prep_tris(e);
    if (rc) return rc;
    CU(cudaStreamSynchronize(e->stream), "triangle upload");
    return 0;
}
int rtb_bind_triangles_device(rtb_engine* e, const void* d_tris, uint32_t T) {
    if (!e) return fail(-1, "null engine");
    if (T && !d_tris) return fail(-1, "rtb_bind_triangles_device: null array");
    if ((uintptr_t)d_tris & 15) return fail(-1, "rtb_bind_triangles_device: array must be 16-byte aligned");
    CU(cudaSetDevice(e->device), "cudaSetDevice");
    e->d_tris = (const unsigned char*)d_tris;
    e->T = T;
    tris_changed(e);
    return prep_tris(e);
}

int rtb_upload_images(rtb_engine* e, int which, const uint8_t* rgba, uint32_t w, uint32_t h, uint32_t layers) {
    if (!e) return fail(-1, "null engine");
    if (which != 0 && which != 1) return fail(-1, "rtb_upload_images: which must be 0 (textures) or 1 (skybox)");
    if (!rgba || !w || !h || !layers) return fail(-1, "rtb_upload_images: empty image array");
    if (which == 1 && layers != 1) return fail(-1, "rtb_upload_images: the skybox is a 1-layer array");
    CU(cudaSetDevice(e->device), "cudaSetDevice");
    CU(cudaStreamSynchronize(e->stream), "sync before image upload");
    size_t n = (size_t)w * h * layers * 4;
    DevBuf& b = which ? e->sky : e->tex;
    if (b.ensure(n)) return -2;
    CU(cudaMemcpy(b.p, rgba, n, cudaMemcpyHostToDevice), "H2D images");
    if (which) { e->sw = w; e->sh = h; e->d_sky = b.p; }
    else { e->tw = w; e->th = h; e->tl = layers; e->d_tex = b.p; }
    return 0;
}

int rtb_bind_images_device(rtb_engine* e, int which, const void* d_rgba, uint32_t w, uint32_t h, uint32_t layers) {
    if (!e) return fail(-1, "null engine");
    if (which != 0 && which != 1) return fail(-1, "rtb_bind_images_device: which must be 0 (textures) or 1 (skybox)");
    if (!d_rgba || !w || !h || !layers) return fail(-1, "rtb_bind_images_device: empty image array");
    if (which == 1 && layers != 1) return fail(-1, "rtb_bind_images_device: the skybox is a 1-layer array");
    if ((uintptr_t)d_rgba & 3) return fail(-1, "rtb_bind_images_device: texels must be 4-byte aligned");
    if (which) { e->sw = w; e->sh = h; e->d_sky = d_rgba; }
    else { e->tw = w; e->th = h; e->tl = layers; e->d_tex = d_rgba; }
    return 0;
}

int rtb_set_camera(rtb_engine* e, const float cam[14]) {
    if (!e || !cam) return fail(-1, "rtb_set_camera: null argument");
    memcpy(e->cam, cam, sizeof e->cam);
    e->have_cam = true;
    return 0;
}

uint32_t rtb_partition_rows(uint32_t H, const rtb_partition* p) {
    if (!p || p->parts <= 1) return H;
    if (!p->band_rows || p->part >= p->parts) return 0;
    uint32_t bands = (H + p->band_rows - 1) / p->band_rows;
    uint32_t rows = 0;
    for (uint32_t b = p->part; b < bands; b += p->parts) {
        uint32_t y0 = b * p->band_rows, y1 = y0 + p->band_rows;
        rows += (y1 > H ? H : y1) - y0;
    }
    return rows;
}

static int render_impl(rtb_engine* e, uint32_t W, uint32_t H, uint32_t total, const rtb_partition* part, uint32_t* d_out,
                       uint32_t* d_hit, int want_hit, int mode, uint64_t* counters) {
    if (!e) return fail(-1, "null engine");
    if (!e->have_scene) return fail(-1, "rtb_render: no scene uploaded");
    if (!e->have_cam) return fail(-1, "rtb_render: no camera set");
    if (!e->d_sky) return fail(-1, "rtb_render: no skybox image uploaded (kernel arg 9)");
    if (!W || !H) return fail(-1, "rtb_render: empty frame");
    if ((uint64_t)W * H > 0xffffffffull) return fail(-1, "rtb_render: W*H exceeds the 32-bit pixel id (raygen.cl:10)");
    CU(cudaSetDevice(e->device), "cudaSetDevice");

    rtb_partition whole = {H, 1, 0};
    rtb_partition pp = part ? *part : whole;
    if (pp.parts <= 1) { pp.parts = 1; pp.part = 0; pp.band_rows = H; }
    if (!pp.band_rows || pp.part >= pp.parts) return fail(-1, "rtb_render: bad partition");
    const uint32_t rows = rtb_partition_rows(H, &pp);

    RtFrameParams fp;
    memset(&fp, 0, sizeof fp);
    memcpy(fp.corner, e->cam + 0, 12); memcpy(fp.origin, e->cam + 3, 12);
    memcpy(fp.up, e->cam + 6, 12); memcpy(fp.right, e->cam + 9, 12);
    fp.wf = e->cam[12]; fp.hf = e->cam[13];
    fp.W = W; fp.H = H; fp.total = total;
    fp.band_rows = pp.band_rows; fp.parts = pp.parts; fp.part = pp.part; fp.local_rows = rows;
    fp.S = e->S; fp.P = e->P; fp.L = e->L; fp.lay = e->lay;
    fp.blob = (const float4*)e->blob.p;
    fp.raw_spheres = e->d_spheres; fp.raw_planes = e->d_planes;
    fp.T = e->T; fp.tri = (const float4*)e->tri_rec.p; fp.raw_tris = e->d_tris;
    /* Extension X1, large meshes: EVERY tile goes through the CTA-wide tree kernel, whose lock step lets the triangle records be
     * staged into shared memory in chunks by TMA and shared by 128 rays (rt_trace_tree.cuh: tri_walk_staged) */
    /* N4 over the triangles: built on first use; scenes it cannot take (too few, non-finite, huge) stay on the brute-force build */
    if (e->opt_accel > 0 && mode == 0 && e->T != 0u && !e->tacc_valid) {
        if (int rc = prep_tri_accel(e)) return rc;
        e->tacc_valid = true;
    }
    const bool use_tacc = e->opt_accel > 0 && mode == 0 && e->T != 0u && e->tacc_nc > 0u;
    const bool tri_all = !use_tacc && mode == 0 && e->opt_tree == 1 && e->L <= 16u && e->opt_tri_tree_min > 0 && e->T >= (uint32_t)e->opt_tri_tree_min;
    /* N4: the cluster walk instead of the record walk -- an extension, off unless asked for ("accel"), production trace only */
    const bool use_accel = e->opt_accel > 0 && mode == 0 && (e->T == 0u ? e->acc_nc > 0u : use_tacc);
    if (use_tacc) {
        const size_t b0 = (size_t)e->tacc_nodes * 32, b1 = (size_t)e->tacc_nc * RT_ACCEL_K * 48;
        fp.tacc_node = (const float4*)e->tacc.p;
        fp.tacc_rec = (const float4*)((const char*)e->tacc.p + b0);
        fp.tacc_idx = (const uint32_t*)((const char*)e->tacc.p + b0 + b1);
        fp.tacc_nc = e->tacc_nc; fp.tacc_nl = e->tacc_nl;
        fp.tacc_ntop = e->tacc_ntop;
    }
    if (use_accel && e->acc_nc > 0u) {
        const size_t b0 = (size_t)e->acc_nc * 16, b1 = (size_t)e->acc_nc * RT_ACCEL_K * 16;
        fp.acc_cl = (const float4*)e->accel.p;
        fp.acc_rec = (const float4*)((const char*)e->accel.p + b0);
        fp.acc_idx = (const uint32_t*)((const char*)e->accel.p + b0 + b1);
        fp.acc_grp = (const float4*)((const char*)e->accel.p + b0 + b1 + ((size_t)e->acc_nc * RT_ACCEL_K * 4 + 15) / 16 * 16);
        fp.acc_nc = e->acc_nc;
    }
    /* a scene whose planes carry texture ids needs the texture array; without one, ids clamp onto a 1x1 dummy */
    fp.tex = (const uchar4*)(e->d_tex ? e->d_tex : e->d_sky);
    fp.tw = e->d_tex ? (int)e->tw : 1; fp.th = e->d_tex ? (int)e->th : 1; fp.tl = e->d_tex ? (int)e->tl : 1;
    fp.sky = (const uchar4*)e->d_sky; fp.sw = (int)e->sw; fp.sh = (int)e->sh;

    const size_t nbytes = (size_t)rows * W * 4;
    if (!d_out) {
        if (e->frame.ensure(nbytes)) return -2;
        d_out = (uint32_t*)e->frame.p;
    }
    if (!d_hit && want_hit) {
        if (e->hit.ensure(nbytes)) return -2;
        d_hit = (uint32_t*)e->hit.p;
    }
    if (((uintptr_t)d_out & 15) || ((uintptr_t)d_hit & 15)) return fail(-1, "rtb_render: output buffers must be 16-byte aligned");
    fp.out = d_out;
    fp.hit = d_hit;
    fp.mirror = e->mirror;
    fp.mirror_global = e->mirror_global ? 1u : 0u;
    e->mirror = nullptr;
    e->mirror_global = false;
    if ((uintptr_t)fp.mirror & 15) return fail(-1, "rtb_render: mirror buffer must be 16-byte aligned");
    auto bind_sched = [&](int set) { /* cursors, counters: one 256-byte half of `sched` per set of queue buffers */
        char* b = (char*)e->sched.p + 256 * set;
        fp.tile_counter = (unsigned int*)b;
        fp.tree_counter = (unsigned int*)(b + 4);
        fp.done_counter = (unsigned int*)(b + 32);
        fp.counters = (unsigned long long*)(b + 64);
        e->last_sched_off = 256u * (uint32_t)set;
    };
    bind_sched(0);
    fp.done_flag = e->done_flag;
    fp.done_value = e->done_value;
    e->done_flag = nullptr;
    fp.tiles_x = (W + 7) / 8;
    fp.tiles_y = (rows + 3) / 4;

    e->last.tiles = fp.tiles_x * fp.tiles_y;
    if (rows == 0) { e->timed = false; return 0; }

    if (e->look_pending) { /* whatever this frame does with the queue buffers comes after the look-ahead that may be writing them */
        CU(cudaStreamWaitEvent(e->stream, e->ev_look, 0), "wait for the schedule look-ahead");
        e->look_pending = false;
    }
    CU(cudaEventRecord(e->ev0, e->stream), "event record");
    const bool ordered = e->opt_tile_order && e->last.tiles >= 4096 && e->last.tiles < (1u << 24);
    if (!ordered) {
        CU(cudaMemsetAsync(e->sched.p, 0, 256, e->stream), "reset tile counter"); /* else tile_order zeroes it */
        e->look_valid = false;
    }
    bool history_frame = false, same_frame = false;
    int set = 0, launcher = e->T ? (use_tacc ? 3 : 2) : (use_accel ? 1 : 0);
    auto tile_order_launcher = [&](int which) {
        return which == 3 ? rt_launch_tile_order_tacc : (which == 2 ? rt_launch_tile_order_tri : (which == 1 ? rt_launch_tile_order_acc : rt_launch_tile_order));
    };
    fp.order = nullptr;
    fp.cost = nullptr;
    fp.coop_min_cost = (uint32_t)e->opt_coop_min_cost;
    fp.wide_min_records = (uint32_t)e->opt_wide_min_records;
    fp.wide_max_live = e->opt_wide_max_live;
    fp.split_min_cost = 64u;
    fp.split32_min_cost = 64u;
    fp.tree_min_cost = 64u;
    /* queue entries carry the tile index in 24 bits (rt_kernels.cu): beyond 2^24 tiles (> 536 Mpixel) the queue stays row-major */
    if (e->opt_tile_order && e->last.tiles >= 4096 && e->last.tiles < (1u << 24)) { /* small frames: the queue is short enough as it is */
        const size_t nt = e->last.tiles;
        for (int k = 0; k < 2; k++) {
            const bool fresh_hist = e->tile_hist[k].p == nullptr;
            if (e->tile_cost[k].ensure(nt * 4) || e->tile_order[k].ensure(nt * 128) || e->tile_hist[k].ensure(1024)) return -2;
            if (fresh_hist) CU(cudaMemsetAsync(e->tile_hist[k].p, 0, 1024, e->stream), "zero tile histogram"); /* once: tile_order leaves it clean */
        }
        /* scheduling hint: the previous frame's measured tile times when it had this geometry and scene (the camera may
         * have moved: frames of an interactive sweep are coherent), else the skeleton trace.  Pixels never depend on it. */
        const bool same = e->time_valid && e->tk_W == W && e->tk_H == H && e->tk_total == total && e->tk_band == pp.band_rows &&
                          e->tk_parts == pp.parts && e->tk_part == pp.part && e->tk_scene_gen == e->scene_gen + e->T;
        if (e->opt_history && mode == 0) {
            /* each buffer: 16 bytes for the frame's sum, then one word per tile */
            if (e->tile_time[0].ensure(nt * 4 + 16) || e->tile_time[1].ensure(nt * 4 + 16) || e->tile_time[2].ensure(nt * 4 + 16)) return -2;
            history_frame = true;
            same_frame = same;
            fp.tile_time_sum = (unsigned long long*)e->tile_time[e->time_cur].p;
            fp.tile_time = (uint32_t*)((char*)e->tile_time[e->time_cur].p + 16);
            if (same) {
                const int tp = (e->time_cur + 2) % 3; /* the previous frame's */
                fp.tile_time_sum_prev = (const unsigned long long*)e->tile_time[tp].p;
                fp.tile_time_prev = (const uint32_t*)((const char*)e->tile_time[tp].p + 16);
                /* measured costs: a tile is a tail risk when its time is comparable to the time the whole queue needs,
                 * T = mean * tiles / resident warps.  Automatic thresholds (they reproduce the optima of the sweeps in
                 * tools/exp/sweep_split.py at N = 1, 2, 4, 8): four warps from 0.75 T, one pixel per warp from 3 T.
                 * On the relative cost scale (24 + 4*log2(t / mean)) that is 24 + 4*log2(0.75 * tiles / warps) etc. */
                if (mode == 0) {
                    const double per_warp = (double)nt / (double)(e->sm_count * 20);
                    auto thr = [&](double alpha, double floor_cost) { /* never below 16x (64x) the mean tile: splitting costs occupancy */
                        double c = 24.0 + 4.0 * log2(alpha * per_warp) + 0.5;
                        return (uint32_t)(c < floor_cost ? floor_cost : (c > 64.0 ? 64.0 : c));
                    };
                    /* the pixel-tree kernel: not for many-sphere scenes (the WIDE variant's territory: a tree node walks all
                     * records on ONE lane; measured 50 -> 115 ms per N = 8 part on the 1,024-sphere grid) */
                    const bool wide_scene = e->T == 0u && !use_accel && e->lay.Lpad + e->lay.Spad >= (uint32_t)e->opt_wide_min_records;
                    /* nor for N4 over triangles: a walk is a chain of dependent loads and the CTA-wide trees only lose time
                     * on it (measured on the 49,928-triangle mesh at 4K: 40 ms -> 61..74 ms); "tree" >= 2 still forces them (tests) */
                    const bool tree = e->opt_tree > 0 && e->L <= 16u && (!wide_scene || e->opt_tree >= 2) && (!use_tacc || e->opt_tree >= 2);
                    if (tree) {
                        /* with the CTA-wide pixel-tree kernel beside render_frame (tools/exp/sweep_tree.py): tiles longer than
                         * 0.75 T go to it whole (4 warps on the tile's pooled segment trees), tiles longer than 3 T as four
                         * row entries (16 warps); single-pixel entries no longer pay */
                        /* (N4 scenes: a record walk is cheap there, so what is left of a frame is the segment chains of
                         * its glass pixels -- up to 2^15 segments each: trees much earlier, rows and single pixels too) */
                        fp.split_min_cost = e->opt_split_min_cost > 0 ? (uint32_t)e->opt_split_min_cost : (use_accel ? thr(0.75, 40.0) : thr(3.0, 44.0));
                        fp.split32_min_cost = e->opt_split32_min_cost > 0 ? (uint32_t)e->opt_split32_min_cost : (use_accel ? thr(3.0, 48.0) : 64u);
                        /* (tools/exp/sweep_tree_fine.py, slowest part's kernels: N=4 0.785 -> 0.764 ms with trees from 0.55 T
                         * instead of 0.75 T, never below cost 36; halves and whole frames keep 0.75 T, where the tree
                         * kernel only costs: N=2 1.36 ms either way from cost 44, N=1 2.66 -> 3.1 ms from cost 44) */
                        uint32_t t = e->opt_tree_min_cost > 0 ? (uint32_t)e->opt_tree_min_cost
                                     : (use_accel ? thr(0.2, 32.0) : (pp.parts >= 3 ? thr(0.55, 36.0) : thr(0.75, 36.0)));
                        if (fp.split_min_cost < t) t = fp.split_min_cost;
                        if (fp.split32_min_cost < t) t = fp.split32_min_cost;
                        fp.tree_min_cost = t;
                        fp.tree_group_max_cost = e->opt_tree_group_max_cost > 0 ? (uint32_t)e->opt_tree_group_max_cost : 0u; /* pooling four entries per pass lengthens the chains it is there to shorten (tools/exp/sweep_tree.py): off unless asked for */
                        fp.cost_is_history = e->cost_hist_set >= 0 ? 1u : 0u;
                    } else if (wide_scene) {
                        /* many-sphere scenes: a split entry's few rays are walked by all 32 lanes (WIDE variant), so
                         * splitting costs little and the tail of a part is its longest tile: rows from 0.5 T, single
                         * pixels from T (1,024-sphere grid, slowest of 8 parts 50.8 -> 31-34 ms, of 4 parts 69 -> 54 ms) */
                        /* (a whole frame on one GPU keeps the r01 thresholds: its tail is 1/8 of the matter, 183.7 vs 190 ms) */
                        const bool whole = pp.parts <= 1;
                        fp.split_min_cost = e->opt_split_min_cost > 0 ? (uint32_t)e->opt_split_min_cost : (whole ? thr(0.75, 40.0) : thr(0.5, 32.0));
                        fp.split32_min_cost = e->opt_split32_min_cost > 0 ? (uint32_t)e->opt_split32_min_cost : (whole ? thr(3.0, 48.0) : thr(1.0, 36.0));
                    } else {
                        fp.split_min_cost = e->opt_split_min_cost > 0 ? (uint32_t)e->opt_split_min_cost : thr(0.75, 40.0);
                        fp.split32_min_cost = e->opt_split32_min_cost > 0 ? (uint32_t)e->opt_split32_min_cost : thr(3.0, 48.0);
                    }
                }
            }
        } else if (mode == 0) {
            e->time_valid = false;
            e->cost_hist_set = -1;
        } /* accounting / A-B launches (mode != 0) are scheduled by the skeleton trace and leave the history alone */
        if (tri_all) { fp.split_min_cost = 64u; fp.split32_min_cost = 64u; fp.tree_min_cost = 64u; } /* whole tiles, all of them through render_tree */
        /* the queue: already built by the look-ahead of the previous frame (same geometry, scene, thresholds, kernel
         * family; it also cleared this frame's time buffer and its half of `sched`), or built here and now */
        const bool looked = e->look_valid && history_frame && same_frame && e->look_time == e->time_cur && e->look_launcher == launcher &&
                            e->look_W == W && e->look_H == H && e->look_total == total && e->look_band == pp.band_rows &&
                            e->look_parts == pp.parts && e->look_part == pp.part && e->look_scene == e->scene_gen + e->T &&
                            e->look_split == fp.split_min_cost && e->look_split32 == fp.split32_min_cost &&
                            e->look_tree == fp.tree_min_cost && e->look_group == fp.tree_group_max_cost && e->look_cih == fp.cost_is_history;
        e->look_valid = false;
        if (looked) {
            set = e->look_set;
            bind_sched(set);
        } else {
            set = 1 - e->set_cur;
            bind_sched(set);
            if (history_frame && !same_frame) CU(cudaMemsetAsync(e->tile_time[e->time_cur].p, 0, nt * 4 + 16, e->stream), "reset tile times"); /* history frames: tile_cost_history zeroes it */
            fp.cost_prev = e->cost_hist_set >= 0 ? (const uint32_t*)e->tile_cost[e->cost_hist_set].p : nullptr;
            if (!fp.cost_prev) fp.cost_is_history = 0u;
            CU(tile_order_launcher(launcher)(fp, (uint32_t*)e->tile_cost[set].p, (uint32_t*)e->tile_hist[set].p, (uint32_t*)e->tile_order[set].p, e->stream),
               "tile_order launch");
            if (!(history_frame && same_frame) && e->cost_hist_set == set) e->cost_hist_set = -1; /* overwritten by skeleton costs */
        }
        e->set_cur = set;
        if (history_frame) {
            if (same_frame) e->cost_hist_set = set; /* this frame's costs come from measured times: the next one may lean on them */
            else e->cost_hist_set = -1;
            e->time_cur = (e->time_cur + 1) % 3;
            e->time_valid = true;
            e->tk_W = W; e->tk_H = H; e->tk_total = total; e->tk_band = pp.band_rows; e->tk_parts = pp.parts; e->tk_part = pp.part;
            e->tk_scene_gen = e->scene_gen + e->T;
        }
        fp.order = (const uint32_t*)e->tile_order[set].p;
        fp.cost = (const uint32_t*)e->tile_cost[set].p;
        fp.queue_len = (const uint32_t*)e->tile_hist[set].p + 128;
        fp.n_split = (const uint32_t*)e->tile_hist[set].p + 129;
    }
    /* level-wise pixel trees for the heavy queue entries (rt_trace_tree.cuh): production trace only, and only while a
     * node's 1 + 2L addition slots stay small.  One slab per CTA of the persistent grid. */
    fp.tree_mode = 0;
    if (mode == 0 && e->opt_tree > 0 && e->L <= 16u && ((fp.order && fp.tree_min_cost < 64u) || e->opt_tree >= 2 || tri_all)) {
        const uint32_t cap = use_accel && e->opt_tree_cap == 8192 ? 32768u : (uint32_t)e->opt_tree_cap; /* N4 scenes: whole glass pixels */
        const uint32_t slots = (uint32_t)e->sm_count * (uint32_t)(tri_all && e->opt_tree_ctas_per_sm == 3 ? 4 : e->opt_tree_ctas_per_sm); /* = the render_tree grid: one slab per CTA */
        const size_t records = rt_tree_slab_records(cap, e->L);
        if (e->tree_slab.ensure(records * 16 * slots) == 0) {
            fp.tree_slab = (float4*)e->tree_slab.p;
            fp.tree_records = records;
            fp.tree_cap = cap;
            fp.tree_slots = slots;
            fp.tree_mode = tri_all ? 2u : (uint32_t)e->opt_tree;
            if (tri_all) fp.tree_group_max_cost = 64u; /* four tiles per CTA pass: 128 rays per level fill the 128 threads */
            fp.jump = (const uint32_t*)e->jump.p;
        } else {
            (void)cudaGetLastError(); /* no room for the slabs: the cooperative walk handles everything */
        }
    }
    RtLaunchShape shape{};
    CU(cudaEventRecord(e->evk, e->stream), "event record");
    CU((e->T ? (use_tacc ? rt_launch_render_tacc : rt_launch_render_tri) : (use_accel ? rt_launch_render_acc : rt_launch_render))(fp, mode, e->sm_count, e->stream, &shape), "render_frame launch");
    CU(cudaEventRecord(e->ev1, e->stream), "event record");
    /* schedule look-ahead: the NEXT frame's queue, from the PREVIOUS frame's tile times (this frame's are being measured),
     * on the second stream while this frame renders: tile_cost_history + tile_order leave the critical path (they were
     * ~25 us of a 0.48 ms frame on one of 8 GPUs).  The hints are one frame staler; pixels never depend on them. */
    if (e->opt_lookahead && ordered && history_frame && same_frame && mode == 0 && !counters && !tri_all && fp.tile_time_prev) {
        RtFrameParams fq = fp;
        const int nset = 1 - set, tn = e->time_cur; /* the other set (last used by the frame before this one); the buffer the next frame writes */
        fq.tile_time_sum = (unsigned long long*)e->tile_time[tn].p;
        fq.tile_time = (uint32_t*)((char*)e->tile_time[tn].p + 16);
        char* b = (char*)e->sched.p + 256 * nset;
        fq.tile_counter = (unsigned int*)b;
        fq.cost_prev = (const uint32_t*)e->tile_cost[set].p;
        fq.cost_is_history = fp.cost_is_history; /* 1 in the pixel-tree regime (hysteresis on), else 0: what the next frame will ask for */
        CU(cudaStreamWaitEvent(e->stream2, e->evk, 0), "look-ahead: after the frames before this one"); /* evk: recorded before this frame's kernels */
        CU(tile_order_launcher(launcher)(fq, (uint32_t*)e->tile_cost[nset].p, (uint32_t*)e->tile_hist[nset].p, (uint32_t*)e->tile_order[nset].p, e->stream2),
           "tile_order launch (look-ahead)");
        CU(cudaEventRecord(e->ev_look, e->stream2), "event record");
        e->look_pending = true;
        e->look_valid = true;
        e->look_set = nset; e->look_time = tn; e->look_launcher = launcher;
        e->look_W = W; e->look_H = H; e->look_total = total; e->look_band = pp.band_rows; e->look_parts = pp.parts; e->look_part = pp.part;
        e->look_scene = e->scene_gen + e->T;
        e->look_cih = fq.cost_is_history;
        e->look_split = fp.split_min_cost; e->look_split32 = fp.split32_min_cost; e->look_tree = fp.tree_min_cost; e->look_group = fp.tree_group_max_cost;
    }
    e->timed = true;
    e->last.grid = shape.grid; e->last.block = shape.block; e->last.smem_bytes = shape.smem_bytes;
    e->last.resident_scene = shape.resident;
    e->last.tree_grid = shape.tree_grid;
    if (counters) {
        unsigned long long host[RTB_NCOUNTERS];
        CU(cudaMemcpyAsync(host, fp.counters, sizeof host, cudaMemcpyDeviceToHost, e->stream), "D2H counters");
        CU(cudaStreamSynchronize(e->stream), "render_frame (counted)");
        for (int k = 0; k < RTB_NCOUNTERS; k++) counters[k] = host[k];
    }
    return 0;
}

int rtb_render(rtb_engine* e, uint32_t W, uint32_t H, uint32_t total, const rtb_partition* part, uint32_t* d_out,
               uint32_t* d_hit, int want_hit) {
    if (!e) return fail(-1, "null engine");
    return render_impl(e, W, H, total, part, d_out, d_hit, want_hit, e->opt_trace == 1 ? 1 : (e->opt_trace == 2 ? 3 : 0), nullptr);
}
int rtb_render_counted(rtb_engine* e, uint32_t W, uint32_t H, uint32_t total, const rtb_partition* part, uint32_t* d_out,
                       uint32_t* d_hit, int want_hit, int executed, uint64_t* counters) {
    if (!counters) return fail(-1, "rtb_render_counted: null counters");
    return render_impl(e, W, H, total, part, d_out, d_hit, want_hit, executed ? 2 : 1, counters);
}

int rtb_assemble_bands(rtb_engine* e, const uint32_t* d_compact, size_t stride_words, uint32_t* d_frame, uint32_t W,
                       uint32_t H, uint32_t band_rows, uint32_t parts) {
    if (!e || !d_compact || !d_frame) return fail(-1, "rtb_assemble_bands: null argument");
    if (!W || !H || !band_rows || !parts) return fail(-1, "rtb_assemble_bands: bad geometry");
    rtb_partition p0 = {band_rows, parts, 0};
    if ((size_t)rtb_partition_rows(H, &p0) * W > stride_words) return fail(-1, "rtb_assemble_bands: stride smaller than a part");
    CU(cudaSetDevice(e->device), "cudaSetDevice");
    CU(rt_launch_assemble_bands(d_compact, stride_words, d_frame, W, H, band_rows, parts, e->stream), "assemble_bands launch");
    return 0;
}

int rtb_set_option(rtb_engine* e, const char* name, int value) {
    if (!e || !name) return fail(-1, "rtb_set_option: null argument");
    if (!strcmp(name, "trace")) { e->opt_trace = (value == 1 || value == 2) ? value : 0; return 0; }
    if (!strcmp(name, "history")) { e->opt_history = value ? 1 : 0; e->time_valid = false; return 0; }
    if (!strcmp(name, "split32_min_cost")) { e->opt_split32_min_cost = value < 0 ? 0 : value; return 0; }
    if (!strcmp(name, "split_min_cost")) { e->opt_split_min_cost = value < 0 ? 0 : value; return 0; }
    if (!strcmp(name, "coop_min_cost")) { e->opt_coop_min_cost = value < 0 ? 0 : value; return 0; }
    if (!strcmp(name, "wide_min_records")) { e->opt_wide_min_records = value < 0 ? 0x7fffffff : value; return 0; }
    if (!strcmp(name, "wide_max_live")) { e->opt_wide_max_live = value > 32 ? 32 : value; return 0; }
    if (!strcmp(name, "tile_order")) { e->opt_tile_order = value ? 1 : 0; return 0; }
    if (!strcmp(name, "lookahead")) { e->opt_lookahead = value ? 1 : 0; return 0; }
    if (!strcmp(name, "tree")) { e->opt_tree = value < 0 ? 0 : (value > 3 ? 3 : value); return 0; }
    if (!strcmp(name, "accel")) { e->opt_accel = value ? 1 : 0; e->time_valid = false; return 0; }
    if (!strcmp(name, "tri_tree_min")) { e->opt_tri_tree_min = value < 0 ? 0 : value; return 0; }
    if (!strcmp(name, "tree_ctas_per_sm")) { e->opt_tree_ctas_per_sm = value < 1 ? 1 : (value > 8 ? 8 : value); e->tree_slab.release(); return 0; }
    if (!strcmp(name, "tree_group_max_cost")) { e->opt_tree_group_max_cost = value < 0 ? 0 : value; return 0; }
    if (!strcmp(name, "tree_min_cost")) { e->opt_tree_min_cost = value < 0 ? 0 : value; return 0; }
    if (!strcmp(name, "tree_cap")) { e->opt_tree_cap = value < 64 ? 64 : (value > 32768 ? 32768 : value); e->tree_slab.release(); return 0; }
    return fail(-1, "rtb_set_option: unknown option");
}

int rtb_selftest_arith(rtb_engine* e, uint32_t seed, uint32_t per_thread, uint32_t out[3]) {
    if (!e || !out) return fail(-1, "rtb_selftest_arith: null argument");
    CU(cudaSetDevice(e->device), "cudaSetDevice");
    uint32_t* d = (uint32_t*)((char*)e->sched.p + 128);
    CU(cudaMemsetAsync(d, 0, 12, e->stream), "memset");
    CU(rt_launch_arith_selftest(seed, per_thread, d, e->stream), "arith_selftest launch");
    CU(cudaMemcpyAsync(out, d, 12, cudaMemcpyDeviceToHost, e->stream), "D2H");
    CU(cudaStreamSynchronize(e->stream), "arith_selftest");
    return 0;
}

int rtb_sync(rtb_engine* e) {
    if (!e) return fail(-1, "null engine");
    CU(cudaSetDevice(e->device), "cudaSetDevice");
    CU(cudaStreamSynchronize(e->stream), "stream synchronize (a kernel may have faulted)");
    return 0;
}

static int read_buf(rtb_engine* e, DevBuf& b, void* host, size_t nbytes, const char* what) {
    if (!e || !host) return fail(-1, "null argument");
    if (!b.p || nbytes > b.cap) return fail(-1, what);
    CU(cudaSetDevice(e->device), "cudaSetDevice");
    CU(cudaMemcpyAsync(host, b.p, nbytes, cudaMemcpyDeviceToHost, e->stream), "D2H");
    CU(cudaStreamSynchronize(e->stream), "D2H sync");
    return 0;
}
int rtb_read_frame(rtb_engine* e, void* host, size_t nbytes) { return read_buf(e, e->frame, host, nbytes, "rtb_read_frame: no frame of that size"); }
int rtb_read_hit(rtb_engine* e, void* host, size_t nbytes) { return read_buf(e, e->hit, host, nbytes, "rtb_read_hit: no hit buffer of that size"); }
uint32_t* rtb_frame_ptr(rtb_engine* e) { return e ? (uint32_t*)e->frame.p : nullptr; }
uint32_t* rtb_hit_ptr(rtb_engine* e) { return e ? (uint32_t*)e->hit.p : nullptr; }

float rtb_last_render_ms(rtb_engine* e) {
    if (!e || !e->timed) return -1.0f;
    if (cudaSetDevice(e->device) != cudaSuccess) return -1.0f;
    if (cudaEventSynchronize(e->ev1) != cudaSuccess) return -1.0f;
    float ms = -1.0f;
    if (cudaEventElapsedTime(&ms, e->ev0, e->ev1) != cudaSuccess) return -1.0f;
    return ms;
}
float rtb_last_kernel_ms(rtb_engine* e) {
    if (!e || !e->timed) return -1.0f;
    if (cudaSetDevice(e->device) != cudaSuccess) return -1.0f;
    if (cudaEventSynchronize(e->ev1) != cudaSuccess) return -1.0f;
    float ms = -1.0f;
    if (cudaEventElapsedTime(&ms, e->evk, e->ev1) != cudaSuccess) return -1.0f;
    return ms;
}
void* rtb_stream(rtb_engine* e) { return e ? (void*)e->stream : nullptr; }

int rtb_dev_alloc(rtb_engine* e, void** dptr, size_t nbytes) {
    if (!e || !dptr) return fail(-1, "null argument");
    CU(cudaSetDevice(e->device), "cudaSetDevice");
    CU(cudaMalloc(dptr, nbytes ? nbytes : 16), "cudaMalloc");
    return 0;
}
int rtb_dev_free(rtb_engine* e, void* dptr) {
    if (!e) return fail(-1, "null engine");
    CU(cudaSetDevice(e->device), "cudaSetDevice");
    CU(cudaFree(dptr), "cudaFree");
    return 0;
}
int rtb_dev_write(rtb_engine* e, void* dptr, const void* host, size_t nbytes) {
    if (!e || !dptr || !host) return fail(-1, "null argument");
    CU(cudaSetDevice(e->device), "cudaSetDevice");
    CU(cudaMemcpyAsync(dptr, host, nbytes, cudaMemcpyHostToDevice, e->stream), "H2D");
    CU(cudaStreamSynchronize(e->stream), "H2D sync");
    return 0;
}
int rtb_dev_read(rtb_engine* e, void* host, const void* dptr, size_t nbytes) {
    if (!e || !dptr || !host) return fail(-1, "null argument");
    CU(cudaSetDevice(e->device), "cudaSetDevice");
    CU(cudaMemcpyAsync(host, dptr, nbytes, cudaMemcpyDeviceToHost, e->stream), "D2H");
    CU(cudaStreamSynchronize(e->stream), "D2H sync");
    return 0;
}
int rtb_set_mirror(rtb_engine* e, uint32_t* mirror) {
    if (!e) return fail(-1, "null engine");
    e->mirror = mirror;
    e->mirror_global = false;
    return 0;
}
int rtb_set_mirror_frame(rtb_engine* e, uint32_t* frame) {
    if (!e) return fail(-1, "null engine");
    e->mirror = frame;
    e->mirror_global = true;
    return 0;
}
int rtb_set_completion(rtb_engine* e, uint32_t* flag, uint32_t value) {
    if (!e) return fail(-1, "null engine");
    if ((uintptr_t)flag & 3) return fail(-1, "rtb_set_completion: flag must be 4-byte aligned");
    e->done_flag = flag;
    e->done_value = value;
    return 0;
}
int rtb_wait_flags(rtb_engine* e, const uint32_t* d_flags, uint32_t stride_words, uint32_t n, uint32_t value) {
    if (!e || !d_flags) return fail(-1, "rtb_wait_flags: null argument");
    if (n > 1024) return fail(-1, "rtb_wait_flags: at most 1024 flags");
    CU(cudaSetDevice(e->device), "cudaSetDevice");
    if (!e->wait_word.p) {
        if (e->wait_word.ensure(16)) return -2;
        CU(cudaMemsetAsync(e->wait_word.p, 0, 16, e->stream), "cudaMemsetAsync");
    }
    CU(rt_launch_wait_flags(d_flags, stride_words, n, value, (uint32_t*)e->wait_word.p, e->stream), "wait_flags launch");
    return 0;
}
int rtb_wait_timeouts(rtb_engine* e, uint32_t* out) {
    if (!e || !out) return fail(-1, "rtb_wait_timeouts: null argument");
    *out = 0;
    if (!e->wait_word.p) return 0;
    CU(cudaSetDevice(e->device), "cudaSetDevice");
    CU(cudaMemcpyAsync(out, e->wait_word.p, 4, cudaMemcpyDeviceToHost, e->stream), "D2H");
    CU(cudaStreamSynchronize(e->stream), "rtb_wait_timeouts");
    return 0;
}
int rtb_ipc_export(rtb_engine* e, const void* dptr, unsigned char handle[64]) {
    if (!e || !dptr || !handle) return fail(-1, "rtb_ipc_export: null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t");
    CU(cudaSetDevice(e->device), "cudaSetDevice");
    cudaIpcMemHandle_t h;
    CU(cudaIpcGetMemHandle(&h, const_cast<void*>(dptr)), "cudaIpcGetMemHandle");
    memcpy(handle, &h, 64);
    return 0;
}
int rtb_ipc_open(rtb_engine* e, const unsigned char handle[64], void** dptr) {
    if (!e || !dptr || !handle) return fail(-1, "rtb_ipc_open: null argument");
    CU(cudaSetDevice(e->device), "cudaSetDevice");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    CU(cudaIpcOpenMemHandle(dptr, h, cudaIpcMemLazyEnablePeerAccess), "cudaIpcOpenMemHandle");
    return 0;
}
int rtb_ipc_close(rtb_engine* e, void* dptr) {
    if (!e || !dptr) return fail(-1, "rtb_ipc_close: null argument");
    CU(cudaSetDevice(e->device), "cudaSetDevice");
    CU(cudaIpcCloseMemHandle(dptr), "cudaIpcCloseMemHandle");
    return 0;
}
int rtb_host_map(void* host, size_t nbytes, void** dev_view, int* registered_here) {
    if (!host || !dev_view) return fail(-1, "rtb_host_map: null argument");
    if (registered_here) *registered_here = 0;
    void* d = nullptr;
    if (cudaHostGetDevicePointer(&d, host, 0) != cudaSuccess) { /* not page-locked yet */
        (void)cudaGetLastError();
        cudaError_t ce = cudaHostRegister(host, nbytes, cudaHostRegisterMapped | cudaHostRegisterPortable);
        if (ce != cudaSuccess) { (void)cudaGetLastError(); return fail(-2, "cudaHostRegister", ce); }
        if (registered_here) *registered_here = 1;
        ce = cudaHostGetDevicePointer(&d, host, 0);
        if (ce != cudaSuccess) { (void)cudaGetLastError(); cudaHostUnregister(host); if (registered_here) *registered_here = 0; return fail(-2, "cudaHostGetDevicePointer", ce); }
    }
    *dev_view = d;
    return 0;
}
int rtb_host_register(void* host, size_t nbytes) {
    CU(cudaHostRegister(host, nbytes, cudaHostRegisterDefault), "cudaHostRegister");
    return 0;
}
int rtb_host_unregister(void* host) {
    cudaError_t ce = cudaHostUnregister(host);
    if (ce != cudaSuccess) { (void)cudaGetLastError(); return fail(-2, "cudaHostUnregister", ce); }
    return 0;
}

/* Host-only views of what the engine builds at upload time (no GPU needed): for the CPU tests. */
int rtb_debug_jump_table(uint32_t L, uint32_t out[480]) {
    if (!out) return fail(-1, "rtb_debug_jump_table: null argument");
    uint32_t tab[15][32];
    build_jump_table(L, tab);
    memcpy(out, tab, sizeof tab);
    return 0;
}
int rtb_debug_accel_build(const void* spheres, uint32_t S, float* clusters, float* records, uint32_t* index, float* groups,
                          uint32_t* n_clusters, uint32_t* n_groups) {
    if (!n_clusters || !n_groups) return fail(-1, "rtb_debug_accel_build: null argument");
    AccelHost h;
    const int built = build_accel_host((const unsigned char*)spheres, S, h);
    *n_clusters = h.nc;
    *n_groups = h.ng;
    if (built == 1) {
        if (clusters) memcpy(clusters, h.cl.data(), h.cl.size() * 4);
        if (records) memcpy(records, h.rec.data(), h.rec.size() * 4);
        if (index) memcpy(index, h.idx.data(), h.idx.size() * 4);
        if (groups) memcpy(groups, h.grp.data(), h.grp.size() * 4);
    }
    return built;
}
int rtb_debug_tri_accel_build(const float* records, uint32_t T, float* nodes, float* sorted_records, uint32_t* index,
                              uint32_t* n_clusters, uint32_t* n_levels, uint32_t level_offset[RT_TACC_LEVELS], uint32_t* n_nodes) {
    if (!n_clusters || !n_levels || !level_offset || !n_nodes) return fail(-1, "rtb_debug_tri_accel_build: null argument");
    TriAccelHost h;
    const int built = build_tri_accel_host(records, T, h);
    *n_clusters = h.nc;
    *n_levels = h.nl;
    *n_nodes = (uint32_t)(h.node.size() / 8);
    memcpy(level_offset, h.off, sizeof h.off);
    if (built == 1) {
        if (nodes) memcpy(nodes, h.node.data(), h.node.size() * 4);
        if (sorted_records) memcpy(sorted_records, h.rec.data(), h.rec.size() * 4);
        if (index) memcpy(index, h.idx.data(), h.idx.size() * 4);
    }
    return built;
}
int rtb_debug_counters(rtb_engine* e, uint64_t out[RTB_NCOUNTERS]) {
    if (!e || !out) return fail(-1, "rtb_debug_counters: null argument");
    CU(cudaSetDevice(e->device), "cudaSetDevice");
    CU(cudaMemcpyAsync(out, (char*)e->sched.p + e->last_sched_off + 64, RTB_NCOUNTERS * 8, cudaMemcpyDeviceToHost, e->stream), "D2H counters");
    CU(cudaStreamSynchronize(e->stream), "rtb_debug_counters");
    return 0;
}
int rtb_last_launch(rtb_engine* e, rtb_launch_info* info) {
    if (!e || !info) return fail(-1, "null argument");
    *info = e->last;
    return 0;
}

double rtb_fp32_peak_tflops(rtb_engine* e, int iters) {
    if (!e || iters < 1) return -1.0;
    if (cudaSetDevice(e->device) != cudaSuccess) return -1.0;
    float* sink = (float*)((char*)e->sched.p + 192);
    const int threads = 256, blocks = e->sm_count * 8;
    cudaEvent_t a, b;
    if (cudaEventCreate(&a) != cudaSuccess || cudaEventCreate(&b) != cudaSuccess) return -1.0;
    double best = -1.0;
    for (int rep = 0; rep < 4; rep++) { /* first repetition warms up */
        cudaEventRecord(a, e->stream);
        if (rt_launch_fma_peak(sink, iters, blocks, threads, e->stream) != cudaSuccess) { best = -1.0; break; }
        cudaEventRecord(b, e->stream);
        if (cudaEventSynchronize(b) != cudaSuccess) { best = -1.0; break; }
        float ms = 0.0f;
        cudaEventElapsedTime(&ms, a, b);
        double flop = 2.0 * 64.0 * (double)iters * (double)threads * (double)blocks;
        double tf = flop / ((double)ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
    }
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    return best;
}

/* ---------------------------------------------------------------------------------------------------------------------
 * rtb_render_noalloc -- one frame without a single allocation (no counterpart in the reference; added
 * for callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops, harnesses that
 * forbid it).
 *
 * The render kernel's scheduling scratch -- the 256-byte cursor block of render_frame, the histogram / cursors of
 * tile_order, one cost word and one queue entry per 8x4 tile -- lives in ONE file-scope device array instead of the
 * engine's cudaMalloc'ed buffers.  Every GPU that loads the library holds its 2.0 MiB until exit.  In host code the
 * array's NAME is a host-side stand-in, not a device address: the address comes from cudaGetSymbolAddress, once per
 * engine, with the engine's device current.  Every call on a device shares the array: calls that overlap on the GPU
 * (two engines of one device, two streams) RACE; the header words that must start at zero are cleared by a kernel at the
 * head of every call (the array keeps the last call's contents).
 * ------------------------------------------------------------------------------------------------------------------- */
#define RTB_NOALLOC_MAX_TILES 262144u /* 8x4-pixel tiles: 8,388,608 pixels in whole tiles (3840x2160 = 259,200 tiles fits) */
#define RTB_NOALLOC_HEADER_WORDS 320u /* 64 words cursor block (tile / tree / done counters, work counters) + 256 words tile_order scratch */
static __device__ __align__(256) uint32_t g_noalloc_scratch[RTB_NOALLOC_HEADER_WORDS + 2u * RTB_NOALLOC_MAX_TILES];

/* device code may name the array directly: clear what must start at zero */
static __global__ void noalloc_reset() {
    for (uint32_t i = threadIdx.x; i < RTB_NOALLOC_HEADER_WORDS; i += blockDim.x) g_noalloc_scratch[i] = 0u;
}

int rtb_render_noalloc(rtb_engine* e, const float cam[14], uint32_t W, uint32_t H, uint32_t total, const rtb_partition* part,
                       uint32_t* d_out, uint32_t* d_hit) {
    if (!e) return fail(-1, "null engine");
    if (!e->have_scene) return fail(-1, "rtb_render_noalloc: no scene uploaded");
    if (!cam && !e->have_cam) return fail(-1, "rtb_render_noalloc: no camera given and none set");
    if (!e->d_sky) return fail(-1, "rtb_render_noalloc: no skybox image uploaded (kernel arg 9)");
    if (!W || !H) return fail(-1, "rtb_render_noalloc: empty frame");
    if ((uint64_t)W * H > 0xffffffffull) return fail(-1, "rtb_render_noalloc: W*H exceeds the 32-bit pixel id (raygen.cl:10)");
    if (!d_out) return fail(-1, "rtb_render_noalloc: the caller passes the output buffer (this call allocates nothing)");
    if (((uintptr_t)d_out & 15) || ((uintptr_t)d_hit & 15)) return fail(-1, "rtb_render_noalloc: output buffers must be 16-byte aligned");
    rtb_partition whole = {H, 1, 0};
    rtb_partition pp = part ? *part : whole;
    if (pp.parts <= 1) { pp.parts = 1; pp.part = 0; pp.band_rows = H; }
    if (!pp.band_rows || pp.part >= pp.parts) return fail(-1, "rtb_render_noalloc: bad partition");
    const uint32_t rows = rtb_partition_rows(H, &pp);
    const uint64_t tiles = (uint64_t)((W + 7) / 8) * ((rows + 3) / 4);
    if (tiles > RTB_NOALLOC_MAX_TILES) {
        snprintf(g_err, sizeof g_err, "rtb_render_noalloc: %llu tiles of 8x4 pixels exceed the static scratch (%u tiles = 8,388,608 pixels; "
                 "3840x2160 fits): use rtb_render, or split the frame with an rtb_partition", (unsigned long long)tiles, RTB_NOALLOC_MAX_TILES);
        return -1;
    }
    CU(cudaSetDevice(e->device), "cudaSetDevice");
    if (!e->noalloc_scratch) { /* once per engine (= per device), with that device current */
        uint32_t* ptr = nullptr;
        CU(cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch), "cudaGetSymbolAddress(g_noalloc_scratch)");
        if (!ptr) return fail(-2, "rtb_render_noalloc: cudaGetSymbolAddress returned a null address");
        e->noalloc_scratch = ptr;
    }
    if (rows == 0) return 0;
    const float* cm = cam ? cam : e->cam;

    RtFrameParams fp;
    memset(&fp, 0, sizeof fp);
    memcpy(fp.corner, cm + 0, 12); memcpy(fp.origin, cm + 3, 12);
    memcpy(fp.up, cm + 6, 12); memcpy(fp.right, cm + 9, 12);
    fp.wf = cm[12]; fp.hf = cm[13];
    fp.W = W; fp.H = H; fp.total = total;
    fp.band_rows = pp.band_rows; fp.parts = pp.parts; fp.part = pp.part; fp.local_rows = rows;
    fp.S = e->S; fp.P = e->P; fp.L = e->L; fp.lay = e->lay;
    fp.blob = (const float4*)e->blob.p;
    fp.raw_spheres = e->d_spheres; fp.raw_planes = e->d_planes;
    fp.T = e->T; fp.tri = (const float4*)e->tri_rec.p; fp.raw_tris = e->d_tris;
    fp.tex = (const uchar4*)(e->d_tex ? e->d_tex : e->d_sky);
    fp.tw = e->d_tex ? (int)e->tw : 1; fp.th = e->d_tex ? (int)e->th : 1; fp.tl = e->d_tex ? (int)e->tl : 1;
    fp.sky = (const uchar4*)e->d_sky; fp.sw = (int)e->sw; fp.sh = (int)e->sh;
    fp.out = d_out;
    fp.hit = d_hit;
    fp.tiles_x = (W + 7) / 8;
    fp.tiles_y = (rows + 3) / 4;
    /* scratch: [0, 64) cursor block, [64, 320) tile_order's histogram / cursors / queue length, then costs, then the queue */
    uint32_t* const base = e->noalloc_scratch;
    uint32_t* const hist = base + 64;
    uint32_t* const cost = base + RTB_NOALLOC_HEADER_WORDS;
    uint32_t* const order = cost + RTB_NOALLOC_MAX_TILES;
    fp.tile_counter = (unsigned int*)base;
    fp.tree_counter = (unsigned int*)(base + 1);
    fp.done_counter = (unsigned int*)(base + 8);
    fp.counters = (unsigned long long*)(base + 16);
    fp.coop_min_cost = (uint32_t)e->opt_coop_min_cost;
    fp.wide_min_records = (uint32_t)e->opt_wide_min_records;
    fp.wide_max_live = e->opt_wide_max_live;
    /* the skeleton-trace schedule of a first frame: every tile once, heaviest first; no measured-time history, no split
     * entries, no pixel-tree kernel (those need per-frame buffers) -- a scheduling difference only, the words are the same */
    fp.split_min_cost = 64u;
    fp.split32_min_cost = 64u;
    fp.tree_min_cost = 64u;
    fp.tree_mode = 0;
    noalloc_reset<<<1, 320, 0, e->stream>>>();
    CU(cudaGetLastError(), "noalloc_reset launch");
    CU((e->T ? rt_launch_tile_order_tri : rt_launch_tile_order)(fp, cost, hist, order, e->stream), "tile_order launch");
    fp.order = order;
    fp.cost = cost;
    fp.queue_len = hist + 128;
    fp.n_split = hist + 129;
    RtLaunchShape shape{};
    CU((e->T ? rt_launch_render_tri : rt_launch_render)(fp, 0, e->sm_count, e->stream, &shape), "render_frame launch");
    return 0;
}

}  // extern "C"
