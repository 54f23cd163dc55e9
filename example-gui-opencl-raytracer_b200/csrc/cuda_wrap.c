/* cuda_wrap.c -- the cl_wrap_* entry points of include/cuda_wrap.h over the B200 engine (rtb200.h).
 *
 * Host-side mirror of /root/reference/src/opencl_wrap.c: same entry points, argument meaning, slot
 * bookkeeping (`buffers`, `buffers_ids`, `buffers_num`) and error convention; the OpenCL runtime calls
 * are replaced by the engine's C ABI.  Kernel argument positions follow the reference kernels'
 * signatures: raygen (src/cl/raygen.cl:5-8) and raytracer (src/cl/raytracing.cl:14-21).
 */
#include "../../include/cuda_wrap.h"

#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/rtb200.h"
#include "png_io.h"

enum { K_NONE = 0, K_RAYGEN = 1, K_RAYTRACER = 2 };

/* raygen args (raygen.cl:5-8) */
enum { RG_CORNER = 0, RG_ORIGIN = 1, RG_UP = 2, RG_RIGHT = 3, RG_WF = 4, RG_HF = 5, RG_W = 6, RG_H = 7, RG_RAYS = 8 };
/* raytracer args (raytracing.cl:14-21) */
enum { RT_RAYS = 0, RT_SPHERES = 1, RT_PLANES = 2, RT_LIGHTS = 3, RT_NS = 4, RT_NP = 5, RT_NL = 6, RT_TOTAL = 7,
       RT_TEX = 8, RT_SKY = 9, RT_OUT = 10,
       RT_TRIS = 11, RT_NT = 12 /* EXTENSION X1 (optional): rtriangle buffer and its cl_uint count */ };

typedef struct {
    unsigned char bytes[16];
    size_t size; /* 0 = never set */
} scalar_cell;

typedef struct {
    size_t size;               /* bytes of the global buffer / image array in this slot (0 = none) */
    uint32_t iw, ih, il;       /* image array dimensions (il = 0: not an image) */
    int placeholder;           /* 1: handle only, no storage behind it (raygen's ray buffer) */
} slot_info;

typedef struct {
    scalar_cell scalars[__MAX_KERNELS][__MAX_BUFFERS];
    slot_info slots[__MAX_KERNELS][__MAX_BUFFERS];
    /* raygen state latched by running kernel "raygen" */
    int rays_valid;
    float cam[14];
    uint32_t W, H;
    /* what the engine currently has bound */
    const void *bound_s, *bound_p, *bound_l;
    uint32_t bound_ns, bound_np, bound_nl;
    int scene_dirty;
    const void *bound_tex, *bound_sky;
    const void* bound_t; uint32_t bound_nt;
    /* host read-back buffers seen by cl_wrap_output (N1: frame mirrored into the caller's buffer by the kernel) */
    struct {
        void* host;      /* caller's pointer */
        size_t size;
        void* dev_view;  /* device-visible alias once mapped, else NULL */
        int registered;  /* page-locked by this wrapper (to undo in cl_wrap_release) */
        int uses, failed;
    } hostbuf[4];
    int mirror_mode;     /* 0 auto (map on the 2nd use of a buffer), 1 never, 2 from the first use */
    unsigned long mirrored_frames, copied_frames;
} wrap_state;

#define CANARY(k) (0xA5C30000u | (uint32_t)(k)) /* top byte != 0: never a 0x00RRGGBB pixel word (raytracing.cl:194) */
enum { N_CANARY = 8 };

static void die(const char* msg) {
    printf("ERROR:\t%s\n", msg);
    fflush(stdout);
    exit(1);
}
static void die_engine(const char* what) {
    printf("ERROR:\t%s (%s)\n", what, rtb_last_error());
    fflush(stdout);
    exit(1);
}

static rtb_engine* eng(cl_wrap* w) { return (rtb_engine*)w->context; }
static wrap_state* st(cl_wrap* w) { return (wrap_state*)w->program; }

static void check_kernel(cl_wrap* wrap, cl_uint kernel_id) {
    if (kernel_id >= wrap->kernels_num) die("Wrong kernel index given");
}
/* opencl_wrap.c:136-141 */
static void check_unused(cl_wrap* wrap, cl_uint kernel_id, cl_uint arg_id) {
    for (cl_uint i = 0; i < wrap->buffers_num[kernel_id]; i++)
        if (wrap->buffers_ids[kernel_id][i] == arg_id) die("Given kernel argument already in use");
}

void cl_wrap_init(cl_wrap* wrap, cl_device_type type, ...) {
    va_list vars;
    const char* current_source_file;
    (void)type; /* the reference drivers hard-code CL_DEVICE_TYPE_GPU (raypng.c:32); any type selects the B200 */

    wrap->kernels_num = 0;
    memset(wrap->buffers, 0, sizeof wrap->buffers);
    memset(wrap->buffers_num, 0, sizeof wrap->buffers_num);

    if (rtb_device_count() < 1) die("Cannot find a device of the given type");
    int device = 0;
    const char* env = getenv("RTB_DEVICE"); /* one process per GPU: the launcher picks the ordinal */
    if (env && *env) device = atoi(env);
    rtb_engine* e = NULL;
    if (rtb_create(&e, device) != 0) die_engine("Could not create a CUDA context from device");
    wrap->context = (cl_context)e;
    wrap->device = (cl_device_id)(size_t)(device + 1);
    wrap->queue = (cl_command_queue)rtb_stream(e);
    wrap_state* s = (wrap_state*)calloc(1, sizeof(wrap_state));
    if (s) {
        const char* ro = getenv("RTB_HOST_OUTPUT"); /* copy | mirror | auto (default) */
        s->mirror_mode = ro && !strcmp(ro, "copy") ? 1 : (ro && !strcmp(ro, "mirror") ? 2 : 0);
    }
    if (!s) die("Out of host memory");
    wrap->program = (cl_program)s;

    va_start(vars, type);
    current_source_file = va_arg(vars, const char*);
    while (current_source_file) {
        const char* name = va_arg(vars, const char*);
        if (!name) { /* opencl_wrap.c:47-51 */
            va_end(vars);
            die("Source file was not followed by kernel name");
        }
        if (wrap->kernels_num >= __MAX_KERNELS) { va_end(vars); die("Too many kernels"); }
        int kind = K_NONE;
        if (!strcmp(name, "raygen")) kind = K_RAYGEN;
        else if (!strcmp(name, "raytracer")) kind = K_RAYTRACER;
        if (kind == K_NONE) { /* opencl_wrap.c:109-113 */
            va_end(vars);
            printf("ERROR:\tCouldn't create the CL kernel from: %s\n", name);
            exit(1);
        }
        wrap->kernels[wrap->kernels_num] = (cl_kernel)(size_t)kind;
        wrap->buffers_num[wrap->kernels_num] = 0;
        wrap->kernels_num++;
        current_source_file = va_arg(vars, const char*);
    }
    va_end(vars);
}

static int kind_of(cl_wrap* wrap, cl_uint kernel_id) { return (int)(size_t)wrap->kernels[kernel_id]; }

void cl_wrap_load_global_data(cl_wrap* wrap, cl_uint kernel_id, cl_uint arg_id, const void* data, size_t size,
                              cl_mem_flags mem_flags) {
    (void)mem_flags;
    check_kernel(wrap, kernel_id);
    check_unused(wrap, kernel_id, arg_id);
    if (arg_id >= __MAX_BUFFERS) die("Wrong kernel ID given"); /* opencl_wrap.c:143-146 */

    wrap_state* s = st(wrap);
    slot_info* si = &s->slots[kernel_id][arg_id];
    void* dptr = NULL;
    /* raygen's ray buffer (64 B/pixel, raypng.c:59) is fused away: hand out a handle, keep no storage */
    int placeholder = (kind_of(wrap, kernel_id) == K_RAYGEN && arg_id == RG_RAYS && data == NULL);
    if (rtb_dev_alloc(eng(wrap), &dptr, placeholder ? 16 : size) != 0) die_engine("Couldn't allocate device memory");
    if (data && size) {
        if (rtb_dev_write(eng(wrap), dptr, data, size) != 0) die_engine("Couldn't transfer the data from host to the device");
    }
    wrap->buffers[kernel_id][arg_id] = (cl_mem)dptr;
    si->size = size;
    si->il = 0;
    si->placeholder = placeholder;
    if (kind_of(wrap, kernel_id) == K_RAYTRACER && arg_id >= RT_SPHERES && arg_id <= RT_LIGHTS) s->scene_dirty = 1;
    wrap->buffers_ids[kernel_id][wrap->buffers_num[kernel_id]++] = arg_id;
}

void cl_wrap_load_single_data(cl_wrap* wrap, cl_uint kernel_id, cl_uint arg_id, const void* data, size_t obj_size) {
    check_kernel(wrap, kernel_id);
    check_unused(wrap, kernel_id, arg_id);
    if (arg_id >= __MAX_BUFFERS || !data || obj_size == 0 || obj_size > 16) die("Couldn't pass the data argument to the kernel");
    scalar_cell* c = &st(wrap)->scalars[kernel_id][arg_id];
    memset(c->bytes, 0, sizeof c->bytes);
    memcpy(c->bytes, data, obj_size);
    c->size = obj_size;
    if (kind_of(wrap, kernel_id) == K_RAYTRACER && arg_id >= RT_NS && arg_id <= RT_NL) st(wrap)->scene_dirty = 1;
}

void cl_wrap_load_images(cl_wrap* wrap, cl_uint kernel_id, cl_uint arg_id, cl_mem_flags mem_flags, cl_uint image_num, ...) {
    (void)mem_flags;
    va_list vars;
    uint8_t* images = NULL;
    uint32_t prev_w = 0, prev_h = 0;

    check_kernel(wrap, kernel_id);
    check_unused(wrap, kernel_id, arg_id);
    if (arg_id >= __MAX_BUFFERS) die("Wrong kernel ID given");

    va_start(vars, image_num);
    for (cl_uint i = 0; i < image_num; i++) {
        const char* filename = va_arg(vars, const char*);
        uint32_t w = 0, h = 0;
        int depth = 0, ctype = 0, interlace = 0;
        int rc = rtpng_info(filename, &w, &h, &depth, &ctype, &interlace);
        if (rc == RTPNG_EOPEN) { printf("ERROR:\tCannot open file \"%s\"\n", filename); exit(1); }
        if (rc != RTPNG_OK) { printf("ERROR:\t\"%s\" is not a PNG file\n", filename); exit(1); }
        if (prev_w == 0 && prev_h == 0) {
            prev_w = w;
            prev_h = h;
            images = (uint8_t*)malloc((size_t)4 * w * h * image_num);
            if (!images) die("Out of host memory");
        }
        if (prev_w != w || prev_h != h) { free(images); die("All images must have same dimensions"); }
        if (depth != 8 || ctype != 2) { /* opencl_wrap.c:299-307 */
            free(images);
            printf("ERROR:\t\"%s\" must have a depth of 8 bits and be RGB\n", filename);
            exit(1);
        }
        /* RTB_IMAGE_CACHE=<dir>: decoded texels are kept there and re-read on the next start (N3) */
        rc = rtpng_read_rgb8_to_rgba_cached(filename, images + (size_t)i * 4 * w * h, w, h, getenv("RTB_IMAGE_CACHE"), NULL);
        if (rc != RTPNG_OK) {
            free(images);
            printf("ERROR:\tCould not decode \"%s\" (%d)\n", filename, rc);
            exit(1);
        }
    }
    va_end(vars);
    if (!images) die("Couldn't create an image array 0");

    size_t nbytes = (size_t)4 * prev_w * prev_h * image_num;
    void* dptr = NULL;
    if (rtb_dev_alloc(eng(wrap), &dptr, nbytes) != 0) { free(images); die_engine("Couldn't create an image array"); }
    if (rtb_dev_write(eng(wrap), dptr, images, nbytes) != 0) { free(images); die_engine("Couldn't create an image array"); }
    free(images);

    wrap->buffers[kernel_id][arg_id] = (cl_mem)dptr;
    slot_info* si = &st(wrap)->slots[kernel_id][arg_id];
    si->size = nbytes; si->iw = prev_w; si->ih = prev_h; si->il = image_num; si->placeholder = 0;
    wrap->buffers_ids[kernel_id][wrap->buffers_num[kernel_id]++] = arg_id;
}

/* N1: a device-visible alias of the caller's read-back buffer, or NULL to use the plain D2H copy.
 * Page-locking costs ~0.25 ms/MB once, so in auto mode a buffer is mapped on its second use (a one-frame
 * raypng run never pays it; the interactive loop pays it once). */
static void* mirror_view(cl_wrap* wrap, void* host, size_t size) {
    wrap_state* s = st(wrap);
    if (s->mirror_mode == 1 || ((uintptr_t)host & 15)) return NULL;
    int free_i = -1;
    for (int i = 0; i < 4; i++) {
        if (s->hostbuf[i].host == host && s->hostbuf[i].size == size) {
            if (s->hostbuf[i].failed) return NULL;
            if (s->hostbuf[i].dev_view) return s->hostbuf[i].dev_view;
            free_i = i;
            break;
        }
        if (!s->hostbuf[i].host && free_i < 0) free_i = i;
    }
    if (free_i < 0) return NULL; /* table full: keep copying */
    s->hostbuf[free_i].host = host;
    s->hostbuf[free_i].size = size;
    if (++s->hostbuf[free_i].uses < 2 && s->mirror_mode != 2) return NULL;
    void* view = NULL;
    int reg = 0;
    if (rtb_host_map(host, size, &view, &reg) != 0 || !view) { s->hostbuf[free_i].failed = 1; return NULL; }
    s->hostbuf[free_i].dev_view = view;
    s->hostbuf[free_i].registered = reg;
    return view;
}
static void mirror_drop(cl_wrap* wrap, void* host) {
    wrap_state* s = st(wrap);
    for (int i = 0; i < 4; i++)
        if (s->hostbuf[i].host == host) {
            if (s->hostbuf[i].registered) rtb_host_unregister(host);
            s->hostbuf[i].dev_view = NULL;
            s->hostbuf[i].registered = 0;
            s->hostbuf[i].failed = 1;
        }
}

static uint32_t scalar_count(const scalar_cell* c) { /* cl_uchar in the reference (raytracing.cl:17); cl_uint accepted */
    if (c->size == 1) return c->bytes[0];
    if (c->size == 2) { uint16_t v; memcpy(&v, c->bytes, 2); return v; }
    uint32_t v;
    memcpy(&v, c->bytes, 4);
    return v;
}

static void run_raygen(cl_wrap* wrap, size_t array_size) {
    wrap_state* s = st(wrap);
    int k = -1;
    for (cl_uint i = 0; i < wrap->kernels_num; i++) if (kind_of(wrap, i) == K_RAYGEN) { k = (int)i; break; }
    const scalar_cell* c = s->scalars[k];
    static const size_t want[8] = {16, 16, 16, 16, 4, 4, 4, 4};
    for (int a = 0; a < 8; a++)
        if (c[a].size != want[a]) die("Couldn't run the kernel");  /* CL_INVALID_KERNEL_ARGS */
    memcpy(s->cam + 0, c[RG_CORNER].bytes, 12);
    memcpy(s->cam + 3, c[RG_ORIGIN].bytes, 12);
    memcpy(s->cam + 6, c[RG_UP].bytes, 12);
    memcpy(s->cam + 9, c[RG_RIGHT].bytes, 12);
    memcpy(s->cam + 12, c[RG_WF].bytes, 4);
    memcpy(s->cam + 13, c[RG_HF].bytes, 4);
    memcpy(&s->W, c[RG_W].bytes, 4);
    memcpy(&s->H, c[RG_H].bytes, 4);
    (void)array_size;
    s->rays_valid = 1;
}

static uint64_t run_raytracer(cl_wrap* wrap, cl_uint k, size_t array_size, void* host_output, size_t output_size, int* mirrored) {
    wrap_state* s = st(wrap);
    rtb_engine* e = eng(wrap);
    const scalar_cell* c = s->scalars[k];
    if (!s->rays_valid) die("Couldn't run the kernel"); /* rays were never generated */
    if (c[RT_NS].size == 0 || c[RT_NP].size == 0 || c[RT_NL].size == 0 || c[RT_TOTAL].size != 4 || c[RT_RAYS].size != sizeof(cl_mem))
        die("Couldn't run the kernel");
    uint32_t ns = scalar_count(&c[RT_NS]), np = scalar_count(&c[RT_NP]), nl = scalar_count(&c[RT_NL]);
    uint32_t total;
    memcpy(&total, c[RT_TOTAL].bytes, 4);
    const void* ds = wrap->buffers[k][RT_SPHERES];
    const void* dp = wrap->buffers[k][RT_PLANES];
    const void* dl = wrap->buffers[k][RT_LIGHTS];
    const slot_info* sl = s->slots[k];
    if ((size_t)ns * 96 > sl[RT_SPHERES].size || (size_t)np * 96 > sl[RT_PLANES].size || (size_t)nl * 48 > sl[RT_LIGHTS].size)
        die("Couldn't run the kernel"); /* counts exceed the uploaded arrays */
    if (sl[RT_TEX].il == 0 || sl[RT_SKY].il == 0 || sl[RT_OUT].size == 0) die("Couldn't run the kernel");

    if (s->scene_dirty || ds != s->bound_s || dp != s->bound_p || dl != s->bound_l || ns != s->bound_ns ||
        np != s->bound_np || nl != s->bound_nl) {
        if (rtb_bind_scene_device(e, ds, ns, dp, np, dl, nl) != 0) die_engine("Couldn't pass the data argument to the kernel");
        s->bound_s = ds; s->bound_p = dp; s->bound_l = dl;
        s->bound_ns = ns; s->bound_np = np; s->bound_nl = nl;
        s->scene_dirty = 0;
    }
    { /* extension X1: optional args 11 (rtriangle buffer) and 12 (count); absent = no triangles */
        const void* dt = wrap->buffers[k][RT_TRIS];
        uint32_t nt = (dt && c[RT_NT].size) ? scalar_count(&c[RT_NT]) : 0;
        if ((size_t)nt * 112 > sl[RT_TRIS].size) die("Couldn't run the kernel");
        if (dt != s->bound_t || nt != s->bound_nt) {
            if (rtb_bind_triangles_device(e, dt, nt) != 0) die_engine("Couldn't pass the data argument to the kernel");
            s->bound_t = dt; s->bound_nt = nt;
        }
    }
    if (rtb_set_camera(e, s->cam) != 0) die_engine("Couldn't pass the data argument to the kernel");

    /* work-items beyond array_size do not exist; those beyond total_size return early (raytracing.cl:23-26) */
    uint64_t n = (uint64_t)s->W * s->H;
    if (n > array_size) n = array_size;
    if (n > total) n = total;
    if ((size_t)n * 4 > sl[RT_OUT].size) die("The device kernel failed"); /* would write past the output buffer */
    *mirrored = 0;
    if (host_output && output_size == (size_t)n * 4 && n >= 64) {
        uint32_t* view = (uint32_t*)mirror_view(wrap, host_output, output_size);
        if (view) {
            volatile uint32_t* h = (volatile uint32_t*)host_output;
            for (int c = 0; c < N_CANARY; c++) h[(size_t)c * (n - 1) / (N_CANARY - 1)] = CANARY(c);
            if (rtb_set_mirror(e, view) != 0) die_engine("Couldn't run the kernel");
            *mirrored = 1;
        }
    }
    if (rtb_render(e, s->W, s->H, (uint32_t)n, NULL, (uint32_t*)wrap->buffers[k][RT_OUT], NULL, 0) != 0)
        die_engine("Couldn't run the kernel");
    return n;
}

/* image arrays are bound lazily (their slots can be filled in any order before the first launch) */
static void bind_images(cl_wrap* wrap, cl_uint k);

void cl_wrap_output(cl_wrap* wrap, size_t array_size, size_t output_size, cl_uint kernel_run_id, cl_uint kernel_id,
                    cl_int arg_id, void* host_output) {
    check_kernel(wrap, kernel_run_id);
    int mirrored = 0;
    uint64_t n = 0;
    switch (kind_of(wrap, kernel_run_id)) {
        case K_RAYGEN: run_raygen(wrap, array_size); break;
        case K_RAYTRACER: {
            bind_images(wrap, kernel_run_id);
            /* the frame can be mirrored into host_output only when the read-back names this launch's output slot */
            int same = host_output && kernel_id == kernel_run_id && arg_id == RT_OUT && wrap->buffers[kernel_id][RT_OUT];
            n = run_raytracer(wrap, kernel_run_id, array_size, same ? host_output : NULL, output_size, &mirrored);
            break;
        }
        default: die("Couldn't run the kernel");
    }
    if (rtb_sync(eng(wrap)) != 0) die_engine("The device kernel failed"); /* clFinish, opencl_wrap.c:380-385 */

    if (!host_output) return;
    if (mirrored) {
        /* every canary must have been overwritten by a pixel word; if one survives, the mapping no longer
         * aliases the caller's pages (buffer freed and re-allocated at the same address): drop it and copy */
        const volatile uint32_t* h = (const volatile uint32_t*)host_output;
        int ok = 1;
        for (int c = 0; c < N_CANARY; c++) if ((h[(size_t)c * (n - 1) / (N_CANARY - 1)] >> 24) != 0) ok = 0;
        if (ok) { st(wrap)->mirrored_frames++; return; }
        mirror_drop(wrap, host_output);
    }
    st(wrap)->copied_frames++;
    check_kernel(wrap, kernel_id);
    if (arg_id < 0 || arg_id >= __MAX_BUFFERS || !wrap->buffers[kernel_id][arg_id]) die("Failed to transfer device memory to host");
    const slot_info* si = &st(wrap)->slots[kernel_id][arg_id];
    if (si->placeholder || output_size > si->size) die("Failed to transfer device memory to host");
    if (rtb_dev_read(eng(wrap), host_output, wrap->buffers[kernel_id][arg_id], output_size) != 0)
        die_engine("Failed to transfer device memory to host");
}

static void bind_images(cl_wrap* wrap, cl_uint k) {
    wrap_state* s = st(wrap);
    const slot_info* sl = s->slots[k];
    if (sl[RT_TEX].il && wrap->buffers[k][RT_TEX] != s->bound_tex) {
        if (rtb_bind_images_device(eng(wrap), 0, wrap->buffers[k][RT_TEX], sl[RT_TEX].iw, sl[RT_TEX].ih, sl[RT_TEX].il) != 0)
            die_engine("Couldn't pass the image array to the kernel");
        s->bound_tex = wrap->buffers[k][RT_TEX];
    }
    if (sl[RT_SKY].il && wrap->buffers[k][RT_SKY] != s->bound_sky) {
        if (rtb_bind_images_device(eng(wrap), 1, wrap->buffers[k][RT_SKY], sl[RT_SKY].iw, sl[RT_SKY].ih, sl[RT_SKY].il) != 0)
            die_engine("Couldn't pass the image array to the kernel");
        s->bound_sky = wrap->buffers[k][RT_SKY];
    }
}

void cl_wrap_release(cl_wrap* wrap) {
    if (!wrap->context) return;
    rtb_sync(eng(wrap));
    for (int i = 0; i < 4; i++)
        if (st(wrap)->hostbuf[i].registered) rtb_host_unregister(st(wrap)->hostbuf[i].host);
    for (cl_uint k = 0; k < wrap->kernels_num; k++) {
        for (cl_uint i = 0; i < wrap->buffers_num[k]; i++) {
            cl_uint a = wrap->buffers_ids[k][i];
            if (wrap->buffers[k][a]) rtb_dev_free(eng(wrap), wrap->buffers[k][a]);
            wrap->buffers[k][a] = NULL;
        }
        wrap->buffers_num[k] = 0;
    }
    rtb_destroy(eng(wrap));
    free(wrap->program);
    wrap->context = NULL;
    wrap->program = NULL;
    wrap->queue = NULL;
    wrap->kernels_num = 0;
}

struct rtb_engine* cuda_wrap_engine(cl_wrap* wrap) { return wrap ? eng(wrap) : NULL; }
void cuda_wrap_readback_stats(cl_wrap* wrap, unsigned long* mirrored, unsigned long* copied) {
    if (mirrored) *mirrored = wrap && wrap->program ? st(wrap)->mirrored_frames : 0;
    if (copied) *copied = wrap && wrap->program ? st(wrap)->copied_frames : 0;
}
float cuda_wrap_last_kernel_ms(cl_wrap* wrap) { return wrap && wrap->context ? rtb_last_render_ms(eng(wrap)) : -1.0f; }
