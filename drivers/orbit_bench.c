/* drivers/orbit_bench.c -- BASELINE.json config 5: the interactive camera-orbit sweep at an arbitrary size.
 *
 * The reference's rayinteractive.c fixes WIDTH/HEIGHT at 800x600 with unconditional #defines
 * (rayinteractive.c:13-14), so the 2560x1440 sweep cannot go through that file unchanged.  This driver
 * replays ITS call sequence -- set-up as rayinteractive.c:132-177, per frame the keyboard-callback update
 * of :86-103 (RIGHT = look around, D = strafe) followed by the two cl_wrap_output calls of :186-189 --
 * against the same cl_wrap_* boundary, using the reference's own host helpers (src/cpu_ray.c
 * rinit_camera/rgen_perspective, src/cpu_obj.c extract_robj) linked in unchanged.
 *
 *   orbit_bench [W H [frames [last_frame.png|- [count_every]]]]   (run from a directory holding scenes/ and assets/)
 * Prints one JSON line with frame-time p50/p99 (host clock around each frame: the API is synchronous).
 * count_every > 0: after every count_every-th frame (outside the timed bracket) the same camera is rendered once more
 * by the counted literal launch (rtb_render_counted: the reference's own work counts), so the line also carries the
 * rays and algorithmic flop of those sample frames next to their render_frame kernel times (the roofline of config 5).
 */
#define _GNU_SOURCE
#include <CL/opencl.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "opencl_wrap.h"
#include "rtb200.h"
#include "cpu_ray.h"
#include "cpu_obj.h"

#define CAMERA_SPEED 0.05f /* rayinteractive.c:16 */
#define MOVE_SPEED 0.1f    /* rayinteractive.c:17 */

static double now_ms(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return 1e3 * (double)ts.tv_sec + 1e-6 * (double)ts.tv_nsec;
}
static int cmp_d(const void* a, const void* b) { double x = *(const double*)a, y = *(const double*)b; return (x > y) - (x < y); }

int main(int argc, char** argv) {
    cl_uint W = argc > 2 ? (cl_uint)atoi(argv[1]) : 2560, H = argc > 2 ? (cl_uint)atoi(argv[2]) : 1440;
    int frames = argc > 3 ? atoi(argv[3]) : 240;
    const char* last_png = argc > 4 && strcmp(argv[4], "-") ? argv[4] : NULL;
    const int count_every = argc > 5 ? atoi(argv[5]) : 0;
    if (W == 0 || H == 0 || frames < 1) { fprintf(stderr, "usage: orbit_bench [W H [frames [last.png]]]\n"); return 2; }

    rcamera camera = rinit_camera((cl_float3){.x = 0.8f, .y = 2.5f, .z = -8.0f}, (cl_float3){.x = 0.0f, .y = 0.0f, .z = 1.0f},
                                  90.0f, 1.0f);
    float X_ROT = (float)M_PI_2, Y_ROT = (float)M_PI_2;

    cl_uchar ns, np, nl;
    rsphere* spheres; rplane* planes; rlight* lights;
    extract_robj("scenes/render.map", &spheres, &ns, &planes, &np, &lights, &nl);

    cl_wrap wrap;
    cl_wrap_init(&wrap, CL_DEVICE_TYPE_GPU, "src/cl/raygen.cl", "raygen", "src/cl/raytracing.cl", "raytracer", NULL);

    cl_float3 corner, origin, up, right;
    cl_float wf, hf;
    rgen_perspective(&camera, &corner, &origin, &up, &right, &wf, &hf, W, H);
    cl_uint pixels = W * H;
    size_t ray_size = (size_t)sizeof(rray) * pixels, buffer_size = (size_t)pixels * sizeof(cl_uint);
    cl_uint* buffer = malloc(buffer_size);

    cl_wrap_load_single_data(&wrap, 0, 0, &corner, sizeof(cl_float3));
    cl_wrap_load_single_data(&wrap, 0, 1, &origin, sizeof(cl_float3));
    cl_wrap_load_single_data(&wrap, 0, 2, &up, sizeof(cl_float3));
    cl_wrap_load_single_data(&wrap, 0, 3, &right, sizeof(cl_float3));
    cl_wrap_load_single_data(&wrap, 0, 4, &wf, sizeof(cl_float));
    cl_wrap_load_single_data(&wrap, 0, 5, &hf, sizeof(cl_float));
    cl_wrap_load_single_data(&wrap, 0, 6, &W, sizeof(cl_uint));
    cl_wrap_load_single_data(&wrap, 0, 7, &H, sizeof(cl_uint));
    cl_wrap_load_global_data(&wrap, 0, 8, NULL, ray_size, CL_MEM_READ_WRITE);
    cl_wrap_load_single_data(&wrap, 1, 0, &wrap.buffers[0][8], sizeof(cl_mem));
    cl_wrap_load_global_data(&wrap, 1, 1, spheres, sizeof(rsphere) * ns, CL_MEM_READ_ONLY);
    cl_wrap_load_global_data(&wrap, 1, 2, planes, sizeof(rplane) * np, CL_MEM_READ_ONLY);
    cl_wrap_load_global_data(&wrap, 1, 3, lights, sizeof(rlight) * nl, CL_MEM_READ_ONLY);
    cl_wrap_load_single_data(&wrap, 1, 4, &ns, sizeof(cl_uchar));
    cl_wrap_load_single_data(&wrap, 1, 5, &np, sizeof(cl_uchar));
    cl_wrap_load_single_data(&wrap, 1, 6, &nl, sizeof(cl_uchar));
    cl_wrap_load_single_data(&wrap, 1, 7, &pixels, sizeof(cl_uint));
    cl_wrap_load_images(&wrap, 1, 8, CL_MEM_COPY_HOST_PTR, 4, "assets/cobblestone.png", "assets/sand.png", "assets/check.png",
                        "assets/grass.png");
    cl_wrap_load_images(&wrap, 1, 9, CL_MEM_COPY_HOST_PTR, 1, "assets/bg/stormydays.png");
    cl_wrap_load_global_data(&wrap, 1, 10, NULL, buffer_size, CL_MEM_WRITE_ONLY);

    double* dt = malloc(sizeof(double) * (size_t)frames);
    double kernel_ms = 0.0, sampled_kernel_ms = 0.0;
    unsigned long long checksum = 0, sampled_rays = 0, sampled_flop = 0;
    int sampled = 0;
    double total_ms = 0.0; /* sum of the frame brackets (the accounting launches are outside them) */
    for (int f = 0; f < frames; f++) {
        const double t0 = now_ms();
        /* KB_KEY_RIGHT then KB_KEY_D, each followed by the callback's tail (rayinteractive.c:86-103) */
        for (int ev = 0; ev < 2; ev++) {
            if (ev == 0) Y_ROT += CAMERA_SPEED;
            else {
                camera.pos_dir.origin.x += MOVE_SPEED * right.x;
                camera.pos_dir.origin.y += MOVE_SPEED * right.y;
                camera.pos_dir.origin.z += MOVE_SPEED * right.z;
            }
            cl_float3 nd;
            nd.x = sin(X_ROT) * cos(Y_ROT);
            nd.z = sin(X_ROT) * sin(Y_ROT);
            nd.y = cos(X_ROT);
            camera.pos_dir.dir = nd;
            rgen_perspective(&camera, &corner, &origin, &up, &right, &wf, &hf, W, H);
            cl_wrap_load_single_data(&wrap, 0, 0, &corner, sizeof(cl_float3));
            cl_wrap_load_single_data(&wrap, 0, 1, &origin, sizeof(cl_float3));
            cl_wrap_load_single_data(&wrap, 0, 2, &up, sizeof(cl_float3));
            cl_wrap_load_single_data(&wrap, 0, 3, &right, sizeof(cl_float3));
            cl_wrap_load_single_data(&wrap, 0, 4, &wf, sizeof(cl_float));
            cl_wrap_load_single_data(&wrap, 0, 5, &hf, sizeof(cl_float));
        }
        cl_wrap_output(&wrap, (size_t)W * H, 0, 0, 0, 0, NULL);
        cl_wrap_output(&wrap, (size_t)W * H, buffer_size, 1, 1, 10, buffer);
        dt[f] = now_ms() - t0;
        total_ms += dt[f];
        const double this_kernel_ms = cuda_wrap_last_kernel_ms(&wrap);
        kernel_ms += this_kernel_ms;
        checksum = checksum * 1099511628211ull + buffer[((size_t)f * 7919u) % pixels];
        if (count_every > 0 && f % count_every == count_every - 1) { /* untimed accounting launch, same camera */
            uint64_t cnt[RTB_NCOUNTERS];
            if (rtb_render_counted(cuda_wrap_engine(&wrap), W, H, pixels, NULL, NULL, NULL, 0, 0, cnt) == 0) {
                sampled++;
                sampled_kernel_ms += this_kernel_ms;
                sampled_rays += cnt[RTB_SEGMENTS] + cnt[RTB_SHADOW_RAYS];
                sampled_flop += 21ull * cnt[RTB_SPHERE_TESTS] + 5ull * cnt[RTB_SPHERE_DPOS] + 14ull * cnt[RTB_PLANE_TESTS];
            }
        }
    }
    if (last_png) png_dump(last_png, buffer, (cl_int)W, (cl_int)H);
    cl_wrap_release(&wrap);

    qsort(dt, (size_t)frames, sizeof(double), cmp_d);
    printf("{\"workload\": \"orbit\", \"width\": %u, \"height\": %u, \"frames\": %d, \"frames_per_s\": %.3f, "
           "\"frame_ms_p50\": %.4f, \"frame_ms_p99\": %.4f, \"frame_ms_max\": %.4f, \"kernel_ms_mean\": %.4f, "
           "\"d2h_bytes_per_frame\": %zu, \"checksum\": \"%016llx\", \"sampled_frames\": %d, \"sampled_rays\": %llu, "
           "\"sampled_flop\": %llu, \"sampled_kernel_ms\": %.4f}\n",
           W, H, frames, frames / (total_ms * 1e-3), dt[frames / 2], dt[(int)(0.99 * (frames - 1))], dt[frames - 1],
           kernel_ms / frames, buffer_size, checksum, sampled, sampled_rays, sampled_flop, sampled_kernel_ms);
    free(dt); free(spheres); free(planes); free(lights); free(buffer);
    return 0;
}
