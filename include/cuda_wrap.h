/* include/cuda_wrap.h -- the drop-in device-wrapper boundary, CUDA edition.
 *
 * Replaces /root/reference/src/opencl_wrap.h:1-42 (implementation src/opencl_wrap.c:11-416) with the
 * SAME type name, field names, entry-point names, signatures, call order and error behaviour
 * ("ERROR:\t<msg>\n" on stdout, then exit(1); opencl_wrap.h:9), so raypng.c:31-98 and
 * rayinteractive.c:28,98-103,132-202 compile and run against it untouched.  include/shim/opencl_wrap.h
 * forwards the reference's include name to this file.
 *
 * What changes underneath (documented in DESIGN.md):
 *   - kernel index k names one of two built-in CUDA entry points, chosen by the kernel NAME passed to
 *     cl_wrap_init ("raygen", "raytracer"); the .cl source paths are accepted and ignored;
 *   - raygen is fused into the raytracer launch: running kernel "raygen" latches its scalar arguments
 *     (camera basis, width, height), running "raytracer" renders; the ray buffer requested for raygen
 *     arg 8 is given a handle (drivers alias it into raytracer arg 0) but never materialised;
 *   - read-back: when cl_wrap_output's host_output names the launch's own output slot, the
 *     kernel also stores the frame straight into that host buffer (page-locked and mapped on its second use;
 *     RTB_HOST_OUTPUT=copy|mirror|auto) so no D2H copy follows the render; the device slot still gets the frame;
 *   - cl_mem is a CUDA device pointer; cl_wrap.context holds the rtb_engine (include/rtb200.h),
 *     cl_wrap.program the wrapper's private slot table.
 */
#ifndef CUDA_WRAP_H
#define CUDA_WRAP_H

#include <CL/opencl.h>

#define __MAX_KERNELS 16 /* opencl_wrap.h:6 */
#define __MAX_BUFFERS 32 /* opencl_wrap.h:7 */

#ifdef __cplusplus
extern "C" {
#endif

/* opencl_wrap.h:11-23 -- drivers read exactly one field, buffers[0][8] (raypng.c:61). */
typedef struct {
    cl_context context;     /* rtb_engine* */
    cl_device_id device;    /* CUDA device ordinal + 1 */
    cl_program program;     /* private slot table */
    cl_command_queue queue; /* the engine's cudaStream_t */

    cl_uint kernels_num;
    cl_kernel kernels[__MAX_KERNELS]; /* 1 = raygen, 2 = raytracer */

    cl_mem buffers[__MAX_KERNELS][__MAX_BUFFERS]; /* CUDA device pointers */
    cl_uint buffers_ids[__MAX_KERNELS][__MAX_BUFFERS];
    cl_uint buffers_num[__MAX_KERNELS];
} cl_wrap;

/* opencl_wrap.h:29 / opencl_wrap.c:11-127.  varargs: (source path, kernel name)* NULL. */
void cl_wrap_init(cl_wrap* wrap, cl_device_type type, ...);
/* opencl_wrap.h:32-33 / opencl_wrap.c:129-170.  Allocates `size` bytes for slot (kernel, arg); blocking H2D if data. */
void cl_wrap_load_global_data(cl_wrap* wrap, cl_uint kernel_id, cl_uint arg_id, const void* data, size_t size,
                              cl_mem_flags mem_flags);
/* opencl_wrap.h:34-35 / opencl_wrap.c:172-187.  By-value argument; may be repeated per frame. */
void cl_wrap_load_single_data(cl_wrap* wrap, cl_uint kernel_id, cl_uint arg_id, const void* data, size_t obj_size);
/* opencl_wrap.h:36-37 / opencl_wrap.c:189-349.  varargs: image_num PNG paths (8-bit RGB, equal size). */
void cl_wrap_load_images(cl_wrap* wrap, cl_uint kernel_id, cl_uint arg_id, cl_mem_flags mem_flags, cl_uint image_num, ...);
/* opencl_wrap.h:39-41 / opencl_wrap.c:351-398.  Run kernel_run_id over array_size work-items, wait, then
 * (host_output != NULL) blocking D2H of output_size bytes from slot (kernel_id, arg_id). */
void cl_wrap_output(cl_wrap* wrap, size_t array_size, size_t output_size, cl_uint kernel_run_id, cl_uint kernel_id,
                    cl_int arg_id, void* host_output);
/* opencl_wrap.h:42 / opencl_wrap.c:400-416 */
void cl_wrap_release(cl_wrap* wrap);

/* ---- extensions (not in the reference) ---- */
/* The engine behind a wrapper, for the hit-id AOV, counters and timing (include/rtb200.h). */
struct rtb_engine* cuda_wrap_engine(cl_wrap* wrap);
/* How many read-backs were mirrored into the caller's buffer by the kernel / done as a D2H copy. */
void cuda_wrap_readback_stats(cl_wrap* wrap, unsigned long* mirrored, unsigned long* copied);
/* Device milliseconds of the last raytracer launch. */
float cuda_wrap_last_kernel_ms(cl_wrap* wrap);

#ifdef __cplusplus
}
#endif
#endif
