/* include/shim/CL/opencl.h -- build shim, NOT an OpenCL implementation.
 *
 * The reference's host files (raypng.c, rayinteractive.c, src/cpu_ray.[ch], src/cpu_obj.[ch]) include
 * <CL/opencl.h> only for the cl_* scalar/vector TYPES and a handful of enums; no OpenCL headers are
 * installed in this image.  This header supplies exactly those names with the
 * Khronos-specified sizes, alignments and enum values so that those files compile untouched against
 * the CUDA backend (include/cuda_wrap.h).  On a machine with real OpenCL headers use those instead.
 */
#ifndef RTB_SHIM_CL_OPENCL_H
#define RTB_SHIM_CL_OPENCL_H
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h> /* the reference's cpu_obj.c calls malloc without including it */

#ifdef __cplusplus
extern "C" {
#endif

typedef int8_t cl_char;
typedef uint8_t cl_uchar;
typedef int16_t cl_short;
typedef uint16_t cl_ushort;
typedef int32_t cl_int;
typedef uint32_t cl_uint;
typedef int64_t cl_long;
typedef uint64_t cl_ulong;
typedef float cl_float;
typedef double cl_double;
typedef cl_uint cl_bool;
typedef cl_ulong cl_bitfield;
typedef cl_bitfield cl_device_type;
typedef cl_bitfield cl_mem_flags;

#define CL_FALSE 0
#define CL_TRUE 1

#define CL_DEVICE_TYPE_DEFAULT (1 << 0)
#define CL_DEVICE_TYPE_CPU (1 << 1)
#define CL_DEVICE_TYPE_GPU (1 << 2)
#define CL_DEVICE_TYPE_ACCELERATOR (1 << 3)
#define CL_DEVICE_TYPE_ALL 0xFFFFFFFF

#define CL_MEM_READ_WRITE (1 << 0)
#define CL_MEM_WRITE_ONLY (1 << 1)
#define CL_MEM_READ_ONLY (1 << 2)
#define CL_MEM_USE_HOST_PTR (1 << 3)
#define CL_MEM_ALLOC_HOST_PTR (1 << 4)
#define CL_MEM_COPY_HOST_PTR (1 << 5)

/* cl_float4 / cl_float3: 16 bytes, 16-byte aligned, .s[4] and .x .y .z .w views (cl_platform.h) */
typedef union {
    cl_float __attribute__((aligned(16))) s[4];
    __extension__ struct { cl_float x, y, z, w; };
    __extension__ struct { cl_float s0, s1, s2, s3; };
} cl_float4;
typedef cl_float4 cl_float3;

/* opaque handles */
typedef struct _cl_platform_id* cl_platform_id;
typedef struct _cl_device_id* cl_device_id;
typedef struct _cl_context* cl_context;
typedef struct _cl_command_queue* cl_command_queue;
typedef struct _cl_mem* cl_mem;
typedef struct _cl_program* cl_program;
typedef struct _cl_kernel* cl_kernel;

#ifdef __cplusplus
}
#endif
#endif
