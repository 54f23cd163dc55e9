/* rt_kernels_tri_accel.cu -- SURVEY.md 8(f) N4 applied to extension X1: the triangle build of the render kernels with the
 * triangle walks (and the sphere walks, when the scene has enough spheres) going through the bounding-sphere hierarchy
 * of rt_trace_accel.cuh (namespace rt_tacc, launchers rt_launch_render_tacc / rt_launch_tile_order_tacc).  Only the
 * production variant (mode 0) is compiled.  Off unless rtb_set_option("accel", 1); brute force stays the default. */
#define RT_TRIANGLES 1
#define RT_ACCEL 1
#define RT_NS rt_tacc
#include "rt_kernels.cu"
