/* rt_kernels_accel.cu -- SURVEY.md 8(f) N4 build of the render kernels: the same source with the sphere walks going
 * through the cluster structure of rt_trace_accel.cuh (namespace rt_acc, launchers rt_launch_render_acc /
 * rt_launch_tile_order_acc).  Scenes rendered by brute force -- the default, as in the reference -- never run this
 * code: the plain build (rt_kernels.cu) has no trace of it. */
#define RT_ACCEL 1
#define RT_NS rt_acc
#include "rt_kernels.cu"
