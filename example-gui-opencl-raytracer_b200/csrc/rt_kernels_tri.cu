/* rt_kernels_tri.cu -- EXTENSION X1 build of the render kernels: the same source with the triangle loops compiled in
 * (namespace rt_tri, launchers rt_launch_render_tri / rt_launch_tile_order_tri / rt_launch_tri_prep).  Scenes without
 * triangles never run this code: the plain build (rt_kernels.cu) has no trace of it. */
#define RT_TRIANGLES 1
#define RT_NS rt_tri
#include "rt_kernels.cu"
