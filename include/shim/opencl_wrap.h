/* include/shim/opencl_wrap.h -- forwards the reference's `#include "opencl_wrap.h"` (raypng.c:3,
 * rayinteractive.c:6) to the CUDA backend.  Put -Iinclude/shim before the reference's own src/. */
#pragma once
#include "../cuda_wrap.h"
