# Builds the product library in-tree (it travels to the GPU box with the snapshot):
#   example-gui-opencl-raytracer_b200/librtb200.so   engine C ABI (include/rtb200.h) + cl_wrap_* drop-in (include/cuda_wrap.h)
# The render kernels are compiled WITHOUT contraction (-fmad=false): parity contract, see rt_kernels.cu.
PKG   := example-gui-opencl-raytracer_b200
CSRC  := $(PKG)/csrc
NVCC  ?= nvcc
ARCH  := -gencode arch=compute_100a,code=sm_100a
NVFLAGS := $(ARCH) -O3 -lineinfo -std=c++17 --extended-lambda -Xcompiler -fPIC -Xptxas -v
KFLAGS  := -fmad=false -prec-div=true -prec-sqrt=true -ftz=false
CFLAGS  := -O2 -std=gnu11 -fPIC -Wall -Wextra -ffp-contract=off -Iinclude -Iinclude/shim

OBJS := $(CSRC)/rt_kernels.o $(CSRC)/rt_kernels_tri.o $(CSRC)/rt_kernels_accel.o $(CSRC)/rt_kernels_tri_accel.o $(CSRC)/rt_engine.o $(CSRC)/cuda_wrap.o $(CSRC)/png_io.o $(CSRC)/map_io.o

all: $(PKG)/librtb200.so

$(CSRC)/rt_kernels.o: $(CSRC)/rt_kernels.cu $(CSRC)/rt_kernels.h $(CSRC)/rt_device.cuh $(CSRC)/rt_trace_fast.cuh $(CSRC)/rt_trace_coop.cuh $(CSRC)/rt_trace_tree.cuh $(CSRC)/rt_trace_literal.cuh include/rtb200.h
	$(NVCC) $(NVFLAGS) $(KFLAGS) -c $< -o $@ 2> $(CSRC)/rt_kernels.ptxas.log || (cat $(CSRC)/rt_kernels.ptxas.log; false)
	@grep -E "registers|spill" $(CSRC)/rt_kernels.ptxas.log | sort | uniq -c | sort -rn | head -20

KDEPS := $(CSRC)/rt_kernels.cu $(CSRC)/rt_kernels.h $(CSRC)/rt_device.cuh $(CSRC)/rt_trace_fast.cuh $(CSRC)/rt_trace_coop.cuh $(CSRC)/rt_trace_tree.cuh $(CSRC)/rt_trace_literal.cuh include/rtb200.h
$(CSRC)/rt_kernels_tri.o: $(CSRC)/rt_kernels_tri.cu $(KDEPS)
	$(NVCC) $(NVFLAGS) $(KFLAGS) -c $< -o $@ 2> $(CSRC)/rt_kernels_tri.ptxas.log || (cat $(CSRC)/rt_kernels_tri.ptxas.log; false)

$(CSRC)/rt_kernels_accel.o: $(CSRC)/rt_kernels_accel.cu $(KDEPS) $(CSRC)/rt_trace_accel.cuh
	$(NVCC) $(NVFLAGS) $(KFLAGS) -c $< -o $@ 2> $(CSRC)/rt_kernels_accel.ptxas.log || (cat $(CSRC)/rt_kernels_accel.ptxas.log; false)

$(CSRC)/rt_kernels_tri_accel.o: $(CSRC)/rt_kernels_tri_accel.cu $(KDEPS) $(CSRC)/rt_trace_accel.cuh
	$(NVCC) $(NVFLAGS) $(KFLAGS) -c $< -o $@ 2> $(CSRC)/rt_kernels_tri_accel.ptxas.log || (cat $(CSRC)/rt_kernels_tri_accel.ptxas.log; false)

$(CSRC)/rt_engine.o: $(CSRC)/rt_engine.cu $(CSRC)/rt_kernels.h include/rtb200.h
	$(NVCC) $(ARCH) -O2 -lineinfo -std=c++17 -Xcompiler -fPIC -c $< -o $@

$(CSRC)/cuda_wrap.o: $(CSRC)/cuda_wrap.c include/cuda_wrap.h include/rtb200.h $(CSRC)/png_io.h
	gcc $(CFLAGS) -c $< -o $@

$(CSRC)/png_io.o: $(CSRC)/png_io.c $(CSRC)/png_io.h
	gcc $(CFLAGS) -c $< -o $@

$(CSRC)/map_io.o: $(CSRC)/map_io.c include/rtb200.h
	gcc $(CFLAGS) -D_POSIX_C_SOURCE=200809L -c $< -o $@

$(PKG)/librtb200.so: $(OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -lz

# diagnostic build: per-tile wall time in the hit-id AOV (tools/exp/tile_times.py, RTB_LIB=...)
timing: all
	$(NVCC) $(NVFLAGS) $(KFLAGS) -DRT_TILE_TIMING -c $(CSRC)/rt_kernels.cu -o /tmp/rt_kernels_timing.o 2> /dev/null
	$(NVCC) $(ARCH) -shared -o $(PKG)/librtb200_timing.so /tmp/rt_kernels_timing.o $(CSRC)/rt_kernels_tri.o $(CSRC)/rt_kernels_accel.o $(CSRC)/rt_kernels_tri_accel.o $(CSRC)/rt_engine.o $(CSRC)/cuda_wrap.o $(CSRC)/png_io.o $(CSRC)/map_io.o -lz

# A/B build of the render kernels with extra defines: make variant NAME=narrow DEFS=-DRT_COOP_WIDE=0 ; RTB_LIB=.../librtb200_narrow.so
variant: all
	$(NVCC) $(NVFLAGS) $(KFLAGS) $(DEFS) -c $(CSRC)/rt_kernels.cu -o /tmp/rt_kernels_$(NAME).o 2> /dev/null
	$(NVCC) $(NVFLAGS) $(KFLAGS) $(DEFS) -c $(CSRC)/rt_kernels_tri.cu -o /tmp/rt_kernels_tri_$(NAME).o 2> /dev/null
	$(NVCC) $(ARCH) -shared -o $(PKG)/librtb200_$(NAME).so /tmp/rt_kernels_$(NAME).o /tmp/rt_kernels_tri_$(NAME).o $(CSRC)/rt_kernels_accel.o $(CSRC)/rt_kernels_tri_accel.o $(CSRC)/rt_engine.o $(CSRC)/cuda_wrap.o $(CSRC)/png_io.o $(CSRC)/map_io.o -lz

oracle:
	$(MAKE) -C oracle

clean:
	rm -f $(OBJS) $(PKG)/librtb200.so $(CSRC)/*.ptxas.log
.PHONY: all oracle clean timing variant
