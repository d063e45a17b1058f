# Builds libb200df.so (sm_100a only) in-tree; `python __graft_entry__.py` / build() calls this.
NVCC ?= /usr/local/cuda/bin/nvcc
SRC_DIR := deepfakedetection_b200/csrc
OUT := deepfakedetection_b200/libb200df.so
SRCS := $(SRC_DIR)/engine.cu $(SRC_DIR)/stem.cu $(SRC_DIR)/stem_tc.cu $(SRC_DIR)/dwconv.cu $(SRC_DIR)/dwconv_tma.cu $(SRC_DIR)/se.cu \
        $(SRC_DIR)/pwconv_simt.cu $(SRC_DIR)/conv3x3.cu $(SRC_DIR)/conv3x3_tc.cu $(SRC_DIR)/attn.cu $(SRC_DIR)/attn_mma.cu $(SRC_DIR)/vit_ops.cu $(SRC_DIR)/cam_bwd.cu $(SRC_DIR)/pwconv_tc.cu $(SRC_DIR)/tail.cu $(SRC_DIR)/metrics.cu $(SRC_DIR)/overlay.cu
OBJS := $(SRCS:.cu=.o)
HDRS := $(wildcard $(SRC_DIR)/*.h) $(wildcard $(SRC_DIR)/*.cuh) $(wildcard $(SRC_DIR)/*.inc) include/b200df.h
NVFLAGS := -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC \
           -Xcompiler -fvisibility=hidden --expt-relaxed-constexpr -cudart static

all: $(OUT)

%.o: %.cu $(HDRS)
	$(NVCC) $(NVFLAGS) -Xptxas -v -c $< -o $@ 2> $@.ptxas.log || (cat $@.ptxas.log; exit 1)

$(OUT): $(OBJS)
	$(NVCC) -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o $@ $(OBJS)

# dev build with per-role wait-cycle counters in the tcgen05 GEMM (tools/tc_prof.py)
PROF_OUT := deepfakedetection_b200/libb200df_prof.so
prof: $(OBJS)
	$(NVCC) $(NVFLAGS) -DTC_PROF -c $(SRC_DIR)/pwconv_tc.cu -o $(SRC_DIR)/pwconv_tc_prof.o
	$(NVCC) -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o $(PROF_OUT) \
	    $(filter-out $(SRC_DIR)/pwconv_tc.o,$(OBJS)) $(SRC_DIR)/pwconv_tc_prof.o

clean:
	rm -f $(OBJS) $(SRC_DIR)/*.ptxas.log $(SRC_DIR)/pwconv_tc_prof.o $(OUT) $(PROF_OUT)

.PHONY: all clean prof
