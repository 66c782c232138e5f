# Builds libgsb.so (sm_100a only) in-tree so it travels to the GPU box.
NVCC ?= /usr/local/cuda/bin/nvcc
SRC_DIR := gslam_b200/csrc
SRCS := $(wildcard $(SRC_DIR)/*.cu)
OBJS := $(patsubst $(SRC_DIR)/%.cu,build/%.o,$(SRCS))
NVFLAGS := -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 \
           -Xcompiler -fPIC,-fvisibility=hidden -Xptxas -v --expt-relaxed-constexpr
LIB := gslam_b200/libgsb.so
# make TRACE=1: per-unit timing trace of the compositing kernels (tools/trace_units.py)
ifdef TRACE
NVFLAGS += -DGSB_TRACE
endif

all: $(LIB)

build/%.o: $(SRC_DIR)/%.cu $(wildcard $(SRC_DIR)/*.cuh) include/gsb.h
	@mkdir -p build
	$(NVCC) $(NVFLAGS) -c $< -o $@ 2> build/$*.ptxas.log || (cat build/$*.ptxas.log; exit 1)

$(LIB): $(OBJS)
	$(NVCC) -shared -o $@ $(OBJS)

clean:
	rm -rf build $(LIB)
.PHONY: all clean
