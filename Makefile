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

# make variant NAME=x EXTRA="-DFWD_PPT=1 ...": an A/B build of the kernels -> build/var_x/libgsb_x.so, used with
# GSB_LIB=build/var_x/libgsb_x.so (gslam_b200/_lib.py); the product library is untouched
variant:
	@mkdir -p build/var_$(NAME)
	@for f in $(SRCS); do b=$$(basename $$f .cu); \
	  $(NVCC) $(NVFLAGS) $(EXTRA) -c $$f -o build/var_$(NAME)/$$b.o 2> build/var_$(NAME)/$$b.ptxas.log || (cat build/var_$(NAME)/$$b.ptxas.log; exit 1); done
	$(NVCC) -shared -o build/var_$(NAME)/libgsb_$(NAME).so build/var_$(NAME)/*.o


build/%.o: $(SRC_DIR)/%.cu $(wildcard $(SRC_DIR)/*.cuh) include/gsb.h
	@mkdir -p build
	$(NVCC) $(NVFLAGS) -c $< -o $@ 2> build/$*.ptxas.log || (cat build/$*.ptxas.log; exit 1)

$(LIB): $(OBJS)
	$(NVCC) -shared -o $@ $(OBJS)

clean:
	rm -rf build $(LIB)
.PHONY: all clean variant
