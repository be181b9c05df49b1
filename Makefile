# Native build of b200bench for sm_100a.  `make` -> b200bench/lib/libb200bench.so (in-tree, so it
# travels to the GPU box) and the standalone kernel harnesses under build/bin/.
NVCC      ?= nvcc
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVCCFLAGS := $(ARCH) -O3 -std=c++17 -lineinfo --use_fast_math -Xcompiler -fPIC -Xcompiler -Wall
SRCS      := $(wildcard csrc/*.cu)
OBJS      := $(patsubst csrc/%.cu,build/obj/%.o,$(SRCS))
HDRS      := $(wildcard csrc/*.h csrc/*.cuh include/*.h)
LIB       := b200bench/lib/libb200bench.so
TESTS     := build/bin/test_gemm build/bin/test_attn build/bin/test_sanitize

all: $(LIB) $(TESTS)

build/obj/%.o: csrc/%.cu $(HDRS)
	@mkdir -p build/obj
	$(NVCC) $(NVCCFLAGS) -c $< -o $@

$(LIB): $(OBJS)
	@mkdir -p b200bench/lib
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -lcudart

build/bin/%: csrc/tests/%.cu $(OBJS) $(wildcard csrc/tests/*.cu csrc/tests/*.h)
	@mkdir -p build/bin
	$(NVCC) $(NVCCFLAGS) $< $(OBJS) -o $@ -ldl -lpthread

ptxas-info:
	@for f in $(SRCS); do echo "== $$f"; $(NVCC) $(NVCCFLAGS) -Xptxas -v -c $$f -o /dev/null 2>&1 | grep -E "registers|spill" ; done

# Debug build of the collectives: a timed-out flag wait prints rank / CTA / flag plane / expected and
# last seen value before it traps.  Separate library, the default build is unchanged:
#   make debug-comm && B200BENCH_LIB=b200bench/lib/libb200bench_debug.so python ...
debug-comm: $(OBJS)
	@mkdir -p build/obj_debug b200bench/lib
	$(NVCC) $(NVCCFLAGS) -DB200_COMM_DEBUG -c csrc/comm.cu -o build/obj_debug/comm.o
	$(NVCC) $(ARCH) -shared -o b200bench/lib/libb200bench_debug.so build/obj_debug/comm.o $(filter-out build/obj/comm.o,$(OBJS)) -lcudart

# compute-sanitizer targets (need a GPU): every single-GPU kernel at small shapes, see csrc/tests/test_sanitize.cu
SANITIZER ?= compute-sanitizer
memcheck: build/bin/test_sanitize
	$(SANITIZER) --tool memcheck --error-exitcode 9 build/bin/test_sanitize
synccheck: build/bin/test_sanitize
	$(SANITIZER) --tool synccheck --error-exitcode 9 build/bin/test_sanitize
initcheck: build/bin/test_sanitize
	$(SANITIZER) --tool initcheck --error-exitcode 9 build/bin/test_sanitize

clean:
	rm -rf build b200bench/lib

.PHONY: all clean ptxas-info memcheck synccheck initcheck debug-comm
