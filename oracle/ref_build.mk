# Builds the REFERENCE GPU library (vendored tfhe-cuda-backend, the library Concrete's runtime
# links: compilers/concrete-compiler/compiler/lib/Runtime/CMakeLists.txt:35) from its sources
# where they lie under /root/reference, into oracle/_ref/ (git-ignored, not gpurun-ignored).
# Test/bench infrastructure only: a second parity witness (its keyswitch must equal ours bit for
# bit; its PBS must decrypt like ours) and "the kernel to beat" timed on the same box.
# Only the classic-PBS / keyswitch / device translation units are compiled (no multi-bit PBS, no
# integer ops); flags follow TCB/CMakeLists.txt:60-92 with the arch pinned to sm_100.
# The reference's CPU path (concrete-cpu, Rust) cannot be built in this image.
REF   := /root/reference/third_party/tfhe-rs/backends/tfhe-cuda-backend/cuda
OUT   := _ref
NVCC  ?= /usr/local/cuda/bin/nvcc
SRCS  := device.cu crypto/keyswitch.cu pbs/programmable_bootstrap_classic.cu \
         pbs/programmable_bootstrap_amortized.cu pbs/bootstraping_key.cu \
         pbs/programmable_bootstrap.cu fft/twiddles.cu utils/helper_multi_gpu.cu \
         linearalgebra/addition.cu linearalgebra/multiplication.cu linearalgebra/negation.cu
OBJS  := $(patsubst %.cu,$(OUT)/%.o,$(subst /,_,$(SRCS)))
FLAGS := -gencode arch=compute_100,code=sm_100 -O3 -std=c++17 --no-exceptions \
         --expt-relaxed-constexpr -rdc=true --use_fast_math -Xcompiler -fPIC -Xcompiler -fopenmp \
         -DCUDA_ARCH=1000 -I$(REF)/include -I$(REF)/src -ccbin /usr/bin/g++ -w

all: $(OUT)/libtfhe_cuda_backend_ref.so

$(OUT)/libtfhe_cuda_backend_ref.so: $(OBJS)
	$(NVCC) -shared -gencode arch=compute_100,code=sm_100 -Xcompiler -fPIC -o $@ $(OBJS) \
	    -Xlinker -Bsymbolic -lgomp -ccbin /usr/bin/g++

define RULE
$(OUT)/$(subst /,_,$(1:.cu=.o)): $(REF)/src/$(1)
	@mkdir -p $(OUT)
	$(NVCC) $(FLAGS) -c $$< -o $$@
endef
$(foreach s,$(SRCS),$(eval $(call RULE,$(s))))

clean:
	rm -rf $(OUT)
