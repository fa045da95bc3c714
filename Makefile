# Build for NVIDIA B200 (sm_100a) only.  Mirrors the reference's Makefile targets
# (/root/reference/Makefile:22-33: build/libflash_attention.so + build/tests) with nvcc
# instead of g++/Poplar.  No Triton, no multi-arch fatbins, no CPU fallback.
NVCC    ?= nvcc
CC      ?= gcc
ARCH    ?= -gencode arch=compute_100a,code=sm_100a
OUT     ?= build
NVFLAGS ?= -std=c++17 -O3 -lineinfo $(ARCH) -Xcompiler -fPIC,-Wall -Xptxas -v
CSRC    := flash_attention_ipu_b200/csrc
HDRS    := $(wildcard $(CSRC)/*.cuh) include/fa_b200.h

all: $(OUT)/libflash_attention.so $(OUT)/libfa_torch_ops.so $(OUT)/liboracle.so $(OUT)/tests $(OUT)/umma_probe $(OUT)/umma2_probe $(OUT)/umma_rate $(OUT)/exp_rate

$(OUT)/libflash_attention.so: $(CSRC)/fa_api.cu $(HDRS)
	@mkdir -p $(OUT)
	$(NVCC) $(NVFLAGS) -shared -o $@ $< 2> $(OUT)/ptxas_fa_api.log || (cat $(OUT)/ptxas_fa_api.log; false)
	@grep -E "error|spill" $(OUT)/ptxas_fa_api.log | sort | uniq -c | head -20 || true

# the torch.library ops (schema, autograd, CUDA implementation over the C ABI), plain g++: no kernels
PYTHON    ?= python
TORCH_INC := $(shell $(PYTHON) -c "import torch.utils.cpp_extension as c; print(' '.join('-isystem ' + p for p in c.include_paths()))" 2>/dev/null)
TORCH_LIB := $(shell $(PYTHON) -c "import os, torch; print(os.path.join(os.path.dirname(torch.__file__), 'lib'))" 2>/dev/null)
CUDA_HOME ?= /usr/local/cuda
$(OUT)/libfa_torch_ops.so: $(CSRC)/torch_ops.cpp include/fa_b200.h $(OUT)/libflash_attention.so
	g++ -std=c++17 -O2 -fPIC -shared -Wall -D_GLIBCXX_USE_CXX11_ABI=1 $(TORCH_INC) -isystem $(CUDA_HOME)/include \
	    -o $@ $< -L$(OUT) -lflash_attention -L$(TORCH_LIB) -ltorch -ltorch_cpu -lc10 -lc10_cuda \
	    -Wl,-rpath,'$$ORIGIN' -Wl,-rpath,$(TORCH_LIB)

# the CPU oracle (test infrastructure, never linked into the product library)
$(OUT)/liboracle.so: oracle/vanilla_attention.c
	@mkdir -p $(OUT)
	$(CC) -O2 -fPIC -shared -Wall -Wextra -o $@ $< -lm

# The C++ test binary (reference: Makefile:25-33 builds build/tests from Catch2's amalgamated
# sources).  CATCH2_DIR=<a Catch2 v3 checkout> builds against the real thing; without it the
# TEST_CASE / REQUIRE / "[tag]" surface comes from tests/cuda/catch_shim.hpp.
CATCH2_DIR ?=
ifneq ($(CATCH2_DIR),)
CATCH_FLAGS := -DFA_USE_CATCH2 -I$(CATCH2_DIR)/extras
CATCH_SRC   := $(CATCH2_DIR)/extras/catch_amalgamated.cpp
endif
$(OUT)/tests: tests/cuda/test_flash_attention.cu tests/cuda/catch_shim.hpp oracle/vanilla_attention.c $(OUT)/libflash_attention.so
	$(CC) -O2 -c -o $(OUT)/vanilla_attention.o oracle/vanilla_attention.c
	$(NVCC) -std=c++17 -O2 $(ARCH) -Iinclude $(CATCH_FLAGS) -o $@ tests/cuda/test_flash_attention.cu $(CATCH_SRC) \
	    $(OUT)/vanilla_attention.o -L$(OUT) -lflash_attention -Xlinker -rpath -Xlinker '$$ORIGIN'

$(OUT)/umma_probe: tests/cuda/umma_probe.cu $(CSRC)/sm100_ptx.cuh
	@mkdir -p $(OUT)
	$(NVCC) -std=c++17 -O2 -lineinfo $(ARCH) -o $@ $<

# CTA-pair (cta_group::2) probes: accumulator layouts, operand splits, issue rates
$(OUT)/umma2_probe: tests/cuda/umma2_probe.cu $(CSRC)/sm100_ptx.cuh
	@mkdir -p $(OUT)
	$(NVCC) -std=c++17 -O2 -lineinfo $(ARCH) -o $@ $<

# microbenchmarks behind DESIGN.md section 4.5 (tensor-pipe issue rate, exponential throughput)
$(OUT)/umma_rate: tests/cuda/umma_rate.cu $(CSRC)/sm100_ptx.cuh
	@mkdir -p $(OUT)
	$(NVCC) -std=c++17 -O2 -lineinfo $(ARCH) -o $@ $<

$(OUT)/exp_rate: tests/cuda/exp_rate.cu $(CSRC)/sm100_ptx.cuh
	@mkdir -p $(OUT)
	$(NVCC) -std=c++17 -O3 -lineinfo $(ARCH) -o $@ $<

# SASS evidence that the hot kernels are tcgen05 / TMEM / TMA code (mnemonics: B200_PROFILING.md)
sass-summary: $(OUT)/libflash_attention.so
	@mkdir -p profiles
	@{ echo "# cuobjdump -sass $(OUT)/libflash_attention.so | grep -c <mnemonic>   (make sass-summary)"; \
	   cuobjdump -sass $(OUT)/libflash_attention.so > $(OUT)/lib.sass; \
	   for m in UTCHMMA UTCBAR LDTM STTM UTMALDG UTMASTG UTMAREDG UTMAPF UBLKCP SYNCS.ARRIVE USETMAXREG FFMA2 FADD2 FMUL2 MUFU.EX2; do \
	     printf "%-14s %s\n" $$m $$(grep -c "$$m" $(OUT)/lib.sass); done; \
	   echo "# kernels:"; cuobjdump -sass $(OUT)/libflash_attention.so | grep -oE "Function : [A-Za-z0-9_]+" | sed 's/Function : //' | c++filt | grep -E "sm100|prep|convert" | sort | uniq -c; \
	 } > profiles/sass_summary.txt; cat profiles/sass_summary.txt

clean:
	rm -rf $(OUT)

.PHONY: all clean sass-summary
