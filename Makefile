# Convenience targets; python __graft_entry__.py does the same builds.
NVCC ?= /usr/local/cuda/bin/nvcc
LIB  := xlayer_prover_b200/libglprover.so
SRC  := $(wildcard xlayer_prover_b200/csrc/*) include/glprover.h

.PHONY: all lib oracle emu test test-gpu bench clean
all: lib oracle
lib: $(LIB)
$(LIB): $(SRC)
	$(NVCC) -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -shared -o $@ xlayer_prover_b200/csrc/capi.cu
oracle:
	$(MAKE) -C oracle all
emu:
	sh tests/cuemu/build_emu.sh
test: lib oracle
	python -m pytest tests -q -m "not gpu"
test-gpu: lib oracle
	python -m pytest tests -q -m gpu
bench: lib oracle
	python bench.py
clean:
	rm -f $(LIB) tests/cuemu/*.so; $(MAKE) -C oracle clean
