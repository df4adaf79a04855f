# Top-level build: the CUDA library (sm_100a only), the C host side and the CPU checkers.
# Same commands as image_compression_b200/build.py / __graft_entry__.build(); nvcc cross-compiles without a GPU.
#   make            libgpu_codec.so + host/ (awi_cli, libawi_gpu_host.so)
#   make oracle     CPU checkers (liboracle.so; the unmodified reference build where /root/reference exists)
#   make test       CPU test suite      make test-gpu   GPU parity suite (needs a B200)
NVCC      ?= nvcc
NVCCFLAGS := -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 --fmad=false -std=c++17 -Xcompiler -fPIC -Xcompiler -O2
CSRC      := image_compression_b200/csrc
LIB       := image_compression_b200/libgpu_codec.so
OBJS      := $(CSRC)/awi_encode.o $(CSRC)/awi_decode.o $(CSRC)/gpu_codec.o
HDRS      := $(CSRC)/codec_math.cuh $(CSRC)/codec_tables.h $(CSRC)/colour_fix_tables.h $(CSRC)/awi_kernels.h include/gpu_codec.h

all: $(LIB) host

$(CSRC)/codec_tables.h $(CSRC)/colour_fix_tables.h: tools/gen_tables.py
	python tools/gen_tables.py

$(CSRC)/%.o: $(CSRC)/%.cu $(HDRS)
	$(NVCC) $(NVCCFLAGS) -c $< -o $@

$(LIB): $(OBJS)
	$(NVCC) -shared -o $@ $(OBJS) -lcudart

host: $(LIB)
	$(MAKE) -C host

oracle:
	$(MAKE) -C oracle

test: all oracle
	python -m pytest tests -x -q -m "not gpu"

test-gpu: all oracle
	python -m pytest tests -x -q -m gpu

clean:
	rm -f $(OBJS) $(LIB)
	$(MAKE) -C host clean

.PHONY: all host oracle test test-gpu clean
