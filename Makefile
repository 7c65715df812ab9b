# Builds the CUDA library in-tree (sm_100a only) and the C oracle pieces.
NVCC ?= nvcc
NVCCFLAGS = -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC
LIB = vega_b200/libvega_b200.so
SRC = vega_b200/csrc/vega_b200.cu
HDR = $(wildcard vega_b200/csrc/*.cuh) include/vega_b200.h

all: $(LIB)

$(LIB): $(SRC) $(HDR)
	$(NVCC) $(NVCCFLAGS) -shared -o $@ $(SRC)

clean:
	rm -f $(LIB)
