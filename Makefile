# Drop-in build: `make` produces ./p2p_matrix, launched as `mpirun -n N p2p_matrix`
# exactly like the reference (reference Makefile:1-2, README.md:5).
#
#   make            engine library + p2p_matrix (+ the MPI shim if the host has no MPI)
#   make oracle     test oracles and baselines under oracle/_ref (needs /root/reference for
#                   the unmodified reference binary; the rest builds anywhere)
#   make tests      helper binaries the pytest suite drives
#   make sass       SASS listings of the engine kernels into profiles/
#   make clean

CUDA_HOME ?= /usr/local/cuda
NVCC      ?= $(CUDA_HOME)/bin/nvcc
CXX       ?= g++
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVCCFLAGS := $(ARCH) -std=c++17 -O3 -lineinfo --extended-lambda -Xcompiler -fPIC,-fvisibility=hidden,-Wall
PKG       := test-nccl-p2p_b200
CSRC      := $(PKG)/csrc
LIBDIR    := $(PKG)/lib
BUILD     := build

# ---- MPI: prefer a real one, fall back to the in-repo shim (this image has none) ----
MPICXX_FOUND := $(shell command -v mpicxx 2>/dev/null)
ifneq ($(MPICXX_FOUND),)
  MPI_INC   := $(shell mpicxx --showme:compile 2>/dev/null || mpicxx -compile_info 2>/dev/null | cut -d' ' -f2-)
  MPI_LIB   := $(shell mpicxx --showme:link 2>/dev/null || mpicxx -link_info 2>/dev/null | cut -d' ' -f2-)
  MPI_DEP   :=
  MPIRUN    := mpirun
  MPI_LABEL := system MPI ($(MPICXX_FOUND))
else ifneq ($(MPI_HOME),)
  MPI_INC   := -I$(MPI_HOME)/include
  MPI_LIB   := -L$(MPI_HOME)/lib -lmpi
  MPI_DEP   :=
  MPIRUN    := $(MPI_HOME)/bin/mpirun
  MPI_LABEL := MPI_HOME=$(MPI_HOME)
else
  MPI_INC   := -Ithird_party/minimpi
  MPI_LIB   := -Lthird_party/minimpi -lmpi -lpthread
  MPI_DEP   := third_party/minimpi/libmpi.a third_party/minimpi/bin/mpirun
  MPIRUN    := third_party/minimpi/bin/mpirun
  MPI_LABEL := minimpi (in-repo shim)
endif
export MPI_INC MPI_LIB MPIRUN

ENGINE_SRCS := $(CSRC)/p2p_engine.cu $(CSRC)/p2p_kernels.cu $(CSRC)/p2p_symm.cu $(CSRC)/p2p_fdpass.cu
ENGINE_HDRS := include/p2p.h $(CSRC)/p2p_debug.h $(CSRC)/p2p_internal.h $(CSRC)/p2p_device.cuh $(CSRC)/p2p_tma.cuh
ENGINE_OBJS := $(BUILD)/p2p_engine.o $(BUILD)/p2p_kernels.o $(BUILD)/p2p_symm.o $(BUILD)/p2p_fdpass.o

all: $(LIBDIR)/libb200p2p.so p2p_matrix mpirun-link
	@echo "built with: $(MPI_LABEL)"

$(BUILD)/%.o: $(CSRC)/%.cu $(ENGINE_HDRS)
	@mkdir -p $(BUILD)
	$(NVCC) $(NVCCFLAGS) -Iinclude -I$(CSRC) -c $< -o $@

$(LIBDIR)/libb200p2p.so: $(ENGINE_OBJS)
	@mkdir -p $(LIBDIR)
	$(NVCC) $(ARCH) -shared -o $@ $^ -Xlinker -soname=libb200p2p.so

$(BUILD)/libb200p2p.a: $(ENGINE_OBJS)
	ar rcs $@ $^

third_party/minimpi/libmpi.a third_party/minimpi/bin/mpirun: third_party/minimpi/minimpi.c third_party/minimpi/mpirun.c third_party/minimpi/mpi.h third_party/minimpi/minimpi_shm.h
	$(MAKE) -C third_party/minimpi

# the benchmark: host code is plain C++ over MPI, linked against the engine objects
p2p_matrix: $(PKG)/host/p2p_matrix.cc $(PKG)/host/matrix_driver.hpp include/p2p.h include/p2p_mpi.h $(BUILD)/libb200p2p.a $(MPI_DEP)
	$(NVCC) $(ARCH) -std=c++17 -O2 -Iinclude -I$(PKG)/host $(MPI_INC) $< $(BUILD)/libb200p2p.a $(MPI_LIB) -o $@

# a stable path to whichever launcher this build uses
mpirun-link: $(MPI_DEP)
	@mkdir -p $(BUILD)
	@if [ -x "$(MPIRUN)" ]; then ln -sf "$(abspath $(MPIRUN))" $(BUILD)/mpirun; else ln -sf "$$(command -v $(MPIRUN))" $(BUILD)/mpirun; fi

oracle: $(MPI_DEP) $(BUILD)/libb200p2p.a
	$(MAKE) -C oracle

tests: $(BUILD)/minimpi_selftest $(BUILD)/driver_unit $(BUILD)/mpi_glue_nogpu $(BUILD)/fdpass_unit

$(BUILD)/fdpass_unit: tests/csrc/fdpass_unit.cu $(CSRC)/p2p_internal.h $(BUILD)/libb200p2p.a
	@mkdir -p $(BUILD)
	$(NVCC) $(ARCH) -std=c++17 -O1 -Iinclude -I$(CSRC) $< $(BUILD)/libb200p2p.a -o $@

$(BUILD)/mpi_glue_nogpu: tests/csrc/mpi_glue_nogpu.cc include/p2p.h include/p2p_mpi.h $(BUILD)/libb200p2p.a $(MPI_DEP)
	@mkdir -p $(BUILD)
	$(NVCC) $(ARCH) -std=c++17 -O1 -Iinclude $(MPI_INC) $< $(BUILD)/libb200p2p.a $(MPI_LIB) -o $@

$(BUILD)/driver_unit: tests/csrc/driver_unit.cc $(PKG)/host/matrix_driver.hpp $(MPI_DEP)
	@mkdir -p $(BUILD)
	$(NVCC) -std=c++17 -O1 -I$(PKG)/host $(MPI_INC) $< $(MPI_LIB) -o $@

$(BUILD)/minimpi_selftest: tests/csrc/minimpi_selftest.c third_party/minimpi/libmpi.a
	@mkdir -p $(BUILD)
	gcc -O2 -Wall -Ithird_party/minimpi $< third_party/minimpi/libmpi.a -lpthread -o $@

sass: $(LIBDIR)/libb200p2p.so
	@mkdir -p profiles
	$(CUDA_HOME)/bin/cuobjdump -sass $(LIBDIR)/libb200p2p.so > profiles/sass_libb200p2p.txt
	@grep -cE "UBLKCP|SYNCS|STG.E.*(128|256)|LDG.E.*(128|256)|MEMBAR.ALL.SYS|STRONG.SYS" profiles/sass_libb200p2p.txt

clean:
	rm -rf $(BUILD) $(LIBDIR) p2p_matrix
	$(MAKE) -C third_party/minimpi clean
	-$(MAKE) -C oracle clean

.PHONY: all oracle tests sass clean mpirun-link experiments

# experiments and profiling harnesses (not part of the product; see scripts/README.md)
experiments: $(BUILD)/exp_probe $(BUILD)/exp_peer_bw $(BUILD)/prof_peer $(BUILD)/exp_ce_chain
$(BUILD)/exp_probe: scripts/exp_probe.cu
	@mkdir -p $(BUILD)
	$(NVCC) $(ARCH) -std=c++17 -O3 -o $@ $< -L$(CUDA_HOME)/lib64/stubs -lcuda
$(BUILD)/exp_peer_bw: scripts/exp_peer_bw.cu
	@mkdir -p $(BUILD)
	$(NVCC) $(ARCH) -std=c++20 -O3 -o $@ $<
$(BUILD)/prof_peer: scripts/prof_peer.cu $(BUILD)/libb200p2p.a $(ENGINE_HDRS)
	@mkdir -p $(BUILD)
	$(NVCC) $(ARCH) -std=c++17 -O3 -lineinfo -Iinclude -I$(CSRC) -o $@ $< $(BUILD)/libb200p2p.a
$(BUILD)/exp_ce_chain: scripts/exp_ce_chain.cu $(BUILD)/libb200p2p.a $(MPI_DEP)
	@mkdir -p $(BUILD)
	$(NVCC) $(ARCH) -std=c++17 -O3 --extended-lambda -Iinclude $(MPI_INC) -o $@ $< $(BUILD)/libb200p2p.a $(MPI_LIB) -L$(CUDA_HOME)/lib64/stubs -lcuda
