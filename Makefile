# Build of the flextree-b200 native libraries and tools.
#   make host     -> build/libflextree_host.so  (planner + CPU world; g++ only, runs anywhere)
#   make cuda     -> allreduce-over-mpi_b200/lib/libflextree.so (everything; nvcc, sm_100a)
#   make tools    -> build/ft_plan, build/ft_bench (--comm-type nccl is the NCCL harness), build/ftrun
# Capability parity with the reference build files: allreduce_over_mpi/CMakeLists.txt (flags, version
# stamping), allreduce_over_mpi/Makefile (build/clean), cost_model/CMakeLists.txt.
NVCC ?= nvcc
# The image exports CXX=/opt/gcc/bin/g++, a wrapper that links libstdc++ statically; a second copy of
# libstdc++ inside a dlopen()ed library crashes once numpy/torch have loaded the system one.
CXX := $(shell test -x /usr/bin/g++ && echo /usr/bin/g++ || echo g++)
PKG := allreduce-over-mpi_b200
LIBDIR := $(PKG)/lib
B := build
ARCH := -gencode arch=compute_100a,code=sm_100a

GIT_DESCRIBE := $(shell git describe --always --dirty 2>/dev/null || echo unknown)
GIT_DATE := $(shell git log -1 --format=%cd --date=short 2>/dev/null || echo unknown)
GIT_HASH := $(shell git rev-parse --short HEAD 2>/dev/null || echo unknown)
VERDEFS := -DFT_GIT_DESCRIBE='"$(GIT_DESCRIBE)"' -DFT_GIT_DATE='"$(GIT_DATE)"' -DFT_GIT_HASH='"$(GIT_HASH)"'

CXXFLAGS := -std=c++17 -O2 -fPIC -Wall -Wextra -Wno-unknown-pragmas -Wno-maybe-uninitialized -Wno-uninitialized -Iinclude $(VERDEFS)
NVFLAGS := -ccbin /usr/bin/g++ -std=c++17 -O3 $(ARCH) -lineinfo -Xcompiler -fPIC -Xcompiler -Wall -Iinclude $(VERDEFS) --expt-relaxed-constexpr

PLANNER_SRC := csrc/planner/factorize.cc csrc/planner/cost_model.cc csrc/planner/schedule.cc csrc/planner/c_api.cc
HOST_SRC := csrc/host/host_world.cc csrc/host/host_comm.cc csrc/host/host_run_all.cc $(wildcard csrc/host/host_run_g*.cc)
CUDA_SRC := $(wildcard csrc/kernels/*.cu) $(wildcard csrc/runtime/*.cu)
RT_CC_SRC := $(wildcard csrc/runtime/*.cc)

PLANNER_OBJ := $(PLANNER_SRC:%.cc=$(B)/%.o)
HOST_OBJ := $(HOST_SRC:%.cc=$(B)/%.o)
CUDA_OBJ := $(CUDA_SRC:%.cu=$(B)/%.cu.o)
RT_CC_OBJ := $(RT_CC_SRC:%.cc=$(B)/%.o)

.PHONY: all host cuda tools clean tsan asan fdtest memcheck racecheck dropin
all: host cuda tools
host: $(B)/libflextree_host.so
cuda: $(LIBDIR)/libflextree.so $(LIBDIR)/_ftfast.so
tools: $(B)/ft_plan $(B)/ft_bench $(B)/ftrun

$(B)/%.o: %.cc
	@mkdir -p $(dir $@)
	$(CXX) $(CXXFLAGS) -MMD -MP -c $< -o $@

# the CPU data plane: -O3 so that the bulk-tile loops (csrc/host/cpu_backend.h) are vectorised (at -O2 gcc's
# "very-cheap" cost model skips every loop that needs an epilogue); baseline x86-64 ISA only - the library
# travels to other machines
$(B)/csrc/host/host_run_g%.o: csrc/host/host_run_g%.cc
	@mkdir -p $(dir $@)
	$(CXX) $(CXXFLAGS) -O3 -MMD -MP -c $< -o $@

$(B)/%.cu.o: %.cu
	@mkdir -p $(dir $@)
	$(NVCC) $(NVFLAGS) -MMD -MP -c $< -o $@

$(B)/libflextree_host.so: $(PLANNER_OBJ) $(HOST_OBJ)
	$(CXX) -shared -o $@ $^ -lpthread

$(LIBDIR)/libflextree.so: $(PLANNER_OBJ) $(HOST_OBJ) $(RT_CC_OBJ) $(CUDA_OBJ)
	@mkdir -p $(LIBDIR)
	$(NVCC) -ccbin /usr/bin/g++ $(ARCH) -shared -o $@ $^ -lpthread -ldl

# the per-call path of the Python API as a CPython extension (csrc/python/ftfast.c); links the CUDA library
PYINC := $(shell python3 -c "import sysconfig;print(sysconfig.get_paths()['include'])" 2>/dev/null)
$(LIBDIR)/_ftfast.so: csrc/python/ftfast.c $(LIBDIR)/libflextree.so
	/usr/bin/gcc -O2 -fPIC -shared -Wall -Iinclude -I$(PYINC) -o $@ $< -L$(LIBDIR) -lflextree -Wl,-rpath,'$$ORIGIN'

$(B)/ft_plan: csrc/tools/ft_plan.cc $(PLANNER_OBJ)
	$(CXX) $(CXXFLAGS) -o $@ $^

$(B)/ftrun: csrc/tools/ftrun.cc
	@mkdir -p $(B)
	$(CXX) $(CXXFLAGS) -o $@ $<

$(B)/ft_bench: csrc/tools/ft_bench.cu $(LIBDIR)/libflextree.so
	$(NVCC) $(NVFLAGS) -o $@ $< -L$(LIBDIR) -lflextree -ldl -Xlinker -rpath -Xlinker '$$ORIGIN/../$(LIBDIR)'

# ThreadSanitizer run of the algorithm templates on the CPU world (the reference has no
# sanitizer targets at all) and the SCM_RIGHTS fd-channel test.
tsan:
	$(CXX) -std=c++17 -O1 -g -fsanitize=thread -Wno-unknown-pragmas -Wno-tsan -Iinclude -DFT_ONLY_F32 -DFT_WITH_F16 -o $(B)/tsan_world \
	  tests/cpp/tsan_world.cc csrc/host/host_world.cc csrc/host/host_run_all.cc csrc/host/host_run_g1.cc csrc/host/host_run_g2.cc \
	  csrc/planner/factorize.cc csrc/planner/cost_model.cc -lpthread
	$(B)/tsan_world
# AddressSanitizer + UBSan over the same CPU world and over the planner CLI
asan:
	$(CXX) -std=c++17 -O1 -g -fsanitize=address,undefined -fno-sanitize-recover=undefined -Wno-unknown-pragmas -Iinclude -DFT_ONLY_F32 -DFT_WITH_F16 \
	  -o $(B)/asan_world tests/cpp/tsan_world.cc csrc/host/host_world.cc csrc/host/host_run_all.cc csrc/host/host_run_g1.cc \
	  csrc/host/host_run_g2.cc csrc/planner/factorize.cc csrc/planner/cost_model.cc -lpthread
	$(B)/asan_world
	$(CXX) -std=c++17 -O1 -g -fsanitize=address,undefined -fno-sanitize-recover=undefined -Iinclude $(VERDEFS) -o $(B)/asan_plan csrc/tools/ft_plan.cc $(PLANNER_SRC)
	$(B)/asan_plan --sweep --max 300 --out /dev/null > /dev/null
	$(B)/asan_plan --candidates 7 > /dev/null
	$(B)/asan_plan --print-ops --widths 2,3,2 --rank 5 > /dev/null
	$(B)/asan_plan --print-ops --widths 2,3 --rank 6 --lonely 1 > /dev/null
	$(B)/asan_plan --legacy 24 --chunk 100 > /dev/null
	$(B)/asan_plan --choose 8 --bytes 1048576 --nvls > /dev/null
	$(B)/asan_plan --enumerate 64 --legacy-flat > /dev/null
# The reference's unmodified benchmark.cpp (and a plain MPI test program) built against compat/mpi_mod.hpp +
# libflextree_host.so + the in-repo MPI subset -> baseline/_ref/bin/{allreduce_over_mpi_dropin,dropin_user}
dropin: host
	bash baseline/build_ref.sh
fdtest:
	$(CXX) -std=c++17 -O1 -o $(B)/fd_channel_test tests/cpp/fd_channel_test.cc
	$(B)/fd_channel_test 8
# compute-sanitizer on the single-GPU emulation (one tool per invocation, see B200_PROFILING.md)
memcheck:
	compute-sanitizer --tool memcheck python bench/sanitize_target.py
racecheck:
	compute-sanitizer --tool racecheck python bench/sanitize_target.py

clean:
	rm -rf $(B) $(LIBDIR)/*.so

-include $(PLANNER_OBJ:.o=.d) $(HOST_OBJ:.o=.d) $(CUDA_OBJ:.o=.d) $(RT_CC_OBJ:.o=.d)
