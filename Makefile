# Build of the native pieces.  `make` = everything; `make binaries` = control-plane only (no
# CUDA needed); `make kernels` = libpsb_kernels.so for sm_100a.  CMakeLists.txt is equivalent.
CXX      ?= g++
NVCC     ?= nvcc
CXXFLAGS ?= -std=c++17 -O2 -g -Wall -Wextra -Wno-unused-parameter -pthread
SAN      ?=
BIN      := build/bin
OBJ      := build/obj$(if $(SAN),-$(SAN),)
ifneq ($(SAN),)
CXXFLAGS += -fsanitize=$(SAN) -fno-omit-frame-pointer
BIN      := build/bin-$(SAN)
endif

WIRE_SRC := csrc/wire/hpack.cpp csrc/wire/h2.cpp csrc/wire/grpc_lite.cpp
COORD_SRC := csrc/control/coordinator_core.cpp csrc/control/coordinator_service.cpp
PS_SRC := csrc/ps/ps_core.cpp csrc/ps/ps_device.cpp csrc/ps/ps_sparse.cpp csrc/ps/ps_service.cpp
WORKER_SRC := csrc/worker/worker.cpp
HDRS := $(wildcard csrc/*/*.h) include/psb_kernels.h

obj = $(patsubst csrc/%.cpp,$(OBJ)/%.o,$(1))

all: binaries kernels
CONTROL_LIB := parameter_server_distributed_b200/_native/libpsb_control$(if $(SAN),-$(SAN),).so
binaries: $(BIN)/coordinator $(BIN)/parameter_server $(BIN)/worker_main $(BIN)/psb_selftest $(BIN)/psb_pbtool $(CONTROL_LIB)

# native coordinator client of the Python package (C ABI: include/psb_control.h); no CUDA
$(CONTROL_LIB): $(WIRE_SRC) csrc/control/client.cpp $(HDRS) include/psb_control.h
	@mkdir -p $(dir $@)
	$(CXX) $(CXXFLAGS) -fPIC -shared $(WIRE_SRC) csrc/control/client.cpp -o $@

$(OBJ)/%.o: csrc/%.cpp $(HDRS)
	@mkdir -p $(dir $@)
	$(CXX) $(CXXFLAGS) -c $< -o $@

$(BIN)/coordinator: $(call obj,$(WIRE_SRC) $(COORD_SRC) csrc/control/coordinator_main.cpp)
	@mkdir -p $(BIN)
	$(CXX) $(CXXFLAGS) $^ -o $@

$(BIN)/parameter_server: $(call obj,$(WIRE_SRC) $(PS_SRC) csrc/ps/ps_main.cpp)
	@mkdir -p $(BIN)
	$(CXX) $(CXXFLAGS) $^ -ldl -o $@

$(BIN)/worker_main: $(call obj,$(WIRE_SRC) $(WORKER_SRC) csrc/worker/worker_main.cpp)
	@mkdir -p $(BIN)
	$(CXX) $(CXXFLAGS) $^ -ldl -o $@

$(BIN)/psb_selftest: $(call obj,$(WIRE_SRC) $(COORD_SRC) $(PS_SRC) $(WORKER_SRC) csrc/selftest/selftest.cpp)
	@mkdir -p $(BIN)
	$(CXX) $(CXXFLAGS) $^ -ldl -o $@

$(BIN)/psb_pbtool: $(call obj,csrc/selftest/pbtool.cpp)
	@mkdir -p $(BIN)
	$(CXX) $(CXXFLAGS) $^ -o $@

kernels:
	python -c "from parameter_server_distributed_b200 import _build; print(_build.build_kernels())"

# The sanitizer runtimes (libtsan / libasan) ship with the distribution's compiler; a g++ from
# elsewhere on PATH or in $CXX may not have them, so these targets prefer /usr/bin/g++.
SAN_CXX ?= $(if $(wildcard /usr/bin/g++),/usr/bin/g++,$(CXX))
tsan:
	$(MAKE) SAN=thread CXX=$(SAN_CXX) binaries
asan:
	$(MAKE) SAN=address CXX=$(SAN_CXX) binaries

clean:
	rm -rf build parameter_server_distributed_b200/_native/*.so parameter_server_distributed_b200/_native/*.stamp

.PHONY: all binaries kernels tsan asan clean
