// ref_harness.cpp — C ABI around the UNMODIFIED reference classes, for bench.py --impl reference.
//
// Links the reference's own translation units, compiled as they are from /root/reference:
//   src/parameter_server.cpp  (ParameterServerCore: staging, mean, p -= g, serve)
//   src/nccl_manager.cpp      (NCCLManager: clique init, ncclAllReduce float sum)   -DHAVE_NCCL
//   src/coordinator.cpp       (CoordinatorCore)
// The three gRPC binaries cannot be built on this box (no gRPC C++ / protobuf libraries), so the
// worker's call sequence is driven from here through the classes' public methods.  The
// multi-GPU aggregation follows what the reference worker does per gradient tensor
// (src/worker.cpp:408-448): allocate a buffer on every GPU, upload the same host gradient to
// each, one all-reduce per GPU, download from GPU 0, divide by the GPU count on the host, free.
// REF_NCCL_GROUP=1 wraps the per-GPU calls in ncclGroupStart/End (the reference omits the
// group, which can block on NCCL >= 2.22 — SURVEY D4); the variant used is reported.
#include <cuda_runtime.h>
#include <nccl.h>

#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "coordinator.h"
#include "nccl_manager.h"
#include "parameter_server.h"

extern "C" {

void* ref_ps_create(int total_workers) { return new ParameterServerCore(total_workers); }
void ref_ps_destroy(void* ps) { delete static_cast<ParameterServerCore*>(ps); }

// tensors are described by parallel arrays; shapes are flattened ([ndim0 dims..., ndim1 dims...])
static std::vector<tensor> make_tensors(int n, const char** names, const int* ndims,
                                        const int* dims, float** data, const long long* sizes) {
  std::vector<tensor> out(n);
  int d = 0;
  for (int i = 0; i < n; ++i) {
    out[i].name = names[i];
    out[i].shape.assign(dims + d, dims + d + ndims[i]);
    d += ndims[i];
    out[i].dtype = 0;
    out[i].data.assign(data[i], data[i] + sizes[i]);
  }
  return out;
}

void ref_ps_init(void* ps, int n, const char** names, const int* ndims, const int* dims,
                 float** data, const long long* sizes) {
  static_cast<ParameterServerCore*>(ps)->initialize_parameters(
      make_tensors(n, names, ndims, dims, data, sizes));
}

int ref_ps_push(void* ps, int worker_id, int iteration, int n, const char** names,
                const int* ndims, const int* dims, float** data, const long long* sizes) {
  return static_cast<ParameterServerCore*>(ps)->receive_gradients(
             worker_id, iteration, make_tensors(n, names, ndims, dims, data, sizes))
             ? 1
             : 0;
}

// copies the served parameters into caller buffers; returns the number of tensors
int ref_ps_pull(void* ps, int iteration, int n, float** out, const long long* sizes) {
  std::vector<tensor> params = static_cast<ParameterServerCore*>(ps)->serve_parameters(iteration);
  int m = static_cast<int>(params.size()) < n ? static_cast<int>(params.size()) : n;
  for (int i = 0; i < m; ++i) {
    size_t k = params[i].data.size() < static_cast<size_t>(sizes[i]) ? params[i].data.size()
                                                                     : static_cast<size_t>(sizes[i]);
    memcpy(out[i], params[i].data.data(), k * sizeof(float));
  }
  return static_cast<int>(params.size());
}

int ref_ps_sync_status(void* ps, int iteration, int* received) {
  int32_t r = 0;
  bool ok = static_cast<ParameterServerCore*>(ps)->check_sync_status(iteration, r);
  *received = r;
  return ok ? 1 : 0;
}

int ref_detect_num_gpus() { return detect_num_gpus(); }
void* ref_nccl_create() { return new NCCLManager(); }
int ref_nccl_init(void* m, int n) { return static_cast<NCCLManager*>(m)->initialize(n) ? 1 : 0; }
void ref_nccl_destroy(void* m) { delete static_cast<NCCLManager*>(m); }

int ref_nccl_aggregate(void* mgr, float* host, long long count, int use_group) {
  NCCLManager* m = static_cast<NCCLManager*>(mgr);
  const int n = m->get_num_gpus();
  if (!m->is_initialized() || n <= 1) return 0;
  std::vector<float*> bufs(n);
  for (int g = 0; g < n; ++g) {
    cudaSetDevice(g);
    if (cudaMalloc(&bufs[g], count * sizeof(float)) != cudaSuccess) return -1;
    cudaMemcpy(bufs[g], host, count * sizeof(float), cudaMemcpyHostToDevice);
  }
  if (use_group) ncclGroupStart();
  for (int g = 0; g < n; ++g) {
    cudaSetDevice(g);
    if (!m->allreduce_float(bufs[g], static_cast<size_t>(count), g)) return -2;
  }
  if (use_group) ncclGroupEnd();
  cudaSetDevice(0);
  cudaMemcpy(host, bufs[0], count * sizeof(float), cudaMemcpyDeviceToHost);
  for (long long i = 0; i < count; ++i) host[i] /= n;
  for (int g = 0; g < n; ++g) {
    cudaSetDevice(g);
    cudaFree(bufs[g]);
  }
  return 1;
}

// CoordinatorCore smoke (registry semantics oracle for tests)
void* ref_coord_create(const char* host, int port) { return new CoordinatorCore(host, port); }
void ref_coord_destroy(void* c) { delete static_cast<CoordinatorCore*>(c); }
int ref_coord_register(void* c, int worker_id, char* ps_addr_out, int cap) {
  WorkerRegistryEntry e;
  e.worker_id = worker_id;
  e.address = "localhost";
  e.port = 0;
  e.hostname = "worker-" + std::to_string(worker_id);
  e.status = 0;
  std::string addr;
  int32_t total = 0;
  static_cast<CoordinatorCore*>(c)->register_worker(e, addr, total);
  strncpy(ps_addr_out, addr.c_str(), cap - 1);
  ps_addr_out[cap - 1] = 0;
  return total;
}

}  // extern "C"
