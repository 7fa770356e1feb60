// Flag-protocol litmus test ("race detection"): the message-passing pattern every
// cross-GPU exchange of the engine relies on.  A producer kernel on GPU A writes a payload into
// GPU B's memory with plain peer stores, then publishes it with fence + st.release.sys of an epoch
// flag; a consumer kernel on GPU B spins with ld.acquire.sys and checks that the payload it then
// reads belongs to the epoch it saw.  Any reordering shows up as a non-zero error count.
#include "common.cuh"
#include "../include/jordan_b200.h"

using namespace jb;

namespace {

constexpr int LIT_WORDS = 64;  // payload: 64 doubles = 512 bytes per epoch

__global__ void k_litmus_producer(double* payload_remote, uint32_t* flag_remote,
                                  const uint32_t* ack_local, int iters) {
  for (int it = 1; it <= iters; ++it) {
    // wait until the consumer has acknowledged the previous epoch (so the slot may be reused)
    if (threadIdx.x == 0) wait_flag_ge(ack_local, (uint32_t)(it - 1));
    __syncthreads();
    if (threadIdx.x < LIT_WORDS) payload_remote[threadIdx.x] = (double)it * 1000.0 + threadIdx.x;
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) st_release_sys_u32(flag_remote, (uint32_t)it);
  }
}

__global__ void k_litmus_consumer(const double* payload_local, const uint32_t* flag_local,
                                  uint32_t* ack_remote, int iters, unsigned long long* errors) {
  for (int it = 1; it <= iters; ++it) {
    if (threadIdx.x == 0) wait_flag_ge(flag_local, (uint32_t)it);
    __syncthreads();
    if (threadIdx.x < LIT_WORDS) {
      const double v = *(volatile const double*)&payload_local[threadIdx.x];
      if (v != (double)it * 1000.0 + threadIdx.x) atomicAdd(errors, 1ull);
    }
    __syncthreads();
    if (threadIdx.x == 0) st_release_sys_u32(ack_remote, (uint32_t)it);
  }
}

}  // namespace

// Runs the litmus between two devices of THIS process (peer access is enabled here).
// Returns JB_OK and the number of ordering violations in *errors_out.
extern "C" int jb_flag_litmus(int dev_producer, int dev_consumer, int iters,
                              unsigned long long* errors_out) {
  int can = 0;
  JB_CHECK(cudaDeviceCanAccessPeer(&can, dev_producer, dev_consumer));
  if (!can) {
    jb_set_error("jb_flag_litmus: devices %d and %d are not peers", dev_producer, dev_consumer);
    return JB_ERR_ARG;
  }
  double* payload = nullptr;      // lives on the consumer
  uint32_t *flag = nullptr, *ack = nullptr;
  unsigned long long* errs = nullptr;
  JB_CHECK(cudaSetDevice(dev_consumer));
  cudaDeviceEnablePeerAccess(dev_producer, 0);
  cudaGetLastError();
  JB_CHECK(cudaMalloc((void**)&payload, LIT_WORDS * 8));
  JB_CHECK(cudaMalloc((void**)&flag, 256));
  JB_CHECK(cudaMalloc((void**)&errs, 8));
  JB_CHECK(cudaMemset(payload, 0, LIT_WORDS * 8));
  JB_CHECK(cudaMemset(flag, 0, 256));
  JB_CHECK(cudaMemset(errs, 0, 8));
  JB_CHECK(cudaSetDevice(dev_producer));
  cudaDeviceEnablePeerAccess(dev_consumer, 0);
  cudaGetLastError();
  JB_CHECK(cudaMalloc((void**)&ack, 256));
  JB_CHECK(cudaMemset(ack, 0, 256));
  JB_CHECK(cudaDeviceSynchronize());
  JB_CHECK(cudaSetDevice(dev_consumer));
  JB_CHECK(cudaDeviceSynchronize());
  k_litmus_consumer<<<1, 128>>>(payload, flag, ack, iters, errs);
  JB_CHECK(cudaGetLastError());
  JB_CHECK(cudaSetDevice(dev_producer));
  k_litmus_producer<<<1, 128>>>(payload, flag, ack, iters);
  JB_CHECK(cudaGetLastError());
  JB_CHECK(cudaDeviceSynchronize());
  JB_CHECK(cudaSetDevice(dev_consumer));
  JB_CHECK(cudaDeviceSynchronize());
  JB_CHECK(cudaMemcpy(errors_out, errs, 8, cudaMemcpyDeviceToHost));
  cudaFree(payload); cudaFree(flag); cudaFree(errs);
  JB_CHECK(cudaSetDevice(dev_producer));
  cudaFree(ack);
  return JB_OK;
}

// ---------------------------------------------------------------------------------------------
// Store bandwidth of the paths the SCALE items use to hand the pivot row to the other GPUs
// (measured NVLink roof for profiles/): 16-byte stores to a peer-mapped address, or multimem.st to
// the NVLS multicast alias.  mode 0: st.global.v2.f64 (dst may be local or a peer mapping),
// mode 1: multimem.st (dst = multicast alias), mode 2: ld.global.v2.f64 (peer pull).
__global__ void __launch_bounds__(256) k_store_bw(double* dst, size_t n_pairs, int mode, double* sink) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  double acc = 0.0;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_pairs; i += stride) {
    double* p = dst + 2 * i;
    if (mode == 0) {
      *reinterpret_cast<double2*>(p) = make_double2((double)i, 1.0);
    } else if (mode == 1) {
      jb::multimem_st_v2f64(p, (double)i, 1.0);
    } else {
      const double2 v = *reinterpret_cast<const double2*>(p);
      acc += v.x + v.y;
    }
  }
  if (mode == 2 && acc == 123.456) *sink = acc;  // keep the loads
}

// Returns the average milliseconds of `iters` launches over `bytes` (after one warm-up launch).
extern "C" int jb_store_bandwidth(void* dst, size_t bytes, int mode, int iters, float* ms_out) {
  if (!dst || bytes < 16 || iters < 1 || mode < 0 || mode > 2 || !ms_out) {
    jb_set_error("jb_store_bandwidth: bad arguments");
    return JB_ERR_ARG;
  }
  double* sink = nullptr;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  float ms = 0;
  cudaError_t ce = cudaMalloc((void**)&sink, 8);
  if (ce == cudaSuccess) ce = cudaEventCreate(&e0);
  if (ce == cudaSuccess) ce = cudaEventCreate(&e1);
  const size_t pairs = bytes / 16;
  if (ce == cudaSuccess) {
    k_store_bw<<<148 * 8, 256>>>((double*)dst, pairs, mode, sink);
    ce = cudaDeviceSynchronize();
  }
  if (ce == cudaSuccess) ce = cudaEventRecord(e0);
  if (ce == cudaSuccess) {
    for (int i = 0; i < iters; ++i) k_store_bw<<<148 * 8, 256>>>((double*)dst, pairs, mode, sink);
    ce = cudaEventRecord(e1);
  }
  if (ce == cudaSuccess) ce = cudaEventSynchronize(e1);
  if (ce == cudaSuccess) ce = cudaGetLastError();
  if (ce == cudaSuccess) ce = cudaEventElapsedTime(&ms, e0, e1);
  if (e0) cudaEventDestroy(e0);
  if (e1) cudaEventDestroy(e1);
  cudaFree(sink);
  if (ce != cudaSuccess) {
    jb_set_error("jb_store_bandwidth: %s", cudaGetErrorString(ce));
    return JB_ERR_CUDA;
  }
  *ms_out = ms / iters;
  return JB_OK;
}
