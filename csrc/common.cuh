// Common device/host helpers for the B200 (sm_100a) block Gauss-Jordan engine.
// PTX wrappers: mbarrier, TMA (cp.async.bulk.tensor), FP64 tensor-core MMA (DMMA),
// system-scope acquire/release for cross-GPU flags, NVLS multimem stores.
#pragma once
#include <cuda_runtime.h>
#include <cuda.h>
#include <stdint.h>
#include <stdio.h>

#define JB_CHECK(expr)                                                              \
  do {                                                                              \
    cudaError_t _e = (expr);                                                        \
    if (_e != cudaSuccess) {                                                        \
      jb_set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #expr,               \
                   cudaGetErrorString(_e));                                         \
      return JB_ERR_CUDA;                                                           \
    }                                                                               \
  } while (0)

enum {
  JB_OK = 0,
  JB_ERR_CUDA = -1,
  JB_ERR_SINGULAR = -2,   // same code the reference's Jordan() returns (main.cpp:1082)
  JB_ERR_NOMEM = -3,      // same code as main.cpp:989
  JB_ERR_ARG = -4,
  JB_ERR_IO = -5,
  JB_ERR_TIMEOUT = -6,  // a cross-GPU spin-wait gave up (failure detection)
};

void jb_set_error(const char* fmt, ...);

#ifdef __CUDACC__

namespace jb {

// ---------------------------------------------------------------- misc
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

// ---------------------------------------------------------------- mbarrier
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}

// ---------------------------------------------------------------- TMA
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* m) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
// 2-D tiled load global -> shared, completion on an mbarrier (SASS: UTMALDG).
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* m, int x, int y,
                                            uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%2, %3}], [%4];" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(x), "r"(y), "r"(smem_u32(bar))
      : "memory");
}
// generic-proxy <-> async-proxy ordering (data written with st/remote st, then read by TMA)
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async;" ::: "memory");
}

// ---------------------------------------------------------------- FP64 tensor-core MMA
// D(8x8) = A(8x4, row) * B(4x8, col) + C.  SASS: DMMA.8x8x4.
// lane l: a = A[l>>2][l&3], b = B[l&3][l>>2], c/d = C[l>>2][2*(l&3) + {0,1}].
__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(d0), "+d"(d1)
               : "d"(a), "d"(b));
}

// ---------------------------------------------------------------- system-scope flags
__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_relaxed_sys_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void red_release_sys_add_u32(uint32_t* p, uint32_t v) {
  asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void fence_acq_rel_sys() {
  asm volatile("fence.acq_rel.sys;" ::: "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
// spin until *p >= want (monotonic epoch counters).  Failure detection (the reference has none: a
// dead rank hangs the job): with a watchdog (abort != nullptr, timeout_ns > 0) the wait gives up
// after timeout_ns, raises *abort so that every other wait and every later kernel drains, and
// returns false; the host reports JB_ERR_TIMEOUT.
__device__ __forceinline__ bool wait_flag_ge(const uint32_t* p, uint32_t want, int* abort = nullptr,
                                             unsigned long long timeout_ns = 0) {
  unsigned long long t0 = 0;
  unsigned spins = 0;
  while ((int32_t)(ld_acquire_sys_u32(p) - want) < 0) {
    __nanosleep(20);
    if (abort && ((++spins & 1023u) == 0)) {
      if (*(volatile int*)abort) return false;
      if (timeout_ns) {
        const unsigned long long now = globaltimer_ns();
        if (!t0) t0 = now;
        else if (now - t0 > timeout_ns) {
          *(volatile int*)abort = 1;
          __threadfence_system();
          return false;
        }
      }
    }
  }
  return true;
}

// ---------------------------------------------------------------- NVLS multicast (multimem)
// 16-byte store replicated by the NVSwitch into every GPU bound to the multicast object.
__device__ __forceinline__ void multimem_st_v2f64(void* mc_addr, double x, double y) {
  uint32_t a = (uint32_t)__double2loint(x), b = (uint32_t)__double2hiint(x);
  uint32_t c = (uint32_t)__double2loint(y), d = (uint32_t)__double2hiint(y);
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc_addr),
               "f"(__uint_as_float(a)), "f"(__uint_as_float(b)), "f"(__uint_as_float(c)),
               "f"(__uint_as_float(d))
               : "memory");
}
__device__ __forceinline__ void multimem_red_release_add_u32(void* mc_addr, uint32_t v) {
  asm volatile("multimem.red.release.sys.global.add.u32 [%0], %1;" ::"l"(mc_addr), "r"(v)
               : "memory");
}

// named barrier over a subset of warps
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

}  // namespace jb
#endif  // __CUDACC__
