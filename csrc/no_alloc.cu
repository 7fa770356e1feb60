// jb_block_inverse_noalloc — batched block inversion for callers that must not allocate.
//
// Who needs this: code that runs after the application has given all device memory to its own
// pool (so a cudaMalloc inside a library call would fail or eat into somebody else's budget), and
// latency-critical loops that cannot afford the device-wide synchronisation cudaMalloc/cudaFree
// imply.  jb_block_inverse (engine.cu) allocates and frees its staging area on every call; this
// variant keeps the staging area in a fixed device array that is part of the library image.
//
// Cost / limits: the array below (JB_NOALLOC_SCRATCH_BYTES, ~4 MiB) is resident on every GPU that
// loads the library, until the process exits.  A request is accepted only if
// nblocks * Kp * Kp <= JB_NOALLOC_MAX_DOUBLES, Kp = m rounded up to 16, and m <= 256.
//
// Sharing: there is ONE array per device.  Every call reuses it and finds the previous call's
// contents, so each call clears the parts that have to start at zero (control block, mailbox tags,
// the staging it is about to use).  Two calls that overlap on the same GPU (different streams or
// threads) race on it - serialise them.
//
// Set-up: the first call on a device resolves the device address of the array
// (cudaGetSymbolAddress with that device current) and remembers it for the life of the process.
// Make that first call during initialisation, before any stream capture is started.  The
// remembered address does not survive cudaDeviceReset (the module is reloaded): do not reset a
// device between calls.
//
// Outputs: d_inv holds nblocks packed m x m inverses, d_norms / d_ok nblocks entries each; a block
// found singular (d_ok = 0) comes back as zeros with norm 0.
#include <mutex>
#include <string.h>
#include "common.cuh"
#include "engine_types.h"
#include "kernels.h"
#include "../include/jordan_b200.h"

#define JB_NOALLOC_CTRL_BYTES 8192
#define JB_NOALLOC_MAX_DOUBLES (512 * 1024) /* 4 MiB of inverse staging */
#define JB_NOALLOC_SCRATCH_BYTES (JB_NOALLOC_CTRL_BYTES + JB_NOALLOC_MAX_DOUBLES * 8)
#define JB_NOALLOC_MAX_DEVICES 64

// The name below is a host-side handle only: host code must never pass `g_noalloc_scratch` itself
// to a kernel, only the address obtained from cudaGetSymbolAddress.
static __device__ __align__(256) unsigned char g_noalloc_scratch[JB_NOALLOC_SCRATCH_BYTES];

namespace {

std::mutex g_mu;
unsigned char* g_dev_ptr[JB_NOALLOC_MAX_DEVICES] = {nullptr};

__global__ void k_noalloc_set_norm(JbCtl* ctl, double norm_a) { ctl->norm_a = norm_a; }

int scratch_for_current_device(unsigned char** out) {
  int dev = -1;
  JB_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= JB_NOALLOC_MAX_DEVICES) {
    jb_set_error("jb_block_inverse_noalloc: device ordinal %d not supported (max %d)", dev,
                 JB_NOALLOC_MAX_DEVICES);
    return JB_ERR_ARG;
  }
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_dev_ptr[dev]) {
    void* p = nullptr;
    JB_CHECK(cudaGetSymbolAddress(&p, g_noalloc_scratch));  // per device: `dev` is current here
    g_dev_ptr[dev] = static_cast<unsigned char*>(p);
  }
  *out = g_dev_ptr[dev];
  return JB_OK;
}

}  // namespace

extern "C" int jb_block_inverse_noalloc(const double* d_blocks, int64_t nblocks, int32_t m,
                                        double norm_a, double* d_inv, double* d_norms,
                                        int32_t* d_ok, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (!d_blocks || !d_inv || !d_norms || !d_ok || nblocks <= 0 || m <= 0) {
    jb_set_error("jb_block_inverse_noalloc: bad arguments");
    return JB_ERR_ARG;
  }
  if (m > 256) {
    jb_set_error("jb_block_inverse_noalloc: m=%d exceeds the supported maximum of 256", m);
    return JB_ERR_ARG;
  }
  const int64_t Kp = (m + JB_BK - 1) / JB_BK * JB_BK;
  const int64_t need = nblocks * Kp * Kp;
  if (need > JB_NOALLOC_MAX_DOUBLES) {
    jb_set_error("jb_block_inverse_noalloc: %lld blocks of m=%d need %lld doubles of staging, the "
                 "static scratch holds %d (use jb_block_inverse, or fewer blocks per call)",
                 (long long)nblocks, m, (long long)need, JB_NOALLOC_MAX_DOUBLES);
    return JB_ERR_ARG;
  }
  unsigned char* scratch = nullptr;
  int rc = scratch_for_current_device(&scratch);
  if (rc) return rc;

  // everything that must start from zero: control block + mailbox + step record, the staging this
  // call will read back, and the verdict flags
  JB_CHECK(cudaMemsetAsync(scratch, 0, JB_NOALLOC_CTRL_BYTES, st));
  JB_CHECK(cudaMemsetAsync(scratch + JB_NOALLOC_CTRL_BYTES, 0, (size_t)need * 8, st));
  JB_CHECK(cudaMemsetAsync(d_ok, 0, (size_t)nblocks * 4, st));

  K1Params K;
  memset(&K, 0, sizeof(K));
  K.W = d_blocks;
  K.Lbuf = nullptr;  // no panel-copy CTAs are launched (panel_ctas = 0)
  K.invbuf = reinterpret_cast<double*>(scratch + JB_NOALLOC_CTRL_BYTES);
  K.ctl = reinterpret_cast<JbCtl*>(scratch);
  K.mail_peers[0] = reinterpret_cast<JbMail*>(scratch + 256);
  K.stepinfo = reinterpret_cast<JbStepInfo*>(scratch + 256 + 2 * JB_MAX_RANKS * sizeof(JbMail));
  K.normbuf = d_norms;
  K.okbuf = d_ok;
  K.ldw = m; K.m = m; K.Kp = (int)Kp; K.p = 1; K.rank = 0; K.t = 0; K.rloc = (int)nblocks;
  K.s0 = 0; K.ncand = (int)nblocks; K.ngroups = (int)nblocks; K.tag = 1;
  k_noalloc_set_norm<<<1, 1, 0, st>>>(K.ctl, norm_a);
  JB_CHECK(cudaGetLastError());
  JB_CHECK(jb_launch_k1(K, 0, st));
  // one launch packs all blocks (pitch Kp -> m); a loop of per-block copies would cost up to
  // thousands of launches for small m, which defeats a latency-critical caller
  JB_CHECK(jb_launch_compact_blocks(K.invbuf, (int)Kp, d_inv, m, nblocks, st));
  return JB_OK;
}
