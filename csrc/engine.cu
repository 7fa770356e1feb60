// Host-side engine: owns the per-rank buffers, builds the TMA descriptors once, and enqueues the
// whole elimination (Nr steps x {K1 pivot search + min-loc, fused scale/broadcast/update}) on one
// stream with no host synchronisation: the pivot choice of every step stays in device memory.
// This is the B200 replacement of solve()/Jordan() (main.cpp:343-519, 953-1204).
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <cmath>
#include <vector>
#include "common.cuh"
#include "engine_types.h"
#include "kernels.h"
#include "../include/jordan_b200.h"

static thread_local char g_err[512] = "";
void jb_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
extern "C" const char* jb_last_error(void) { return g_err; }

namespace {

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

// 2-D fp64 tensor map: inner dimension `cols` (contiguous), `rows` rows of pitch ld doubles.
int make_map(CUtensorMap* map, const void* base, uint64_t cols, uint64_t rows, uint64_t ld,
             uint32_t box_cols, uint32_t box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) {
    jb_set_error("cuTensorMapEncodeTiled entry point not available");
    return JB_ERR_CUDA;
  }
  cuuint64_t gdim[2] = {cols ? cols : 1, rows ? rows : 1};
  cuuint64_t gstr[1] = {ld * sizeof(double)};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<void*>(base), gdim, gstr, box,
                  estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    jb_set_error("cuTensorMapEncodeTiled failed: %d (cols=%llu rows=%llu ld=%llu box=%ux%u)", (int)r,
                 (unsigned long long)cols, (unsigned long long)rows, (unsigned long long)ld, box_cols,
                 box_rows);
    return JB_ERR_CUDA;
  }
  return JB_OK;
}

}  // namespace

struct JbEngine {
  JbConfig cfg;
  int64_t n, Nr, ldw, rows_loc;
  int m, p, rank, Kp, rloc, nsub, CT, sm_count;
  // local buffers
  double *W = nullptr, *normbuf = nullptr, *Acopy = nullptr;
  // K1(t+2) may already run while step t is being applied and A(t+1) reads its own panel: the L
  // panel of step g lives in Lbuf[g % 3], the candidate inverses of step g in invbuf[g & 1]
  double* Lbuf[3] = {nullptr, nullptr, nullptr};
  double* invbuf[2] = {nullptr, nullptr};
  uint32_t* counters = nullptr;          // one work-item counter per step-kernel launch
  cudaStream_t hstream = nullptr;        // high-priority stream running K1(t+1) beside step_B(t)
  cudaStream_t bstream = nullptr;        // carries B2(t) so it can move in while B1(t) drains
  cudaEvent_t evA = nullptr, evK[2] = {nullptr, nullptr}, evG[2] = {nullptr, nullptr};
  cudaEvent_t evB1[2] = {nullptr, nullptr}, evB2[2] = {nullptr, nullptr};
  uint32_t* b1sync = nullptr;            // {B1 CTA exit counter, B1 done tag} (look-ahead, see StepParams::b1sync)
  bool lookahead = true;
  bool k1_early = true;    // K1(t+2) pending right behind A(t+1) (JB_K1_GATED=1: released by the end of B2(t))
  int sched = 0;           // JB_SCHED: 1 = phase A in line (between B2(t-1) and B1(t)), else A beside B2(t-1)
  // B2(t) starts behind B1(t).  Letting it move in while B1 drains (JB_B2_OVERLAP=1: no event wait,
  // only the per-tile flags) measured no gain (0.3099 s against 0.3084-0.3123 s at 8 GPUs) and gave
  // WRONG inverses at 8 GPUs, n = 32768.  Cause (found with tools/schedule_model.py): phase A(t+1)
  // then starts while B1(t) may still run, and although the TMA producer thread waits for the flag
  // of B1's early tile, the MMA warps read the C tile of that item from W as soon as it is handed
  // to them - before B1's update of it.  Behind a finished B1 the pair cannot occur.  The switch is
  // kept for debugging only; it needs that extra wait in the MMA warps before it may be used.
  bool b2_overlap = false;
  // tcgen05 split-precision update (precision = 1)
  bool tc05 = false;
  int tc_slices = 8;
  int8_t* Lq[2][2] = {{nullptr, nullptr}, {nullptr, nullptr}};  // [step parity][k-block]
  int8_t* Rq[2] = {nullptr, nullptr};                              // [k-block]
  double* rscale[2][2] = {{nullptr, nullptr}, {nullptr, nullptr}};
  double* cscale[2] = {nullptr, nullptr};
  int tc_kb = 1;
  int64_t cols_pad = 0;
  int* okbuf = nullptr;
  JbCtl* ctl = nullptr;
  JbStepInfo* stepinfo = nullptr;
  double *Xp = nullptr, *rowsum = nullptr, *scratch_max = nullptr, *Rslab = nullptr;
  // symmetric region
  size_t off_rowbuf[2], off_swapbuf[2], off_rowflag[2], off_swapflag[2], off_mail, off_stepdone, symm_bytes;
  uint32_t* prowflag = nullptr;  // [CT][nsub] local flags of the pivot-row look-ahead (phase B2)
  double b1_frac = -1.0;         // share of the non-priority column tiles given to phase B1 (<0: auto)
  unsigned char* symm_peers[JB_MAX_RANKS] = {nullptr};
  unsigned char* symm_mc = nullptr;
  bool own_symm = false, symm_ready = false;
  bool peers_opened[3] = {false, false, false};
  const double* W_peers[JB_MAX_RANKS] = {nullptr};
  const double* R_peers[JB_MAX_RANKS] = {nullptr};  // refinement residual slabs (K7)
  StepMaps maps[2][6];  // [SCALE context is the next step][global step % 6]
  bool maps_ready = false;
  uint32_t gstep = 0;  // global step counter (same on every rank)
  int64_t launches = 0;
  bool norm_set = false;
  // optional per-kernel timing (JB_TIME_KERNELS=1 or jb_engine_set_profiling)
  bool profile = false;
  std::vector<cudaEvent_t> ev;
  // per-step segments summed over the inversion.  look-ahead: [0] phase A, [1] phase B1 (K1(t+1)
  // beside it), [2] main stream still waiting for K1(t+1), [3] phase B2.  serial / tcgen05: [0] step
  // kernel (tcgen05: scale/broadcast/swap), [1] tcgen05 update, [2] K1(t+1), [3] -
  double prof_ms[4] = {0, 0, 0, 0};
  double prof_tail_ms = 0;
};

namespace {

int build_maps(JbEngine* e) {
  unsigned char* loc = e->symm_peers[e->rank];
  for (int nx = 0; nx < 2; ++nx)
    for (int g = 0; g < 6; ++g) {
      StepMaps& M = e->maps[nx][g];
      const int par = g & 1;
      int rc;
      if ((rc = make_map(&M.tmW, e->W, e->ldw, e->rows_loc, e->ldw, JB_RS, JB_BK))) return rc;
      if ((rc = make_map(&M.tmRow, loc + e->off_rowbuf[par], e->ldw, e->Kp, e->ldw, JB_RS, JB_BK)))
        return rc;
      if ((rc = make_map(&M.tmL, e->Lbuf[g % 3], e->Kp, e->rows_loc, e->Kp, JB_LS, JB_BM))) return rc;
      // inverse of the pivot block of the step the SCALE items of the launch belong to
      if ((rc = make_map(&M.tmInv, e->invbuf[(g + nx) & 1], e->Kp, (uint64_t)e->rloc * e->Kp, e->Kp, JB_LS,
                         JB_BM)))
        return rc;
      if ((rc = make_map(&M.tmSwap, loc + e->off_swapbuf[par], e->ldw, e->Kp, e->ldw, JB_LS, JB_BM)))
        return rc;
    }
  e->maps_ready = true;
  return JB_OK;
}

}  // namespace

extern "C" {

int jb_engine_create(const JbConfig* cfg, JbEngine** out) {
  if (!cfg || !out || cfg->n <= 0 || cfg->m <= 0 || cfg->p <= 0 || cfg->p > JB_MAX_RANKS ||
      cfg->rank < 0 || cfg->rank >= cfg->p) {
    jb_set_error("jb_engine_create: bad configuration");
    return JB_ERR_ARG;
  }
  if (jb_k1_cluster_size(cfg->m) == 0) {
    jb_set_error("block size m=%d does not fit distributed shared memory (max ~440)", cfg->m);
    return JB_ERR_ARG;
  }
  JB_CHECK(cudaSetDevice(cfg->device));
  JbEngine* e = new JbEngine();
  e->cfg = *cfg;
  e->n = cfg->n;
  e->m = cfg->m;
  e->p = cfg->p;
  e->rank = cfg->rank;
  e->Nr = jb_num_block_rows(e->n, e->m);
  e->rloc = (int)jb_rows_per_process(e->Nr, e->p, e->rank);
  e->rows_loc = (int64_t)e->rloc * e->m;
  e->ldw = jb_padded_ld(e->n, e->m);
  e->Kp = (int)align_up((size_t)e->m, JB_BK);
  e->nsub = (e->m + JB_BM - 1) / JB_BM;
  e->CT = (int)((e->ldw + JB_BN - 1) / JB_BN);
  cudaDeviceProp prop;
  JB_CHECK(cudaGetDeviceProperties(&prop, cfg->device));
  e->sm_count = prop.multiProcessorCount;

  auto dmalloc = [&](void** ptr, size_t bytes) -> cudaError_t {
    cudaError_t r = cudaMalloc(ptr, bytes ? bytes : 256);
    if (r == cudaSuccess) r = cudaMemset(*ptr, 0, bytes ? bytes : 256);
    return r;
  };
  const size_t rl = (size_t)std::max<int64_t>(e->rows_loc, 1);
  cudaError_t r = cudaSuccess;
  if (r == cudaSuccess) r = dmalloc((void**)&e->W, rl * e->ldw * 8);
  for (int i = 0; i < 3 && r == cudaSuccess; ++i) r = dmalloc((void**)&e->Lbuf[i], (rl + JB_BM) * e->Kp * 8);
  if (r == cudaSuccess) r = dmalloc((void**)&e->counters, (size_t)(4 * e->Nr + 4) * 4);
  if (r == cudaSuccess) r = dmalloc((void**)&e->prowflag, (size_t)e->CT * e->nsub * 4);
  if (r == cudaSuccess) r = dmalloc((void**)&e->b1sync, 256);
  if (r == cudaSuccess)
    for (int i = 0; i < 2 && r == cudaSuccess; ++i)
      r = dmalloc((void**)&e->invbuf[i], ((size_t)std::max(e->rloc, 1) * e->Kp + JB_BM) * e->Kp * 8);
  if (r == cudaSuccess) r = dmalloc((void**)&e->normbuf, (size_t)std::max(e->rloc, 1) * 8);
  if (r == cudaSuccess) r = dmalloc((void**)&e->okbuf, (size_t)std::max(e->rloc, 1) * 4);
  if (r == cudaSuccess) r = dmalloc((void**)&e->ctl, sizeof(JbCtl));
  if (r == cudaSuccess) r = dmalloc((void**)&e->stepinfo, (size_t)e->Nr * sizeof(JbStepInfo));
  if (r == cudaSuccess) r = dmalloc((void**)&e->rowsum, rl * 8);
  if (r == cudaSuccess) r = dmalloc((void**)&e->scratch_max, 64);
  if (r == cudaSuccess && cfg->keep_copy) r = dmalloc((void**)&e->Acopy, rl * e->ldw * 8);
  if (cfg->precision == 1) {
    if (e->m % 128 != 0) {
      jb_set_error("precision=1 (tcgen05 update) needs a block size that is a multiple of 128");
      jb_engine_destroy(e);
      return JB_ERR_ARG;
    }
    e->tc05 = true;
    if (const char* sl = getenv("JB_TC05_SLICES")) e->tc_slices = std::max(1, std::min(8, atoi(sl)));
    e->cols_pad = (e->ldw + 63) / 64 * 64;
    e->tc_kb = (e->m + 127) / 128;
    if (e->tc_kb > 2) {
      jb_set_error("precision=1 supports m = 128 or 256");
      jb_engine_destroy(e);
      return JB_ERR_ARG;
    }
    for (int kb = 0; kb < e->tc_kb && r == cudaSuccess; ++kb) {
      for (int i = 0; i < 2 && r == cudaSuccess; ++i) {
        r = dmalloc((void**)&e->Lq[i][kb], jb_tc05_lq_bytes(rl, 128, 8));
        if (r == cudaSuccess) r = dmalloc((void**)&e->rscale[i][kb], rl * 8);
      }
      if (r == cudaSuccess) r = dmalloc((void**)&e->Rq[kb], jb_tc05_rq_bytes(e->cols_pad, 128, 8));
      if (r == cudaSuccess) r = dmalloc((void**)&e->cscale[kb], (size_t)e->cols_pad * 8);
    }
  }
  if (r != cudaSuccess) {
    jb_set_error("Not enough memory! (%s)", cudaGetErrorString(r));  // main.cpp:375
    jb_engine_destroy(e);
    return JB_ERR_NOMEM;
  }
  // symmetric region layout
  size_t off = 0;
  const size_t rb = align_up((size_t)e->Kp * e->ldw * 8, 256);
  e->off_rowbuf[0] = off; off += rb;
  e->off_rowbuf[1] = off; off += rb;
  // swap buffer and its flags are double-buffered by step parity like the pivot-row buffer: with
  // look-ahead the ranks drift by up to one phase, so ROWT(t+1) may push while a slower pivot owner
  // still reads what ROWT(t) pushed
  e->off_swapbuf[0] = off; off += rb;
  e->off_swapbuf[1] = off; off += rb;
  const size_t fb = align_up((size_t)e->CT * 4, 256);
  e->off_rowflag[0] = off; off += fb;
  e->off_rowflag[1] = off; off += fb;
  const size_t sfb = align_up((size_t)e->CT * e->nsub * 4, 256);  // one flag per 64-row chunk
  e->off_swapflag[0] = off; off += sfb;
  e->off_swapflag[1] = off; off += sfb;
  e->off_mail = off; off += align_up(2 * JB_MAX_RANKS * sizeof(JbMail), 256);
  e->off_stepdone = off; off += align_up(JB_MAX_RANKS * 4, 256);
  e->symm_bytes = align_up(off, 4096);
  e->W_peers[e->rank] = e->W;
  {  // watchdog for cross-GPU spin-waits: default 30 s, JB_SPIN_TIMEOUT_MS=0 disables it
    unsigned long long ns = 30ull * 1000 * 1000 * 1000;
    if (const char* t = getenv("JB_SPIN_TIMEOUT_MS")) ns = strtoull(t, nullptr, 10) * 1000000ull;
    if (cudaMemcpy(&e->ctl->spin_timeout_ns, &ns, 8, cudaMemcpyHostToDevice) != cudaSuccess) {
      jb_set_error("jb_engine_create: cannot initialise the control block (%s)",
                   cudaGetErrorString(cudaGetLastError()));
      jb_engine_destroy(e);
      return JB_ERR_CUDA;
    }
  }
  if (const char* f = getenv("JB_B1_FRAC")) e->b1_frac = std::max(0.0, std::min(1.0, atof(f)));
  if (getenv("JB_TIME_KERNELS")) e->profile = true;
  if (getenv("JB_NO_LOOKAHEAD")) e->lookahead = false;
  if (getenv("JB_B2_OVERLAP")) e->b2_overlap = true;
  if (getenv("JB_K1_GATED")) e->k1_early = false;
  if (const char* sc = getenv("JB_SCHED")) e->sched = atoi(sc);
  {
    int lo = 0, hi = 0;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);
    if (cudaStreamCreateWithPriority(&e->hstream, cudaStreamNonBlocking, hi) != cudaSuccess) e->lookahead = false;
    if (cudaStreamCreateWithPriority(&e->bstream, cudaStreamNonBlocking, lo) != cudaSuccess) e->lookahead = false;
    for (int i = 0; i < 2 && e->lookahead; ++i) {
      if ((i == 0 && cudaEventCreateWithFlags(&e->evA, cudaEventDisableTiming) != cudaSuccess) ||
          cudaEventCreateWithFlags(&e->evK[i], cudaEventDisableTiming) != cudaSuccess ||
          cudaEventCreateWithFlags(&e->evG[i], cudaEventDisableTiming) != cudaSuccess ||
          cudaEventCreateWithFlags(&e->evB1[i], cudaEventDisableTiming) != cudaSuccess ||
          cudaEventCreateWithFlags(&e->evB2[i], cudaEventDisableTiming) != cudaSuccess)
        e->lookahead = false;
    }
  }
  *out = e;
  return JB_OK;
}

void jb_engine_destroy(JbEngine* e) {
  if (!e) return;
  cudaSetDevice(e->cfg.device);
  cudaDeviceSynchronize();
  for (int which = 0; which < 3; ++which)
    if (e->peers_opened[which])
      for (int r = 0; r < e->p; ++r)
        if (r != e->rank) {
          void* ptr = which == 0 ? (void*)e->symm_peers[r]
                                 : (which == 1 ? (void*)e->W_peers[r] : (void*)e->R_peers[r]);
          if (ptr) cudaIpcCloseMemHandle(ptr);
        }
  if (e->own_symm && e->symm_peers[e->rank]) cudaFree(e->symm_peers[e->rank]);
  if (e->hstream) cudaStreamDestroy(e->hstream);
  if (e->bstream) cudaStreamDestroy(e->bstream);
  for (int i = 0; i < 2; ++i) {
    if (i == 0 && e->evA) cudaEventDestroy(e->evA);
    if (e->evK[i]) cudaEventDestroy(e->evK[i]);
    if (e->evG[i]) cudaEventDestroy(e->evG[i]);
    if (e->evB1[i]) cudaEventDestroy(e->evB1[i]);
    if (e->evB2[i]) cudaEventDestroy(e->evB2[i]);
  }
  cudaFree(e->counters);
  cudaFree(e->prowflag);
  cudaFree(e->b1sync);
  for (int i = 0; i < 2; ++i) {
    for (int kb = 0; kb < 2; ++kb) { cudaFree(e->Lq[i][kb]); cudaFree(e->rscale[i][kb]); }
    cudaFree(e->Rq[i]); cudaFree(e->cscale[i]);
  }
  cudaFree(e->W); cudaFree(e->Lbuf[0]); cudaFree(e->Lbuf[1]); cudaFree(e->Lbuf[2]); cudaFree(e->invbuf[0]);
  cudaFree(e->invbuf[1]); cudaFree(e->normbuf); cudaFree(e->okbuf);
  cudaFree(e->ctl); cudaFree(e->stepinfo); cudaFree(e->rowsum); cudaFree(e->scratch_max);
  cudaFree(e->Acopy); cudaFree(e->Xp); cudaFree(e->Rslab);
  delete e;
}

size_t jb_engine_symm_bytes(const JbEngine* e) { return e->symm_bytes; }

int jb_engine_alloc_symm(JbEngine* e) {
  JB_CHECK(cudaSetDevice(e->cfg.device));
  void* ptr = nullptr;
  if (cudaMalloc(&ptr, e->symm_bytes) != cudaSuccess) {
    jb_set_error("Not enough memory! (symmetric region, %zu bytes)", e->symm_bytes);
    return JB_ERR_NOMEM;
  }
  JB_CHECK(cudaMemset(ptr, 0, e->symm_bytes));
  e->symm_peers[e->rank] = (unsigned char*)ptr;
  e->own_symm = true;
  if (e->p == 1) {
    e->symm_ready = true;
    return build_maps(e);
  }
  return JB_OK;
}

int jb_engine_bind_symm(JbEngine* e, void* const* peer_ptrs, void* multicast_ptr) {
  JB_CHECK(cudaSetDevice(e->cfg.device));
  for (int r = 0; r < e->p; ++r) {
    if (!peer_ptrs[r]) {
      jb_set_error("jb_engine_bind_symm: null peer pointer for rank %d", r);
      return JB_ERR_ARG;
    }
    e->symm_peers[r] = (unsigned char*)peer_ptrs[r];
  }
  e->symm_mc = (unsigned char*)multicast_ptr;
  e->symm_ready = true;
  return build_maps(e);
}

int jb_engine_ipc_handle(JbEngine* e, int which, void* handle64) {
  JB_CHECK(cudaSetDevice(e->cfg.device));
  cudaIpcMemHandle_t h;
  if (which == 2 && !e->Rslab) {  // refinement slabs are allocated on first use
    const size_t slab = (size_t)std::max<int64_t>(e->rows_loc, 1) * e->ldw * 8;
    if (cudaMalloc((void**)&e->Rslab, 2 * slab) != cudaSuccess) {
      jb_set_error("Not enough memory! (refinement slabs)");
      return JB_ERR_NOMEM;
    }
    JB_CHECK(cudaMemset(e->Rslab, 0, 2 * slab));
    e->R_peers[e->rank] = e->Rslab;
  }
  void* ptr = which == 0 ? (void*)e->symm_peers[e->rank] : (which == 1 ? (void*)e->W : (void*)e->Rslab);
  if (!ptr) {
    jb_set_error("jb_engine_ipc_handle: region %d not allocated", which);
    return JB_ERR_ARG;
  }
  JB_CHECK(cudaIpcGetMemHandle(&h, ptr));
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(handle64, &h, 64);
  return JB_OK;
}

int jb_engine_open_peers(JbEngine* e, int which, const void* handles) {
  JB_CHECK(cudaSetDevice(e->cfg.device));
  for (int r = 0; r < e->p; ++r) {
    if (r == e->rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, (const char*)handles + 64 * r, 64);
    void* ptr = nullptr;
    JB_CHECK(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
    if (which == 0) e->symm_peers[r] = (unsigned char*)ptr;
    else if (which == 1) e->W_peers[r] = (const double*)ptr;
    else e->R_peers[r] = (const double*)ptr;
  }
  e->peers_opened[which] = true;
  if (which == 0) {
    e->symm_ready = true;
    return build_maps(e);
  }
  return JB_OK;
}

int jb_engine_reset_sync(JbEngine* e) {
  JB_CHECK(cudaSetDevice(e->cfg.device));
  JB_CHECK(cudaDeviceSynchronize());
  unsigned char* loc = e->symm_peers[e->rank];
  if (loc) JB_CHECK(cudaMemset(loc + e->off_rowflag[0], 0, e->symm_bytes - e->off_rowflag[0]));
  JB_CHECK(cudaMemset(e->ctl, 0, offsetof(JbCtl, norm_a)));
  JB_CHECK(cudaMemset(e->prowflag, 0, (size_t)e->CT * e->nsub * 4));
  JB_CHECK(cudaDeviceSynchronize());
  e->gstep = 0;
  return JB_OK;
}

int jb_engine_fill_formula(JbEngine* e, int formula, void* stream) {
  JB_CHECK(cudaSetDevice(e->cfg.device));
  cudaStream_t st = (cudaStream_t)stream;
  JB_CHECK(cudaMemsetAsync(e->scratch_max, 0, 8, st));
  FillParams P;
  P.W = e->W; P.ldw = e->ldw; P.n = e->n; P.m = e->m; P.p = e->p; P.rank = e->rank;
  P.rloc = e->rloc; P.formula = formula; P.rowsum_max = e->scratch_max;
  JB_CHECK(jb_launch_fill(P, st));
  e->launches += (e->rows_loc > 0);
  if (e->Acopy)
    JB_CHECK(cudaMemcpyAsync(e->Acopy, e->W, (size_t)e->rows_loc * e->ldw * 8,
                             cudaMemcpyDeviceToDevice, st));
  return JB_OK;
}

int jb_engine_upload_rows(JbEngine* e, const double* host_rows, int64_t local_row0, int64_t nrows,
                          void* stream) {
  JB_CHECK(cudaSetDevice(e->cfg.device));
  if (local_row0 < 0 || local_row0 + nrows > e->rows_loc) {
    jb_set_error("jb_engine_upload_rows: rows out of range");
    return JB_ERR_ARG;
  }
  JB_CHECK(cudaMemcpy2DAsync(e->W + local_row0 * e->ldw, (size_t)e->ldw * 8, host_rows,
                             (size_t)e->n * 8, (size_t)e->n * 8, (size_t)nrows,
                             cudaMemcpyHostToDevice, (cudaStream_t)stream));
  return JB_OK;
}

// identity padding of rows / columns >= n (formula 0 = keep what was uploaded)
__global__ void k_pad_identity(double* W, int64_t ldw, int64_t n, int m, int p, int rank,
                               int64_t rows) {
  for (int64_t lr = blockIdx.x; lr < rows; lr += gridDim.x) {
    const int64_t gi = ((lr / m) * p + rank) * m + lr % m;
    const int64_t j0 = (gi >= n) ? 0 : n;
    for (int64_t j = j0 + threadIdx.x; j < ldw; j += blockDim.x)
      W[lr * ldw + j] = (gi == j) ? 1.0 : 0.0;
  }
}

int jb_engine_finish_upload(JbEngine* e, void* stream) {
  JB_CHECK(cudaSetDevice(e->cfg.device));
  cudaStream_t st = (cudaStream_t)stream;
  if (e->rows_loc > 0) {
    int grid = (int)std::min<int64_t>(e->rows_loc, 148 * 16);
    k_pad_identity<<<grid, 256, 0, st>>>(e->W, e->ldw, e->n, e->m, e->p, e->rank, e->rows_loc);
    JB_CHECK(cudaGetLastError());
    e->launches++;
  }
  JB_CHECK(cudaMemsetAsync(e->scratch_max, 0, 8, st));
  // ||A||inf over the real rows only: padded rows have row sum 1 and are excluded by n-wide sums
  // of rows gi < n; rows gi >= n contribute |1| which is harmless only if ||A|| >= 1, so mask them.
  int64_t real_rows = 0;
  for (int64_t lr = 0; lr < e->rows_loc; ++lr)
    if (jb_local_to_global_row(lr, e->m, e->p, e->rank) < e->n) real_rows = lr + 1;
  JB_CHECK(jb_launch_rowsum_max(e->W, real_rows, e->n, e->ldw, e->scratch_max, st));
  e->launches += (real_rows > 0);
  if (e->Acopy)
    JB_CHECK(cudaMemcpyAsync(e->Acopy, e->W, (size_t)e->rows_loc * e->ldw * 8,
                             cudaMemcpyDeviceToDevice, st));
  return JB_OK;
}

int jb_engine_local_norm(JbEngine* e, double* out) {
  JB_CHECK(cudaSetDevice(e->cfg.device));
  JB_CHECK(cudaDeviceSynchronize());
  JB_CHECK(cudaMemcpy(out, e->scratch_max, 8, cudaMemcpyDeviceToHost));
  return JB_OK;
}

int jb_engine_set_norm(JbEngine* e, double global_norm) {
  JB_CHECK(cudaSetDevice(e->cfg.device));
  JB_CHECK(cudaMemcpy(&e->ctl->norm_a, &global_norm, 8, cudaMemcpyHostToDevice));
  e->norm_set = true;
  return JB_OK;
}

int jb_engine_invert(JbEngine* e, void* stream) {
  JB_CHECK(cudaSetDevice(e->cfg.device));
  if (!e->symm_ready || !e->maps_ready) {
    jb_set_error("jb_engine_invert: symmetric region not bound (alloc_symm/bind_symm/open_peers)");
    return JB_ERR_ARG;
  }
  if (!e->norm_set) {
    jb_set_error("jb_engine_invert: jb_engine_set_norm was not called");
    return JB_ERR_ARG;
  }
  cudaStream_t st = (cudaStream_t)stream;
  JB_CHECK(cudaMemsetAsync(e->ctl, 0, offsetof(JbCtl, norm_a), st));

  JB_CHECK(cudaMemsetAsync(e->counters, 0, (size_t)(4 * e->Nr + 4) * 4, st));
  JB_CHECK(cudaMemsetAsync(e->b1sync, 0, 256, st));
  K1Params K;
  K.W = e->W; K.normbuf = e->normbuf; K.okbuf = e->okbuf;
  K.ctl = e->ctl; K.stepinfo = e->stepinfo;
  for (int r = 0; r < JB_MAX_RANKS; ++r)
    K.mail_peers[r] = r < e->p ? (JbMail*)(e->symm_peers[r] + e->off_mail) : nullptr;
  K.ldw = e->ldw; K.m = e->m; K.Kp = e->Kp; K.p = e->p; K.rank = e->rank; K.rloc = e->rloc;

  StepParams S;
  memset(&S, 0, sizeof(S));
  S.W = e->W; S.stepinfo = e->stepinfo; S.ctl = e->ctl;
  S.ldw = e->ldw; S.m = e->m; S.Kp = e->Kp; S.p = e->p; S.rank = e->rank;
  S.rloc = e->rloc; S.nsub = e->nsub; S.CT = e->CT; S.prowflag = e->prowflag;
  const bool la = e->lookahead && e->Nr > 1 && !e->tc05;
  S.b1sync = la ? e->b1sync : nullptr;
  S.early = e->b1sync + 16;
  const int grid = 2 * e->sm_count;
  const int panel_ctas = (int)std::max<int64_t>(1, std::min<int64_t>(32, (e->rows_loc + 15) / 16));
  const size_t nev = (size_t)(8 * e->Nr + 2);
  if (e->profile && e->ev.size() < nev) {
    size_t old = e->ev.size();
    e->ev.resize(nev);
    for (size_t i = old; i < e->ev.size(); ++i) JB_CHECK(cudaEventCreate(&e->ev[i]));
  }
  const uint32_t g0 = e->gstep;
  e->gstep += (uint32_t)e->Nr;
  // Share of the non-priority column tiles that phase B1 updates while K1(t+1) runs beside it.  B1
  // should last about as long as K1 (~1.5 us per elimination column, 2.6 us with a 4-CTA cluster)
  // plus the skew between ranks; everything after it (phase B2) already knows the next pivot.
  double frac = e->b1_frac;
  if (frac < 0) {
    const double waves = std::max(1.0, std::ceil(e->rloc / 128.0));  // candidate CTAs per free SM
    const double k1_us = ((e->m > 128 ? 2.6 : 1.5) * e->m + 20.0) * waves;
    const double b_us = 2.0 * (double)e->rows_loc * (double)e->ldw * e->m / 30e6;
    frac = std::max(0.03, std::min(0.85, 1.3 * k1_us / std::max(b_us, 1.0)));
  }
  auto pri_lo = [&](int64_t t) { return (int)((t * e->m) / JB_BN); };
  auto pri_hi = [&](int64_t t) { return (int)std::min<int64_t>(e->CT - 1, ((t + 2) * e->m - 1) / JB_BN); };

  // K1 of step t: candidates are the local block rows with global index >= t (main.cpp:1028)
  auto launch_k1 = [&](int64_t t, cudaStream_t s) -> cudaError_t {
    const uint32_t g = g0 + (uint32_t)t;
    const int64_t s0 = t > e->rank ? (t - e->rank + e->p - 1) / e->p : 0;
    K.t = (int)t;
    K.s0 = (int)s0;
    K.ncand = (int)std::max<int64_t>(0, e->rloc - s0);
    K.ngroups = std::max(K.ncand, 1);
    K.tag = g + 1;
    K.Lbuf = e->Lbuf[g % 3];
    K.invbuf = e->invbuf[g & 1];
    e->launches += 1;
    cudaError_t ce = jb_launch_k1(K, panel_ctas, s);
    if (ce == cudaSuccess && e->tc05) {
      // integer slices of the L panel for the tcgen05 update of this step
      for (int kb = 0; kb < e->tc_kb && ce == cudaSuccess; ++kb) {
        ce = jb_tc05_slice_rows(e->Lbuf[g % 3] + kb * 128, e->Kp, e->rows_loc, std::min(128, e->m - kb * 128),
                                e->tc_slices, e->Lq[g & 1][kb], e->rows_loc, e->rscale[g & 1][kb], s);
        e->launches += 1;
      }
    }
    return ce;
  };
  // `wide` = every column tile is a priority tile (A then is ROWT + UPD of the whole row)
  auto launch_step = [&](int64_t t, int phase, cudaStream_t s, int use_grid, bool wide = false) -> cudaError_t {
    const uint32_t g = g0 + (uint32_t)t;
    const int par = g & 1;
    S.t = (int)t;
    S.t_slot = (t % e->p == e->rank) ? (int)(t / e->p) : -1;
    S.row_expect = ((g >> 1) + 1) * (uint32_t)e->nsub;
    S.tag = g + 1;
    // SCALE items of phase B2 produce the pivot row of step t + 1 (other buffer parity)
    const int nx = (phase == JB_PHASE_B2) ? 1 : 0;
    S.sc_t = (int)t + nx;
    S.has_next = (phase == JB_PHASE_B2 && t + 1 < e->Nr) ? 1 : 0;
    const int spar = par ^ nx;
    S.invbuf = e->invbuf[spar];
    for (int r = 0; r < JB_MAX_RANKS; ++r) {
      unsigned char* base = r < e->p ? e->symm_peers[r] : nullptr;
      S.rowbuf_peers[r] = base ? (double*)(base + e->off_rowbuf[par]) : nullptr;
      S.rowflag_peers[r] = base ? (uint32_t*)(base + e->off_rowflag[par]) : nullptr;
      S.swapbuf_peers[r] = base ? (double*)(base + e->off_swapbuf[par]) : nullptr;
      S.swapflag_peers[r] = base ? (uint32_t*)(base + e->off_swapflag[par]) : nullptr;
      S.sc_rowbuf_peers[r] = base ? (double*)(base + e->off_rowbuf[spar]) : nullptr;
      S.sc_rowflag_peers[r] = base ? (uint32_t*)(base + e->off_rowflag[spar]) : nullptr;
      S.stepdone_peers[r] = base ? (uint32_t*)(base + e->off_stepdone) : nullptr;
    }
    S.rowbuf_mc = e->symm_mc ? (double*)(e->symm_mc + e->off_rowbuf[par]) : nullptr;
    S.rowflag_mc = e->symm_mc ? (uint32_t*)(e->symm_mc + e->off_rowflag[par]) : nullptr;
    S.sc_rowbuf_mc = e->symm_mc ? (double*)(e->symm_mc + e->off_rowbuf[spar]) : nullptr;
    S.sc_rowflag_mc = e->symm_mc ? (uint32_t*)(e->symm_mc + e->off_rowflag[spar]) : nullptr;
    S.phase = phase;
    // priority column tiles = the ones covering block column t+1 (what K1(t+1) reads)
    // ... plus block column t itself: a swapped-in row takes its L operand from the swap buffer's
    // column block t, so the ROWT pushes of those tiles must be in the same phase
    S.pct0 = pri_lo(t);
    S.pct1 = pri_hi(t);
    // phase A of step t > 0 runs beside B2(t-1) / the tail of B1(t-1): tiles outside the previous
    // priority range are B1(t-1)'s early tiles and are waited for through their flags
    S.prev_pct0 = t > 0 ? pri_lo(t - 1) : 0;
    S.prev_pct1 = t > 0 ? pri_hi(t - 1) : e->CT - 1;
    S.prev_tag = t > 0 ? g : 0;
    if (wide) {  // everything before this launch has completed: no early-tile flags to wait for
      S.pct0 = 0;
      S.pct1 = e->CT - 1;
      S.prev_tag = 0;
    }
    if (e->tc05) {  // no priority tiles: phase A2 = scale/broadcast/swap, phase B = swapped-in row
      S.pct0 = 0;
      S.pct1 = -1;
    }
    const int nnp = e->CT - (S.pct1 >= S.pct0 ? S.pct1 - S.pct0 + 1 : 0);
    // B1's early tiles: what the priority range of step t+1 adds to the one of step t (block
    // column t+2); B1 always covers them
    S.nearly = std::max(0, std::min({nnp, 8, pri_hi(t + 1) - S.pct1}));
    S.nb1 = std::min(nnp, std::max(S.nearly, (int)(frac * nnp + 0.999)));
    if (phase == JB_PHASE_B1 && t + 1 >= e->Nr) S.nb1 = nnp;  // last step: B1 is all that is left
    int cslot = 0;
    switch (phase) {
      case JB_PHASE_B: case JB_PHASE_B1: cslot = 1; break;
      case JB_PHASE_A2: case JB_PHASE_B2: cslot = 2; break;
      case JB_PHASE_S: cslot = 3; break;
      default: cslot = 0;
    }
    S.counter = e->counters + 4 * t + cslot;
    S.only_swapped = (e->tc05 && phase == JB_PHASE_B) ? 1 : 0;
    e->launches += 1;
    return jb_launch_step(S, e->maps[nx][g % 6], use_grid, s);
  };
  // tcgen05 mode: phase B = slice the (negated) pivot row, int8 tensor-core update of every block
  // row except row t and the swapped-in row, which stays on the fp64 kernel (only_swapped)
  auto launch_tc05_b = [&](int64_t t, cudaStream_t s) -> int {
    const uint32_t g = g0 + (uint32_t)t;
    const int par = g & 1;
    const double* rowbuf = (const double*)(e->symm_peers[e->rank] + e->off_rowbuf[par]);
    const uint32_t* rowflag = (const uint32_t*)(e->symm_peers[e->rank] + e->off_rowflag[par]);
    for (int kb = 0; kb < e->tc_kb; ++kb) {
      const int kk = std::min(128, e->m - kb * 128);
      JB_CHECK(jb_tc05_slice_cols(rowbuf + (int64_t)kb * 128 * e->ldw, e->ldw, e->ldw, kk, e->tc_slices,
                                  e->Rq[kb], e->cols_pad, e->cscale[kb], rowflag,
                                  ((g >> 1) + 1) * (uint32_t)e->nsub, e->ctl, s));
      Tc05Launch a;
      memset(&a, 0, sizeof(a));
      a.C = e->W; a.ldc = e->ldw; a.rows = e->rows_loc; a.cols = e->ldw; a.Lq = e->Lq[par][kb];
      a.Rq = e->Rq[kb]; a.rows_pad = e->rows_loc; a.cols_pad = e->cols_pad; a.rscale = e->rscale[par][kb];
      a.cscale = e->cscale[kb]; a.K = kk; a.S = e->tc_slices; a.grid = e->sm_count;
      a.skip0 = (t % e->p == e->rank) ? (t / e->p) * e->m : -1;
      a.skiplen = (t % e->p == e->rank) ? e->m : 0;
      a.ctl = e->ctl; a.stepinfo = e->stepinfo; a.t = (int)t; a.m = e->m; a.rank = e->rank;
      a.col_t0 = (int)(t * e->m); a.col_t1 = (int)((t + 1) * e->m);
      a.zero_slot = (kb == 0);
      if (kb > 0) { a.col_t0 = -1; a.col_t1 = -1; }  // v1: only the first k-block starts from zero
      e->launches += 3;
      int rc5 = jb_tc05_launch(a, s);
      if (rc5) return rc5;
    }
    return JB_OK;
  };

  // ---------------------------------------------------------------- look-ahead schedule
  // Three streams, per step t:
  //   caller's stream : B1(t)   ROWT(t) of the non-priority tiles + UPD(t) of the first few of them
  //                             (starting with block column t+2, flagged when done)
  //   bstream         : B2(t)   launched when K1(t+1) and B1(t) have finished.  Next pivot row first
  //                             (UPDP), SCALE(t+1) -> broadcast for step t+1, then the bulk of UPD(t)
  //   hstream (high)  : A(t+1)  launched together with B2(t) on a small grid: ROWT(t+1) + UPD(t+1)
  //                             of block columns t+1, t+2 as soon as their pivot-row tiles land;
  //                     K1(t+2) right behind it: pending at high priority, so its CTAs (which need
  //                             whole SMs) are placed first when B2(t) retires, before B1(t+1)
  // The only place where the GPU drains is the end of B2(t); the critical chain
  // K1 -> SCALE -> A -> K1 runs a full step ahead of the bulk update.
  const int a_items = (e->rloc * e->nsub) * (pri_hi(1) - pri_lo(1) + 1);
  // Default: phase A beside B2 of the previous step.  At 8 GPUs A is a fraction of a wave and, run
  // in line between two bulk kernels, left most SMs idle for ~40 us per step (0.310 s -> 0.301 s);
  // at 1 GPU the two schedules measure the same.  JB_SCHED=1 selects the in-line variant.
  const bool a_inline = e->sched == 1;
  if (la && a_inline) {
    // hstream: A(t) -> K1(t+1) back to back (K1's CTAs are placed the moment A(t) retires; B1(t)
    // waits for A(t) across streams, so it cannot fill the SMs first); caller's stream: B1(t);
    // bstream: B2(t) behind K1(t+1), moving in while B1(t) drains.
    JB_CHECK(launch_k1(0, st));
    JB_CHECK(launch_step(0, JB_PHASE_S, st, grid));
    JB_CHECK(cudaEventRecord(e->evB1[1], st));
    JB_CHECK(cudaEventRecord(e->evB2[1], st));
    for (int64_t t = 0; t + 1 < e->Nr; ++t) {
      JB_CHECK(cudaStreamWaitEvent(e->hstream, e->evB1[(t + 1) & 1], 0));
      JB_CHECK(cudaStreamWaitEvent(e->hstream, e->evB2[(t + 1) & 1], 0));
      if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * t], e->hstream));
      JB_CHECK(launch_step(t, JB_PHASE_A, e->hstream, grid));
      JB_CHECK(cudaEventRecord(e->evG[t & 1], e->hstream));
      if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * t + 1], e->hstream));
      JB_CHECK(launch_k1(t + 1, e->hstream));
      JB_CHECK(cudaEventRecord(e->evK[(t + 1) & 1], e->hstream));
      if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * t + 2], e->hstream));
      JB_CHECK(cudaStreamWaitEvent(st, e->evG[t & 1], 0));
      if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * t + 3], st));
      JB_CHECK(launch_step(t, JB_PHASE_B1, st, grid));
      JB_CHECK(cudaEventRecord(e->evB1[t & 1], st));
      if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * t + 4], st));
      JB_CHECK(cudaStreamWaitEvent(e->bstream, e->evK[(t + 1) & 1], 0));
      if (!e->b2_overlap) JB_CHECK(cudaStreamWaitEvent(e->bstream, e->evB1[t & 1], 0));
      if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * t + 5], e->bstream));
      JB_CHECK(launch_step(t, JB_PHASE_B2, e->bstream, grid));
      JB_CHECK(cudaEventRecord(e->evB2[t & 1], e->bstream));
      if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * t + 6], e->bstream));
    }
    {  // last step: A over the whole width, nothing to look ahead to
      const int64_t t = e->Nr - 1;
      JB_CHECK(cudaStreamWaitEvent(st, e->evB2[(t + 1) & 1], 0));
      if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * t], st));
      JB_CHECK(launch_step(t, JB_PHASE_A, st, grid, /*wide=*/true));
      if (e->profile)
        for (int k = 1; k <= 6; ++k) JB_CHECK(cudaEventRecord(e->ev[8 * t + k], st));
    }
  } else if (la) {
    const int a_grid = std::max(32, std::min(e->sm_count, a_items / 2));
    JB_CHECK(launch_k1(0, st));
    JB_CHECK(launch_step(0, JB_PHASE_S, st, grid));  // pivot row of step 0: scale + broadcast
    JB_CHECK(cudaEventRecord(e->evG[1], st));
    // A(0) -> K1(1) back to back on hstream, B1(0) behind A(0): K1(1) gets its SMs first
    JB_CHECK(cudaStreamWaitEvent(e->hstream, e->evG[1], 0));
    if (e->profile) JB_CHECK(cudaEventRecord(e->ev[0], e->hstream));
    JB_CHECK(launch_step(0, JB_PHASE_A, e->hstream, grid));
    JB_CHECK(cudaEventRecord(e->evA, e->hstream));
    if (e->profile) JB_CHECK(cudaEventRecord(e->ev[1], e->hstream));
    JB_CHECK(launch_k1(1, e->hstream));
    JB_CHECK(cudaEventRecord(e->evK[1], e->hstream));
    if (e->profile) JB_CHECK(cudaEventRecord(e->ev[2], e->hstream));
    JB_CHECK(cudaStreamWaitEvent(st, e->evA, 0));
    JB_CHECK(cudaStreamWaitEvent(e->bstream, e->evG[1], 0));
    for (int64_t t = 0; t + 1 < e->Nr; ++t) {
      // B1(t) (for t > 0: behind B2(t-1), which ends the previous step on this rank)
      if (t > 0) JB_CHECK(cudaStreamWaitEvent(st, e->evB2[(t + 1) & 1], 0));
      if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * t + 3], st));
      JB_CHECK(launch_step(t, JB_PHASE_B1, st, grid));
      JB_CHECK(cudaEventRecord(e->evB1[t & 1], st));
      if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * t + 4], st));
      // B2(t): needs pivot(t+1); B1(t-1) is long finished (explicit only for the record)
      JB_CHECK(cudaStreamWaitEvent(e->bstream, e->evK[(t + 1) & 1], 0));
      if (t > 0) JB_CHECK(cudaStreamWaitEvent(e->bstream, e->evB1[(t + 1) & 1], 0));
      if (!e->b2_overlap) JB_CHECK(cudaStreamWaitEvent(e->bstream, e->evB1[t & 1], 0));
      JB_CHECK(cudaEventRecord(e->evG[t & 1], e->bstream));
      if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * t + 5], e->bstream));
      JB_CHECK(launch_step(t, JB_PHASE_B2, e->bstream, grid));
      JB_CHECK(cudaEventRecord(e->evB2[t & 1], e->bstream));
      if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * t + 6], e->bstream));
      // A(t+1) beside B2(t), K1(t+2) behind it
      JB_CHECK(cudaStreamWaitEvent(e->hstream, e->evG[t & 1], 0));
      if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * (t + 1)], e->hstream));
      JB_CHECK(launch_step(t + 1, JB_PHASE_A, e->hstream, a_grid));
      if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * (t + 1) + 1], e->hstream));
      if (t + 2 < e->Nr) {
        // K1(t+2) stays pending on the high-priority stream until B2(t) retires and SMs drain:
        // its CTAs (which need whole SMs) are then placed before B1(t+1)'s.  (Releasing it by the
        // end-of-B2 event instead loses that race: 0.320 s against 0.301 s at 8 GPUs.)
        if (!e->k1_early) JB_CHECK(cudaStreamWaitEvent(e->hstream, e->evB2[t & 1], 0));
        JB_CHECK(launch_k1(t + 2, e->hstream));
        JB_CHECK(cudaEventRecord(e->evK[t & 1], e->hstream));
      } else {
        JB_CHECK(cudaEventRecord(e->evA, e->hstream));
      }
      if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * (t + 1) + 2], e->hstream));
    }
    {  // last step: its phase A ran beside B2(Nr-2); B1 takes all that is left, then join
      const int64_t t = e->Nr - 1;
      JB_CHECK(cudaStreamWaitEvent(st, e->evB2[(t + 1) & 1], 0));
      JB_CHECK(cudaStreamWaitEvent(st, e->evA, 0));
      if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * t + 3], st));
      JB_CHECK(launch_step(t, JB_PHASE_B1, st, grid));
      if (e->profile)
        for (int k = 4; k <= 6; ++k) JB_CHECK(cudaEventRecord(e->ev[8 * t + k], st));
    }
  } else {
    JB_CHECK(launch_k1(0, st));
  }
  for (int64_t t = 0; t < e->Nr && !la; ++t) {
    if (e->tc05) {
      // A2 with an EMPTY priority range = scale + broadcast + swap only; then the tcgen05 update of
      // everything, then the fp64 kernel for the swapped-in row (if any)
      if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * t], st));
      JB_CHECK(launch_step(t, JB_PHASE_A2, st, grid));
      if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * t + 1], st));
      if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * t + 3], st));
      int rc5 = launch_tc05_b(t, st);
      if (rc5) return rc5;
      JB_CHECK(launch_step(t, JB_PHASE_B, st, grid));
      if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * t + 4], st));
      if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * t + 5], st));
      if (t + 1 < e->Nr) JB_CHECK(launch_k1(t + 1, st));
      if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * t + 6], st));
    } else {
      if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * t], st));
      JB_CHECK(launch_step(t, JB_PHASE_ALL, st, grid));
      if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * t + 1], st));
      if (e->profile)
        for (int k = 3; k <= 5; ++k) JB_CHECK(cudaEventRecord(e->ev[8 * t + k], st));
      if (t + 1 < e->Nr) JB_CHECK(launch_k1(t + 1, st));
      if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * t + 6], st));
    }
  }
  if (e->profile) JB_CHECK(cudaEventRecord(e->ev[8 * e->Nr], st));
  JB_CHECK(jb_launch_unpermute(e->W, e->rows_loc, e->ldw, e->m, (int)e->Nr, e->stepinfo, st));
  e->launches += (e->rows_loc > 0);
  if (e->profile) {
    JB_CHECK(cudaEventRecord(e->ev[8 * e->Nr + 1], st));
    JB_CHECK(cudaStreamSynchronize(st));
    // kernel durations summed over the steps (event pairs on the stream of each kernel):
    //   look-ahead: [0] A(t), [1] B1(t), [2] B2(t), [3] K1(t+1) incl. its wait for free SMs
    //   serial:     [0] fused step kernel, [2] K1(t+1)
    //   tcgen05:    [0] scale/broadcast/swap, [1] update, [2] K1(t+1)
    for (int k = 0; k < 4; ++k) e->prof_ms[k] = 0;
    for (int64_t t = 0; t < e->Nr; ++t) {
      float x = 0;
      JB_CHECK(cudaEventElapsedTime(&x, e->ev[8 * t], e->ev[8 * t + 1]));
      e->prof_ms[0] += x;
      JB_CHECK(cudaEventElapsedTime(&x, e->ev[8 * t + 3], e->ev[8 * t + 4]));
      e->prof_ms[1] += x;
      JB_CHECK(cudaEventElapsedTime(&x, e->ev[8 * t + 5], e->ev[8 * t + 6]));
      e->prof_ms[2] += x;
      if (la && t + 1 < e->Nr) {
        JB_CHECK(cudaEventElapsedTime(&x, e->ev[8 * t + 1], e->ev[8 * t + 2]));
        e->prof_ms[3] += x;
      }
    }
    float c = 0;
    JB_CHECK(cudaEventElapsedTime(&c, e->ev[8 * e->Nr], e->ev[8 * e->Nr + 1]));
    e->prof_tail_ms = c;
  }
  return JB_OK;
}

int jb_engine_set_profiling(JbEngine* e, int on) {
  e->profile = on != 0;
  return JB_OK;
}

int jb_engine_phase_times(JbEngine* e, double* seg_ms /* [4] */, double* tail_ms) {
  for (int k = 0; k < 4; ++k) seg_ms[k] = e->prof_ms[k];
  *tail_ms = e->prof_tail_ms;
  return JB_OK;
}

int jb_engine_status(JbEngine* e, void* stream) {
  JB_CHECK(cudaSetDevice(e->cfg.device));
  JB_CHECK(cudaStreamSynchronize((cudaStream_t)stream));
  JbCtl c;
  JB_CHECK(cudaMemcpy(&c, e->ctl, sizeof(c), cudaMemcpyDeviceToHost));
  if (c.timeout_flag) {
    jb_set_error("a cross-GPU wait timed out (a peer rank died or fell out of step); results are invalid");
    return JB_ERR_TIMEOUT;
  }
  if (c.singular) {
    jb_set_error("singular matrix (step %d)", c.singular_step);  // main.cpp:438
    return JB_ERR_SINGULAR;
  }
  return JB_OK;
}

int jb_engine_pivots(JbEngine* e, int32_t* piv_out) {
  JB_CHECK(cudaSetDevice(e->cfg.device));
  std::vector<JbStepInfo> si((size_t)e->Nr);
  JB_CHECK(cudaMemcpy(si.data(), e->stepinfo, si.size() * sizeof(JbStepInfo), cudaMemcpyDeviceToHost));
  for (int64_t t = 0; t < e->Nr; ++t) piv_out[t] = si[t].piv_g;
  return JB_OK;
}

int64_t jb_engine_launch_count(const JbEngine* e) { return e->launches; }
void* jb_engine_w_ptr(JbEngine* e) { return e->W; }
int64_t jb_engine_ldw(const JbEngine* e) { return e->ldw; }
int64_t jb_engine_local_rows(const JbEngine* e) { return e->rows_loc; }

int jb_engine_download(JbEngine* e, int which, int64_t local_row0, int64_t nrows, int64_t col0,
                       int64_t ncols, double* host, int64_t ldh) {
  JB_CHECK(cudaSetDevice(e->cfg.device));
  const double* src = which == 0 ? e->W : e->Acopy;
  if (!src || local_row0 < 0 || local_row0 + nrows > e->rows_loc || col0 + ncols > e->ldw) {
    jb_set_error("jb_engine_download: bad region");
    return JB_ERR_ARG;
  }
  if (nrows == 0 || ncols == 0) return JB_OK;
  JB_CHECK(cudaMemcpy2D(host, (size_t)ldh * 8, src + local_row0 * e->ldw + col0, (size_t)e->ldw * 8,
                        (size_t)ncols * 8, (size_t)nrows, cudaMemcpyDeviceToHost));
  return JB_OK;
}

static int ensure_panel(JbEngine* e) {
  if (!e->Xp) {
    if (cudaMalloc((void**)&e->Xp, (size_t)e->n * 1024 * 8) != cudaSuccess) {
      jb_set_error("Not enough memory! (X panel)");
      return JB_ERR_NOMEM;
    }
  }
  return JB_OK;
}

// generic panel loop: for each 1024-wide column panel gather the right operand from the peers'
// slabs `right_peers` and multiply with the left operand.
static int panel_gemm(JbEngine* e, const double* const* right_peers, const double* left,
                      int formula, int mode, double* C, const double* Cin, cudaStream_t st) {
  int rc = ensure_panel(e);
  if (rc) return rc;
  const int PW = 1024;
  for (int64_t col0 = 0; col0 < e->n; col0 += PW) {
    const int pw = (int)std::min<int64_t>(PW, e->ldw - col0);
    GatherParams G;
    for (int r = 0; r < JB_MAX_RANKS; ++r) G.X_peers[r] = r < e->p ? right_peers[r] : nullptr;
    G.ldx = e->ldw; G.Xp = e->Xp; G.ldxp = PW; G.n = e->n; G.col0 = col0; G.m = e->m; G.p = e->p;
    G.pw = pw;
    JB_CHECK(jb_launch_gather_panel(G, st));
    DistGemmParams D;
    D.A = left; D.lda = e->ldw; D.formula = formula; D.Xp = e->Xp; D.ldxp = PW; D.col0 = col0;
    D.pw = pw; D.C = C; D.ldc = e->ldw; D.Cin = Cin; D.ldcin = e->ldw; D.rowsum = e->rowsum;
    D.n = e->n; D.m = e->m; D.p = e->p; D.rank = e->rank; D.rloc = e->rloc; D.mode = mode;
    JB_CHECK(jb_launch_dist_gemm(D, st));
    e->launches += 1 + (e->rows_loc > 0);
  }
  return JB_OK;
}

int jb_engine_residual(JbEngine* e, int formula, double* out_local, void* stream) {
  JB_CHECK(cudaSetDevice(e->cfg.device));
  cudaStream_t st = (cudaStream_t)stream;
  for (int r = 0; r < e->p; ++r)
    if (!e->W_peers[r]) {
      jb_set_error("jb_engine_residual: peer slabs not opened (jb_engine_open_peers(which=1))");
      return JB_ERR_ARG;
    }
  const double* left = nullptr;
  if (formula < 0) {
    if (!e->Acopy) {
      jb_set_error("jb_engine_residual: no kept copy of A (keep_copy=0)");
      return JB_ERR_ARG;
    }
    left = e->Acopy;
  }
  JB_CHECK(cudaMemsetAsync(e->rowsum, 0, (size_t)std::max<int64_t>(e->rows_loc, 1) * 8, st));
  JB_CHECK(cudaMemsetAsync(e->scratch_max, 0, 8, st));
  int rc = panel_gemm(e, e->W_peers, left, formula, JB_GEMM_RESIDUAL, nullptr, nullptr, st);
  if (rc) return rc;
  JB_CHECK(jb_launch_max_reduce(e->rowsum, e->rows_loc, e->scratch_max, st));
  e->launches += (e->rows_loc > 0);
  JB_CHECK(cudaStreamSynchronize(st));
  JB_CHECK(cudaMemcpy(out_local, e->scratch_max, 8, cudaMemcpyDeviceToHost));
  return JB_OK;
}

// Newton-Schulz refinement X <- X + X (I - A X) in fp64 (K7), split in two collective phases so the
// caller can place the cross-rank barriers:
//   phase 0: Rslab = I - A X       (right operand X pulled from every rank's W)
//   -- barrier --
//   phase 1: X = X + X * R         (right operand R pulled from every rank's Rslab)
//   -- barrier --
// For p > 1 exchange the IPC handles of the residual slabs first (jb_engine_ipc_handle /
// jb_engine_open_peers with which = 2).
int jb_engine_refine_phase(JbEngine* e, int formula, int phase, void* stream) {
  JB_CHECK(cudaSetDevice(e->cfg.device));
  cudaStream_t st = (cudaStream_t)stream;
  const size_t slab = (size_t)std::max<int64_t>(e->rows_loc, 1) * e->ldw * 8;
  if (!e->Rslab) {
    if (cudaMalloc((void**)&e->Rslab, 2 * slab) != cudaSuccess) {
      jb_set_error("Not enough memory! (refinement slabs)");
      return JB_ERR_NOMEM;
    }
    JB_CHECK(cudaMemset(e->Rslab, 0, 2 * slab));
    e->R_peers[e->rank] = e->Rslab;
  }
  for (int r = 0; r < e->p; ++r)
    if (!e->W_peers[r] || !e->R_peers[r]) {
      jb_set_error("jb_engine_refine_phase: peer slabs not opened (which = 1 and which = 2)");
      return JB_ERR_ARG;
    }
  double* R = e->Rslab;
  double* Xnew = e->Rslab + slab / 8;
  const double* left = formula < 0 ? e->Acopy : nullptr;
  if (formula < 0 && !left) {
    jb_set_error("jb_engine_refine_phase: no kept copy of A");
    return JB_ERR_ARG;
  }
  if (phase == 0) {
    return panel_gemm(e, e->W_peers, left, formula, JB_GEMM_STORE_R, R, nullptr, st);
  } else {
    int rc = panel_gemm(e, e->R_peers, e->W, 0, JB_GEMM_XPLUS, Xnew, e->W, st);
    if (rc) return rc;
    JB_CHECK(cudaMemcpyAsync(e->W, Xnew, slab, cudaMemcpyDeviceToDevice, st));
    return JB_OK;
  }
}

// Stand-alone batched block inversion (tests, tools): blocks stacked vertically, [nblocks*m][m].
int jb_block_inverse(const double* d_blocks, int64_t nblocks, int32_t m, double norm_a,
                     double* d_inv, double* d_norms, int32_t* d_ok, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (jb_k1_cluster_size(m) == 0 || nblocks <= 0) {
    jb_set_error("jb_block_inverse: unsupported m=%d or nblocks", m);
    return JB_ERR_ARG;
  }
  const int Kp = (int)align_up((size_t)m, JB_BK);
  unsigned char* scratch = nullptr;
  const size_t inv_bytes = (size_t)nblocks * Kp * Kp * 8;
  const size_t lbuf_bytes = (size_t)nblocks * m * Kp * 8;
  const size_t total = inv_bytes + lbuf_bytes + 4096 + (size_t)sizeof(JbStepInfo) * 2;
  if (cudaMalloc((void**)&scratch, total) != cudaSuccess) {
    jb_set_error("Not enough memory! (block inversion staging, %zu bytes)", total);
    return JB_ERR_NOMEM;
  }
  // every exit below goes through `done`, so the staging area is released on errors too
  cudaError_t ce = cudaSuccess;
  K1Params K;
  memset(&K, 0, sizeof(K));
  K.W = d_blocks;
  K.invbuf = (double*)scratch;
  K.Lbuf = (double*)(scratch + inv_bytes);
  unsigned char* tail = scratch + inv_bytes + lbuf_bytes;
  K.ctl = (JbCtl*)tail;
  K.mail_peers[0] = (JbMail*)(tail + 256);
  K.stepinfo = (JbStepInfo*)(tail + 256 + 2 * JB_MAX_RANKS * sizeof(JbMail));
  K.normbuf = d_norms;
  K.okbuf = d_ok;
  K.ldw = m; K.m = m; K.Kp = Kp; K.p = 1; K.rank = 0; K.t = 0; K.rloc = (int)nblocks;
  K.s0 = 0; K.ncand = (int)nblocks; K.ngroups = (int)nblocks; K.tag = 1;
  do {
    if ((ce = cudaMemsetAsync(scratch, 0, total, st)) != cudaSuccess) break;
    if ((ce = cudaMemcpyAsync(&K.ctl->norm_a, &norm_a, 8, cudaMemcpyHostToDevice, st)) != cudaSuccess) break;
    if ((ce = cudaMemsetAsync(d_ok, 0, (size_t)nblocks * 4, st)) != cudaSuccess) break;
    if ((ce = jb_launch_k1(K, 1, st)) != cudaSuccess) break;
    if ((ce = jb_launch_compact_blocks(K.invbuf, Kp, d_inv, m, nblocks, st)) != cudaSuccess) break;
    ce = cudaStreamSynchronize(st);
  } while (0);
  cudaFree(scratch);
  if (ce != cudaSuccess) {
    jb_set_error("jb_block_inverse: %s", cudaGetErrorString(ce));
    return JB_ERR_CUDA;
  }
  return JB_OK;
}

}  // extern "C"
