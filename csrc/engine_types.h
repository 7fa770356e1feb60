// Shared host/device control structures of the engine (not part of the public C ABI).
#pragma once
#include <stdint.h>

#define JB_EPS 1e-15          // relative singularity threshold, reference main.cpp:7 / :782
#define JB_MAX_RANKS 16
#define JB_BM 64              // update tile rows
#define JB_BN 128             // update tile cols
#define JB_BK 16              // K chunk per pipeline stage
#define JB_LS 20              // smem row stride (doubles) of an L chunk  (16 + 4 pad -> conflict-free)
#define JB_RS 132             // smem row stride (doubles) of an R chunk  (128 + 4 pad -> conflict-free)
#define JB_STAGES 4

// One pivot candidate as exchanged between ranks. Mirrors the reference's
// struct PivotMin {double min; int k; int min_row; int non_sing;} (main.cpp:35-42) plus a
// step tag that makes the slot self-validating (no separate flag store).
struct __attribute__((aligned(32))) JbMail {
  double norm;
  int32_t rank;
  int32_t slot;       // LOCAL block-row index on `rank`
  int32_t nonsing;
  uint32_t tag;       // written last with st.release.sys
  uint32_t pad[2];
};

// Result of the pivot search of step t, left in device memory for the kernels of the step.
struct JbStepInfo {
  int32_t piv_g;       // global block row chosen as pivot
  int32_t piv_rank;    // owner rank of that row
  int32_t piv_slot;    // local slot on the owner
  int32_t do_swap;     // piv_g != t
  uint32_t swap_expect;  // value swapflag[] must reach before swapped-in tiles may be read
  uint32_t pad[3];
};

// Device-resident control block.
struct JbCtl {
  int32_t singular;        // set when no rank had a non-singular candidate (main.cpp:1075)
  int32_t singular_step;
  uint32_t ticket;         // K1 arrival counter (last candidate CTA performs the min-loc)
  uint32_t swap_count;     // number of block-row swaps performed so far (monotonic)
  int32_t timeout_flag;    // a spin-wait gave up (peer dead / protocol error): results are invalid
  int32_t pad0;
  double norm_a;           // GLOBAL ||A||_inf (the reference uses the rank-local one)
  double residual_bits;    // scratch for atomic max of |row sums|
  unsigned long long spin_timeout_ns;  // watchdog for cross-GPU spin-waits (0 = wait forever)
};
