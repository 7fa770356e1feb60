// Block-cyclic index math, 64-bit.  Closed forms equivalent to the reference's
// rows_p_process (main.cpp:95-116), local_to_global (118-123), num_block_rows (124-127) and
// find_sender (526-532).
#include "../../include/jordan_b200.h"
#include "../kernels_formula.h"

extern "C" {

int64_t jb_num_block_rows(int64_t n, int64_t m) { return (n + m - 1) / m; }

// rank owning the last block row: (Nr % p - 1) mod p
int64_t jb_find_sender(int64_t Nr, int64_t p) {
  int64_t s = Nr % p - 1;
  if (s < 0) s += p;
  return s;
}

// number of block rows I with I % p == k
int64_t jb_rows_per_process(int64_t Nr, int64_t p, int64_t k) {
  if (k >= p || k < 0) return 0;
  return Nr / p + ((k < Nr % p) ? 1 : 0);
}

int64_t jb_local_to_global_row(int64_t i, int64_t m, int64_t p, int64_t k) {
  return ((i / m) * p + k) * m + i % m;
}

int64_t jb_global_to_local_row(int64_t gi, int64_t m, int64_t p) {
  return ((gi / m) / p) * m + gi % m;
}

int64_t jb_owner_of_row(int64_t gi, int64_t m, int64_t p) { return (gi / m) % p; }

// leading dimension: whole blocks, rounded up to 16 doubles (128 B) so every row is TMA-aligned
int64_t jb_padded_ld(int64_t n, int64_t m) {
  int64_t w = jb_num_block_rows(n, m) * m;
  return (w + 15) / 16 * 16;
}

double jb_formula(int id, int64_t i, int64_t j, int64_t n) { return jb_formula_eval(id, i, j, n); }

}  // extern "C"
