// Text matrix input and corner formatting.
// Format (read_matrix, main.cpp:209-282): n*n whitespace-separated decimal numbers, row-major, no
// header; trailing data ignored.  Errors: -1 cannot open, -2 cannot read (we fail fast instead of
// scattering garbage).  Output (print_row, main.cpp:284-295): "%.2f\t" per value.
#include <errno.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>
#include "../../include/jordan_b200.h"

namespace {

// Buffered strtod scanner: ~10x faster than fscanf("%lf") and same accepted syntax.
struct Scanner {
  FILE* f;
  std::vector<char> buf;
  size_t pos = 0, len = 0;
  bool eof = false;
  explicit Scanner(FILE* fp) : f(fp), buf(1 << 20) {}
  bool refill() {
    size_t keep = len - pos;
    if (keep) memmove(buf.data(), buf.data() + pos, keep);
    pos = 0;
    len = keep;
    if (eof) return false;
    size_t got = fread(buf.data() + len, 1, buf.size() - len - 1, f);
    if (got == 0) eof = true;
    len += got;
    buf[len] = 0;
    return got > 0;
  }
  // returns 1 ok, 0 end of data, -1 garbage
  int next(double* out) {
    for (;;) {
      while (pos < len && (buf[pos] == ' ' || buf[pos] == '\n' || buf[pos] == '\t' ||
                           buf[pos] == '\r' || buf[pos] == '\f' || buf[pos] == '\v'))
        ++pos;
      // make sure a whole token is in the buffer
      if (len - pos < 64 && !eof) {
        refill();
        continue;
      }
      if (pos >= len) return 0;
      char* end = nullptr;
      double v = strtod(buf.data() + pos, &end);
      if (end == buf.data() + pos) return -1;
      pos = (size_t)(end - buf.data());
      *out = v;
      return 1;
    }
  }
};

}  // namespace

extern "C" {

int jb_read_matrix_rows(const char* path, int64_t n, int64_t row0, int64_t nrows, double* out) {
  FILE* f = fopen(path, "r");
  if (!f) return -1;
  Scanner sc(f);
  sc.buf[0] = 0;
  double v;
  for (int64_t i = 0; i < row0 * n; ++i) {
    if (sc.next(&v) != 1) {
      fclose(f);
      return -2;
    }
  }
  for (int64_t i = 0; i < nrows * n; ++i) {
    if (sc.next(&v) != 1) {
      fclose(f);
      return -2;
    }
    out[i] = v;
  }
  fclose(f);
  return 0;
}

// Streaming reader: the file is scanned once, a caller-chosen number of values at a time (the
// reference reads one block row of m*n numbers, sends it on, and re-uses the buffer,
// main.cpp:249-263) - the reading rank never holds more than one block row.
struct JbMatrixReader {
  FILE* f;
  Scanner sc;
  explicit JbMatrixReader(FILE* fp) : f(fp), sc(fp) { sc.buf[0] = 0; }
};

int jb_matrix_reader_open(const char* path, JbMatrixReader** out) {
  *out = nullptr;
  FILE* f = fopen(path, "r");
  if (!f) return -1;
  *out = new JbMatrixReader(f);
  return 0;
}

int jb_matrix_reader_next(JbMatrixReader* r, int64_t count, double* out) {
  if (!r) return -1;
  for (int64_t i = 0; i < count; ++i)
    if (r->sc.next(out + i) != 1) return -2;
  return 0;
}

void jb_matrix_reader_close(JbMatrixReader* r) {
  if (!r) return;
  fclose(r->f);
  delete r;
}

int jb_read_matrix_full(const char* path, int64_t n, double* out) {
  return jb_read_matrix_rows(path, n, 0, n, out);
}

int64_t jb_format_corner(const double* a, int64_t rows, int64_t cols, int64_t ld, char* buf,
                         int64_t buflen) {
  int64_t w = 0;
  for (int64_t i = 0; i < rows; ++i) {
    for (int64_t j = 0; j < cols; ++j) {
      int k = snprintf(buf + w, (size_t)(buflen - w), "%.2f\t", a[i * ld + j]);
      if (k < 0 || w + k >= buflen) return -1;
      w += k;
    }
    if (w + 1 >= buflen) return -1;
    buf[w++] = '\n';
    buf[w] = 0;
  }
  return w;
}

}  // extern "C"
