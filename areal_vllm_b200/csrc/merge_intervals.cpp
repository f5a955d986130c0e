// Host-side merge of touching intervals, exported through the same C-ABI library.
// ref: csrc/interval_op/interval_op.cpp:4-25 — neighbours [a,b),[b,c) collapse to [a,c);
// the list is taken in the order given (no sort), everything else is kept as is.
#include <stddef.h>
#include <stdint.h>

#include "../../include/b200rl.h"

extern "C" size_t b200rl_merge_intervals(const uint64_t* in, size_t n, uint64_t* out) {
  if (n == 0 || in == nullptr || out == nullptr) return 0;
  size_t m = 0;  // index of the interval currently being grown
  uint64_t lo = in[0], hi = in[1];
  for (size_t i = 1; i < n; ++i) {
    const uint64_t s = in[2 * i], e = in[2 * i + 1];
    if (s == hi) {
      hi = e;
    } else {
      out[2 * m] = lo;
      out[2 * m + 1] = hi;
      ++m;
      lo = s;
      hi = e;
    }
  }
  out[2 * m] = lo;
  out[2 * m + 1] = hi;
  return m + 1;
}
