/* CPU restatement (plain C) of the reference's GAE and interval algorithms.
 * TEST INFRASTRUCTURE — never linked into or called by the product library.
 * Build: see oracle/Makefile (gcc -O2 -ffp-contract=off; contraction off so that every fp32
 * operation rounds separately, exactly like the scalar torch ops of the python reference).
 */
#include <stdint.h>
#include <string.h>
#include <math.h>

/* GAE, python-reference semantics.
 * ref: realhf/impl/model/utils/ppo_functional.py:292-323 (pygae1d_nolp_misalign):
 *   delta   = rewards[t] + gamma * nextvalues - values[t]      (:316, left to right, fp32 each)
 *   lastgae = delta + gamma * lam * lastgae                    (:317; `gamma * lam` is a python
 *             double product, rounded to fp32 when it meets the fp32 tensor)
 *   nextvalues *= bootstrap[i] at the last step               (:313-315; here applied to a copy —
 *             the reference mutates `values` through a view)
 * values of sequence i start at cu[i] + i (:300-301). */
void oracle_gae_py(const float* rewards, const float* values, const int32_t* cu,
                   const float* bootstrap, int batch, double gamma, double lam, float* adv,
                   float* ret) {
  const float g = (float)gamma;
  const float c = (float)(gamma * lam);
  for (int i = batch - 1; i >= 0; --i) {
    const int rs = cu[i], re = cu[i + 1];
    const long vs = (long)rs + i;
    float last = 0.0f;
    for (int t = re - rs - 1; t >= 0; --t) {
      float nv = values[vs + t + 1];
      if (t == re - rs - 1) nv = nv * bootstrap[i];
      float d = g * nv;
      d = rewards[rs + t] + d;
      d = d - values[vs + t];
      float m = c * last;
      last = d + m;
      adv[rs + t] = last;
      ret[rs + t] = last + values[vs + t];
    }
  }
}

/* GAE, CUDA-reference semantics.
 * ref: csrc/cugae/gae.cu:10-28 compiled by nvcc with its default -fmad=true:
 *   nex_v  = last && !bootstrap ? 0 : values[...]          (:21, a select, not a product)
 *   delta  = fma(gamma, nex_v, rewards) - values           (:22)
 *   lastgae= fma(gamma*lmbda, lastgae, delta)              (:23, gamma*lmbda in fp32)
 *   ret    = lastgae + values                              (:25) */
void oracle_gae_cuda(const float* rewards, const float* values, const int32_t* cu,
                     const uint8_t* bootstrap, int batch, float gamma, float lam, float* adv,
                     float* ret) {
  const float c = gamma * lam;
  for (int i = 0; i < batch; ++i) {
    const int rs = cu[i], re = cu[i + 1];
    const long vs = (long)rs + i;
    float last = 0.0f;
    for (int t = re - rs - 1; t >= 0; --t) {
      const float nv = (t == re - rs - 1 && !bootstrap[i]) ? 0.0f : values[vs + t + 1];
      const float d = fmaf(gamma, nv, rewards[rs + t]) - values[vs + t];
      last = fmaf(c, last, d);
      adv[rs + t] = last;
      ret[rs + t] = last + values[vs + t];
    }
  }
}

/* GAE in double precision (error budgeting only). */
void oracle_gae_f64(const float* rewards, const float* values, const int32_t* cu,
                    const uint8_t* bootstrap, int batch, double gamma, double lam, double* adv,
                    double* ret) {
  const double c = gamma * lam;
  for (int i = 0; i < batch; ++i) {
    const int rs = cu[i], re = cu[i + 1];
    const long vs = (long)rs + i;
    double last = 0.0;
    for (int t = re - rs - 1; t >= 0; --t) {
      const double nv = (t == re - rs - 1 && !bootstrap[i]) ? 0.0 : (double)values[vs + t + 1];
      const double d = (double)rewards[rs + t] + gamma * nv - (double)values[vs + t];
      last = d + c * last;
      adv[rs + t] = last;
      ret[rs + t] = last + (double)values[vs + t];
    }
  }
}

/* interval gather: out = cat(src[s:e] for (s,e) in intervals).
 * ref: realhf/impl/model/nn/flatten_param.py:65-74 (_slice_intervals_py). Returns elements written. */
int64_t oracle_slice_intervals(const void* src, int64_t src_numel, void* dst,
                               const int64_t* intervals, int64_t n, int elem_size) {
  int64_t off = 0;
  for (int64_t i = 0; i < n; ++i) {
    const int64_t s = intervals[2 * i], e = intervals[2 * i + 1];
    if (s < 0 || e > src_numel || e < s) return -1; /* :68-69 asserts */
    memcpy((char*)dst + off * elem_size, (const char*)src + s * elem_size, (size_t)(e - s) * elem_size);
    off += e - s;
  }
  return off;
}

/* interval scatter: dst[s:e] = src[off:off+e-s].
 * ref: realhf/impl/model/nn/flatten_param.py:77-90 (_set_intervals_py). Returns elements read. */
int64_t oracle_set_intervals(const void* src, void* dst, int64_t dst_numel,
                             const int64_t* intervals, int64_t n, int elem_size) {
  int64_t off = 0;
  for (int64_t i = 0; i < n; ++i) {
    const int64_t s = intervals[2 * i], e = intervals[2 * i + 1];
    if (s < 0 || e > dst_numel || e < s) return -1; /* :86-87 asserts */
    memcpy((char*)dst + s * elem_size, (const char*)src + off * elem_size, (size_t)(e - s) * elem_size);
    off += e - s;
  }
  return off; /* caller asserts off == len(src), :90 */
}

/* ref: csrc/interval_op/interval_op.cpp:4-25 (merge_intervals). out has room for n pairs. */
int64_t oracle_merge_intervals(const uint64_t* in, int64_t n, uint64_t* out) {
  int64_t m = 0;
  for (int64_t i = 0; i < n; ++i) {
    if (m > 0 && out[2 * (m - 1) + 1] == in[2 * i]) {
      out[2 * (m - 1) + 1] = in[2 * i + 1];
    } else {
      out[2 * m] = in[2 * i];
      out[2 * m + 1] = in[2 * i + 1];
      ++m;
    }
  }
  return m;
}


/* Padded 2-D GAE, csrc/cugae/gae.cu:62-92 (gae_kernel_2d_olp), one (row, segment) at a time,
 * with nvcc's FMA contraction written out. */
void oracle_gae_2d_olp(const float* rewards, const float* values, const uint8_t* done,
                       const int32_t* done_y, const int32_t* cu_dones, const uint8_t* is_truncate,
                       int batch, int horizon, float gamma, float lam, float* adv, float* ret) {
  const float c = gamma * lam;
  for (int b = 0; b < batch; ++b) {
    const int c0 = cu_dones[b];
    const int breadth = cu_dones[b + 1] - c0 + 1;
    const float* r = rewards + (long)b * horizon;
    const float* v = values + (long)b * (horizon + 1);
    const uint8_t* dn = done + (long)b * (horizon + 1);
    for (int s = 0; s < breadth; ++s) {
      const int t_start = s == 0 ? 0 : done_y[c0 + s - 1];
      const int t_end = s == breadth - 1 ? horizon : done_y[c0 + s];
      const int abandon = s != breadth - 1 && is_truncate[c0 + s];
      float last = 0.0f;
      for (int t = t_end - 1; t >= t_start; --t) {
        const float nv = (t == t_end - 1 && dn[t + 1]) ? 0.0f : v[t + 1];
        const float d = fmaf(gamma, nv, r[t]) - v[t];
        last = fmaf(c, last, d);
        if (t == t_end - 1 && abandon) last = 0.0f;
        adv[(long)b * horizon + t] = last;
        ret[(long)b * horizon + t] = last + v[t];
      }
    }
  }
}
