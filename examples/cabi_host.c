/* A torch-free host for the C ABI (include/b200rl.h): plain C, CUDA runtime only.
 * It does what a pybind / cgo / JNI binding of the reference's extension modules would do:
 *   - csrc/cugae/gae.cu:30-60 (gae_1d_nolp_misalign)      -> b200rl_gae_1d_nolp_misalign
 *   - csrc/interval_op/interval_op.cu:28-105 (slice / set) -> b200rl_slice_intervals / b200rl_set_intervals
 * and checks the results against a scalar restatement written here (the reference's loops,
 * ppo_functional.py:292-323 and flatten_param.py:65-90). Exit code 0 = all equal within the bars
 * (GAE 1e-5, intervals bit-exact).
 *
 * Build (tests/test_cabi_c_host.py does this):
 *   gcc -O2 -std=c11 -Iinclude -I/usr/local/cuda/include examples/cabi_host.c \
 *       -Lareal_vllm_b200 -lb200rl -L/usr/local/cuda/lib64 -lcudart -lm -o examples/cabi_host
 */
#include <cuda_runtime_api.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "b200rl.h"

#define CK(call)                                                                         \
  do {                                                                                   \
    cudaError_t e_ = (call);                                                             \
    if (e_ != cudaSuccess) {                                                             \
      fprintf(stderr, "%s:%d %s: %s\n", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
      return 2;                                                                          \
    }                                                                                    \
  } while (0)
#define RC(call)                                                                   \
  do {                                                                             \
    int rc_ = (call);                                                              \
    if (rc_ != 0) {                                                                \
      fprintf(stderr, "%s:%d %s -> %d: %s\n", __FILE__, __LINE__, #call, rc_, b200rl_last_error()); \
      return 3;                                                                    \
    }                                                                              \
  } while (0)

static uint32_t rng_state = 12345u;
static float frand(void) { /* uniform in [-1, 1) */
  rng_state = rng_state * 1664525u + 1013904223u;
  return (float)((rng_state >> 8) & 0xffffff) / 8388608.0f - 1.0f;
}

static int check_gae(int mode, float gamma, float lam) {
  enum { B = 37 };
  int32_t cu[B + 1];
  uint8_t boot[B];
  cu[0] = 0;
  for (int i = 0; i < B; i++) {
    rng_state = rng_state * 1664525u + 1013904223u;
    cu[i + 1] = cu[i] + 1 + (int)((rng_state >> 10) % 700u);
    boot[i] = (uint8_t)((rng_state >> 5) & 1u);
  }
  const int total = cu[B];
  float* r = (float*)malloc(sizeof(float) * total);
  float* v = (float*)malloc(sizeof(float) * (total + B));
  float* adv_ref = (float*)malloc(sizeof(float) * total);
  float* ret_ref = (float*)malloc(sizeof(float) * total);
  float* adv = (float*)malloc(sizeof(float) * total);
  float* ret = (float*)malloc(sizeof(float) * total);
  for (int t = 0; t < total; t++) r[t] = frand();
  for (int t = 0; t < total + B; t++) v[t] = frand();
  /* ppo_functional.py:292-323, in double so that the 1e-5 bar is against exact arithmetic */
  for (int i = 0; i < B; i++) {
    const int rs = cu[i], n = cu[i + 1] - cu[i], vs = cu[i] + i;
    double last = 0.0;
    for (int t = n - 1; t >= 0; t--) {
      const double nv = (t == n - 1 && !boot[i]) ? 0.0 : (double)v[vs + t + 1];
      const double d = (double)r[rs + t] + (double)gamma * nv - (double)v[vs + t];
      last = d + (double)gamma * (double)lam * last;
      adv_ref[rs + t] = (float)last;
      ret_ref[rs + t] = (float)(last + (double)v[vs + t]);
    }
  }
  float *dr, *dv, *dadv, *dret;
  int32_t* dcu;
  uint8_t* dboot;
  void* ws;
  const size_t ws_bytes = b200rl_gae_workspace_bytes(B, total);
  CK(cudaMalloc((void**)&dr, sizeof(float) * total));
  CK(cudaMalloc((void**)&dv, sizeof(float) * (total + B)));
  CK(cudaMalloc((void**)&dadv, sizeof(float) * total));
  CK(cudaMalloc((void**)&dret, sizeof(float) * total));
  CK(cudaMalloc((void**)&dcu, sizeof(cu)));
  CK(cudaMalloc((void**)&dboot, sizeof(boot)));
  CK(cudaMalloc(&ws, ws_bytes));
  CK(cudaMemcpy(dr, r, sizeof(float) * total, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dv, v, sizeof(float) * (total + B), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dcu, cu, sizeof(cu), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dboot, boot, sizeof(boot), cudaMemcpyHostToDevice));
  cudaStream_t st;
  CK(cudaStreamCreate(&st));
  RC(b200rl_gae_1d_nolp_misalign(dr, dv, dcu, dboot, dadv, dret, B, total, gamma, lam, mode, ws, ws_bytes, st));
  CK(cudaMemcpyAsync(adv, dadv, sizeof(float) * total, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(ret, dret, sizeof(float) * total, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  double worst = 0.0;
  for (int t = 0; t < total; t++) {
    const double ea = fabs((double)adv[t] - adv_ref[t]) / (1.0 + fabs((double)adv_ref[t]));
    const double er = fabs((double)ret[t] - ret_ref[t]) / (1.0 + fabs((double)ret_ref[t]));
    if (ea > worst) worst = ea;
    if (er > worst) worst = er;
  }
  printf("gae mode=%d gamma=%.2f lam=%.2f tokens=%d worst scaled error %.3g\n", mode, gamma, lam, total, worst);
  cudaFree(dr); cudaFree(dv); cudaFree(dadv); cudaFree(dret); cudaFree(dcu); cudaFree(dboot); cudaFree(ws);
  cudaStreamDestroy(st);
  free(r); free(v); free(adv_ref); free(ret_ref); free(adv); free(ret);
  return worst < 2e-5 ? 0 : 1;
}

static int check_intervals(void) {
  enum { N = 3000, ES = 2 };
  const int64_t big = 4000000;
  int64_t* iv = (int64_t*)malloc(sizeof(int64_t) * 2 * N);
  int64_t pos = 3, total = 0;
  for (int i = 0; i < N; i++) {
    rng_state = rng_state * 1664525u + 1013904223u;
    pos += (int64_t)((rng_state >> 12) % 50u);
    const int64_t len = (int64_t)((rng_state >> 4) % 1200u);
    iv[2 * i] = pos;
    iv[2 * i + 1] = pos + len;
    pos += len;
    total += len;
  }
  if (pos > big) return 4;
  /* shuffle the order of the intervals (the packed order follows the list) */
  for (int i = N - 1; i > 0; i--) {
    rng_state = rng_state * 1664525u + 1013904223u;
    const int j = (int)((rng_state >> 8) % (uint32_t)(i + 1));
    int64_t a = iv[2 * i], b = iv[2 * i + 1];
    iv[2 * i] = iv[2 * j]; iv[2 * i + 1] = iv[2 * j + 1];
    iv[2 * j] = a; iv[2 * j + 1] = b;
  }
  uint16_t* src = (uint16_t*)malloc(ES * big);
  uint16_t* packed_ref = (uint16_t*)malloc(ES * total);
  uint16_t* packed = (uint16_t*)malloc(ES * total);
  uint16_t* dst_ref = (uint16_t*)calloc(big, ES);
  uint16_t* dst = (uint16_t*)malloc(ES * big);
  for (int64_t k = 0; k < big; k++) {
    rng_state = rng_state * 1664525u + 1013904223u;
    src[k] = (uint16_t)(rng_state >> 13);
  }
  int64_t off = 0;  /* flatten_param.py:65-74 then :77-90 */
  for (int i = 0; i < N; i++) {
    const int64_t len = iv[2 * i + 1] - iv[2 * i];
    memcpy(packed_ref + off, src + iv[2 * i], ES * len);
    memcpy(dst_ref + iv[2 * i], packed_ref + off, ES * len);
    off += len;
  }
  void *dsrc, *dpacked, *ddst, *ws;
  int64_t* div_;
  const size_t ws_bytes = b200rl_interval_workspace_bytes(N, total, ES);
  CK(cudaMalloc(&dsrc, ES * big));
  CK(cudaMalloc(&dpacked, ES * total));
  CK(cudaMalloc(&ddst, ES * big));
  CK(cudaMalloc((void**)&div_, sizeof(int64_t) * 2 * N));
  CK(cudaMalloc(&ws, ws_bytes));
  CK(cudaMemcpy(dsrc, src, ES * big, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(div_, iv, sizeof(int64_t) * 2 * N, cudaMemcpyHostToDevice));
  CK(cudaMemset(ddst, 0, ES * big));
  RC(b200rl_slice_intervals(dsrc, big, dpacked, total, div_, N, ES, ws, ws_bytes, 0, NULL));
  RC(b200rl_set_intervals(dpacked, total, ddst, big, div_, N, ES, ws, ws_bytes, 1, NULL)); /* same plan */
  CK(cudaMemcpy(packed, dpacked, ES * total, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(dst, ddst, ES * big, cudaMemcpyDeviceToHost));
  const int same = memcmp(packed, packed_ref, ES * total) == 0 && memcmp(dst, dst_ref, ES * big) == 0;
  /* host-side merge of touching neighbours (interval_op.cpp:4-25) */
  uint64_t in[8] = {0, 4, 4, 9, 20, 30, 30, 31}, out[8];
  const size_t m = b200rl_merge_intervals(in, 4, out);
  const int merged_ok = m == 2 && out[0] == 0 && out[1] == 9 && out[2] == 20 && out[3] == 31;
  printf("intervals n=%d elems=%lld bit-exact=%d merge=%d\n", N, (long long)total, same, merged_ok);
  cudaFree(dsrc); cudaFree(dpacked); cudaFree(ddst); cudaFree(div_); cudaFree(ws);
  free(iv); free(src); free(packed_ref); free(packed); free(dst_ref); free(dst);
  return same && merged_ok ? 0 : 1;
}

int main(void) {
  printf("libb200rl ABI version %d\n", b200rl_version());
  int bad = 0;
  bad |= check_gae(B200RL_GAE_SERIAL, 1.0f, 1.0f);
  bad |= check_gae(B200RL_GAE_SCAN, 0.99f, 0.95f);
  bad |= check_gae(B200RL_GAE_AUTO, 0.9f, 0.9f);
  bad |= check_intervals();
  /* argument validation returns codes, never throws or aborts */
  if (b200rl_gae_1d_nolp_misalign(NULL, NULL, NULL, NULL, NULL, NULL, 1, 5, 1.f, 1.f, 0, NULL, 0, NULL) == 0) bad |= 1;
  printf(bad ? "FAILED\n" : "OK\n");
  return bad;
}
