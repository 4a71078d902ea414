// K3-pre: QAM query/key frame similarity, once per cascade level.
// Replaces PPMStereo.compute_qk_similarity (reference models/core/ppmstereo.py:397-423):
//   x_[b,t,:]   = mean_c AdaptiveMaxPool2d((h//4, w//4))(x[b,:,t])          for x = query, key
//   sim[b,i,j]  = <k_[b,i], q_[b,j]> / (max(|k_[b,i]|, eps) * max(|q_[b,j]|, eps)),  eps = 1e-8
// fp32 arithmetic (inputs fp32 or bf16), fixed reduction orders => run-to-run deterministic, which matters
// because sim feeds the top-k frame pick.  dtype rule: for bf16 inputs (the reference's autocast regime) the pooled
// channel mean is rounded to bf16 -- `q_.mean(dim=1)` of a bf16 tensor returns bf16 (:409), only
// F.cosine_similarity up-casts; for fp32 inputs nothing is rounded.  HBM-bound: both tensors are read exactly once.
//   pool kernel  : CTA = (pooled row, frame); thread = pooled column; loops over the 128 channels and the
//                  adaptive window (start = floor(i*n/o), end = ceil((i+1)*n/o)); consecutive threads read
//                  consecutive windows of an image row, so requests are coalesced.
//   cosine kernel: CTA = (i, j, b); block-wide tree reduction of the dot product and the two norms.
#include <math.h>

#include "../../include/ppm_b200.h"
#include "common.cuh"

namespace ppm {

constexpr int kSimPoolThreads = 128;
constexpr int kSimCosThreads = 256;

__device__ __forceinline__ float sim_ld(const float* p) { return __ldg(p); }
__device__ __forceinline__ float sim_ld(const __nv_bfloat16* p) { return __bfloat162float(__ldg(p)); }

template <typename T>
__global__ void __launch_bounds__(kSimPoolThreads)
sim_pool_kernel(const T* __restrict__ x, long long sb, long long sc, long long st, int Tn, int C, int h, int w,
                int oh, int ow, float* __restrict__ pooled) {
  const int oy = blockIdx.x;
  const int frame = blockIdx.y;                          // b*T + t
  const int bb = frame / Tn, t = frame - bb * Tn;
  const int r0 = (oy * h) / oh, r1 = ((oy + 1) * h + oh - 1) / oh;
  const T* src = x + bb * sb + t * st;
  for (int ox = threadIdx.x; ox < ow; ox += kSimPoolThreads) {
    const int c0 = (ox * w) / ow, c1 = ((ox + 1) * w + ow - 1) / ow;
    float acc = 0.f;
    for (int c = 0; c < C; ++c) {
      const T* pc = src + c * sc;
      float m = -INFINITY;
      for (int r = r0; r < r1; ++r)
        for (int q = c0; q < c1; ++q) m = fmaxf(m, sim_ld(pc + r * w + q));
      acc += m;
    }
    const float m = __fmul_rn(acc, 1.0f / static_cast<float>(C));
    pooled[(static_cast<long long>(frame) * oh + oy) * ow + ox] = sizeof(T) == 2 ? bf16_round_f(m) : m;
  }
}

__device__ __forceinline__ float sim_block_sum(float v, float* sh) {
  #pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();                                        // sh may still be read from the previous call
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = threadIdx.x < kSimCosThreads / 32 ? sh[threadIdx.x] : 0.f;
  #pragma unroll
  for (int o = 4; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  return __shfl_sync(0xffffffffu, t, 0);                 // warp 0's lanes hold the total; others unused
}

__global__ void __launch_bounds__(kSimCosThreads)
sim_cosine_kernel(const float* __restrict__ q_, const float* __restrict__ k_, int Tn, int P, float* __restrict__ sim) {
  __shared__ float sh[kSimCosThreads / 32];
  __shared__ float res[3];
  const int j = blockIdx.x, i = blockIdx.y, bb = blockIdx.z;
  const float* kv = k_ + (static_cast<long long>(bb) * Tn + i) * P;
  const float* qv = q_ + (static_cast<long long>(bb) * Tn + j) * P;
  float d = 0.f, nk = 0.f, nq = 0.f;
  for (int p = threadIdx.x; p < P; p += kSimCosThreads) {
    const float a = kv[p], c = qv[p];
    d = fmaf(a, c, d); nk = fmaf(a, a, nk); nq = fmaf(c, c, nq);
  }
  const float sd = sim_block_sum(d, sh);
  if (threadIdx.x == 0) res[0] = sd;
  const float sk = sim_block_sum(nk, sh);
  if (threadIdx.x == 0) res[1] = sk;
  const float sq = sim_block_sum(nq, sh);
  if (threadIdx.x == 0) {
    const float a = fmaxf(sqrtf(res[1]), 1e-8f), c = fmaxf(sqrtf(sq), 1e-8f);
    sim[(static_cast<long long>(bb) * Tn + i) * Tn + j] = res[0] / (a * c);
  }
}

template <typename T>
static int launch_pool(const void* x, long long sb, long long sc, long long st, int b, int Tn, int h, int w,
                       int oh, int ow, float* pooled, cudaStream_t s) {
  sim_pool_kernel<T><<<dim3(oh, b * Tn), kSimPoolThreads, 0, s>>>(static_cast<const T*>(x), sb, sc, st, Tn, 128, h, w,
                                                                  oh, ow, pooled);
  PPM_CUDA(cudaGetLastError());
  return PPM_OK;
}

}  // namespace ppm

extern "C" int ppm_qk_similarity(const void* query, const void* key, int dtype, int64_t sb, int64_t sc, int64_t st,
                                 int b, int T, int h, int w, float* pooled_ws, float* sim, void* stream) {
  using namespace ppm;
  PPM_CHECK(query && key && pooled_ws && sim, PPM_ERR_ARG, "ppm_qk_similarity: null pointer");
  PPM_CHECK(b > 0 && T > 0 && h >= 4 && w >= 4, PPM_ERR_ARG, "ppm_qk_similarity: bad shape (h, w must be >= 4)");
  PPM_CHECK(dtype == PPM_F32 || dtype == PPM_BF16, PPM_ERR_ARG, "ppm_qk_similarity: bad dtype");
  PPM_CHECK(b * T <= 65535 && T <= 65535 && b <= 65535, PPM_ERR_UNSUPPORTED, "ppm_qk_similarity: b*T > 65535");
  const int oh = h / 4, ow = w / 4, P = oh * ow;
  float* q_ = pooled_ws;
  float* k_ = pooled_ws + static_cast<long long>(b) * T * P;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  int rc;
  if (dtype == PPM_BF16) {
    if ((rc = launch_pool<__nv_bfloat16>(query, sb, sc, st, b, T, h, w, oh, ow, q_, s))) return rc;
    if ((rc = launch_pool<__nv_bfloat16>(key, sb, sc, st, b, T, h, w, oh, ow, k_, s))) return rc;
  } else {
    if ((rc = launch_pool<float>(query, sb, sc, st, b, T, h, w, oh, ow, q_, s))) return rc;
    if ((rc = launch_pool<float>(key, sb, sc, st, b, T, h, w, oh, ow, k_, s))) return rc;
  }
  sim_cosine_kernel<<<dim3(T, T, b), kSimCosThreads, 0, s>>>(q_, k_, T, P, sim);
  PPM_CUDA(cudaGetLastError());
  return PPM_OK;
}
