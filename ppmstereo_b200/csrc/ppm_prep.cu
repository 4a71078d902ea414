// K3a/K3b: QAM frame scoring + top-k pick, and operand packing for the memory read.
// Replaces reference models/core/ppmstereo.py:501-513 (score, pick, strive update),
// :518-548 (query + PE, masked_select gather, "play" modulation, flatten/permute/cast).
// All HBM-bound helper kernels; together ~0.5 % of the memory-read time at 720p.
#include <math.h>
#include <mutex>

#include "../../include/ppm_b200.h"
#include "common.cuh"

namespace ppm {

__device__ __forceinline__ float ld_as_float(const float* p) { return __ldg(p); }
__device__ __forceinline__ float ld_as_float(const __nv_bfloat16* p) { return __bfloat162float(__ldg(p)); }

// ------------------------------------------------------------------ frame confidence
// conf[f] = mean over hw of uncertainty[f, :]   (ppmstereo.py:506).  One CTA per frame,
// fixed reduction tree => deterministic.  dtype rule (include/ppm_b200.h, ppm_qam_pick): torch.mean is not an
// autocast-fp32 op, so for the bf16 uncertainty map of the reference's autocast regime `.mean(-1)` returns a
// bf16 tensor -- fp32 accumulation, sum * (1/hw) (ATen MeanOps), ONE rounding to bf16.  Mirrored here.
template <typename T>
__global__ void __launch_bounds__(512) frame_mean_kernel(const T* __restrict__ unc, int hw,
                                                         float* __restrict__ conf) {
  __shared__ float part[16];
  const T* src = unc + static_cast<long long>(blockIdx.x) * hw;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  int i = threadIdx.x;
  for (; i + 3 * 512 < hw; i += 4 * 512) {
    acc[0] += ld_as_float(src + i);
    acc[1] += ld_as_float(src + i + 512);
    acc[2] += ld_as_float(src + i + 1024);
    acc[3] += ld_as_float(src + i + 1536);
  }
  for (; i < hw; i += 512) acc[0] += ld_as_float(src + i);
  float s = (acc[0] + acc[1]) + (acc[2] + acc[3]);
  #pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = threadIdx.x < 16 ? part[threadIdx.x] : 0.f;
    #pragma unroll
    for (int o = 8; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if (threadIdx.x == 0) {
      const float m = __fmul_rn(t, 1.0f / static_cast<float>(hw));
      conf[blockIdx.x] = sizeof(T) == 2 ? bf16_round_f(m) : m;
    }
  }
}

// ------------------------------------------------------------------ score + pick
// One CTA; warp w handles rows (b,i) = w, w + nwarps, ...; lane j holds frames j and j + 32.
__global__ void __launch_bounds__(256) qam_pick_kernel(const float* __restrict__ sim, float* __restrict__ strive,
                                                       const float* __restrict__ conf, int b, int T, int top_k,
                                                       float* __restrict__ frame_score, int32_t* __restrict__ picked,
                                                       float* __restrict__ score_norm, uint8_t* __restrict__ mask) {
  extern __shared__ float sel[];                     // (b*T, top_k) picked scores
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  const int K = min(top_k, T);
  for (int row = warp; row < b * T; row += nw) {
    const int bb = row / T;
    const long long base = static_cast<long long>(row) * T;
    float S[2], sc[2];
    bool has[2];
    float ssum = 0.f;
    #pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int j = lane + 32 * u;
      has[u] = j < T;
      S[u] = has[u] ? strive[base + j] : 0.f;
      ssum += S[u];
    }
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) ssum += __shfl_xor_sync(0xffffffffu, ssum, o);
    #pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int j = lane + 32 * u;
      sc[u] = -INFINITY;
      if (has[u]) {
        const float pen = expf(-S[u] / (ssum + static_cast<float>(T)));        // :505
        sc[u] = __fadd_rn(__fmul_rn(pen, sim[base + j]), conf[bb * T + j]);    // :507
        frame_score[base + j] = sc[u];
      }
    }
    // rank by counting: frames that beat j (larger score, or equal score and lower index)
    int rank[2] = {0, 0};
    for (int jj = 0; jj < T; ++jj) {
      const float o = __shfl_sync(0xffffffffu, jj < 32 ? sc[0] : sc[1], jj & 31);
      #pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int j = lane + 32 * u;
        rank[u] += (o > sc[u]) || (o == sc[u] && jj < j);
      }
    }
    const bool pk0 = has[0] && rank[0] < K, pk1 = has[1] && rank[1] < K;
    const unsigned m0 = __ballot_sync(0xffffffffu, pk0), m1 = __ballot_sync(0xffffffffu, pk1);
    const unsigned lt = (1u << lane) - 1u;
    const int pos[2] = {__popc(m0 & lt), __popc(m0) + __popc(m1 & lt)};
    const bool pk[2] = {pk0, pk1};
    if (lane < top_k) { picked[static_cast<long long>(row) * top_k + lane] = -1; sel[row * top_k + lane] = 0.f; }
    __syncwarp();
    #pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int j = lane + 32 * u;
      if (has[u]) {
        if (mask) mask[base + j] = pk[u] ? 1 : 0;
        if (pk[u] && pos[u] < K) {
          picked[static_cast<long long>(row) * top_k + pos[u]] = j;
          sel[row * top_k + pos[u]] = sc[u];
          strive[base + j] = S[u] + 1.f;                                         // :513
        }
      }
    }
  }
  __syncthreads();
  // score_norm[b,i,n] = sel / mean over (b, n) of the picked scores of query frame i   (:532-533)
  for (int i = warp; i < T; i += nw) {
    float s = 0.f;
    for (int e = lane; e < b * K; e += 32) s += sel[((e / K) * T + i) * top_k + (e % K)];
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float mean = __fmul_rn(s, 1.0f / static_cast<float>(b * K));   // ATen CUDA mean: sum * (1/n)
    for (int e = lane; e < b * top_k; e += 32) {
      const int bb = e / top_k, n = e % top_k;
      const int idx = (bb * T + i) * top_k + n;
      score_norm[idx] = n < K ? sel[idx] / mean : 0.f;
    }
  }
}

// ------------------------------------------------------------------ packing
// Query operand: transposes a [128 channels][32 pixels] slab (pixels contiguous in the NCHW source) into
// [32 pixels][128 channels] bf16 rows, + pe (once per level; the per-iteration key operand has its own kernel below).
template <typename T>
__global__ void __launch_bounds__(256) pack_rows_kernel(const T* __restrict__ src, long long sb, long long sc,
                                                        long long st, const float* __restrict__ pe, int Tn, int hw,
                                                        __nv_bfloat16* __restrict__ dst) {
  __shared__ float tile[128][33];
  const int p0 = blockIdx.x * 32;
  const int item = blockIdx.y;                    // b*T + t
  const int bb = item / Tn, t = item % Tn;
  __nv_bfloat16* out = dst + (static_cast<long long>(item) * hw + p0) * 128;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const T* base = src + bb * sb + t * st + p0;
  #pragma unroll
  for (int k = 0; k < 16; ++k) {
    const int c = k * 8 + w;
    float v = 0.f;
    if (p0 + lane < hw) v = __fadd_rn(ld_as_float(base + c * sc + lane), __ldg(pe + t * 128 + c));
    tile[c][lane] = v;
  }
  __syncthreads();
  #pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int px = k * 4 + (threadIdx.x >> 6), cp = threadIdx.x & 63;
    if (p0 + px < hw)
      reinterpret_cast<uint32_t*>(out)[px * 64 + cp] = pack_bf16x2(tile[2 * cp][px], tile[2 * cp + 1][px]);
  }
}

// Modulated keys, organised by SOURCE frame: K'[(b*T + i)*Kp + n][p][c] = bf16(key[b, c, j, p] * s_in + pe[j][c]) with
// j = picked[b, i, n].  A CTA stages one [128 channels][64 pixels] slab of key frame j ONCE (128-byte row reads for bf16)
// and writes it for every (i, n) that picked j -- each key frame is read once instead of once per pick (x Kp on average),
// and every store is a 16-byte piece of a 256-byte row.  Arithmetic: separate fp32 multiply and add, one rounding (what
// torch's `key * s + pe` followed by `.to(bf16)` does).  Slots with picked < 0 (T < top_k) are zero-filled by the CTAs of frame 0.
template <typename T>
__global__ void __launch_bounds__(256) pack_key_kernel(const T* __restrict__ src, long long sb, long long sc, long long st,
                                                       const float* __restrict__ pe, const int32_t* __restrict__ picked,
                                                       const float* __restrict__ score_norm, int Tn, int Kp, int hw,
                                                       __nv_bfloat16* __restrict__ dst) {
  __shared__ float tile[128][65];
  __shared__ float s_pe[128];
  const int p0 = blockIdx.x * 64;
  const int bb = blockIdx.y / Tn, j = blockIdx.y % Tn;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int32_t* pk = picked + static_cast<long long>(bb) * Tn * Kp;
  bool any = false;
  for (int e = lane; e < Tn * Kp; e += 32) { const int v = pk[e]; any |= (v == j) || (j == 0 && v < 0); }
  if (!__any_sync(0xffffffffu, any)) return;       // nobody picked this frame (every warp sees the same answer)
  const T* base = src + bb * sb + j * st + p0;
  #pragma unroll 4
  for (int k = 0; k < 16; ++k) {
    const int c = k * 8 + w;
    tile[c][lane] = (p0 + lane < hw) ? ld_as_float(base + c * sc + lane) : 0.f;
    tile[c][lane + 32] = (p0 + lane + 32 < hw) ? ld_as_float(base + c * sc + lane + 32) : 0.f;
  }
  if (threadIdx.x < 128) s_pe[threadIdx.x] = __ldg(pe + j * 128 + threadIdx.x);
  __syncthreads();
  // a warp covers 8 pixels x 4 channel groups: shared-memory bank = (channel + pixel) % 32 is distinct for all 32 lanes,
  // and each pixel row gets a 64-byte (two full sectors) piece per store instruction
  const int cg = ((lane & 3) + 4 * (w & 3)) * 8, pr = (w >> 2) * 8 + (lane >> 2);
  for (int e = 0; e < Tn * Kp; ++e) {
    const int v = pk[e];
    const bool zero = v < 0 && j == 0;
    if (v != j && !zero) continue;
    const long long item = static_cast<long long>(bb) * Tn * Kp + e;
    const float s = zero ? 0.f : __ldg(score_norm + item);
    __nv_bfloat16* out = dst + (item * hw + p0) * 128;
    #pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int px = q * 16 + pr;
      if (p0 + px >= hw) continue;
      uint4 o;
      if (zero) o = make_uint4(0u, 0u, 0u, 0u);
      else {
        float f[8];
        #pragma unroll
        for (int u = 0; u < 8; ++u) f[u] = __fadd_rn(__fmul_rn(tile[cg + u][px], s), s_pe[cg + u]);
        o.x = pack_bf16x2(f[0], f[1]); o.y = pack_bf16x2(f[2], f[3]); o.z = pack_bf16x2(f[4], f[5]); o.w = pack_bf16x2(f[6], f[7]);
      }
      *reinterpret_cast<uint4*>(out + px * 128 + cg) = o;
    }
  }
}

// V[(b*T+t), c, p_pad] = bf16(value[b,c,t,p]); pad columns zeroed.
template <typename T>
__global__ void __launch_bounds__(256) pack_value_kernel(const T* __restrict__ src, long long sb, long long sc,
                                                         long long st, int Tn, int hw, int hw_pad,
                                                         __nv_bfloat16* __restrict__ dst) {
  const int c = blockIdx.y, f = blockIdx.z;
  const int bb = f / Tn, t = f % Tn;
  const T* s = src + bb * sb + t * st + c * sc;
  __nv_bfloat16* d = dst + (static_cast<long long>(f) * 128 + c) * hw_pad;
  const int p = (blockIdx.x * 256 + threadIdx.x) * 2;
  if (p < hw_pad) {
    const float a = p < hw ? ld_as_float(s + p) : 0.f;
    const float b2 = p + 1 < hw ? ld_as_float(s + p + 1) : 0.f;
    *reinterpret_cast<uint32_t*>(d + p) = pack_bf16x2(a, b2);
  }
}

}  // namespace ppm

using namespace ppm;

extern "C" int ppm_qam_pick(const float* sim, float* strive, const void* uncertainty, int unc_dtype, int b,
                            int T, int hw, int top_k, float* frame_score, int32_t* picked,
                            float* score_norm, uint8_t* mask, float* conf_ws, void* stream) {
  PPM_CHECK(sim && strive && uncertainty && frame_score && picked && score_norm && conf_ws, PPM_ERR_ARG,
            "ppm_qam_pick: null pointer");
  PPM_CHECK(b > 0 && T > 0 && hw > 0 && top_k > 0, PPM_ERR_ARG, "ppm_qam_pick: bad shape");
  PPM_CHECK(T <= 64, PPM_ERR_UNSUPPORTED, "ppm_qam_pick: T = %d > 64", T);
  PPM_CHECK(top_k <= 16, PPM_ERR_UNSUPPORTED, "ppm_qam_pick: top_k = %d > 16", top_k);
  PPM_CHECK(static_cast<size_t>(b) * T * top_k * 4 <= 48 * 1024, PPM_ERR_UNSUPPORTED,
            "ppm_qam_pick: b*T*top_k too large");
  PPM_CHECK(unc_dtype == PPM_F32 || unc_dtype == PPM_BF16, PPM_ERR_ARG, "ppm_qam_pick: bad dtype");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (unc_dtype == PPM_BF16)
    frame_mean_kernel<__nv_bfloat16><<<b * T, 512, 0, st>>>(static_cast<const __nv_bfloat16*>(uncertainty), hw, conf_ws);
  else
    frame_mean_kernel<float><<<b * T, 512, 0, st>>>(static_cast<const float*>(uncertainty), hw, conf_ws);
  PPM_CUDA(cudaGetLastError());
  qam_pick_kernel<<<1, 256, static_cast<size_t>(b) * T * top_k * 4, st>>>(sim, strive, conf_ws, b, T, top_k,
                                                                            frame_score, picked, score_norm, mask);
  PPM_CUDA(cudaGetLastError());
  return PPM_OK;
}

// ------------------------------------------------------------------ allocation-free variant of ppm_qam_pick
// No counterpart in the reference: added for callers that cannot provide (or
// allocate) the `conf_ws` scratch -- code under CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc.  The frame-confidence scratch lives in a file-scope device array of PPM_QAM_STATIC_MAX_FRAMES floats
// (16 KB; every GPU that loads the library holds it until exit).  In HOST code the array's name is a host-side
// stand-in, not a device address: its device address is taken once per device with cudaGetSymbolAddress (that device
// current) and THAT pointer is handed to the kernels.  Nothing in the array has to start at zero: frame_mean_kernel
// overwrites conf[0 .. b*T) before qam_pick_kernel reads exactly that range, so the previous call's contents (which
// the array keeps) are never observed.  Every call shares the one array per device: calls that overlap on the GPU
// (different streams, no ordering between them) race on it.
namespace ppm {

static __device__ __align__(256) float g_qam_conf_scratch[PPM_QAM_STATIC_MAX_FRAMES];

constexpr int kQamMaxDevices = 64;

static int qam_static_scratch(float** out) {
  static std::mutex mu;
  static float* addr[kQamMaxDevices] = {};
  int dev = -1;
  PPM_CUDA(cudaGetDevice(&dev));
  PPM_CHECK(dev >= 0 && dev < kQamMaxDevices, PPM_ERR_UNSUPPORTED, "ppm_qam_pick_static: device ordinal %d >= %d", dev,
            kQamMaxDevices);
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {                                     // set-up, once per device, with that device current
    void* p = nullptr;
    PPM_CUDA(cudaGetSymbolAddress(&p, g_qam_conf_scratch));
    PPM_CHECK(p != nullptr, PPM_ERR_CUDA, "ppm_qam_pick_static: cudaGetSymbolAddress returned a null address");
    addr[dev] = static_cast<float*>(p);
  }
  *out = addr[dev];
  return PPM_OK;
}

}  // namespace ppm

extern "C" int ppm_qam_pick_static(const float* sim, float* strive, const void* uncertainty, int unc_dtype, int b,
                                   int T, int hw, int top_k, float* frame_score, int32_t* picked,
                                   float* score_norm, uint8_t* mask, void* stream) {
  PPM_CHECK(sim && strive && uncertainty && frame_score && picked && score_norm, PPM_ERR_ARG,
            "ppm_qam_pick_static: null pointer");
  PPM_CHECK(b > 0 && T > 0 && hw > 0 && top_k > 0, PPM_ERR_ARG, "ppm_qam_pick_static: bad shape");
  PPM_CHECK(static_cast<long long>(b) * T <= PPM_QAM_STATIC_MAX_FRAMES, PPM_ERR_UNSUPPORTED,
            "ppm_qam_pick_static: b*T = %lld frames exceeds the static scratch of %d frames "
            "(PPM_QAM_STATIC_MAX_FRAMES); use ppm_qam_pick with a caller-provided conf_ws",
            static_cast<long long>(b) * T, PPM_QAM_STATIC_MAX_FRAMES);
  PPM_CHECK(T <= 64, PPM_ERR_UNSUPPORTED, "ppm_qam_pick_static: T = %d > 64", T);
  PPM_CHECK(top_k <= 16, PPM_ERR_UNSUPPORTED, "ppm_qam_pick_static: top_k = %d > 16", top_k);
  PPM_CHECK(static_cast<size_t>(b) * T * top_k * 4 <= 48 * 1024, PPM_ERR_UNSUPPORTED,
            "ppm_qam_pick_static: b*T*top_k too large");
  PPM_CHECK(unc_dtype == PPM_F32 || unc_dtype == PPM_BF16, PPM_ERR_ARG, "ppm_qam_pick_static: bad dtype");
  float* conf = nullptr;
  if (int rc = qam_static_scratch(&conf)) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (unc_dtype == PPM_BF16)
    frame_mean_kernel<__nv_bfloat16><<<b * T, 512, 0, st>>>(static_cast<const __nv_bfloat16*>(uncertainty), hw, conf);
  else
    frame_mean_kernel<float><<<b * T, 512, 0, st>>>(static_cast<const float*>(uncertainty), hw, conf);
  PPM_CUDA(cudaGetLastError());
  qam_pick_kernel<<<1, 256, static_cast<size_t>(b) * T * top_k * 4, st>>>(sim, strive, conf, b, T, top_k, frame_score,
                                                                            picked, score_norm, mask);
  PPM_CUDA(cudaGetLastError());
  return PPM_OK;
}

extern "C" int ppm_pack_bytes(int b, int T, int Kp, int hw, size_t* q_bytes, size_t* k_bytes, size_t* v_bytes) {
  PPM_CHECK(b > 0 && T > 0 && Kp > 0 && hw > 0, PPM_ERR_ARG, "ppm_pack_bytes: bad shape");
  const size_t hw_pad = (static_cast<size_t>(hw) + 7) / 8 * 8;
  if (q_bytes) *q_bytes = static_cast<size_t>(b) * T * hw * 128 * 2;
  if (k_bytes) *k_bytes = static_cast<size_t>(b) * T * Kp * hw * 128 * 2;
  if (v_bytes) *v_bytes = static_cast<size_t>(b) * T * 128 * hw_pad * 2;
  return PPM_OK;
}

static int check_pack(const char* who, const void* src, int dt, int b, int T, int hw, const void* out) {
  PPM_CHECK(src && out, PPM_ERR_ARG, "%s: null pointer", who);
  PPM_CHECK(b > 0 && T > 0 && hw > 0, PPM_ERR_ARG, "%s: bad shape", who);
  PPM_CHECK(dt == PPM_F32 || dt == PPM_BF16, PPM_ERR_ARG, "%s: bad dtype", who);
  PPM_CHECK((reinterpret_cast<uintptr_t>(out) & 15) == 0, PPM_ERR_ARG, "%s: output not 16-byte aligned", who);
  return PPM_OK;
}

extern "C" int ppm_pack_query(const void* query, int src_dtype, int64_t sb, int64_t sc, int64_t st,
                              const float* pe, int b, int T, int hw, void* q_out, void* stream) {
  if (int rc = check_pack("ppm_pack_query", query, src_dtype, b, T, hw, q_out)) return rc;
  PPM_CHECK(pe, PPM_ERR_ARG, "ppm_pack_query: null pe");
  PPM_CHECK(b * T <= 65535, PPM_ERR_UNSUPPORTED, "ppm_pack_query: b*T > 65535");
  dim3 grid((hw + 31) / 32, b * T);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  auto* d = static_cast<__nv_bfloat16*>(q_out);
  if (src_dtype == PPM_BF16)
    pack_rows_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>(static_cast<const __nv_bfloat16*>(query), sb, sc, st, pe, T, hw, d);
  else
    pack_rows_kernel<float><<<grid, 256, 0, s>>>(static_cast<const float*>(query), sb, sc, st, pe, T, hw, d);
  PPM_CUDA(cudaGetLastError());
  return PPM_OK;
}

extern "C" int ppm_pack_key(const void* key, int src_dtype, int64_t sb, int64_t sc, int64_t st, const float* pe,
                            const int32_t* picked, const float* score_norm, int b, int T, int Kp, int hw,
                            void* k_out, void* stream) {
  if (int rc = check_pack("ppm_pack_key", key, src_dtype, b, T, hw, k_out)) return rc;
  PPM_CHECK(pe && picked && score_norm && Kp > 0, PPM_ERR_ARG, "ppm_pack_key: null pointer or bad Kp");
  PPM_CHECK(b * T <= 65535, PPM_ERR_UNSUPPORTED, "ppm_pack_key: b*T > 65535");
  PPM_CHECK((reinterpret_cast<uintptr_t>(k_out) & 15) == 0, PPM_ERR_ARG, "ppm_pack_key: k_out must be 16-byte aligned");
  dim3 grid((hw + 63) / 64, b * T);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  auto* d = static_cast<__nv_bfloat16*>(k_out);
  if (src_dtype == PPM_BF16)
    pack_key_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>(static_cast<const __nv_bfloat16*>(key), sb, sc, st, pe, picked,
                                                        score_norm, T, Kp, hw, d);
  else
    pack_key_kernel<float><<<grid, 256, 0, s>>>(static_cast<const float*>(key), sb, sc, st, pe, picked, score_norm, T, Kp,
                                                hw, d);
  PPM_CUDA(cudaGetLastError());
  return PPM_OK;
}

extern "C" int ppm_pack_value(const void* value, int src_dtype, int64_t sb, int64_t sc, int64_t st, int b, int T,
                              int hw, void* v_out, void* stream) {
  if (int rc = check_pack("ppm_pack_value", value, src_dtype, b, T, hw, v_out)) return rc;
  PPM_CHECK(b * T <= 65535, PPM_ERR_UNSUPPORTED, "ppm_pack_value: b*T > 65535");
  const int hw_pad = (hw + 7) / 8 * 8;
  dim3 grid((hw_pad / 2 + 255) / 256, 128, b * T);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  auto* d = static_cast<__nv_bfloat16*>(v_out);
  if (src_dtype == PPM_BF16)
    pack_value_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>(static_cast<const __nv_bfloat16*>(value), sb, sc, st, T, hw,
                                                          hw_pad, d);
  else
    pack_value_kernel<float><<<grid, 256, 0, s>>>(static_cast<const float*>(value), sb, sc, st, T, hw, hw_pad, d);
  PPM_CUDA(cudaGetLastError());
  return PPM_OK;
}
