// K3c: query-adaptive softmax aggregation over the picked memory frames (the PPM "read").
// Replaces the flash_attn_func call and residual at reference models/core/ppmstereo.py:550-555.
//
//   hidden[p, :] = softmax_{p'}(scale * Q'[p] . K'[p']) V[p'],  p' over all pixels of the picked frames
//   out[c, p]    = motion_features[c, p] + beta * bf16(hidden[p, c])
//
// FlashAttention-style streaming softmax on tcgen05 tensor cores, one 128-query tile per CTA:
//   warp 0     TMA producer: Q tile once, then K'/V tiles of the picked frames through two
//              2-stage mbarrier rings; the picked-frame list drives the TMA coordinates, so the
//              reference's masked_select gather copies never exist.
//   warp 1     single-thread tcgen05.mma issue.  S = Q K'^T (SS, both K-major SWIZZLE_128B)
//              into a double-buffered TMEM accumulator; O += P V with P read straight from TMEM
//              (TS form) or from shared memory (SS form, kPInTmem = false) and V^T taken as it
//              lies in NCHW (pixels contiguous = K-major B operand, no transpose).
//   warps 2-5  softmax: one query row per thread (TMEM lane == row, so row max / row sum need no
//              shuffles), exp2 with the scale folded in, P rounded to bf16 (FlashAttention-2
//              rounding) and written back over S; lazy rescale of O only when the running max
//              grows by more than 2^8; final normalise, bf16 round, residual add, NCHW store.
// TMEM columns: S0 [0,128) S1 [128,256) O [256,384); P aliases the first 64 columns of its S.
#include <math.h>

#include "../../include/ppm_b200.h"
#include "common.cuh"

namespace ppm {

constexpr int kAtThreads = 192;
constexpr int kAtTile = 128;                      // queries per CTA, keys per step, head dim
constexpr int kAtTileBytes = kAtTile * kAtTile * 2;   // 32 KB: two [128 rows][128 B] halves
constexpr int kAtHalfBytes = kAtTileBytes / 2;
constexpr int kAtKStages = 2, kAtVStages = 2;
constexpr uint32_t kTmemS = 0, kTmemO = 256, kTmemCols = 512;
constexpr float kRescaleThreshold = 8.0f;         // log2 units (default; see AttendParams::rescale_thr)

struct AttendParams {
  const int32_t* picked;     // (b*T, Kp)
  const void* mf;            // motion features, indexed b*mb + c*mc + i*mt + p
  long long mb, mc, mt;
  const float* beta;
  float* out;                // (b*T, 128, hw)
  float scale_log2;          // softmax_scale * log2(e)
  float rescale_thr;         // lazy-rescale threshold, log2 units
  int T, Kp, hw;
};

__device__ __forceinline__ float at_ld(const float* p) { return __ldg(p); }
__device__ __forceinline__ float at_ld(const __nv_bfloat16* p) { return __bfloat162float(__ldg(p)); }

template <bool kPInTmem, typename MfT>
__global__ void __launch_bounds__(kAtThreads, 1)
ppm_attend_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_k,
                  const __grid_constant__ CUtensorMap map_v, const AttendParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + kAtTileBytes;
  uint8_t* sV = sK + kAtKStages * kAtTileBytes;
  uint8_t* sP = sV + kAtVStages * kAtTileBytes;   // only used when !kPInTmem
  __shared__ uint64_t bar_q, bar_kf[kAtKStages], bar_ke[kAtKStages], bar_vf[kAtVStages], bar_ve[kAtVStages];
  __shared__ uint64_t bar_sf[2], bar_pf[2], bar_o[2];
  __shared__ uint32_t tmem_slot;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int frame = blockIdx.y;                   // b*T + i
  const int bb = frame / p.T;
  const int q0 = blockIdx.x * kAtTile;
  const int tiles_per_frame = (p.hw + kAtTile - 1) / kAtTile;
  int n_frames = 0;
  for (int n = 0; n < p.Kp; ++n) n_frames += (__ldg(p.picked + frame * p.Kp + n) >= 0);
  const int n_tiles = n_frames * tiles_per_frame;

  if (threadIdx.x == 0) {
    mbar_init(&bar_q, 1);
    for (int s = 0; s < kAtKStages; ++s) { mbar_init(&bar_kf[s], 1); mbar_init(&bar_ke[s], 1); }
    for (int s = 0; s < kAtVStages; ++s) { mbar_init(&bar_vf[s], 1); mbar_init(&bar_ve[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&bar_sf[s], 1); mbar_init(&bar_pf[s], 4); }
    mbar_init(&bar_o[0], 1); mbar_init(&bar_o[1], 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<kTmemCols>(&tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;

  if (warp == 0) {
    // ---------------------------------------------------------------- TMA producer
    if (elect_one() && n_tiles > 0) {
      tma_prefetch_desc(&map_q); tma_prefetch_desc(&map_k); tma_prefetch_desc(&map_v);
      mbar_expect_tx(&bar_q, kAtTileBytes);
      tma_load_3d(sQ, &map_q, &bar_q, 0, q0, frame);
      tma_load_3d(sQ + kAtHalfBytes, &map_q, &bar_q, 64, q0, frame);
      int j = 0;
      for (int n = 0; n < p.Kp; ++n) {
        const int src = __ldg(p.picked + frame * p.Kp + n);
        if (src < 0) continue;
        for (int t = 0; t < tiles_per_frame; ++t, ++j) {
          const int ks = j % kAtKStages, vs = j % kAtVStages;
          mbar_wait(&bar_ke[ks], ((j / kAtKStages) & 1) ^ 1);
          mbar_expect_tx(&bar_kf[ks], kAtTileBytes);
          uint8_t* dk = sK + ks * kAtTileBytes;
          tma_load_3d(dk, &map_k, &bar_kf[ks], 0, t * kAtTile, frame * p.Kp + n);
          tma_load_3d(dk + kAtHalfBytes, &map_k, &bar_kf[ks], 64, t * kAtTile, frame * p.Kp + n);
          mbar_wait(&bar_ve[vs], ((j / kAtVStages) & 1) ^ 1);
          mbar_expect_tx(&bar_vf[vs], kAtTileBytes);
          uint8_t* dv = sV + vs * kAtTileBytes;
          tma_load_3d(dv, &map_v, &bar_vf[vs], t * kAtTile, 0, bb * p.T + src);
          tma_load_3d(dv + kAtHalfBytes, &map_v, &bar_vf[vs], t * kAtTile + 64, 0, bb * p.T + src);
        }
      }
    }
  } else if (warp == 1) {
    // ---------------------------------------------------------------- MMA issuer
    if (elect_one() && n_tiles > 0) {
      const uint32_t idesc = umma_idesc(1, 0, 0, kAtTile, kAtTile);     // bf16, K-major A and B
      const uint32_t q_addr = smem_u32(sQ);
      auto issue_qk = [&](int j) {
        const int ks = j % kAtKStages;
        mbar_wait(&bar_kf[ks], (j / kAtKStages) & 1);
        tc_fence_after();
        const uint32_t k_addr = smem_u32(sK + ks * kAtTileBytes);
        #pragma unroll
        for (int k = 0; k < 8; ++k) {
          const uint32_t off = (k >> 2) * kAtHalfBytes + (k & 3) * 32;
          umma_f16_ss(tmem + kTmemS + (j & 1) * 128, umma_desc_sw128(q_addr + off, 16, 1024),
                      umma_desc_sw128(k_addr + off, 16, 1024), idesc, k != 0);
        }
        umma_commit(&bar_sf[j & 1]);
        umma_commit(&bar_ke[ks]);
      };
      mbar_wait(&bar_q, 0);
      issue_qk(0);
      for (int j = 0; j < n_tiles; ++j) {
        if (j + 1 < n_tiles) issue_qk(j + 1);
        const int vs = j % kAtVStages;
        mbar_wait(&bar_pf[j & 1], (j >> 1) & 1);
        mbar_wait(&bar_vf[vs], (j / kAtVStages) & 1);
        tc_fence_after();
        const uint32_t v_addr = smem_u32(sV + vs * kAtTileBytes);
        #pragma unroll
        for (int k = 0; k < 8; ++k) {
          const uint32_t off = (k >> 2) * kAtHalfBytes + (k & 3) * 32;
          const uint64_t vd = umma_desc_sw128(v_addr + off, 16, 1024);
          if constexpr (kPInTmem)
            umma_f16_ts(tmem + kTmemO, tmem + kTmemS + (j & 1) * 128 + k * 8, vd, idesc, (j | k) != 0);
          else
            umma_f16_ss(tmem + kTmemO, umma_desc_sw128(smem_u32(sP) + off, 16, 1024), vd, idesc, (j | k) != 0);
        }
        umma_commit(&bar_ve[vs]);
        umma_commit(&bar_o[j & 1]);
      }
    }
  } else {
    // ---------------------------------------------------------------- softmax / correction / epilogue
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;                 // query row in the tile == TMEM lane
    const uint32_t lane_base = static_cast<uint32_t>(quarter * 32) << 16;
    const float c1 = p.scale_log2;
    const float thr = p.rescale_thr / c1;            // in raw-score units
    float m_used = -INFINITY, l_sum = 0.f;
    for (int j = 0; j < n_tiles; ++j) {
      const int t = j % tiles_per_frame;
      const int valid = min(kAtTile, p.hw - t * kAtTile);     // keys of this tile inside the frame
      mbar_wait(&bar_sf[j & 1], (j >> 1) & 1);
      tc_fence_after();
      const uint32_t s_addr = tmem + lane_base + kTmemS + (j & 1) * 128;
      uint32_t s[4][32];
      #pragma unroll
      for (int c = 0; c < 4; ++c) tmem_ld_x32(s_addr + c * 32, s[c]);
      tmem_ld_wait();
      float mx = -INFINITY;
      #pragma unroll
      for (int c = 0; c < 4; ++c)
        #pragma unroll
        for (int e = 0; e < 32; ++e)
          if (c * 32 + e < valid) mx = fmaxf(mx, __uint_as_float(s[c][e]));
      const bool grow = mx > m_used + thr;               // also true on the first tile (-inf)
      // "P.V of tile j-1 has landed in O" is waited for exactly once per tile by every warp (early
      // when O must be rescaled, late otherwise), so no phase of bar_o is ever skipped.
      bool o_waited = (j == 0);
      if (__any_sync(0xffffffffu, grow)) {
        float alpha = 1.f;
        if (grow) {
          alpha = (m_used == -INFINITY) ? 0.f : ex2_approx((m_used - mx) * c1);
          m_used = mx;
          l_sum *= alpha;
        }
        if (j > 0) {
          mbar_wait(&bar_o[(j - 1) & 1], ((j - 1) >> 1) & 1);
          o_waited = true;
          tc_fence_after();
          #pragma unroll
          for (int c = 0; c < 4; ++c) {
            uint32_t o[32];
            tmem_ld_x32(tmem + lane_base + kTmemO + c * 32, o);
            tmem_ld_wait();
            #pragma unroll
            for (int e = 0; e < 32; ++e) o[e] = __float_as_uint(__uint_as_float(o[e]) * alpha);
            tmem_st_x32(tmem + lane_base + kTmemO + c * 32, o);
          }
        }
      }
      const float mneg = -m_used * c1;
      uint32_t pk[64];
      #pragma unroll
      for (int c = 0; c < 4; ++c)
        #pragma unroll
        for (int e = 0; e < 32; e += 2) {
          float p0 = ex2_approx(fmaf(__uint_as_float(s[c][e]), c1, mneg));
          float p1 = ex2_approx(fmaf(__uint_as_float(s[c][e + 1]), c1, mneg));
          if (c * 32 + e >= valid) p0 = 0.f;
          if (c * 32 + e + 1 >= valid) p1 = 0.f;
          l_sum += p0 + p1;
          pk[(c * 32 + e) >> 1] = pack_bf16x2(p0, p1);
        }
      if constexpr (kPInTmem) {
        tmem_st_x32(s_addr, *reinterpret_cast<const uint32_t(*)[32]>(&pk[0]));
        tmem_st_x32(s_addr + 32, *reinterpret_cast<const uint32_t(*)[32]>(&pk[32]));
        tmem_st_wait();
      } else {
        if (!o_waited) { mbar_wait(&bar_o[(j - 1) & 1], ((j - 1) >> 1) & 1); o_waited = true; }   // sP is free again
        #pragma unroll
        for (int ch = 0; ch < 16; ++ch) {                 // 16-byte chunks of this row, SWIZZLE_128B
          uint8_t* dst = sP + (ch >> 3) * kAtHalfBytes + row * 128 + (((ch & 7) ^ (row & 7)) << 4);
          *reinterpret_cast<uint4*>(dst) = make_uint4(pk[4 * ch], pk[4 * ch + 1], pk[4 * ch + 2], pk[4 * ch + 3]);
        }
        fence_proxy_async_smem();
        tmem_st_wait();
      }
      if (!o_waited) mbar_wait(&bar_o[(j - 1) & 1], ((j - 1) >> 1) & 1);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar_pf[j & 1]);
    }
    // ---- epilogue: out = mf + beta * bf16(O / l)
    if (n_tiles > 0) {
      mbar_wait(&bar_o[(n_tiles - 1) & 1], ((n_tiles - 1) >> 1) & 1);
      tc_fence_after();
    }
    const float inv = n_tiles > 0 ? 1.f / l_sum : 0.f;
    const float beta = __ldg(p.beta);
    const int pix = q0 + row;
    const MfT* mf = static_cast<const MfT*>(p.mf) + bb * p.mb + (frame - bb * p.T) * p.mt + pix;
    float* dst = p.out + static_cast<long long>(frame) * kAtTile * p.hw + pix;
    #pragma unroll
    for (int c = 0; c < 4; ++c) {
      uint32_t o[32];
      if (n_tiles > 0) {
        tmem_ld_x32(tmem + lane_base + kTmemO + c * 32, o);
        tmem_ld_wait();
      } else {
        #pragma unroll
        for (int e = 0; e < 32; ++e) o[e] = 0u;
      }
      if (pix < p.hw) {
        #pragma unroll
        for (int e = 0; e < 32; ++e) {
          const int ch = c * 32 + e;
          const float hid = bf16_round_f(__uint_as_float(o[e]) * inv);
          dst[static_cast<long long>(ch) * p.hw] = __fadd_rn(at_ld(mf + ch * p.mc), __fmul_rn(beta, hid));
        }
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<kTmemCols>(tmem);
  }
}

template <bool kPInTmem, typename MfT>
static int launch_attend(const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv,
                         const AttendParams& p, int n_qtiles, int n_frames, cudaStream_t st) {
  constexpr int kSmem = (1 + kAtKStages + kAtVStages + (kPInTmem ? 0 : 1)) * kAtTileBytes + 1024;
  auto kern = ppm_attend_kernel<kPInTmem, MfT>;
  PPM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
  kern<<<dim3(n_qtiles, n_frames), kAtThreads, kSmem, st>>>(mq, mk, mv, p);
  PPM_CUDA(cudaGetLastError());
  return PPM_OK;
}

int attend2_dispatch(const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv, const int32_t* picked,
                     const void* mf, int mf_dtype, long long mb, long long mc, long long mt, const float* beta,
                     float scale_log2, float rescale_thr, int b, int T, int Kp, int hw, float* out,
                     cudaStream_t st);

int attend3_dispatch(const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv, const int32_t* picked,
                     const void* mf, int mf_dtype, long long mb, long long mc, long long mt, const float* beta,
                     float scale_log2, float rescale_thr, int emu, int b, int T, int Kp, int hw, int q_begin,
                     int q_count, float* out, cudaStream_t st);

// 3 = two query tiles, chunked P hand-off (ppm_attend3.cu); 2 = two query tiles (ppm_attend2.cu);
// 1 = one tile (this file).  2 and 1 are kept for cross-checks.
static int g_attend_version = 3;
static int g_exp_emulation = 0;    // v3: 4 = every other exponential by polynomial on the FMA pipe (measured slower)
static int g_p_in_tmem = 1;
static float g_rescale_thr = kRescaleThreshold;

}  // namespace ppm

// Test hook: choose how P reaches the P.V MMA (1 = TMEM operand, 0 = shared memory operand).
extern "C" int ppm_debug_set_p_in_tmem(int v) { ppm::g_p_in_tmem = v ? 1 : 0; return PPM_OK; }
// Test hook: kernel generation (3 = default; 2 and 1 = earlier kernels kept for cross-checks).
extern "C" int ppm_debug_set_attend_version(int v) { ppm::g_attend_version = (v >= 1 && v <= 3) ? v : 3; return PPM_OK; }
// Test hook: how many of every 8 exponentials the v3 kernel evaluates by polynomial (0 or 4).
extern "C" int ppm_debug_set_exp_emulation(int n) { ppm::g_exp_emulation = n; return PPM_OK; }
// Test hook: lazy-rescale threshold in log2 units (0 = rescale whenever the running max grows).
extern "C" int ppm_debug_set_rescale_threshold(float v) { ppm::g_rescale_thr = v; return PPM_OK; }

extern "C" int ppm_memory_attend(const void* Q, const void* K, const void* V, const int32_t* picked, const void* mf,
                                 int mf_dtype, int64_t mb, int64_t mc, int64_t mt, const float* beta, float scale,
                                 int b, int T, int Kp, int hw, float* out, void* stream) {
  return ppm_memory_attend_band(Q, K, V, picked, mf, mf_dtype, mb, mc, mt, beta, scale, b, T, Kp, hw, 0, hw, out, stream);
}

extern "C" int ppm_memory_attend_band(const void* Q, const void* K, const void* V, const int32_t* picked,
                                      const void* mf, int mf_dtype, int64_t mb, int64_t mc, int64_t mt,
                                      const float* beta, float scale, int b, int T, int Kp, int hw, int q_begin,
                                      int q_count, float* out, void* stream) {
  using namespace ppm;
  PPM_CHECK(q_begin >= 0 && q_count >= 0 && q_begin + q_count <= hw, PPM_ERR_ARG,
            "ppm_memory_attend_band: query band [%d, %d) outside [0, %d)", q_begin, q_begin + q_count, hw);
  if (q_count == 0) return PPM_OK;
  PPM_CHECK(Q && K && V && picked && mf && beta && out, PPM_ERR_ARG, "ppm_memory_attend: null pointer");
  PPM_CHECK(b > 0 && T > 0 && Kp > 0 && hw > 0, PPM_ERR_ARG, "ppm_memory_attend: bad shape");
  PPM_CHECK(mf_dtype == PPM_F32 || mf_dtype == PPM_BF16, PPM_ERR_ARG, "ppm_memory_attend: bad dtype");
  PPM_CHECK(b * T <= 65535, PPM_ERR_UNSUPPORTED, "ppm_memory_attend: b*T > 65535");
  PPM_CHECK(((reinterpret_cast<uintptr_t>(Q) | reinterpret_cast<uintptr_t>(K) | reinterpret_cast<uintptr_t>(V)) & 15) == 0,
            PPM_ERR_ARG, "ppm_memory_attend: packed operands must be 16-byte aligned");
  const int hw_pad = (hw + 7) / 8 * 8;
  CUtensorMap mq, mk, mv;
  const uint32_t box[3] = {64, 128, 1};
  {
    const uint64_t dims[3] = {128, (uint64_t)hw, (uint64_t)b * T};
    const uint64_t str[2] = {256, (uint64_t)hw * 256};
    if (int rc = make_tensor_map(&mq, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, Q, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
  }
  {
    const uint64_t dims[3] = {128, (uint64_t)hw, (uint64_t)b * T * Kp};
    const uint64_t str[2] = {256, (uint64_t)hw * 256};
    if (int rc = make_tensor_map(&mk, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, K, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
  }
  {
    const uint64_t dims[3] = {(uint64_t)hw_pad, 128, (uint64_t)b * T};
    const uint64_t str[2] = {(uint64_t)hw_pad * 2, (uint64_t)hw_pad * 256};
    if (int rc = make_tensor_map(&mv, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, V, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
  }
  AttendParams p{};
  p.picked = picked; p.mf = mf; p.mb = mb; p.mc = mc; p.mt = mt; p.beta = beta; p.out = out;
  p.scale_log2 = scale * 1.4426950408889634f;
  p.rescale_thr = g_rescale_thr;
  p.T = T; p.Kp = Kp; p.hw = hw;
  const int n_qtiles = (hw + kAtTile - 1) / kAtTile;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (g_attend_version == 3 && g_p_in_tmem)
    return attend3_dispatch(mq, mk, mv, picked, mf, mf_dtype, mb, mc, mt, beta, p.scale_log2, p.rescale_thr,
                            g_exp_emulation, b, T, Kp, hw, q_begin, q_count, out, st);
  PPM_CHECK(q_begin == 0 && q_count == hw, PPM_ERR_UNSUPPORTED,
            "ppm_memory_attend_band: only the default (v3) kernel supports a partial query band");
  if (g_attend_version >= 2 && g_p_in_tmem)
    return attend2_dispatch(mq, mk, mv, picked, mf, mf_dtype, mb, mc, mt, beta, p.scale_log2, p.rescale_thr, b, T,
                            Kp, hw, out, st);
  if (g_p_in_tmem) {
    if (mf_dtype == PPM_BF16) return launch_attend<true, __nv_bfloat16>(mq, mk, mv, p, n_qtiles, b * T, st);
    return launch_attend<true, float>(mq, mk, mv, p, n_qtiles, b * T, st);
  }
  if (mf_dtype == PPM_BF16) return launch_attend<false, __nv_bfloat16>(mq, mk, mv, p, n_qtiles, b * T, st);
  return launch_attend<false, float>(mq, mk, mv, p, n_qtiles, b * T, st);
}
