// K3c (v2): PPM memory read with TWO 128-query tiles per CTA (256 queries), the layout that keeps
// the tensor pipe fed: every K'/V tile fetched from L2 is used by both query tiles (half the
// L2->SM traffic per FLOP of the one-tile kernel), and while the softmax warpgroup of one query
// tile works, the tensor core runs the MMAs of the other.
// Replaces the flash_attn_func call and residual at reference models/core/ppmstereo.py:550-555.
//
//   warps 0-3   control group (registers released with setmaxnreg): warp 0 = TMA producer,
//               warp 1 = tcgen05.mma issuer, warps 2-3 idle
//   warps 4-7   softmax / correction / epilogue for query tile A (TMEM lane = query row)
//   warps 8-11  same for query tile B
// Tensor-pipe issue order per key tile j:  PV_A(j)  QK_A(j+1)  PV_B(j)  QK_B(j+1)
// so the softmax of tile A for step j+1 has the whole PV_B + QK_B window to finish, and vice versa.
// TMEM columns: S_A 0-127, S_B 128-255, O_A 256-383, O_B 384-511; P (bf16) overwrites the first
// 64 columns of its S.
#include <math.h>

#include "../../include/ppm_b200.h"
#include "common.cuh"

namespace ppm {

constexpr int kA2Threads = 384;
constexpr int kA2Tile = 128;
constexpr int kA2TileBytes = kA2Tile * kA2Tile * 2;       // 32 KB
constexpr int kA2HalfBytes = kA2TileBytes / 2;
constexpr int kA2Stages = 2;
constexpr uint32_t kA2TmemCols = 512;

struct Attend2Params {
  const int32_t* picked;
  const void* mf;
  long long mb, mc, mt;
  const float* beta;
  float* out;
  float scale_log2, rescale_thr;
  int T, Kp, hw;
};

__device__ __forceinline__ float a2_ld(const float* p) { return __ldg(p); }
__device__ __forceinline__ float a2_ld(const __nv_bfloat16* p) { return __bfloat162float(__ldg(p)); }
__device__ __forceinline__ float max3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}

template <typename MfT>
__global__ void __launch_bounds__(kA2Threads, 1)
ppm_attend2_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_k,
                   const __grid_constant__ CUtensorMap map_v, const Attend2Params p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;                                     // 2 tiles
  uint8_t* sK = sQ + 2 * kA2TileBytes;
  uint8_t* sV = sK + kA2Stages * kA2TileBytes;
  __shared__ uint64_t bar_q, bar_kf[kA2Stages], bar_ke[kA2Stages], bar_vf[kA2Stages], bar_ve[kA2Stages];
  __shared__ uint64_t bar_sf[2], bar_pf[2], bar_o[2][2];
  __shared__ uint32_t tmem_slot;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int frame = blockIdx.y;
  const int bb = frame / p.T;
  const int q0 = blockIdx.x * 2 * kA2Tile;
  const int tiles_per_frame = (p.hw + kA2Tile - 1) / kA2Tile;
  int n_frames = 0;
  for (int n = 0; n < p.Kp; ++n) n_frames += (__ldg(p.picked + frame * p.Kp + n) >= 0);
  const int n_tiles = n_frames * tiles_per_frame;

  if (threadIdx.x == 0) {
    mbar_init(&bar_q, 1);
    for (int s = 0; s < kA2Stages; ++s) {
      mbar_init(&bar_kf[s], 1); mbar_init(&bar_ke[s], 1); mbar_init(&bar_vf[s], 1); mbar_init(&bar_ve[s], 1);
    }
    for (int x = 0; x < 2; ++x) {
      mbar_init(&bar_sf[x], 1); mbar_init(&bar_pf[x], 4);
      mbar_init(&bar_o[x][0], 1); mbar_init(&bar_o[x][1], 1);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<kA2TmemCols>(&tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;

  if (warp < 4) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 88;");
    if (warp == 0) {
      // -------------------------------------------------------------- TMA producer
      if (elect_one() && n_tiles > 0) {
        tma_prefetch_desc(&map_q); tma_prefetch_desc(&map_k); tma_prefetch_desc(&map_v);
        mbar_expect_tx(&bar_q, 2 * kA2TileBytes);
        #pragma unroll
        for (int x = 0; x < 2; ++x) {
          tma_load_3d(sQ + x * kA2TileBytes, &map_q, &bar_q, 0, q0 + x * kA2Tile, frame);
          tma_load_3d(sQ + x * kA2TileBytes + kA2HalfBytes, &map_q, &bar_q, 64, q0 + x * kA2Tile, frame);
        }
        int j = 0;
        for (int n = 0; n < p.Kp; ++n) {
          const int src = __ldg(p.picked + frame * p.Kp + n);
          if (src < 0) continue;
          for (int t = 0; t < tiles_per_frame; ++t, ++j) {
            const int st = j % kA2Stages;
            const uint32_t ph = ((j / kA2Stages) & 1) ^ 1;
            mbar_wait(&bar_ke[st], ph);
            mbar_expect_tx(&bar_kf[st], kA2TileBytes);
            uint8_t* dk = sK + st * kA2TileBytes;
            tma_load_3d(dk, &map_k, &bar_kf[st], 0, t * kA2Tile, frame * p.Kp + n);
            tma_load_3d(dk + kA2HalfBytes, &map_k, &bar_kf[st], 64, t * kA2Tile, frame * p.Kp + n);
            mbar_wait(&bar_ve[st], ph);
            mbar_expect_tx(&bar_vf[st], kA2TileBytes);
            uint8_t* dv = sV + st * kA2TileBytes;
            tma_load_3d(dv, &map_v, &bar_vf[st], t * kA2Tile, 0, bb * p.T + src);
            tma_load_3d(dv + kA2HalfBytes, &map_v, &bar_vf[st], t * kA2Tile + 64, 0, bb * p.T + src);
          }
        }
      }
    } else if (warp == 1) {
      // -------------------------------------------------------------- MMA issuer
      if (elect_one() && n_tiles > 0) {
        const uint32_t idesc = umma_idesc(1, 0, 0, kA2Tile, kA2Tile);
        auto issue_qk = [&](int x, int j) {                 // S_x = Q_x K'_j^T
          const uint32_t q_addr = smem_u32(sQ + x * kA2TileBytes);
          const uint32_t k_addr = smem_u32(sK + (j % kA2Stages) * kA2TileBytes);
          #pragma unroll
          for (int k = 0; k < 8; ++k) {
            const uint32_t off = (k >> 2) * kA2HalfBytes + (k & 3) * 32;
            umma_f16_ss(tmem + x * 128, umma_desc_sw128(q_addr + off, 16, 1024),
                        umma_desc_sw128(k_addr + off, 16, 1024), idesc, k != 0);
          }
          umma_commit(&bar_sf[x]);
        };
        auto issue_pv = [&](int x, int j) {                 // O_x += P_x V_j
          const uint32_t v_addr = smem_u32(sV + (j % kA2Stages) * kA2TileBytes);
          #pragma unroll
          for (int k = 0; k < 8; ++k) {
            const uint32_t off = (k >> 2) * kA2HalfBytes + (k & 3) * 32;
            umma_f16_ts(tmem + 256 + x * 128, tmem + x * 128 + k * 8, umma_desc_sw128(v_addr + off, 16, 1024),
                        idesc, (j | k) != 0);
          }
          umma_commit(&bar_o[x][j & 1]);
        };
        mbar_wait(&bar_q, 0);
        mbar_wait(&bar_kf[0], 0);
        tc_fence_after();
        issue_qk(0, 0);
        issue_qk(1, 0);
        umma_commit(&bar_ke[0]);
        for (int j = 0; j < n_tiles; ++j) {
          const int st = j % kA2Stages, sn = (j + 1) % kA2Stages;
          const bool more = j + 1 < n_tiles;
          mbar_wait(&bar_pf[0], j & 1);
          mbar_wait(&bar_vf[st], (j / kA2Stages) & 1);
          tc_fence_after();
          issue_pv(0, j);
          if (more) {
            mbar_wait(&bar_kf[sn], ((j + 1) / kA2Stages) & 1);
            tc_fence_after();
            issue_qk(0, j + 1);
          }
          mbar_wait(&bar_pf[1], j & 1);
          tc_fence_after();
          issue_pv(1, j);
          umma_commit(&bar_ve[st]);
          if (more) {
            issue_qk(1, j + 1);
            umma_commit(&bar_ke[sn]);
          }
        }
      }
    }
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 208;");
    // ---------------------------------------------------------------- softmax / correction / epilogue
    const int x = (warp - 4) >> 2;                          // query tile of this warpgroup
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const uint32_t lane_base = static_cast<uint32_t>(quarter * 32) << 16;
    const uint32_t s_addr = tmem + lane_base + x * 128;
    const uint32_t o_addr = tmem + lane_base + 256 + x * 128;
    const float c1 = p.scale_log2;
    const float thr = p.rescale_thr / c1;
    float m_used = -INFINITY, l0 = 0.f, l1 = 0.f, l2 = 0.f, l3 = 0.f;
    for (int j = 0; j < n_tiles; ++j) {
      const int t = j % tiles_per_frame;
      const int valid = min(kA2Tile, p.hw - t * kA2Tile);
      mbar_wait(&bar_sf[x], j & 1);
      tc_fence_after();
      uint32_t s[4][32];
      #pragma unroll
      for (int c = 0; c < 4; ++c) tmem_ld_x32(s_addr + c * 32, s[c]);
      tmem_ld_wait();
      if (valid < kA2Tile) {                                 // tail tile of a frame: mask to -inf
        #pragma unroll
        for (int c = 0; c < 4; ++c)
          #pragma unroll
          for (int e = 0; e < 32; ++e)
            if (c * 32 + e >= valid) s[c][e] = 0xff800000u;
      }
      float mx[4];
      #pragma unroll
      for (int c = 0; c < 4; ++c) {
        mx[c] = __uint_as_float(s[c][0]);
        #pragma unroll
        for (int e = 1; e < 31; e += 2)
          mx[c] = max3(mx[c], __uint_as_float(s[c][e]), __uint_as_float(s[c][e + 1]));
        mx[c] = fmaxf(mx[c], __uint_as_float(s[c][31]));
      }
      const float m_tile = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
      const bool grow = m_tile > m_used + thr;
      bool o_waited = (j == 0);
      if (__any_sync(0xffffffffu, grow)) {
        float alpha = 1.f;
        if (grow) {
          alpha = (m_used == -INFINITY) ? 0.f : ex2_approx((m_used - m_tile) * c1);
          m_used = m_tile;
          l0 *= alpha; l1 *= alpha; l2 *= alpha; l3 *= alpha;
        }
        if (j > 0) {
          mbar_wait(&bar_o[x][(j - 1) & 1], ((j - 1) >> 1) & 1);
          o_waited = true;
          tc_fence_after();
          #pragma unroll
          for (int c = 0; c < 4; ++c) {
            uint32_t o[32];
            tmem_ld_x32(o_addr + c * 32, o);
            tmem_ld_wait();
            #pragma unroll
            for (int e = 0; e < 32; ++e) o[e] = __float_as_uint(__uint_as_float(o[e]) * alpha);
            tmem_st_x32(o_addr + c * 32, o);
          }
        }
      }
      const float mneg = -m_used * c1;
      #pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t pk[16];
        #pragma unroll
        for (int e = 0; e < 32; e += 4) {
          const float p0 = ex2_approx(fmaf(__uint_as_float(s[c][e]), c1, mneg));
          const float p1 = ex2_approx(fmaf(__uint_as_float(s[c][e + 1]), c1, mneg));
          const float p2 = ex2_approx(fmaf(__uint_as_float(s[c][e + 2]), c1, mneg));
          const float p3 = ex2_approx(fmaf(__uint_as_float(s[c][e + 3]), c1, mneg));
          l0 += p0; l1 += p1; l2 += p2; l3 += p3;
          pk[e >> 1] = pack_bf16x2(p0, p1);
          pk[(e >> 1) + 1] = pack_bf16x2(p2, p3);
        }
        tmem_st_x16(s_addr + c * 16, pk);                    // P columns [16c, 16c+16) of this S buffer
      }
      tmem_st_wait();
      if (!o_waited) mbar_wait(&bar_o[x][(j - 1) & 1], ((j - 1) >> 1) & 1);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar_pf[x]);
    }
    if (n_tiles > 0) {
      mbar_wait(&bar_o[x][(n_tiles - 1) & 1], ((n_tiles - 1) >> 1) & 1);
      tc_fence_after();
    }
    const float l_sum = (l0 + l1) + (l2 + l3);
    const float inv = n_tiles > 0 ? 1.f / l_sum : 0.f;
    const float beta = __ldg(p.beta);
    const int pix = q0 + x * kA2Tile + row;
    const MfT* mf = static_cast<const MfT*>(p.mf) + bb * p.mb + (frame - bb * p.T) * p.mt + pix;
    float* dst = p.out + static_cast<long long>(frame) * kA2Tile * p.hw + pix;
    #pragma unroll
    for (int c = 0; c < 4; ++c) {
      uint32_t o[32];
      if (n_tiles > 0) {
        tmem_ld_x32(o_addr + c * 32, o);
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
          dst[static_cast<long long>(ch) * p.hw] = __fadd_rn(a2_ld(mf + ch * p.mc), __fmul_rn(beta, hid));
        }
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<kA2TmemCols>(tmem);
  }
}

template <typename MfT>
int launch_attend2(const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv, const Attend2Params& p,
                   int n_qtiles, int n_frames, cudaStream_t st) {
  constexpr int kSmem = (2 + 2 * kA2Stages) * kA2TileBytes + 1024;
  auto kern = ppm_attend2_kernel<MfT>;
  PPM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
  kern<<<dim3(n_qtiles, n_frames), kA2Threads, kSmem, st>>>(mq, mk, mv, p);
  PPM_CUDA(cudaGetLastError());
  return PPM_OK;
}

// Entry used by ppm_memory_attend (ppm_attend.cu).
int attend2_dispatch(const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv, const int32_t* picked,
                     const void* mf, int mf_dtype, long long mb, long long mc, long long mt, const float* beta,
                     float scale_log2, float rescale_thr, int b, int T, int Kp, int hw, float* out,
                     cudaStream_t st) {
  Attend2Params p{};
  p.picked = picked; p.mf = mf; p.mb = mb; p.mc = mc; p.mt = mt; p.beta = beta; p.out = out;
  p.scale_log2 = scale_log2; p.rescale_thr = rescale_thr; p.T = T; p.Kp = Kp; p.hw = hw;
  const int n_qtiles = (hw + 2 * kA2Tile - 1) / (2 * kA2Tile);
  if (mf_dtype == PPM_BF16) return launch_attend2<__nv_bfloat16>(mq, mk, mv, p, n_qtiles, b * T, st);
  return launch_attend2<float>(mq, mk, mv, p, n_qtiles, b * T, st);
}

}  // namespace ppm
