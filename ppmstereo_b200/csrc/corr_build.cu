// K1: 1-D all-pairs correlation pyramid on tcgen05 tensor cores with TMA-staged feature tiles
// and the avg-pool pyramid fused into the epilogue.
// Replaces CorrBlock1D.__init__ / CorrBlock1D.corr (reference models/core/corr.py:56-72, :96-104).
//
// Per image row (b,h) the volume is the GEMM  C[i,j] = sum_d F1[d,i] * F2[d,j]  with
// M = W1, N = W2, K = D.  The operands are NCHW, i.e. for a fixed (b,h) both are [D][W]
// matrices with W contiguous: MN-major for A and for B.  TMA boxes of [KC channels][128 bytes
// of W] land in shared memory as SWIZZLE_128B MN-major tiles that tcgen05.mma consumes
// directly (no transpose pass, no repack).  bf16 inputs use kind::f16, fp32 inputs kind::tf32.
//
// One CTA computes a 128 x NT tile (NT <= 256) of one row problem:
//   warp 0    TMA producer (3-stage mbarrier ring over the channel axis)
//   warp 1    TMEM allocation + single-thread tcgen05.mma issue, accumulator in TMEM
//   warps 2-5 epilogue: tcgen05.ld (one output row per thread, 32 columns at a time), scale by
//             1/sqrt(D), form levels 1..4 by pair-averaging IN REGISTERS (a thread owns whole
//             row segments, so pooling along j needs no communication), transpose through a
//             per-warp shared-memory slab and write every level with row-contiguous stores.
// The volume is written exactly once and never re-read: 1.9375 x W1 x W2 elements per row.
#include <math.h>

#include "../../include/ppm_b200.h"
#include "common.cuh"

namespace ppm {

constexpr int kCbM = 128;            // pixels of W1 per CTA (UMMA M)
constexpr int kCbStages = 3;
constexpr int kCbThreads = 192;
constexpr int kCbMaxN = 256;
constexpr int kCbTmemCols = 256;
constexpr int kCbLevels = 5;
// per-warp epilogue slab: [32 rows][n_l + 1] floats for n_l = 32,16,8,4,2
__host__ __device__ constexpr int slab_off(int l) {
  int o = 0;
  for (int i = 0; i < l; ++i) o += 32 * ((32 >> i) + 1);
  return o;
}
constexpr int kSlabFloats = 32 * (33 + 17 + 9 + 5 + 3);

struct CorrBuildParams {
  void* level[kCbLevels];
  int width[kCbLevels];
  int n_levels;                      // pooled levels to write (<= 4)
  int D, W1, W2, NT, n_ntiles;
  float sqrt_d;
};

template <typename ElemT> struct CbTraits;
template <> struct CbTraits<__nv_bfloat16> {
  static constexpr int kChunkW = 64;    // elements in 128 bytes
  static constexpr int kKC = 32;        // channels per stage
  static constexpr int kUK = 16;        // channels per MMA
  static constexpr uint32_t kFmt = 1;   // BF16
  static constexpr uint32_t kLayout = 2;     // SWIZZLE_128B (16-byte atoms), 8 channel rows per swizzle group
  static constexpr uint32_t kSbo = 1024;
  static constexpr CUtensorMapSwizzle kTmaSwizzle = CU_TENSOR_MAP_SWIZZLE_128B;
};
template <> struct CbTraits<float> {
  static constexpr int kChunkW = 32;
  static constexpr int kKC = 16;
  static constexpr int kUK = 8;
  static constexpr uint32_t kFmt = 2;   // TF32
  // MN-major 32-bit operands must use the 32-byte-atom flavour of the 128-byte swizzle
  // (UMMA SWIZZLE_128B_BASE32B == TMA SWIZZLE_128B_ATOM_32B): 4 channel rows per swizzle group.
  static constexpr uint32_t kLayout = 1;
  static constexpr uint32_t kSbo = 512;
  static constexpr CUtensorMapSwizzle kTmaSwizzle = CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B;
};

__device__ __forceinline__ float cb_round(float v, float) { return v; }
__device__ __forceinline__ float cb_round(float v, __nv_bfloat16) { return bf16_round_f(v); }

template <typename ElemT, typename OutT>
__global__ void __launch_bounds__(kCbThreads)
corr_build_kernel(const __grid_constant__ CUtensorMap map1, const __grid_constant__ CUtensorMap map2,
                  const CorrBuildParams p) {
  using Tr = CbTraits<ElemT>;
  constexpr int kChunkBytes = Tr::kKC * 128;                    // one [KC][128 B] box
  constexpr int kAChunks = kCbM / Tr::kChunkW;
  constexpr int kBChunksMax = kCbMaxN / Tr::kChunkW;
  constexpr int kABytes = kAChunks * kChunkBytes;               // 8 KB
  constexpr int kStageBytes = kABytes + kBChunksMax * kChunkBytes;   // 24 KB

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  float* slabs = reinterpret_cast<float*>(smem + kCbStages * kStageBytes);
  __shared__ uint64_t bar_full[kCbStages], bar_empty[kCbStages], bar_acc;
  __shared__ uint32_t tmem_slot;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int mt = blockIdx.x / p.n_ntiles, nt = blockIdx.x % p.n_ntiles;
  const int h = blockIdx.y, b = blockIdx.z;
  const int m0 = mt * kCbM, n0 = nt * p.NT;
  const int n_bchunks = (p.NT + Tr::kChunkW - 1) / Tr::kChunkW;
  const int n_k = (p.D + Tr::kKC - 1) / Tr::kKC;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kCbStages; ++s) { mbar_init(&bar_full[s], 1); mbar_init(&bar_empty[s], 1); }
    mbar_init(&bar_acc, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<kCbTmemCols>(&tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (elect_one()) {
      tma_prefetch_desc(&map1);
      tma_prefetch_desc(&map2);
      const uint32_t bytes = (kAChunks + n_bchunks) * kChunkBytes;
      for (int kc = 0; kc < n_k; ++kc) {
        const int s = kc % kCbStages;
        mbar_wait(&bar_empty[s], ((kc / kCbStages) & 1) ^ 1);
        mbar_expect_tx(&bar_full[s], bytes);
        uint8_t* st = smem + s * kStageBytes;
        #pragma unroll
        for (int c = 0; c < kAChunks; ++c)
          tma_load_4d(st + c * kChunkBytes, &map1, &bar_full[s], m0 + c * Tr::kChunkW, h, kc * Tr::kKC, b);
        for (int c = 0; c < n_bchunks; ++c)
          tma_load_4d(st + kABytes + c * kChunkBytes, &map2, &bar_full[s], n0 + c * Tr::kChunkW, h,
                      kc * Tr::kKC, b);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    if (elect_one()) {
      const uint32_t idesc = umma_idesc(Tr::kFmt, /*a MN-major*/ 1, /*b MN-major*/ 1, kCbM, p.NT);
      for (int kc = 0; kc < n_k; ++kc) {
        const int s = kc % kCbStages;
        mbar_wait(&bar_full[s], (kc / kCbStages) & 1);
        tc_fence_after();
        const uint32_t a0 = smem_u32(smem + s * kStageBytes);
        const uint32_t b0 = a0 + kABytes;
        #pragma unroll
        for (int k = 0; k < Tr::kKC / Tr::kUK; ++k) {
          // MN-major: LBO = stride between 128-byte MN chunks, SBO = one swizzle group of channel rows
          const uint64_t ad = umma_desc_sw128(a0 + k * Tr::kUK * 128, kChunkBytes, Tr::kSbo, Tr::kLayout);
          const uint64_t bd = umma_desc_sw128(b0 + k * Tr::kUK * 128, kChunkBytes, Tr::kSbo, Tr::kLayout);
          if constexpr (Tr::kFmt == 2) umma_tf32_ss(tmem_base, ad, bd, idesc, (kc | k) != 0);
          else                         umma_f16_ss(tmem_base, ad, bd, idesc, (kc | k) != 0);
        }
        umma_commit(&bar_empty[s]);
      }
      umma_commit(&bar_acc);
    }
  } else {
    // ------------------------------------------------------------ epilogue
    const int q = warp & 3;                         // TMEM lane quarter this warp may access
    float* slab = slabs + (warp - 2) * kSlabFloats;
    const long long bh = static_cast<long long>(b) * gridDim.y + h;
    const long long row_base = bh * p.W1 + m0 + q * 32;          // global row of slab row 0
    const int rows_valid = min(32, p.W1 - (m0 + q * 32));        // may be <= 0
    mbar_wait(&bar_acc, 0);
    tc_fence_after();
    const int n_chunks = p.NT / 32;
    for (int cc = 0; cc < n_chunks; ++cc) {
      uint32_t raw[32];
      tmem_ld_x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + cc * 32, raw);
      tmem_ld_wait();
      float v[32];
      #pragma unroll
      for (int j = 0; j < 32; ++j) {
        float a = cb_round(__uint_as_float(raw[j]), OutT());
        v[j] = cb_round(__fdiv_rn(a, p.sqrt_d), OutT());
      }
      __syncwarp();                                  // previous chunk's write-out is done
      #pragma unroll
      for (int j = 0; j < 32; ++j) slab[slab_off(0) + lane * 33 + j] = v[j];
      #pragma unroll
      for (int l = 1; l < kCbLevels; ++l) {
        const int n = 32 >> l;
        #pragma unroll
        for (int j = 0; j < n; ++j) {
          v[j] = cb_round(__fmul_rn(__fadd_rn(v[2 * j], v[2 * j + 1]), 0.5f), OutT());
          slab[slab_off(l) + lane * (n + 1) + j] = v[j];
        }
      }
      __syncwarp();
      // write-out: row-contiguous stores, level by level
      #pragma unroll
      for (int l = 0; l < kCbLevels; ++l) {
        if (l > p.n_levels) break;
        const int n = 32 >> l;
        const int Wl = p.width[l];
        const int col0 = (n0 + cc * 32) >> l;
        OutT* dst = static_cast<OutT*>(p.level[l]);
        const float* src = slab + slab_off(l);
        if (sizeof(OutT) == 2 && (Wl & 1) == 0) {
          const int u = n / 2;                       // 4-byte units per row
          for (int idx = lane; idx < 32 * u; idx += 32) {
            const int r = idx / u, c = 2 * (idx % u);
            if (r < rows_valid && col0 + c < Wl) {
              const uint32_t pk = pack_bf16x2(src[r * (n + 1) + c], src[r * (n + 1) + c + 1]);
              *reinterpret_cast<uint32_t*>(reinterpret_cast<__nv_bfloat16*>(dst) +
                                           (row_base + r) * Wl + col0 + c) = pk;
            }
          }
        } else {
          for (int idx = lane; idx < 32 * n; idx += 32) {
            const int r = idx / n, c = idx % n;
            if (r < rows_valid && col0 + c < Wl) dst[(row_base + r) * Wl + col0 + c] = OutT(src[r * (n + 1) + c]);
          }
        }
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<kCbTmemCols>(tmem_base);
  }
}

template <typename ElemT, typename OutT>
static int launch_corr_build(const void* f1, const void* f2, int B, int D, int H, int W1, int W2,
                             int ld1, int ld2, void* const* levels, int n_levels, cudaStream_t st) {
  using Tr = CbTraits<ElemT>;
  constexpr int kStageBytes = (kCbM / Tr::kChunkW + kCbMaxN / Tr::kChunkW) * Tr::kKC * 128;
  constexpr int kSmem = kCbStages * kStageBytes + 4 * kSlabFloats * 4 + 1024;
  CorrBuildParams p{};
  p.n_levels = n_levels;
  for (int l = 0; l <= n_levels; ++l) { p.level[l] = levels[l]; p.width[l] = W2 >> l; }
  p.D = D; p.W1 = W1; p.W2 = W2;
  p.n_ntiles = (W2 + kCbMaxN - 1) / kCbMaxN;
  p.NT = (((W2 + p.n_ntiles - 1) / p.n_ntiles) + 31) / 32 * 32;
  p.sqrt_d = sqrtf(static_cast<float>(D));
  const CUtensorMapDataType dt = sizeof(ElemT) == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                                                    : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  CUtensorMap m1, m2;
  const uint32_t box[4] = {static_cast<uint32_t>(Tr::kChunkW), 1, static_cast<uint32_t>(Tr::kKC), 1};
  {
    const uint64_t dims[4] = {(uint64_t)W1, (uint64_t)H, (uint64_t)D, (uint64_t)B};
    const uint64_t str[3] = {(uint64_t)ld1 * sizeof(ElemT), (uint64_t)ld1 * H * sizeof(ElemT),
                             (uint64_t)ld1 * H * D * sizeof(ElemT)};
    int rc = make_tensor_map(&m1, dt, 4, f1, dims, str, box, Tr::kTmaSwizzle);
    if (rc) return rc;
  }
  {
    const uint64_t dims[4] = {(uint64_t)W2, (uint64_t)H, (uint64_t)D, (uint64_t)B};
    const uint64_t str[3] = {(uint64_t)ld2 * sizeof(ElemT), (uint64_t)ld2 * H * sizeof(ElemT),
                             (uint64_t)ld2 * H * D * sizeof(ElemT)};
    int rc = make_tensor_map(&m2, dt, 4, f2, dims, str, box, Tr::kTmaSwizzle);
    if (rc) return rc;
  }
  auto kern = corr_build_kernel<ElemT, OutT>;
  PPM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
  dim3 grid(((W1 + kCbM - 1) / kCbM) * p.n_ntiles, H, B);
  kern<<<grid, kCbThreads, kSmem, st>>>(m1, m2, p);
  PPM_CUDA(cudaGetLastError());
  return PPM_OK;
}

}  // namespace ppm

extern "C" int ppm_corr_build(const void* fmap1, const void* fmap2, int in_dtype, int B, int D, int H,
                              int W1, int W2, int ld1, int ld2, void* const* levels, int n_levels,
                              int out_dtype, void* stream) {
  using namespace ppm;
  PPM_CHECK(fmap1 && fmap2 && levels, PPM_ERR_ARG, "ppm_corr_build: null pointer");
  PPM_CHECK(B > 0 && D > 0 && H > 0 && W1 > 0 && W2 > 0, PPM_ERR_ARG, "ppm_corr_build: bad shape");
  PPM_CHECK(n_levels >= 0 && n_levels <= 4, PPM_ERR_UNSUPPORTED, "ppm_corr_build: n_levels %d not in [0,4]", n_levels);
  PPM_CHECK((W2 >> n_levels) >= 1, PPM_ERR_ARG,
            "ppm_corr_build: W2 = %d too small for %d pooled levels (reference: avg_pool2d output size is too small)",
            W2, n_levels);
  PPM_CHECK((in_dtype == PPM_F32 || in_dtype == PPM_BF16) && (out_dtype == PPM_F32 || out_dtype == PPM_BF16),
            PPM_ERR_ARG, "ppm_corr_build: bad dtype");
  const int wmul = in_dtype == PPM_BF16 ? 8 : 4;
  PPM_CHECK(ld1 >= W1 && ld2 >= W2 && ld1 % wmul == 0 && ld2 % wmul == 0, PPM_ERR_ARG,
            "ppm_corr_build: row pitches %d, %d must be >= W and multiples of %d (TMA 16-byte pitch)", ld1, ld2, wmul);
  PPM_CHECK((reinterpret_cast<uintptr_t>(fmap1) & 15) == 0 && (reinterpret_cast<uintptr_t>(fmap2) & 15) == 0,
            PPM_ERR_ARG, "ppm_corr_build: feature maps must be 16-byte aligned");
  PPM_CHECK(H <= 65535 && B <= 65535, PPM_ERR_UNSUPPORTED, "ppm_corr_build: H or B > 65535");
  for (int l = 0; l <= n_levels; ++l)
    PPM_CHECK(levels[l] && (reinterpret_cast<uintptr_t>(levels[l]) & 3) == 0, PPM_ERR_ARG,
              "ppm_corr_build: level %d pointer null or misaligned", l);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (in_dtype == PPM_BF16) {
    if (out_dtype == PPM_BF16)
      return launch_corr_build<__nv_bfloat16, __nv_bfloat16>(fmap1, fmap2, B, D, H, W1, W2, ld1, ld2, levels, n_levels, st);
    return launch_corr_build<__nv_bfloat16, float>(fmap1, fmap2, B, D, H, W1, W2, ld1, ld2, levels, n_levels, st);
  }
  if (out_dtype == PPM_BF16)
    return launch_corr_build<float, __nv_bfloat16>(fmap1, fmap2, B, D, H, W1, W2, ld1, ld2, levels, n_levels, st);
  return launch_corr_build<float, float>(fmap1, fmap2, B, D, H, W1, W2, ld1, ld2, levels, n_levels, st);
}
