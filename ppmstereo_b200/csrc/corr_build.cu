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
// One work item is a 128 x NT tile (NT <= 256) of one row problem:
//   warp 0    TMA producer (mbarrier ring of 64-channel stages over the channel axis; every byte of shared
//             memory that is left goes to this ring -- the load path is latency-bound by the bytes in flight)
//   warp 1    TMEM allocation + tcgen05.mma issue (the warp walks the loop convergently, one elected lane
//             issues; descriptors formed with one add; the next stage's barrier probed ahead of the MMAs)
//   warps 2-9 epilogue, every warp on its own: tcgen05.ld (one output row per thread, 32 columns at a
//             time), hand the accumulator back right after the warp's last load, scale by 1/sqrt(D), form
//             levels 1..4 by pair-averaging IN REGISTERS (a thread owns whole row segments, so pooling
//             along j needs no communication).  A warp takes units of two adjacent chunks aligned to 64 columns
//             of the image row: level 0 of a unit goes into the warp's private, XOR-swizzled [32 rows][128 B]
//             staging box and leaves through one TMA store (a thread's own row segments written with LSU stores
//             touch 32 different lines per instruction); level 1 (one 32-byte sector per thread and chunk) and
//             levels 2-4 (one aligned 32 / 16 / 8-byte piece per thread and UNIT) go straight from registers.
//             No CTA-wide barrier anywhere in the epilogue.
// Persistent: one CTA per SM loops over work items; the smem ring runs on across items and the
// accumulator is double-buffered in TMEM (2 x 256 columns), so the TMA loads and MMAs of tile i+1
// overlap the epilogue of tile i.  When an image row has several M tiles, a work item is (image
// row, N tile): its F2 tile stays in shared memory while the F1 tiles stream past it.  When the row
// has exactly two N tiles (W2 = 320 -> 2 x 160), a 2-CTA cluster owns the image row (CTA rank = N tile)
// and every F1 tile is fetched once and multicast to both CTAs (each CTA issues half of the boxes; the
// ring's empty barriers take one tcgen05.commit arrival from each CTA).
// Measured at the 720p 1/4-res shape (2.67 GB algorithmic, 0.41 ms at the HBM copy peak), CUDA-graph
// replay: 0.465 ms (0.88); history 2.49 -> 0.84 -> 0.60 -> 0.58 -> 0.50 -> 0.465.  Ablations (PPM_CB_ABLATE bit mask,
// same box): MMA + barrier skeleton 0.25 ms, + loads 0.31, + epilogue arithmetic (no stores) 0.27,
// + stores (no loads) 0.48 -- the store side (TMA box rows + scattered LSU lines) is what is left.
// The volume is written exactly once and never re-read: 1.9375 x W1 x W2 elements per row.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/ppm_b200.h"
#include "common.cuh"

namespace ppm {

constexpr int kCbM = 128;            // pixels of W1 per CTA (UMMA M)
constexpr int kCbStages = 6;         // streaming mode: at most this many [F1 | F2] stages
constexpr int kCbMaxRing = 16;       // barriers: ring stages (either mode) and resident F2 channel chunks
constexpr int kCbThreads = 320;      // TMA warp, MMA warp, 8 epilogue warps
constexpr int kCbMaxN = 256;
constexpr int kCbTmemCols = 512;     // two 256-column accumulators: MMA of tile i+1 overlaps the epilogue of tile i
constexpr int kCbEpiWarps = 8;
constexpr int kCbLevels = 5;

// Timing ablations and tuning switches (PPM_CB_* environment variables, CorrBuildParams::ablate) exist only in an
// experiments build (-DPPM_EXPERIMENTS, scripts/ only).  The library build reads no environment variable and the
// ablation branches fold away at compile time.
#ifdef PPM_EXPERIMENTS
#define CB_ABLATE(p) ((p).ablate)
static inline const char* cb_env(const char* name) { return getenv(name); }
#else
#define CB_ABLATE(p) 0
static inline const char* cb_env(const char*) { return nullptr; }
#endif

struct CorrBuildParams {
  void* level[kCbLevels];
  int width[kCbLevels];
  int n_levels;                      // pooled levels to write (<= 4)
  int D, H, W1, W2, NT, n_ntiles, n_mtiles;
  long long n_tiles, n_items;        // n_items = image rows x N tiles (resident mode)
  float sqrt_d;
  float rcp_sqrt_d;                  // RN(1 / sqrt_d)
  float inv_sqrt_d;                  // 1/sqrt(D) when that is exact (sqrt(D) a power of two), else 0
  uint32_t inv_bf16x2;               // the same as a (bf16, bf16) pair
  int pair_vec[kCbLevels];           // levels 2-4: rows allow one aligned store per 64-column unit (bf16 pyramid)
  int vec_ok[kCbLevels];             // level rows allow the full-width vector store
  int res_b;                         // 1: the F2 tile of a work item stays in smem for all its M tiles
  int a_stages;                      // res_b: stages of the F1 ring
  int tma_out;                       // levels 0..2 leave through TMA stores from a staged smem tile
  int ablate;                        // experiments only (PPM_CB_ABLATE bit mask): 1 no stores, 2 no TMA-store issue, 4 no loads, 8 no epilogue body, 16 tmem_ld only
  int pair;                          // resident mode, 2 N tiles: a 2-CTA cluster owns an image row (CTA rank = N tile)
                                     // and every F1 tile is fetched once and multicast to both CTAs
};

// Level-0 store maps: [0] = [32 rows (one epilogue warp)][two 32-column chunks = 128 B], SWIZZLE_128B;
// [1] = [32 rows][one chunk = 64 B], SWIZZLE_64B (the tail unit of a tile with an odd number of chunks).  The swizzle
// spreads the warp's row-per-thread 16-byte writes over all banks.
struct CbStoreMaps { CUtensorMap m[2]; };
constexpr int kCbWarpStage = 4096;                // per-warp staging box
// Byte offset of 16-byte piece `u` of row `row` inside a staged [128][bp] box (bp = 128, 64, 32 or 16), following
// the TMA swizzle of that width: the piece index is XORed with the address bits [7, 7 + log2(bp/16)).
__device__ __forceinline__ int cb_stage_off(int row, int u, int bp) {
  const int sw = ((row * bp) >> 7) & ((bp >> 4) - 1);
  return row * bp + ((u ^ sw) << 4);
}

template <typename ElemT> struct CbTraits;
template <> struct CbTraits<__nv_bfloat16> {
  static constexpr int kChunkW = 64;    // elements in 128 bytes
  static constexpr int kKC = 64;        // channels per stage
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


// packed bf16 helpers of the fast epilogue (HMUL2.BF16 / HFMA2.BF16)
__device__ __forceinline__ uint32_t cb_mul_bf16x2(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}
__device__ __forceinline__ uint32_t cb_avg_bf16x2(uint32_t a, uint32_t b) {     // bf16(a/2 + b/2), one rounding
  const uint32_t half2 = 0x3F003F00u;                                            // (0.5, 0.5)
  uint32_t hb, d;
  asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(hb) : "r"(b), "r"(half2));            // exact
  asm("fma.rn.bf16x2 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(half2), "r"(hb));
  return d;
}

// ---- TMA store (shared -> global), bulk-group completion
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* m, const void* src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(src)), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}
__device__ __forceinline__ void tma_load_4d_mc(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2,
                                               int c3, uint16_t cta_mask) {   // lands at the same offset in every CTA of the mask
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster"
      " [%0], [%1, {%3, %4, %5, %6}], [%2], %7;"
      ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3),
      "h"(cta_mask)
      : "memory");
}
__device__ __forceinline__ void umma_commit_mc(uint64_t* bar, uint16_t cta_mask) {   // arrive on `bar` of every CTA in the mask
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"(cta_mask) : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int kPending>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(kPending) : "memory");
}
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// ---- fast epilogue for bf16 output with an exact power-of-two 1/sqrt(D)
// Rounding goes through cvt.rn.bf16x2.f32 (F2FP, ALU pipe): the scalar cvt.rn.bf16.f32 is an XU-pipe
// instruction (16 lanes/clk/SM) and made the first epilogue conversion-bound.  Scaling by a power of
// two commutes with rounding, so bf16(bf16(acc) / sqrt(D)) == bf16(acc * 2^-k) and the packed word
// is both the value stored and (unpacked) the input of the next pyramid level.
__device__ __forceinline__ void cb_st256(void* p, const uint32_t* w) {
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
               ::"l"(p), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7])
               : "memory");
}
// words: n/2 packed bf16 pairs of one level chunk (n = 32 >> l elements); vec 2 = 32-byte aligned rows,
// 1 = aligned to min(16, chunk) bytes, 0 = element stores.
__device__ __forceinline__ void cb_store_words(__nv_bfloat16* dst, const uint32_t* w, int n, int valid, int vec) {
  if (vec && valid >= n) {
    if (n == 32) {
      if (vec == 2) { cb_st256(dst, w); cb_st256(dst + 16, w + 8); }
      else {
        #pragma unroll
        for (int j = 0; j < 4; ++j) reinterpret_cast<uint4*>(dst)[j] = make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
      }
    } else if (n == 16) {
      if (vec == 2) cb_st256(dst, w);
      else {
        #pragma unroll
        for (int j = 0; j < 2; ++j) reinterpret_cast<uint4*>(dst)[j] = make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
      }
    } else if (n == 8) {
      *reinterpret_cast<uint4*>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
    } else if (n == 4) {
      *reinterpret_cast<uint2*>(dst) = make_uint2(w[0], w[1]);
    } else {
      *reinterpret_cast<uint32_t*>(dst) = w[0];
    }
  } else {
    #pragma unroll
    for (int j = 0; j < 32; ++j)
      if (j < n && j < valid)
        reinterpret_cast<uint16_t*>(dst)[j] = static_cast<uint16_t>((j & 1) ? (w[j >> 1] >> 16) : (w[j >> 1] & 0xFFFFu));
  }
}

// fp32 pyramid: store v[0..n) (n = 32, 16, 8, 4 or 2, compile-time at every call after unrolling) to dst[0..n);
// only the first `valid` elements exist in the row.  vec: dst is aligned to min(16, n*4) bytes.
__device__ __forceinline__ void cb_store_row(float* dst, const float (&v)[32], int n, int valid, bool vec) {
  if (vec && valid >= n) {
    if (n >= 4) {
      #pragma unroll
      for (int j = 0; j < 8; ++j)
        if (4 * j < n) reinterpret_cast<float4*>(dst)[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    } else {
      *reinterpret_cast<float2*>(dst) = make_float2(v[0], v[1]);
    }
  } else {
    #pragma unroll
    for (int j = 0; j < 32; ++j)
      if (j < n && j < valid) dst[j] = v[j];
  }
}

template <typename ElemT, typename OutT>
__global__ void __launch_bounds__(kCbThreads, 1)
corr_build_kernel(const __grid_constant__ CUtensorMap map1, const __grid_constant__ CUtensorMap map2,
                  const __grid_constant__ CbStoreMaps omap, const CorrBuildParams p) {
  using Tr = CbTraits<ElemT>;
  constexpr int kChunkBytes = Tr::kKC * 128;                    // one [KC][128 B] box
  constexpr int kAChunks = kCbM / Tr::kChunkW;
  constexpr int kBChunksMax = kCbMaxN / Tr::kChunkW;
  constexpr int kABytes = kAChunks * kChunkBytes;               // 8 KB
  constexpr int kStageBytes = kABytes + kBChunksMax * kChunkBytes;   // 24 KB

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar_full[kCbMaxRing], bar_empty[kCbMaxRing], bar_bfull[kCbMaxRing], bar_bempty[kCbMaxRing];
  __shared__ uint64_t bar_acc_full[2], bar_acc_empty[2];
  __shared__ uint32_t tmem_slot;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_bchunks = (p.NT + Tr::kChunkW - 1) / Tr::kChunkW;
  const int n_k = (p.D + Tr::kKC - 1) / Tr::kKC;
  // smem: streaming mode  [a_stages x (F1 | F2)] [staging]
  //       resident mode   [a_stages x F1 8 KB] [n_k x F2 channel chunk] [staging]
  const int b_chunk_bytes = n_bchunks * kChunkBytes;
  const int n_ring = p.a_stages;
  uint8_t* b_res = smem + p.a_stages * kABytes;
  uint8_t* stage0 = smem + (p.res_b ? p.a_stages * kABytes + n_k * b_chunk_bytes : p.a_stages * kStageBytes);
  // Work order.  Streaming: tile t = blockIdx.x + i*grid -> (bh, mt, nt).  Resident: item = blockIdx.x +
  // (i / n_mtiles)*grid -> (bh, nt), and the item's M tiles follow each other on the same CTA.
  const long long n_rows_bh = p.n_items / p.n_ntiles;
  const uint32_t crank = p.pair ? cluster_ctarank() : 0;
  const long long my_tiles = p.pair ? ((n_rows_bh - (blockIdx.x >> 1) + (gridDim.x >> 1) - 1) / (gridDim.x >> 1)) * p.n_mtiles
      : p.res_b
      ? ((p.n_items - blockIdx.x + gridDim.x - 1) / gridDim.x) * p.n_mtiles
      : (p.n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x;
  auto item_of = [&](long long j) -> long long {          // j-th work item of this CTA -> global item (bh * n_ntiles + nt)
    if (p.pair) return ((blockIdx.x >> 1) + j * (gridDim.x >> 1)) * 2 + crank;
    return blockIdx.x + j * gridDim.x;
  };
  auto decode = [&](long long i, int& bh, int& m0, int& n0, int& mt) {
    if (p.res_b) {
      const long long item = item_of(i / p.n_mtiles);
      mt = static_cast<int>(i % p.n_mtiles);
      bh = static_cast<int>(item / p.n_ntiles);
      n0 = static_cast<int>(item % p.n_ntiles) * p.NT;
    } else {
      const long long t = blockIdx.x + i * gridDim.x;
      const int r = static_cast<int>(t % (p.n_mtiles * p.n_ntiles));
      bh = static_cast<int>(t / (p.n_mtiles * p.n_ntiles));
      mt = r / p.n_ntiles;
      n0 = (r % p.n_ntiles) * p.NT;
    }
    m0 = mt * kCbM;
  };

  if (threadIdx.x == 0) {
    for (int s = 0; s < kCbMaxRing; ++s) {
      mbar_init(&bar_full[s], 1); mbar_init(&bar_empty[s], p.pair ? 2 : 1); mbar_init(&bar_bfull[s], 1); mbar_init(&bar_bempty[s], 1);
    }
    for (int s = 0; s < 2; ++s) { mbar_init(&bar_acc_full[s], 1); mbar_init(&bar_acc_empty[s], kCbEpiWarps); }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<kCbTmemCols>(&tmem_slot);
  tc_fence_before();
  __syncthreads();
  if (p.pair) cluster_sync_all();          // the peer's barriers exist before anything is multicast into them
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (elect_one()) {
      tma_prefetch_desc(&map1);
      tma_prefetch_desc(&map2);
      int s = 0;
      uint32_t ph = 1;
      for (long long i = 0; i < my_tiles; ++i) {
        int bh, m0, n0, mt;
        decode(i, bh, m0, n0, mt);
        const int b = bh / p.H, h = bh % p.H;
        const uint32_t bph = static_cast<uint32_t>((i / p.n_mtiles) & 1) ^ 1;     // resident: item parity
        for (int kc = 0; kc < n_k; ++kc) {
          if (p.res_b && mt == 0) {                      // the item's F2 tile, once, chunk by chunk
            mbar_wait(&bar_bempty[kc], bph);
            if ((CB_ABLATE(p) & 4)) mbar_arrive(&bar_bfull[kc]); else
            mbar_expect_tx(&bar_bfull[kc], b_chunk_bytes);
            for (int c = 0; c < n_bchunks && !(CB_ABLATE(p) & 4); ++c)
              tma_load_4d(b_res + kc * b_chunk_bytes + c * kChunkBytes, &map2, &bar_bfull[kc], n0 + c * Tr::kChunkW, h,
                          kc * Tr::kKC, b);
          }
          mbar_wait(&bar_empty[s], ph);
          if ((CB_ABLATE(p) & 4)) { mbar_arrive(&bar_full[s]); if (++s == n_ring) { s = 0; ph ^= 1; } continue; }
          mbar_expect_tx(&bar_full[s], p.res_b ? kABytes : kABytes + b_chunk_bytes);
          uint8_t* st = smem + s * (p.res_b ? kABytes : kStageBytes);
          #pragma unroll
          for (int c = 0; c < kAChunks; ++c) {
            if (!p.pair)
              tma_load_4d(st + c * kChunkBytes, &map1, &bar_full[s], m0 + c * Tr::kChunkW, h, kc * Tr::kKC, b);
            else if ((c & 1) == static_cast<int>(crank))          // each CTA fetches half of the tile for both
              tma_load_4d_mc(st + c * kChunkBytes, &map1, &bar_full[s], m0 + c * Tr::kChunkW, h, kc * Tr::kKC, b, 3);
          }
          if (!p.res_b)
            for (int c = 0; c < n_bchunks; ++c)
              tma_load_4d(st + kABytes + c * kChunkBytes, &map2, &bar_full[s], n0 + c * Tr::kChunkW, h,
                          kc * Tr::kKC, b);
          if (++s == n_ring) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    // The whole warp walks the loop (uniform control flow: ring indices and descriptor words stay in
    // uniform registers, no per-MMA R2UR traffic); one elected lane issues the tcgen05 instructions.  A
    // descriptor is (hi, lo) with hi constant and lo = base | (smem address >> 4): another K slice, ring
    // stage or resident chunk is one integer add.  The next stage's barrier is probed (non-blocking) BEFORE
    // this stage's MMAs are issued: a probe's answer takes ~100 cycles even on a completed phase and MMA
    // issue runs only a couple of instructions ahead of execution, so a blocking wait in front of every
    // pair of MMAs left the tensor pipe idle most of the time (measured: the MMA + barrier skeleton alone
    // took 0.39 ms of the 0.41 ms HBM budget at the 720p 1/4-res shape).
    {
      const bool leader = elect_one();
      const uint32_t idesc = (CB_ABLATE(p) & 32) ? umma_idesc(Tr::kFmt, 0, 0, kCbM, p.NT)      // timing experiment only
                                            : umma_idesc(Tr::kFmt, /*a MN-major*/ 1, /*b MN-major*/ 1, kCbM, p.NT);
      // MN-major: LBO = stride between 128-byte MN chunks, SBO = one swizzle group of channel rows
      const uint64_t d0 = umma_desc_sw128(0, kChunkBytes, Tr::kSbo, Tr::kLayout);
      const uint32_t desc_hi = static_cast<uint32_t>(d0 >> 32);
      // (& 0x3FFF: inside a cluster the shared-window address of rank 1 carries the CTA rank above bit 18)
      const uint32_t a_lo0 = static_cast<uint32_t>(d0) + ((smem_u32(smem) >> 4) & 0x3FFF);
      const uint32_t b_lo0 = static_cast<uint32_t>(d0) + (((p.res_b ? smem_u32(b_res) : smem_u32(smem) + kABytes) >> 4) & 0x3FFF);
      const uint32_t a_step = static_cast<uint32_t>(p.res_b ? kABytes : kStageBytes) >> 4;
      const uint32_t b_step = p.res_b ? static_cast<uint32_t>(b_chunk_bytes) >> 4 : a_step;
      constexpr uint32_t kKStep = (Tr::kUK * 128) >> 4;
      auto mk = [&](uint32_t lo) -> uint64_t {
        uint64_t d;
        asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"(desc_hi));
        return d;
      };
      asm volatile(".reg .pred ppm_cb_pr0, ppm_cb_pr1;");
      auto probe0 = [&](uint64_t* bar, uint32_t parity) {
        asm volatile("mbarrier.test_wait.parity.shared::cta.b64 ppm_cb_pr0, [%0], %1;" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
      };
      auto probe1 = [&](uint64_t* bar, uint32_t parity) {
        asm volatile("mbarrier.test_wait.parity.shared::cta.b64 ppm_cb_pr1, [%0], %1;" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
      };
      auto probed0 = [&]() -> bool { uint32_t ok; asm volatile("selp.b32 %0, 1, 0, ppm_cb_pr0;" : "=r"(ok)::"memory"); return ok != 0; };
      auto probed1 = [&]() -> bool { uint32_t ok; asm volatile("selp.b32 %0, 1, 0, ppm_cb_pr1;" : "=r"(ok)::"memory"); return ok != 0; };
      int s = 0;
      uint32_t ph = 0;
      if (my_tiles > 0) probe0(&bar_full[0], 0);
      for (long long it = 0; it < my_tiles; ++it) {
        const int buf = static_cast<int>(it & 1);
        const int mt = p.res_b ? static_cast<int>(it % p.n_mtiles) : 0;
        const bool first_m = p.res_b && mt == 0, last_m = p.res_b && mt == p.n_mtiles - 1;
        const uint32_t bph = static_cast<uint32_t>((it / p.n_mtiles) & 1);
        if (first_m) probe1(&bar_bfull[0], bph);
        mbar_wait(&bar_acc_empty[buf], static_cast<uint32_t>((it >> 1) & 1) ^ 1);   // epilogue has drained this accumulator
        const uint32_t acc = tmem_base + buf * kCbMaxN;
        for (int kc = 0; kc < n_k; ++kc) {
          if (!probed0()) mbar_wait(&bar_full[s], ph);
          if (first_m && (!probed1())) mbar_wait(&bar_bfull[kc], bph);
          tc_fence_after();
          const uint32_t a_lo = a_lo0 + s * a_step;
          const uint32_t b_lo = b_lo0 + (p.res_b ? kc : s) * b_step;
          const int s_cur = s;
          if (++s == n_ring) { s = 0; ph ^= 1; }
          probe0(&bar_full[s], ph);                                   // next stage (of this or the next tile)
          if (first_m && kc + 1 < n_k) probe1(&bar_bfull[kc + 1], bph);
          if (leader) {
            #pragma unroll
            for (int k = 0; k < Tr::kKC / Tr::kUK; ++k) {
              if constexpr (Tr::kFmt == 2) umma_tf32_ss(acc, mk(a_lo + k * kKStep), mk(b_lo + k * kKStep), idesc, (kc | k) != 0);
              else                         umma_f16_ss(acc, mk(a_lo + k * kKStep), mk(b_lo + k * kKStep), idesc, (kc | k) != 0);
            }
            if (p.pair) umma_commit_mc(&bar_empty[s_cur], 3); else umma_commit(&bar_empty[s_cur]);
            if (last_m) umma_commit(&bar_bempty[kc]);                 // last M tile of the item: chunk is free
          }
          __syncwarp();
        }
        if (leader) umma_commit(&bar_acc_full[buf]);
        __syncwarp();
      }
    }
  } else {
    // ------------------------------------------------------------ epilogue
    // Thread = one output row (TMEM lane); two warps share a lane quarter and take alternate
    // 32-column chunks.  A thread holds whole row segments, so the pooled levels come from
    // pair-averaging in registers and every level is stored straight from registers with the
    // widest aligned vector store (a 32-column chunk is 64 contiguous bytes of a bf16 row).
    const int q = warp & 3;                         // TMEM lane quarter this warp may access
    const int half = (warp - 2) >> 2;               // which chunks this warp takes
    const int n_chunks = p.NT / 32;
    for (long long it = 0; it < my_tiles; ++it) {
      int bh, m0, n0, mt_unused;
      decode(it, bh, m0, n0, mt_unused);
      const int m = m0 + q * 32 + lane;             // pixel of W1 owned by this thread
      const long long row = static_cast<long long>(bh) * p.W1 + m;
      const bool row_ok = m < p.W1;
      const int buf = static_cast<int>(it & 1);
      mbar_wait(&bar_acc_full[buf], static_cast<uint32_t>((it >> 1) & 1));
      tc_fence_after();
      const uint32_t acc = tmem_base + buf * kCbMaxN + (static_cast<uint32_t>(q * 32) << 16);
      // A warp takes units of two adjacent chunks whose first column is a multiple of 64 in the IMAGE row (global
      // chunk index g = n0/32 + cc, unit = g >> 1), so that a unit is one 128-byte line of level 0 (one TMA store box),
      // one 32-byte sector of level 2, 16 bytes of level 3 and 8 of level 4 whatever n0 is (n0 = 160: units {5}, {6,7},
      // {8,9}).  The two warps of a lane quarter take alternate units and swap from tile to tile.
      const int hsel = (half ^ static_cast<int>(it & 1)) & 1;
      const int g_first = n0 >> 5, g_last = g_first + n_chunks - 1;
      const int u_first = g_first >> 1, u_last = g_last >> 1;
      int last_cc = -1;
      for (int ug = u_last; ug >= u_first; --ug)
        if (((ug - u_first) & 1) == hsel) { last_cc = (2 * ug + 1 < g_last ? 2 * ug + 1 : g_last) - g_first; break; }
      bool released = false;
      uint8_t* const wstg = stage0 + (warp - 2) * kCbWarpStage;   // this warp's private staging box
      // a warp whose 32 rows all lie past W1 (the partial last M tile) has nothing to store: it only hands the
      // accumulator back
      const bool warp_has_rows = m0 + q * 32 < p.W1;
      for (int ug = u_first + hsel; ug <= u_last && warp_has_rows && !(CB_ABLATE(p) & 8); ug += 2) {
      const int ga = 2 * ug > g_first ? 2 * ug : g_first, gb = 2 * ug + 1 < g_last ? 2 * ug + 1 : g_last;
      const bool single = ga == gb;                              // edge unit of one chunk: [32][64 B] box
      const bool pair_full = !single && 64 * ug + 64 <= p.W2;    // levels 2-4 of the unit leave in one store each
      uint32_t keep[7];                                          // levels 2 | 3 | 4 of the unit's first chunk
      for (int g = ga; g <= gb; ++g) {
        const int cc = g - g_first;
        uint32_t raw[32];
        tmem_ld_x32(acc + cc * 32, raw);
        tmem_ld_wait();
        if (cc == last_cc) {                          // last read of this accumulator by this warp: hand it back now
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&bar_acc_empty[buf]);
          released = true;
        }
        if (CB_ABLATE(p) & 16) {
          uint32_t x = 0;
          #pragma unroll
          for (int j = 0; j < 32; ++j) x ^= raw[j];
          if (x == 0x12345678u) *reinterpret_cast<uint32_t*>(p.level[0]) = x;
          continue;
        }
        if constexpr (sizeof(OutT) == 2) {
          {
            // bf16 pyramid: packed bf16x2 arithmetic
            //   level 0   bf16(acc) * 2^-k            (F2FP pack, HMUL2.BF16 -- exact) when sqrt(D) is a power of two,
            //             else bf16(bf16(acc) / sqrt(D)), the quotient correctly rounded in fp32 (D = 128: the reference's
            //             bf16 / fp32-scalar division)
            //   level l   bf16(lo/2 + hi/2) per word  (2 PRMT + HMUL2 + HFMA2 per two outputs; one rounding,
            //             the same value as the reference's fp32 average rounded to bf16)
            uint32_t w[16];
            if (p.inv_sqrt_d != 0.f) {
              #pragma unroll
              for (int j = 0; j < 16; ++j)
                w[j] = cb_mul_bf16x2(pack_bf16x2(__uint_as_float(raw[2 * j]), __uint_as_float(raw[2 * j + 1])), p.inv_bf16x2);
            } else {
              #pragma unroll
              for (int j = 0; j < 16; ++j) {
                const uint32_t pk = pack_bf16x2(__uint_as_float(raw[2 * j]), __uint_as_float(raw[2 * j + 1]));
                // x / s as q = x*r, q += (x - q*s)*r with r = RN(1/s): the correctly rounded quotient (Markstein), 3 instructions
                const float x0 = __uint_as_float(pk << 16), x1 = __uint_as_float(pk & 0xFFFF0000u);
                float q0 = __fmul_rn(x0, p.rcp_sqrt_d), q1 = __fmul_rn(x1, p.rcp_sqrt_d);
                q0 = __fmaf_rn(__fmaf_rn(-q0, p.sqrt_d, x0), p.rcp_sqrt_d, q0);
                q1 = __fmaf_rn(__fmaf_rn(-q1, p.sqrt_d, x1), p.rcp_sqrt_d, q1);
                w[j] = pack_bf16x2(q0, q1);
              }
            }
            #pragma unroll
            for (int l = 0; l < kCbLevels; ++l) {
              const int n = 32 >> l;
              if (l > 0) {                              // pool the (lo, hi) pair of every word of the previous level
                #pragma unroll
                for (int j = 0; j < n / 2; ++j) {
                  const uint32_t lo = __byte_perm(w[2 * j], w[2 * j + 1], 0x5410);   // (w0.lo, w1.lo)
                  const uint32_t hi = __byte_perm(w[2 * j], w[2 * j + 1], 0x7632);   // (w0.hi, w1.hi)
                  w[j] = cb_avg_bf16x2(lo, hi);
                }
              }
              if (CB_ABLATE(p) & 1) { if (w[0] == 0x12345678u && l == 4) *reinterpret_cast<uint32_t*>(p.level[0]) = w[0]; continue; }
              if (p.tma_out && l == 0) {                 // level 0: this thread's 64-byte row segment -> the warp's staging box
                if (g == ga) {                           // first chunk of the unit: the previous unit's store has read the box
                  if (lane == 0) bulk_wait_read<0>();
                  __syncwarp();
                }
                #pragma unroll
                for (int j = 0; j < 4; ++j)
                  *reinterpret_cast<uint4*>(wstg + (single ? cb_stage_off(lane, j, 64) : cb_stage_off(lane, (g & 1) * 4 + j, 128))) =
                      make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
              } else if (l <= p.n_levels && row_ok) {
                const int Wl = p.width[l];
                if (l >= 2 && pair_full && p.pair_vec[l]) {
                  // levels 2-4: the unit's two chunks are adjacent halves of one aligned 32 / 16 / 8-byte piece: keep
                  // the first chunk's words and store once (every LSU request of this kernel touches 32 sectors, one
                  // per lane; half-filled sectors made l1tex the busiest unit at 65 %)
                  const int ko = l == 2 ? 0 : (l == 3 ? 4 : 6);
                  if (g == ga) {
                    #pragma unroll
                    for (int j = 0; j < n / 2; ++j) keep[ko + j] = w[j];
                  } else {
                    __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(p.level[l]) + row * Wl + ((64 * ug) >> l);
                    if (l == 2) {
                      const uint32_t w8[8] = {keep[0], keep[1], keep[2], keep[3], w[0], w[1], w[2], w[3]};
                      cb_st256(dst, w8);
                    } else if (l == 3) {
                      *reinterpret_cast<uint4*>(dst) = make_uint4(keep[4], keep[5], w[0], w[1]);
                    } else {
                      *reinterpret_cast<uint2*>(dst) = make_uint2(keep[6], w[0]);
                    }
                  }
                } else {
                  const int col0 = (n0 + cc * 32) >> l;
                  cb_store_words(reinterpret_cast<__nv_bfloat16*>(p.level[l]) + row * Wl + col0, w, n, Wl - col0, p.vec_ok[l]);
                }
              }
            }
            if (p.tma_out && !(CB_ABLATE(p) & 3) && g == gb) {     // unit complete
              fence_proxy_async_smem();                 // staged rows -> visible to the TMA engine
              __syncwarp();
              if (lane == 0) {                          // rows past W1 and columns past W2 are clipped by the TMA unit
                tma_store_3d(&omap.m[single ? 1 : 0], wstg, single ? 32 * g : 64 * ug, m0 + q * 32, bh);
                bulk_commit();
              }
            }
          }
        } else {
        // fp32 pyramid (generic epilogue): unpacked arithmetic, element / vector stores from registers
        float v[32];
        if (p.inv_sqrt_d != 0.f) {                  // sqrt(D) is a power of two: multiplying is exact
          #pragma unroll
          for (int j = 0; j < 32; ++j)
            v[j] = __fmul_rn(__uint_as_float(raw[j]), p.inv_sqrt_d);
        } else {
          #pragma unroll
          for (int j = 0; j < 32; ++j)
            v[j] = __fdiv_rn(__uint_as_float(raw[j]), p.sqrt_d);
        }
        #pragma unroll
        for (int l = 0; l < kCbLevels; ++l) {
          const int n = 32 >> l;                     // compile-time after unrolling
          if (l > 0) {
            #pragma unroll
            for (int j = 0; j < n; ++j)
              v[j] = __fmul_rn(__fadd_rn(v[2 * j], v[2 * j + 1]), 0.5f);
          }
          if (l <= p.n_levels && row_ok) {
            const int Wl = p.width[l];
            const int col0 = (n0 + cc * 32) >> l;
            cb_store_row(static_cast<float*>(p.level[l]) + row * Wl + col0, v, n, Wl - col0, p.vec_ok[l] != 0);
          }
        }
        }
      }
      }
      if (!released) {                              // (a warp without a chunk in this tile, or the ablation paths)
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar_acc_empty[buf]);
      }
    }
    if (p.tma_out && lane == 0) bulk_wait_all();
  }
  __syncthreads();
  if (p.pair) cluster_sync_all();          // no exit while the peer may still signal this CTA's barriers
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
  CorrBuildParams p{};
  p.n_levels = n_levels;
  for (int l = 0; l <= n_levels; ++l) { p.level[l] = levels[l]; p.width[l] = W2 >> l; }
  p.D = D; p.W1 = W1; p.W2 = W2;
  p.n_ntiles = (W2 + kCbMaxN - 1) / kCbMaxN;
  p.n_mtiles = (W1 + kCbM - 1) / kCbM;
  p.H = H;
  p.n_tiles = static_cast<long long>(B) * H * p.n_mtiles * p.n_ntiles;
  p.NT = (((W2 + p.n_ntiles - 1) / p.n_ntiles) + 31) / 32 * 32;
  p.sqrt_d = sqrtf(static_cast<float>(D));
  p.rcp_sqrt_d = 1.f / p.sqrt_d;
  {
    int e = 0;
    const float mant = frexpf(p.sqrt_d, &e);
    p.inv_sqrt_d = (mant == 0.5f && p.sqrt_d * p.sqrt_d == static_cast<float>(D)) ? 1.f / p.sqrt_d : 0.f;
    uint32_t bits;
    memcpy(&bits, &p.inv_sqrt_d, 4);
    p.inv_bf16x2 = (bits >> 16) | (bits & 0xFFFF0000u);       // a power of two is exact in bf16
  }
  for (int l = 0; l <= n_levels; ++l) {
    const size_t esz = sizeof(OutT);
    const size_t chunk = static_cast<size_t>(32 >> l) * esz;
    const size_t need = chunk < 16 ? chunk : 16;
    p.vec_ok[l] = ((static_cast<size_t>(W2 >> l) * esz) % need == 0) &&
                  (reinterpret_cast<uintptr_t>(levels[l]) % 16 == 0);
    if (p.vec_ok[l] && chunk >= 32 && (static_cast<size_t>(W2 >> l) * esz) % 32 == 0 &&
        reinterpret_cast<uintptr_t>(levels[l]) % 32 == 0)
      p.vec_ok[l] = 2;                               // 256-bit stores (bf16 fast path)
    const size_t unit = 2 * chunk;                   // bytes of a 64-column unit at this level
    p.pair_vec[l] = (esz == 2 && l >= 2 && (static_cast<size_t>(W2 >> l) * esz) % unit == 0 &&
                     reinterpret_cast<uintptr_t>(levels[l]) % unit == 0) ? 1 : 0;
  }
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
  // Output path: level 0 of a bf16 pyramid leaves through TMA stores, one [32 rows][128 B] box
  // per epilogue warp and pair of chunks, from that warp's private staging; needs a 16-byte row pitch in global memory.
  CbStoreMaps om{};
  p.tma_out = 0;
  int stage_bytes = 0;
  if (sizeof(OutT) == 2 && !cb_env("PPM_CB_NOTMA")) {
    bool ok = ((static_cast<size_t>(W2) * 2) % 16 == 0) && (reinterpret_cast<uintptr_t>(levels[0]) % 16 == 0);
    if (ok) {
      const uint64_t dims[3] = {(uint64_t)W2, (uint64_t)W1, (uint64_t)B * H};
      const uint64_t str[2] = {(uint64_t)W2 * 2, (uint64_t)W1 * W2 * 2};
      const uint32_t box2[3] = {64, 32, 1}, box1[3] = {32, 32, 1};
      ok = make_tensor_map(&om.m[0], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, levels[0], dims, str, box2, CU_TENSOR_MAP_SWIZZLE_128B) == PPM_OK &&
           make_tensor_map(&om.m[1], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, levels[0], dims, str, box1, CU_TENSOR_MAP_SWIZZLE_64B) == PPM_OK;
    }
    if (ok) {
      // Only level 0 goes through TMA: a TMA store costs ~2.7 clk per box row whatever its width (measured, same box:
      // levels 0-2 / 0-1 / 0 through TMA = 0.536 / 0.522 / 0.513 ms), so the narrow rows of levels 1-4 leave from registers.
      p.tma_out = 1;
      stage_bytes = kCbEpiWarps * kCbWarpStage;          // 32 KB
    }
  }
  // Operand path: when an image row has several M tiles, keep the F2 tile of a work item (image row x
  // N tile) in shared memory for all of them and stream only F1 -- the streaming form re-reads F2 once
  // per M tile and is bound by L2->SM traffic (measured: 3.2 GB of operand reads for 1.2 GB of features).
  // Everything left of the 227 KB goes to the F1 ring: the load path is latency-bound by the bytes in
  // flight (measured at the 720p 1/4-res shape, loads + MMAs only: 0.89 / 0.58 / 0.53 / 0.33 ms with
  // 1 / 2 / 3 / 8 stages of 16 KB).
  const int kSmemMax = 227 * 1024 - 1024;
  const int n_k = (D + Tr::kKC - 1) / Tr::kKC;
  const int b_bytes = n_k * ((p.NT + Tr::kChunkW - 1) / Tr::kChunkW) * Tr::kKC * 128;
  constexpr int kABytes = (kCbM / Tr::kChunkW) * Tr::kKC * 128;
  p.res_b = 0;
  p.n_items = static_cast<long long>(B) * H * p.n_ntiles;
  {
    const int st_max = (kSmemMax - stage_bytes) / kStageBytes;
    p.a_stages = st_max > kCbStages ? kCbStages : st_max;
  }
  int ring_bytes = p.a_stages * kStageBytes;
  if (p.n_mtiles >= 2 && n_k <= kCbMaxRing && b_bytes + 3 * kABytes + stage_bytes <= kSmemMax) {
    p.res_b = 1;
    int a_st = (kSmemMax - b_bytes - stage_bytes) / kABytes;
    p.a_stages = a_st > kCbMaxRing ? kCbMaxRing : a_st;
    if (cb_env("PPM_CB_ASTAGES") && atoi(cb_env("PPM_CB_ASTAGES")) < p.a_stages) p.a_stages = atoi(cb_env("PPM_CB_ASTAGES"));
    ring_bytes = p.a_stages * kABytes + b_bytes;
  }
  const int smem_bytes = ring_bytes + stage_bytes + 1024;
  auto kern = corr_build_kernel<ElemT, OutT>;
  PPM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  // persistent: one CTA per SM (it owns all 512 TMEM columns), static round-robin over the tiles
  int dev = 0, n_sm = 148;
  PPM_CUDA(cudaGetDevice(&dev));
  PPM_CUDA(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
  p.ablate = cb_env("PPM_CB_ABLATE") ? atoi(cb_env("PPM_CB_ABLATE")) : 0;
  p.pair = (p.res_b && p.n_ntiles == 2 && !cb_env("PPM_CB_NO_PAIR")) ? 1 : 0;
  if (p.pair) {
    cudaLaunchConfig_t cfg{};
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.gridDim = dim3(static_cast<unsigned>(n_sm & ~1)); cfg.blockDim = dim3(kCbThreads);
    cfg.dynamicSmemBytes = smem_bytes; cfg.stream = st; cfg.attrs = at; cfg.numAttrs = 1;
    int n_cl = 0;
    PPM_CUDA(cudaOccupancyMaxActiveClusters(&n_cl, kern, &cfg));     // clusters that can be co-resident (persistent kernel)
    const long long rows = p.n_items / 2;
    if (n_cl > rows) n_cl = static_cast<int>(rows);
    if (cb_env("PPM_CB_VERBOSE")) fprintf(stderr, "corr_build: %d co-resident 2-CTA clusters\n", n_cl);
    if (n_cl >= 1) {
      cfg.gridDim = dim3(static_cast<unsigned>(2 * n_cl));
      PPM_CUDA(cudaLaunchKernelEx(&cfg, kern, m1, m2, om, p));
      PPM_CUDA(cudaGetLastError());
      return PPM_OK;
    }
    p.pair = 0;
  }
  const long long units = p.res_b ? p.n_items : p.n_tiles;
  const int grid = static_cast<int>(units < n_sm ? units : n_sm);
  kern<<<grid, kCbThreads, smem_bytes, st>>>(m1, m2, om, p);
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
