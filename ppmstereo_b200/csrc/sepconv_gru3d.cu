// K5: SKSepConvGRU3D -- the recurrent unit of SequenceUpdateBlock3D.
// Replaces SKSepConvGRU3D.forward (reference models/core/ppmtereo_update.py:254-312), called once per update iteration
// from SequenceUpdateBlock3D.forward (:971-1003).  Three gated stages, each a set of 1-D convolutions over ONE axis of a
// (b, C, T, H, W) tensor (W with 15- and 5-tap kernels, then H and T with 5 taps), C = 128 hidden + 384 input channels:
//     z = sigmoid(convz(cat[h, x]));  r = sigmoid(convr(cat[h, x]));  q = tanh(convq(cat[r*h, x]));  h = (1-z)*h + z*q
// 8.85 MFLOP per position (10.4 TFLOP per call at 720p 1/4 res x 20 frames): tensor-core bound.  The reference runs nine
// cuDNN conv3d calls on NCDHW bf16 tensors plus ~25 elementwise launches (measured: 16 % of the bf16 peak).
//
// Formulation: a 1-D convolution is a sum of shifted GEMMs,  out[p, :] = bias + sum_d  A[p + d*e_axis, :] . W_d^T,
// with A the activations in a CHANNELS-LAST bf16 layout [position][channel] (K-major operand rows) and W_d the
// [Cout][Cin] slice of the kernel at tap d.  One persistent tcgen05 kernel runs every convolution of the unit:
//   * M tile = 128 positions = a (bw x bh) patch of one frame (bw * bh = 128, bw in {16, 32, 64} chosen to waste the
//     fewest columns of W); its operand rows come from a 5-D TMA box [64 channels][bw][bh][1][1] of the activation
//     tensor (dims C, W, H, T, b).  The shift of tap d is a COORDINATE offset of the box along the convolution's axis and
//     the zero padding of the convolution is TMA's out-of-bounds zero fill -- W, H and T convolutions are the same code;
//   * K loop = taps x (Cin / 64) steps; a step stages [128 positions][64 ch] of A (16 KB) and [N][64 ch] of W_d
//     (N = 256: the z and r convolutions of a stage share their input and run as ONE GEMM with N = 256) through a
//     4-stage mbarrier ring; 4 tcgen05.mma (K = 16) per step into a 128 x N fp32 accumulator in TMEM, double-buffered
//     (2 x 256 columns) so that the epilogue of tile i overlaps the MMAs of tile i + 1;
//   * the gate arithmetic is fused into the epilogues (8 warps: TMEM lane = position, 32 channels per tcgen05.ld), with
//     the reference's autocast rounding: every conv result, gelu / sigmoid / tanh, r*h, 1-z, (1-z)*h, z*q and their sum
//     are bf16 tensors in the reference (conv3d is an autocast-to-bf16 op and elementwise ops of bf16 tensors return
//     bf16) -- each is rounded once here, in the same order.  A conv result is TWO roundings, bf16(bf16(acc) + bias):
//     cuDNN stores the bf16 convolution and torch adds the (bf16) bias as a separate step -- identified on the box
//     (scripts/exp/gru_rounding_probe.py: 99.8-99.97 % of cuDNN's outputs reproduced, vs 73 % for a fused fp32 bias add;
//     the gate chain, sigmoid, gelu (erf form) and tanh reproduce torch's bf16 results 100 %).
// A operand channels come from up to two tensors (h or r*h: 128 channels; x: 384 channels), so x is packed once per call.
#include <math.h>
#include <mutex>

#include "../../include/ppm_b200.h"
#include "common.cuh"

namespace ppm {

constexpr int kGrEpiWarps = 16;                 // 4 TMEM lane quarters x 4 column groups
constexpr int kGrThreads = 64 + 32 * kGrEpiWarps;   // TMA warp, MMA warp, epilogue warps (576: <= 112 registers each)
constexpr int kGrM = 128;
constexpr int kGrKC = 64;                       // channels per K step (128 B of bf16)
constexpr int kGrABytes = kGrM * kGrKC * 2;     // 16 KB
constexpr int kGrStages = 4;
constexpr int kGrTmemCols = 512;
constexpr int kGrHidden = 128;

enum GruEpilogue : int {
  GR_EPI_GELU = 0,     // out[p, n] = bf16(gelu(c)), c = bf16(bf16(acc) + bias)                     -> dst0 [P, N]
  GR_EPI_Z = 1,        // z = bf16(sigmoid(bf16(acc + bias)))                           -> dst0 [P, 128]
  GR_EPI_R = 2,        // r likewise; rh = bf16(r * h)                                   -> dst1 [P, 128]
  GR_EPI_ZR = 3,       // N = 256: columns 0-127 as GR_EPI_Z, 128-255 as GR_EPI_R
  GR_EPI_Q = 4,        // q = bf16(tanh(bf16(acc + bias))); h = bf16(bf16(bf16(1-z)*h) + bf16(z*q)) -> hbuf (+ NCTHW out)
  GR_EPI_RELU = 5,     // out = relu(c)   -> dst0 [P, dst_ld] at channel dst_c0 (channels-last bf16) and / or NCTHW `out`
  GR_EPI_LINEAR = 6,   // out = c         -> likewise          (c = bf16(bf16(acc) + bias) in every mode)
  GR_EPI_SIGMOID = 7,  // out = bf16(sigmoid(c))  -> likewise
  GR_EPI_RAW32 = 8     // out = acc (fp32, no bias, no rounding) -> strided `out` only (partial sums for a later pass)
};

struct GruGemmParams {
  const float* bias;                 // [N] (bf16-representable values)
  __nv_bfloat16* dst0;               // GELU: G [P, N]; Z / ZR: Z [P, 128]
  __nv_bfloat16* dst1;               // R / ZR: RH [P, 128]
  __nv_bfloat16* hbuf;               // [P, 128] bf16(hidden state): the conv operand (rewritten by Q)
  float* hstate;                     // [P, 128] the hidden state itself, fp32 storage (read by R / ZR / Q, rewritten by Q)
  int h_fp32;                        // 1: the caller's h is fp32 -> products / sums with h are fp32 ops (see below)
  const __nv_bfloat16* zbuf;         // Q: Z [P, 128]
  void* out;                         // Q (last stage) / RELU / LINEAR: strided (b, C, T, H, W) result, or nullptr:
  long long out_sb, out_sc, out_st;  //   element (bb, ch, t, y, x) at bb*out_sb + ch*out_sc + t*out_st + y*W + x
  float out_scale;                   // LINEAR: results multiplied by this (a power of two: exact) before the store
  void* out2;                        // RELU / LINEAR: channels >= split_c go here instead (channel index - split_c)
  long long out2_sb, out2_sc, out2_st;
  int out2_dtype, split_c;
  int out_dtype;
  int N;                             // GEMM N: 64, 128, 192 or 256 (output channels padded with zero weights)
  int n_valid;                       // RELU / LINEAR: output channels actually stored (<= N)
  int dst_ld, dst_c0;                // RELU / LINEAR: row pitch (channels) and channel offset of dst0 (free concatenation)
  int kw, kh, kt;                    // kernel footprint; tap d = (it*kh + ih)*kw + iw, offsets centred ((k-1)/2 padding)
  int seg0_chunks, seg0_c0;          // A channels: seg0_chunks 64-channel chunks of map_a0 starting at channel seg0_c0,
  int seg1_chunks;                   //             then seg1_chunks chunks of map_a1 starting at channel 0
  int b, T, H, W, bw, bh;
  int tiles_x, tiles_y;
  int t_fastest;                     // tile order: 0 = x, y, t, b;  1 = t, x, y, b (T-axis convolutions: the tap window
                                     // slides over frames that are still in L2 -- ncu: 6.3 GB -> DRAM reads of a 5-tap T conv)
  long long n_tiles;
  int mode;
};

__device__ __forceinline__ void tma_load_5d(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2,
                                            int c3, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
      ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3),
      "r"(c4)
      : "memory");
}

__device__ __forceinline__ float gr_ld(const float* p) { return __ldg(p); }
__device__ __forceinline__ float gr_ld(const __nv_bfloat16* p) { return __bfloat162float(__ldg(p)); }

// ------------------------------------------------------------------------------------------------ packing
// (b, C, T, H, W) strided source (pixels contiguous) -> channels-last bf16 [P][dst_c] at channel offset c_off.
// CTA = 32 pixels x 64 channels, transposed through shared memory; reads are 128-byte rows of pixels, writes are
// 128-byte rows of channels.
template <typename SrcT>
__global__ void __launch_bounds__(256) gru_pack_kernel(const SrcT* __restrict__ src, long long sb, long long sc,
                                                       long long st, int T, int HW, int C, int dst_c, int c_off,
                                                       __nv_bfloat16* __restrict__ dst, float* __restrict__ dst32) {
  __shared__ float tile[64][33];
  const int p0 = blockIdx.x * 32, c0 = blockIdx.y * 64, frame = blockIdx.z;
  const int bb = frame / T, t = frame - bb * T;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const SrcT* base = src + bb * sb + t * st + p0;
  #pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int c = k * 8 + w;
    tile[c][lane] = (p0 + lane < HW && c0 + c < C) ? gr_ld(base + static_cast<long long>(c0 + c) * sc + lane) : 0.f;
  }
  __syncthreads();
  const int px = threadIdx.x >> 3, cq = (threadIdx.x & 7) * 8;       // 32 pixels x 8 groups of 8 channels (16 B)
  if (p0 + px < HW && c0 + cq < C) {
    uint4 v;
    v.x = pack_bf16x2(tile[cq + 0][px], tile[cq + 1][px]);
    v.y = pack_bf16x2(tile[cq + 2][px], tile[cq + 3][px]);
    v.z = pack_bf16x2(tile[cq + 4][px], tile[cq + 5][px]);
    v.w = pack_bf16x2(tile[cq + 6][px], tile[cq + 7][px]);
    __nv_bfloat16* d = dst + (static_cast<long long>(frame) * HW + p0 + px) * dst_c + c_off + c0 + cq;
    *reinterpret_cast<uint4*>(d) = v;
    if (dst32) {                                           // unrounded copy (the fp32 hidden state)
      float* d32 = dst32 + (static_cast<long long>(frame) * HW + p0 + px) * dst_c + c_off + c0 + cq;
      *reinterpret_cast<float4*>(d32) = make_float4(tile[cq + 0][px], tile[cq + 1][px], tile[cq + 2][px], tile[cq + 3][px]);
      *reinterpret_cast<float4*>(d32 + 4) = make_float4(tile[cq + 4][px], tile[cq + 5][px], tile[cq + 6][px], tile[cq + 7][px]);
    }
  }
}

// 7 x 7 neighbourhood of the two flow channels as a K-major GEMM operand: row p = pixel, column k = c*49 + ih*7 + iw
// (the order of Conv2d.weight.reshape(Cout, 98)), columns 98..127 zero; values bf16(flow) (autocast's cast of the conv
// input), zero outside the image (the conv's padding).  One thread = one pixel x 8 columns (a 16-byte store).
__global__ void __launch_bounds__(256) flow_im2col_kernel(const float* __restrict__ flow, int H, int W, long long P,
                                                          __nv_bfloat16* __restrict__ dst) {
  const long long i = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x;
  const long long pix = i >> 4;
  if (pix >= P) return;
  const int g = static_cast<int>(i & 15);
  const int HW = H * W;
  const int frame = static_cast<int>(pix / HW), r = static_cast<int>(pix - static_cast<long long>(frame) * HW);
  const int y = r / W, x = r - y * W;
  const float* f = flow + static_cast<long long>(frame) * 2 * HW;
  float v[8];
  #pragma unroll
  for (int e = 0; e < 8; ++e) {
    const int k = g * 8 + e;
    float val = 0.f;
    if (k < 98) {
      const int c = k >= 49, kk = k - c * 49, ih = kk / 7, iw = kk - ih * 7;
      const int yy = y + ih - 3, xx = x + iw - 3;
      if (yy >= 0 && yy < H && xx >= 0 && xx < W) val = __ldg(f + c * HW + yy * W + xx);
    }
    v[e] = val;
  }
  uint4 o;
  o.x = pack_bf16x2(v[0], v[1]); o.y = pack_bf16x2(v[2], v[3]); o.z = pack_bf16x2(v[4], v[5]); o.w = pack_bf16x2(v[6], v[7]);
  *reinterpret_cast<uint4*>(dst + pix * 128 + g * 8) = o;
}

// ------------------------------------------------------------------------------------------------ epilogue math
// fp32 evaluation + one rounding, as torch's CUDA elementwise kernels do for bf16 tensors (opmath = float).
__device__ __forceinline__ float gr_gelu(float x) { return 0.5f * x * (1.f + erff(x * 0.70710678118654752440f)); }
__device__ __forceinline__ float gr_sigmoid(float x) { return 1.f / (1.f + expf(-x)); }

__device__ __forceinline__ void gr_unpack8(const uint4& v, float (&f)[8]) {
  f[0] = __uint_as_float(v.x << 16); f[1] = __uint_as_float(v.x & 0xFFFF0000u);
  f[2] = __uint_as_float(v.y << 16); f[3] = __uint_as_float(v.y & 0xFFFF0000u);
  f[4] = __uint_as_float(v.z << 16); f[5] = __uint_as_float(v.z & 0xFFFF0000u);
  f[6] = __uint_as_float(v.w << 16); f[7] = __uint_as_float(v.w & 0xFFFF0000u);
}
__device__ __forceinline__ uint4 gr_pack8(const float (&f)[8]) {
  uint4 v;
  v.x = pack_bf16x2(f[0], f[1]); v.y = pack_bf16x2(f[2], f[3]);
  v.z = pack_bf16x2(f[4], f[5]); v.w = pack_bf16x2(f[6], f[7]);
  return v;
}

// ------------------------------------------------------------------------------------------------ the GEMM
__device__ __forceinline__ void gr_decode_tile(const GruGemmParams& p, long long tile, int& bx, int& by, int& t, int& bb) {
  long long r = tile;
  if (p.t_fastest) { t = static_cast<int>(r % p.T); r /= p.T; }
  bx = static_cast<int>(r % p.tiles_x); r /= p.tiles_x;
  by = static_cast<int>(r % p.tiles_y); r /= p.tiles_y;
  if (!p.t_fastest) { t = static_cast<int>(r % p.T); r /= p.T; }
  bb = static_cast<int>(r);
}

__global__ void __launch_bounds__(kGrThreads, 1)
gru_gemm_kernel(const __grid_constant__ CUtensorMap map_a0, const __grid_constant__ CUtensorMap map_a1,
                const __grid_constant__ CUtensorMap map_w, const GruGemmParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar_full[kGrStages], bar_empty[kGrStages], bar_acc_full[2], bar_acc_empty[2];
  __shared__ uint32_t tmem_slot;
  __shared__ float s_bias[256];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x < p.N) s_bias[threadIdx.x] = __ldg(p.bias + threadIdx.x);
  const int b_bytes = p.N * 128;
  const int stage_bytes = kGrABytes + b_bytes;
  const int n_chunks = p.seg0_chunks + p.seg1_chunks;
  const int taps = p.kw * p.kh * p.kt;
  const int n_steps = taps * n_chunks;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kGrStages; ++s) { mbar_init(&bar_full[s], 1); mbar_init(&bar_empty[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&bar_acc_full[a], 1); mbar_init(&bar_acc_empty[a], kGrEpiWarps); }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<kGrTmemCols>(&tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------------------------ TMA producer
    if (elect_one()) {
      tma_prefetch_desc(&map_a0); tma_prefetch_desc(&map_a1); tma_prefetch_desc(&map_w);
      int stage = 0;
      uint32_t ph = 1;                                    // parity to wait for on bar_empty[stage]
      for (long long tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
        int bx, by, t, bb;
        gr_decode_tile(p, tile, bx, by, t, bb);
        for (int d = 0; d < taps; ++d) {
          const int iw = d % p.kw, ih = (d / p.kw) % p.kh, it = d / (p.kw * p.kh);
          const int cx = bx * p.bw + iw - (p.kw - 1) / 2;
          const int cy = by * p.bh + ih - (p.kh - 1) / 2;
          const int ct = t + it - (p.kt - 1) / 2;
          for (int kc = 0; kc < n_chunks; ++kc) {
            mbar_wait(&bar_empty[stage], ph);
            mbar_expect_tx(&bar_full[stage], stage_bytes);
            uint8_t* dst = smem + stage * stage_bytes;
            if (kc < p.seg0_chunks) tma_load_5d(dst, &map_a0, &bar_full[stage], p.seg0_c0 + kc * kGrKC, cx, cy, ct, bb);
            else tma_load_5d(dst, &map_a1, &bar_full[stage], (kc - p.seg0_chunks) * kGrKC, cx, cy, ct, bb);
            tma_load_3d(dst + kGrABytes, &map_w, &bar_full[stage], kc * kGrKC, 0, d);
            if (++stage == kGrStages) { stage = 0; ph ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------------------ MMA issuer
    const bool leader = elect_one();
    const uint32_t idesc = umma_idesc(1, 0, 0, kGrM, p.N);
    const uint32_t desc_hi = static_cast<uint32_t>(umma_desc_sw128(0, 16, 1024) >> 32);
    const uint32_t desc_lo0 = static_cast<uint32_t>(umma_desc_sw128(0, 16, 1024));
    const uint32_t base_lo = desc_lo0 + (smem_u32(smem) >> 4);
    auto mk = [&](uint32_t lo) -> uint64_t {
      uint64_t d;
      asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"(desc_hi));
      return d;
    };
    int stage = 0;
    uint32_t ph = 0;
    int i = 0;
    for (long long tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++i) {
      const int buf = i & 1, use = i >> 1;
      mbar_wait(&bar_acc_empty[buf], (use & 1) ^ 1);      // epilogue has drained this accumulator
      tc_fence_after();
      const uint32_t d_tmem = tmem + buf * 256;
      for (int step = 0; step < n_steps; ++step) {
        mbar_wait(&bar_full[stage], ph);
        tc_fence_after();
        if (leader) {
          const uint32_t a_lo = base_lo + ((stage * stage_bytes) >> 4);
          const uint32_t b_lo = a_lo + (kGrABytes >> 4);
          #pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_f16_ss(d_tmem, mk(a_lo + 2 * k), mk(b_lo + 2 * k), idesc, (step | k) != 0);
          umma_commit(&bar_empty[stage]);                 // frees the ring stage once these MMAs have read it
        }
        __syncwarp();
        if (++stage == kGrStages) { stage = 0; ph ^= 1; }
      }
      if (leader) umma_commit(&bar_acc_full[buf]);
      __syncwarp();
    }
  } else {
    // ------------------------------------------------------------------------------ epilogue (16 warps)
    // The K = 2560 GEMMs are paced by this code, not by the MMAs (sigmoid / tanh / erf evaluated exactly, state and gate
    // loads): four warps per scheduler, each owning a quarter of the N columns in 16-column pieces, hide that latency.
    const int quarter = warp & 3;                         // TMEM lanes 32*quarter .. +31 (hardware rule: warp_id % 4)
    const int half = (warp - 2) >> 2;                     // which quarter of the N columns
    const int row = quarter * 32 + lane;
    const int cols_per_warp = p.N >> 2;                   // 16, 32, 48 or 64
    const int HW = p.H * p.W;
    int i = 0;
    for (long long tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++i) {
      const int buf = i & 1, use = i >> 1;
      int bx, by, t, bb;
      gr_decode_tile(p, tile, bx, by, t, bb);
      const int x = bx * p.bw + row % p.bw, y = by * p.bh + row / p.bw;
      const bool valid = x < p.W && y < p.H;
      const long long pos = (static_cast<long long>(bb) * p.T + t) * HW + static_cast<long long>(y) * p.W + x;
      mbar_wait(&bar_acc_full[buf], use & 1);
      tc_fence_after();
      const uint32_t t_addr = tmem + (static_cast<uint32_t>(quarter * 32) << 16) + buf * 256 + half * cols_per_warp;
      const int n_chunks_e = cols_per_warp >> 4;
      for (int c = 0; c < n_chunks_e; ++c) {
        uint32_t acc[16];
        tmem_ld_x16(t_addr + c * 16, acc);
        const int col0 = half * cols_per_warp + c * 16;   // column of acc[0] in the GEMM's N
        int mode = p.mode, ch0 = col0;                    // ch0: channel inside the 128-wide gate tensors
        if (mode == GR_EPI_ZR) { mode = col0 < kGrHidden ? GR_EPI_Z : GR_EPI_R; ch0 = col0 & (kGrHidden - 1); }
        // the chunk's state / gate operands are requested before the accumulator wait
        float4 hreg[4];
        uint4 zreg[2];
        if (valid && (mode == GR_EPI_R || mode == GR_EPI_Q)) {
          const float4* hp = reinterpret_cast<const float4*>(p.hstate + pos * kGrHidden + ch0);
          #pragma unroll
          for (int e = 0; e < 4; ++e) hreg[e] = hp[e];
          if (mode == GR_EPI_Q) {
            const uint4* zp = reinterpret_cast<const uint4*>(p.zbuf + pos * kGrHidden + ch0);
            #pragma unroll
            for (int e = 0; e < 2; ++e) zreg[e] = zp[e];
          }
        }
        tmem_ld_wait();
        if (c == n_chunks_e - 1) {                        // this warp's last read of the accumulator
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&bar_acc_empty[buf]);
        }
        if (!valid) continue;
        #pragma unroll
        for (int g = 0; g < 2; ++g) {                     // 2 groups of 8 channels = one 16-byte store each
          float v[8];
          #pragma unroll
          for (int e = 0; e < 8; ++e)
            v[e] = mode == GR_EPI_RAW32 ? __uint_as_float(acc[g * 8 + e])
                                        : bf16_round_f(bf16_round_f(__uint_as_float(acc[g * 8 + e])) + s_bias[col0 + g * 8 + e]);
          const int ch = ch0 + g * 8;
          if (mode == GR_EPI_GELU) {
            #pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = gr_gelu(v[e]);
            *reinterpret_cast<uint4*>(p.dst0 + pos * p.N + col0 + g * 8) = gr_pack8(v);
          } else if (mode == GR_EPI_Z) {
            #pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = gr_sigmoid(v[e]);
            *reinterpret_cast<uint4*>(p.dst0 + pos * kGrHidden + ch) = gr_pack8(v);
          } else if (mode == GR_EPI_R) {
            // r * h: a bf16 op for a bf16 h; for an fp32 h an fp32 product that the next conv's autocast cast rounds to
            // bf16 -- fp32(r*h) then bf16 in both cases (the fp32 product of two bf16 values is exact)
            const float4 h0 = hreg[2 * g], h1 = hreg[2 * g + 1];
            const float h[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
            #pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = __fmul_rn(bf16_round_f(gr_sigmoid(v[e])), h[e]);
            *reinterpret_cast<uint4*>(p.dst1 + pos * kGrHidden + ch) = gr_pack8(v);
          } else if (mode == GR_EPI_RELU || mode == GR_EPI_LINEAR || mode == GR_EPI_SIGMOID || mode == GR_EPI_RAW32) {
            if (col0 + g * 8 < p.n_valid) {
              if (mode == GR_EPI_RELU) {
                #pragma unroll
                for (int e = 0; e < 8; ++e) v[e] = fmaxf(v[e], 0.f);
              } else if (mode == GR_EPI_SIGMOID) {
                #pragma unroll
                for (int e = 0; e < 8; ++e) v[e] = bf16_round_f(gr_sigmoid(v[e]));
              } else if (mode == GR_EPI_LINEAR) {
                #pragma unroll
                for (int e = 0; e < 8; ++e) v[e] *= p.out_scale;
              }
              if (p.dst0) {
                if (col0 + g * 8 + 8 <= p.n_valid) {
                  *reinterpret_cast<uint4*>(p.dst0 + pos * p.dst_ld + p.dst_c0 + col0 + g * 8) = gr_pack8(v);
                } else {
                  #pragma unroll
                  for (int e = 0; e < 8; ++e)
                    if (col0 + g * 8 + e < p.n_valid)
                      p.dst0[pos * p.dst_ld + p.dst_c0 + col0 + g * 8 + e] = __float2bfloat16_rn(v[e]);
                }
              }
              if (p.out) {
                const long long sp = static_cast<long long>(y) * p.W + x;
                #pragma unroll
                for (int e = 0; e < 8; ++e) {
                  const int col = col0 + g * 8 + e;
                  if (col < p.n_valid) {
                    if (col < p.split_c) {
                      const long long o = bb * p.out_sb + col * p.out_sc + t * p.out_st + sp;
                      if (p.out_dtype == PPM_BF16) static_cast<__nv_bfloat16*>(p.out)[o] = __float2bfloat16_rn(v[e]);
                      else static_cast<float*>(p.out)[o] = v[e];
                    } else {
                      const long long o = bb * p.out2_sb + (col - p.split_c) * p.out2_sc + t * p.out2_st + sp;
                      if (p.out2_dtype == PPM_BF16) static_cast<__nv_bfloat16*>(p.out2)[o] = __float2bfloat16_rn(v[e]);
                      else static_cast<float*>(p.out2)[o] = v[e];
                    }
                  }
                }
              }
            }
          } else {                                        // GR_EPI_Q
            float z[8];
            float* hs = p.hstate + pos * kGrHidden + ch;
            const float4 h0 = hreg[2 * g], h1 = hreg[2 * g + 1];
            const float h[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
            gr_unpack8(zreg[g], z);
            #pragma unroll
            for (int e = 0; e < 8; ++e) {
              // (1 - z) * h + z * q: 1 - z and z * q are bf16 ops (z, q bf16).  bf16 h: the product and the sum are bf16
              // ops too.  fp32 h (what the model hands over: the SST blocks return fp32): type promotion makes both fp32.
              const float q = bf16_round_f(tanhf(v[e]));
              float a = __fmul_rn(bf16_round_f(1.f - z[e]), h[e]);
              const float bq = bf16_round_f(z[e] * q);
              if (!p.h_fp32) a = bf16_round_f(a);
              float hn = __fadd_rn(a, bq);
              if (!p.h_fp32) hn = bf16_round_f(hn);
              v[e] = hn;
            }
            *reinterpret_cast<float4*>(hs) = make_float4(v[0], v[1], v[2], v[3]);
            *reinterpret_cast<float4*>(hs + 4) = make_float4(v[4], v[5], v[6], v[7]);
            *reinterpret_cast<uint4*>(p.hbuf + pos * kGrHidden + ch) = gr_pack8(v);
            if (p.out) {
              const long long o = bb * p.out_sb + ch * p.out_sc + t * p.out_st + static_cast<long long>(y) * p.W + x;
              const long long cs = p.out_sc;
              if (p.out_dtype == PPM_BF16) {
                __nv_bfloat16* ob = static_cast<__nv_bfloat16*>(p.out) + o;
                #pragma unroll
                for (int e = 0; e < 8; ++e) ob[e * cs] = __float2bfloat16_rn(v[e]);
              } else {
                float* of = static_cast<float*>(p.out) + o;
                #pragma unroll
                for (int e = 0; e < 8; ++e) of[e * cs] = v[e];
              }
            }
          }
        }
      }
    }
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<kGrTmemCols>(tmem);
  }
}

// ------------------------------------------------------------------------------------------------ flow head, second layer
// FlowHead3D.conv2 is 256 -> 2 channels over a 3 x 3 x 3 footprint: as a shifted GEMM it would be 27 x 4 K-steps on an N
// tile (64 is the narrowest) that is 97 % padding.  Instead the contraction over the 256 channels runs ONCE per position for
// all 27 x 2 (offset, output) pairs -- a plain K = 256, N = 54 GEMM whose fp32 results are stored as 54 planes --
//   Y[tap*2 + c][p] = sum_ci W2[c, ci, tap] * hd[p, ci]
// and this kernel adds the 27 shifted planes per output: out[p, c] = bf16(bf16(sum_tap Y[tap*2 + c][p + off(tap)]) + b[c]),
// positions outside the volume contributing nothing (the convolution's zero padding).  fp32 throughout until cuDNN's two
// roundings; the planes are read along x, coalesced.
__global__ void __launch_bounds__(256) flow_head_gather_kernel(const float* __restrict__ Y, const float* __restrict__ bias,
                                                               int b, int T, int H, int W, void* __restrict__ out,
                                                               long long out_sb, long long out_sc, long long out_st,
                                                               int out_dtype) {
  const long long P = static_cast<long long>(b) * T * H * W;
  const long long p = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x;
  if (p >= P) return;
  const int x = static_cast<int>(p % W);
  long long r = p / W;
  const int y = static_cast<int>(r % H); r /= H;
  const int t = static_cast<int>(r % T);
  const int bb = static_cast<int>(r / T);
  float s0 = 0.f, s1 = 0.f;
  #pragma unroll
  for (int it = 0; it < 3; ++it) {
    const int tt = t + it - 1;
    if (tt < 0 || tt >= T) continue;
    #pragma unroll
    for (int ih = 0; ih < 3; ++ih) {
      const int yy = y + ih - 1;
      if (yy < 0 || yy >= H) continue;
      #pragma unroll
      for (int iw = 0; iw < 3; ++iw) {
        const int xx = x + iw - 1;
        if (xx < 0 || xx >= W) continue;
        const int tap = (it * 3 + ih) * 3 + iw;
        const long long q = ((static_cast<long long>(bb) * T + tt) * H + yy) * W + xx;
        s0 = __fadd_rn(s0, __ldg(Y + (2 * tap) * P + q));
        s1 = __fadd_rn(s1, __ldg(Y + (2 * tap + 1) * P + q));
      }
    }
  }
  const float o0 = bf16_round_f(bf16_round_f(s0) + __ldg(bias));
  const float o1 = bf16_round_f(bf16_round_f(s1) + __ldg(bias + 1));
  const long long o = bb * out_sb + t * out_st + static_cast<long long>(y) * W + x;
  if (out_dtype == PPM_BF16) {
    static_cast<__nv_bfloat16*>(out)[o] = __float2bfloat16_rn(o0);
    static_cast<__nv_bfloat16*>(out)[o + out_sc] = __float2bfloat16_rn(o1);
  } else {
    static_cast<float*>(out)[o] = o0;
    static_cast<float*>(out)[o + out_sc] = o1;
  }
}

// ------------------------------------------------------------------------------------------------ host
struct GruShape {
  int b, T, H, W, bw, bh, tiles_x, tiles_y;
  long long P, n_tiles;
};

static GruShape gru_shape(int b, int T, int H, int W) {
  GruShape s{};
  s.b = b; s.T = T; s.H = H; s.W = W;
  int best = 64;
  long long best_cols = -1;
  for (int bw = 64; bw >= 16; bw >>= 1) {                 // fewest wasted columns; ties -> the wider box
    const long long cols = static_cast<long long>((W + bw - 1) / bw) * bw;
    if (best_cols < 0 || cols < best_cols) { best = bw; best_cols = cols; }
  }
  s.bw = best; s.bh = kGrM / best;
  s.tiles_x = (W + s.bw - 1) / s.bw;
  s.tiles_y = (H + s.bh - 1) / s.bh;
  s.P = static_cast<long long>(b) * T * H * W;
  s.n_tiles = static_cast<long long>(s.tiles_x) * s.tiles_y * T * b;
  return s;
}

static int gru_act_map(CUtensorMap* m, const void* base, int C, const GruShape& s) {
  const uint64_t dims[5] = {(uint64_t)C, (uint64_t)s.W, (uint64_t)s.H, (uint64_t)s.T, (uint64_t)s.b};
  const uint64_t e = static_cast<uint64_t>(C) * 2;
  const uint64_t str[4] = {e, e * s.W, e * s.W * s.H, e * s.W * s.H * s.T};
  const uint32_t box[5] = {(uint32_t)kGrKC, (uint32_t)s.bw, (uint32_t)s.bh, 1, 1};
  return make_tensor_map(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, base, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B);
}

static int gru_w_map(CUtensorMap* m, const void* base, int Cin, int N, int taps) {
  const uint64_t dims[3] = {(uint64_t)Cin, (uint64_t)N, (uint64_t)taps};
  const uint64_t str[2] = {(uint64_t)Cin * 2, (uint64_t)Cin * 2 * N};
  const uint32_t box[3] = {(uint32_t)kGrKC, (uint32_t)N, 1};
  return make_tensor_map(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, base, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B);
}

static int gru_launch(const CUtensorMap& a0, const CUtensorMap& a1, const CUtensorMap& w, GruGemmParams p,
                      const GruShape& s, cudaStream_t st) {
  p.b = s.b; p.T = s.T; p.H = s.H; p.W = s.W; p.bw = s.bw; p.bh = s.bh;
  p.tiles_x = s.tiles_x; p.tiles_y = s.tiles_y; p.n_tiles = s.n_tiles;
  p.t_fastest = p.kt > 1 && p.kw == 1 && p.kh == 1 ? 1 : 0;
  if (p.out && p.out_sc == 0) {                           // default: contiguous (b, C, T, H, W)
    const long long HW = static_cast<long long>(s.H) * s.W;
    const int C = (p.mode == GR_EPI_Q) ? kGrHidden : p.n_valid;
    p.out_st = HW; p.out_sc = HW * s.T; p.out_sb = HW * s.T * C;
  }
  if (p.out_scale == 0.f) p.out_scale = 1.f;
  if (!p.out2) p.split_c = 1 << 30;
  const int smem = kGrStages * (kGrABytes + p.N * 128) + 1024;
  static std::mutex mu;
  static bool attr_done[64] = {};
  int dev = 0, n_sm = 148;
  PPM_CUDA(cudaGetDevice(&dev));
  {
    std::lock_guard<std::mutex> lock(mu);
    if (dev < 0 || dev >= 64 || !attr_done[dev]) {
      PPM_CUDA(cudaFuncSetAttribute(gru_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    kGrStages * (kGrABytes + 256 * 128) + 1024));
      if (dev >= 0 && dev < 64) attr_done[dev] = true;
    }
  }
  PPM_CUDA(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
  const int grid = static_cast<int>(s.n_tiles < n_sm ? s.n_tiles : n_sm);
  gru_gemm_kernel<<<grid, kGrThreads, smem, st>>>(a0, a1, w, p);
  PPM_CUDA(cudaGetLastError());
  return PPM_OK;
}

template <typename SrcT>
static int gru_pack(const void* src, long long sb, long long sc, long long st_, const GruShape& s, int C, int dst_c,
                    int c_off, __nv_bfloat16* dst, float* dst32, cudaStream_t st) {
  const int HW = s.H * s.W;
  dim3 grid((HW + 31) / 32, (C + 63) / 64, s.b * s.T);
  gru_pack_kernel<SrcT><<<grid, 256, 0, st>>>(static_cast<const SrcT*>(src), sb, sc, st_, s.T, HW, C, dst_c, c_off, dst, dst32);
  PPM_CUDA(cudaGetLastError());
  return PPM_OK;
}

// The eight GEMMs of one unit on packed operands (X, Hb, Hs filled by the caller).  out strides 0 = contiguous (b,128,T,H,W).
static int gru_unit(const GruShape& s, __nv_bfloat16* X, __nv_bfloat16* Hb, __nv_bfloat16* RH, __nv_bfloat16* Z,
                    __nv_bfloat16* G, float* Hs, const void* const* weights, const float* const* biases, int h_fp32,
                    void* out, int out_dtype, long long out_sb, long long out_sc, long long out_st, cudaStream_t st) {
  int rc;
  CUtensorMap mX, mH, mRH, mG, mW[8];
  if ((rc = gru_act_map(&mX, X, 384, s))) return rc;
  if ((rc = gru_act_map(&mH, Hb, 128, s))) return rc;
  if ((rc = gru_act_map(&mRH, RH, 128, s))) return rc;
  if ((rc = gru_act_map(&mG, G, 256, s))) return rc;
  const int w_cin[8] = {512, 128, 128, 512, 512, 512, 512, 512};
  const int w_n[8] = {256, 128, 128, 128, 256, 128, 256, 128};
  const int w_taps[8] = {15, 5, 5, 5, 5, 5, 5, 5};
  for (int i = 0; i < 8; ++i)
    if ((rc = gru_w_map(&mW[i], weights[i], w_cin[i], w_n[i], w_taps[i]))) return rc;

  GruGemmParams p{};
  p.hbuf = Hb; p.hstate = Hs; p.h_fp32 = h_fp32; p.zbuf = Z; p.out = nullptr; p.out_dtype = out_dtype;
  // ---- stage 1: along W.  z / r = Conv(15) -> GELU -> Conv(5) (ppmtereo_update.py:257-266)
  p.bias = biases[0]; p.N = 256; p.kw = 15; p.kh = 1; p.kt = 1; p.mode = GR_EPI_GELU;
  p.seg0_chunks = 2; p.seg0_c0 = 0; p.seg1_chunks = 6; p.dst0 = G; p.dst1 = nullptr;
  if ((rc = gru_launch(mH, mX, mW[0], p, s, st))) return rc;
  p.bias = biases[1]; p.N = 128; p.kw = 5; p.mode = GR_EPI_Z; p.seg0_chunks = 2; p.seg0_c0 = 0; p.seg1_chunks = 0; p.dst0 = Z;
  if ((rc = gru_launch(mG, mG, mW[1], p, s, st))) return rc;
  p.bias = biases[2]; p.mode = GR_EPI_R; p.seg0_c0 = 128; p.dst0 = nullptr; p.dst1 = RH;
  if ((rc = gru_launch(mG, mG, mW[2], p, s, st))) return rc;
  p.bias = biases[3]; p.mode = GR_EPI_Q; p.seg0_c0 = 0; p.seg0_chunks = 2; p.seg1_chunks = 6;
  if ((rc = gru_launch(mRH, mX, mW[3], p, s, st))) return rc;
  // ---- stages 2 (along H) and 3 (along T): z | r as one N = 256 GEMM, then q
  for (int stage = 2; stage <= 3; ++stage) {
    const int wi = stage == 2 ? 4 : 6;
    p.kw = 1; p.kh = stage == 2 ? 5 : 1; p.kt = stage == 3 ? 5 : 1;
    p.bias = biases[wi]; p.N = 256; p.mode = GR_EPI_ZR; p.seg0_chunks = 2; p.seg0_c0 = 0; p.seg1_chunks = 6;
    p.dst0 = Z; p.dst1 = RH; p.out = nullptr;
    if ((rc = gru_launch(mH, mX, mW[wi], p, s, st))) return rc;
    p.bias = biases[wi + 1]; p.N = 128; p.mode = GR_EPI_Q; p.out = stage == 3 ? out : nullptr;
    p.out_sb = out_sb; p.out_sc = out_sc; p.out_st = out_st;
    if ((rc = gru_launch(mRH, mX, mW[wi + 1], p, s, st))) return rc;
  }
  return PPM_OK;
}

// Flow head on a channels-last bf16 state `a_in` [P,128]; Hd = [P,256] scratch.  out strides 0 = contiguous (b,2,T,H,W).
static int flow_head_unit(const GruShape& s, const __nv_bfloat16* a_in, __nv_bfloat16* Hd, float* Y, const void* w1,
                          const float* b1, const void* w2, const float* b2, void* out, int out_dtype, long long out_sb,
                          long long out_sc, long long out_st, cudaStream_t st) {
  int rc;
  CUtensorMap mA, mH, mW1, mW2;
  if ((rc = gru_act_map(&mA, a_in, 128, s))) return rc;
  if ((rc = gru_act_map(&mH, Hd, 256, s))) return rc;
  if ((rc = gru_w_map(&mW1, w1, 128, 256, 27))) return rc;
  if ((rc = gru_w_map(&mW2, w2, 256, 64, 1))) return rc;
  GruGemmParams p{};
  p.kw = 3; p.kh = 3; p.kt = 3;
  p.bias = b1; p.N = 256; p.n_valid = 256; p.mode = GR_EPI_RELU; p.seg0_chunks = 2; p.seg0_c0 = 0; p.seg1_chunks = 0;
  p.dst0 = Hd; p.dst_ld = 256; p.dst_c0 = 0; p.out = nullptr;
  if ((rc = gru_launch(mA, mA, mW1, p, s, st))) return rc;
  // conv2 (256 -> 2, 3 x 3 x 3): one K = 256 GEMM for all 27 x 2 (offset, output) pairs -> 54 fp32 planes, then the gather-sum
  const long long HW = static_cast<long long>(s.H) * s.W;
  p.kw = 1; p.kh = 1; p.kt = 1;
  p.bias = b2; p.N = 64; p.n_valid = 54; p.mode = GR_EPI_RAW32; p.seg0_chunks = 4; p.dst0 = nullptr;
  p.out = Y; p.out_dtype = PPM_F32; p.out_sc = s.P; p.out_sb = HW * s.T; p.out_st = HW;     // planes: Y[j][(b, t, y, x)]
  if ((rc = gru_launch(mH, mH, mW2, p, s, st))) return rc;
  if (out_sc == 0) { out_st = HW; out_sc = HW * s.T; out_sb = HW * s.T * 2; }               // contiguous (b, 2, T, H, W)
  flow_head_gather_kernel<<<static_cast<unsigned>((s.P + 255) / 256), 256, 0, st>>>(Y, b2, s.b, s.T, s.H, s.W, out, out_sb,
                                                                                    out_sc, out_st, out_dtype);
  PPM_CUDA(cudaGetLastError());
  return PPM_OK;
}

// mask_2d head (ppmtereo_update.py:~905, used at :996): 0.25 * Conv1x1(relu(Conv3x3(net))), 128 -> 256 -> 144, per frame
// (a 3 x 3 x 1 footprint).  w2 / b2 padded to N = 192.
static int mask_head_unit(const GruShape& s, const __nv_bfloat16* a_in, __nv_bfloat16* Hd, const void* w1, const float* b1,
                          const void* w2, const float* b2, int n_mask, void* out, int out_dtype, long long out_sb,
                          long long out_sc, long long out_st, cudaStream_t st) {
  int rc;
  CUtensorMap mA, mH, mW1, mW2;
  if ((rc = gru_act_map(&mA, a_in, 128, s))) return rc;
  if ((rc = gru_act_map(&mH, Hd, 256, s))) return rc;
  if ((rc = gru_w_map(&mW1, w1, 128, 256, 9))) return rc;
  if ((rc = gru_w_map(&mW2, w2, 256, 192, 1))) return rc;
  GruGemmParams p{};
  p.kw = 3; p.kh = 3; p.kt = 1; p.out_dtype = out_dtype;
  p.bias = b1; p.N = 256; p.n_valid = 256; p.mode = GR_EPI_RELU; p.seg0_chunks = 2; p.seg0_c0 = 0; p.seg1_chunks = 0;
  p.dst0 = Hd; p.dst_ld = 256; p.dst_c0 = 0; p.out = nullptr;
  if ((rc = gru_launch(mA, mA, mW1, p, s, st))) return rc;
  p.kw = 1; p.kh = 1;
  p.bias = b2; p.N = 192; p.n_valid = n_mask; p.mode = GR_EPI_LINEAR; p.seg0_chunks = 4; p.dst0 = nullptr; p.out = out;
  p.out_scale = 0.25f; p.out_sb = out_sb; p.out_sc = out_sc; p.out_st = out_st;
  return gru_launch(mH, mH, mW2, p, s, st);
}

}  // namespace ppm

extern "C" int ppm_sepconv_gru3d_workspace(int b, int T, int H, int W, size_t* bytes) {
  PPM_CHECK(b > 0 && T > 0 && H > 0 && W > 0 && bytes, PPM_ERR_ARG, "ppm_sepconv_gru3d_workspace: bad argument");
  const size_t P = static_cast<size_t>(b) * T * H * W;
  *bytes = P * ((384 + 128 + 128 + 128 + 256) * 2 + 128 * 4);   // X, H, RH, Z, G (bf16) + hidden state (fp32), channels-last
  return PPM_OK;
}

extern "C" int ppm_sepconv_gru3d(const void* h, int h_dtype, int64_t hsb, int64_t hsc, int64_t hst, const void* x,
                                 int x_dtype, int64_t xsb, int64_t xsc, int64_t xst, const void* const* weights,
                                 const float* const* biases, int b, int T, int H, int W, void* workspace,
                                 size_t workspace_bytes, void* out, int out_dtype, void* stream) {
  using namespace ppm;
  PPM_CHECK(h && x && weights && biases && workspace && out, PPM_ERR_ARG, "ppm_sepconv_gru3d: null pointer");
  PPM_CHECK(b > 0 && T > 0 && H > 0 && W > 0, PPM_ERR_ARG, "ppm_sepconv_gru3d: bad shape");
  PPM_CHECK(b <= 65535 && T <= 65535 && static_cast<long long>(b) * T <= 65535, PPM_ERR_UNSUPPORTED,
            "ppm_sepconv_gru3d: b*T > 65535");
  PPM_CHECK((h_dtype == PPM_F32 || h_dtype == PPM_BF16) && (x_dtype == PPM_F32 || x_dtype == PPM_BF16) &&
                (out_dtype == PPM_F32 || out_dtype == PPM_BF16),
            PPM_ERR_ARG, "ppm_sepconv_gru3d: bad dtype");
  size_t need = 0;
  ppm_sepconv_gru3d_workspace(b, T, H, W, &need);
  PPM_CHECK(workspace_bytes >= need, PPM_ERR_ARG, "ppm_sepconv_gru3d: workspace of %zu bytes < %zu needed", workspace_bytes, need);
  PPM_CHECK((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, PPM_ERR_ARG, "ppm_sepconv_gru3d: workspace must be 256-byte aligned");
  for (int i = 0; i < 8; ++i)
    PPM_CHECK(weights[i] && biases[i] && (reinterpret_cast<uintptr_t>(weights[i]) & 15) == 0, PPM_ERR_ARG,
              "ppm_sepconv_gru3d: weight / bias %d null or misaligned", i);
  const GruShape s = gru_shape(b, T, H, W);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  __nv_bfloat16* X = static_cast<__nv_bfloat16*>(workspace);
  __nv_bfloat16* Hb = X + s.P * 384;
  __nv_bfloat16* RH = Hb + s.P * 128;
  __nv_bfloat16* Z = RH + s.P * 128;
  __nv_bfloat16* G = Z + s.P * 128;
  float* Hs = reinterpret_cast<float*>(G + s.P * 256);
  int rc;
  // channels-last bf16 operands (the cast conv3d's autocast performs on its input)
  if (x_dtype == PPM_BF16) rc = gru_pack<__nv_bfloat16>(x, xsb, xsc, xst, s, 384, 384, 0, X, nullptr, st);
  else rc = gru_pack<float>(x, xsb, xsc, xst, s, 384, 384, 0, X, nullptr, st);
  if (rc) return rc;
  if (h_dtype == PPM_BF16) rc = gru_pack<__nv_bfloat16>(h, hsb, hsc, hst, s, 128, 128, 0, Hb, Hs, st);
  else rc = gru_pack<float>(h, hsb, hsc, hst, s, 128, 128, 0, Hb, Hs, st);
  if (rc) return rc;
  return gru_unit(s, X, Hb, RH, Z, G, Hs, weights, biases, h_dtype == PPM_F32 ? 1 : 0, out, out_dtype, 0, 0, 0, st);
}

// ------------------------------------------------------------------------------------------------ FlowHead3D
// FlowHead3D.forward (reference models/core/ppmtereo_update.py:670-678): conv2(relu(conv1(x))), two 3x3x3 Conv3d
// (128 -> 256 -> 2), on the SAME shifted-GEMM kernel: a 3x3x3 footprint is 27 coordinate offsets of the TMA box.
// conv2's two output channels are padded to N = 64 with zero weight rows (the GEMM is bound by its A traffic, not by MMA).
extern "C" int ppm_sepconv_gru3d_state_offset(int b, int T, int H, int W, size_t* offset) {
  PPM_CHECK(b > 0 && T > 0 && H > 0 && W > 0 && offset, PPM_ERR_ARG, "ppm_sepconv_gru3d_state_offset: bad argument");
  *offset = static_cast<size_t>(b) * T * H * W * 384 * 2;       // the bf16 channels-last state [P, 128] follows X
  return PPM_OK;
}

extern "C" int ppm_flow_head3d_workspace(int b, int T, int H, int W, size_t* bytes) {
  PPM_CHECK(b > 0 && T > 0 && H > 0 && W > 0 && bytes, PPM_ERR_ARG, "ppm_flow_head3d_workspace: bad argument");
  *bytes = static_cast<size_t>(b) * T * H * W * (128 + 256) * 2;
  return PPM_OK;
}

extern "C" int ppm_flow_head3d(const void* x, int x_dtype, int64_t xsb, int64_t xsc, int64_t xst, const void* x_cl,
                               const void* w1, const float* b1, const void* w2, const float* b2, int b, int T, int H,
                               int W, void* workspace, size_t workspace_bytes, void* out, int out_dtype, void* stream) {
  using namespace ppm;
  PPM_CHECK((x || x_cl) && w1 && b1 && w2 && b2 && workspace && out, PPM_ERR_ARG, "ppm_flow_head3d: null pointer");
  PPM_CHECK(b > 0 && T > 0 && H > 0 && W > 0 && static_cast<long long>(b) * T <= 65535, PPM_ERR_ARG, "ppm_flow_head3d: bad shape");
  PPM_CHECK((x_cl || x_dtype == PPM_F32 || x_dtype == PPM_BF16) && (out_dtype == PPM_F32 || out_dtype == PPM_BF16), PPM_ERR_ARG,
            "ppm_flow_head3d: bad dtype");
  size_t need = 0;
  ppm_flow_head3d_workspace(b, T, H, W, &need);
  PPM_CHECK(workspace_bytes >= need, PPM_ERR_ARG, "ppm_flow_head3d: workspace of %zu bytes < %zu needed", workspace_bytes, need);
  PPM_CHECK((reinterpret_cast<uintptr_t>(workspace) & 255) == 0 && (reinterpret_cast<uintptr_t>(x_cl) & 15) == 0 &&
                (reinterpret_cast<uintptr_t>(w1) & 15) == 0 && (reinterpret_cast<uintptr_t>(w2) & 15) == 0,
            PPM_ERR_ARG, "ppm_flow_head3d: misaligned pointer");
  const GruShape s = gru_shape(b, T, H, W);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  __nv_bfloat16* Xc = static_cast<__nv_bfloat16*>(workspace);
  __nv_bfloat16* Hd = Xc + s.P * 128;
  int rc;
  const __nv_bfloat16* a_in = static_cast<const __nv_bfloat16*>(x_cl);
  if (!a_in) {
    if (x_dtype == PPM_BF16) rc = gru_pack<__nv_bfloat16>(x, xsb, xsc, xst, s, 128, 128, 0, Xc, nullptr, st);
    else rc = gru_pack<float>(x, xsb, xsc, xst, s, 128, 128, 0, Xc, nullptr, st);
    if (rc) return rc;
    a_in = Xc;
  }
  // the 54 fp32 planes (256 B per position) reuse the Xc region: conv1 has consumed it by the time conv2's GEMM writes
  return flow_head_unit(s, a_in, Hd, reinterpret_cast<float*>(Xc), w1, b1, w2, b2, out, out_dtype, 0, 0, 0, st);
}

// ------------------------------------------------------------------------------------------------ whole update block
// SequenceUpdateBlock3D.forward (reference models/core/ppmtereo_update.py:971-1003) for a block WITHOUT update-time /
// update-space attention (update_block08 / update_block04 of PPMStereo) and with the 2-D convex mask, as one call:
//   inp_tensor = cat[inp, motion_features, motion_features_global]       -> never materialised: the three tensors are
//                                                                           packed straight into the GRU's operand
//   net = gru(net, inp_tensor); delta_flow = flow_head(net); mask = 0.25 * mask_2d(net)
// All tensors frame-major (b*T, C, H, W) contiguous, as the reference holds them outside the unit; results are written
// frame-major as well, so neither the `rearrange` copies around the unit nor the 384-channel cat exist.
extern "C" int ppm_update_block3d_workspace(int b, int T, int H, int W, size_t* bytes) {
  size_t g = 0;
  if (int rc = ppm_sepconv_gru3d_workspace(b, T, H, W, &g)) return rc;
  PPM_CHECK(bytes, PPM_ERR_ARG, "ppm_update_block3d_workspace: null pointer");
  *bytes = g + static_cast<size_t>(b) * T * H * W * 256 * 2;
  return PPM_OK;
}

extern "C" int ppm_update_block3d(const void* net, int net_dtype, const void* const* x_parts, const int* x_dtypes,
                                  const void* const* gru_weights, const float* const* gru_biases, const void* flow_w1,
                                  const float* flow_b1, const void* flow_w2, const float* flow_b2, const void* mask_w1,
                                  const float* mask_b1, const void* mask_w2, const float* mask_b2, int n_mask, int b, int T,
                                  int H, int W, void* workspace, size_t workspace_bytes, void* net_out, void* mask_out,
                                  void* dflow_out, void* stream) {
  using namespace ppm;
  PPM_CHECK(net && x_parts && x_dtypes && gru_weights && gru_biases && flow_w1 && flow_b1 && flow_w2 && flow_b2 && mask_w1 &&
                mask_b1 && mask_w2 && mask_b2 && workspace && net_out && mask_out && dflow_out,
            PPM_ERR_ARG, "ppm_update_block3d: null pointer");
  PPM_CHECK(b > 0 && T > 0 && H > 0 && W > 0 && static_cast<long long>(b) * T <= 65535, PPM_ERR_ARG, "ppm_update_block3d: bad shape");
  PPM_CHECK(n_mask > 0 && n_mask <= 192, PPM_ERR_UNSUPPORTED, "ppm_update_block3d: n_mask = %d not in (0, 192]", n_mask);
  PPM_CHECK(net_dtype == PPM_F32 || net_dtype == PPM_BF16, PPM_ERR_ARG, "ppm_update_block3d: bad dtype");
  for (int i = 0; i < 3; ++i)
    PPM_CHECK(x_parts[i] && (x_dtypes[i] == PPM_F32 || x_dtypes[i] == PPM_BF16), PPM_ERR_ARG,
              "ppm_update_block3d: input part %d null or of a bad dtype", i);
  for (int i = 0; i < 8; ++i)
    PPM_CHECK(gru_weights[i] && gru_biases[i] && (reinterpret_cast<uintptr_t>(gru_weights[i]) & 15) == 0, PPM_ERR_ARG,
              "ppm_update_block3d: GRU weight / bias %d null or misaligned", i);
  size_t need = 0;
  ppm_update_block3d_workspace(b, T, H, W, &need);
  PPM_CHECK(workspace_bytes >= need, PPM_ERR_ARG, "ppm_update_block3d: workspace of %zu bytes < %zu needed", workspace_bytes, need);
  PPM_CHECK((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, PPM_ERR_ARG, "ppm_update_block3d: workspace must be 256-byte aligned");
  const GruShape s = gru_shape(b, T, H, W);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long HW = static_cast<long long>(H) * W;
  __nv_bfloat16* X = static_cast<__nv_bfloat16*>(workspace);
  __nv_bfloat16* Hb = X + s.P * 384;
  __nv_bfloat16* RH = Hb + s.P * 128;
  __nv_bfloat16* Z = RH + s.P * 128;
  __nv_bfloat16* G = Z + s.P * 128;
  float* Hs = reinterpret_cast<float*>(G + s.P * 256);
  __nv_bfloat16* Hd = reinterpret_cast<__nv_bfloat16*>(Hs + s.P * 128);
  int rc;
  // frame-major (b*T, 128, H, W) seen as (b, 128, T, H, W): sb = T*128*HW, sc = HW, st = 128*HW
  const long long sb = T * 128 * HW, sc = HW, stt = 128 * HW;
  for (int i = 0; i < 3; ++i) {
    if (x_dtypes[i] == PPM_BF16) rc = gru_pack<__nv_bfloat16>(x_parts[i], sb, sc, stt, s, 128, 384, 128 * i, X, nullptr, st);
    else rc = gru_pack<float>(x_parts[i], sb, sc, stt, s, 128, 384, 128 * i, X, nullptr, st);
    if (rc) return rc;
  }
  if (net_dtype == PPM_BF16) rc = gru_pack<__nv_bfloat16>(net, sb, sc, stt, s, 128, 128, 0, Hb, Hs, st);
  else rc = gru_pack<float>(net, sb, sc, stt, s, 128, 128, 0, Hb, Hs, st);
  if (rc) return rc;
  if ((rc = gru_unit(s, X, Hb, RH, Z, G, Hs, gru_weights, gru_biases, net_dtype == PPM_F32 ? 1 : 0, net_out, net_dtype, sb, sc,
                     stt, st)))
    return rc;
  // (the flow head's 54 fp32 planes go where the GRU's packed input was: the GRU is done with it)
  if ((rc = flow_head_unit(s, Hb, Hd, reinterpret_cast<float*>(X), flow_w1, flow_b1, flow_w2, flow_b2, dflow_out, PPM_BF16,
                           T * 2 * HW, HW, 2 * HW, st)))
    return rc;
  return mask_head_unit(s, Hb, Hd, mask_w1, mask_b1, mask_w2, mask_b2, n_mask, mask_out, PPM_BF16,
                        static_cast<long long>(T) * n_mask * HW, HW, static_cast<long long>(n_mask) * HW, st);
}

// ------------------------------------------------------------------------------------------------ motion encoder (K7)
// BasicMotionEncoder_v2.forward (reference models/core/ppmtereo_update.py:458-482) with motion_hidden_state given
// (init_conv, first iteration only, stays with the caller):
//   flo1 = relu(convf1(flow))             7x7, 2 -> 128         im2col pack + one K = 128 GEMM
//   cor = gelu(convc1(lookup))            K6 up to the 54-channel hidden vector + a K = 64 GEMM with the GELU epilogue
//   cor = relu(convc2(cor))               3x3, 256 -> 192       shifted GEMM, 9 offsets  -> cat[:,   0:192]
//   flo = relu(convf2(flo1))              3x3, 128 ->  64       -> cat[:, 192:256]
//   cat[:, 256:320] = motion_hidden_state (packed)              -- torch.cat([cor, flo, mhs]) is never built separately:
//   out = relu(final_conv(cat))           3x3, 320 -> 190          the three producers write one channels-last buffer
//   motion_features[:, 0:126] = out[:, :126] (fp32: the reference's cat with the fp32 flow promotes),  mhs = out[:, 126:190]
extern "C" int ppm_motion_encoder_workspace(int B, int H, int W, size_t* bytes) {
  PPM_CHECK(B > 0 && H > 0 && W > 0 && bytes, PPM_ERR_ARG, "ppm_motion_encoder_workspace: bad argument");
  const size_t P = static_cast<size_t>(B) * H * W;
  *bytes = P * (36 * 4 + (256 + 128 + 320) * 2);          // x2 (fp32), cor, flo1, cat (bf16 channels-last)
  return PPM_OK;
}

extern "C" int ppm_motion_encoder(const void* level0, const float* flow, const float* corr, int W2, const float* front_a,
                                  const float* front_b, const void* wd, const void* wf1, const float* bf1, const void* mhs,
                                  int mhs_dtype, const void* wc2, const float* bc2, const void* wf2, const float* bf2, const void* wfin,
                                  const float* bfin, int B, int H, int W, void* workspace, size_t workspace_bytes,
                                  float* motion_features, void* mhs_out, void* stream) {
  using namespace ppm;
  PPM_CHECK(flow && front_b && wd && wf1 && bf1 && mhs && wc2 && bc2 && wf2 && bf2 && wfin && bfin && workspace && motion_features && mhs_out,
            PPM_ERR_ARG, "ppm_motion_encoder: null pointer");
  PPM_CHECK(B > 0 && H > 0 && W > 0 && B <= 65535, PPM_ERR_ARG, "ppm_motion_encoder: bad shape");
  PPM_CHECK(mhs_dtype == PPM_F32 || mhs_dtype == PPM_BF16, PPM_ERR_ARG, "ppm_motion_encoder: bad dtype");
  size_t need = 0;
  ppm_motion_encoder_workspace(B, H, W, &need);
  PPM_CHECK(workspace_bytes >= need, PPM_ERR_ARG, "ppm_motion_encoder: workspace of %zu bytes < %zu needed", workspace_bytes, need);
  PPM_CHECK((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, PPM_ERR_ARG, "ppm_motion_encoder: workspace must be 256-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const GruShape s = gru_shape(1, B, H, W);                 // B frames of one "batch": 2-D convolutions, kt = 1
  const long long HW = static_cast<long long>(H) * W;
  __nv_bfloat16* cor = static_cast<__nv_bfloat16*>(workspace);
  __nv_bfloat16* flo = cor + s.P * 256;
  __nv_bfloat16* cat = flo + s.P * 128;
  float* x2 = reinterpret_cast<float*>(cat + s.P * 320);
  __nv_bfloat16* col = cor;                                 // the flow branch runs first: its im2col operand borrows cor
  int rc;
  CUtensorMap mCor, mFlo, mCat, mCol, mW;
  if ((rc = gru_act_map(&mCor, cor, 256, s))) return rc;
  if ((rc = gru_act_map(&mFlo, flo, 128, s))) return rc;
  if ((rc = gru_act_map(&mCat, cat, 320, s))) return rc;
  if ((rc = gru_act_map(&mCol, col, 128, s))) return rc;
  GruGemmParams p{};
  p.mode = GR_EPI_RELU; p.seg0_c0 = 0; p.seg1_chunks = 0;
  // flo = relu(convf1(flow)): 7x7 on 2 channels = one K = 98 (padded 128) GEMM on the im2col operand
  {
    const long long threads = s.P * 16;
    flow_im2col_kernel<<<static_cast<unsigned>((threads + 255) / 256), 256, 0, st>>>(flow, H, W, s.P, col);
    PPM_CUDA(cudaGetLastError());
  }
  if ((rc = gru_w_map(&mW, wf1, 128, 128, 1))) return rc;
  p.kw = 1; p.kh = 1; p.kt = 1; p.bias = bf1; p.N = 128; p.n_valid = 128; p.seg0_chunks = 2;
  p.dst0 = flo; p.dst_ld = 128; p.dst_c0 = 0;
  if ((rc = gru_launch(mCol, mCol, mW, p, s, st))) return rc;
  // convf2: 128 -> 64 -> cat[:, 192:256]
  p.kw = 3; p.kh = 3; p.dst0 = cat; p.dst_ld = 320;
  if ((rc = gru_w_map(&mW, wf2, 128, 64, 9))) return rc;
  p.bias = bf2; p.N = 64; p.n_valid = 64; p.seg0_chunks = 2; p.dst_c0 = 192;
  if ((rc = gru_launch(mFlo, mFlo, mW, p, s, st))) return rc;
  // frame-major (B, C, H, W) seen as (1, C, B, H, W): sc = HW, st = C*HW
  if (mhs_dtype == PPM_BF16) rc = gru_pack<__nv_bfloat16>(mhs, 0, HW, 64 * HW, s, 64, 320, 256, cat, nullptr, st);
  else rc = gru_pack<float>(mhs, 0, HW, 64 * HW, s, 64, 320, 256, cat, nullptr, st);
  if (rc) return rc;
  // cor = gelu(convc1(lookup)): K6 up to the block's 54-channel hidden vector ([P,64], in the flo buffer, free by now), then
  // its last layer (54 -> 256) + the outer GELU as one K = 64 GEMM -> cor channels-last
  __nv_bfloat16* hid = flo;
  if ((rc = motion_front_launch(level0, level0 ? flow : nullptr, corr, front_a, front_b, B, H, W, W2, x2, hid, 2, st))) return rc;
  CUtensorMap mHid;
  if ((rc = gru_act_map(&mHid, hid, 64, s))) return rc;
  if ((rc = gru_w_map(&mW, wd, 64, 256, 1))) return rc;
  p.kw = 1; p.kh = 1; p.mode = GR_EPI_GELU; p.bias = front_b + kMotionFrontBiasD; p.N = 256; p.n_valid = 256; p.seg0_chunks = 1;
  p.dst0 = cor; p.dst_ld = 256; p.dst_c0 = 0;
  if ((rc = gru_launch(mHid, mHid, mW, p, s, st))) return rc;
  // convc2: 256 -> 192 -> cat[:, 0:192]
  p.kw = 3; p.kh = 3; p.mode = GR_EPI_RELU; p.dst0 = cat; p.dst_ld = 320;
  if ((rc = gru_w_map(&mW, wc2, 256, 192, 9))) return rc;
  p.bias = bc2; p.N = 192; p.n_valid = 192; p.seg0_chunks = 4; p.dst_c0 = 0;
  if ((rc = gru_launch(mCor, mCor, mW, p, s, st))) return rc;
  // final_conv: 320 -> 190, split into motion_features[:, :126] (fp32) and the new hidden state (bf16)
  if ((rc = gru_w_map(&mW, wfin, 320, 192, 9))) return rc;
  p.bias = bfin; p.N = 192; p.n_valid = 190; p.seg0_chunks = 5; p.dst0 = nullptr;
  p.out = motion_features; p.out_dtype = PPM_F32; p.out_sb = 0; p.out_sc = HW; p.out_st = 128 * HW;
  p.out2 = mhs_out; p.out2_dtype = PPM_BF16; p.out2_sb = 0; p.out2_sc = HW; p.out2_st = 64 * HW; p.split_c = 126;
  return gru_launch(mCat, mCat, mW, p, s, st);
}

// ------------------------------------------------------------------------------------------------ value + uncertainty (K8)
// The two lines of the update iteration between the motion encoder and the memory read (reference
// models/core/ppmstereo.py:492-495 -> ppmtereo_update.py:945-950 and :936-938):
//   value = aggregator.to_v(motion_features)                       1x1, 128 -> 128, no bias
//   uncertainty = sigmoid(conv1x1(relu(conv3x3(cat[net, value]))))  256 -> 128 -> 1
// value is written twice by its GEMM's epilogue: NCHW bf16 (what the memory read takes in place as its V operand) and
// channels-last (the second A segment of the 3x3 convolution: the cat is never built).
extern "C" int ppm_value_uncertainty_workspace(int B, int H, int W, size_t* bytes) {
  PPM_CHECK(B > 0 && H > 0 && W > 0 && bytes, PPM_ERR_ARG, "ppm_value_uncertainty_workspace: bad argument");
  *bytes = static_cast<size_t>(B) * H * W * (4 * 128 * 2);          // motion features, net, value, hidden: [P,128] bf16 each
  return PPM_OK;
}

extern "C" int ppm_value_uncertainty(const void* motion_features, int mf_dtype, const void* net, int net_dtype,
                                     const void* wv, const float* bv, const void* wu1, const float* bu1, const void* wu2,
                                     const float* bu2, int B, int H, int W, void* workspace, size_t workspace_bytes,
                                     void* value, void* uncertainty, void* stream) {
  using namespace ppm;
  PPM_CHECK(motion_features && net && wv && bv && wu1 && bu1 && wu2 && bu2 && workspace && value && uncertainty, PPM_ERR_ARG,
            "ppm_value_uncertainty: null pointer");
  PPM_CHECK(B > 0 && H > 0 && W > 0 && B <= 65535, PPM_ERR_ARG, "ppm_value_uncertainty: bad shape");
  PPM_CHECK((mf_dtype == PPM_F32 || mf_dtype == PPM_BF16) && (net_dtype == PPM_F32 || net_dtype == PPM_BF16), PPM_ERR_ARG,
            "ppm_value_uncertainty: bad dtype");
  size_t need = 0;
  ppm_value_uncertainty_workspace(B, H, W, &need);
  PPM_CHECK(workspace_bytes >= need, PPM_ERR_ARG, "ppm_value_uncertainty: workspace of %zu bytes < %zu needed", workspace_bytes, need);
  PPM_CHECK((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, PPM_ERR_ARG, "ppm_value_uncertainty: workspace must be 256-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const GruShape s = gru_shape(1, B, H, W);
  const long long HW = static_cast<long long>(H) * W;
  __nv_bfloat16* mfc = static_cast<__nv_bfloat16*>(workspace);
  __nv_bfloat16* netc = mfc + s.P * 128;
  __nv_bfloat16* valc = netc + s.P * 128;
  __nv_bfloat16* hid = valc + s.P * 128;
  int rc;
  if (mf_dtype == PPM_BF16) rc = gru_pack<__nv_bfloat16>(motion_features, 0, HW, 128 * HW, s, 128, 128, 0, mfc, nullptr, st);
  else rc = gru_pack<float>(motion_features, 0, HW, 128 * HW, s, 128, 128, 0, mfc, nullptr, st);
  if (rc) return rc;
  if (net_dtype == PPM_BF16) rc = gru_pack<__nv_bfloat16>(net, 0, HW, 128 * HW, s, 128, 128, 0, netc, nullptr, st);
  else rc = gru_pack<float>(net, 0, HW, 128 * HW, s, 128, 128, 0, netc, nullptr, st);
  if (rc) return rc;
  CUtensorMap mMf, mNet, mVal, mHid, mW;
  if ((rc = gru_act_map(&mMf, mfc, 128, s))) return rc;
  if ((rc = gru_act_map(&mNet, netc, 128, s))) return rc;
  if ((rc = gru_act_map(&mVal, valc, 128, s))) return rc;
  if ((rc = gru_act_map(&mHid, hid, 128, s))) return rc;
  GruGemmParams p{};
  p.kt = 1; p.seg0_c0 = 0; p.dst_ld = 128; p.dst_c0 = 0;
  // value = to_v(motion_features)
  if ((rc = gru_w_map(&mW, wv, 128, 128, 1))) return rc;
  p.kw = 1; p.kh = 1; p.mode = GR_EPI_LINEAR; p.bias = bv; p.N = 128; p.n_valid = 128; p.seg0_chunks = 2; p.seg1_chunks = 0;
  p.dst0 = valc; p.out = value; p.out_dtype = PPM_BF16; p.out_sb = 0; p.out_sc = HW; p.out_st = 128 * HW;
  if ((rc = gru_launch(mMf, mMf, mW, p, s, st))) return rc;
  // hidden = relu(conv3x3(cat[net, value]))
  if ((rc = gru_w_map(&mW, wu1, 256, 128, 9))) return rc;
  p.kw = 3; p.kh = 3; p.mode = GR_EPI_RELU; p.bias = bu1; p.seg0_chunks = 2; p.seg1_chunks = 2; p.dst0 = hid; p.out = nullptr;
  if ((rc = gru_launch(mNet, mVal, mW, p, s, st))) return rc;
  // uncertainty = sigmoid(conv1x1(hidden))
  if ((rc = gru_w_map(&mW, wu2, 128, 64, 1))) return rc;
  p.kw = 1; p.kh = 1; p.mode = GR_EPI_SIGMOID; p.bias = bu2; p.N = 64; p.n_valid = 1; p.seg0_chunks = 2; p.seg1_chunks = 0;
  p.dst0 = nullptr; p.out = uncertainty; p.out_dtype = PPM_BF16; p.out_sb = 0; p.out_sc = HW; p.out_st = HW;
  return gru_launch(mHid, mHid, mW, p, s, st);
}
