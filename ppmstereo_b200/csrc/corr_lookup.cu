// K2: radius-r bilinear lookup into the 1-D correlation pyramid (HBM-bound gather).
// Replaces CorrBlock1D.__call__ + bilinear_sampler (reference models/core/corr.py:74-94, :10-27).
//
// Every pixel (b,y,x) owns a private row of each pyramid level, so there is no reuse between
// pixels: the job is to touch each needed 32-byte sector once, keep many loads in flight, and
// write the (B, L*(2r+1), H, W) output with full 128-byte lines -- with as few instructions as
// possible (the first version of this kernel was issue-bound, not memory-bound).
//   * one thread per pixel, all levels unrolled: per level ONE 256-bit load per 32-byte sector
//     that holds taps (LDG.E.256 on sector-aligned addresses; the second sector only when the
//     2r+2 taps spill into it).  Narrower loads cost one L2 request each even when they fall in
//     the same sector: the 4 x 8-byte version of this kernel read 620 MB from HBM for 94 MB of
//     taps; the loads of a pixel are independent and issued back to back;
//   * the window is shifted into place in registers (word select + funnel shift), taps that
//     fall outside the pixel's row are zeroed (grid_sample zeros padding);
//   * lanes are consecutive pixels, so every store writes 128 contiguous bytes of one channel.
// Interpolation: grid_sample(align_corners=True, zeros padding) on a height-1 image, evaluated with the SAME fp32
// operation sequence the reference executes on a GPU (lk_sample below), so that the sample positions -- and with
// them the interpolation weights -- are bit-identical to the reference's, tap by tap.  (Round 1 used the closed form
// out_k = (1-a) V[f+k] + a V[f+k+1]; the reference's normalise / un-normalise round trip moves each tap's position
// by up to an ulp of x, 2e-5 of |out|, and the recurrent update amplifies that to ~0.01 px of final disparity.)
#include <atomic>

#include "../../include/ppm_b200.h"
#include "common.cuh"

namespace ppm {

constexpr int kLookupThreads = 128;
constexpr int kMaxLevels = 4;
constexpr int kMaxTaps = 9;

struct LookupParams {
  const void* level[kMaxLevels];
  int width[kMaxLevels];             // W2 >> l
  long long total[kMaxLevels];       // elements in level l
  const float* flow;                 // (B,2,H,W1)
  float* out;                        // (B, L*(2r+1), H, W1)
  int HW, W1, radius, n_levels;
};

__device__ __forceinline__ float lk_load1(const float* p) { return __ldg(p); }
__device__ __forceinline__ float lk_load1(const __nv_bfloat16* p) { return __bfloat162float(__ldg(p)); }

// One request per 32-byte sector: 256-bit loads (LDG.E.256, sm_100) on sector-aligned addresses.
struct alignas(32) LkSector { uint32_t w[8]; };
__device__ __forceinline__ LkSector lk_ldg256(const void* p) {
  LkSector r;
  asm volatile("ld.global.nc.v8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]), "=r"(r.w[4]), "=r"(r.w[5]),
                 "=r"(r.w[6]), "=r"(r.w[7])
               : "l"(p));
  return r;
}
// Elements the vector path may touch past the sector-aligned start of the window.
__device__ __forceinline__ constexpr int lk_span(const __nv_bfloat16*) { return 32; }
__device__ __forceinline__ constexpr int lk_span(const float*) { return 24; }

// Fetch elements [g, g + 10) of `base` into v[0..9] (as fp32).  The caller guarantees that the
// sector-aligned span [g & ~(spe-1), +lk_span) lies inside the allocation (spe = elements per sector).
__device__ __forceinline__ void lk_window(const __nv_bfloat16* base, long long g, float (&v)[kMaxTaps + 1]) {
  const long long a = g & ~15LL;                         // 16 bf16 per sector
  const int o = static_cast<int>(g - a);                 // 0..15
  uint32_t w[16];
  const LkSector s0 = lk_ldg256(base + a);
  #pragma unroll
  for (int i = 0; i < 8; ++i) w[i] = s0.w[i];
  if (o > 6) {                                           // taps spill into the next sector
    const LkSector s1 = lk_ldg256(base + a + 16);
    #pragma unroll
    for (int i = 0; i < 8; ++i) w[8 + i] = s1.w[i];
  } else {
    #pragma unroll
    for (int i = 0; i < 8; ++i) w[8 + i] = 0u;
  }
  const int wi = o >> 1;                                 // first word, 0..7
  uint32_t s1v[9], s2v[7], s3v[6];
  #pragma unroll
  for (int i = 0; i < 9; ++i) s1v[i] = (wi & 4) ? w[i + 4] : w[i];
  #pragma unroll
  for (int i = 0; i < 7; ++i) s2v[i] = (wi & 2) ? s1v[i + 2] : s1v[i];
  #pragma unroll
  for (int i = 0; i < 6; ++i) s3v[i] = (wi & 1) ? s2v[i + 1] : s2v[i];
  const int sh = (o & 1) * 16;
  #pragma unroll
  for (int i = 0; i < 5; ++i) {
    const uint32_t x = __funnelshift_r(s3v[i], s3v[i + 1], sh);
    v[2 * i] = __uint_as_float(x << 16);
    v[2 * i + 1] = __uint_as_float(x & 0xFFFF0000u);
  }
}
__device__ __forceinline__ void lk_window(const float* base, long long g, float (&v)[kMaxTaps + 1]) {
  const long long a = g & ~7LL;                          // 8 floats per sector
  const int o = static_cast<int>(g - a);                 // 0..7
  uint32_t w[24];
  const LkSector s0 = lk_ldg256(base + a), s1 = lk_ldg256(base + a + 8);
  #pragma unroll
  for (int i = 0; i < 8; ++i) { w[i] = s0.w[i]; w[8 + i] = s1.w[i]; }
  if (o > 6) {
    const LkSector s2 = lk_ldg256(base + a + 16);
    #pragma unroll
    for (int i = 0; i < 8; ++i) w[16 + i] = s2.w[i];
  } else {
    #pragma unroll
    for (int i = 0; i < 8; ++i) w[16 + i] = 0u;
  }
  uint32_t s1v[14], s2v[12];
  #pragma unroll
  for (int i = 0; i < 14; ++i) s1v[i] = (o & 4) ? w[i + 4] : w[i];
  #pragma unroll
  for (int i = 0; i < 12; ++i) s2v[i] = (o & 2) ? s1v[i + 2] : s1v[i];
  #pragma unroll
  for (int i = 0; i < kMaxTaps + 1; ++i) v[i] = __uint_as_float((o & 1) ? s2v[i + 1] : s2v[i]);
}

// One tap of one level, following the reference op by op (fp32, every line one rounding):
//   corr.py:84      x0    = dx + coords / 2**l                      (coords / 2**l is exact)
//   corr.py:14      xgrid = 2 * x0 / (W_l - 1) - 1                  ATen's CUDA div-by-scalar multiplies by the
//                                                                   reciprocal 1.0f / (W_l - 1) (BinaryDivTrueKernel.cu)
//   unnormalise (align_corners):     ix = ((xgrid + 1) / 2) * (W_l - 1)
//   bilinear weights and blend: the reference's F.grid_sample call (4-D input, bilinear, zeros padding,
//   align_corners=True) is dispatched by torch to cuDNN's spatial-transformer sampler, not to ATen's own kernel.
//   cuDNN is closed source; its arithmetic was identified on the B200 box against 11 candidate forms
//   (scripts/exp/gridsample_probe.py, gridsample_probe2.py):
//       i0 = floor(ix);  w0 = (i0 + 1) - ix;  w1 = 1 - w0;  out = fma(V[i0], w0, V[i0 + 1] * w1)
//   reproduces cuDNN bit for bit for every sample with ix >= 0 (400 000 samples, incl. |ix| < 1 where the weights
//   carry fine mantissa bits); for -1 < ix < 0 (only element 0 in range) cuDNN's weight differs from 1 + ix by one
//   ulp in half of the cases -- an unidentified rounding, 0.07 % of random taps, <= 6e-8 relative.  (ATen's own
//   kernel, used when cuDNN is disabled, is w1 = ix - i0 and out = fma(V[i0+1], w1, V[i0] * w0): 100 % identified,
//   but it is not what the reference runs.)  Out-of-row taps contribute 0; non-finite ix -> 0.
// v[0..9] holds the row elements first .. first + 9 (zeroed outside the row), first = floor(x_l) - r; tap k nominally
// reads v[k], v[k+1]; when the rounding of ix moves floor(ix) by one the neighbours are taken instead (at k = 0 /
// k = 2r the element one beyond the window is replaced by the window's edge element: its weight is <= 1 ulp of ix).
template <int K>
__device__ __forceinline__ float lk_sample(const float (&v)[kMaxTaps + 1], int first, float cl, int r, float inv_wm1,
                                           float wm1) {
  const float x0 = __fadd_rn(static_cast<float>(K - r), cl);
  const float xg = __fsub_rn(__fmul_rn(2.f * x0, inv_wm1), 1.f);
  const float ix = __fmul_rn(__fadd_rn(xg, 1.f) * 0.5f, wm1);
  const float fi = floorf(ix);
  const float nw = __fsub_rn(fi + 1.f, ix), ne = __fsub_rn(1.f, nw);
  const float dj = fi - static_cast<float>(first + K);       // 0 nominally; -1 / +1 when ix rounded across an integer
  float lo = v[K], hi = v[K + 1];
  if (dj < 0.f) { hi = v[K]; lo = v[K >= 1 ? K - 1 : 0]; }
  if (dj > 0.f) { lo = v[K + 1]; hi = v[K + 2 <= kMaxTaps ? K + 2 : kMaxTaps]; }
  const float o = __fmaf_rn(lo, nw, __fmul_rn(hi, ne));
  return (fabsf(ix) <= 3.0e9f) ? o : 0.f;                    // NaN / inf / beyond int range: reference yields 0
}

template <typename VolT>
__global__ void __launch_bounds__(kLookupThreads)
corr_lookup_kernel(const LookupParams p) {
  const int b = blockIdx.y;
  const int q = blockIdx.x * kLookupThreads + threadIdx.x;
  if (q >= p.HW) return;
  const int r = p.radius, taps = 2 * r + 1;
  const float x = __fadd_rn(static_cast<float>(q % p.W1), __ldg(p.flow + static_cast<long long>(b) * 2 * p.HW + q));
  const long long pixel = static_cast<long long>(b) * p.HW + q;
  float* dst = p.out + static_cast<long long>(b) * p.n_levels * taps * p.HW + q;

  float v[kMaxLevels][kMaxTaps + 1];
  int first[kMaxLevels];                                   // in-row index of v[l][0]
  // ---- gather: all loads of all levels are issued before anything is consumed
  #pragma unroll
  for (int l = 0; l < kMaxLevels; ++l) {
    if (l < p.n_levels) {
      const int Wl = p.width[l];
      const float xl = x * (1.f / static_cast<float>(1 << l));          // exact: power of two
      float fl = fminf(fmaxf(floorf(xl), -1048576.f), 1048576.f);
      fl = (fl == fl) ? fl : 0.f;
      first[l] = static_cast<int>(fl) - r;
      const long long g = pixel * Wl + first[l];
      const VolT* base = static_cast<const VolT*>(p.level[l]);
      const long long spe = 32 / static_cast<long long>(sizeof(VolT));        // elements per sector
      if (g >= 0 && (g & ~(spe - 1)) + lk_span(base) <= p.total[l]) {
        lk_window(base, g, v[l]);
      } else {                                              // first / last rows of the whole volume
        #pragma unroll
        for (int i = 0; i < kMaxTaps + 1; ++i) {
          const long long e = g + i;
          v[l][i] = (e >= 0 && e < p.total[l]) ? lk_load1(base + e) : 0.f;
        }
      }
    }
  }
  // ---- zero padding, interpolate, store
  #pragma unroll
  for (int l = 0; l < kMaxLevels; ++l) {
    if (l < p.n_levels) {
      const int Wl = p.width[l];
      if (first[l] < 0 || first[l] + 2 * r + 1 >= Wl) {
        #pragma unroll
        for (int i = 0; i < kMaxTaps + 1; ++i)
          if (static_cast<unsigned>(first[l] + i) >= static_cast<unsigned>(Wl)) v[l][i] = 0.f;
      }
      const float wm1 = static_cast<float>(Wl - 1), inv = __frcp_rn(wm1);
      const float cl = x * (1.f / static_cast<float>(1 << l));
      float* d = dst + static_cast<long long>(l) * taps * p.HW;
      float o[kMaxTaps];
      o[0] = lk_sample<0>(v[l], first[l], cl, r, inv, wm1); o[1] = lk_sample<1>(v[l], first[l], cl, r, inv, wm1);
      o[2] = lk_sample<2>(v[l], first[l], cl, r, inv, wm1); o[3] = lk_sample<3>(v[l], first[l], cl, r, inv, wm1);
      o[4] = lk_sample<4>(v[l], first[l], cl, r, inv, wm1); o[5] = lk_sample<5>(v[l], first[l], cl, r, inv, wm1);
      o[6] = lk_sample<6>(v[l], first[l], cl, r, inv, wm1); o[7] = lk_sample<7>(v[l], first[l], cl, r, inv, wm1);
      o[8] = lk_sample<8>(v[l], first[l], cl, r, inv, wm1);
      #pragma unroll
      for (int k = 0; k < kMaxTaps; ++k)
        if (k < taps) d[static_cast<long long>(k) * p.HW] = o[k];
    }
  }
}


// ------------------------------------------------------------------------------------------------
// K2b: the same lookup for a bf16 pyramid, sampling ALL four levels from one span of level 0.
// On B200 a random 32-byte read costs a whole 128-byte line of DRAM traffic (scripts/exp/gather_exp.cu), so
// the four separate windows of the kernel above cost ~4.6 lines per pixel.  The level-3 window of a pixel
// covers 80 consecutive level-0 elements (160 bytes, ~2.2 lines) that contain the windows of levels 0-2 as
// well, so this kernel reads that span once and rebuilds the pooled taps in registers with exactly the
// arithmetic ppm_corr_build uses (bf16(lo/2 + hi/2) per level: HMUL2/HFMA2.BF16).  With f = floor(x),
// the span starts at level-0 index 8*((f>>3) - 4) and the windows start at fixed offsets plus a few low
// bits of f:  level 0: 28 + (f&7), level 1: 12 + ((f>>1)&3), level 2: 4 + ((f>>2)&1), level 3: 0.
// Valid for pyramids whose levels 1-3 ARE the cascaded poolings of level 0 (those built by ppm_corr_build),
// radius 4, W2 % 8 == 0.
__device__ __forceinline__ uint32_t lk_avg_bf16x2(uint32_t w0, uint32_t w1) {       // (avg(w0.lo,w0.hi), avg(w1.lo,w1.hi))
  const uint32_t lo = __byte_perm(w0, w1, 0x5410), hi = __byte_perm(w0, w1, 0x7632);
  const uint32_t half2 = 0x3F003F00u;
  uint32_t hb, d;
  asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(hb) : "r"(hi), "r"(half2));
  asm("fma.rn.bf16x2 %0, %1, %2, %3;" : "=r"(d) : "r"(lo), "r"(half2), "r"(hb));
  return d;
}
// ten taps starting `odd` elements into words u[0..5]
__device__ __forceinline__ void lk_taps(const uint32_t (&u)[6], int odd, float (&v)[kMaxTaps + 1]) {
  const int sh = odd * 16;
  #pragma unroll
  for (int i = 0; i < 5; ++i) {
    const uint32_t x = __funnelshift_r(u[i], u[i + 1], sh);
    v[2 * i] = __uint_as_float(x << 16);
    v[2 * i + 1] = __uint_as_float(x & 0xFFFF0000u);
  }
}

struct LookupL0Params {
  const __nv_bfloat16* level0;
  long long total;                   // elements in level 0
  const float* flow;
  float* out;
  int HW, W1, W2;
};

// The 36 outputs of pixel (b, q), delivered level by level to `sink.put(level, o[9])` (K2b stores them, the fused
// motion-encoder front end K6a keeps them in registers).
template <typename Sink>
__device__ __forceinline__ void lookup_l0_pixel(const LookupL0Params& p, int b, int q, Sink& sink) {
  const float x = __fadd_rn(static_cast<float>(q % p.W1), __ldg(p.flow + static_cast<long long>(b) * 2 * p.HW + q));
  float fl = fminf(fmaxf(floorf(x), -1048576.f), 1048576.f);
  fl = (fl == fl) ? fl : 0.f;
  const int f = static_cast<int>(fl);
  const int s0 = 8 * ((f >> 3) - 4);                      // in-row level-0 index of span element 0
  if (s0 >= p.W2 || s0 + 80 <= 0) {                      // every tap of every level is outside the row
    const float zero[kMaxTaps] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    #pragma unroll
    for (int l = 0; l < kMaxLevels; ++l) sink.put(l, zero);
    return;
  }
  const long long g0 = (static_cast<long long>(b) * p.HW + q) * p.W2 + s0;      // multiple of 8
  const long long gs = g0 & ~15LL;
  uint32_t w0[40];                                        // the 80-element span, two bf16 per word
  if (gs >= 0 && gs + 96 <= p.total) {
    uint32_t w[48];
    #pragma unroll
    for (int i = 0; i < 6; ++i) {
      const LkSector sct = lk_ldg256(p.level0 + gs + 16 * i);
      #pragma unroll
      for (int j = 0; j < 8; ++j) w[8 * i + j] = sct.w[j];
    }
    const bool hi = g0 != gs;                            // span starts 8 elements into the first sector
    #pragma unroll
    for (int i = 0; i < 40; ++i) w0[i] = hi ? w[i + 4] : w[i];
  } else {                                                // first / last rows of the whole volume
    const uint16_t* h16 = reinterpret_cast<const uint16_t*>(p.level0);
    #pragma unroll
    for (int i = 0; i < 40; ++i) {
      const long long e = g0 + 2 * i;
      const uint32_t lo = (e >= 0 && e < p.total) ? __ldg(h16 + e) : 0u;
      const uint32_t hi = (e + 1 >= 0 && e + 1 < p.total) ? __ldg(h16 + e + 1) : 0u;
      w0[i] = lo | (hi << 16);
    }
  }
  uint32_t w1[20], w2[10], w3[5];
  #pragma unroll
  for (int i = 0; i < 20; ++i) w1[i] = lk_avg_bf16x2(w0[2 * i], w0[2 * i + 1]);
  #pragma unroll
  for (int i = 0; i < 10; ++i) w2[i] = lk_avg_bf16x2(w1[2 * i], w1[2 * i + 1]);
  #pragma unroll
  for (int i = 0; i < 5; ++i) w3[i] = lk_avg_bf16x2(w2[2 * i], w2[2 * i + 1]);

  float v[kMaxLevels][kMaxTaps + 1];
  {                                                       // level 0: element offset 28 + (f & 7)
    const int r = f & 7;
    uint32_t t[8], u[6];
    #pragma unroll
    for (int i = 0; i < 8; ++i) t[i] = (r & 4) ? w0[16 + i] : w0[14 + i];
    #pragma unroll
    for (int i = 0; i < 6; ++i) u[i] = (r & 2) ? t[i + 1] : t[i];
    lk_taps(u, r & 1, v[0]);
  }
  {                                                       // level 1: element offset 12 + ((f >> 1) & 3)
    const int r = (f >> 1) & 3;
    uint32_t u[6];
    #pragma unroll
    for (int i = 0; i < 6; ++i) u[i] = (r & 2) ? w1[7 + i] : w1[6 + i];
    lk_taps(u, r & 1, v[1]);
  }
  {                                                       // level 2: element offset 4 + ((f >> 2) & 1)
    uint32_t u[6];
    #pragma unroll
    for (int i = 0; i < 6; ++i) u[i] = w2[2 + i];
    lk_taps(u, (f >> 2) & 1, v[2]);
  }
  {                                                       // level 3: element offset 0
    const uint32_t u[6] = {w3[0], w3[1], w3[2], w3[3], w3[4], 0u};
    lk_taps(u, 0, v[3]);
  }
  #pragma unroll
  for (int l = 0; l < kMaxLevels; ++l) {
    const int Wl = p.W2 >> l;
    const int first = (f >> l) - 4;
    if (first < 0 || first + 9 >= Wl) {
      #pragma unroll
      for (int i = 0; i < kMaxTaps + 1; ++i)
        if (static_cast<unsigned>(first + i) >= static_cast<unsigned>(Wl)) v[l][i] = 0.f;
    }
    const float cl = x * (1.f / static_cast<float>(1 << l));
    const float wm1 = static_cast<float>(Wl - 1), inv = __frcp_rn(wm1);
    float o[kMaxTaps];
    o[0] = lk_sample<0>(v[l], first, cl, 4, inv, wm1); o[1] = lk_sample<1>(v[l], first, cl, 4, inv, wm1);
    o[2] = lk_sample<2>(v[l], first, cl, 4, inv, wm1); o[3] = lk_sample<3>(v[l], first, cl, 4, inv, wm1);
    o[4] = lk_sample<4>(v[l], first, cl, 4, inv, wm1); o[5] = lk_sample<5>(v[l], first, cl, 4, inv, wm1);
    o[6] = lk_sample<6>(v[l], first, cl, 4, inv, wm1); o[7] = lk_sample<7>(v[l], first, cl, 4, inv, wm1);
    o[8] = lk_sample<8>(v[l], first, cl, 4, inv, wm1);
    sink.put(l, o);
  }
}

struct LookupStoreSink {
  float* dst;
  long long HW;
  __device__ __forceinline__ void put(int l, const float (&o)[kMaxTaps]) {
    float* d = dst + static_cast<long long>(l) * 9 * HW;
    #pragma unroll
    for (int k = 0; k < kMaxTaps; ++k) d[static_cast<long long>(k) * HW] = o[k];
  }
};

__global__ void __launch_bounds__(kLookupThreads)
corr_lookup_l0_kernel(const LookupL0Params p) {
  const int b = blockIdx.y;
  const int q = blockIdx.x * kLookupThreads + threadIdx.x;
  if (q >= p.HW) return;
  LookupStoreSink sink{p.out + static_cast<long long>(b) * 36 * p.HW + q, p.HW};
  lookup_l0_pixel(p, b, q, sink);
}



// ================================================================================================================
// K6: lookup fused into the motion encoder's front end.
// Replaces `cor = F.gelu(self.convc1(corr))` of BasicMotionEncoder_v2.forward (reference
// models/core/ppmtereo_update.py:466) with convc1 = PCBlock4_Deep_nopool_res(36, 256, k_conv=[1, 7]) (:91-116), and the
// lookup that produced `corr` (corr.py:74-94): the (B, 36, h, w) fp32 lookup result never reaches HBM.
//   K6a (this kernel), per pixel:  x = lookup;  x = gelu(x + ffn1(x));  x = gelu(x + dw1(x))           -> x2 (B,36,h,w) fp32
//   K6b, 32 x 16 pixel tiles:      x = gelu(x + dw7(x));  x = gelu(x + pw(x));  cor = gelu(ffn2(x))    -> (B,256,h,w) bf16
// Rounding flow = torch / cuDNN under bf16 autocast, identified on the box (scripts/exp/pcblock_rounding_probe.py):
// a 1x1 conv is bf16(bf16(acc) + bias) (99.999 % reproduced), a depthwise conv bf16(acc + bias) (99.995-100 %), GELU of a
// bf16 tensor is bf16, and the residual stream x + branch with its GELUs stays fp32 (fp32 + bf16 promotes to fp32).
// Weights arrive as one fp32 array of bf16-representable values (layout: MfA / MfB offsets below).
namespace mf {
constexpr int kC = 36, kHid = 54, kOut = 256;
// K6a block
constexpr int kWa = 0;                      // ffn1.0 weight [54][36]
constexpr int kBa = kWa + kHid * kC;        // ffn1.0 bias [54] (+2 pad)
constexpr int kWb = kBa + 56;               // ffn1.2 weight [36][56] (rows padded from 54)
constexpr int kBb = kWb + kC * 56;          // ffn1.2 bias [36]
constexpr int kW1 = kBb + kC;               // conv_list.0 (depthwise k = 1) weight [36]
constexpr int kB1 = kW1 + kC;               // conv_list.0 bias [36]
constexpr int kASize = kB1 + kC;            // 4124 floats
// K6b block
constexpr int kW7 = 0;                      // conv_list.1 (depthwise 7x7) weight [36][49]
constexpr int kB7 = kW7 + kC * 49;          // bias [36]
constexpr int kWp = kB7 + kC;               // pw weight [36][36]
constexpr int kBp = kWp + kC * kC;          // pw bias [36]
constexpr int kWc = kBp + kC;               // ffn2.0 weight [54][36]
constexpr int kBc = kWc + kHid * kC;        // ffn2.0 bias [54] (+2 pad)
constexpr int kWd = kBc + 56;               // ffn2.2 weight TRANSPOSED [54][256]
constexpr int kBd = kWd + kHid * kOut;      // ffn2.2 bias [256]
constexpr int kBSize = kBd + kOut;          // 19212 floats
static_assert(kBd == ppm::kMotionFrontBiasD, "common.cuh's copy of the ffn2.2 bias offset is stale");
}  // namespace mf

__device__ __forceinline__ float mf_gelu(float x) { return 0.5f * x * (1.f + erff(x * 0.70710678118654752440f)); }
__device__ __forceinline__ float mf_r(float x) { return bf16_round_f(x); }

struct LookupRegSink {
  float* x;                                                // 36 registers
  __device__ __forceinline__ void put(int l, const float (&o)[kMaxTaps]) {
    #pragma unroll
    for (int k = 0; k < kMaxTaps; ++k) x[l * kMaxTaps + k] = o[k];
  }
};

struct MotionFrontAParams {
  LookupL0Params lk;               // fused lookup (lk.out unused), or
  const float* corr_in;            // a materialised (B, 36, h, w) fp32 lookup result (foreign pyramids); nullptr = fused
  const float* wts;                // mf::kASize floats
  float* x2;                       // (B, 36, h, w) fp32
};

__global__ void __launch_bounds__(kLookupThreads, 3)      // <= 168 registers: 12 warps / SM for the gather's latency
motion_front_a_kernel(const MotionFrontAParams p) {
  __shared__ __align__(16) float sw[mf::kASize];
  for (int i = threadIdx.x; i < mf::kASize; i += kLookupThreads) sw[i] = __ldg(p.wts + i);
  __syncthreads();
  const int b = blockIdx.y, HW = p.lk.HW;
  const int q = blockIdx.x * kLookupThreads + threadIdx.x;
  if (q >= HW) return;
  float x[mf::kC];
  if (p.corr_in) {
    #pragma unroll
    for (int c = 0; c < mf::kC; ++c) x[c] = __ldg(p.corr_in + (static_cast<long long>(b) * mf::kC + c) * HW + q);
  } else {
    LookupRegSink sink{x};
    lookup_l0_pixel(p.lk, b, q, sink);
  }
  float xr[mf::kC];
  #pragma unroll
  for (int c = 0; c < mf::kC; ++c) xr[c] = mf_r(x[c]);                   // the conv's autocast cast of its input
  // ffn1: 36 -> 54 -> GELU -> 36
  float hid[mf::kHid];
  #pragma unroll
  for (int o = 0; o < mf::kHid; ++o) {
    const float4* w = reinterpret_cast<const float4*>(sw + mf::kWa + o * mf::kC);
    float acc = 0.f;
    #pragma unroll
    for (int c4 = 0; c4 < mf::kC / 4; ++c4) {
      const float4 w4 = w[c4];
      acc = fmaf(w4.x, xr[4 * c4], acc); acc = fmaf(w4.y, xr[4 * c4 + 1], acc);
      acc = fmaf(w4.z, xr[4 * c4 + 2], acc); acc = fmaf(w4.w, xr[4 * c4 + 3], acc);
    }
    hid[o] = mf_r(mf_gelu(mf_r(mf_r(acc) + sw[mf::kBa + o])));
  }
  float* dst = p.x2 + static_cast<long long>(b) * mf::kC * HW + q;
  #pragma unroll
  for (int o = 0; o < mf::kC; ++o) {
    const float2* w = reinterpret_cast<const float2*>(sw + mf::kWb + o * 56);
    float acc = 0.f;
    #pragma unroll
    for (int c2 = 0; c2 < mf::kHid / 2; ++c2) {
      const float2 w2 = w[c2];
      acc = fmaf(w2.x, hid[2 * c2], acc); acc = fmaf(w2.y, hid[2 * c2 + 1], acc);
    }
    const float x1 = mf_gelu(x[o] + mf_r(mf_r(acc) + sw[mf::kBb + o]));          // fp32 residual stream
    const float d1 = mf_r(fmaf(sw[mf::kW1 + o], mf_r(x1), sw[mf::kB1 + o]));     // depthwise k = 1: bf16(w * x + b)
    dst[static_cast<long long>(o) * HW] = mf_gelu(x1 + d1);
  }
}

// K6b: CTA = 32 x 16 pixels of one frame, thread = pixel.  The bf16 copy of x2 the depthwise conv reads (its autocast
// cast) is staged with a 3-pixel halo in shared memory, zero outside the image (the conv's zero padding); all weights of
// the tail sit in shared memory as fp32 (77 KB) and are read as warp-wide broadcasts.
constexpr int kMfTileW = 32, kMfTileH = 16, kMfHalo = 3;
constexpr int kMfTW = kMfTileW + 2 * kMfHalo + 2;            // 40: row pitch of the tile (38 used)
constexpr int kMfTH = kMfTileH + 2 * kMfHalo;                // 22
constexpr int kMfBThreads = kMfTileW * kMfTileH;             // 512
constexpr int kMfBSmem = mf::kBSize * 4 + mf::kC * kMfTH * kMfTW * 2;

__global__ void __launch_bounds__(kMfBThreads, 1)
motion_front_b_kernel(const float* __restrict__ x2, const float* __restrict__ wts, __nv_bfloat16* __restrict__ cor,
                      int H, int W, int channels_last) {
  extern __shared__ __align__(16) uint8_t mf_smem[];
  float* sw = reinterpret_cast<float*>(mf_smem);
  __nv_bfloat16* tile = reinterpret_cast<__nv_bfloat16*>(mf_smem + mf::kBSize * 4);
  const int HW = H * W;
  const int x0 = blockIdx.x * kMfTileW, y0 = blockIdx.y * kMfTileH, frame = blockIdx.z;
  const float* src = x2 + static_cast<long long>(frame) * mf::kC * HW;
  for (int i = threadIdx.x; i < mf::kBSize; i += kMfBThreads) sw[i] = __ldg(wts + i);
  for (int i = threadIdx.x; i < mf::kC * kMfTH * kMfTW; i += kMfBThreads) {
    const int tx = i % kMfTW, ty = (i / kMfTW) % kMfTH, c = i / (kMfTW * kMfTH);
    const int gx = x0 + tx - kMfHalo, gy = y0 + ty - kMfHalo;
    float v = 0.f;
    if (tx < kMfTileW + 2 * kMfHalo && gx >= 0 && gx < W && gy >= 0 && gy < H) v = __ldg(src + static_cast<long long>(c) * HW + gy * W + gx);
    tile[i] = __float2bfloat16_rn(v);
  }
  __syncthreads();
  const int tx = threadIdx.x & (kMfTileW - 1), ty = threadIdx.x / kMfTileW;
  const int gx = x0 + tx, gy = y0 + ty;
  if (gx >= W || gy >= H) return;
  const long long pix = static_cast<long long>(gy) * W + gx;
  // ---- x3 = gelu(x2 + dw7(x2))
  float x3[mf::kC];
  #pragma unroll
  for (int c = 0; c < mf::kC; ++c) {
    const __nv_bfloat16* t = tile + (c * kMfTH + ty) * kMfTW + tx;
    const float* w = sw + mf::kW7 + c * 49;
    float acc = 0.f;
    #pragma unroll
    for (int ky = 0; ky < 7; ++ky)
      #pragma unroll
      for (int kx = 0; kx < 7; ++kx) acc = fmaf(w[ky * 7 + kx], __bfloat162float(t[ky * kMfTW + kx]), acc);
    const float d7 = mf_r(acc + sw[mf::kB7 + c]);                          // depthwise: bf16(acc + b)
    x3[c] = mf_gelu(__ldg(src + static_cast<long long>(c) * HW + pix) + d7);
  }
  // ---- x4 = gelu(x3 + pw(x3))
  float xr[mf::kC], x4r[mf::kC];
  #pragma unroll
  for (int c = 0; c < mf::kC; ++c) xr[c] = mf_r(x3[c]);
  #pragma unroll
  for (int o = 0; o < mf::kC; ++o) {
    const float4* w = reinterpret_cast<const float4*>(sw + mf::kWp + o * mf::kC);
    float acc = 0.f;
    #pragma unroll
    for (int c4 = 0; c4 < mf::kC / 4; ++c4) {
      const float4 w4 = w[c4];
      acc = fmaf(w4.x, xr[4 * c4], acc); acc = fmaf(w4.y, xr[4 * c4 + 1], acc);
      acc = fmaf(w4.z, xr[4 * c4 + 2], acc); acc = fmaf(w4.w, xr[4 * c4 + 3], acc);
    }
    x4r[o] = mf_r(mf_gelu(x3[o] + mf_r(mf_r(acc) + sw[mf::kBp + o])));    // bf16(x4): ffn2.0's input cast
  }
  // ---- u = bf16(gelu(ffn2.0(x4)))
  float u[mf::kHid];
  #pragma unroll
  for (int o = 0; o < mf::kHid; ++o) {
    const float4* w = reinterpret_cast<const float4*>(sw + mf::kWc + o * mf::kC);
    float acc = 0.f;
    #pragma unroll
    for (int c4 = 0; c4 < mf::kC / 4; ++c4) {
      const float4 w4 = w[c4];
      acc = fmaf(w4.x, x4r[4 * c4], acc); acc = fmaf(w4.y, x4r[4 * c4 + 1], acc);
      acc = fmaf(w4.z, x4r[4 * c4 + 2], acc); acc = fmaf(w4.w, x4r[4 * c4 + 3], acc);
    }
    u[o] = mf_r(mf_gelu(mf_r(mf_r(acc) + sw[mf::kBc + o])));
  }
  if (channels_last == 2) {
    // hidden vector only: [pixel][64] bf16 (54 values + zero padding) = the K-major operand of the 54 -> 256 convolution,
    // which then runs as a tensor-core GEMM with the GELU in its epilogue (ppm_motion_encoder, K7)
    uint4* d = reinterpret_cast<uint4*>(cor + (static_cast<long long>(frame) * HW + pix) * 64);
    #pragma unroll
    for (int g = 0; g < 8; ++g) {
      float e[8];
      #pragma unroll
      for (int j = 0; j < 8; ++j) e[j] = (g * 8 + j < mf::kHid) ? u[(g * 8 + j < mf::kHid) ? g * 8 + j : 0] : 0.f;
      uint4 v;
      v.x = pack_bf16x2(e[0], e[1]); v.y = pack_bf16x2(e[2], e[3]); v.z = pack_bf16x2(e[4], e[5]); v.w = pack_bf16x2(e[6], e[7]);
      d[g] = v;
    }
    return;
  }
  // ---- cor = bf16(gelu(ffn2.2(u))), 16 output channels at a time (weights transposed: [c][o])
  // NCHW (the reference's layout) or channels-last [pixel][256] (operand of the shifted-GEMM convolutions, K5/K7)
  __nv_bfloat16* dst = channels_last ? cor + (static_cast<long long>(frame) * HW + pix) * mf::kOut
                                     : cor + static_cast<long long>(frame) * mf::kOut * HW + pix;
  for (int o0 = 0; o0 < mf::kOut; o0 += 16) {
    float acc[16];
    #pragma unroll
    for (int j = 0; j < 16; ++j) acc[j] = 0.f;
    #pragma unroll
    for (int c = 0; c < mf::kHid; ++c) {
      const float4* w = reinterpret_cast<const float4*>(sw + mf::kWd + c * mf::kOut + o0);
      #pragma unroll
      for (int j4 = 0; j4 < 4; ++j4) {
        const float4 w4 = w[j4];
        acc[4 * j4] = fmaf(w4.x, u[c], acc[4 * j4]); acc[4 * j4 + 1] = fmaf(w4.y, u[c], acc[4 * j4 + 1]);
        acc[4 * j4 + 2] = fmaf(w4.z, u[c], acc[4 * j4 + 2]); acc[4 * j4 + 3] = fmaf(w4.w, u[c], acc[4 * j4 + 3]);
      }
    }
    #pragma unroll
    for (int j = 0; j < 16; ++j)
      acc[j] = mf_gelu(mf_r(mf_r(acc[j]) + sw[mf::kBd + o0 + j]));                   // the encoder's outer F.gelu
    if (channels_last) {
      uint4 v0, v1;
      v0.x = pack_bf16x2(acc[0], acc[1]); v0.y = pack_bf16x2(acc[2], acc[3]);
      v0.z = pack_bf16x2(acc[4], acc[5]); v0.w = pack_bf16x2(acc[6], acc[7]);
      v1.x = pack_bf16x2(acc[8], acc[9]); v1.y = pack_bf16x2(acc[10], acc[11]);
      v1.z = pack_bf16x2(acc[12], acc[13]); v1.w = pack_bf16x2(acc[14], acc[15]);
      *reinterpret_cast<uint4*>(dst + o0) = v0;
      *reinterpret_cast<uint4*>(dst + o0 + 8) = v1;
    } else {
      #pragma unroll
      for (int j = 0; j < 16; ++j) dst[static_cast<long long>(o0 + j) * HW] = __float2bfloat16_rn(acc[j]);
    }
  }
}

}  // namespace ppm

extern "C" int ppm_corr_lookup(const void* const* levels, int vol_dtype, const float* flow, int B,
                               int H, int W1, int W2, int radius, int n_levels, float* out,
                               void* stream) {
  using namespace ppm;
  PPM_CHECK(levels && flow && out, PPM_ERR_ARG, "ppm_corr_lookup: null pointer");
  PPM_CHECK(B > 0 && H > 0 && W1 > 0 && W2 > 0, PPM_ERR_ARG, "ppm_corr_lookup: bad shape");
  PPM_CHECK(radius >= 0 && radius <= 4, PPM_ERR_UNSUPPORTED, "ppm_corr_lookup: radius %d > 4", radius);
  PPM_CHECK(n_levels >= 1 && n_levels <= kMaxLevels, PPM_ERR_UNSUPPORTED,
            "ppm_corr_lookup: n_levels %d not in [1,4]", n_levels);
  PPM_CHECK(vol_dtype == PPM_F32 || vol_dtype == PPM_BF16, PPM_ERR_ARG, "ppm_corr_lookup: bad dtype");
  PPM_CHECK((W2 >> (n_levels - 1)) >= 1, PPM_ERR_ARG, "ppm_corr_lookup: W2 %d too small", W2);
  PPM_CHECK(B <= 65535, PPM_ERR_UNSUPPORTED, "ppm_corr_lookup: B %d > 65535", B);
  LookupParams p{};
  const long long rows = static_cast<long long>(B) * H * W1;
  for (int l = 0; l < n_levels; ++l) {
    PPM_CHECK(levels[l] && (reinterpret_cast<uintptr_t>(levels[l]) & 31) == 0, PPM_ERR_ARG,
              "ppm_corr_lookup: level %d pointer null or not 32-byte aligned", l);
    p.level[l] = levels[l];
    p.width[l] = W2 >> l;
    p.total[l] = rows * (W2 >> l);
  }
  p.flow = flow; p.out = out; p.HW = H * W1; p.W1 = W1; p.radius = radius; p.n_levels = n_levels;
  dim3 grid((p.HW + kLookupThreads - 1) / kLookupThreads, B);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (vol_dtype == PPM_BF16)
    corr_lookup_kernel<__nv_bfloat16><<<grid, kLookupThreads, 0, st>>>(p);
  else
    corr_lookup_kernel<float><<<grid, kLookupThreads, 0, st>>>(p);
  PPM_CUDA(cudaGetLastError());
  return PPM_OK;
}

extern "C" int ppm_corr_lookup_l0(const void* level0, const float* flow, int B, int H, int W1, int W2, float* out,
                                  void* stream) {
  using namespace ppm;
  PPM_CHECK(level0 && flow && out, PPM_ERR_ARG, "ppm_corr_lookup_l0: null pointer");
  PPM_CHECK(B > 0 && H > 0 && W1 > 0 && W2 > 0, PPM_ERR_ARG, "ppm_corr_lookup_l0: bad shape");
  PPM_CHECK(W2 % 8 == 0 && (W2 >> 3) >= 1, PPM_ERR_UNSUPPORTED, "ppm_corr_lookup_l0: W2 = %d must be a multiple of 8", W2);
  PPM_CHECK((reinterpret_cast<uintptr_t>(level0) & 31) == 0, PPM_ERR_ARG, "ppm_corr_lookup_l0: level 0 must be 32-byte aligned");
  PPM_CHECK(B <= 65535, PPM_ERR_UNSUPPORTED, "ppm_corr_lookup_l0: B %d > 65535", B);
  LookupL0Params p{};
  p.level0 = static_cast<const __nv_bfloat16*>(level0);
  p.HW = H * W1; p.W1 = W1; p.W2 = W2;
  p.total = static_cast<long long>(B) * H * W1 * W2;
  p.flow = flow; p.out = out;
  dim3 grid((p.HW + kLookupThreads - 1) / kLookupThreads, B);
  corr_lookup_l0_kernel<<<grid, kLookupThreads, 0, static_cast<cudaStream_t>(stream)>>>(p);
  PPM_CUDA(cudaGetLastError());
  return PPM_OK;
}


// K6 entry: see the section header above.  level0 / flow as ppm_corr_lookup_l0 (fused lookup), or corr != NULL to start
// from a materialised (B, 36, H, W1) fp32 lookup result.  weights_a / weights_b: fp32 arrays (bf16-representable values)
// in the MfA / MfB layouts; x2_ws: B*36*H*W1 floats of scratch; cor: (B, 256, H, W1) bf16.
namespace ppm {
// Both K6 kernels; `cor` NCHW (channels_last = 0) or [pixel][256] (channels_last = 1).  Declared in common.cuh.
int motion_front_launch(const void* level0, const float* flow, const float* corr, const float* weights_a,
                        const float* weights_b, int B, int H, int W1, int W2, float* x2_ws, void* cor, int channels_last,
                        cudaStream_t st) {
  PPM_CHECK((corr || (level0 && flow)) && weights_a && weights_b && x2_ws && cor, PPM_ERR_ARG,
            "ppm_motion_corr_front: null pointer");
  PPM_CHECK(B > 0 && H > 0 && W1 > 0 && B <= 65535, PPM_ERR_ARG, "ppm_motion_corr_front: bad shape");
  MotionFrontAParams a{};
  a.lk.HW = H * W1; a.lk.W1 = W1; a.lk.W2 = W2;
  if (!corr) {
    PPM_CHECK(W2 > 0 && W2 % 8 == 0, PPM_ERR_UNSUPPORTED, "ppm_motion_corr_front: W2 = %d must be a multiple of 8", W2);
    PPM_CHECK((reinterpret_cast<uintptr_t>(level0) & 31) == 0, PPM_ERR_ARG, "ppm_motion_corr_front: level 0 must be 32-byte aligned");
    a.lk.level0 = static_cast<const __nv_bfloat16*>(level0);
    a.lk.total = static_cast<long long>(B) * H * W1 * W2;
    a.lk.flow = flow;
  }
  a.corr_in = corr; a.wts = weights_a; a.x2 = x2_ws;
  motion_front_a_kernel<<<dim3((a.lk.HW + kLookupThreads - 1) / kLookupThreads, B), kLookupThreads, 0, st>>>(a);
  PPM_CUDA(cudaGetLastError());
  static std::atomic<uint64_t> attr_done{0};
  int dev = 0;
  PPM_CUDA(cudaGetDevice(&dev));
  if (dev >= 64 || !((attr_done.load(std::memory_order_acquire) >> dev) & 1)) {
    PPM_CUDA(cudaFuncSetAttribute(motion_front_b_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kMfBSmem));
    if (dev < 64) attr_done.fetch_or(1ull << dev, std::memory_order_release);
  }
  PPM_CHECK((H + kMfTileH - 1) / kMfTileH <= 65535, PPM_ERR_UNSUPPORTED, "ppm_motion_corr_front: H too large");
  dim3 grid((W1 + kMfTileW - 1) / kMfTileW, (H + kMfTileH - 1) / kMfTileH, B);
  motion_front_b_kernel<<<grid, kMfBThreads, kMfBSmem, st>>>(x2_ws, weights_b, static_cast<__nv_bfloat16*>(cor), H, W1,
                                                              channels_last);
  PPM_CUDA(cudaGetLastError());
  return PPM_OK;
}
}  // namespace ppm

extern "C" int ppm_motion_corr_front(const void* level0, const float* flow, const float* corr, const float* weights_a,
                                     const float* weights_b, int B, int H, int W1, int W2, float* x2_ws, void* cor,
                                     void* stream) {
  return ppm::motion_front_launch(level0, flow, corr, weights_a, weights_b, B, H, W1, W2, x2_ws, cor, 0,
                                  static_cast<cudaStream_t>(stream));
}
