// K2: radius-r bilinear lookup into the 1-D correlation pyramid (HBM-bound gather).
// Replaces CorrBlock1D.__call__ + bilinear_sampler (reference models/core/corr.py:74-94, :10-27).
//
// Every pixel (b,y,x) owns a private row of each pyramid level, so there is no data reuse
// between pixels: the job is to touch each needed 32-byte sector exactly once and to write the
// (B, L*(2r+1), H, W) output with full 128-byte lines.
//   phase 1: 4 lanes per (pixel, level) fetch an aligned 16-element window that covers the
//            2r+3 candidate taps with one vector load each (8 B for bf16, 16 B for fp32
//            volumes), zero the elements that fall outside the pixel's row, and stage the
//            window in shared memory as fp32.  8 independent loads in flight per thread.
//   phase 2: one thread per (pixel, level) interpolates the 2r+1 outputs from shared memory;
//            lanes are consecutive pixels, so every store instruction writes 128 contiguous
//            bytes of one output channel.
// The sample position goes through the same normalise / un-normalise round trip as the
// reference's grid_sample call (align_corners=True) so that results agree to ~1e-6.
#include "../../include/ppm_b200.h"
#include "common.cuh"

namespace ppm {

constexpr int kLookupPix = 128;      // pixels per CTA
constexpr int kLookupThreads = 256;
constexpr int kWin = 16;             // staged window, elements
constexpr int kWinStride = 17;       // +1 pad: conflict-free phase-2 reads
constexpr int kMaxLevels = 4;

struct LookupParams {
  const void* level[kMaxLevels];
  int width[kMaxLevels];             // W2 >> l
  long long total[kMaxLevels];       // elements in level l
  const float* flow;                 // (B,2,H,W1)
  float* out;                        // (B, L*(2r+1), H, W1)
  int HW, W1, radius, n_levels;
};

// Select from a kernel-parameter array without forcing a local-memory copy of it.
template <typename T>
__device__ __forceinline__ T pick(const T (&a)[kMaxLevels], int l) {
  T v = a[0];
  #pragma unroll
  for (int i = 1; i < kMaxLevels; ++i) v = (l == i) ? a[i] : v;
  return v;
}

__device__ __forceinline__ void load4(const float* p, float (&v)[4]) {
  float4 t = __ldg(reinterpret_cast<const float4*>(p));
  v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
__device__ __forceinline__ void load4(const __nv_bfloat16* p, float (&v)[4]) {
  uint2 t = __ldg(reinterpret_cast<const uint2*>(p));
  v[0] = __uint_as_float(t.x << 16); v[1] = __uint_as_float(t.x & 0xFFFF0000u);
  v[2] = __uint_as_float(t.y << 16); v[3] = __uint_as_float(t.y & 0xFFFF0000u);
}
__device__ __forceinline__ float load1(const float* p) { return __ldg(p); }
__device__ __forceinline__ float load1(const __nv_bfloat16* p) {
  return __bfloat162float(__ldg(p));
}

// floor((x + flow) / 2^l), clamped so that the int conversion is defined for any input.
__device__ __forceinline__ float level_center(float xf, float fl, int l) {
  return __fdiv_rn(__fadd_rn(xf, fl), static_cast<float>(1 << l));
}
__device__ __forceinline__ int safe_floor(float v) {
  v = fminf(fmaxf(floorf(v), -1048576.f), 1048576.f);
  return (v == v) ? static_cast<int>(v) : 0;
}

template <typename VolT>
__global__ void __launch_bounds__(kLookupThreads)
corr_lookup_kernel(const LookupParams p) {
  __shared__ float win[kMaxLevels * kLookupPix * kWinStride];
  const int tid = threadIdx.x;
  const int b = blockIdx.y;
  const int q0 = blockIdx.x * kLookupPix;
  const int r = p.radius;
  const float* flow_b = p.flow + static_cast<long long>(b) * 2 * p.HW;

  // ---------------- phase 1: gather windows
  const int lane_tasks = p.n_levels * kLookupPix * 4;
  #pragma unroll
  for (int it = 0; it < kMaxLevels * kLookupPix * 4 / kLookupThreads; ++it) {
    const int u = it * kLookupThreads + tid;
    if (u < lane_tasks) {
      const int l = it / (kLookupPix * 4 / kLookupThreads);   // compile-time after unrolling
      const int pix = (u >> 2) & (kLookupPix - 1);
      const int sub = u & 3;
      const int q = q0 + pix;
      float v[4] = {0.f, 0.f, 0.f, 0.f};
      if (q < p.HW) {
        const int Wl = p.width[l];
        const float xl = level_center(static_cast<float>(q % p.W1), __ldg(flow_b + q), l);
        const long long row0 = (static_cast<long long>(b) * p.HW + q) * Wl;
        const long long g = row0 + (safe_floor(xl) - r - 1);
        const long long e0 = (g & ~3LL) + 4 * sub;            // first element of this lane
        const VolT* base = static_cast<const VolT*>(p.level[l]);
        const long long total = p.total[l];
        if (e0 >= 0 && e0 + 3 < total) {
          load4(base + e0, v);
        } else {
          #pragma unroll
          for (int m = 0; m < 4; ++m)
            if (e0 + m >= 0 && e0 + m < total) v[m] = load1(base + e0 + m);
        }
        #pragma unroll
        for (int m = 0; m < 4; ++m) {
          const long long in_row = e0 + m - row0;
          if (in_row < 0 || in_row >= Wl) v[m] = 0.f;           // zeros padding
        }
      }
      float* w = win + (l * kLookupPix + pix) * kWinStride + 4 * sub;
      w[0] = v[0]; w[1] = v[1]; w[2] = v[2]; w[3] = v[3];
    }
  }
  __syncthreads();

  // ---------------- phase 2: interpolate and store
  const int pix = tid & (kLookupPix - 1);
  const int q = q0 + pix;
  if (q >= p.HW) return;
  const float xf = static_cast<float>(q % p.W1);
  const float fl = __ldg(flow_b + q);
  const int taps = 2 * r + 1;
  const int C = p.n_levels * taps;
  for (int l = tid / kLookupPix; l < p.n_levels; l += kLookupThreads / kLookupPix) {
    const int Wl = pick(p.width, l);
    const float xl = level_center(xf, fl, l);
    const int f0 = safe_floor(xl) - r - 1;                      // in-row index of window tap 0
    const long long g = (static_cast<long long>(b) * p.HW + q) * Wl + f0;
    const int o = static_cast<int>(g - (g & ~3LL));
    const float* w = win + (l * kLookupPix + pix) * kWinStride;
    const float wm1 = static_cast<float>(Wl - 1);
    float* dst = p.out + (static_cast<long long>(b) * C + l * taps) * p.HW + q;
    for (int k = 0; k < taps; ++k) {
      // reference: x0 = dx + x/2^l; xn = 2*x0/(W-1) - 1; grid_sample: ix = ((xn+1)/2)*(W-1)
      const float x0 = __fadd_rn(static_cast<float>(k - r), xl);
      float ix = x0;
      if (Wl > 1) {
        const float xn = __fsub_rn(__fdiv_rn(__fmul_rn(2.f, x0), wm1), 1.f);
        ix = __fmul_rn(__fdiv_rn(__fadd_rn(xn, 1.f), 2.f), wm1);
      }
      const float i0f = floorf(ix);
      const int i0 = safe_floor(ix);
      const float w1 = __fsub_rn(ix, i0f);
      const float w0 = __fsub_rn(__fadd_rn(i0f, 1.f), ix);
      int j = i0 - f0 + o;
      j = min(max(j, 0), kWin - 2);
      dst[static_cast<long long>(k) * p.HW] = __fadd_rn(__fmul_rn(w[j], w0), __fmul_rn(w[j + 1], w1));
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
    PPM_CHECK(levels[l] && (reinterpret_cast<uintptr_t>(levels[l]) & 15) == 0, PPM_ERR_ARG,
              "ppm_corr_lookup: level %d pointer null or not 16-byte aligned", l);
    p.level[l] = levels[l];
    p.width[l] = W2 >> l;
    p.total[l] = rows * (W2 >> l);
  }
  p.flow = flow; p.out = out; p.HW = H * W1; p.W1 = W1; p.radius = radius; p.n_levels = n_levels;
  dim3 grid((p.HW + kLookupPix - 1) / kLookupPix, B);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (vol_dtype == PPM_BF16)
    corr_lookup_kernel<__nv_bfloat16><<<grid, kLookupThreads, 0, st>>>(p);
  else
    corr_lookup_kernel<float><<<grid, kLookupThreads, 0, st>>>(p);
  PPM_CUDA(cudaGetLastError());
  return PPM_OK;
}
