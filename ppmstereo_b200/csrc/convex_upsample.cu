// K4: convex upsampling of the flow field (first glue op moved onto the GPU path).
// Replaces PPMStereo.convex_upsample (reference models/core/ppmstereo.py:185-197), called once per update
// iteration at :577:
//   mask.view(N, 1, 9, r, r, H, W) -> softmax over the 9 taps -> sum_k p_k * (r * flow)[3x3 neighbour k]
//   -> (N, 2, r*H, r*W), zero padding outside the image.
// The reference spends ~7 passes over (N, 9*r*r, H, W) tensors (view/contiguous, softmax, unfold, mul, sum,
// permute/contiguous); here every mask value is read once and every output written once.
// HBM-bound: 9*r*r*e_mask + 2*r*r*4 (+ 8 for the flow) bytes per coarse pixel.
//   thread = one coarse pixel (n, y, x); lanes = consecutive x, so each of the 9*r*r mask channels is read with
//   a full-width coalesced request and every store instruction writes r*4 contiguous bytes per lane;
//   per fine row i: softmax of the r columns' 9 taps in fp32, then the two flow channels.  A bf16 mask is a STORAGE
//   type only: the reference reaches this op under torch.autocast (ppmstereo.py:617), where torch.softmax is an
//   autocast-to-fp32 op -- the weights are fp32 and are never rounded to bf16.
#include <math.h>

#include "../../include/ppm_b200.h"
#include "common.cuh"

namespace ppm {

constexpr int kCuThreads = 128;
constexpr int kCuRate = 4;

__device__ __forceinline__ float cu_ld(const float* p) { return __ldg(p); }
__device__ __forceinline__ float cu_ld(const __nv_bfloat16* p) { return __bfloat162float(__ldg(p)); }

template <typename MaskT>
__global__ void __launch_bounds__(kCuThreads)
convex_upsample_kernel(const float* __restrict__ flow, const MaskT* __restrict__ mask, float* __restrict__ out,
                       int H, int W) {
  const int n = blockIdx.y;
  const int HW = H * W;
  const int q = blockIdx.x * kCuThreads + threadIdx.x;
  if (q >= HW) return;
  const int y = q / W, x = q - y * W;
  // 3x3 neighbourhood of r * flow, zero padded (F.unfold(rate * flow, [3, 3], padding=1))
  float f[2][9];
  const float* fl = flow + static_cast<long long>(n) * 2 * HW;
  #pragma unroll
  for (int k = 0; k < 9; ++k) {
    const int yy = y + k / 3 - 1, xx = x + k % 3 - 1;
    const bool in = yy >= 0 && yy < H && xx >= 0 && xx < W;
    #pragma unroll
    for (int c = 0; c < 2; ++c)
      f[c][k] = in ? __fmul_rn(static_cast<float>(kCuRate), __ldg(fl + c * HW + yy * W + xx)) : 0.f;
  }
  const MaskT* mk = mask + static_cast<long long>(n) * 9 * kCuRate * kCuRate * HW + q;
  float* o = out + static_cast<long long>(n) * 2 * kCuRate * kCuRate * HW + static_cast<long long>(y) * kCuRate * kCuRate * W +
             x * kCuRate;
  #pragma unroll
  for (int i = 0; i < kCuRate; ++i) {
    float m[9][kCuRate];
    #pragma unroll
    for (int k = 0; k < 9; ++k)
      #pragma unroll
      for (int j = 0; j < kCuRate; ++j)
        m[k][j] = cu_ld(mk + static_cast<long long>((k * kCuRate + i) * kCuRate + j) * HW);
    float r0[kCuRate], r1[kCuRate];
    #pragma unroll
    for (int j = 0; j < kCuRate; ++j) {
      float mx = m[0][j];
      #pragma unroll
      for (int k = 1; k < 9; ++k) mx = fmaxf(mx, m[k][j]);
      // exp through ex2.approx and a reciprocal instead of expf / IEEE division: relative error ~2e-7 per
      // weight, far inside the 1e-5 parity bar (the accurate forms made this kernel instruction-bound)
      float e[9], s = 0.f;
      #pragma unroll
      for (int k = 0; k < 9; ++k) { e[k] = ex2_approx((m[k][j] - mx) * 1.4426950408889634f); s += e[k]; }
      const float inv = __frcp_rn(s);
      float a0 = 0.f, a1 = 0.f;
      #pragma unroll
      for (int k = 0; k < 9; ++k) {
        const float pk = e[k] * inv;
        a0 = __fadd_rn(a0, __fmul_rn(pk, f[0][k]));
        a1 = __fadd_rn(a1, __fmul_rn(pk, f[1][k]));
      }
      r0[j] = a0; r1[j] = a1;
    }
    const long long row = static_cast<long long>(i) * kCuRate * W;
    *reinterpret_cast<float4*>(o + row) = make_float4(r0[0], r0[1], r0[2], r0[3]);
    *reinterpret_cast<float4*>(o + static_cast<long long>(kCuRate) * kCuRate * HW + row) = make_float4(r1[0], r1[1], r1[2], r1[3]);
  }
}

// bf16 masks, even W: a thread takes the pixel pair (x, x+1) so that every mask channel is read with 4-byte loads
// (the one-pixel form issues 144 two-byte loads per thread and is load-instruction-bound at half the HBM rate).
// Per pixel the arithmetic and its order are those of the kernel above: results are bit-identical.
__global__ void __launch_bounds__(kCuThreads)
convex_upsample_bf16x2_kernel(const float* __restrict__ flow, const __nv_bfloat16* __restrict__ mask,
                              float* __restrict__ out, int H, int W) {
  const int n = blockIdx.y;
  const int HW = H * W;
  const int q = 2 * (blockIdx.x * kCuThreads + threadIdx.x);
  if (q >= HW) return;
  const int y = q / W, x = q - y * W;                       // x even, x + 1 < W
  float fw[2][3][4];                                        // rows y-1..y+1, columns x-1..x+2 of r * flow, zero padded
  const float* fl = flow + static_cast<long long>(n) * 2 * HW;
  #pragma unroll
  for (int dy = 0; dy < 3; ++dy)
    #pragma unroll
    for (int dx = 0; dx < 4; ++dx) {
      const int yy = y + dy - 1, xx = x + dx - 1;
      const bool in = yy >= 0 && yy < H && xx >= 0 && xx < W;
      #pragma unroll
      for (int c = 0; c < 2; ++c)
        fw[c][dy][dx] = in ? __fmul_rn(static_cast<float>(kCuRate), __ldg(fl + c * HW + yy * W + xx)) : 0.f;
    }
  const uint32_t* mk = reinterpret_cast<const uint32_t*>(mask + static_cast<long long>(n) * 9 * kCuRate * kCuRate * HW + q);
  const long long ch_words = HW / 2;
  float* o = out + static_cast<long long>(n) * 2 * kCuRate * kCuRate * HW + static_cast<long long>(y) * kCuRate * kCuRate * W +
             x * kCuRate;
  #pragma unroll
  for (int i = 0; i < kCuRate; ++i) {
    uint32_t mw[9][kCuRate];
    #pragma unroll
    for (int k = 0; k < 9; ++k)
      #pragma unroll
      for (int j = 0; j < kCuRate; ++j)
        mw[k][j] = __ldg(mk + static_cast<long long>((k * kCuRate + i) * kCuRate + j) * ch_words);
    float r[2][2][kCuRate];                                 // [pixel][channel][fine column]
    #pragma unroll
    for (int px = 0; px < 2; ++px)
      #pragma unroll
      for (int j = 0; j < kCuRate; ++j) {
        float m[9];
        #pragma unroll
        for (int k = 0; k < 9; ++k) m[k] = __uint_as_float(px ? (mw[k][j] & 0xFFFF0000u) : (mw[k][j] << 16));
        float mx = m[0];
        #pragma unroll
        for (int k = 1; k < 9; ++k) mx = fmaxf(mx, m[k]);
        float e[9], sum = 0.f;
        #pragma unroll
        for (int k = 0; k < 9; ++k) { e[k] = ex2_approx((m[k] - mx) * 1.4426950408889634f); sum += e[k]; }
        const float inv = __frcp_rn(sum);
        float pw[9];
        #pragma unroll
        for (int k = 0; k < 9; ++k) pw[k] = e[k] * inv;
        float a0 = 0.f, a1 = 0.f;
        #pragma unroll
        for (int k = 0; k < 9; ++k) {
          a0 = __fadd_rn(a0, __fmul_rn(pw[k], fw[0][k / 3][k % 3 + px]));
          a1 = __fadd_rn(a1, __fmul_rn(pw[k], fw[1][k / 3][k % 3 + px]));
        }
        r[px][0][j] = a0; r[px][1][j] = a1;
      }
    const long long row = static_cast<long long>(i) * kCuRate * W;
    #pragma unroll
    for (int c = 0; c < 2; ++c) {
      float* oc = o + static_cast<long long>(c) * kCuRate * kCuRate * HW + row;
      *reinterpret_cast<float4*>(oc) = make_float4(r[0][c][0], r[0][c][1], r[0][c][2], r[0][c][3]);
      *reinterpret_cast<float4*>(oc + kCuRate) = make_float4(r[1][c][0], r[1][c][1], r[1][c][2], r[1][c][3]);
    }
  }
}

}  // namespace ppm

extern "C" int ppm_convex_upsample(const float* flow, const void* mask, int mask_dtype, int N, int H, int W,
                                   int rate, float* out, void* stream) {
  using namespace ppm;
  PPM_CHECK(flow && mask && out, PPM_ERR_ARG, "ppm_convex_upsample: null pointer");
  PPM_CHECK(N > 0 && H > 0 && W > 0, PPM_ERR_ARG, "ppm_convex_upsample: bad shape");
  PPM_CHECK(rate == kCuRate, PPM_ERR_UNSUPPORTED, "ppm_convex_upsample: rate %d (only 4, the reference's call site)", rate);
  PPM_CHECK(mask_dtype == PPM_F32 || mask_dtype == PPM_BF16, PPM_ERR_ARG, "ppm_convex_upsample: bad dtype");
  PPM_CHECK(N <= 65535, PPM_ERR_UNSUPPORTED, "ppm_convex_upsample: N > 65535");
  PPM_CHECK((reinterpret_cast<uintptr_t>(out) & 15) == 0, PPM_ERR_ARG, "ppm_convex_upsample: out must be 16-byte aligned");
  dim3 grid((H * W + kCuThreads - 1) / kCuThreads, N);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (mask_dtype == PPM_BF16 && W % 2 == 0 && (reinterpret_cast<uintptr_t>(mask) & 3) == 0) {
    dim3 grid2((H * W / 2 + kCuThreads - 1) / kCuThreads, N);
    convex_upsample_bf16x2_kernel<<<grid2, kCuThreads, 0, st>>>(flow, static_cast<const __nv_bfloat16*>(mask), out, H, W);
  } else if (mask_dtype == PPM_BF16)
    convex_upsample_kernel<__nv_bfloat16><<<grid, kCuThreads, 0, st>>>(flow, static_cast<const __nv_bfloat16*>(mask), out, H, W);
  else
    convex_upsample_kernel<float><<<grid, kCuThreads, 0, st>>>(flow, static_cast<const float*>(mask), out, H, W);
  PPM_CUDA(cudaGetLastError());
  return PPM_OK;
}
