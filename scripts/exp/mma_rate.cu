// Experiment: sustained rate of back-to-back tcgen05.mma (bf16, M=128, K=16) issued by one thread per SM:
// SS vs TS form, N = 64 / 128 / 256, with and without concurrent TMEM loads by other warps.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I ppmstereo_b200/csrc scripts/exp/mma_rate.cu -o scripts/exp/mma_rate
#include <cstdio>
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>
#include "common.cuh"
using namespace ppm;
__device__ int g_random_data = 0;
__device__ __forceinline__ uint32_t rnd_bf16x2(uint32_t i) {
  uint32_t h = i * 2654435761u; h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
  // two bf16 in [-2, 2): sign + exponent 0x3f/0x3e..., random mantissa
  return (h & 0x807f807fu) | 0x3f003f00u | ((h >> 3) & 0x00800080u);
}

// mode: 0 = SS, 1 = TS.  smem operands: garbage but valid addresses (A 16 KB half-tile, B up to 32 KB).
template <int MODE, int N>
__global__ void __launch_bounds__(256, 1) rate_kernel(long long* out, int iters, int ld_warps) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  if (warp == 0) tmem_alloc<512>(&tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  // zero smem so the tensor pipe sees quiet data (power)
  for (int i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = g_random_data ? rnd_bf16x2(i + blockIdx.x * 7919u) : 0x3c003c00u;
  fence_proxy_async_smem();
  __syncthreads();
  if (warp == 0) {
    if (elect_one()) {
      const uint32_t idesc = umma_idesc(1, 0, 0, 128, N);
      const uint64_t a0 = umma_desc_sw128(smem_u32(smem), 16, 1024);
      const uint64_t b0 = umma_desc_sw128(smem_u32(smem + 32768), 16, 1024);
      long long t0 = clock64();
      for (int i = 0; i < iters; ++i) {
        #pragma unroll
        for (int k = 0; k < 8; ++k) {
          const uint64_t koff = (k & 3) * 2 + (k >> 2) * 1024;       // 32 B per K step, second half 16 KB on
          if (MODE == 0) umma_f16_ss(tmem + 256, a0 + koff, b0 + koff, idesc, 1);
          else umma_f16_ts(tmem + 256, tmem + k * 8, b0 + koff, idesc, 1);
        }
      }
      umma_commit(&bar);
      long long t1 = clock64();
      mbar_wait(&bar, 0);
      long long t2 = clock64();
      if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
    }
    __syncwarp();
  } else if (warp >= 4 && warp < 4 + ld_warps) {
    // concurrent TMEM reads (the softmax side's S loads) from columns 0..127
    uint32_t r[32];
    uint32_t acc = 0;
    const uint32_t base = tmem + (static_cast<uint32_t>((warp & 3) * 32) << 16);
    for (int i = 0; i < iters * 2; ++i) {
      #pragma unroll
      for (int c = 0; c < 4; ++c) {
        tmem_ld_x32(base + c * 32, r);
        tmem_ld_wait();
        acc += r[lane & 31];
      }
    }
    if (acc == 0x12345678u) out[7] = acc;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc<512>(tmem); }
}

// The attend kernel's P.V issue pattern: whole warp walks the loop; per chunk wait on an (already completed)
// mbarrier, tcgen05.fence::after_thread_sync, elected lane issues 2 TS MMAs.  VAR: 0 = wait + fence, 1 = wait only,
// 2 = neither, 3 = wait + fence, 4 MMAs per wait.
template <int VAR>
__global__ void __launch_bounds__(384, 1) chunk_kernel(long long* out, int iters, int poll_warps, int poll_lanes) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar, bar_done[4], bar_never;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    mbar_init(&bar_never, 1);
    for (int i = 0; i < 4; ++i) mbar_init(&bar_done[i], 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc<512>(&tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  for (int i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = g_random_data ? rnd_bf16x2(i + blockIdx.x * 7919u) : 0x3c003c00u;
  fence_proxy_async_smem();
  if (threadIdx.x == 0) for (int i = 0; i < 4; ++i) mbar_arrive(&bar_done[i]);     // phase 0 complete
  __syncthreads();
  if (warp == 0) {
    const bool leader = elect_one();
    const uint32_t idesc = umma_idesc(1, 0, 0, 128, 128);
    const uint64_t b0 = umma_desc_sw128(smem_u32(smem + 32768), 16, 1024);
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      #pragma unroll
      for (int c = 0; c < 4; ++c) {
        if (VAR == 0 || VAR == 1 || (VAR == 3 && (c & 1) == 0)) mbar_wait(&bar_done[c], 0);
        if (VAR == 0 || (VAR == 3 && (c & 1) == 0)) tc_fence_after();
        if (leader) {
          #pragma unroll
          for (int kk = 0; kk < 2; ++kk) {
            const int k = 2 * c + kk;
            const uint64_t koff = (k & 3) * 2 + (k >> 2) * 1024;
            umma_f16_ts(tmem + 256, tmem + k * 8, b0 + koff, idesc, 1);
          }
        }
        __syncwarp();
      }
    }
    if (leader) umma_commit(&bar);
    __syncwarp();
    long long t1 = clock64();
    mbar_wait(&bar, 0);
    long long t2 = clock64();
    if (blockIdx.x == 0 && leader) { out[0] = t1 - t0; out[1] = t2 - t0; }
    if (leader) mbar_arrive(&bar_never);                 // release the pollers
    __syncwarp();
  } else if (warp >= 4 && warp < 4 + poll_warps) {
    if ((threadIdx.x & 31) < poll_lanes) mbar_wait(&bar_never, 0);
    __syncwarp();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc<512>(tmem); }
}

template <int VAR>
void run_chunk(const char* name, long long* d_out, int poll_warps = 0, int poll_lanes = 32) {
  const int iters = 2000, smem = 100 * 1024;
  cudaFuncSetAttribute(chunk_kernel<VAR>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  chunk_kernel<VAR><<<148, 384, smem>>>(d_out, iters, poll_warps, poll_lanes);
  cudaError_t err = cudaDeviceSynchronize();
  long long h[2];
  cudaMemcpy(h, d_out, sizeof(h), cudaMemcpyDeviceToHost);
  printf("%-44s pollers %d warps x %2d lanes: issue %.1f clk/MMA, complete %.1f clk/MMA (ideal 64) [%s]\n", name, poll_warps, poll_lanes,
         h[0] / (iters * 8.0), h[1] / (iters * 8.0), cudaGetErrorString(err));
}

// One "step" of the attend kernel's issue order: PV_A QK_A PV_B QK_B, 8 MMAs each, with its commits.
// flags: 1 = commit after each group (4/step), 2 = two more commits (ring slots), 4 = per-chunk wait+fence in PV
__global__ void __launch_bounds__(128, 1) step_kernel(long long* out, int iters, int flags) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar, bar_done[4], bar_sink[6];
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    for (int i = 0; i < 4; ++i) mbar_init(&bar_done[i], 1);
    for (int i = 0; i < 6; ++i) mbar_init(&bar_sink[i], 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc<512>(&tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  for (int i = threadIdx.x; i < 192 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  fence_proxy_async_smem();
  if (threadIdx.x == 0) for (int i = 0; i < 4; ++i) mbar_arrive(&bar_done[i]);
  __syncthreads();
  if (warp == 0) {
    const bool leader = elect_one();
    const uint32_t idesc = umma_idesc(1, 0, 0, 128, 128);
    const uint64_t q0 = umma_desc_sw128(smem_u32(smem), 16, 1024);               // Q tiles: 2 x 32 KB
    const uint64_t r0 = umma_desc_sw128(smem_u32(smem + 65536), 16, 1024);       // ring: 4 x 32 KB
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      #pragma unroll
      for (int x = 0; x < 2; ++x) {
        #pragma unroll
        for (int c = 0; c < 4; ++c) {                                            // P.V of tile x
          if (flags & 4) { mbar_wait(&bar_done[c], 0); tc_fence_after(); }
          if (leader) {
            #pragma unroll
            for (int kk = 0; kk < 2; ++kk) {
              const int k = 2 * c + kk;
              const uint64_t koff = (k & 3) * 2 + (k >> 2) * 1024;
              umma_f16_ts(tmem + 256 + x * 128, tmem + x * 128 + k * 8, r0 + ((i & 1) * 2 + 1) * 2048 + koff, idesc, 1);
            }
          }
          __syncwarp();
        }
        if (leader && (flags & 1)) umma_commit(&bar_sink[x]);
        if (leader && (flags & 2) && x == 1) umma_commit(&bar_sink[4]);
        if (leader) {                                                             // Q K^T of tile x
          #pragma unroll
          for (int k = 0; k < 8; ++k) {
            const uint64_t koff = (k & 3) * 2 + (k >> 2) * 1024;
            umma_f16_ss(tmem + x * 128, q0 + x * 2048 + koff, r0 + ((i & 1) * 2) * 2048 + koff, idesc, k != 0);
          }
          if (flags & 1) umma_commit(&bar_sink[2 + x]);
          if ((flags & 2) && x == 1) umma_commit(&bar_sink[5]);
        }
        __syncwarp();
      }
    }
    if (leader) umma_commit(&bar);
    __syncwarp();
    long long t1 = clock64();
    mbar_wait(&bar, 0);
    long long t2 = clock64();
    if (blockIdx.x == 0 && leader) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc<512>(tmem); }
}

// The same step with the real barrier protocol: 8 "softmax" warps wait for S (bar_sf, committed by the issuer after
// each Q K^T) and arrive on the four P barriers (count 4); the issuer waits for them per chunk.
// FLAGS: 2 = one wait per P.V group (chunk 3's barrier), 4 = ring barriers + producer warp, 8 = setmaxnreg,
//        16 = two issuer warps (warp 1 -> tile A, warp 2 -> tile B; ring-empty barriers count 2)
template <int FLAGS>
__global__ void __launch_bounds__(384, 1) proto_kernel(long long* out, int iters) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar[2], bar_sf[2], bar_pf[2][4], bar_full[5], bar_empty[5];
  __shared__ uint32_t tmem_slot;
  constexpr bool kRing = FLAGS & 4, kTwo = FLAGS & 16;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(&bar[0], 1); mbar_init(&bar[1], 1);
    for (int i = 0; i < 5; ++i) { mbar_init(&bar_full[i], 1); mbar_init(&bar_empty[i], kTwo ? 2 : 1); }
    for (int x = 0; x < 2; ++x) {
      mbar_init(&bar_sf[x], 1);
      for (int c = 0; c < 4; ++c) mbar_init(&bar_pf[x][c], 4);
    }
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc<512>(&tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  for (int i = threadIdx.x; i < 192 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  fence_proxy_async_smem();
  __syncthreads();
  if (FLAGS & 8) {
    if (warp < 4) asm volatile("setmaxnreg.dec.sync.aligned.u32 88;");
    else asm volatile("setmaxnreg.inc.sync.aligned.u32 208;");
  }
  if (warp == 0 && kRing) {
    if (elect_one()) {
      int slot = 0; uint32_t ph = 1;
      for (int n = 0; n < 2 * iters + 1; ++n) {
        if (!(FLAGS & 64)) mbar_wait(&bar_empty[slot], ph);
        mbar_arrive(&bar_full[slot]);
        if (++slot == 5) { slot = 0; ph ^= 1; }
      }
    }
    __syncwarp();
  }
  if (warp == 1 || (kTwo && warp == 2)) {
    asm volatile(".reg .pred ppm_pr0, ppm_pr1;");
    const bool leader = elect_one();
    const uint32_t idesc = umma_idesc(1, 0, 0, 128, 128);
    const uint64_t q0 = umma_desc_sw128(smem_u32(smem), 16, 1024);
    const uint64_t r0 = umma_desc_sw128(smem_u32(smem + 65536), 16, 1024);
    int slot = 0; uint32_t ph = 0;
    long long t_pf = 0, t_full = 0, t_commit = 0, t_qk = 0, t_pv = 0, tt = 0;
    constexpr bool kTime = FLAGS & 128;
#define TIC() do { if (kTime) tt = clock64(); } while (0)
#define TOC(acc) do { if (kTime) acc += clock64() - tt; } while (0)
    auto advance = [&]() { if (++slot == 5) { slot = 0; ph ^= 1; } };
    auto issue_qk = [&](int x, int ks) {
      if (leader) {
        #pragma unroll
        for (int k = 0; k < 8; ++k) {
          const uint64_t koff = (k & 3) * 2 + (k >> 2) * 1024;
          umma_f16_ss(tmem + x * 128, q0 + x * 2048 + koff, r0 + (ks >= 3 ? ks - 3 : ks) * 2048 + koff, idesc, k != 0);
        }
        umma_commit(&bar_sf[x]);
      }
      __syncwarp();
    };
    auto probe = [&](uint64_t* b, uint32_t parity, int reg) {      // non-blocking; result stays in a predicate
      const uint32_t a = smem_u32(b);
      if (reg == 0) asm volatile("mbarrier.test_wait.parity.shared::cta.b64 ppm_pr0, [%0], %1;" ::"r"(a), "r"(parity) : "memory");
      else asm volatile("mbarrier.test_wait.parity.shared::cta.b64 ppm_pr1, [%0], %1;" ::"r"(a), "r"(parity) : "memory");
    };
    auto probed = [&](int reg) -> bool {
      uint32_t ok;
      if (reg == 0) asm volatile("selp.b32 %0, 1, 0, ppm_pr0;" : "=r"(ok));
      else asm volatile("selp.b32 %0, 1, 0, ppm_pr1;" : "=r"(ok));
      return ok != 0;
    };
    auto issue_pv = [&](int x, int vs, int j) {
      if (FLAGS & 256) {
        probe(&bar_pf[x][0], j & 1, 0);
        #pragma unroll
        for (int c = 0; c < 4; ++c) {
          if (!probed(c & 1)) mbar_wait(&bar_pf[x][c], j & 1);
          tc_fence_after();
          if (c < 3) probe(&bar_pf[x][c + 1], j & 1, (c + 1) & 1);
          if (leader) {
            #pragma unroll
            for (int kk = 0; kk < 2; ++kk) {
              const int k = 2 * c + kk;
              const uint64_t koff = (k & 3) * 2 + (k >> 2) * 1024;
              umma_f16_ts(tmem + 256 + x * 128, tmem + x * 128 + k * 8, r0 + (vs >= 3 ? vs - 3 : vs) * 2048 + koff, idesc, 1);
            }
          }
          __syncwarp();
        }
        return;
      }
      #pragma unroll
      for (int c = 0; c < 4; ++c) {
        TIC();
        if (!(FLAGS & 2)) { mbar_wait(&bar_pf[x][c], j & 1); tc_fence_after(); }
        else if (c == 0) { mbar_wait(&bar_pf[x][3], j & 1); tc_fence_after(); }
        TOC(t_pf);
        TIC();
        if (leader) {
          #pragma unroll
          for (int kk = 0; kk < 2; ++kk) {
            const int k = 2 * c + kk;
            const uint64_t koff = (k & 3) * 2 + (k >> 2) * 1024;
            umma_f16_ts(tmem + 256 + x * 128, tmem + x * 128 + k * 8, r0 + (vs >= 3 ? vs - 3 : vs) * 2048 + koff, idesc, 1);
          }
        }
        __syncwarp();
        TOC(t_pv);
      }
    };
    const int x0 = kTwo ? warp - 1 : 0, x1 = kTwo ? warp - 1 : 1;
    long long t0 = clock64();
    if (kRing && !(FLAGS & 32)) { mbar_wait(&bar_full[0], 0); tc_fence_after(); }
    for (int x = x0; x <= x1; ++x) issue_qk(x, 0);
    if (kRing && !(FLAGS & 64)) { if (leader) umma_commit(&bar_empty[0]); __syncwarp(); }
    advance();
    for (int j = 0; j < iters; ++j) {
      const int vs = slot;
      TIC();
      if (kRing && !(FLAGS & 32)) mbar_wait(&bar_full[vs], ph);
      TOC(t_full);
      advance();
      const int ks = slot; const uint32_t kph = ph;
      for (int x = x0; x <= x1; ++x) {
        issue_pv(x, vs, j);
        TIC();
        if (kRing && !(FLAGS & 64) && x == x1) { if (leader) umma_commit(&bar_empty[vs]); __syncwarp(); }
        TOC(t_commit);
        TIC();
        if (kRing && !(FLAGS & 32) && x == x0) { mbar_wait(&bar_full[ks], kph); tc_fence_after(); }
        TOC(t_full);
        TIC();
        issue_qk(x, ks);
        TOC(t_qk);
        TIC();
        if (kRing && !(FLAGS & 64) && x == x1) { if (leader) umma_commit(&bar_empty[ks]); __syncwarp(); }
        TOC(t_commit);
      }
      advance();
    }
    if (leader) umma_commit(&bar[warp - 1]);
    __syncwarp();
    long long t1 = clock64();
    mbar_wait(&bar[warp - 1], 0);
    long long t2 = clock64();
    if (blockIdx.x == 0 && leader && warp == 1) {
      out[0] = t1 - t0; out[1] = t2 - t0;
      out[2] = t_pf; out[3] = t_full; out[4] = t_commit; out[5] = t_qk; out[6] = t_pv;
    }
  } else if (warp >= 4) {
    const int x = (warp - 4) >> 2;
    for (int j = 0; j < iters + 1; ++j) {
      mbar_wait(&bar_sf[x], j & 1);
      tc_fence_after();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        #pragma unroll
        for (int c = 0; c < 4; ++c) mbar_arrive(&bar_pf[x][c]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc<512>(tmem); }
}

template <int FLAGS>
void run_proto(const char* name, long long* d_out) {
  const int iters = 1000, smem = 200 * 1024;
  cudaFuncSetAttribute(proto_kernel<FLAGS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  proto_kernel<FLAGS><<<148, 384, smem>>>(d_out, iters);
  cudaError_t err = cudaDeviceSynchronize();
  long long h[7];
  cudaMemcpy(h, d_out, sizeof(h), cudaMemcpyDeviceToHost);
  printf("%-64s: %.0f clk/step issue, %.0f complete (ideal 2048) [%s]\n", name, h[0] / double(iters), h[1] / double(iters),
         cudaGetErrorString(err));
  if (FLAGS & 128)
    printf("      per step: P waits %.0f, ring-full waits %.0f, ring-empty commits %.0f, QK issue %.0f, PV issue %.0f\n",
           h[2] / double(iters), h[3] / double(iters), h[4] / double(iters), h[5] / double(iters), h[6] / double(iters));
}

void run_step(const char* name, long long* d_out, int flags) {
  const int iters = 1000, smem = 200 * 1024;
  cudaFuncSetAttribute(step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  step_kernel<<<148, 128, smem>>>(d_out, iters, flags);
  cudaError_t err = cudaDeviceSynchronize();
  long long h[2];
  cudaMemcpy(h, d_out, sizeof(h), cudaMemcpyDeviceToHost);
  printf("%-60s: %.0f clk/step issue, %.0f complete (ideal 2048) [%s]\n", name, h[0] / double(iters), h[1] / double(iters),
         cudaGetErrorString(err));
}

template <int MODE, int N>
void run(const char* name, long long* d_out, int ld_warps) {
  const int iters = 2000, smem = 100 * 1024;
  cudaFuncSetAttribute(rate_kernel<MODE, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  rate_kernel<MODE, N><<<148, 256, smem>>>(d_out, iters, ld_warps);
  cudaDeviceSynchronize();
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  rate_kernel<MODE, N><<<148, 256, smem>>>(d_out, iters, ld_warps);
  cudaEventRecord(e1);
  cudaError_t err = cudaDeviceSynchronize();
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long h[2];
  cudaMemcpy(h, d_out, sizeof(h), cudaMemcpyDeviceToHost);
  const double n_mma = iters * 8.0;
  const double ideal = 128.0 * N * 16 * 2 / 8192.0;
  printf("%-28s ld_warps %d: issue %.1f clk/MMA, complete %.1f clk/MMA (ideal %.0f) -> %.0f TFLOP/s chip  [%s]\n", name,
         ld_warps, h[0] / n_mma, h[1] / n_mma, ideal, 148 * n_mma * 128.0 * N * 16 * 2 / (ms * 1e-3) / 1e12,
         cudaGetErrorString(err));
}

int main() {
  setvbuf(stdout, nullptr, _IONBF, 0);
  long long* d_out;
  cudaMalloc(&d_out, 128);
  for (int rd = 0; rd < 1; ++rd) {
  cudaMemcpyToSymbol(g_random_data, &rd, sizeof(int));
  printf("---- operand data: %s\n", rd ? "random bf16 in (-2, 2)" : "constant");
  for (int ld = 0; ld <= 0; ld += 4) {
    run<0, 64>("SS M128 N64 K16", d_out, ld);
    run<0, 128>("SS M128 N128 K16", d_out, ld);
    run<0, 256>("SS M128 N256 K16", d_out, ld);
    run<1, 64>("TS M128 N64 K16", d_out, ld);
    run<1, 128>("TS M128 N128 K16", d_out, ld);
    run<1, 256>("TS M128 N256 K16", d_out, ld);
  }
  run_proto<0>("protocol, per-chunk waits", d_out);
  run_proto<2>("protocol, one wait per P.V group", d_out);
  run_proto<4>("protocol + ring, per-chunk waits", d_out);
  run_proto<6>("protocol + ring, one wait per group", d_out);
  run_proto<4 + 256>("protocol + ring, per-chunk waits probed one chunk ahead", d_out);
  run_proto<0 + 256>("protocol, per-chunk waits probed one chunk ahead", d_out);
  run_proto<4 + 128>("protocol + ring, per-chunk waits, timed", d_out);
  run_proto<0 + 128>("protocol, per-chunk waits, timed", d_out);
  run_proto<6 + 128>("protocol + ring, one wait per group, timed", d_out);
  run_proto<4 + 32>("protocol + ring, issuer skips the full waits", d_out);
  run_proto<12>("protocol + ring + setmaxnreg, per-chunk waits", d_out);
  run_proto<20>("protocol + ring, TWO issuer warps, per-chunk waits", d_out);
  run_proto<22>("protocol + ring, TWO issuer warps, one wait per group", d_out);
  run_proto<16>("protocol, TWO issuer warps, per-chunk waits", d_out);
  run_step("step PV_A QK_A PV_B QK_B, no commits", d_out, 0);
  run_step("step, commit after each group", d_out, 1);
  run_step("step, group commits + 2 ring commits", d_out, 3);
  run_step("step, all commits + per-chunk wait/fence", d_out, 7);
  run_step("step, per-chunk wait/fence only", d_out, 4);
  run_chunk<2>("chunk loop, no wait, no fence", d_out);
  run_chunk<1>("chunk loop, mbar wait per 2 MMAs", d_out);
  run_chunk<0>("chunk loop, mbar wait + fence per 2 MMAs", d_out);
  run_chunk<3>("chunk loop, mbar wait + fence per 4 MMAs", d_out);
  run_chunk<0>("chunk loop, mbar wait + fence per 2 MMAs", d_out, 8, 32);
  run_chunk<0>("chunk loop, mbar wait + fence per 2 MMAs", d_out, 8, 1);
  run_chunk<2>("chunk loop, no wait, no fence", d_out, 8, 32);
  run_chunk<1>("chunk loop, mbar wait per 2 MMAs", d_out, 8, 32);
  }
  return 0;
}
