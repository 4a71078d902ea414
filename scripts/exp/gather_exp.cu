// Experiment: DRAM/L2 traffic of random one-sector gathers under different load flavours.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 scripts/exp/gather_exp.cu -o scripts/exp/gather_exp
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
struct alignas(32) S8 { uint32_t w[8]; };
template <int MODE> __device__ __forceinline__ S8 ld(const void* p) {
  S8 r;
  if (MODE == 0) asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(r.w[0]),"=r"(r.w[1]),"=r"(r.w[2]),"=r"(r.w[3]),"=r"(r.w[4]),"=r"(r.w[5]),"=r"(r.w[6]),"=r"(r.w[7]) : "l"(p));
  if (MODE == 1) asm volatile("ld.global.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(r.w[0]),"=r"(r.w[1]),"=r"(r.w[2]),"=r"(r.w[3]),"=r"(r.w[4]),"=r"(r.w[5]),"=r"(r.w[6]),"=r"(r.w[7]) : "l"(p));
  if (MODE == 2) asm volatile("ld.global.cg.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(r.w[0]),"=r"(r.w[1]),"=r"(r.w[2]),"=r"(r.w[3]),"=r"(r.w[4]),"=r"(r.w[5]),"=r"(r.w[6]),"=r"(r.w[7]) : "l"(p));
  if (MODE == 3) asm volatile("ld.global.nc.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(r.w[0]),"=r"(r.w[1]),"=r"(r.w[2]),"=r"(r.w[3]),"=r"(r.w[4]),"=r"(r.w[5]),"=r"(r.w[6]),"=r"(r.w[7]) : "l"(p));
  if (MODE == 4) asm volatile("ld.global.cs.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(r.w[0]),"=r"(r.w[1]),"=r"(r.w[2]),"=r"(r.w[3]),"=r"(r.w[4]),"=r"(r.w[5]),"=r"(r.w[6]),"=r"(r.w[7]) : "l"(p));
  if (MODE == 5) asm volatile("ld.global.lu.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(r.w[0]),"=r"(r.w[1]),"=r"(r.w[2]),"=r"(r.w[3]),"=r"(r.w[4]),"=r"(r.w[5]),"=r"(r.w[6]),"=r"(r.w[7]) : "l"(p));
  return r;
}
// thread i reads the sector at row i (pitch bytes) + pseudo-random sector inside the row
template <int MODE> __global__ void gather(const uint8_t* base, long long n, int pitch, uint32_t* out) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t h = (uint32_t)i * 2654435761u; h ^= h >> 15;
  int sec = h % (pitch / 32);
  S8 v = ld<MODE>(base + i * pitch + sec * 32);
  uint32_t s = 0;
  for (int k = 0; k < 8; ++k) s ^= v.w[k];
  if (s == 0x12345678u) out[0] = s;
}
int main() {
  const long long n = 1177600; const int pitch = 640;
  uint8_t* d; uint32_t* o;
  cudaMalloc(&d, n * pitch); cudaMalloc(&o, 4); cudaMemset(d, 1, n * pitch);
  for (int rep = 0; rep < 2; ++rep) {
    gather<0><<<(n + 127) / 128, 128>>>(d, n, pitch, o);
    gather<1><<<(n + 127) / 128, 128>>>(d, n, pitch, o);
    gather<2><<<(n + 127) / 128, 128>>>(d, n, pitch, o);
    gather<3><<<(n + 127) / 128, 128>>>(d, n, pitch, o);
    gather<4><<<(n + 127) / 128, 128>>>(d, n, pitch, o);
    gather<5><<<(n + 127) / 128, 128>>>(d, n, pitch, o);
  }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
