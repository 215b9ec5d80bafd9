// micro-benchmark: cycles per warp-level op for the primitives the LZ team encoder could be built on
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define FULL 0xffffffffu
__global__ void k(uint32_t* out, int mode, int iters, int warps_active)
{
  __shared__ uint32_t tab32[8192];
  uint16_t* tab16 = reinterpret_cast<uint16_t*>(tab32);
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 8192; i += blockDim.x) tab32[i] = 0;
  __syncthreads();
  if ((int)warp >= warps_active) return;
  uint32_t x = (threadIdx.x * 2654435761u) ^ (blockIdx.x * 40503u), acc = 0;
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
    x = x * 1664525u + 1013904223u;
    const uint32_t h = x >> 19; // 13 bits
    if (mode == 0) { // match_any on distinct-ish hashes
      acc += __match_any_sync(FULL, h);
    } else if (mode == 1) { // u16 table: probe, store, sync, read back, vote (the current window step)
      uint16_t* s = tab16 + h;
      uint32_t t = *s;
      *s = (uint16_t)i;
      __syncwarp();
      bool lost = *reinterpret_cast<volatile uint16_t*>(s) < (uint16_t)i;
      acc += t + __any_sync(FULL, lost);
    } else if (mode == 2) { // u32 table: probe, sync, atomicMax (no return), sync
      uint32_t t = tab32[h];
      __syncwarp();
      atomicMax(&tab32[h], (uint32_t)i * 32 + lane);
      __syncwarp();
      acc += t;
    } else if (mode == 3) { // u16 table: probe + plain store, no conflict handling (lower bound)
      uint16_t* s = tab16 + h;
      acc += *s;
      __syncwarp();
      *s = (uint16_t)i;
      __syncwarp();
    } else if (mode == 4) { // match_any with few distinct values
      acc += __match_any_sync(FULL, h & 3u);
    } else if (mode == 5) { // ballot + shfl baseline
      acc += __ballot_sync(FULL, h & 1) + __shfl_sync(FULL, h, (i & 31));
    }
  }
  long long t1 = clock64();
  if (lane == 0) out[blockIdx.x * 32 + warp] = (uint32_t)((t1 - t0) * 100 / iters);
  if (acc == 0x12345) out[0] = acc;
}
int main()
{
  uint32_t* d;
  cudaMalloc(&d, 1 << 20);
  const char* names[] = {"match_any distinct", "u16 probe+store+readback+vote", "u32 probe+atomicMax", "u16 probe+store", "match_any 4 values", "ballot+shfl"};
  for (int mode = 0; mode < 6; ++mode)
    for (int w : {1, 4, 8, 16, 32}) {
      k<<<148, 1024>>>(d, mode, 4096, w);
      cudaDeviceSynchronize();
      uint32_t h[32];
      cudaMemcpy(h, d, sizeof h, cudaMemcpyDeviceToHost);
      printf("%-32s warps/SM %2d: %7.2f cycles per op per warp  -> %.2f cycles per op per SM\n", names[mode], w, h[0] / 100.0, h[0] / 100.0 / w);
    }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
