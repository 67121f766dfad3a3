// Micro-benchmark: streaming write bandwidth on B200 for the page kernel's store patterns.
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void wr(float4* out, size_t n4) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  float4 v = make_float4(1.f, 2.f, 3.f, (float)i);
  if (MODE == 0) { for (; i < n4; i += stride) out[i] = v; }                 // coalesced, default
  if (MODE == 1) { for (; i < n4; i += stride) __stcs(out + i, v); }         // coalesced, .cs
  if (MODE == 2) { for (; 2*i+1 < n4; i += stride) { __stcs(out + 2*i, v); __stcs(out + 2*i + 1, v); } } // 2 x half-sector per thread
  if (MODE == 3) { for (; 2*i+1 < n4; i += stride) { out[2*i] = v; out[2*i+1] = v; } }
  if (MODE == 4) { for (; i < n4; i += stride) __stwt(out + i, v); }
}
template <int MODE>
__global__ void wr_tile(float4* out, int rows, int row4) {
  // page-kernel-like: block (64,4), each thread 8 rows, 2 float4 per row
  int x = (blockIdx.x * 64 + threadIdx.x);
  int y0 = blockIdx.y * 32 + threadIdx.y * 8;
  if (2*x+1 >= row4) return;
  float4 v = make_float4(1.f, 2.f, 3.f, (float)x);
  for (int yy = 0; yy < 8 && y0 + yy < rows; ++yy) {
    float4* r = out + (size_t)(y0 + yy) * row4;
    if (MODE == 0) { __stcs(r + 2*x, v); __stcs(r + 2*x + 1, v); }
    else { int w = (x & ~31) * 2; int l = x & 31; __stcs(r + w + l, v); __stcs(r + w + 32 + l, v); }
  }
}
int main() {
  size_t bytes = 281300000; size_t n4 = bytes / 16;
  float4* d; cudaMalloc(&d, bytes + (1<<20));
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  auto time = [&](auto f, const char* name) {
    for (int i = 0; i < 3; ++i) f();
    float best = 1e9;
    for (int i = 0; i < 10; ++i) { cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b); float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms; }
    printf("%-28s %.3f ms  %.1f GB/s\n", name, best, bytes / best / 1e6);
  };
  int blocks = 148 * 8;
  time([&]{ wr<0><<<blocks, 256>>>(d, n4); }, "coalesced default");
  time([&]{ wr<1><<<blocks, 256>>>(d, n4); }, "coalesced .cs");
  time([&]{ wr<4><<<blocks, 256>>>(d, n4); }, "coalesced .wt");
  time([&]{ wr<2><<<blocks, 256>>>(d, n4); }, "2x half-sector .cs");
  time([&]{ wr<3><<<blocks, 256>>>(d, n4); }, "2x half-sector default");
  time([&]{ wr<0><<<148*32, 256>>>(d, n4); }, "coalesced default 32/SM");
  time([&]{ cudaMemsetAsync(d, 0, bytes); }, "cudaMemset");
  int rows = 7250, row4 = 4850 / 2;
  dim3 blk(64, 4), grd((4850 + 255) / 256, (7250 + 31) / 32);
  time([&]{ wr_tile<0><<<grd, blk>>>(d, rows, row4); }, "tile half-sector .cs");
  time([&]{ wr_tile<1><<<grd, blk>>>(d, rows, row4); }, "tile full-sector .cs");
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
