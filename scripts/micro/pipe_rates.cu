// Micro-benchmark: per-SM issue rates of the instruction classes the line kernels lean on.
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 4096
__device__ __forceinline__ unsigned long long pack2(float a, float b) { return (unsigned long long)__float_as_uint(a) | ((unsigned long long)__float_as_uint(b) << 32); }
__device__ __forceinline__ float lo2(unsigned long long p) { return __uint_as_float((unsigned)p); }
template <int OP>
__global__ void k(double *out, double seed, float fseed) {
  double a0 = seed + threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3;
  float f0 = fseed + threadIdx.x, f1 = f0 + 1.f, f2 = f0 + 2.f, f3 = f0 + 3.f;
  int i0 = threadIdx.x, i1 = i0 + 1, i2 = i0 + 2, i3 = i0 + 3;
  for (int it = 0; it < ITERS; ++it) {
    if (OP == 0) { a0 = __dadd_rn(a0, seed); a1 = __dadd_rn(a1, seed); a2 = __dadd_rn(a2, seed); a3 = __dadd_rn(a3, seed); }
    if (OP == 1) { a0 = __dmul_rn(a0, seed); a1 = __dmul_rn(a1, seed); a2 = __dmul_rn(a2, seed); a3 = __dmul_rn(a3, seed); }
    if (OP == 2) { a0 = (double)f0; a1 = (double)f1; a2 = (double)f2; a3 = (double)f3; f0 = __fadd_rn(f0, (float)0.5f) + (float)(a0 > 1e30); f1 += (float)(a1 > 1e30); f2 += (float)(a2 > 1e30); f3 += (float)(a3 > 1e30); }
    if (OP == 3) { i0 = __double2int_rd(a0); i1 = __double2int_rd(a1); i2 = __double2int_rd(a2); i3 = __double2int_rd(a3); a0 += (i0 > 1 << 30); a1 += (i1 > 1 << 30); a2 += (i2 > 1 << 30); a3 += (i3 > 1 << 30); }
    if (OP == 4) { a0 = (double)i0; a1 = (double)i1; a2 = (double)i2; a3 = (double)i3; i0 += (a0 > 1e30); i1 += (a1 > 1e30); i2 += (a2 > 1e30); i3 += (a3 > 1e30); }
    if (OP == 5) { f0 = (float)a0; f1 = (float)a1; f2 = (float)a2; f3 = (float)a3; a0 += (f0 > 1e30f); a1 += (f1 > 1e30f); a2 += (f2 > 1e30f); a3 += (f3 > 1e30f); }
    if (OP == 6) { i0 = __float2int_rn(f0); i1 = __float2int_rn(f1); i2 = __float2int_rn(f2); i3 = __float2int_rn(f3); f0 += (i0 > 1 << 30); f1 += (i1 > 1 << 30); f2 += (i2 > 1 << 30); f3 += (i3 > 1 << 30); }
    if (OP == 7) { f0 = __fmul_rn(f0, fseed); f1 = __fmul_rn(f1, fseed); f2 = __fmul_rn(f2, fseed); f3 = __fmul_rn(f3, fseed); }
    if (OP == 8) { i0 = i0 * 3 + i1; i1 = i1 * 5 + i2; i2 = i2 * 7 + i3; i3 = i3 * 9 + i0; }
    if (OP == 9) { f0 = __fmaf_rn(f0, fseed, f1); f1 = __fmaf_rn(f1, fseed, f2); f2 = __fmaf_rn(f2, fseed, f3); f3 = __fmaf_rn(f3, fseed, f0); }
    if (OP == 10) {  // packed fp32 pairs (Blackwell FFMA2): two chains of (lo, hi) per instruction
      unsigned long long p0 = pack2(f0, f1), p1 = pack2(f2, f3), p2 = pack2(f1, f2), p3 = pack2(f3, f0), s2 = pack2(fseed, fseed);
      asm volatile("fma.rn.f32x2 %0, %0, %4, %1;\n\tfma.rn.f32x2 %1, %1, %4, %2;\n\tfma.rn.f32x2 %2, %2, %4, %3;\n\tfma.rn.f32x2 %3, %3, %4, %0;"
        : "+l"(p0), "+l"(p1), "+l"(p2), "+l"(p3) : "l"(s2));
      f0 = lo2(p0); f1 = lo2(p1); f2 = lo2(p2); f3 = lo2(p3);
    }
    if (OP == 11) {  // FFMA2 beside integer work: does the pair instruction leave issue slots for the ALU pipe?
      unsigned long long p0 = pack2(f0, f1), p1 = pack2(f2, f3), s2 = pack2(fseed, fseed);
      asm volatile("fma.rn.f32x2 %0, %0, %2, %1;\n\tfma.rn.f32x2 %1, %1, %2, %0;" : "+l"(p0), "+l"(p1) : "l"(s2));
      i0 = (i0 ^ i1) + 3; i1 = (i1 ^ i2) + 5;
      f0 = lo2(p0); f2 = lo2(p1);
    }
    if (OP == 12) { i0 = __dp4a(i0, i1, i2); i1 = __dp4a(i1, i2, i3); i2 = __dp4a(i2, i3, i0); i3 = __dp4a(i3, i0, i1); }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + f0 + f1 + f2 + f3 + i0 + i1 + i2 + i3;
}
template <int OP> void run(const char *name, double *d) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  int blocks = 148 * 4, threads = 512;
  k<OP><<<blocks, threads>>>(d, 1.0000001, 1.0000001f);
  cudaEventRecord(a); k<OP><<<blocks, threads>>>(d, 1.0000001, 1.0000001f); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  double warp_instr = (double)blocks * threads / 32 * ITERS * 4;
  double cyc = ms * 1e-3 * 1.965e9;
  printf("%-22s %.3f ms  %.2f target warp-instr/clk/SM (incl. the dependency-keeping extras)\n", name, ms, warp_instr / cyc / 148);
}
int main() {
  double *d; cudaMalloc(&d, 148 * 4 * 512 * 8);
  run<0>("DADD", d); run<1>("DMUL", d); run<2>("F2F.F64.F32 (+fadd)", d); run<3>("F2I.F64 (+dadd)", d); run<4>("I2F.F64 (+iadd)", d);
  run<5>("F2F.F32.F64 (+dadd)", d); run<6>("F2I.F32 (+fadd)", d); run<7>("FMUL", d); run<8>("IMAD", d);
  run<9>("FFMA", d); run<10>("FFMA2 (pairs)", d); run<11>("2 FFMA2 + 4 ALU", d); run<12>("DP4A", d);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
}
