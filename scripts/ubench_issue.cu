// Issue-rate microbenchmark for the mixture-chain kernel design (sm_100a): how many warp-instructions per clock per
// SM sub-partition do FFMA, packed fma.rn.f32x2, IMAD, LOP3, MUFU and mixes of them sustain?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_issue scripts/ubench_issue.cu && ./ubench_issue
#include <cstdio>
#include <cuda_runtime.h>

constexpr int ITERS = 4096, UNROLL = 8;

__device__ __forceinline__ unsigned long long pack2(float a, float b) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long d;
  asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}

template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, float s, unsigned u, long long* clk) {
  float a[UNROLL];
  unsigned long long p[UNROLL];
  unsigned q[UNROLL];
#pragma unroll
  for (int i = 0; i < UNROLL; ++i) {
    a[i] = threadIdx.x * 0.001f + i;
    p[i] = pack2(a[i], a[i] + 0.5f);
    q[i] = threadIdx.x * 2654435761u + i;
  }
  const unsigned long long ps = pack2(s, s * 0.5f), pc = pack2(0.25f, 0.125f);
  long long t0 = clock64();
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < UNROLL; ++i) {
      if (MODE == 0) a[i] = fmaf(a[i], s, 0.25f * s);                 // FFMA 3-reg
      if (MODE == 1) p[i] = fma2(p[i], ps, pc);                       // FFMA2
      if (MODE == 2) { a[i] = fmaf(a[i], s, s); q[i] = (q[i] ^ u) + (q[i] >> 3); }   // FFMA + 2-3 ALU
      if (MODE == 3) a[i] = __expf(a[i]) * s;                         // MUFU.EX2 + FMUL(s)
      if (MODE == 4) q[i] = q[i] * u + (q[i] >> 7);                   // IMAD + SHF
      if (MODE == 5) { p[i] = fma2(p[i], ps, pc); q[i] = (q[i] ^ u) + (q[i] >> 3); }  // FFMA2 + ALU
      if (MODE == 6) { p[i] = fma2(p[i], ps, pc); q[i] = q[i] * u + 12345u; }         // FFMA2 + IMAD
      if (MODE == 7) { a[i] = fmaf(a[i], s, s); q[i] = q[i] * u + 12345u; }           // FFMA + IMAD
      if (MODE == 9) a[i] = fminf(fmaxf(a[i], -s), s) + 0.0f * s;        // 2 FMNMX (clamp)
      if (MODE == 10) { a[i] = fminf(fmaxf(a[i], -s), s); a[i] = fmaf(a[i], s, s); }   // 2 FMNMX + FFMA
      if (MODE == 11) q[i] = (q[i] & u) ^ (q[i] | 0x5555u);                // LOP3
      if (MODE == 12) { p[i] = fma2(p[i], ps, pc); a[i] = fminf(fmaxf(a[i], -s), s); }   // FFMA2 + 2 FMNMX
      if (MODE == 8) { unsigned long long w = (unsigned long long)q[i] * 0xD2511F53u; q[i] = (unsigned)(w >> 32) ^ (unsigned)w ^ u; }  // Philox-like
    }
  }
  long long t1 = clock64();
  float acc = 0;
#pragma unroll
  for (int i = 0; i < UNROLL; ++i) acc += a[i] + __uint_as_float((unsigned)(p[i] >> 32)) + __uint_as_float((unsigned)p[i]) + q[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}

template <int MODE>
void run(const char* name, int instr_per_inner, int blocks_per_sm) {
  int sms;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  float* out;
  long long* clk;
  cudaMalloc(&out, sizeof(float) * 256 * sms * blocks_per_sm);
  cudaMalloc(&clk, 8);
  k<MODE><<<sms * blocks_per_sm, 256>>>(out, 1.0001f, 0x9E3779B9u, clk);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaEventRecord(e0);
  k<MODE><<<sms * blocks_per_sm, 256>>>(out, 1.0001f, 0x9E3779B9u, clk);
  cudaEventRecord(e1);
  cudaDeviceSynchronize();
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  long long c;
  cudaMemcpy(&c, clk, 8, cudaMemcpyDeviceToHost);
  // warps per SMSP = blocks_per_sm * 8 / 4; inner statements executed per warp = ITERS * UNROLL
  const double warps_per_smsp = blocks_per_sm * 2.0;
  const double stmts = (double)ITERS * UNROLL * warps_per_smsp;
  printf("%-28s blocks/SM=%d  %8.3f ms  clk=%lld  statements/clk/SMSP=%.3f  (x%d SASS instr est. = %.3f warp-inst/clk/SMSP)\n",
         name, blocks_per_sm, ms, c, stmts / c, instr_per_inner, instr_per_inner * stmts / c);
  cudaFree(out);
  cudaFree(clk);
}

int main() {
  for (int b : {2, 4}) {
    run<0>("FFMA", 1, b);
    run<1>("FFMA2 (fma.rn.f32x2)", 1, b);
    run<2>("FFMA + XOR/SHF/ADD", 4, b);
    run<3>("MUFU.EX2 + 2 FMUL", 3, b);
    run<4>("IMAD + SHF", 2, b);
    run<5>("FFMA2 + XOR/SHF/ADD", 4, b);
    run<6>("FFMA2 + IMAD", 2, b);
    run<7>("FFMA + IMAD", 2, b);
    run<8>("IMAD.WIDE + 2 LOP3", 3, b);
    run<9>("2 FMNMX (clamp)", 2, b);
    run<10>("2 FMNMX + FFMA", 3, b);
    run<11>("LOP3", 1, b);
    run<12>("FFMA2 + 2 FMNMX", 3, b);
  }
  return 0;
}
