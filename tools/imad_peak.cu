// Integer-pipe micro-benchmark for B200 (sm_100a): measures the warp-instruction issue rate of the
// instructions the Montgomery multiplier is made of, per SM per clock, with clock64() inside the
// kernel (so DVFS does not matter) and with CUDA events (absolute rate at the clocks the GPU ran at).
// The best 32x32->64 multiply-accumulate rate measured here is the denominator of every
// "fraction of the integer-multiply roofline" figure in bench.py / DESIGN.md.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
static int ITERS = 4096;   // argv[1] overrides (longer runs show the sustained clock)
#define CHK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1;}}while(0)

template<int MODE> __global__ void __launch_bounds__(256) k(uint32_t* out, uint32_t seed, long long* cyc, int ITERS) {
  uint32_t a[8], b = seed ^ threadIdx.x, c = seed * 3 + threadIdx.x;
  #pragma unroll
  for (int i = 0; i < 8; i++) a[i] = seed + i * 7 + threadIdx.x;
  uint64_t w[8];
  #pragma unroll
  for (int i = 0; i < 8; i++) w[i] = (uint64_t)a[i] * 0x9E3779B97F4A7C15ull;
  long long t0 = clock64();
  #pragma unroll 1
  for (int it = 0; it < ITERS; it++) {
    if (MODE == 0) {        // IMAD (mad.lo.u32), 8 independent chains
      #pragma unroll
      for (int i = 0; i < 8; i++) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
    } else if (MODE == 1) { // IMAD.HI
      #pragma unroll
      for (int i = 0; i < 8; i++) asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
    } else if (MODE == 2) { // IMAD.WIDE.U32 (64-bit accumulate), 8 independent chains
      #pragma unroll
      for (int i = 0; i < 8; i++) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"(b), "r"(c));
    } else if (MODE == 3) { // the multiplier's row: carry-chained lo/hi pairs -> 4x IMAD.WIDE.U32(.X) + IADD3.X, two independent rows
      asm volatile("mad.lo.cc.u32 %0, %8, %9, %0; madc.hi.cc.u32 %1, %8, %9, %1;"
                   "madc.lo.cc.u32 %2, %8, %10, %2; madc.hi.cc.u32 %3, %8, %10, %3;"
                   "madc.lo.cc.u32 %4, %9, %10, %4; madc.hi.cc.u32 %5, %9, %10, %5;"
                   "madc.lo.cc.u32 %6, %10, %10, %6; madc.hi.u32 %7, %10, %10, %7;"
                   : "+r"(a[0]), "+r"(a[1]), "+r"(a[2]), "+r"(a[3]), "+r"(a[4]), "+r"(a[5]), "+r"(a[6]), "+r"(a[7]) : "r"(b), "r"(c), "r"(seed));
      uint32_t* q = (uint32_t*)w;
      asm volatile("mad.lo.cc.u32 %0, %8, %9, %0; madc.hi.cc.u32 %1, %8, %9, %1;"
                   "madc.lo.cc.u32 %2, %8, %10, %2; madc.hi.cc.u32 %3, %8, %10, %3;"
                   "madc.lo.cc.u32 %4, %9, %10, %4; madc.hi.cc.u32 %5, %9, %10, %5;"
                   "madc.lo.cc.u32 %6, %10, %10, %6; madc.hi.u32 %7, %10, %10, %7;"
                   : "+r"(q[0]), "+r"(q[1]), "+r"(q[2]), "+r"(q[3]), "+r"(q[4]), "+r"(q[5]), "+r"(q[6]), "+r"(q[7]) : "r"(b), "r"(c), "r"(seed));
    } else if (MODE == 4) { // IADD3 (ALU pipe)
      #pragma unroll
      for (int i = 0; i < 8; i++) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
    } else if (MODE == 5) { // 8 IMAD.WIDE + 8 IADD3 interleaved: do the two pipes co-issue?
      uint32_t* q = (uint32_t*)(w + 4);
      #pragma unroll
      for (int i = 0; i < 4; i++) {
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"(b), "r"(c));
        asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"(c), "r"(b));
        asm volatile("add.u32 %0, %0, %1;" : "+r"(q[i]) : "r"(c));
      }
    } else if (MODE == 6) { // 8 IMAD.WIDE + 4 IADD3
      #pragma unroll
      for (int i = 0; i < 4; i++) {
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"(b), "r"(c));
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i+4]) : "r"(c), "r"(b));
        asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
      }
    } else if (MODE == 7) { // carry adds IADD3.X chain
      asm volatile("add.cc.u32 %0, %0, %8; addc.cc.u32 %1, %1, %8; addc.cc.u32 %2, %2, %8; addc.cc.u32 %3, %3, %8;"
                   "addc.cc.u32 %4, %4, %8; addc.cc.u32 %5, %5, %8; addc.cc.u32 %6, %6, %8; addc.u32 %7, %7, %8;"
                   : "+r"(a[0]), "+r"(a[1]), "+r"(a[2]), "+r"(a[3]), "+r"(a[4]), "+r"(a[5]), "+r"(a[6]), "+r"(a[7]) : "r"(b));
    } else if (MODE == 9) { // what ptxas makes of the 29-bit column sums: IMAD.WIDE.U32 with RZ accumulate + 3-input IADD3 / IADD3.X
      #pragma unroll
      for (int i = 0; i < 4; i++) {
        uint64_t p0, p1;
        asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(p0) : "r"(b + i), "r"(c));
        asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(p1) : "r"(c + i), "r"(b));
        w[i] += p0 + p1;
      }
    } else if (MODE == 10) { // products only (RZ accumulate), folded with LOP3 so they stay live
      #pragma unroll
      for (int i = 0; i < 8; i++) {
        uint64_t p0;
        asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(p0) : "r"(b + i), "r"(c));
        w[i] ^= p0;
      }
    } else if (MODE == 11) { // DFMA, 8 independent chains
      double* d = (double*)w; double x = __longlong_as_double(0x3ff0000000000001ll + b), y = 1e-9 * c;
      #pragma unroll
      for (int i = 0; i < 8; i++) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(x), "d"(y));
    } else if (MODE == 12) { // 4 DFMA + 8 IMAD.WIDE: do the FP64 and integer-multiply pipes overlap?
      double* d = (double*)(w + 4); double x = __longlong_as_double(0x3ff0000000000001ll + b), y = 1e-9 * c;
      #pragma unroll
      for (int i = 0; i < 4; i++) {
        asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(x), "d"(y));
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"(b), "r"(c));
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"(c), "r"(b));
      }
    } else if (MODE == 13) { // IMAD.WIDE with accumulate, distinct operand registers per chain (no operand-reuse cache help)
      #pragma unroll
      for (int i = 0; i < 8; i++) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"(a[i]), "r"(a[(i + 3) & 7]));
    } else if (MODE == 14) { // 16 products per 8 accumulators, RZ form + adds, distinct operands (closest to acc_mac)
      #pragma unroll
      for (int i = 0; i < 8; i++) {
        uint64_t p0, p1;
        asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(p0) : "r"(a[i]), "r"(a[(i + 3) & 7]));
        asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(p1) : "r"(a[(i + 1) & 7]), "r"(a[(i + 5) & 7]));
        w[i] += p0 + p1;
      }
    } else if (MODE == 8) { // 8 IMAD.WIDE + 16 IADD3
      uint32_t* q = (uint32_t*)(w + 4);
      #pragma unroll
      for (int i = 0; i < 4; i++) {
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"(b), "r"(c));
        asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
        asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i+4]) : "r"(b));
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"(c), "r"(b));
        asm volatile("add.u32 %0, %0, %1;" : "+r"(q[i]) : "r"(c));
        asm volatile("add.u32 %0, %0, %1;" : "+r"(q[i+4]) : "r"(c));
      }
    }
  }
  long long t1 = clock64();
  uint32_t r = 0;
  #pragma unroll
  for (int i = 0; i < 8; i++) r ^= a[i] ^ (uint32_t)w[i] ^ (uint32_t)(w[i] >> 32);
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template<int MODE> int run(const char* name, int instr_per_iter, int mul_per_iter, int ctas_per_sm, int nsm) {
  int grid = nsm * ctas_per_sm, block = 256;
  uint32_t* out; long long* cyc;
  CHK(cudaMalloc(&out, (size_t)grid * block * 4)); CHK(cudaMalloc(&cyc, grid * 8));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<grid, block>>>(out, 12345u, cyc, ITERS); CHK(cudaDeviceSynchronize());
  CHK(cudaEventRecord(e0));
  k<MODE><<<grid, block>>>(out, 12345u, cyc, ITERS);
  CHK(cudaEventRecord(e1)); CHK(cudaDeviceSynchronize());
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long* h = (long long*)malloc(grid * 8); cudaMemcpy(h, cyc, grid * 8, cudaMemcpyDeviceToHost);
  double avg = 0; for (int i = 0; i < grid; i++) avg += h[i]; avg /= grid;
  double warps_per_sm = ctas_per_sm * (block / 32.0);
  double winstr_per_clk_sm = warps_per_sm * (double)ITERS * instr_per_iter / avg;
  double total_mul = (double)grid * block * (double)ITERS * mul_per_iter;
  printf("{\"mode\": \"%s\", \"ctas_per_sm\": %d, \"warp_instr_per_clk_per_sm\": %.3f, \"lane_ops_per_clk_per_sm\": %.1f, \"mul32x32_per_s\": %.4e, \"ms\": %.3f, \"implied_sm_mhz\": %.0f}\n",
         name, ctas_per_sm, winstr_per_clk_sm, winstr_per_clk_sm * 32, total_mul / (ms * 1e-3), ms, avg / (ms * 1e-3) / 1e6);
  cudaFree(out); cudaFree(cyc); free(h); return 0;
}
int main(int argc, char** argv) {
  if (argc > 1) ITERS = atoi(argv[1]);
  cudaDeviceProp p; CHK(cudaGetDeviceProperties(&p, 0));
  int nsm = p.multiProcessorCount; printf("{\"gpu\": \"%s\", \"sms\": %d, \"clock_khz\": %d}\n", p.name, nsm, p.clockRate);
  for (int c = 1; c <= 8; c *= 2) {
    run<0>("imad_lo", 8, 8, c, nsm); run<1>("imad_hi", 8, 8, c, nsm); run<2>("imad_wide", 8, 8, c, nsm);
    run<3>("imad_wide_carry_rows", 8, 8, c, nsm); run<4>("iadd3", 8, 0, c, nsm);
    run<5>("imad_wide8+iadd8", 16, 8, c, nsm); run<6>("imad_wide8+iadd4", 12, 8, c, nsm); run<7>("iadd3x_chain", 8, 0, c, nsm);
    run<8>("imad_wide8+iadd16", 24, 8, c, nsm);
    run<9>("imad_wide_rz8+iadd3x4pairs", 16, 8, c, nsm); run<10>("imad_wide_rz8+lop", 24, 8, c, nsm); run<11>("dfma8", 8, 8, c, nsm);
    run<12>("dfma4+imad_wide8", 12, 8, c, nsm); run<13>("imad_wide_acc_distinct_ops", 8, 8, c, nsm); run<14>("imad_wide_rz16+iadd3x8pairs", 32, 16, c, nsm);
  }
  return 0;
}
