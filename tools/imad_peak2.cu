// Integer-multiply roofline for B200 (sm_100a), second version.  tools/imad_peak.cu turned out to measure the wrong
// thing: with loop-invariant operands ptxas hoists the product and leaves 64-bit ADDS in the loop, and PTX
// `mad.wide.u32` with a live addend is split into IMAD.WIDE.U32 (RZ) + IADD3 + IADD3.X.  Here every mode's operands
// change each iteration (they are taken from the accumulators themselves, which costs no instruction), and the SASS
// of each loop is checked by tools/check_imad_sass.py: the loop of mode M must contain exactly the instruction mix
// named below.  Rates are warp-instructions per clock per SM (clock64 inside the kernel) and absolute per second
// (CUDA events over a run long enough for the clocks to settle).
//   0 wide_acc8     8 x IMAD.WIDE.U32 Rd, Ra, Rb, Rd   (accumulating form; 8 independent chains)
//   1 wide_acc16   16 x the same (16 chains)
//   2 mac81_acc    the 9x9-limb product block of fr29.cuh (81 accumulating IMAD.WIDE, 17 column accumulators)
//   3 mac81_split  the same block as ptxas emits it for inline-asm mad.wide.u32: IMAD.WIDE (RZ) + IADD3 / IADD3.X
//   4 imad_lo8      8 x IMAD (32-bit)
//   5 8 accumulating IMAD.WIDE + 16 IMAD.IADD (ptxas moves plain adds to the multiplier pipe) + 8 LOP3
//   6 wide_acc8+lds     8 accumulating IMAD.WIDE + 1 LDS.128 + 1 uniform LDG.128: the dot-product loop's memory mix
//   7 dfma8         8 x DFMA
//   8 dfma4+wide8   4 DFMA + 8 accumulating IMAD.WIDE: FP64 and integer-multiply pipes in parallel?
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
#define CHK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1;}}while(0)
#define LO(x) ((uint32_t)(x))
#define HI(x) ((uint32_t)((x) >> 32))

__device__ __forceinline__ uint64_t madw_asm(uint32_t a, uint32_t b, uint64_t c) { uint64_t r; asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(r) : "r"(a), "r"(b), "l"(c)); return r; }

template<int MODE> __global__ void __launch_bounds__(256) k(uint32_t* out, const uint32_t* in, long long* cyc, int iters) {
  __shared__ uint4 sh[256];
  uint64_t w[17];
#pragma unroll
  for (int i = 0; i < 17; i++) w[i] = reinterpret_cast<const uint64_t*>(in)[(threadIdx.x + 31 * i) & 511];
  uint32_t a[9], b[9];
#pragma unroll
  for (int i = 0; i < 9; i++) { a[i] = in[(threadIdx.x + 40 + i) & 1023] & 0x1fffffff; b[i] = in[(threadIdx.x + 80 + i) & 1023] & 0x1fffffff; }
  sh[threadIdx.x] = make_uint4(a[0], a[1], a[2], a[3]);
  uint32_t d32[8];
#pragma unroll
  for (int i = 0; i < 8; i++) d32[i] = in[(threadIdx.x + 200 + i) & 1023];
  double d[4]; 
#pragma unroll
  for (int i = 0; i < 4; i++) d[i] = 1.0 + 1e-9 * a[i];
  double dx = 1.0 + 1e-12 * b[0], dy = 1e-9 * b[1];
  const uint4* gq = reinterpret_cast<const uint4*>(in);
  __syncthreads();
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < iters; it++) {
    if (MODE == 0) {          // 16 accumulating IMAD.WIDE; x[] changes every iteration (8 LOP3 on the ALU pipe) so nothing can be hoisted
#pragma unroll
      for (int i = 0; i < 8; i++) { w[i] += (uint64_t)a[i] * b[i]; w[i + 8] += (uint64_t)a[i] * b[(i + 1) & 7]; a[i] ^= it; }
    } else if (MODE == 1) {   // 16 accumulating IMAD.WIDE, 4 LOP3
#pragma unroll
      for (int i = 0; i < 16; i++) w[i] += (uint64_t)a[i & 3] * b[i >> 2 | (i & 3) << 1];
#pragma unroll
      for (int i = 0; i < 4; i++) a[i] ^= it;
    } else if (MODE == 2) {   // the dot-product loop as the Poseidon kernel runs it: per-thread operand from shared memory, warp-uniform operand through L1
      uint4 s0 = sh[(threadIdx.x + it) & 255], s1 = sh[(threadIdx.x + it + 7) & 255]; uint32_t s2 = reinterpret_cast<uint32_t*>(sh)[(threadIdx.x + it * 5) & 1023];
      uint4 g0 = __ldg(gq + ((3 * it) & 255)), g1 = __ldg(gq + ((3 * it + 1) & 255)); uint32_t g2 = __ldg(in + ((12 * it + 8) & 1023));
      a[0] = s0.x; a[1] = s0.y; a[2] = s0.z; a[3] = s0.w; a[4] = s1.x; a[5] = s1.y; a[6] = s1.z; a[7] = s1.w; a[8] = s2;
      b[0] = g0.x; b[1] = g0.y; b[2] = g0.z; b[3] = g0.w; b[4] = g1.x; b[5] = g1.y; b[6] = g1.z; b[7] = g1.w; b[8] = g2;
#pragma unroll
      for (int i = 0; i < 9; i++)
#pragma unroll
        for (int j = 0; j < 9; j++) w[i + j] += (uint64_t)a[i] * b[j];
    } else if (MODE == 3) {   // same loop, multiply-accumulate spelled as inline-asm mad.wide.u32 (what round 1a shipped)
      uint4 s0 = sh[(threadIdx.x + it) & 255], s1 = sh[(threadIdx.x + it + 7) & 255]; uint32_t s2 = reinterpret_cast<uint32_t*>(sh)[(threadIdx.x + it * 5) & 1023];
      uint4 g0 = __ldg(gq + ((3 * it) & 255)), g1 = __ldg(gq + ((3 * it + 1) & 255)); uint32_t g2 = __ldg(in + ((12 * it + 8) & 1023));
      a[0] = s0.x; a[1] = s0.y; a[2] = s0.z; a[3] = s0.w; a[4] = s1.x; a[5] = s1.y; a[6] = s1.z; a[7] = s1.w; a[8] = s2;
      b[0] = g0.x; b[1] = g0.y; b[2] = g0.z; b[3] = g0.w; b[4] = g1.x; b[5] = g1.y; b[6] = g1.z; b[7] = g1.w; b[8] = g2;
#pragma unroll
      for (int i = 0; i < 9; i++)
#pragma unroll
        for (int j = 0; j < 9; j++) w[i + j] = madw_asm(a[i], b[j], w[i + j]);
    } else if (MODE == 4) {
#pragma unroll
      for (int i = 0; i < 8; i++) a[i] = a[i] * b[i] + a[(i + 1) & 7];
    } else if (MODE == 5) {
#pragma unroll
      for (int i = 0; i < 8; i++) { w[i] += (uint64_t)a[i] * b[i]; a[i] ^= it; asm volatile("add.u32 %0, %0, %1;" : "+r"(b[i]) : "r"(a[i])); asm volatile("add.u32 %0, %0, %1;" : "+r"(d32[i]) : "r"(a[i])); }
    } else if (MODE == 6) {
      uint4 s = sh[(threadIdx.x + it) & 255];
      uint4 g = __ldg(gq + (it & 63));
      w[0] += (uint64_t)s.x * g.x; w[1] += (uint64_t)s.y * g.y; w[2] += (uint64_t)s.z * g.z; w[3] += (uint64_t)s.w * g.w;
      w[4] += (uint64_t)s.x * g.y; w[5] += (uint64_t)s.y * g.z; w[6] += (uint64_t)s.z * g.w; w[7] += (uint64_t)s.w * g.x;
      w[8] += (uint64_t)s.x * g.z; w[9] += (uint64_t)s.y * g.w; w[10] += (uint64_t)s.z * g.x; w[11] += (uint64_t)s.w * g.y;
      w[12] += (uint64_t)s.x * g.w; w[13] += (uint64_t)s.y * g.x; w[14] += (uint64_t)s.z * g.y; w[15] += (uint64_t)s.w * g.z;
    } else if (MODE == 7) {
#pragma unroll
      for (int i = 0; i < 4; i++) { d[i] = fma(d[i], dx, dy); }
#pragma unroll
      for (int i = 0; i < 4; i++) { d[i] = fma(d[i], dy, dx); }
    } else if (MODE == 8) {
#pragma unroll
      for (int i = 0; i < 4; i++) { d[i] = fma(d[i], dx, dy); w[2 * i] += (uint64_t)a[i] * b[2 * i]; w[2 * i + 1] += (uint64_t)a[i] * b[2 * i + 1]; a[i] ^= it; }
    }
  }
  long long t1 = clock64();
  uint32_t r = 0;
#pragma unroll
  for (int i = 0; i < 17; i++) r ^= LO(w[i]) ^ HI(w[i]);
#pragma unroll
  for (int i = 0; i < 9; i++) r ^= a[i] ^ b[i];
#pragma unroll
  for (int i = 0; i < 4; i++) r ^= (uint32_t)__double_as_longlong(d[i]);
#pragma unroll
  for (int i = 0; i < 8; i++) r ^= d32[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template<int MODE> int run(const char* name, int instr_per_iter, int mul_per_iter, int ctas_per_sm, int nsm, int iters, const uint32_t* in) {
  int grid = nsm * ctas_per_sm, block = 256;
  uint32_t* out; long long* cyc;
  CHK(cudaMalloc(&out, (size_t)grid * block * 4)); CHK(cudaMalloc(&cyc, grid * 8));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<grid, block>>>(out, in, cyc, iters); CHK(cudaDeviceSynchronize());
  CHK(cudaEventRecord(e0));
  k<MODE><<<grid, block>>>(out, in, cyc, iters);
  CHK(cudaEventRecord(e1)); CHK(cudaDeviceSynchronize());
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long* h = (long long*)malloc(grid * 8); cudaMemcpy(h, cyc, grid * 8, cudaMemcpyDeviceToHost);
  double avg = 0; for (int i = 0; i < grid; i++) avg += h[i]; avg /= grid;
  double warps_per_sm = ctas_per_sm * (block / 32.0);
  double w = warps_per_sm * (double)iters * instr_per_iter / avg;
  double total_mul = (double)grid * block * (double)iters * mul_per_iter;
  printf("{\"mode\": \"%s\", \"warps_per_sm\": %d, \"counted_instr_per_iter\": %d, \"warp_instr_per_clk_per_sm\": %.3f, \"mul_per_clk_per_sm\": %.1f, \"mul32x32_per_s\": %.4e, \"ms\": %.3f, \"implied_sm_mhz\": %.0f}\n",
         name, (int)warps_per_sm, instr_per_iter, w, warps_per_sm * 32.0 * iters * mul_per_iter / avg, total_mul / (ms * 1e-3), ms, avg / (ms * 1e-3) / 1e6);
  cudaFree(out); cudaFree(cyc); free(h); return 0;
}
int main(int argc, char** argv) {
  int iters = argc > 1 ? atoi(argv[1]) : 16384;
  cudaDeviceProp p; CHK(cudaGetDeviceProperties(&p, 0));
  int nsm = p.multiProcessorCount; printf("{\"gpu\": \"%s\", \"sms\": %d, \"iters\": %d}\n", p.name, nsm, iters);
  uint32_t* in; CHK(cudaMalloc(&in, 4096 * 4));
  uint32_t hin[4096]; for (int i = 0; i < 4096; i++) hin[i] = 0x9E3779B9u * (i + 1) ^ (i << 7);
  CHK(cudaMemcpy(in, hin, sizeof(hin), cudaMemcpyHostToDevice));
  for (int c = 1; c <= 8; c *= 2) {
    run<0>("wide_acc16+lop8", 24, 16, c, nsm, iters, in); run<1>("wide_acc16+lop4", 20, 16, c, nsm, iters, in);
    run<2>("dotloop_mac81_acc", 81, 81, c, nsm, iters / 8, in); run<3>("dotloop_mac81_split", 81, 81, c, nsm, iters / 8, in);
    run<4>("imad_lo8", 8, 8, c, nsm, iters, in); run<5>("wide_acc8+imad_iadd16+lop8", 32, 8, c, nsm, iters, in);
    run<6>("wide_acc16+lds128+ldg128", 18, 16, c, nsm, iters, in); run<7>("dfma8", 8, 8, c, nsm, iters, in); run<8>("dfma4+wide8+lop4", 16, 8, c, nsm, iters, in);
  }
  return 0;
}
