// Is a 5 x 52-bit limb field product on the FP64 pipe (DFMA hi/lo split, integer accumulation of the raw bit patterns:
// Emmart, Zheng, Weems, ARITH 2018) faster on B200 than the 9 x 29-bit IMAD.WIDE block of fr29.cuh?  One "product" = the
// full schoolbook partial-product block of one Fr multiplication, accumulated lazily into column accumulators:
//   mode 0  imad81   : 81 accumulating IMAD.WIDE.U32 into 17 64-bit columns            (what poseidon.cuh runs today)
//   mode 1  dfma25   : 25 x (DFMA.RZ hi, DADD, DFMA.RZ lo) + 50 64-bit integer adds into 10 columns
// Both loops fetch a per-thread operand from shared memory and a warp-uniform operand through L1 each iteration, as the
// interpreter's term loop does.  Reported: products per clock per SM (clock64) and per second (CUDA events).
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
#define CHK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1;}}while(0)

template<int MODE> __global__ void __launch_bounds__(160) k(uint32_t* out, const uint32_t* in, long long* cyc, int iters) {
  __shared__ uint4 sh[160 * 3];
  for (int i = threadIdx.x; i < 480; i += blockDim.x) sh[i] = reinterpret_cast<const uint4*>(in)[i & 255];
  __syncthreads();
  const uint4* gq = reinterpret_cast<const uint4*>(in);
  uint32_t r = 0;
  long long t0 = clock64();
  if (MODE == 0) {
    uint64_t w[17];
#pragma unroll
    for (int i = 0; i < 17; i++) w[i] = i;
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
      uint4 s0 = sh[(threadIdx.x + it) % 160], s1 = sh[160 + (threadIdx.x + it + 7) % 160]; uint32_t s2 = reinterpret_cast<uint32_t*>(sh)[(threadIdx.x + it * 5) & 1023];
      uint4 g0 = __ldg(gq + ((3 * it) & 255)), g1 = __ldg(gq + ((3 * it + 1) & 255)); uint32_t g2 = __ldg(in + ((12 * it + 8) & 1023));
      uint32_t a[9] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w, s2}, b[9] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w, g2};
#pragma unroll
      for (int i = 0; i < 9; i++)
#pragma unroll
        for (int j = 0; j < 9; j++) w[i + j] += (uint64_t)a[i] * b[j];
    }
#pragma unroll
    for (int i = 0; i < 17; i++) r ^= (uint32_t)w[i] ^ (uint32_t)(w[i] >> 32);
  } else {
    long long w[10];
#pragma unroll
    for (int i = 0; i < 10; i++) w[i] = i;
    const double c1 = 20282409603651670423947251286016.0;          // 2^104
    const double c2 = 20282409603651674927546878656512.0;          // 2^104 + 2^52
    const double* shd = reinterpret_cast<const double*>(sh);
    const double2* gd = reinterpret_cast<const double2*>(in);
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
      // operands: 5 doubles each (limbs < 2^52 held as exact doubles); the masks keep the loaded bit patterns finite and < 2^52
      double a[5], b[5];
      const int ta = (threadIdx.x + it) % 160;
      double2 a01 = reinterpret_cast<const double2*>(shd)[ta], a23 = reinterpret_cast<const double2*>(shd)[160 + (ta + 7) % 160]; double a4 = shd[640 + (threadIdx.x + it * 5) % 160];
      double2 b01 = __ldg(gd + ((3 * it) & 255)), b23 = __ldg(gd + ((3 * it + 1) & 255)); double b4 = __ldg(reinterpret_cast<const double*>(in) + ((6 * it + 4) & 511));
      a[0] = a01.x; a[1] = a01.y; a[2] = a23.x; a[3] = a23.y; a[4] = a4; b[0] = b01.x; b[1] = b01.y; b[2] = b23.x; b[3] = b23.y; b[4] = b4;
#pragma unroll
      for (int i = 0; i < 5; i++) {   // force the values into [2^51, 2^52) as doubles without a conversion: exponent bits set, mantissa from the data
        a[i] = __longlong_as_double((__double_as_longlong(a[i]) & 0x000fffffffffffffLL) | 0x4320000000000000LL);
        b[i] = __longlong_as_double((__double_as_longlong(b[i]) & 0x000fffffffffffffLL) | 0x4320000000000000LL);
      }
#pragma unroll
      for (int i = 0; i < 5; i++)
#pragma unroll
        for (int j = 0; j < 5; j++) {
          double hi = __fma_rz(a[i], b[j], c1);
          double lo = __fma_rz(a[i], b[j], c2 - hi);
          w[i + j + 1] += __double_as_longlong(hi);
          w[i + j] += __double_as_longlong(lo);
        }
    }
#pragma unroll
    for (int i = 0; i < 10; i++) r ^= (uint32_t)w[i] ^ (uint32_t)(w[i] >> 32);
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template<int MODE> int run(const char* name, int ctas_per_sm, int nsm, int iters, const uint32_t* in) {
  int grid = nsm * ctas_per_sm, block = 160;
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
  double thr = (double)ctas_per_sm * block;
  printf("{\"mode\": \"%s\", \"warps_per_sm\": %d, \"products_per_clk_per_sm\": %.4f, \"products_per_s\": %.4e, \"ms\": %.3f, \"implied_sm_mhz\": %.0f}\n",
         name, ctas_per_sm * block / 32, thr * iters / avg, (double)grid * block * iters / (ms * 1e-3), ms, avg / (ms * 1e-3) / 1e6);
  cudaFree(out); cudaFree(cyc); free(h); return 0;
}
int main(int argc, char** argv) {
  int iters = argc > 1 ? atoi(argv[1]) : 20000;
  cudaDeviceProp p; CHK(cudaGetDeviceProperties(&p, 0));
  int nsm = p.multiProcessorCount; printf("{\"gpu\": \"%s\", \"sms\": %d, \"iters\": %d}\n", p.name, nsm, iters);
  uint32_t* in; CHK(cudaMalloc(&in, 4096 * 4));
  uint32_t hin[4096]; for (int i = 0; i < 4096; i++) hin[i] = (0x9E3779B9u * (i + 1) ^ (i << 7)) & 0x1fffffff;
  CHK(cudaMemcpy(in, hin, sizeof(hin), cudaMemcpyHostToDevice));
  for (int c = 1; c <= 4; c *= 2) { run<0>("imad81", c, nsm, iters, in); run<1>("dfma25", c, nsm, iters, in); }
  return 0;
}
