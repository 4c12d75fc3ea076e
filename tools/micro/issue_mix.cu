// Micro-benchmark (development aid): does the B200 SM sub-partition issue
// ALU / FMA / XU work in the shadow of fp64 instructions?  Each variant runs a
// loop of 8 independent DFMAs, alone or interleaved with 8 instructions of
// another pipe, at 1 and 3 warps per sub-partition.  Prints cycles per loop
// iteration per sub-partition.
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void k(double* out, unsigned* iout, int iters, long long* cyc)
{
  double a0 = threadIdx.x, a1 = 1, a2 = 2, a3 = 3, a4 = 4, a5 = 5, a6 = 6, a7 = 7;
  const double m = 1.0000001, c = 1e-9;
  unsigned i0 = threadIdx.x, i1 = 1, i2 = 2, i3 = 3, i4 = 4, i5 = 5, i6 = 6, i7 = 7;
  float f0 = threadIdx.x, f1 = 1, f2 = 2, f3 = 3, f4 = 4, f5 = 5, f6 = 6, f7 = 7;
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < iters; it++) {
    if (MODE != 3 && MODE != 5 && MODE != 8 && MODE != 9 && MODE != 10) {
      a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
      a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
    }
    if (MODE == 1 || MODE == 3) { // LOP3 (ALU pipe)
      i0 = (i0 ^ 0x9E3779B9u) & i1; i1 = (i1 ^ 0x85EBCA6Bu) | i2; i2 = (i2 ^ i3) & 0xC2B2AE35u; i3 = (i3 ^ 0x27D4EB2Fu) | i4;
      i4 = (i4 ^ 0x165667B1u) & i5; i5 = (i5 ^ i6) | 0x9E3779B9u; i6 = (i6 ^ 0x85EBCA6Bu) & i7; i7 = (i7 ^ i0) | 0xC2B2AE35u;
    }
    if (MODE == 2) { // FFMA
      f0 = fmaf(f0, 1.0001f, 1e-5f); f1 = fmaf(f1, 1.0001f, 1e-5f); f2 = fmaf(f2, 1.0001f, 1e-5f); f3 = fmaf(f3, 1.0001f, 1e-5f);
      f4 = fmaf(f4, 1.0001f, 1e-5f); f5 = fmaf(f5, 1.0001f, 1e-5f); f6 = fmaf(f6, 1.0001f, 1e-5f); f7 = fmaf(f7, 1.0001f, 1e-5f);
    }
    if (MODE == 4 || MODE == 5) { // IMAD.WIDE (Philox multiply)
      unsigned long long p0 = (unsigned long long)i0 * 0xD2511F53u, p1 = (unsigned long long)i1 * 0xCD9E8D57u;
      unsigned long long p2 = (unsigned long long)i2 * 0xD2511F53u, p3 = (unsigned long long)i3 * 0xCD9E8D57u;
      i0 = (unsigned)(p0 >> 32) ^ (unsigned)p1; i1 = (unsigned)(p1 >> 32) ^ (unsigned)p2;
      i2 = (unsigned)(p2 >> 32) ^ (unsigned)p3; i3 = (unsigned)(p3 >> 32) ^ (unsigned)p0;
    }
    if (MODE == 7 || MODE == 8) { // the same products as IMAD (low half) + IMAD.HI
      unsigned l0 = i0 * 0xD2511F53u, h0 = __umulhi(i0, 0xD2511F53u);
      unsigned l1 = i1 * 0xCD9E8D57u, h1 = __umulhi(i1, 0xCD9E8D57u);
      unsigned l2 = i2 * 0xD2511F53u, h2 = __umulhi(i2, 0xD2511F53u);
      unsigned l3 = i3 * 0xCD9E8D57u, h3 = __umulhi(i3, 0xCD9E8D57u);
      asm volatile("" : "+r"(l0), "+r"(h0), "+r"(l1), "+r"(h1));
      asm volatile("" : "+r"(l2), "+r"(h2), "+r"(l3), "+r"(h3));
      i0 = h0 ^ l1; i1 = h1 ^ l2; i2 = h2 ^ l3; i3 = h3 ^ l0;
    }
    if (MODE == 9) { // 8 IMAD (32-bit low product)
      i0 = i0 * 0xD2511F53u + i1; i1 = i1 * 0xCD9E8D57u + i2; i2 = i2 * 0xD2511F53u + i3; i3 = i3 * 0xCD9E8D57u + i4;
      i4 = i4 * 0xD2511F53u + i5; i5 = i5 * 0xCD9E8D57u + i6; i6 = i6 * 0xD2511F53u + i7; i7 = i7 * 0xCD9E8D57u + i0;
    }
    if (MODE == 10) { // 8 IMAD.HI
      i0 = __umulhi(i0, 0xD2511F53u) ^ i1; i1 = __umulhi(i1, 0xCD9E8D57u) ^ i2; i2 = __umulhi(i2, 0xD2511F53u) ^ i3; i3 = __umulhi(i3, 0xCD9E8D57u) ^ i4;
      i4 = __umulhi(i4, 0xD2511F53u) ^ i5; i5 = __umulhi(i5, 0xCD9E8D57u) ^ i6; i6 = __umulhi(i6, 0xD2511F53u) ^ i7; i7 = __umulhi(i7, 0xCD9E8D57u) ^ i0;
    }
    if (MODE == 6) { // MUFU
      f0 = __sinf(f0); f1 = __cosf(f1); f2 = __sinf(f2); f3 = __cosf(f3);
    }
  }
  const long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7 + f0 + f1 + f2 + f3 + f4 + f5 + f6 + f7;
  iout[blockIdx.x * blockDim.x + threadIdx.x] = i0 ^ i1 ^ i2 ^ i3 ^ i4 ^ i5 ^ i6 ^ i7;
  if (threadIdx.x == 0 && blockIdx.x == 0)
    *cyc = t1 - t0;
}

template <int MODE>
void run(const char* name, int warps_per_sm)
{
  double* out; unsigned* iout; long long* cyc;
  cudaMalloc(&out, 148 * 1024 * 8); cudaMalloc(&iout, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
  const int iters = 20000;
  k<MODE><<<148, 32 * warps_per_sm>>>(out, iout, iters, cyc);
  k<MODE><<<148, 32 * warps_per_sm>>>(out, iout, iters, cyc);
  cudaDeviceSynchronize();
  long long h = 0; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("%-34s warps/SMSP %d : %.2f cycles per iteration per warp, %.2f per SMSP-iteration\n", name, warps_per_sm / 4,
         (double)h / iters, (double)h / iters / (warps_per_sm / 4));
  cudaFree(out); cudaFree(iout); cudaFree(cyc);
}

int main()
{
  for (int w : {4, 12, 24}) {
    run<0>("8 DFMA", w);
    run<1>("8 DFMA + 8 LOP3", w);
    run<3>("8 LOP3", w);
    run<2>("8 DFMA + 8 FFMA", w);
    run<4>("8 DFMA + 4 IMAD.WIDE + 4 LOP3", w);
    run<5>("4 IMAD.WIDE + 4 LOP3", w);
    run<6>("8 DFMA + 4 MUFU(+range red.)", w);
    run<7>("8 DFMA + 4 (IMAD + IMAD.HI) + 4 LOP3", w);
    run<8>("4 (IMAD + IMAD.HI) + 4 LOP3", w);
    run<9>("8 IMAD (low product + add)", w);
    run<10>("8 IMAD.HI + 8 LOP3", w);
  }
  return 0;
}
