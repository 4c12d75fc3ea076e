// Device check of the fp64 Box-Muller routines of csrc/pf_rng.cuh: the
// routines written for the generator's arguments (neg2_log_uniform,
// sincos_turn_fraction, sqrt_positive) against the CUDA library forms
// (log, sqrt, sincospi) on 2^26 random word pairs and the edge words.  Prints
// the largest absolute difference of the two normals.  Built and run by
// tests/test_gpu_noise.py on the GPU.
#include "pf_rng.cuh"
#include <cstdio>
using namespace pf;

__global__ void
k_compare(unsigned long long n, double* worst)
{
  double w = 0.;
  const PhiloxKeys keys = philox_keys(0x1234567ull);
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
       i < n; i += (unsigned long long)gridDim.x * blockDim.x) {
    uint4 q = philox4x32_r<10>(i, 7, keys);
    if (i < 64) { // edge words: powers of two and their neighbours, both ends
      const unsigned e = (unsigned)(1ull << (i & 31)) + (unsigned)((i >> 5) ? -1 : 0);
      q.x = e;
      q.y = ~e;
      q.z = i == 0 ? 0u : 0xffffffffu;
      q.w = i == 1 ? 0u : 0xffffffffu;
    }
    const double2 a = box_muller_f64(q.x, q.y), b = box_muller_f64_lib(q.x, q.y);
    const double2 c = box_muller_f64(q.z, q.w), d = box_muller_f64_lib(q.z, q.w);
    w = fmax(w, fmax(fabs(a.x - b.x), fabs(a.y - b.y)));
    w = fmax(w, fmax(fabs(c.x - d.x), fabs(c.y - d.y)));
  }
  for (int o = 16; o > 0; o >>= 1)
    w = fmax(w, __shfl_xor_sync(0xffffffffu, w, o));
  if ((threadIdx.x & 31) == 0)
    atomicMax(reinterpret_cast<unsigned long long*>(worst),
              (unsigned long long)__double_as_longlong(w)); // w >= 0: ordered as integers
}

int main()
{
  double* worst;
  cudaMalloc(&worst, 8);
  cudaMemset(worst, 0, 8);
  k_compare<<<592, 256>>>(1ull << 26, worst);
  double h = -1.;
  if (cudaMemcpy(&h, worst, 8, cudaMemcpyDeviceToHost) != cudaSuccess)
    return 2;
  printf("%.3e\n", h);
  return 0;
}
