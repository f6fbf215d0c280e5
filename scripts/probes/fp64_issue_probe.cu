// How fast can ONE SM sub-partition issue fp64 FMAs, as a function of the warps resident on it and of the independent
// accumulators per thread?  (Design input for k_asm_tiles_ws: its Q2-Q2 moment tiles keep 45 accumulators in one warp
// per scheduler.)   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_issue_probe fp64_issue_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int ACC>
__global__ void k_probe(double* out, long long* cyc, int iters, double m, double c) {
  double a[ACC];
#pragma unroll
  for (int i = 0; i < ACC; ++i) a[i] = threadIdx.x * 1e-3 + i;
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ACC; ++i) a[i] = fma(a[i], m, c);
  }
  const long long t1 = clock64();
  double s = 0;
#pragma unroll
  for (int i = 0; i < ACC; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int ACC>
void run(int warps_per_smsp, double* out, long long* cyc) {
  const int threads = 128 * warps_per_smsp, iters = 2000, blocks = 148;
  k_probe<ACC><<<blocks, threads>>>(out, cyc, 10, 0.999999, 1e-6);
  k_probe<ACC><<<blocks, threads>>>(out, cyc, iters, 0.999999, 1e-6);
  cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  double avg = 0;
  for (int i = 0; i < blocks; ++i) avg += (double)h[i];
  avg /= blocks;
  const double fma_per_warp = (double)iters * ACC;
  printf("acc %2d  warps/SMSP %d : %.2f cycles per warp-DFMA of one warp, %.2f DFMA lanes per cycle per SMSP\n", ACC, warps_per_smsp,
         avg / fma_per_warp, 32.0 * fma_per_warp * warps_per_smsp / avg);
}

int main() {
  double* out; long long* cyc;
  cudaMalloc(&out, sizeof(double) * 148 * 1024); cudaMalloc(&cyc, sizeof(long long) * 148);
  for (int w : {1, 2, 4, 8}) { run<4>(w, out, cyc); run<8>(w, out, cyc); run<16>(w, out, cyc); run<45>(w, out, cyc); }
  cudaError_t e = cudaDeviceSynchronize();
  printf("%s\n", cudaGetErrorString(e));
  return e != cudaSuccess;
}
