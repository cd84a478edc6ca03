// Dependent-chain latency microbenchmark (development aid): cycles per op for the instructions on the
// biquad recurrence's critical path.  nvcc -arch=sm_100a -fmad=false -o lat lat.cu
#include <cstdio>
#include <cuda_runtime.h>
#define N 4096
__global__ void k(double *out, long long *cyc, double b1, double c, float seedf) {
  double q = out[0];
  long long t0, t1;
  // 0: DMUL chain
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; i++) q = q * b1;
  t1 = clock64(); cyc[0] = t1 - t0; out[1] = q; q = out[0];
  // 1: DADD chain
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; i++) q = q + c;
  t1 = clock64(); cyc[1] = t1 - t0; out[2] = q; q = out[0];
  // 2: F2F.F32.F64 + F2F.F64.F32 round trip
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; i++) q = (double)(float)q + 0.0 * 0;  // compiler keeps the two conversions
  t1 = clock64(); cyc[2] = t1 - t0; out[3] = q; q = out[0];
  // 3: full recurrence step: (float)(c - b1*q - b2q2) -> double
  double q2 = out[5];
  t0 = clock64();
#pragma unroll 8
  for (int i = 0; i < N; i++) { float yf = (float)(c - b1 * q - q2); q2 = b1 * q; q = (double)yf; }
  t1 = clock64(); cyc[3] = t1 - t0; out[4] = q + q2; q = out[0];
  // 4: same with the `|| 0` select on the chain
  q2 = out[5];
  t0 = clock64();
#pragma unroll 8
  for (int i = 0; i < N; i++) { float yf = (float)(c - b1 * q - q2); q2 = b1 * q; yf = (yf != yf || yf == 0.0f) ? 0.0f : yf; q = (double)yf; }
  t1 = clock64(); cyc[4] = t1 - t0; out[6] = q + q2; q = out[0];
  // 5: magic-number rounding to 24 bits in double (exponent known): (t + M) - M
  double M = out[7];
  t0 = clock64();
#pragma unroll 8
  for (int i = 0; i < N; i++) { double t = c - b1 * q - q2; q2 = b1 * q; q = (t + M) - M; }
  t1 = clock64(); cyc[5] = t1 - t0; out[8] = q + q2;
  // 6: FSEL chain (float)
  float f = seedf;
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; i++) f = (f != f || f == 0.0f) ? 0.0f : f * 1.0000001f;
  t1 = clock64(); cyc[6] = t1 - t0; out[9] = f;
  // 7: FMUL chain
  f = seedf;
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; i++) f = f * 1.0000001f;
  t1 = clock64(); cyc[7] = t1 - t0; out[10] = f;
}
int main() {
  double *out; long long *cyc;
  cudaMallocManaged(&out, 16 * sizeof(double)); cudaMallocManaged(&cyc, 16 * sizeof(long long));
  out[0] = 0.3; out[5] = 0.01; out[7] = 1.5 * 4503599627370496.0 / 8388608.0 * 0.5;
  for (int r = 0; r < 2; r++) { k<<<1, 32>>>(out, cyc, -0.9371, 0.123, 0.5f); cudaDeviceSynchronize(); }
  const char *names[] = {"DMUL", "DADD", "F2F.F32.F64+F2F.F64.F32(+DADD)", "recurrence step", "recurrence step with ||0", "recurrence magic-round", "FSETP+FSEL+FMUL", "FMUL"};
  for (int i = 0; i < 8; i++) printf("%-36s %.1f cycles/iter\n", names[i], (double)cyc[i] / N);
  return 0;
}
