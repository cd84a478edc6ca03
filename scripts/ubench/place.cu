// Development aid: (1) where do the warps of co-resident 160-thread CTAs land (SM, hardware warp slot -> SM
// sub-partition = %warpid % 4)?  (2) F2F (f32<->f64 conversion, XU pipe) throughput of one SM sub-partition: cycles
// per warp-wide conversion when 1 / 2 / 3 warps of the SAME sub-partition convert at once, against warps spread over
// different sub-partitions.   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o place place.cu && ./place
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <map>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(160, 3) where(int *rec) {
  extern __shared__ float pad[];
  if (threadIdx.x == 0) pad[0] = 1.f;
  unsigned smid, warpid;
  asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
  asm volatile("mov.u32 %0, %%warpid;" : "=r"(warpid));
  if ((threadIdx.x & 31) == 0) { rec[(blockIdx.x * 5 + (threadIdx.x >> 5)) * 2] = (int)smid; rec[(blockIdx.x * 5 + (threadIdx.x >> 5)) * 2 + 1] = (int)warpid; }
  // stay resident long enough that all CTAs of the grid overlap
  long long t0 = clock64();
  while (clock64() - t0 < 2000000) {}
}

// each participating warp runs a chain-free stream of conversions; `mask` says which warps of the CTA take part
__global__ void __launch_bounds__(512) xu(float *out, long long *cyc, unsigned mask, int mode) {
  const int w = threadIdx.x >> 5;
  if (!((mask >> w) & 1)) return;
  float f[8];
  for (int i = 0; i < 8; i++) f[i] = out[threadIdx.x * 8 + i];
  double acc = 0;
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < 2048; it++) {
    if (mode == 0) {
#pragma unroll
      for (int i = 0; i < 8; i++) { double d = (double)f[i]; f[i] = (float)(d * 1.0000001); }  // 2 F2F + 1 DMUL per element
    } else {
#pragma unroll
      for (int i = 0; i < 8; i++) { acc += (double)f[i] * 1.0000001; }                          // 1 F2F + DMUL + DADD
    }
  }
  long long t1 = clock64();
  for (int i = 0; i < 8; i++) out[threadIdx.x * 8 + i] = f[i] + (float)acc;
  if ((threadIdx.x & 31) == 0) cyc[w] = t1 - t0;
}

int main() {
  int *rec;
  const int grid = 417;
  cudaMallocManaged(&rec, grid * 5 * 2 * sizeof(int));
  cudaFuncSetAttribute(where, cudaFuncAttributeMaxDynamicSharedMemorySize, 75664);
  where<<<grid, 160, 75664>>>(rec);
  cudaDeviceSynchronize();
  // per SM: the sub-partitions of the CTAs' warp 4
  std::map<int, std::vector<int>> per_sm;
  int hist[4] = {0, 0, 0, 0};
  for (int b = 0; b < grid; b++) { per_sm[rec[(b * 5 + 4) * 2]].push_back(rec[(b * 5 + 4) * 2 + 1]); }
  int clash = 0, sms = 0;
  for (auto &kv : per_sm) {
    sms++;
    int c[4] = {0, 0, 0, 0};
    for (int wid : kv.second) { c[wid % 4]++; hist[wid % 4]++; }
    for (int k = 0; k < 4; k++) if (c[k] > 1) clash++;
  }
  printf("warp 4 of %d CTAs on %d SMs: sub-partition histogram %d %d %d %d; SMs where two warp-4s share a sub-partition: %d\n", grid, sms,
         hist[0], hist[1], hist[2], hist[3], clash);
  for (int b = 0; b < 6; b++) {
    printf("  CTA %d: sm %d warpids", b, rec[b * 10]);
    for (int w = 0; w < 5; w++) printf(" %d", rec[(b * 5 + w) * 2 + 1]);
    printf("\n");
  }
  int sm0 = rec[0];
  printf("  CTAs on SM %d:", sm0);
  for (int b = 0; b < grid; b++) if (rec[b * 10] == sm0) printf(" %d(wid4=%d)", b, rec[(b * 5 + 4) * 2 + 1]);
  printf("\n");

  float *out; long long *cyc;
  cudaMallocManaged(&out, 512 * 8 * sizeof(float)); cudaMallocManaged(&cyc, 16 * sizeof(long long));
  for (int i = 0; i < 512 * 8; i++) out[i] = 0.5f + i * 1e-4f;
  struct { const char *name; unsigned mask; } cases[] = {
      {"1 warp (w0)", 0x1}, {"2 warps same SMSP (w0,w4)", 0x11}, {"3 warps same SMSP (w0,w4,w8)", 0x111},
      {"4 warps same SMSP (w0,w4,w8,w12)", 0x1111}, {"3 warps on 3 SMSPs (w0,w1,w2)", 0x7}, {"4 warps on 4 SMSPs", 0xF}, {"16 warps", 0xFFFF}};
  for (int mode = 0; mode < 2; mode++)
    for (auto &c : cases) {
      for (int r = 0; r < 2; r++) { xu<<<1, 512>>>(out, cyc, c.mask, mode); cudaDeviceSynchronize(); }
      long long mx = 0; int n = 0;
      for (int w = 0; w < 16; w++) if ((c.mask >> w) & 1) { if (cyc[w] > mx) mx = cyc[w]; n++; }
      const double convs = 2048.0 * 8 * (mode == 0 ? 2 : 1);
      printf("mode %d %-36s %.2f cycles per warp-wide F2F per warp; %.2f cycles per F2F on the busiest SMSP\n", mode, c.name, mx / convs,
             mx / convs / (c.mask == 0x7 || c.mask == 0xF ? 1 : (c.mask == 0xFFFF ? 4 : n)));
    }
  return 0;
}
