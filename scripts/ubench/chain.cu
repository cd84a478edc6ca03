// Development aid: per-sample latency of the biquad recurrence (Filter.js:40-48) in the lane-per-voice warp, for
// several bit-exact formulations of the dependent chain.  c[] (the feed-forward half a0 x + a1 x1 + a2 x2, f64) is
// read from shared memory and y (f32) written back, as in the pipelined kernel.  Every variant is checked against
// the literal form on the same inputs (including runs that hit NaN / -0 / denormals, which must take the redo path).
//   nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false -O3 -o chain chain.cu && ./chain
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>

#define NS 8192          // samples per voice
#define G 4              // samples per group
#define VC 24

__device__ __forceinline__ float or0f(float x) { return fabsf(x) > 0.0f ? x : 0.0f; }

// V0: literal (what ops.cuh c_filterc does today)
template <int V>
__device__ __forceinline__ bool group(const double (&c)[G], float (&y)[G], double b1, double b2, double &q1, double &q2);

template <>
__device__ __forceinline__ bool group<0>(const double (&c)[G], float (&y)[G], double b1, double b2, double &q1, double &q2) {
  double b2q2 = b2 * q2;
#pragma unroll
  for (int s = 0; s < G; s++) {
    const float yf = (float)(c[s] - b1 * q1 - b2q2);
    y[s] = yf;
    b2q2 = b2 * q1;
    q2 = q1; q1 = (double)or0f(yf);
  }
  return true;
}
// V1: `|| 0` off the chain (speculation: a NaN or -0 output is noticed afterwards, the group is redone literally)
template <>
__device__ __forceinline__ bool group<1>(const double (&c)[G], float (&y)[G], double b1, double b2, double &q1, double &q2) {
  double b2q2 = b2 * q2;
  bool odd = false;
#pragma unroll
  for (int s = 0; s < G; s++) {
    const float yf = (float)(c[s] - b1 * q1 - b2q2);
    y[s] = yf;
    b2q2 = b2 * q1;
    q2 = q1; q1 = (double)yf;
    odd |= !(fabsf(yf) > 0.0f) && (__float_as_uint(yf) != 0u);  // NaN or -0
  }
  return !odd;
}
// V2: V1 + f32 -> f64 widening on the integer pipe (normal numbers and +0 only; anything else is "odd")
__device__ __forceinline__ double widen_int(float yf, bool &odd) {
  const unsigned u = __float_as_uint(yf);
  const unsigned e = (u >> 23) & 0xffu;
  odd |= (e == 0u && u != 0u) | (e == 0xffu);          // -0, denormals, Inf, NaN
  const unsigned hi = (u >> 3) + ((u >> 31) ? 0xA8000000u : 0x38000000u);
  const double d = __hiloint2double((int)hi, (int)(u << 29));
  return u == 0u ? 0.0 : d;
}
template <>
__device__ __forceinline__ bool group<2>(const double (&c)[G], float (&y)[G], double b1, double b2, double &q1, double &q2) {
  double b2q2 = b2 * q2;
  bool odd = false;
#pragma unroll
  for (int s = 0; s < G; s++) {
    const float yf = (float)(c[s] - b1 * q1 - b2q2);
    y[s] = yf;
    b2q2 = b2 * q1;
    q2 = q1; q1 = widen_int(yf, odd);
  }
  return !odd;
}
// V3: rounding to 24 bits in the FP64 domain: (t + M) - M with M = 1.5 * 2^(E+29), E = exponent of t.  The rounded
// value stays a double (no conversion on the chain); valid for results in the f32 normal range (or zero).
template <>
__device__ __forceinline__ bool group<3>(const double (&c)[G], float (&y)[G], double b1, double b2, double &q1, double &q2) {
  double b2q2 = b2 * q2;
  bool odd = false;
#pragma unroll
  for (int s = 0; s < G; s++) {
    const double t = c[s] - b1 * q1 - b2q2;
    const unsigned th = (unsigned)__double2hiint(t);
    const double M = __hiloint2double((int)((th & 0x7ff00000u) + 0x01D80000u), 0);  // 1.5 * 2^(E+29)
    const double r = (t + M) - M;
    // f32 normal range 2^-126 <= |r| < 2^128, or r == +0 (t == -0 / NaN / tiny / huge -> redo)
    const unsigned rh = (unsigned)__double2hiint(r) & 0x7fffffffu;
    const bool zero = __double2hiint(r) == 0 && __double2loint(r) == 0 && __double2hiint(t) >= 0 && ((th & 0x7fffffffu) | (unsigned)__double2loint(t)) == 0u;
    odd |= !(zero || (rh >= 0x38100000u && rh < 0x47F00000u));
    b2q2 = b2 * q1;
    q2 = q1; q1 = r;
    y[s] = (float)r;  // exact conversion, off the chain
  }
  return !odd;
}
// V4: no speculation at all: `|| 0` and the f32 -> f64 widening folded into one exact integer sequence.
// Normal f32: hi = (u >> 3) + bias (sign moved from bit 28 back to 31), lo = u << 29.  Zero, -0, NaN -> +0.
// Denormals / Inf keep the conversion instruction (rare, predicated off the common path).
__device__ __forceinline__ double widen_or0(float yf) {
  const unsigned u = __float_as_uint(yf);
  const unsigned a = u & 0x7fffffffu;
  const unsigned e = a >> 23;
  double d = __hiloint2double((int)((u >> 3) + ((u >> 31) ? 0xA8000000u : 0x38000000u)), (int)(u << 29));
  if (e == 0u || e == 0xffu) d = (a == 0u || a > 0x7f800000u) ? 0.0 : (double)yf;  // 0 / -0 / NaN -> 0; denormal / Inf -> F2F
  return d;
}
template <>
__device__ __forceinline__ bool group<4>(const double (&c)[G], float (&y)[G], double b1, double b2, double &q1, double &q2) {
  double b2q2 = b2 * q2;
#pragma unroll
  for (int s = 0; s < G; s++) {
    const float yf = (float)(c[s] - b1 * q1 - b2q2);
    y[s] = yf;
    b2q2 = b2 * q1;
    q2 = q1; q1 = widen_or0(yf);
  }
  return true;
}

template <int V>
__global__ void __launch_bounds__(32) k(const double *cin, float *yout, long long *cyc, const double *coef, int *redo_count) {
  __shared__ double cs[VC][G];
  __shared__ float ys[VC][G];
  const int lane = threadIdx.x;
  const int vl = lane < VC ? lane : VC - 1;
  const double b1 = coef[2 * vl], b2 = coef[2 * vl + 1];
  double q1 = 0.0, q2 = 0.0;
  int redo = 0;
  long long t0 = clock64();
#pragma unroll 1
  for (int g = 0; g < NS / G; g++) {
    // stand-in for the exchange buffer: each lane stages its own group through shared memory
#pragma unroll
    for (int s = 0; s < G; s++) cs[vl][s] = cin[(size_t)vl * NS + g * G + s];
    __syncwarp();
    double c[G];
#pragma unroll
    for (int s = 0; s < G; s++) c[s] = cs[vl][s];
    float y[G];
    const double s1 = q1, s2 = q2;
    if (!group<V>(c, y, b1, b2, q1, q2)) { q1 = s1; q2 = s2; group<0>(c, y, b1, b2, q1, q2); redo++; }
#pragma unroll
    for (int s = 0; s < G; s++) ys[vl][s] = y[s];
    __syncwarp();
    if (lane < VC) {
#pragma unroll
      for (int s = 0; s < G; s++) yout[(size_t)vl * NS + g * G + s] = ys[vl][s];
    }
  }
  long long t1 = clock64();
  if (lane == 0) cyc[0] = t1 - t0;
  atomicAdd(redo_count, lane < VC ? redo : 0);
}

// the same loop with c[] already in registers-from-smem only (no global traffic inside the timed loop): pure chain latency
template <int V>
__global__ void __launch_bounds__(32) kpure(const double *cin, float *yout, long long *cyc, const double *coef) {
  __shared__ double cs[VC][65];
  const int lane = threadIdx.x;
  const int vl = lane < VC ? lane : VC - 1;
  for (int i = 0; i < 64; i++) cs[vl][i] = cin[(size_t)vl * NS + i];
  __syncwarp();
  const double b1 = coef[2 * vl], b2 = coef[2 * vl + 1];
  double q1 = 0.0, q2 = 0.0;
  float acc = 0.f;
  long long t0 = clock64();
#pragma unroll 1
  for (int g = 0; g < NS / G; g++) {
    double c[G];
#pragma unroll
    for (int s = 0; s < G; s++) c[s] = cs[vl][(g * G + s) & 63];
    float y[G];
    const double s1 = q1, s2 = q2;
    if (!group<V>(c, y, b1, b2, q1, q2)) { q1 = s1; q2 = s2; group<0>(c, y, b1, b2, q1, q2); }
#pragma unroll
    for (int s = 0; s < G; s++) acc += y[s];
  }
  long long t1 = clock64();
  if (lane == 0) cyc[0] = t1 - t0;
  yout[lane] = acc;
}

template <int V>
void run(const char *name, const double *d_c, float *d_y, long long *d_cyc, const double *d_coef, int *d_redo,
         const std::vector<float> &ref, std::vector<float> &out, bool check) {
  cudaMemset(d_redo, 0, sizeof(int));
  for (int r = 0; r < 2; r++) { cudaMemset(d_redo, 0, sizeof(int)); k<V><<<1, 32>>>(d_c, d_y, d_cyc, d_coef, d_redo); cudaDeviceSynchronize(); }
  long long cyc; int redo;
  cudaMemcpy(&cyc, d_cyc, sizeof cyc, cudaMemcpyDeviceToHost);
  cudaMemcpy(&redo, d_redo, sizeof redo, cudaMemcpyDeviceToHost);
  cudaMemcpy(out.data(), d_y, out.size() * sizeof(float), cudaMemcpyDeviceToHost);
  size_t bad = 0;
  if (check) for (size_t i = 0; i < out.size(); i++) bad += memcmp(&out[i], &ref[i], 4) != 0;
  for (int r = 0; r < 2; r++) { kpure<V><<<1, 32>>>(d_c, d_y, d_cyc, d_coef); cudaDeviceSynchronize(); }
  long long cyc2;
  cudaMemcpy(&cyc2, d_cyc, sizeof cyc2, cudaMemcpyDeviceToHost);
  printf("%-28s %6.1f cycles/sample staged, %6.1f pure chain; redo groups %d; mismatches vs literal %zu\n", name,
         (double)cyc / NS, (double)cyc2 / NS, redo, bad);
}

static void fill(std::vector<double> &c, std::vector<double> &coef, bool adversarial) {
  srand(7);
  for (int v = 0; v < VC; v++) {
    // LP biquad, cutoff 200..4000 Hz (Filter.js:66-74)
    const double f = 200 + 3800.0 * v / VC, lam = 1.0 / tan(M_PI * f / 44100.0), l2 = lam * lam, a0 = 1.0 / (1 + 2 * lam + l2);
    coef[2 * v] = 2 * a0 * (1 - l2); coef[2 * v + 1] = a0 * (1 - 2 * lam + l2);
    for (int i = 0; i < NS; i++) {
      double x = (rand() / (double)RAND_MAX * 2 - 1) * a0 * 4;
      if (adversarial) {
        if (v == 3 && i > 2000) x = 0.0;                   // silence: the output decays through the denormals to zero
        if (v == 5 && i == 1000) x = NAN;                  // a NaN passes through (`|| 0` on the fed-back value)
        if (v == 7 && i > 100 && i < 200) x = -0.0;
        if (v == 9) x *= 1e-30;                            // tiny signal: f32 denormal outputs
        if (v == 11) x *= 1e36;                            // huge signal: overflow to Inf
      }
      c[(size_t)v * NS + i] = x;
    }
  }
}

int main() {
  std::vector<double> c((size_t)VC * NS), coef(2 * VC);
  double *d_c, *d_coef; float *d_y; long long *d_cyc; int *d_redo;
  cudaMalloc(&d_c, c.size() * 8); cudaMalloc(&d_coef, coef.size() * 8); cudaMalloc(&d_y, (size_t)VC * NS * 4);
  cudaMalloc(&d_cyc, 8); cudaMalloc(&d_redo, 4);
  for (int adv = 0; adv < 2; adv++) {
    fill(c, coef, adv != 0);
    printf(adv ? "--- adversarial inputs (silence -> denormals, NaN, -0, tiny, huge): correctness of the redo path\n" : "--- clean inputs: timing\n");
    cudaMemcpy(d_c, c.data(), c.size() * 8, cudaMemcpyHostToDevice);
    cudaMemcpy(d_coef, coef.data(), coef.size() * 8, cudaMemcpyHostToDevice);
    std::vector<float> ref((size_t)VC * NS), out((size_t)VC * NS);
    run<0>("V0 literal", d_c, d_y, d_cyc, d_coef, d_redo, ref, ref, false);
    run<1>("V1 spec ||0", d_c, d_y, d_cyc, d_coef, d_redo, ref, out, true);
    run<2>("V2 spec + int widening", d_c, d_y, d_cyc, d_coef, d_redo, ref, out, true);
    run<3>("V3 magic rounding in f64", d_c, d_y, d_cyc, d_coef, d_redo, ref, out, true);
    run<4>("V4 literal ||0 + int widening", d_c, d_y, d_cyc, d_coef, d_redo, ref, out, true);
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(e)); return 1; }
  return 0;
}
