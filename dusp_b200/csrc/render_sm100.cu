// render_sm100.cu -- the device interpreter: renders every voice of a batch
// through its flat op program.  Hand-written for sm_100a (B200); compiled with
// -fmad=false so that no multiply-add is contracted: the reference computes
// every expression in IEEE double with a rounding after each operation and
// rounds to binary32 only on the store into a Float32Array.
//
// Execution model ("lane = voice"):
//   * one warp owns 32 voices for the whole render; a lane walks its voice
//     forward in time, so every stateful unit (phase accumulators, biquads,
//     delay lines) sees exactly the reference's sequential order -> bit-exact;
//   * time is cut into sub-blocks of B samples.  For each sub-block the warp
//     runs the program's instructions in tick order; an instruction processes
//     B samples of all 32 voices.  Slots (the SignalChunk replacement) live in
//     shared memory as [slot][sample][lane]: every access is bank-conflict free;
//   * unit state and parameters live in HBM/L2 as f64 words [word][voice]
//     (coalesced across the warp), delay lines as per-voice rings in HBM;
//   * the rendered outlet is stored to HBM with the reference's `x || 0`.
// Warps never communicate, so there is no block-level synchronisation at all.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "vm.h"
#include "launch.h"

namespace {

constexpr double SRD = 44100.0;

// ---------------------------------------------------------------- helpers
__device__ __forceinline__ double js_or0(double x) { return (x != x || x == 0.0) ? 0.0 : x; }  // `x || 0`
__device__ __forceinline__ float js_or0f(float x) { return (x != x || x == 0.0f) ? 0.0f : x; }

// JS `x % 44100` (= C fmod) with exact fast paths for the common range.
__device__ __forceinline__ double fmod_sr(double x) {
  if (x >= 0.0) {
    if (x < SRD) return x;
    if (x < 2 * SRD) return x - SRD;  // exact (Sterbenz)
  } else if (x > -SRD) {
    return x;
  } else if (x > -2 * SRD) {
    return x + SRD;
  }
  return fmod(x, SRD);
}

// typed-array read `table[i]` with a double index: out of range / NaN -> undefined -> NaN
__device__ __forceinline__ double tab_read(const float *__restrict__ tab, double idx) {
  return (idx >= 0.0 && idx <= 44100.0) ? (double)__ldg(tab + (int)idx) : (double)NAN;
}
__device__ __forceinline__ bool ring_valid(double idx, double len) { return idx == floor(idx) && idx >= 0.0 && idx < len; }
__device__ __forceinline__ double ring_read(const float *r, double len, double idx) {
  return ring_valid(idx, len) ? (double)r[(long long)idx] : (double)NAN;
}
// CircleBuffer.js:17-20 index normalisation
__device__ __forceinline__ double cb_index(double t, double len) {
  t = floor(fmod(t, len));
  if (t < 0.0) t += len;  // `while(t < 0)`: fmod result is > -len, one step suffices
  return t;
}

// injected PRNG (replaces Math.random): counter-based splitmix64 -> 53-bit uniform
__device__ __forceinline__ double rand53(uint64_t seed, uint64_t k) {
  uint64_t z = seed + (k + 1) * 0x9E3779B97F4A7C15ULL;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  z ^= z >> 31;
  return (double)(z >> 11) * (1.0 / 9007199254740992.0);
}

template <int B>
struct Lane {
  float *S;              // this lane's column of the warp's slot array: slot k, sample s -> S[(k*B+s)*32]
  double *W;             // this voice's words: word w -> W[w * nv]
  long long nv;
  float *rings;          // this voice's ring arena
  const DuspRing *ring_tab;
  float *stale;          // this voice's stale buffers [k][256]
  const float *tables;

  __device__ __forceinline__ void ld(float (&x)[B], int ref) const {
    if (ref >= 0) {
#pragma unroll
      for (int s = 0; s < B; s++) x[s] = S[(ref * B + s) * 32];
    } else {
      float c = ref == REF_NONE ? 0.0f : (float)W[(long long)(REF_CONST_BASE - ref) * nv];
#pragma unroll
      for (int s = 0; s < B; s++) x[s] = c;
    }
  }
  __device__ __forceinline__ void st(const float (&x)[B], int ref) const {
    if (ref >= 0) {
#pragma unroll
      for (int s = 0; s < B; s++) S[(ref * B + s) * 32] = x[s];
    }
  }
  __device__ __forceinline__ double &w(int p) const { return W[(long long)p * nv]; }
  __device__ __forceinline__ float *ring(int r, double &len) const {
    DuspRing rt = ring_tab[r];
    len = (double)rt.length;
    return rings + rt.offset;
  }
};

// ------------------------------------------------------------ stateless ops
template <int B, typename F>
__device__ __forceinline__ void map1(const Lane<B> &L, const DuspInstr &I, F f) {
  float a[B];
  L.ld(a, I.in[0]);
#pragma unroll
  for (int s = 0; s < B; s++) a[s] = f(a[s]);
  L.st(a, I.out[0]);
}
template <int B, typename F>
__device__ __forceinline__ void map2(const Lane<B> &L, const DuspInstr &I, F f) {
  float a[B], b[B];
  L.ld(a, I.in[0]);
  L.ld(b, I.in[1]);
#pragma unroll
  for (int s = 0; s < B; s++) a[s] = f(a[s], b[s]);
  L.st(a, I.out[0]);
}

__device__ __forceinline__ double dB(double db) { return pow(10.0, db / 20.0); }  // Gain.js:26-28

// ------------------------------------------------------------- stateful ops
template <int B>
__device__ __forceinline__ void op_osc(const Lane<B> &L, const DuspInstr &I, bool negfix) {  // Osc.js:35-47
  float f[B];
  L.ld(f, I.in[0]);
  const float *tab = L.tables + (size_t)I.aux[0] * DUSP_TABLE_LEN;
  double ph = L.w(I.p);
#pragma unroll
  for (int s = 0; s < B; s++) {
    ph += (double)f[s];
    ph = fmod_sr(ph);
    if (negfix && ph < 0.0) ph += SRD;
    double fl = floor(ph), ce = ceil(ph);
    double fr = ph - trunc(ph);  // phase % 1
    f[s] = (float)(tab_read(tab, fl) * (1.0 - fr) + tab_read(tab, ce) * fr);
  }
  L.w(I.p) = ph;
  L.st(f, I.out[0]);
}

template <int B>
__device__ __forceinline__ void op_ramp(const Lane<B> &L, const DuspInstr &I) {  // Ramp.js:25-40
  double t = L.w(I.p), playing = L.w(I.p + 1);
  const double dur = L.w(I.p + 2), y0 = L.w(I.p + 3), y1 = L.w(I.p + 4);
  float o[B];
#pragma unroll
  for (int s = 0; s < B; s++) {
    if (playing != 0.0) {
      t += 1.0;
      if (t > dur) { playing = 0.0; t = dur; }
      if (t < 0.0) { playing = 0.0; t = 0.0; }
    }
    o[s] = (float)(y0 + (t / dur) * (y1 - y0));
  }
  L.w(I.p) = t;
  L.w(I.p + 1) = playing;
  L.st(o, I.out[0]);
}

template <int B>
__device__ __forceinline__ void op_shape(const Lane<B> &L, const DuspInstr &I) {  // Shape/index.js:28-59
  float d[B], mn[B], mx[B];
  L.ld(d, I.in[0]); L.ld(mn, I.in[1]); L.ld(mx, I.in[2]);
  const float *tab = L.tables + (size_t)I.aux[0] * DUSP_TABLE_LEN;
  double t = L.w(I.p);
  const double playing = L.w(I.p + 1);
  const double le = (I.aux[1] & 1) ? (double)__ldg(tab) : L.w(I.p + 2);
  const double re = (I.aux[1] & 2) ? (double)__ldg(tab + DUSP_SR) : L.w(I.p + 3);
#pragma unroll
  for (int s = 0; s < B; s++) {
    if (playing != 0.0) t += 1.0 / (double)d[s];
    const double lo = mn[s], hi = mx[s];
    double y;
    if (t <= 0.0) y = le * (hi - lo) + lo;
    else if (t > SRD) y = re * (hi - lo) + lo;
    else {
      double fr = t - trunc(t);  // t % 1
      y = lo + (hi - lo) * (tab_read(tab, ceil(t)) * fr + tab_read(tab, floor(t)) * (1.0 - fr));
    }
    d[s] = (float)y;
  }
  L.w(I.p) = t;
  L.st(d, I.out[0]);
}

template <int B>
__device__ __forceinline__ void op_ahd(const Lane<B> &L, const DuspInstr &I) {  // AHD.js:35-76
  float a[B], h[B], d[B];
  L.ld(a, I.in[0]); L.ld(h, I.in[1]); L.ld(d, I.in[2]);
  double t = L.w(I.p);
  int state = (int)L.w(I.p + 1);
  bool playing = L.w(I.p + 2) != 0.0;
  const double sp = 1.0 / SRD;
#pragma unroll
  for (int s = 0; s < B; s++) {
    float y = 0.0f;
    if (state == 1) {
      y = (float)t;
      if (playing) { t += sp / (double)a[s]; if (t >= 1.0) { state++; t -= 1.0; } }
    } else if (state == 2) {
      y = 1.0f;
      if (playing) { t += sp / (double)h[s]; if (t >= 1.0) { state++; t -= 1.0; } }
    } else if (state == 3) {
      y = (float)(1.0 - t);
      if (playing) { t += sp / (double)d[s]; if (t >= 1.0) { state = 0; playing = false; } }
    }
    a[s] = y;
  }
  L.w(I.p) = t;
  L.w(I.p + 1) = (double)state;
  L.w(I.p + 2) = playing ? 1.0 : 0.0;
  L.st(a, I.out[0]);
}

template <int B>
__device__ __forceinline__ void op_timer(const Lane<B> &L, const DuspInstr &I) {  // Timer.js:37-42
  double t = L.w(I.p);
  const double sp = 1.0 / SRD;
  float o[B];
#pragma unroll
  for (int s = 0; s < B; s++) { t += sp; o[s] = (float)t; }
  L.w(I.p) = t;
  L.st(o, I.out[0]);
}

template <int B>
__device__ __forceinline__ void op_noise(const Lane<B> &L, const DuspInstr &I) {  // Noise.js:16-28
  float f[B];
  L.ld(f, I.in[0]);
  double ph = L.w(I.p), y = L.w(I.p + 1), draw = L.w(I.p + 2);
  const uint64_t seed = (uint64_t)L.w(I.p + 3) | ((uint64_t)L.w(I.p + 4) << 32);
#pragma unroll
  for (int s = 0; s < B; s++) {
    ph += (double)f[s];
    if (ph >= SRD) {
      ph = 0.0;
      y = 2.0 * rand53(seed, (uint64_t)draw) - 1.0;
      draw += 1.0;
    }
    f[s] = (float)y;
  }
  L.w(I.p) = ph; L.w(I.p + 1) = y; L.w(I.p + 2) = draw;
  L.st(f, I.out[0]);
}

// Filter.js:66-84
__device__ __forceinline__ void filter_coeffs(int hp, double f, double &a0, double &a1, double &a2, double &b1, double &b2) {
  const double PI = 3.141592653589793;
  if (!hp) {
    double lamda = 1.0 / tan(PI * f / SRD), l2 = lamda * lamda;
    a0 = 1.0 / (1.0 + 2.0 * lamda + l2); a1 = 2.0 * a0; a2 = a0;
    b1 = 2.0 * a0 * (1.0 - l2); b2 = a0 * (1.0 - 2.0 * lamda + l2);
  } else {
    double lamda = tan(PI * f / SRD), l2 = lamda * lamda;
    a0 = 1.0 / (1.0 + 2.0 * lamda + l2); a1 = 0.0; a2 = -a0;
    b1 = 2.0 * a0 * (l2 - 1.0); b2 = a0 * (1.0 - 2.0 * lamda + l2);
  }
}

template <int B>
__device__ __forceinline__ void op_filter(const Lane<B> &L, const DuspInstr &I) {  // Filter.js:27-51
  float x[B], f[B];
  L.ld(x, I.in[0]); L.ld(f, I.in[1]);
  const int p = I.p;
  double a0 = L.w(p), a1 = L.w(p + 1), a2 = L.w(p + 2), b1 = L.w(p + 3), b2 = L.w(p + 4);
  double lastf = L.w(p + 5);
  bool lastdef = L.w(p + 6) != 0.0;
  double x1 = L.w(p + 7), x2 = L.w(p + 8), y1 = L.w(p + 9), y2 = L.w(p + 10);
#pragma unroll
  for (int s = 0; s < B; s++) {
    const double ft = f[s];
    if (!lastdef || ft != lastf) { lastdef = true; lastf = ft; filter_coeffs(I.aux[0], ft, a0, a1, a2, b1, b2); }
    const double xin = x[s];
    const double y = a0 * xin + a1 * js_or0(x1) + a2 * js_or0(x2) - b1 * js_or0(y1) - b2 * js_or0(y2);
    const float yf = (float)y;
    x[s] = yf;
    y2 = js_or0(y1); y1 = (double)yf; x2 = js_or0(x1); x1 = xin;
  }
  L.w(p) = a0; L.w(p + 1) = a1; L.w(p + 2) = a2; L.w(p + 3) = b1; L.w(p + 4) = b2;
  L.w(p + 5) = lastf; L.w(p + 6) = lastdef ? 1.0 : 0.0;
  L.w(p + 7) = x1; L.w(p + 8) = x2; L.w(p + 9) = y1; L.w(p + 10) = y2;
  L.st(x, I.out[0]);
}

// ----------------------------------------------------------------- rings
template <int B>
__device__ __forceinline__ void op_fixdelay(const Lane<B> &L, const DuspInstr &I) {  // FixedDelay.js:14-20 Comb :11-17 AllPass :8-15
  float x[B], g[B];
  L.ld(x, I.in[0]);
  if (I.op != OP_FIXDELAY) L.ld(g, I.in[1]);
  double len;
  float *buf = L.ring(I.aux[0], len);
  double tb = L.w(I.p);
#pragma unroll
  for (int s = 0; s < B; s++) {
    tb = fmod(tb + 1.0, len);
    const long long p = (long long)tb;
    const float d = buf[p];
    if (I.op == OP_FIXDELAY) { buf[p] = x[s]; x[s] = d; }
    else if (I.op == OP_COMB) { buf[p] = (float)((double)x[s] + (double)d * (double)g[s]); x[s] = d; }
    else { buf[p] = (float)((double)x[s] + (double)d * (double)g[s]); x[s] = (float)((double)d - (double)x[s] * (double)g[s]); }
  }
  L.w(I.p) = tb;
  L.st(x, I.out[0]);
}

template <int B>
__device__ __forceinline__ void op_delay(const Lane<B> &L, const DuspInstr &I, long long t0) {  // Delay.js:20-41 / MonoDelay.js:19-30
  float x[B], dl[B];
  L.ld(x, I.in[0]); L.ld(dl, I.in[1]);
  double len;
  float *buf = L.ring(I.aux[0], len);
  const bool mono = I.op == OP_MONODELAY;
#pragma unroll
  for (int s = 0; s < B; s++) {
    const double tb = fmod((double)(t0 + s), len);
    float y = 0.0f;
    if (!mono) { y = (float)ring_read(buf, len, tb); if (ring_valid(tb, len)) buf[(long long)tb] = 0.0f; }
    const double tw = fmod(tb + (double)dl[s], len);
    const double fr = tw - trunc(tw);
    const double fl = floor(tw);
    const double ce = mono ? fmod(ceil(tw), len) : ceil(tw);
    if (ring_valid(fl, len)) buf[(long long)fl] = (float)((double)buf[(long long)fl] + (double)x[s] * (1.0 - fr));
    if (ring_valid(ce, len)) buf[(long long)ce] = (float)((double)buf[(long long)ce] + (double)x[s] * fr);
    if (mono) { y = (float)ring_read(buf, len, tb); if (ring_valid(tb, len)) buf[(long long)tb] = 0.0f; }
    x[s] = y;
  }
  L.st(x, I.out[0]);
}

template <int B>
__device__ __forceinline__ void op_readback(const Lane<B> &L, const DuspInstr &I) {  // ReadBackDelay.js:23-43
  float x[B], dl[B];
  L.ld(x, I.in[0]); L.ld(dl, I.in[1]);
  double len;
  float *buf = L.ring(I.aux[0], len);
  double t = L.w(I.p);
#pragma unroll
  for (int s = 0; s < B; s++) {
    const double wi = fmod(t + len, len);
    if (ring_valid(wi, len)) buf[(long long)wi] = x[s];
    x[s] = (float)ring_read(buf, len, fmod(t - (double)dl[s] + len, len));
    t += 1.0;
  }
  L.w(I.p) = t;
  L.st(x, I.out[0]);
}

template <int B>
__device__ __forceinline__ void op_cbwrite(const Lane<B> &L, const DuspInstr &I, long long t0) {  // CircleBufferWriter.js:12-25
  float off[B], x[B];
  L.ld(off, I.in[0]);
  const bool has_in = I.in[1] != REF_NONE;
  if (has_in) L.ld(x, I.in[1]);
  double len;
  float *buf = L.ring(I.aux[0], len);
  const double tu = L.w(I.p);
#pragma unroll
  for (int s = 0; s < B; s++) {
    const double tw = tu + (double)(t0 + s) + SRD * (double)off[s];
    const double i = cb_index(tw, len);
    if (ring_valid(i, len)) {
      float v = buf[(long long)i];
      if (I.aux[1]) v = 0.0f;                                   // preWipe: write(c, tWrite, 0)
      if (has_in) v = (float)((double)v + (double)x[s]);        // mix: +=
      buf[(long long)i] = v;
    }
  }
}

template <int B>
__device__ __forceinline__ void op_cbread(const Lane<B> &L, const DuspInstr &I, long long t0) {  // CircleBufferReader.js:12-25
  float off[B];
  L.ld(off, I.in[0]);
  double len;
  float *buf = L.ring(I.aux[0], len);
  const double tu = L.w(I.p);
#pragma unroll
  for (int s = 0; s < B; s++) {
    const double tr = tu + (double)(t0 + s) - SRD * (double)off[s];
    const double i = cb_index(tr, len);
    float y = NAN;
    if (ring_valid(i, len)) { y = buf[(long long)i]; if (I.aux[1]) buf[(long long)i] = 0.0f; }
    off[s] = y;
  }
  L.st(off, I.out[0]);
}

// --------------------------------------------------------------- the kernel
template <int B>
__global__ void __launch_bounds__(DUSP_MAX_CTA_THREADS) dusp_render_lane_per_voice(const DuspKernelParams kp) {
  extern __shared__ float smem[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int warp = blockIdx.x * (blockDim.x >> 5) + wib;
  const long long v = (long long)warp * 32 + lane;  // voice (allocations are padded to a multiple of 32)
  if ((long long)warp * 32 >= kp.n_voices) return;
  const bool active = v < kp.n_voices;

  Lane<B> L;
  L.S = smem + (size_t)wib * kp.n_slots * B * 32 + lane;
  L.W = kp.words + v;
  L.nv = kp.nv_stride;
  L.rings = kp.rings + v * kp.ring_stride;
  L.ring_tab = kp.ring_tab;
  L.stale = kp.stale + v * (long long)kp.n_stale * DUSP_CHUNK;
  L.tables = kp.tables;

  const long long n_proc = (kp.n_samples + DUSP_CHUNK - 1) / DUSP_CHUNK * DUSP_CHUNK;  // whole chunks, like Circuit.tick
  const long long vrow = (long long)(kp.voice0 + v) * kp.out_nch;

  for (long long tl = 0; tl < n_proc; tl += B) {
    const long long t0 = kp.clock0 + tl;   // absolute sample index (the reference's clock + t)
    const int cpos = (int)(tl & (DUSP_CHUNK - 1));
    for (int pc = 0; pc < kp.n_instr; pc++) {
      DuspInstr I;
      {
        const int4 *src = reinterpret_cast<const int4 *>(kp.prog + pc);
        int4 *dst = reinterpret_cast<int4 *>(&I);
        dst[0] = __ldg(src); dst[1] = __ldg(src + 1); dst[2] = __ldg(src + 2);
      }
      switch (I.op) {
        case OP_COPY: map1<B>(L, I, [](float a) { return a; }); break;
        case OP_NEG: map1<B>(L, I, [](float a) { return -a; }); break;
        case OP_ABS: map1<B>(L, I, [](float a) { return fabsf(a); }); break;
        case OP_S2S: map1<B>(L, I, [](float a) { return (float)((double)a * SRD); }); break;
        case OP_SEMI: map1<B>(L, I, [](float a) { return (float)pow(2.0, (double)a / 12.0); }); break;
        case OP_DB2S: map1<B>(L, I, [](float a) { return (float)pow(10.0, (double)a / 20.0); }); break;
        case OP_MIDI2F: map1<B>(L, I, [](float a) { return (float)(pow(2.0, ((double)a - 69.0) / 12.0) * 440.0); }); break;
        case OP_FIXMUL: { const double sf = L.w(I.p); map1<B>(L, I, [sf](float a) { return (float)((double)a * sf); }); break; }
        // a single IEEE operation on binary32 operands: double evaluation + one rounding == native f32 op
        case OP_MUL: map2<B>(L, I, [](float a, float b) { return __fmul_rn(a, b); }); break;
        case OP_DIV: map2<B>(L, I, [](float a, float b) { return __fdiv_rn(a, b); }); break;
        case OP_ADD: map2<B>(L, I, [](float a, float b) { return __fadd_rn(a, b); }); break;
        case OP_SUB: map2<B>(L, I, [](float a, float b) { return __fsub_rn(a, b); }); break;
        case OP_POW: map2<B>(L, I, [](float a, float b) { return (float)pow((double)a, (double)b); }); break;
        case OP_GAIN: map2<B>(L, I, [](float a, float g) { return (float)(dB((double)g) * (double)a); }); break;
        case OP_CLIP: map2<B>(L, I, [](float a, float th) { return fabsf(a) > fabsf(th) ? th : a; }); break;
        case OP_HCA: map2<B>(L, I, [](float a, float th) { return a > th ? th : a; }); break;
        case OP_HCB: map2<B>(L, I, [](float a, float th) { return a < th ? th : a; }); break;
        case OP_RESCALE: {  // Rescale.js:33-34
          float x[B], il[B], iu[B], ol[B], ou[B];
          L.ld(x, I.in[0]); L.ld(il, I.in[1]); L.ld(iu, I.in[2]); L.ld(ol, I.in[3]); L.ld(ou, I.in[4]);
#pragma unroll
          for (int s = 0; s < B; s++)
            x[s] = (float)(((double)x[s] - (double)il[s]) / ((double)iu[s] - (double)il[s]) * ((double)ou[s] - (double)ol[s]) + (double)ol[s]);
          L.st(x, I.out[0]);
          break;
        }
        case OP_XFADE: {  // CrossFader.js:26
          float a[B], b[B], d[B];
          L.ld(a, I.in[0]); L.ld(b, I.in[1]); L.ld(d, I.in[2]);
#pragma unroll
          for (int s = 0; s < B; s++) a[s] = (float)((1.0 - (double)d[s]) * (double)a[s] + (double)d[s] * (double)b[s]);
          L.st(a, I.out[0]);
          break;
        }
        case OP_PAN: {  // Pan.js:18-24
          float x[B], p[B], o1[B];
          L.ld(x, I.in[0]); L.ld(p, I.in[1]);
          const double cdb = L.w(I.p);
#pragma unroll
          for (int s = 0; s < B; s++) {
            const double pan = p[s], in = x[s];
            const double comp = dB((1.0 - fabs(pan)) * cdb);
            x[s] = (float)(in * (1.0 - pan) / 2.0 * comp);
            o1[s] = (float)(in * (1.0 + pan) / 2.0 * comp);
          }
          L.st(x, I.out[0]); L.st(o1, I.out[1]);
          break;
        }
        case OP_VECMAG: {  // VectorMagnitude.js:19-25
          double acc[B];
#pragma unroll
          for (int s = 0; s < B; s++) acc[s] = 0.0;
          for (int c = 0; c < I.aux[0]; c++) {
            float x[B];
            L.ld(x, I.in[c]);
#pragma unroll
            for (int s = 0; s < B; s++) acc[s] += (double)x[s] * (double)x[s];
          }
          float o[B];
#pragma unroll
          for (int s = 0; s < B; s++) o[s] = (float)sqrt(acc[s]);
          L.st(o, I.out[0]);
          break;
        }
        case OP_OSC: op_osc<B>(L, I, true); break;
        case OP_MCOSC: op_osc<B>(L, I, false); break;
        case OP_RAMP: op_ramp<B>(L, I); break;
        case OP_SHAPE: op_shape<B>(L, I); break;
        case OP_AHD: op_ahd<B>(L, I); break;
        case OP_TIMER: op_timer<B>(L, I); break;
        case OP_NOISE: op_noise<B>(L, I); break;
        case OP_FILTER: op_filter<B>(L, I); break;
        case OP_FIXDELAY: case OP_COMB: case OP_ALLPASS: op_fixdelay<B>(L, I); break;
        case OP_DELAY: case OP_MONODELAY: op_delay<B>(L, I, t0); break;
        case OP_READBACK: op_readback<B>(L, I); break;
        case OP_CBWRITE: op_cbwrite<B>(L, I, t0); break;
        case OP_CBREAD: op_cbread<B>(L, I, t0); break;
        case OP_LDSTALE: {
          float x[B];
          const float *src = L.stale + (size_t)I.aux[0] * DUSP_CHUNK + cpos;
#pragma unroll
          for (int s = 0; s < B; s++) x[s] = src[s];
          L.st(x, I.out[0]);
          break;
        }
        case OP_STSTALE: {
          float x[B];
          L.ld(x, I.in[0]);
          float *dst = L.stale + (size_t)I.aux[0] * DUSP_CHUNK + cpos;
#pragma unroll
          for (int s = 0; s < B; s++) dst[s] = x[s];
          break;
        }
        case OP_OUT: {  // renderChannelData.js:42-44
          float x[B];
          L.ld(x, I.in[0]);
          if (kp.out_mode == 0) {
            if (active) {
              float *dst = kp.out + (vrow + I.aux[0]) * kp.out_len;
#pragma unroll
              for (int s = 0; s < B; s++)
                if (tl + s < kp.n_samples) dst[tl + s] = js_or0f(x[s]);
            }
          } else {
            // stereo mixdown: mix[c] += out_v[c % nch_v]; warp tree-sum then one atomic per sample
            const int ch = I.aux[0];
            if (ch < 2) {
#pragma unroll
              for (int s = 0; s < B; s++) {
                float y = active ? js_or0f(x[s]) : 0.0f;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) y += __shfl_xor_sync(0xffffffffu, y, o);
                if (lane == 0 && tl + s < kp.n_samples) {
                  atomicAdd(kp.out + (long long)ch * kp.out_len + tl + s, y);
                  if (kp.prog_nch == 1) atomicAdd(kp.out + kp.out_len + tl + s, y);
                }
              }
            }
          }
          break;
        }
        default: break;
      }
    }
  }
}

}  // namespace

// --------------------------------------------------------------- launcher
extern "C" int dusp_launch_render(const DuspKernelParams *kp, int sub_block, int warps_per_cta, void *stream,
                                  int *grid_out, int *smem_out) {
  const int n_warps = (kp->n_voices + 31) / 32;
  const int grid = (n_warps + warps_per_cta - 1) / warps_per_cta;
  const int block = warps_per_cta * 32;
  const size_t smem = (size_t)warps_per_cta * kp->n_slots * sub_block * 32 * sizeof(float);
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e;
  if (grid_out) *grid_out = grid;
  if (smem_out) *smem_out = (int)smem;
#define LAUNCH(BV)                                                                                         \
  do {                                                                                                     \
    e = cudaFuncSetAttribute(dusp_render_lane_per_voice<BV>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                             (int)smem);                                                                   \
    if (e != cudaSuccess) return (int)e;                                                                   \
    dusp_render_lane_per_voice<BV><<<grid, block, smem, st>>>(*kp);                                       \
  } while (0)
  switch (sub_block) {
    case 4: LAUNCH(4); break;
    case 8: LAUNCH(8); break;
    case 16: LAUNCH(16); break;
    default: return (int)cudaErrorInvalidValue;
  }
#undef LAUNCH
  return (int)cudaGetLastError();
}
