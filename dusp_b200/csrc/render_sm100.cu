// render_sm100.cu -- the device interpreter: renders every voice of a batch
// through its flat op program.  Hand-written for sm_100a (B200); compiled with
// -fmad=false so that no multiply-add is contracted: the reference computes
// every expression in IEEE double with a rounding after each operation and
// rounds to binary32 only on the store into a Float32Array.
//
// Execution model: a warp renders a TILE of V voices x T time-lanes (V*T = 32).
// Lane (vi, tl) walks B consecutive samples of voice vi; the T lanes of a voice
// cover T*B consecutive samples (T*B divides the reference's 256-sample chunk).
//   * T = 1 ("lane = voice"): every stateful unit runs in exactly the
//     reference's sequential order.  Used when the batch alone fills the GPU.
//   * T > 1: the batch has too few voices to fill 148 SMs, so time is spread
//     across lanes as well.  Stateless units are unchanged.  Phase accumulators
//     (Osc, Ramp, Noise draws, ring indices) become an exclusive warp-shuffle
//     scan of per-lane partial sums -- exact whenever the f64 partial sums are
//     exact, which holds for f32 increments of audio-range magnitude.  The biquad
//     splits into a parallel feed-forward part and a short serial recurrence that
//     the T lanes run in turn, handing (y1, y2) on by shuffle: same operations in
//     the same order as the reference, so still bit-exact.  Units with no safe
//     parallel form run lane after lane ("serial turns").
//   * Slots (the SignalChunk replacement) live in shared memory as
//     [slot][B/4][lane] float4: every access is a conflict-free LDS.128/STS.128.
//   * Unit state and parameters are f64 words [word][voice] in HBM/L2; delay
//     lines are per-voice rings in HBM; the rendered outlet goes to HBM with the
//     reference's `x || 0`.
// Warps never communicate with each other: no block-level synchronisation.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "vm.h"
#include "launch.h"
#include "ops.cuh"  // the unit cores shared with the specialised kernels

namespace {

using dops::SRD;

// ---------------------------------------------------------------- helpers
// (`x || 0`, the mixdown reductions, 10^(dB/20) and every unit's arithmetic come from ops.cuh)
using dops::js_or0f;
using dops::mix_add;
using dops::dB;

template <int B, int T>
struct Lane {
  float4 *S;             // this lane's column of the warp's slot array: slot k, quad q -> S[(k*(B/4)+q)*32]
  double *W;             // this voice's words: word w -> W[w * nv]
  unsigned nv;
  int tl;                // time-lane 0..T-1
  int lane;
  float *rings;          // this voice's ring arena
  const DuspRing *ring_tab;
  float *stale;          // this voice's stale buffers [k][256]
  const float *tables;

  __device__ __forceinline__ void ld(float (&x)[B], int ref) const {
    if (ref >= 0) {
      const float4 *s = S + ref * (B / 4) * 32;
#pragma unroll
      for (int q = 0; q < B / 4; q++) {
        const float4 v = s[q * 32];
        x[4 * q] = v.x; x[4 * q + 1] = v.y; x[4 * q + 2] = v.z; x[4 * q + 3] = v.w;
      }
    } else {
      const float c = ref == REF_NONE ? 0.0f : (float)W[(size_t)((unsigned)(REF_CONST_BASE - ref) * nv)];
#pragma unroll
      for (int s = 0; s < B; s++) x[s] = c;
    }
  }
  __device__ __forceinline__ void st(const float (&x)[B], int ref) const {
    if (ref >= 0) {
      float4 *s = S + ref * (B / 4) * 32;
#pragma unroll
      for (int q = 0; q < B / 4; q++) s[q * 32] = make_float4(x[4 * q], x[4 * q + 1], x[4 * q + 2], x[4 * q + 3]);
    }
  }
  __device__ __forceinline__ double &w(int p) const { return W[(size_t)((unsigned)p * nv)]; }
  __device__ __forceinline__ float *ring(int r, long long &len) const {
    const DuspRing rt = ring_tab[r];
    len = rt.length;
    return rings + rt.offset;
  }
  // source lane of time-lane `t` of this lane's voice
  __device__ __forceinline__ int peer(int t) const { return (lane & ~(T - 1)) | t; }
  // exclusive prefix sum over the T time-lanes of a voice (exact when the adds are exact)
  __device__ __forceinline__ double excl_scan(double x) const {
    double incl = x;
#pragma unroll
    for (int o = 1; o < T; o <<= 1) {
      const double y = __shfl_up_sync(0xffffffffu, incl, o, T);
      if (tl >= o) incl += y;
    }
    const double up = __shfl_up_sync(0xffffffffu, incl, 1, T);
    return tl == 0 ? 0.0 : up;
  }
  // serial turns: the T lanes of every voice run `body` one after the other (exact sequential order)
  template <typename F>
  __device__ __forceinline__ void serial(F body) const {
    if (T == 1) { body(); return; }
#pragma unroll 1
    for (int turn = 0; turn < T; turn++) {
      if (tl == turn) body();
      __syncwarp();
    }
  }
  __device__ __forceinline__ void sync() const { if (T > 1) __syncwarp(); }
};

// ------------------------------------------------------------ stateless ops
template <int B, int T, typename F>
__device__ __forceinline__ void map1(const Lane<B, T> &L, const DuspDevInstr &I, F f) {
  float a[B];
  L.ld(a, I.in[0]);
#pragma unroll
  for (int s = 0; s < B; s++) a[s] = f(a[s]);
  L.st(a, I.out[0]);
}
template <int B, int T, typename F>
__device__ __forceinline__ void map2(const Lane<B, T> &L, const DuspDevInstr &I, F f) {
  float a[B], b[B];
  L.ld(a, I.in[0]);
  L.ld(b, I.in[1]);
#pragma unroll
  for (int s = 0; s < B; s++) a[s] = f(a[s], b[s]);
  L.st(a, I.out[0]);
}


// ------------------------------------------------------------- stateful ops
// One copy of the arithmetic: every op below loads its operands and state words, calls the unit's core in ops.cuh (the
// same code the specialised kernels are generated from) and stores the state back.  A core returns with the END-of-tile
// state on every lane of the voice; the last time-lane writes it.
#define CX dops::Cx<B, T> cx; cx.tl = L.tl; cx.lane = L.lane;
#define TABLE(id) (L.tables + (size_t)(id) * DUSP_TABLE_LEN)

template <int B, int T>
__device__ __forceinline__ void op_osc(const Lane<B, T> &L, const DuspDevInstr &I, bool negfix) {  // Osc.js:35-47 / MultiChannelOsc.js:21-38
  float f[B];
  L.ld(f, I.in[0]);
  double ph = L.w(I.p);
  CX
  if (negfix) dops::c_osc<B, T>(cx, f, ph, TABLE(I.aux0));
  else dops::c_mcosc<B, T>(cx, f, ph, TABLE(I.aux0));
  L.sync();
  if (L.tl == T - 1) L.w(I.p) = ph;
  L.st(f, I.out[0]);
}

template <int B, int T>
__device__ __forceinline__ void op_ramp(const Lane<B, T> &L, const DuspDevInstr &I) {  // Ramp.js:25-40
  double t = L.w(I.p), playing = L.w(I.p + 1);
  float o[B];
  CX
  dops::c_ramp<B, T>(cx, o, t, playing, L.w(I.p + 2), L.w(I.p + 3), L.w(I.p + 4));
  L.sync();
  if (L.tl == T - 1) { L.w(I.p) = t; L.w(I.p + 1) = playing; }
  L.st(o, I.out[0]);
}

template <int B, int T>
__device__ __forceinline__ void op_shape(const Lane<B, T> &L, const DuspDevInstr &I) {  // Shape/index.js:28-59
  float d[B], mn[B], mx[B];
  L.ld(d, I.in[0]); L.ld(mn, I.in[1]); L.ld(mx, I.in[2]);
  const float *tab = TABLE(I.aux0);
  double t = L.w(I.p);
  const double le = (I.aux1 & 1) ? (double)__ldg(tab) : L.w(I.p + 2);
  const double re = (I.aux1 & 2) ? (double)__ldg(tab + DUSP_SR) : L.w(I.p + 3);
  CX
  dops::c_shape<B, T>(cx, d, mn, mx, t, L.w(I.p + 1), le, re, tab);
  L.sync();
  if (L.tl == T - 1) L.w(I.p) = t;
  L.st(d, I.out[0]);
}

template <int B, int T>
__device__ __forceinline__ void op_ahd(const Lane<B, T> &L, const DuspDevInstr &I) {  // AHD.js:35-76
  float a[B], h[B], d[B];
  L.ld(a, I.in[0]); L.ld(h, I.in[1]); L.ld(d, I.in[2]);
  double t = L.w(I.p), state = L.w(I.p + 1), playing = L.w(I.p + 2);
  CX
  dops::c_ahd<B, T>(cx, a, h, d, t, state, playing);
  L.sync();
  if (L.tl == T - 1) { L.w(I.p) = t; L.w(I.p + 1) = state; L.w(I.p + 2) = playing; }
  L.st(a, I.out[0]);
}

template <int B, int T>
__device__ __forceinline__ void op_timer(const Lane<B, T> &L, const DuspDevInstr &I) {  // Timer.js:37-42
  float o[B];
  double t = L.w(I.p);
  CX
  dops::c_timer<B, T>(cx, o, t);
  L.sync();
  if (L.tl == T - 1) L.w(I.p) = t;
  L.st(o, I.out[0]);
}

template <int B, int T>
__device__ __forceinline__ void op_srredux(const Lane<B, T> &L, const DuspDevInstr &I) {  // SampleRateRedux.js:20-37
  float x[B], m[B];
  L.ld(x, I.in[0]); L.ld(m, I.in[1]);
  double ts = L.w(I.p), val = L.w(I.p + 1);
  CX
  dops::c_srredux<B, T>(cx, x, m, ts, val);
  L.sync();
  if (L.tl == T - 1) { L.w(I.p) = ts; L.w(I.p + 1) = val; }
  L.st(x, I.out[0]);
}

template <int B, int T>
__device__ __forceinline__ void op_noise(const Lane<B, T> &L, const DuspDevInstr &I, bool every_sample) {  // Noise.js:16-28
  double ph = L.w(I.p), y = L.w(I.p + 1), draw = L.w(I.p + 2);
  const uint64_t seed = (uint64_t)L.w(I.p + 3) | ((uint64_t)L.w(I.p + 4) << 32);
  float f[B];
  CX
  if (every_sample) {  // constant F >= sampleRate: a draw per sample, the counter-based stream makes the lanes independent
    dops::c_noise1<B, T>(cx, f, y, draw, seed);
  } else {
    L.ld(f, I.in[0]);
    dops::c_noise<B, T>(cx, f, ph, y, draw, seed);
  }
  L.sync();
  if (L.tl == T - 1) { L.w(I.p) = ph; L.w(I.p + 1) = y; L.w(I.p + 2) = draw; }
  L.st(f, I.out[0]);
}

// Control-rate retriggering.  The decision covers one whole 256-sample chunk and must land before
// the target ticks that chunk: the host only runs these programs with whole-chunk tiles (B * T = 256).
template <int B, int T>
__device__ __forceinline__ void op_retrig(const Lane<B, T> &L, const DuspDevInstr &I) {  // Retriggerer.js:13-22
  float r[B];
  L.ld(r, I.in[0]);
  if (L.tl == 0) L.w(I.p + 1) = 0.0;
  L.sync();
  L.serial([&] {
    double t = L.w(I.p);
    bool fired = false;
#pragma unroll
    for (int s = 0; s < B; s++) {
      t += (double)r[s];
      if (t >= SRD) { fired = true; t -= SRD; }
    }
    L.w(I.p) = t;
    if (fired) L.w(I.p + 1) = 1.0;
  });
}
template <int B, int T>
__device__ __forceinline__ void op_sretrig(const Lane<B, T> &L, const DuspDevInstr &I) {  // SporadicRetrigger.js:12-16
  float r[B];
  L.ld(r, I.in[0]);
  if (L.tl == 0) {
    const double draw = L.w(I.p);
    const uint64_t seed = (uint64_t)L.w(I.p + 1) | ((uint64_t)L.w(I.p + 2) << 32);
    const bool fired = dops::tick_rand(seed, (uint64_t)draw) < (double)r[0] * (double)DUSP_CHUNK / SRD;
    L.w(I.p) = draw + 1.0;
    L.w(I.p + 3) = fired ? 1.0 : 0.0;
  }
  L.sync();
}
template <int B, int T>
__device__ __forceinline__ void op_pokeif(const Lane<B, T> &L, const DuspDevInstr &I) {  // one assignment of target.trigger()
  if (L.tl == 0 && L.w(I.aux0) != 0.0) L.w(I.p) = L.w(I.aux1);
  L.sync();
}

template <int B, int T>
__device__ __forceinline__ void op_filter(const Lane<B, T> &L, const DuspDevInstr &I) {  // Filter.js:27-51, signal-rate F
  float x[B], f[B];
  L.ld(x, I.in[0]); L.ld(f, I.in[1]);
  const int p = I.p;
  dops::FilterState S = {L.w(p), L.w(p + 1), L.w(p + 2), L.w(p + 3), L.w(p + 4), L.w(p + 5), L.w(p + 6), L.w(p + 7), L.w(p + 8),
                         L.w(p + 9), L.w(p + 10)};
  CX
  dops::c_filter<B, T>(cx, x, f, I.aux0, S);
  L.sync();
  if (L.tl == T - 1) {
    L.w(p) = S.a0; L.w(p + 1) = S.a1; L.w(p + 2) = S.a2; L.w(p + 3) = S.b1; L.w(p + 4) = S.b2; L.w(p + 5) = S.lastf;
    L.w(p + 6) = S.lastdef; L.w(p + 7) = S.x1; L.w(p + 8) = S.x2; L.w(p + 9) = S.y1; L.w(p + 10) = S.y2;
  }
  L.st(x, I.out[0]);
}

template <int B, int T>
__device__ __forceinline__ void op_filterc(const Lane<B, T> &L, const DuspDevInstr &I) {  // Filter.js:27-51, host-derived coefficients
  float x[B];
  L.ld(x, I.in[0]);
  const int p = I.p;
  double x1 = L.w(p + 5), x2 = L.w(p + 6), y1 = L.w(p + 7), y2 = L.w(p + 8);
  CX
  dops::c_filterc<B, T>(cx, x, L.w(p), L.w(p + 1), L.w(p + 2), L.w(p + 3), L.w(p + 4), x1, x2, y1, y2);
  L.sync();
  if (L.tl == T - 1) { L.w(p + 5) = x1; L.w(p + 6) = x2; L.w(p + 7) = y1; L.w(p + 8) = y2; }
  L.st(x, I.out[0]);
}

// ----------------------------------------------------------------- rings
template <int B, int T>
__device__ __forceinline__ void op_fixdelay(const Lane<B, T> &L, const DuspDevInstr &I) {  // FixedDelay.js:14-20, CombFilter.js:11-17, AllPass.js:8-15
  float x[B], g[B];
  L.ld(x, I.in[0]);
  L.ld(g, I.op != OP_FIXDELAY ? I.in[1] : REF_NONE);
  long long n;
  float *buf = L.ring(I.aux0, n);
  double tb = L.w(I.p);
  CX
  dops::c_fixdelay<B, T>(cx, x, g, buf, (int)n, tb, I.op == OP_FIXDELAY ? 0 : I.op == OP_COMB ? 1 : 2);
  L.sync();
  if (L.tl == T - 1) L.w(I.p) = tb;
  L.st(x, I.out[0]);
}

template <int B, int T>
__device__ __forceinline__ void op_delay(const Lane<B, T> &L, const DuspDevInstr &I, long long t0) {  // Delay.js:20-41 / MonoDelay.js:19-30
  float x[B], dl[B];
  L.ld(x, I.in[0]); L.ld(dl, I.in[1]);
  long long n;
  float *buf = L.ring(I.aux0, n);
  CX
  dops::c_delay<B, T>(cx, x, dl, buf, (int)n, t0, I.op == OP_MONODELAY);
  L.st(x, I.out[0]);
}

template <int B, int T>
__device__ __forceinline__ void op_readback(const Lane<B, T> &L, const DuspDevInstr &I) {  // ReadBackDelay.js:23-43
  float x[B], dl[B];
  L.ld(x, I.in[0]); L.ld(dl, I.in[1]);
  long long n;
  float *buf = L.ring(I.aux0, n);
  double t = L.w(I.p);
  CX
  dops::c_readback<B, T>(cx, x, dl, buf, (int)n, t);
  L.sync();
  if (L.tl == T - 1) L.w(I.p) = t;
  L.st(x, I.out[0]);
}

template <int B, int T>
__device__ __forceinline__ void op_cbtap(const Lane<B, T> &L, const DuspDevInstr &I, long long t0) {  // CircleBufferWriter/Reader.js:12-25, signal-rate OFFSET
  float off[B], x[B];
  L.ld(off, I.in[0]);
  const bool writer = I.op == OP_CBWRITE, has_in = writer && I.in[1] != REF_NONE;
  L.ld(x, has_in ? I.in[1] : REF_NONE);
  long long n;
  float *buf = L.ring(I.aux0, n);
  CX
  dops::c_cbtap<B, T>(cx, off, x, buf, (int)n, L.w(I.p), t0, writer, I.aux1 != 0, has_in);
  if (!writer) L.st(off, I.out[0]);
}

template <int B, int T>
__device__ __forceinline__ void op_cbtapc(const Lane<B, T> &L, const DuspDevInstr &I) {  // ... constant OFFSET: the ring index just advances
  float x[B];
  const bool writer = I.op == OP_CBWRITEC, has_in = writer && I.in[1] != REF_NONE;
  L.ld(x, has_in ? I.in[1] : REF_NONE);
  long long n;
  float *buf = L.ring(I.aux0, n);
  double idx = L.w(I.p + 1);
  CX
  dops::c_cbtapc<B, T>(cx, x, buf, (int)n, idx, writer, I.aux1 != 0, has_in);
  L.sync();
  if (L.tl == T - 1) L.w(I.p + 1) = idx;
  if (!writer) L.st(x, I.out[0]);
}
#undef CX
#undef TABLE

// ------------------------------------------------- frame-rate units (components/spectral)
// Whole-chunk tiles only (B = 8, T = 32: one voice per warp, one 256-sample chunk per tile).  Frames are per-voice
// rings that persist between a unit's ticks; the 32 lanes share every loop and meet at __syncwarp between passes.
// Kept out of line so that the register allocation of the sample-rate ops above is unaffected.
struct FrameCtx {
  float *rings;
  const DuspRing *ring_tab;
  double *W;
  unsigned nv;
  int lane;
  const double *ctab;
  __device__ __forceinline__ float *ring(int r, int &len) const { const DuspRing rt = ring_tab[r]; len = (int)rt.length; return rings + rt.offset; }
  __device__ __forceinline__ double &w(int p) const { return W[(size_t)((unsigned)p * nv)]; }
};

// fft.js (the package FFT.js:2 / IFFT.js:2 require; not part of the reference tree): radix-4 passes exactly as the
// oracle restates them (oracle/dusp_oracle.c), one butterfly per lane per step.  Every butterfly of a pass reads and
// writes elements no other butterfly of that pass writes, so the order inside a pass is free; arithmetic is the
// same f64 sequence (no contraction), stores round to the storage type like the JS array they model.
__device__ __forceinline__ int fft_bitrev(int j, int width) {
  int r = 0;
  for (int shift = 0; shift < width; shift += 2) { const int rev = width - shift - 2; if (rev >= 0) r |= ((j >> shift) & 3) << rev; }
  return r;
}
__device__ __forceinline__ int fft_width(int size) {
  int power = 0;
  for (int t = 1; size > t; t <<= 1) power++;
  return power % 2 == 0 ? power - 1 : power;
}
template <typename ST>
__device__ __forceinline__ void fft_butterfly(ST *out, const double *__restrict__ table, int A, int quarterLen, int k, double inv,
                                              double &T0r, double &T0i, double &T1r, double &T1i, double &T2r, double &T2i, double &T3r, double &T3i) {
  const int Bq = A + quarterLen, C = Bq + quarterLen, D = C + quarterLen;
  const double Ar = (double)out[A], Ai = (double)out[A + 1], Br = (double)out[Bq], Bi = (double)out[Bq + 1];
  const double Cr = (double)out[C], Ci = (double)out[C + 1], Dr = (double)out[D], Di = (double)out[D + 1];
  const double tableBr = __ldg(table + k), tableBi = inv * __ldg(table + k + 1);
  const double MBr = Br * tableBr - Bi * tableBi, MBi = Br * tableBi + Bi * tableBr;
  const double tableCr = __ldg(table + 2 * k), tableCi = inv * __ldg(table + 2 * k + 1);
  const double MCr = Cr * tableCr - Ci * tableCi, MCi = Cr * tableCi + Ci * tableCr;
  const double tableDr = __ldg(table + 3 * k), tableDi = inv * __ldg(table + 3 * k + 1);
  const double MDr = Dr * tableDr - Di * tableDi, MDi = Dr * tableDi + Di * tableDr;
  T0r = Ar + MCr; T0i = Ai + MCi; T1r = Ar - MCr; T1i = Ai - MCi;
  T2r = MBr + MDr; T2i = MBi + MDi; T3r = inv * (MBr - MDr); T3i = inv * (MBi - MDi);
}
// _transform4 with this._inv = 1 (inverseTransform), `data` = interleaved complex input
template <typename ST>
__device__ __noinline__ void fft_inverse(ST *out, const float *data, const double *__restrict__ table, int n, int lane) {
  const int size = n << 1, width = fft_width(n);
  int step = 1 << width, len = (size / step) << 1;
  const double inv = -1.0;
  const int nblk = size / len;
  if (len == 4) {
    for (int t = lane; t < nblk; t += 32) {
      const int outOff = t * len, off = fft_bitrev(t, width);
      const double evenR = data[off], evenI = data[off + 1], oddR = data[off + step], oddI = data[off + step + 1];
      out[outOff] = (ST)(evenR + oddR); out[outOff + 1] = (ST)(evenI + oddI); out[outOff + 2] = (ST)(evenR - oddR); out[outOff + 3] = (ST)(evenI - oddI);
    }
  } else {
    for (int t = lane; t < nblk; t += 32) {
      const int outOff = t * len, off = fft_bitrev(t, width), step2 = step * 2, step3 = step * 3;
      const double Ar = data[off], Ai = data[off + 1], Br = data[off + step], Bi = data[off + step + 1];
      const double Cr = data[off + step2], Ci = data[off + step2 + 1], Dr = data[off + step3], Di = data[off + step3 + 1];
      const double T0r = Ar + Cr, T0i = Ai + Ci, T1r = Ar - Cr, T1i = Ai - Ci;
      const double T2r = Br + Dr, T2i = Bi + Di, T3r = inv * (Br - Dr), T3i = inv * (Bi - Di);
      out[outOff] = (ST)(T0r + T2r); out[outOff + 1] = (ST)(T0i + T2i); out[outOff + 2] = (ST)(T1r + T3i); out[outOff + 3] = (ST)(T1i - T3r);
      out[outOff + 4] = (ST)(T0r - T2r); out[outOff + 5] = (ST)(T0i - T2i); out[outOff + 6] = (ST)(T1r - T3i); out[outOff + 7] = (ST)(T1i + T3r);
    }
  }
  __syncwarp();
  for (step >>= 2; step >= 2; step >>= 2) {
    len = (size / step) << 1;
    const int quarterLen = len >> 2, per = quarterLen >> 1, total = size >> 3;
#pragma unroll 1
    for (int idx = lane; idx < total; idx += 32) {
      const int blk = idx / per, j = idx - blk * per;
      const int A = blk * len + 2 * j, k = j * step;
      double T0r, T0i, T1r, T1i, T2r, T2i, T3r, T3i;
      fft_butterfly(out, table, A, quarterLen, k, inv, T0r, T0i, T1r, T1i, T2r, T2i, T3r, T3i);
      const int Bq = A + quarterLen, C = Bq + quarterLen, D = C + quarterLen;
      out[A] = (ST)(T0r + T2r); out[A + 1] = (ST)(T0i + T2i); out[Bq] = (ST)(T1r + T3i); out[Bq + 1] = (ST)(T1i - T3r);
      out[C] = (ST)(T0r - T2r); out[C + 1] = (ST)(T0i - T2i); out[D] = (ST)(T1r - T3i); out[D + 1] = (ST)(T1i + T3r);
    }
    __syncwarp();
  }
}
// realTransform + completeSpectrum, `data` = n reals
template <typename ST>
__device__ __noinline__ void fft_real(ST *out, const float *data, const double *__restrict__ table, int n, int lane) {
  const int size = n << 1, width = fft_width(n);
  int step = 1 << width, len = (size / step) << 1;
  const double inv = 1.0;
  const int nblk = size / len;
  if (len == 4) {
    for (int t = lane; t < nblk; t += 32) {
      const int outOff = t * len, off = fft_bitrev(t, width) >> 1, st = step >> 1;
      const double evenR = data[off], oddR = data[off + st];
      out[outOff] = (ST)(evenR + oddR); out[outOff + 1] = (ST)0.0; out[outOff + 2] = (ST)(evenR - oddR); out[outOff + 3] = (ST)0.0;
    }
  } else {
    for (int t = lane; t < nblk; t += 32) {
      const int outOff = t * len, off = fft_bitrev(t, width) >> 1, st = step >> 1;
      const double Ar = data[off], Br = data[off + st], Cr = data[off + 2 * st], Dr = data[off + 3 * st];
      const double T0r = Ar + Cr, T1r = Ar - Cr, T2r = Br + Dr, T3r = inv * (Br - Dr);
      out[outOff] = (ST)(T0r + T2r); out[outOff + 1] = (ST)0.0; out[outOff + 2] = (ST)T1r; out[outOff + 3] = (ST)(-T3r);
      out[outOff + 4] = (ST)(T0r - T2r); out[outOff + 5] = (ST)0.0; out[outOff + 6] = (ST)T1r; out[outOff + 7] = (ST)T3r;
    }
  }
  __syncwarp();
  for (step >>= 2; step >= 2; step >>= 2) {
    len = (size / step) << 1;
    const int halfLen = len >> 1, quarterLen = halfLen >> 1, hquarterLen = quarterLen >> 1;
    const int per = (hquarterLen >> 1) + 1, total = (size / len) * per;
#pragma unroll 1
    for (int idx = lane; idx < total; idx += 32) {
      const int blk = idx / per, j = idx - blk * per;
      const int outOff = blk * len, i = 2 * j, A = outOff + i, k = j * step;
      double T0r, T0i, T1r, T1i, T2r, T2i, T3r, T3i;
      fft_butterfly(out, table, A, quarterLen, k, inv, T0r, T0i, T1r, T1i, T2r, T2i, T3r, T3i);
      const int Bq = A + quarterLen, C = Bq + quarterLen;
      out[A] = (ST)(T0r + T2r); out[A + 1] = (ST)(T0i + T2i); out[Bq] = (ST)(T1r + T3i); out[Bq + 1] = (ST)(T1i - T3r);
      if (i == 0) { out[C] = (ST)(T0r - T2r); out[C + 1] = (ST)(T0i - T2i); continue; }
      if (i == hquarterLen) continue;
      const double ST0r = T1r, ST0i = -T1i, ST1r = T0r, ST1i = -T0i;
      const double ST2r = -inv * T3i, ST2i = -inv * T3r, ST3r = -inv * T2i, ST3i = -inv * T2r;
      const int SA = outOff + quarterLen - i, SB = outOff + halfLen - i;
      out[SA] = (ST)(ST0r + ST2r); out[SA + 1] = (ST)(ST0i + ST2i); out[SB] = (ST)(ST1r + ST3i); out[SB + 1] = (ST)(ST1i - ST3r);
    }
    __syncwarp();
  }
  const int half = size >> 1;  // completeSpectrum
  for (int i = 2 + 2 * lane; i < half; i += 64) { out[size - i] = out[i]; out[size - i + 1] = (ST)(-(double)out[i + 1]); }
  __syncwarp();
}

__device__ __forceinline__ double js_round_d(double x) { return floor(x + 0.5); }  // Math.round

// `x0`: sample 0 of the op's control inlet (f / threshold / shift), already broadcast to every lane
__device__ __noinline__ void frame_op(const FrameCtx &F, const DuspDevInstr &I, const float (&x)[8], double x0) {
  const int lane = F.lane;
  int n_in, n_out;
  switch (I.op) {
    case OP_HOPPER: {  // Hopper.js:19-37 (tickInterval 256: the whole input chunk enters the circular buffer)
      float *buf = F.ring(I.aux0, n_in);
      float *out = F.ring(I.aux1, n_out);
      long long t = (long long)F.w(I.p);
      const int frame = n_in;
#pragma unroll
      for (int s = 0; s < 8; s++) buf[(int)((t + lane * 8 + s) % frame)] = x[s];
      t += DUSP_CHUNK;
      __syncwarp();
      if (t % I.aux2 == 0) {
        const int rot = (int)(t % frame);
        for (int i = lane; i < frame; i += 32) { int j = i + rot; if (j >= frame) j -= frame; out[i] = buf[j]; }
      }
      if (lane == 0) F.w(I.p) = (double)t;
      break;
    }
    case OP_WINDOW: {  // Windower.js:17-23: one binary32 product per sample
      const float *in = F.ring(I.aux0, n_in);
      float *out = F.ring(I.aux1, n_out);
      const double *env = F.ctab + I.aux2;
      for (int i = lane; i < n_out; i += 32) out[i] = __fmul_rn(in[i], (float)__ldg(env + i));
      break;
    }
    case OP_FFT: {  // FFT.js:19-25, into the outlet's Float32Array
      const float *in = F.ring(I.aux0, n_in);
      float *out = F.ring(I.aux1, n_out);
      fft_real<float>(out, in, F.ctab + I.aux2, n_out >> 1, lane);
      break;
    }
    case OP_IFFT: {  // IFFT.js:21-33, through the untyped (f64) complexOut
      const float *in = F.ring(I.aux0, n_in);
      float *out = F.ring(I.aux1, n_out);
      int nc;
      double *cplx = reinterpret_cast<double *>(F.ring(I.in[0], nc));
      const int n = n_out;
      fft_inverse<double>(cplx, in, F.ctab + I.aux2, n, lane);
      const double size = (double)n;
      for (int i = lane; i < 2 * n; i += 32) cplx[i] = cplx[i] / size;
      __syncwarp();
      for (int t = lane; t < n; t += 32) out[t] = (float)cplx[2 * t];
      break;
    }
    case OP_UNHOP: {  // UnHopper.js:19-44
      const float *in = F.ring(I.aux0, n_in);
      float *out = F.ring(I.aux1, n_out);
      int na;
      double *acc = reinterpret_cast<double *>(F.ring(I.aux2, na));
      const int win = na >> 1, hop = n_out;
      long long t = (long long)F.w(I.p);
      const int rot = (int)(t % win);
      for (int i = lane; i < win; i += 32) { int j = i + rot; if (j >= win) j -= win; acc[j] += (double)in[i]; }
      t += hop;
      __syncwarp();
      if (t > hop) {
        const int r0 = (int)((t - hop) % win);
        for (int i = lane; i < hop; i += 32) { int j = (r0 + i) % win; out[i] = (float)acc[j]; acc[j] = 0.0; }
      }
      if (lane == 0) F.w(I.p) = (double)t;
      break;
    }
    case OP_RECHUNK: {  // ReChunk.js:26-49
      const float *in = F.ring(I.aux0, n_in);
      float *out = F.ring(I.aux1, n_out);
      int bs;
      float *buf = F.ring(I.aux2, bs);
      const int ii = I.in[0], oi = I.in[1];
      long long t = (long long)F.w(I.p);
      if (t % ii == 0) {
        const int rot = (int)(t % bs);
        for (int i = lane; i < ii; i += 32) buf[(rot + i) % bs] = in[i];
      }
      t += I.in[4];
      __syncwarp();
      if (t % oi == 0) {
        const int r0 = (int)((t - oi) % bs);
        for (int i = lane; i < oi; i += 32) out[i] = buf[(r0 + i) % bs];
      }
      if (lane == 0) F.w(I.p) = (double)t;
      break;
    }
    case OP_SPECPASS: {  // HardLowPass.js:20-31 / HardHighPass.js:20-31
      const float *in = F.ring(I.aux0, n_in);
      float *out = F.ring(I.aux1, n_out);
      const double cut = js_round_d(x0 / (SRD / (double)(n_out >> 1))) * 2.0;
      if (cut != cut) break;  // NaN: neither loop of the reference runs, the outlet keeps its frame
      const bool high = I.aux2 != 0;
      for (int i = lane; i < n_out; i += 32) out[i] = (((double)i < cut) != high) ? in[i] : 0.0f;
      break;
    }
    case OP_SPECGATE: {  // SpectralGate.js:15-32
      const float *in = F.ring(I.aux0, n_in);
      float *out = F.ring(I.aux1, n_out);
      const bool invert = I.aux2 != 0;
      for (int b = 2 * lane; b < n_out; b += 64) {
        const double re = in[b], im = in[b + 1], mag = sqrt(re * re + im * im);
        const bool keep = invert ? mag < x0 : mag > x0;
        out[b] = keep ? in[b] : 0.0f; out[b + 1] = keep ? in[b + 1] : 0.0f;
      }
      break;
    }
    case OP_BINSHIFT: {  // BinShift.js:14-26: out[bin + shift] = in[bin] for every bin that lands inside the frame
      const float *in = F.ring(I.aux0, n_in);
      float *out = F.ring(I.aux1, n_out);
      const double shift = js_round_d(x0) * 2.0;
      for (int d = 2 * lane; d < n_out; d += 64) {  // d = destination of the real part (bin + shift - 1)
        const double src = (double)d - shift;       // source real index (bin - 1), bin odd >= 1
        const bool ok = shift == shift && src >= 0.0 && src + 1.0 < (double)n_out;
        out[d] = ok ? in[(int)src] : 0.0f; out[d + 1] = ok ? in[(int)src + 1] : 0.0f;
      }
      break;
    }
    case OP_AUGMENT: {  // Augment.js:14-27: sequential `+=` into a Float32Array; lane 0 keeps the reference's order
      const float *in = F.ring(I.aux0, n_in);
      float *out = F.ring(I.aux1, n_out);
      for (int i = lane; i < n_out; i += 32) out[i] = 0.0f;
      __syncwarp();
      if (lane == 0) {
        const int win = n_out >> 1;
        for (int bin = 0; bin < win; bin++)
          for (int m = 0; m < I.aux2; m++) {
            const double b2 = js_round_d((double)bin * F.w(I.p + 2 * m)) * 2.0, g = F.w(I.p + 2 * m + 1);
            if (!(b2 >= 0.0 && b2 < (double)n_out)) continue;
            const int d = (int)b2;
            out[d] = (float)((double)out[d] + (double)in[bin * 2] * g);
            out[d + 1] = (float)((double)out[d + 1] + (double)in[bin * 2 + 1] * g);
          }
      }
      break;
    }
    case OP_SPECSUM: {  // SpectralSum.js:15-24
      const float *a = F.ring(I.aux0, n_in);
      float *out = F.ring(I.aux1, n_out);
      int nb;
      const float *b = F.ring(I.aux2, nb);
      for (int i = lane; i < n_out; i += 32) out[i] = __fadd_rn(a[i], b[i]);
      break;
    }
    default: break;
  }
  __syncwarp();
}

__device__ __forceinline__ void fetch_instr(DuspDevInstr &I, const DuspDevInstr *p) {
  const int4 *src = reinterpret_cast<const int4 *>(p);
  int4 *dst = reinterpret_cast<int4 *>(&I);
  dst[0] = __ldg(src); dst[1] = __ldg(src + 1); dst[2] = __ldg(src + 2);
}

// --------------------------------------------------------------- the kernel
// registers are capped so that at least 16 (B=8) / 8 (B=16) warps are resident per SM
// D > 0 (heterogeneous batches, one circuit per warp: latency-bound, so residency is what counts): D CTAs of 128 threads
// per SM, i.e. 64 registers per thread at D = 8 (a few hundred bytes of spills, measured faster: DESIGN.md section 5)
template <int B, int T, int D = 0>
__global__ void __launch_bounds__(128, D > 0 ? D : (B <= 8 ? 4 : 2)) dusp_render_tile(const __grid_constant__ DuspKernelParams kp) {
  static_assert(B % 4 == 0 && (DUSP_CHUNK % (B * T)) == 0 && 32 % T == 0, "tile must divide the chunk");
  extern __shared__ float4 smem[];
  constexpr int V = 32 / T;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int warp = blockIdx.x * (blockDim.x >> 5) + wib;
  if ((long long)warp * V >= kp.n_voices) return;
  const int vi = lane / T, tl = lane % T;
  const long long v = (long long)warp * V + vi;  // voice (allocations are padded to a multiple of 96)
  const bool active = v < kp.n_voices;

  Lane<B, T> L;
  L.S = smem + (size_t)wib * kp.n_slots * (B / 4) * 32 + lane;
  L.tl = tl;
  L.lane = lane;
  L.tables = kp.tables;
  const DuspDevInstr *code = kp.prog_global;
  bool use_inline = kp.use_inline != 0;
  int n_instr = kp.n_instr, prog_nch = kp.prog_nch;
  long long row = kp.voice0 + v;
  if (T == 32 && kp.multi) {  // heterogeneous batch: this warp's voice has its own program and arenas
    const DuspWarpDesc d = kp.multi[warp];
    L.W = d.words; L.nv = (unsigned)d.nv_stride; L.rings = d.rings; L.ring_tab = d.ring_tab; L.stale = d.stale;
    code = d.code; use_inline = false; n_instr = d.n_instr; prog_nch = d.prog_nch; row = d.out_row;
  } else {
    L.W = kp.words + v;
    L.nv = (unsigned)kp.nv_stride;
    L.rings = kp.rings + v * kp.ring_stride;
    L.ring_tab = kp.ring_tab;
    L.stale = kp.stale + v * (long long)kp.n_stale * DUSP_CHUNK;
  }

  const long long n_proc = (kp.n_samples + DUSP_CHUNK - 1) / DUSP_CHUNK * DUSP_CHUNK;  // whole chunks, like Circuit.tick
  const long long dbg_t0 = kp.dbg_cycles ? clock64() : 0;
  float *const out_row = kp.out + row * kp.out_nch * kp.out_len;

#pragma unroll 1
  for (long long tile = 0; tile < n_proc; tile += T * B) {
    const long long tl0 = tile + tl * B;   // this lane's first sample within the launch
    const int cpos = (int)(tl0 & (DUSP_CHUNK - 1));
    // programs in global memory: the next instruction's 48 bytes are requested before the current one executes (with
    // a circuit per warp the fetch is an L2 hit more often than an L1 hit: 11 % of the W4 launch's stall samples sat
    // on the switch below, waiting for I.op)
    DuspDevInstr I, Inext;
    if (!use_inline) fetch_instr(Inext, code);
#pragma unroll 1
    for (int pc = 0; pc < n_instr; pc++) {
      if (use_inline) {
        I = kp.prog_inline[pc];
      } else {
        I = Inext;
        fetch_instr(Inext, code + (pc + 1 < n_instr ? pc + 1 : 0));
      }
      switch (I.op) {
        case OP_COPY: map1<B, T>(L, I, [](float a) { return a; }); break;
        case OP_NEG: map1<B, T>(L, I, [](float a) { return -a; }); break;
        case OP_ABS: map1<B, T>(L, I, [](float a) { return fabsf(a); }); break;
        case OP_S2S: map1<B, T>(L, I, [](float a) { return (float)((double)a * SRD); }); break;
        case OP_SEMI: map1<B, T>(L, I, [](float a) { return (float)pow(2.0, (double)a / 12.0); }); break;
        case OP_DB2S: map1<B, T>(L, I, [](float a) { return (float)pow(10.0, (double)a / 20.0); }); break;
        case OP_MIDI2F: map1<B, T>(L, I, [](float a) { return (float)(pow(2.0, ((double)a - 69.0) / 12.0) * 440.0); }); break;
        case OP_FIXMUL: { const double sf = L.w(I.p); map1<B, T>(L, I, [sf](float a) { return (float)((double)a * sf); }); break; }
        // a single IEEE operation on binary32 operands: double evaluation + one rounding == native f32 op
        case OP_MUL: map2<B, T>(L, I, [](float a, float b) { return __fmul_rn(a, b); }); break;
        case OP_DIV: map2<B, T>(L, I, [](float a, float b) { return __fdiv_rn(a, b); }); break;
        case OP_ADD: map2<B, T>(L, I, [](float a, float b) { return __fadd_rn(a, b); }); break;
        case OP_SUB: map2<B, T>(L, I, [](float a, float b) { return __fsub_rn(a, b); }); break;
        case OP_ADD0: map2<B, T>(L, I, [](float a, float b) { return __fadd_rn(__fadd_rn(0.0f, a), b); }); break;
        case OP_POW: map2<B, T>(L, I, [](float a, float b) { return (float)pow((double)a, (double)b); }); break;
        case OP_GAIN: map2<B, T>(L, I, [](float a, float g) { return (float)(dB((double)g) * (double)a); }); break;
        case OP_GAINC: { const double k = L.w(I.p); map1<B, T>(L, I, [k](float a) { return (float)(k * (double)a); }); break; }
        case OP_CLIP: map2<B, T>(L, I, [](float a, float th) { return fabsf(a) > fabsf(th) ? th : a; }); break;
        case OP_HCA: map2<B, T>(L, I, [](float a, float th) { return a > th ? th : a; }); break;
        case OP_HCB: map2<B, T>(L, I, [](float a, float th) { return a < th ? th : a; }); break;
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
        case OP_PANC: {  // Pan.js:21-22 with (1-pan), (1+pan), compensation derived on the host
          float x[B], o1[B];
          L.ld(x, I.in[0]);
          const double kl = L.w(I.p), kr = L.w(I.p + 1), comp = L.w(I.p + 2);
#pragma unroll
          for (int s = 0; s < B; s++) {
            const double in = x[s];
            x[s] = (float)(in * kl / 2.0 * comp);
            o1[s] = (float)(in * kr / 2.0 * comp);
          }
          L.st(x, I.out[0]); L.st(o1, I.out[1]);
          break;
        }
        case OP_VECMAG: {  // VectorMagnitude.js:19-25
          double acc[B];
#pragma unroll
          for (int s = 0; s < B; s++) acc[s] = 0.0;
#pragma unroll
          for (int c = 0; c < 5; c++) {
            if (c < I.aux0) {
              float x[B];
              L.ld(x, I.in[c]);
#pragma unroll
              for (int s = 0; s < B; s++) acc[s] += (double)x[s] * (double)x[s];
            }
          }
          float o[B];
#pragma unroll
          for (int s = 0; s < B; s++) o[s] = (float)sqrt(acc[s]);
          L.st(o, I.out[0]);
          break;
        }
        case OP_PICK: {  // PickChannel.js:19 with a signal-rate index
          float ci[B], o[B];
          L.ld(ci, I.in[0]);
#pragma unroll
          for (int s = 0; s < B; s++) o[s] = NAN;
#pragma unroll
          for (int c = 0; c < 4; c++) {
            if (c < I.aux0) {
              float x[B];
              L.ld(x, I.in[1 + c]);
#pragma unroll
              for (int s = 0; s < B; s++) if (fmod((double)ci[s], (double)I.aux0) == (double)c) o[s] = x[s];
            }
          }
          L.st(o, I.out[0]);
          break;
        }
        case OP_OSC: op_osc<B, T>(L, I, true); break;
        case OP_MCOSC: op_osc<B, T>(L, I, false); break;
        case OP_RAMP: op_ramp<B, T>(L, I); break;
        case OP_SHAPE: op_shape<B, T>(L, I); break;
        case OP_AHD: op_ahd<B, T>(L, I); break;
        case OP_TIMER: op_timer<B, T>(L, I); break;
        case OP_SRREDUX: op_srredux<B, T>(L, I); break;
        case OP_NOISE: op_noise<B, T>(L, I, false); break;
        case OP_NOISE1: op_noise<B, T>(L, I, true); break;
        case OP_RETRIG: if (B * T == DUSP_CHUNK) op_retrig<B, T>(L, I); break;
        case OP_SRETRIG: if (B * T == DUSP_CHUNK) op_sretrig<B, T>(L, I); break;
        case OP_POKEIF: if (B * T == DUSP_CHUNK) op_pokeif<B, T>(L, I); break;
        case OP_FILTER: op_filter<B, T>(L, I); break;
        case OP_FILTERC: op_filterc<B, T>(L, I); break;
        case OP_FIXDELAY: case OP_COMB: case OP_ALLPASS: op_fixdelay<B, T>(L, I); break;
        case OP_DELAY: case OP_MONODELAY: op_delay<B, T>(L, I, kp.clock0 + tl0); break;
        case OP_READBACK: op_readback<B, T>(L, I); break;
        case OP_CBWRITE: case OP_CBREAD: op_cbtap<B, T>(L, I, kp.clock0 + tl0); break;
        case OP_CBWRITEC: case OP_CBREADC: op_cbtapc<B, T>(L, I); break;
        case OP_LDSTALE: {
          float x[B];
          const float4 *src = reinterpret_cast<const float4 *>(L.stale + (size_t)I.aux0 * DUSP_CHUNK + cpos);
#pragma unroll
          for (int q = 0; q < B / 4; q++) { const float4 t = src[q]; x[4 * q] = t.x; x[4 * q + 1] = t.y; x[4 * q + 2] = t.z; x[4 * q + 3] = t.w; }
          L.st(x, I.out[0]);
          break;
        }
        case OP_STSTALE: {
          float x[B];
          L.ld(x, I.in[0]);
          float4 *dst = reinterpret_cast<float4 *>(L.stale + (size_t)I.aux0 * DUSP_CHUNK + cpos);
#pragma unroll
          for (int q = 0; q < B / 4; q++) dst[q] = make_float4(x[4 * q], x[4 * q + 1], x[4 * q + 2], x[4 * q + 3]);
          break;
        }
        case OP_HOPPER: case OP_WINDOW: case OP_FFT: case OP_IFFT: case OP_UNHOP: case OP_RECHUNK: case OP_SPECPASS:
        case OP_SPECGATE: case OP_BINSHIFT: case OP_AUGMENT: case OP_SPECSUM: {
          if constexpr (B == 8 && T == 32) {
            const long long clock = kp.clock0 + tile;
            if (I.in[4] > 0 && clock % I.in[4] != 0) break;  // Circuit.js:35
            float x[8];
            double x0 = 0.0;
            const bool ctl = I.op == OP_HOPPER || I.op == OP_SPECPASS || I.op == OP_SPECGATE || I.op == OP_BINSHIFT;
            if (ctl) { L.ld(x, I.in[0]); x0 = (double)__shfl_sync(0xffffffffu, x[0], 0); }
            FrameCtx F;
            F.rings = L.rings; F.ring_tab = L.ring_tab; F.W = L.W; F.nv = L.nv; F.lane = lane; F.ctab = kp.ctab;
            frame_op(F, I, x, x0);
          }
          break;
        }
        case OP_LDFRAME: {  // a standard unit reads the 256-sample chunk a frame-rate unit left in its outlet
          float x[B];
          long long n;
          const float4 *src = reinterpret_cast<const float4 *>(L.ring(I.aux0, n) + cpos);
#pragma unroll
          for (int q = 0; q < B / 4; q++) { const float4 t = src[q]; x[4 * q] = t.x; x[4 * q + 1] = t.y; x[4 * q + 2] = t.z; x[4 * q + 3] = t.w; }
          L.st(x, I.out[0]);
          break;
        }
        case OP_STFRAME: {
          float x[B];
          L.ld(x, I.in[0]);
          long long n;
          float4 *dst = reinterpret_cast<float4 *>(L.ring(I.aux0, n) + cpos);
#pragma unroll
          for (int q = 0; q < B / 4; q++) dst[q] = make_float4(x[4 * q], x[4 * q + 1], x[4 * q + 2], x[4 * q + 3]);
          L.sync();
          break;
        }
        case OP_OUT: case OP_OUTRING: {  // renderChannelData.js:42-44
          float x[B];
          if (I.op == OP_OUTRING) {  // outlet.chunkSize = aux2 > 256: this chunk's part of the frame the unit left behind
            long long n;
            const float4 *src = reinterpret_cast<const float4 *>(L.ring(I.aux1, n) + (int)((kp.clock0 + tl0) % I.aux2));
#pragma unroll
            for (int q = 0; q < B / 4; q++) { const float4 t = src[q]; x[4 * q] = t.x; x[4 * q + 1] = t.y; x[4 * q + 2] = t.z; x[4 * q + 3] = t.w; }
          } else L.ld(x, I.in[0]);
if (!kp.out_raw) {
#pragma unroll
            for (int s = 0; s < B; s++) x[s] = js_or0f(x[s]);
          }
          if (kp.out_mode == 0) {
            if (active) {
              float *dst = out_row + (long long)I.aux0 * kp.out_len + tl0;
              if (tl0 + B <= kp.n_samples && ((kp.out_len & 3) | (reinterpret_cast<unsigned long long>(kp.out) & 15ull)) == 0) {  // whole sub-block in range, rows 16-byte aligned (else scalar stores)
                float4 *d4 = reinterpret_cast<float4 *>(dst);
#pragma unroll
                for (int q = 0; q < B / 4; q++) __stcs(d4 + q, make_float4(x[4 * q], x[4 * q + 1], x[4 * q + 2], x[4 * q + 3]));
              } else {
#pragma unroll
                for (int s = 0; s < B; s++) if (tl0 + s < kp.n_samples) dst[s] = x[s];
              }
            }
          } else if (I.aux0 < 2) {
            // stereo mixdown: mix[c] += out_v[c % nch_v]; tree-sum over the warp's voices, then one atomic per sample
            float ym[B];
#pragma unroll
            for (int s = 0; s < B; s++) {
              float y = active ? x[s] : 0.0f;
#pragma unroll
              for (int o = 16; o >= T; o >>= 1) y += __shfl_xor_sync(0xffffffffu, y, o);
              ym[s] = y;
            }
            if (vi == 0) {
              mix_add<B>(kp.out + (long long)I.aux0 * kp.out_len + tl0, ym, kp.n_samples - tl0);
              if (prog_nch == 1) mix_add<B>(kp.out + kp.out_len + tl0, ym, kp.n_samples - tl0);
            }
          }
          break;
        }
        default: break;
      }
    }
  }
  if (kp.dbg_cycles && lane == 0) kp.dbg_cycles[warp] = clock64() - dbg_t0;
}

template <int B, int T, int D = 0>
int launch(const DuspKernelParams *kp, int warps_per_cta, cudaStream_t st, int *grid_out, int *smem_out) {
  constexpr int V = 32 / T;
  const int n_warps = (kp->n_voices + V - 1) / V;
  const int grid = (n_warps + warps_per_cta - 1) / warps_per_cta;
  const int block = warps_per_cta * 32;
  const size_t smem = (size_t)warps_per_cta * kp->n_slots * B * 32 * sizeof(float);
  if (grid_out) *grid_out = grid;
  if (smem_out) *smem_out = (int)smem;
  cudaError_t e = cudaFuncSetAttribute(dusp_render_tile<B, T, D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  dusp_render_tile<B, T, D><<<grid, block, smem, st>>>(*kp);
  return (int)cudaGetLastError();
}

}  // namespace

// --------------------------------------------------------------- launcher
// The tile shapes are instantiated in three translation units (the same file compiled with DUSP_INST_GROUP = 0 / 1 / 2,
// see the Makefile) so that a clean build runs them side by side: the switch above is ~4 minutes of ptxas per shape set.
#ifndef DUSP_INST_GROUP
#define DUSP_INST_GROUP 0
#endif
#if DUSP_INST_GROUP == 0
#define DUSP_CONFIGS(X) X(16, 1) X(8, 1) X(8, 4) X(16, 4)
#define DUSP_GROUP_FN dusp_launch_render_g0
#elif DUSP_INST_GROUP == 1
#define DUSP_CONFIGS(X) X(8, 8) X(16, 8) X(16, 16)
#define DUSP_GROUP_FN dusp_launch_render_g1
#else
#define DUSP_CONFIGS(X) X(8, 32) X(4, 32)
#define DUSP_GROUP_FN dusp_launch_render_g2
#endif

// cudaErrorInvalidValue when the shape belongs to another group
extern "C" int DUSP_GROUP_FN(const DuspKernelParams *kp, int sub_block, int time_lanes, int warps_per_cta,
                             void *stream, int *grid_out, int *smem_out) {
#if DUSP_INST_GROUP == 2
  if (kp->multi && time_lanes == 32 && sub_block == 4 && kp->density > 0)  // heterogeneous batch: the register-capped instantiation
    return launch<4, 32, 8>(kp, warps_per_cta, (cudaStream_t)stream, grid_out, smem_out);
#endif
#define X(BV, TV) if (sub_block == BV && time_lanes == TV) return launch<BV, TV>(kp, warps_per_cta, (cudaStream_t)stream, grid_out, smem_out);
  DUSP_CONFIGS(X)
#undef X
  return (int)cudaErrorInvalidValue;
}

#if DUSP_INST_GROUP == 0
extern "C" int dusp_launch_render_g1(const DuspKernelParams *, int, int, int, void *, int *, int *);
extern "C" int dusp_launch_render_g2(const DuspKernelParams *, int, int, int, void *, int *, int *);

extern "C" int dusp_launch_supported(int sub_block, int time_lanes) {
#define X(BV, TV) if (sub_block == BV && time_lanes == TV) return 1;
  X(16, 1) X(8, 1) X(8, 4) X(16, 4) X(8, 8) X(16, 8) X(16, 16) X(8, 32) X(4, 32)
#undef X
  return 0;
}

extern "C" int dusp_launch_render(const DuspKernelParams *kp, int sub_block, int time_lanes, int warps_per_cta,
                                  void *stream, int *grid_out, int *smem_out) {
  if (time_lanes == 32) return dusp_launch_render_g2(kp, sub_block, time_lanes, warps_per_cta, stream, grid_out, smem_out);
  if (time_lanes >= 8) return dusp_launch_render_g1(kp, sub_block, time_lanes, warps_per_cta, stream, grid_out, smem_out);
  return dusp_launch_render_g0(kp, sub_block, time_lanes, warps_per_cta, stream, grid_out, smem_out);
}
#endif
