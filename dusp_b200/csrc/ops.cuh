// ops.cuh -- the per-sample arithmetic of every hot-path unit, as register-level device
// templates ("cores").  One core = one reference `_tick` for ONE channel, processing the B
// consecutive samples a lane owns, with unit state passed by reference.
//
// Two kernels are built from these cores:
//   * the specialised render kernel (fused.cc generates one straight-line kernel per circuit
//     shape that calls the cores with every slot and every state word held in registers; it is
//     compiled at session creation with NVRTC for sm_100a), and
//   * the interpreter in render_sm100.cu (its ops load operands / state words, call the cores, store the state back).
// The file must stay self-contained: NVRTC compiles it without any CUDA header.
//
// Tile model (see render_sm100.cu): a warp renders V voices x T time-lanes, lane (vi, tl) owns
// samples [tl*B, tl*B+B) of the current tile.  On return from a core EVERY lane of a voice holds
// the voice's state as of the END of the tile (cores broadcast it from the last time-lane).
//
// Rounding contract: every expression is evaluated in IEEE double exactly as the
// JavaScript does (no FMA contraction: compile with -fmad=false), rounded to binary32 only where
// the reference stores into a Float32Array.  Reference file:line is cited at each core.
#pragma once

#ifdef __CUDACC_RTC__
typedef unsigned int uint32_t;
typedef int int32_t;
typedef unsigned long long uint64_t;
typedef long long int64_t;
#ifndef NAN
#define NAN __int_as_float(0x7fffffff)
#endif
#else
#include <stdint.h>
#include <math.h>
#endif

#ifndef DUSP_SR
#define DUSP_SR 44100
#define DUSP_CHUNK 256
#define DUSP_TABLE_LEN 44101
#endif

namespace dops {

#define DOPS_DEV __device__ __forceinline__
constexpr double SRD = 44100.0;

// ---------------------------------------------------------------- helpers
DOPS_DEV double js_or0(double x) { return fabs(x) > 0.0 ? x : 0.0; }  // `x || 0`: NaN, +0, -0 -> +0 (one compare: |x| > 0 is false for all three)
DOPS_DEV float js_or0f(float x) { return fabsf(x) > 0.0f ? x : 0.0f; }

// Stereo mixdown accumulation (DUSP_OUT_MIX): y[0..B) += into dst[0..B), `room` = samples left before the requested
// length.  A lane owns B consecutive samples, so the adds go out as 16-byte vector reductions (REDG.ADD.F32x4: one L2
// atomic per four samples) when the destination is aligned and the run is whole; scalar adds otherwise.
template <int B>
DOPS_DEV void mix_add(float *dst, const float (&y)[B], long long room) {
  if (room >= B && (reinterpret_cast<unsigned long long>(dst) & 15ull) == 0) {
#pragma unroll
    for (int q = 0; q < B / 4; q++)
      asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + 4 * q), "f"(y[4 * q]), "f"(y[4 * q + 1]),
                   "f"(y[4 * q + 2]), "f"(y[4 * q + 3]) : "memory");
  } else {
#pragma unroll
    for (int s = 0; s < B; s++) if (s < room) atomicAdd(dst + s, y[s]);
  }
}

// JS `x % 44100` (= C fmod) with exact fast paths for the common range.
DOPS_DEV double fmod_sr(double x) {
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
// x mod 44100 in [0, 44100) for |x| < 1e12: exact quotient correction with one fma
DOPS_DEV double posmod_sr(double x) {
  if (!(fabs(x) < 1.0e12)) { x = fmod(x, SRD); return x < 0.0 ? x + SRD : x; }
  const double q = floor(x / SRD);
  double r = fma(-q, SRD, x);
  if (r < 0.0) r += SRD;
  if (r >= SRD) r -= SRD;
  return r;
}
// typed-array read `table[i]` with a double index: out of range / NaN -> undefined -> NaN
DOPS_DEV double tab_read(const float *__restrict__ tab, double idx) {
  return (idx >= 0.0 && idx <= 44100.0) ? (double)__ldg(tab + (int)idx) : (double)NAN;
}
DOPS_DEV bool ring_valid(double idx, double len) { return idx == floor(idx) && idx >= 0.0 && idx < len; }
DOPS_DEV double ring_read(const float *r, double len, double idx) {
  return ring_valid(idx, len) ? (double)r[(long long)idx] : (double)NAN;
}
// CircleBuffer.js:17-20 index normalisation
DOPS_DEV double cb_index(double t, double len) {
  t = floor(fmod(t, len));
  if (t < 0.0) t += len;
  return t;
}
// Injected PRNG (replaces Math.random inside a tick: Noise.js:10,23, SporadicRetrigger.js:13; include/dusp_b200.h
// dusp_rand): draw k of stream `seed` = lowbias32(low32(k) + low32(seed)) ^ high32(seed), as a fraction of 2^32.
DOPS_DEV uint32_t tick_hash(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du;
  x ^= x >> 15; x *= 0x846ca68bu;
  return x ^ (x >> 16);
}
DOPS_DEV double tick_rand(uint64_t seed, uint64_t k) {
  return (double)(tick_hash((uint32_t)k + (uint32_t)seed) ^ (uint32_t)(seed >> 32)) * (1.0 / 4294967296.0);
}
// `Math.random() * 2 - 1` (Noise.js:10,23) for a 32-bit draw u: (u - 2^31) 2^-31, exactly the signed integer
// int32(u ^ 0x80000000) scaled by a power of two -- exact in double, and its rounding to f32 (the store into the
// outlet's Float32Array) is the integer-to-float conversion's own round-to-nearest-even.
DOPS_DEV double noise_bipolar(uint32_t u) { return (double)(int32_t)(u ^ 0x80000000u) * (1.0 / 2147483648.0); }
DOPS_DEV float noise_bipolar_f32(uint32_t u) { return __int2float_rn((int32_t)(u ^ 0x80000000u)) * (1.0f / 2147483648.0f); }
DOPS_DEV double dB(double db) { return pow(10.0, db / 20.0); }  // Gain.js:26-28, Pan.js:26-28

// ------------------------------------------------------------------- lane context
template <int B, int T>
struct Cx {
  int tl;    // time-lane 0..T-1
  int lane;  // 0..31
  DOPS_DEV int peer(int t) const { return (lane & ~(T - 1)) | t; }
  DOPS_DEV double from(double x, int t) const { return T == 1 ? x : __shfl_sync(0xffffffffu, x, peer(t)); }
  DOPS_DEV float from(float x, int t) const { return T == 1 ? x : __shfl_sync(0xffffffffu, x, peer(t)); }
  // value held by the last time-lane -> every lane of the voice
  DOPS_DEV void last(double &x) const { if (T > 1) x = __shfl_sync(0xffffffffu, x, peer(T - 1)); }
  // exclusive prefix sum over the T time-lanes of a voice (exact when the adds are exact)
  DOPS_DEV double excl_scan(double x) const {
    double incl = x;
#pragma unroll
    for (int o = 1; o < T; o <<= 1) {
      const double y = __shfl_up_sync(0xffffffffu, incl, o, T);
      if (tl >= o) incl += y;
    }
    const double up = __shfl_up_sync(0xffffffffu, incl, 1, T);
    return tl == 0 ? 0.0 : up;
  }
};

// Serial turns: the T lanes of every voice run `body` one after the other, handing the listed
// state variables on by shuffle -- exactly the reference's sequential order.  __syncwarp orders
// ring (global memory) writes of one lane before the next lane's reads.
template <int B, int T, typename F>
DOPS_DEV void serial0(const Cx<B, T> &cx, F body) {
  if (T == 1) { body(); return; }
#pragma unroll 1
  for (int turn = 0; turn < T; turn++) {
    if (cx.tl == turn) body();
    __syncwarp();
  }
}
template <int B, int T, typename F>
DOPS_DEV void serial1(const Cx<B, T> &cx, double &s0, F body) {
  if (T == 1) { body(); return; }
#pragma unroll 1
  for (int turn = 0; turn < T; turn++) {
    if (cx.tl == turn) body();
    __syncwarp();
    s0 = cx.from(s0, turn);
  }
}
template <int B, int T, typename F>
DOPS_DEV void serial3(const Cx<B, T> &cx, double &s0, double &s1, double &s2, F body) {
  if (T == 1) { body(); return; }
#pragma unroll 1
  for (int turn = 0; turn < T; turn++) {
    if (cx.tl == turn) body();
    __syncwarp();
    s0 = cx.from(s0, turn); s1 = cx.from(s1, turn); s2 = cx.from(s2, turn);
  }
}

// ------------------------------------------------------------ oscillators
// Osc.js:35-47 (negfix) / MultiChannelOsc.js:21-38 (!negfix): B samples from phase `ph`.
template <int B>
DOPS_DEV void osc_run(float (&f)[B], double &ph, const float *__restrict__ tab, bool negfix) {
  if (negfix) {
    // Straight-line form for the usual case: while phase + f stays inside (-SR, 2 SR), `% SR` and the negative fix are
    // one exact add or subtract chosen by two compares (Osc.js:40-42), so the B samples run without a branch; whether
    // every phase stayed in range is checked once, and if not the run is redone from the saved phase below.
    double p = ph, q[B];
    bool ok = true;
#pragma unroll
    for (int s = 0; s < B; s++) {
      const double x = p + (double)f[s];
      ok = ok && x > -SRD && x < 2.0 * SRD;
      p = x < 0.0 ? x + SRD : (x >= SRD ? x - SRD : x);
      q[s] = p;
    }
    // (p == SR can still come out of x + SR for a tiny negative x: the literal code reads the table's end there too)
    if (ok) {
#pragma unroll
      for (int s = 0; s < B; s++) ok = ok && q[s] < SRD;
    }
    if (ok) {
#pragma unroll
      for (int s = 0; s < B; s++) {
        const int i = __double2int_rz(q[s]);
        const double fr = q[s] - (double)i;
        const double a = (double)__ldg(tab + i), b = (double)__ldg(tab + i + (fr != 0.0));
        f[s] = (float)(a * (1.0 - fr) + b * fr);
      }
      ph = p;
      return;
    }
  }
#pragma unroll
  for (int s = 0; s < B; s++) {
    ph += (double)f[s];
    ph = fmod_sr(ph);
    if (negfix && ph < 0.0) ph += SRD;
    if (ph >= 0.0 && ph < SRD) {  // the only reachable case for Osc unless f is NaN / Infinity
      const int i = __double2int_rz(ph);
      const double fr = ph - (double)i;  // phase % 1
      const double a = (double)__ldg(tab + i), b = (double)__ldg(tab + i + (fr != 0.0));  // floor / ceil entries
      f[s] = (float)(a * (1.0 - fr) + b * fr);
    } else {
      const double fl = floor(ph), ce = ceil(ph), fr = ph - trunc(ph);
      f[s] = (float)(tab_read(tab, fl) * (1.0 - fr) + tab_read(tab, ce) * fr);
    }
  }
}

// (a + b) mod 44100 for a, b in [0, 44100): exact when a + b is
DOPS_DEV double wrap_add_sr(double a, double b) {
  const double x = a + b;
  return x >= SRD ? x - SRD : x;
}

// Osc.js:35-47.  T > 1: phase at a lane's first sample = (phase + all earlier increments) mod 44100, from an
// exclusive scan over the time-lanes.  The scan works on phases REDUCED to [0, 44100) (per-lane partial sums wrap
// like the oscillator itself does, partial results combine with wrap_add_sr), so every number it forms is below
// 2^17.  It equals the reference's sample-by-sample `phase += f; phase %= SR` exactly when no addition on either
// side rounds: guaranteed here by checking that the start phase and every increment of the tile are multiples of
// 2^-36 (any f32 with |f| >= 2^-13, or 0) and |f| < 44100 -- then all intermediate values are multiples of 2^-36
// below 2^17, i.e. representable.  Tiles that fail the check (sub-audio increments below 1.3e-4 Hz, |f| >= SR,
// NaN / Infinity, a start phase with finer bits) run the lanes one after the other in the reference's order.
template <int B, int T>
DOPS_DEV void c_osc(const Cx<B, T> &cx, float (&f)[B], double &ph, const float *__restrict__ tab) {
  if (T > 1) {
    bool ok = ph >= 0.0 && ph < SRD && (ph + 65536.0) - 65536.0 == ph;  // multiple of 2^-36 (ulp of [2^16, 2^17))
    double part = 0.0;  // this lane's increments, reduced to [0, 44100)
    uint32_t umax = 0u, umin = 0x7fffffffu;  // largest |f| and smallest nonzero |f| of the lane, as bit patterns
#pragma unroll
    for (int s = 0; s < B; s++) {
      const uint32_t u = __float_as_uint(f[s]) & 0x7fffffffu;
      umax = max(umax, u);
      umin = min(umin, u - 1u);  // (0 wraps to 0xffffffff and drops out)
      const double x = part + (double)f[s];
      part = x < 0.0 ? x + SRD : (x >= SRD ? x - SRD : x);
    }
    ok = ok && umax < 0x472C4400u && umin >= 0x38FFFFFFu;  // every f is 0, or 2^-13 <= |f| < 44100 (NaN / Inf fail the first test)
    if (!__all_sync(0xffffffffu, ok)) {
      serial1<B, T>(cx, ph, [&] { osc_run<B>(f, ph, tab, true); });
      return;
    }
    double incl = part;
#pragma unroll
    for (int o = 1; o < T; o <<= 1) {
      const double y = __shfl_up_sync(0xffffffffu, incl, o, T);
      if (cx.tl >= o) incl = wrap_add_sr(incl, y);
    }
    const double before = __shfl_up_sync(0xffffffffu, incl, 1, T);
    if (cx.tl > 0) ph = wrap_add_sr(ph, before);
  }
  osc_run<B>(f, ph, tab, true);
  cx.last(ph);
}

// MultiChannelOsc.js:21-38: truncated-modulo phase has no canonical range -> sequential order only
template <int B, int T>
DOPS_DEV void c_mcosc(const Cx<B, T> &cx, float (&f)[B], double &ph, const float *__restrict__ tab) {
  serial1<B, T>(cx, ph, [&] { osc_run<B>(f, ph, tab, false); });
}

// -------------------------------------------------------------- envelopes
// Ramp.js:25-40
// IEEE a / b for many a and one b (Ramp.js:37: t / duration, B times per lane by the same duration).  With the
// correctly rounded reciprocal y = RN(1 / b):  q = RN(a y),  r = a - b q (one FMA: exact),  RN(q + r y) = RN(a / b)
// (Markstein's theorem; tests/kat_div.c checks 4e8 cases against the hardware's division, among them the significands
// the proof is tight for).  Three FP64 operations instead of the ~25 of a division.  Outside the guarded range (zero,
// subnormal or huge operands, NaN) the literal division runs: signs of zero and overflow behave as IEEE says there.
struct SharedDiv {
  double b, y;
  bool fast;
  DOPS_DEV explicit SharedDiv(double b_) : b(b_), y(1.0 / b_) { const double ab = fabs(b_); fast = ab >= 3.0549363634996047e-151 && ab <= 3.273390607896142e150; }
  DOPS_DEV double operator()(double a) const {
    const double aa = fabs(a);
    if (!(fast && aa >= 3.0549363634996047e-151 && aa <= 3.273390607896142e150)) return a / b;
    const double q = a * y, r = __fma_rn(-q, b, a);
    return __fma_rn(r, y, q);
  }
};

template <int B, int T>
DOPS_DEV void c_ramp(const Cx<B, T> &cx, float (&o)[B], double &t, double &playing, double dur, double y0, double y1) {
  const SharedDiv over_dur(dur);
  if (T > 1 && cx.tl > 0 && playing != 0.0) {
    // the tl*B steps before this lane's first sample: one literal step, then the closed form of
    // "count up by one until t passes duration" (after the first step 0 <= t <= duration holds)
    t += 1.0;
    if (t > dur) { playing = 0.0; t = dur; }
    if (t < 0.0) { playing = 0.0; t = 0.0; }
    if (playing != 0.0) {
      const double rem = (double)(cx.tl * B - 1);
      if (t + rem > dur) { t = dur; playing = 0.0; } else t += rem;
    }
  }
#pragma unroll
  for (int s = 0; s < B; s++) {
    if (playing != 0.0) {
      t += 1.0;
      if (t > dur) { playing = 0.0; t = dur; }
      if (t < 0.0) { playing = 0.0; t = 0.0; }
    }
    o[s] = (float)(y0 + over_dur(t) * (y1 - y0));
  }
  cx.last(t); cx.last(playing);
}

// Shape/index.js:28-59.  d: duration in, output out.  t += 1/duration is an inexact f64 sum: serial.
template <int B, int T>
DOPS_DEV void c_shape(const Cx<B, T> &cx, float (&d)[B], const float (&mn)[B], const float (&mx)[B], double &t,
                      double playing, double le, double re, const float *__restrict__ tab) {
  // only the f64 running sum t += 1/duration is order-dependent: the divisions before it and the table
  // interpolation after it run on every time-lane at once
  double tt[B];
#pragma unroll
  for (int s = 0; s < B; s++) tt[s] = 1.0 / (double)d[s];
  serial1<B, T>(cx, t, [&] {
#pragma unroll
    for (int s = 0; s < B; s++) {
      if (playing != 0.0) t += tt[s];
      tt[s] = t;
    }
  });
#pragma unroll
  for (int s = 0; s < B; s++) {
    const double ts = tt[s], lo = mn[s], hi = mx[s];
    double y;
    if (ts <= 0.0) y = le * (hi - lo) + lo;
    else if (ts > SRD) y = re * (hi - lo) + lo;
    else {
      const double fr = ts - trunc(ts);  // t % 1
      y = lo + (hi - lo) * (tab_read(tab, ceil(ts)) * fr + tab_read(tab, floor(ts)) * (1.0 - fr));
    }
    d[s] = (float)y;
  }
}

// AHD.js:35-76.  a: attack in, output out.
template <int B, int T>
DOPS_DEV void c_ahd(const Cx<B, T> &cx, float (&a)[B], const float (&h)[B], const float (&d)[B], double &t,
                    double &state_d, double &playing_d) {
  serial3<B, T>(cx, t, state_d, playing_d, [&] {
    int state = (int)state_d;
    bool playing = playing_d != 0.0;
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
    state_d = (double)state; playing_d = playing ? 1.0 : 0.0;
  });
}

// Timer.js:37-42
template <int B, int T>
DOPS_DEV void c_timer(const Cx<B, T> &cx, float (&o)[B], double &t) {
  serial1<B, T>(cx, t, [&] {
    const double sp = 1.0 / SRD;
#pragma unroll
    for (int s = 0; s < B; s++) { t += sp; o[s] = (float)t; }
  });
}

// SampleRateRedux.js:20-37, one channel.  x: input in, held value out.
template <int B, int T>
DOPS_DEV void c_srredux(const Cx<B, T> &cx, float (&x)[B], const float (&amt)[B], double &ts, double &val) {
  double unused = 0.0;
  serial3<B, T>(cx, ts, val, unused, [&] {
#pragma unroll
    for (int s = 0; s < B; s++) {
      ts += 1.0;
      if (ts > (double)amt[s]) { val = (double)x[s]; ts = 0.0; }
      x[s] = (float)val;
    }
  });
}

// ------------------------------------------------------------------ noise
// Noise.js:16-28.  f: frequency in, output out.
template <int B, int T>
DOPS_DEV void c_noise(const Cx<B, T> &cx, float (&f)[B], double &ph, double &y, double &draw, uint64_t seed) {
  serial3<B, T>(cx, ph, y, draw, [&] {
#pragma unroll
    for (int s = 0; s < B; s++) {
      ph += (double)f[s];
      if (ph >= SRD) {
        ph = 0.0;
        y = noise_bipolar(tick_hash((uint32_t)(uint64_t)draw + (uint32_t)seed) ^ (uint32_t)(seed >> 32));
        draw += 1.0;
      }
      f[s] = (float)y;
    }
  });
}
// Noise.js:16-28 with constant F >= sampleRate: a new draw on every sample; the counter-based
// stream makes the draws of different time-lanes independent.
template <int B, int T>
DOPS_DEV void c_noise1(const Cx<B, T> &cx, float (&o)[B], double &y, double &draw, uint64_t seed) {
  const uint32_t x0 = (uint32_t)(uint64_t)draw + (uint32_t)(cx.tl * B) + (uint32_t)seed;
  const uint32_t hi = (uint32_t)(seed >> 32);
  uint32_t u = 0;
#pragma unroll
  for (int s = 0; s < B; s++) { u = tick_hash(x0 + (uint32_t)s) ^ hi; o[s] = noise_bipolar_f32(u); }
  y = noise_bipolar(u);  // the unit's held value (Noise.js:23): the last draw, unrounded
  draw += (double)(T * B);
  cx.last(y);
}

// ---------------------------------------------------------------- filters
// Filter.js:66-84
DOPS_DEV void filter_coeffs(int hp, double f, double &a0, double &a1, double &a2, double &b1, double &b2) {
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

struct FilterState {  // Filter.js:17-20 + coefficient cache (:29-32)
  double a0, a1, a2, b1, b2, lastf, lastdef, x1, x2, y1, y2;
};

// Filter.js:27-51 with a signal-rate F: coefficients follow f[t], strictly sequential.
template <int B, int T>
DOPS_DEV void c_filter(const Cx<B, T> &cx, float (&x)[B], const float (&f)[B], int hp, FilterState &S) {
  auto body = [&] {
#pragma unroll
    for (int s = 0; s < B; s++) {
      const double ft = f[s];
      if (S.lastdef == 0.0 || ft != S.lastf) { S.lastdef = 1.0; S.lastf = ft; filter_coeffs(hp, ft, S.a0, S.a1, S.a2, S.b1, S.b2); }
      const double xin = x[s];
      const double y = S.a0 * xin + S.a1 * js_or0(S.x1) + S.a2 * js_or0(S.x2) - S.b1 * js_or0(S.y1) - S.b2 * js_or0(S.y2);
      const float yf = (float)y;
      x[s] = yf;
      S.y2 = js_or0(S.y1); S.y1 = (double)yf; S.x2 = js_or0(S.x1); S.x1 = xin;
    }
  };
  if (T == 1) { body(); return; }
#pragma unroll 1
  for (int turn = 0; turn < T; turn++) {
    if (cx.tl == turn) body();
    S.a0 = cx.from(S.a0, turn); S.a1 = cx.from(S.a1, turn); S.a2 = cx.from(S.a2, turn);
    S.b1 = cx.from(S.b1, turn); S.b2 = cx.from(S.b2, turn); S.lastf = cx.from(S.lastf, turn);
    S.lastdef = cx.from(S.lastdef, turn); S.x1 = cx.from(S.x1, turn); S.x2 = cx.from(S.x2, turn);
    S.y1 = cx.from(S.y1, turn); S.y2 = cx.from(S.y2, turn);
  }
}

// Filter.js:27-51 with host-derived coefficients (constant F).
//   y[n] = ((a0 x + a1 x1) + a2 x2) - b1 y1 - b2 y2
// The bracket depends only on inputs (filterc_ff: any number of samples at once).  The two subtractions and the f32
// rounding are the recurrence (filterc_rec): strictly one sample after the other.
// State convention: x1 raw previous input, x2 / y2 already `|| 0`-normalised, y1 previous output (f32-valued);
// o1 / o2 / q1 / q2 are their normalised forms carried in registers.
template <int B>
DOPS_DEV void filterc_ff(const float (&x)[B], double (&c)[B], double a0, double a1, double a2, double &o1, double &o2) {
#pragma unroll
  for (int s = 0; s < B; s++) {
    const double xin = x[s];
    c[s] = a0 * xin + a1 * o1 + a2 * o2;
    o2 = o1; o1 = fabsf(x[s]) > 0.0f ? xin : 0.0;  // (double)(x || 0) from the one conversion: NaN, +0, -0 -> +0
  }
}
// MAGIC: the f32 rounding of y (the store into the outlet's Float32Array, Filter.js:43) without leaving the FP64 pipe.
// For t with exponent E, (t + M) - M with M = 1.5 * 2^(E+29) rounds t to a multiple of 2^(E-23) -- its 24-bit
// significand, ties to even like the conversion -- and the result is already the double the next sample needs, so
// neither conversion (2 x ~18 cycles on the XU pipe) sits on the recurrence.  Exact whenever every output of the run
// is a normal f32 (then `|| 0` changes nothing either); checked once per run from the min / max of the results'
// exponent words; anything else (zeros, denormals, overflow, NaN) redoes the run in the literal form.  Measured: only
// pays where the warp issues nothing but the recurrence (the pipelined kernel's S-warp after the feed-forward split).
template <int B, bool MAGIC>
DOPS_DEV void filterc_rec(const double (&c)[B], float (&y)[B], double b1, double b2, double &q1, double &q2) {
  if (MAGIC) {
    const double s1 = q1, s2 = q2;
    uint32_t lo_ = 0xffffffffu, hi_ = 0u;
    double b2q2 = b2 * q2;
#pragma unroll
    for (int s = 0; s < B; s++) {
      const double t = c[s] - b1 * q1 - b2q2;
      const double M = __hiloint2double((int)(((uint32_t)__double2hiint(t) & 0x7ff00000u) + 0x01D80000u), 0);
      const double r = (t + M) - M;
      const uint32_t rh = (uint32_t)__double2hiint(r) & 0x7fffffffu;
      lo_ = min(lo_, rh); hi_ = max(hi_, rh);
      b2q2 = b2 * q1;
      q2 = q1; q1 = r;
      y[s] = (float)r;  // exact
    }
    if (lo_ >= 0x38100000u && hi_ < 0x47F00000u) return;  // 2^-126 <= |y| < 2^128 throughout
    q1 = s1; q2 = s2;
  }
  double b2q2 = b2 * q2;
#pragma unroll
  for (int s = 0; s < B; s++) {
    const float yf = (float)(c[s] - b1 * q1 - b2q2);
    y[s] = yf;
    b2q2 = b2 * q1;  // next sample's b2*y2, ready before it is needed
    const double yd = (double)yf;
    q2 = q1; q1 = fabsf(yf) > 0.0f ? yd : 0.0;  // (double)(y || 0): the select runs beside the conversion, not ahead of it
  }
}

// T > 1: the bracket on all lanes at once; the T lanes of a voice run the recurrence in turn, handing (y1, y2) on by
// shuffle -- the same operations in the same order as the reference, so bit-exact.
template <int B, int T>
DOPS_DEV void c_filterc(const Cx<B, T> &cx, float (&x)[B], double a0, double a1, double a2, double b1, double b2,
                        double &x1, double &x2, double &y1, double &y2) {
  double o1 = js_or0(x1), o2 = js_or0(x2), q1 = js_or0(y1), q2 = js_or0(y2);
  if (T > 1) {  // inputs preceding this lane's first sample come from the previous time-lane
    const float pm1 = __shfl_up_sync(0xffffffffu, x[B - 1], 1, T), pm2 = __shfl_up_sync(0xffffffffu, x[B - 2], 1, T);
    if (cx.tl > 0) { o1 = (double)js_or0f(pm1); o2 = (double)js_or0f(pm2); }
  }
  double c[B];
  const double xlast = x[B - 1];
  filterc_ff<B>(x, c, a0, a1, a2, o1, o2);
  if (T == 1) {
#ifdef DOPS_MAGIC
    filterc_rec<B, true>(c, x, b1, b2, q1, q2);
#else
    filterc_rec<B, false>(c, x, b1, b2, q1, q2);
#endif
  } else {
    // T turns.  Every lane runs the recurrence in every turn on its own brackets with the state broadcast from the lane
    // whose turn it was (a warp instruction costs the same with one lane active or all): only the lane whose turn it is
    // keeps the results -- selects instead of a divergent branch, and the state needs no second hand-over.
#pragma unroll 2
    for (int turn = 0; turn < T; turn++) {
      float yt[B];
      double a1 = q1, a2 = q2;
      filterc_rec<B, false>(c, yt, b1, b2, a1, a2);
      const bool mine = cx.tl == turn;
#pragma unroll
      for (int s = 0; s < B; s++) x[s] = mine ? yt[s] : x[s];
      q1 = cx.from(a1, turn); q2 = cx.from(a2, turn);
    }
  }
  x1 = xlast; x2 = o2; y1 = (double)x[B - 1]; y2 = q2;
  cx.last(x1); cx.last(x2); cx.last(y1); cx.last(y2);
}

// ------------------------------------------------------------------ segments of one voice's timeline on different warps
// A batch too small to fill the GPU (W2: 1 000 voices = 1 000 warps) is bound by each warp's own latency.  Where every
// stateful unit of a circuit can be put at an arbitrary sample without running the samples before it, the timeline of a
// launch is cut into segments rendered by different warps (fused.cc, "segmented" kernels).  That holds for Ramp (a
// counter: ramp_jump), Noise at F >= sampleRate (counter-based draws) and Osc -- whose phase at a segment's start is the
// start phase plus the wrapped sum of all earlier increments, exact under c_osc's own conditions: a first kernel
// evaluates only what feeds the oscillators' F inlets and leaves each segment's wrapped total (c_osc_total) and
// whether the conditions held; voices where they did not are rendered in one piece, in the reference's order.
//
// Exact total of a tile's increments, wrapped to [0, 44100) -- the first half of c_osc; ok: c_osc's conditions on f.
template <int B, int T>
DOPS_DEV double c_osc_total(const Cx<B, T> &cx, const float (&f)[B], bool &ok) {
  double part = 0.0;
  uint32_t umax = 0u, umin = 0x7fffffffu;
#pragma unroll
  for (int s = 0; s < B; s++) {
    const uint32_t u = __float_as_uint(f[s]) & 0x7fffffffu;
    umax = max(umax, u);
    umin = min(umin, u - 1u);
    const double x = part + (double)f[s];
    part = x < 0.0 ? x + SRD : (x >= SRD ? x - SRD : x);
  }
  ok = __all_sync(0xffffffffu, umax < 0x472C4400u && umin >= 0x38FFFFFFu);
#pragma unroll
  for (int o = 1; o < T; o <<= 1) part = wrap_add_sr(part, __shfl_xor_sync(0xffffffffu, part, o));
  return part;
}
DOPS_DEV bool osc_phase_exact(double ph) { return ph >= 0.0 && ph < SRD && (ph + 65536.0) - 65536.0 == ph; }
// Ramp.js:25-33 advanced by n samples without output: one literal step, then the closed form of "count up by one
// until t passes duration" (the same two steps c_ramp takes for a lane's earlier samples).
DOPS_DEV void ramp_jump(double &t, double &playing, double dur, long long n) {
  if (n <= 0 || playing == 0.0) return;
  t += 1.0;
  if (t > dur) { playing = 0.0; t = dur; }
  if (t < 0.0) { playing = 0.0; t = 0.0; }
  if (playing != 0.0) {
    const double rem = (double)(n - 1);
    if (t + rem > dur) { t = dur; playing = 0.0; } else t += rem;
  }
}

// ------------------------------------------------------------------ rings
// A warp whose 32 lanes walk ONE ring along time (T = 32): lane tl owns the B consecutive ring positions
// [i0 + tl*B, +B).  The tap offset is arbitrary, so i0 is in general not a multiple of 4 and a lane's run is not
// 16-byte aligned; scalar accesses at a stride of B floats between lanes touch 32 sectors per instruction for 128
// useful bytes (the L1 pipeline of the SM was the bound of the pipelined kernel, profiles/r1_v13_piped_ncu_summary.txt).
// Instead every lane moves the ALIGNED float4s [i0 - m + tl*B, +B) (m = i0 & 3, the same for all lanes) and the m
// elements that belong to a neighbour's run travel by shuffle; only the window's first and last partial float4 use
// scalar accesses (lane 0 / lane 31).  Requires a 16-byte aligned ring base and i0 + 32*B <= n (no wrap inside the
// window): the callers check both (warp-uniform) and otherwise keep the element-wise path.
// Two phases, so that a kernel can ISSUE the loads early and only touch the data where the unit ticks (the shuffles
// are the first consumers: doing them at once exposes the whole memory latency, profiles/r2_c5_piped_roles.txt).
template <int B>
DOPS_DEV void ring_load_issue(float (&f)[B + 3], const float *buf, int i0, int tl) {
  const int m = i0 & 3;
  const float4 *b4 = reinterpret_cast<const float4 *>(buf + (i0 - m)) + tl * (B / 4);
#pragma unroll
  for (int q = 0; q < B / 4; q++) { const float4 t = b4[q]; f[4 * q] = t.x; f[4 * q + 1] = t.y; f[4 * q + 2] = t.z; f[4 * q + 3] = t.w; }
#pragma unroll
  for (int j = 0; j < 3; j++) f[B + j] = 0.0f;
  if (tl == 31 && m > 0) {
#pragma unroll
    for (int j = 0; j < 3; j++) if (j < m) f[B + j] = buf[i0 - m + 32 * B + j];     // (inside the window: index <= i0 + 32 B - 1)
  }
}
template <int B>
DOPS_DEV void ring_load_finish(float (&x)[B], float (&f)[B + 3], int i0, int tl) {
  const int m = i0 & 3;
#pragma unroll
  for (int j = 0; j < 3; j++) {  // the next lane's first elements (lane 31 keeps the window's tail it loaded itself)
    const float nx = __shfl_down_sync(0xffffffffu, f[j], 1);
    if (tl != 31) f[B + j] = nx;
  }
  switch (m) {
    case 0:
#pragma unroll
      for (int s = 0; s < B; s++) x[s] = f[s];
      break;
    case 1:
#pragma unroll
      for (int s = 0; s < B; s++) x[s] = f[s + 1];
      break;
    case 2:
#pragma unroll
      for (int s = 0; s < B; s++) x[s] = f[s + 2];
      break;
    default:
#pragma unroll
      for (int s = 0; s < B; s++) x[s] = f[s + 3];
  }
}
template <int B>
DOPS_DEV void ring_load_shifted(float (&x)[B], const float *buf, int i0, int tl) {
  float f[B + 3];
  ring_load_issue<B>(f, buf, i0, tl);
  ring_load_finish<B>(x, f, i0, tl);
}
template <int B>
DOPS_DEV void ring_store_shifted(const float (&y)[B], float *buf, int i0, int tl) {
  const int m = i0 & 3;
  float p[3];  // the previous lane's last three values
#pragma unroll
  for (int j = 0; j < 3; j++) p[j] = __shfl_up_sync(0xffffffffu, y[B - 3 + j], 1);
  float f[B];  // the aligned run [i0 - m + tl*B, +B): m elements of the previous lane, then this lane's first B - m
  switch (m) {
    case 0:
#pragma unroll
      for (int k = 0; k < B; k++) f[k] = y[k];
      break;
    case 1:
      f[0] = p[2];
#pragma unroll
      for (int k = 1; k < B; k++) f[k] = y[k - 1];
      break;
    case 2:
      f[0] = p[1]; f[1] = p[2];
#pragma unroll
      for (int k = 2; k < B; k++) f[k] = y[k - 2];
      break;
    default:
      f[0] = p[0]; f[1] = p[1]; f[2] = p[2];
#pragma unroll
      for (int k = 3; k < B; k++) f[k] = y[k - 3];
  }
  float *a = buf + (i0 - m) + tl * B;
  if (tl == 0 && m > 0) {  // the window's first float4 is partial: positions before i0 are not ours
#pragma unroll
    for (int k = 1; k < 4; k++) if (k >= m) a[k] = f[k];
  } else {
    *reinterpret_cast<float4 *>(a) = make_float4(f[0], f[1], f[2], f[3]);
  }
#pragma unroll
  for (int q = 1; q < B / 4; q++) reinterpret_cast<float4 *>(a)[q] = make_float4(f[4 * q], f[4 * q + 1], f[4 * q + 2], f[4 * q + 3]);
  if (tl == 31) {          // ... and so is its last one
#pragma unroll
    for (int j = 0; j < 3; j++) if (j < m) a[B + j] = y[B - m + j];
  }
}
// Measured on B200 (W3, 10 000 voices, per launch): 5.53 ms against 5.86 ms for the element-wise path.  (With the
// round-1 splitmix64 noise the same change cost time, 6.42 against 5.87 ms: the P-warps' extra shuffles and selects
// then competed with the lane-per-voice warp for issue slots.)  DOPS_NO_SHIFT (DUSP_FUSED_DEFS) restores the old path.
#ifndef DOPS_NO_SHIFT
DOPS_DEV bool ring_base_aligned(const float *buf) { return (reinterpret_cast<unsigned long long>(buf) & 15ull) == 0; }
#else
DOPS_DEV bool ring_base_aligned(const float *) { return false; }
#endif

// FixedDelay.js:14-20 (KIND 0), CombFilter.js:11-17 (1), AllPass.js:8-15 (2).  x: input in, output out.
// (KIND and the flags of the ring cores below are ordinary arguments: the specialised kernels pass literals, which fold
// away after inlining; the interpreter passes what its instruction says.)
template <int B, int T>
DOPS_DEV void c_fixdelay(const Cx<B, T> &cx, float (&x)[B], const float (&g)[B], float *buf, int n, double &tb, const int KIND) {
  const double len = (double)n;
  auto step = [&](int p, int s) {
    const float d = buf[p];
    if (KIND == 0) { buf[p] = x[s]; x[s] = d; }
    else if (KIND == 1) { buf[p] = (float)((double)x[s] + (double)d * (double)g[s]); x[s] = d; }
    else { buf[p] = (float)((double)x[s] + (double)d * (double)g[s]); x[s] = (float)((double)d - (double)x[s] * (double)g[s]); }
  };
  if (T > 1 && n >= T * B && tb == floor(tb) && tb >= 0.0 && tb < len) {
    // every sample of the tile touches its own ring position: no dependence between time-lanes
    int p = (int)tb + 1 + cx.tl * B;
    if (p >= n) p -= n;
#pragma unroll
    for (int s = 0; s < B; s++) { step(p, s); if (++p == n) p = 0; }
    double e = (double)(p == 0 ? n - 1 : p - 1);
    cx.last(e);
    tb = e;
    __syncwarp();
  } else if (T > 1 && n >= 32 && tb == floor(tb) && tb >= 0.0 && tb < len) {
    // a ring shorter than the tile (e.g. a 2.4 ms AllPass, 106 samples, under 128-sample tiles): sample j of the tile meets
    // what sample j - n left, nothing else -- so the tile runs in ceil(T*B / n) <= 4 rounds of n independent samples
    // instead of T serial turns (which made such circuits the tail of the W4 launch: 20 % of their instructions)
    const int base = (int)tb + 1, j0 = cx.tl * B;
    for (int r0 = 0; r0 < T * B; r0 += n) {
#pragma unroll
      for (int s = 0; s < B; s++) {
        const int j = j0 + s;
        if (j >= r0 && j < r0 + n) step((base + j) % n, s);
      }
      __syncwarp();
    }
    tb = (double)((base + T * B - 1) % n);
  } else {
    serial1<B, T>(cx, tb, [&] {
#pragma unroll
      for (int s = 0; s < B; s++) {
        tb += 1.0;
        if (!(tb < len)) tb = fmod(tb, len);  // (tBuffer + 1) % length
        step((int)tb, s);
      }
    });
  }
}

// Delay.js:20-41 (MONO false) / MonoDelay.js:19-30 (MONO true).  Write-ahead scatter-add: later
// samples read what earlier ones added -> strictly sequential.  t0 = absolute index of x[0].
template <int B, int T>
DOPS_DEV void c_delay(const Cx<B, T> &cx, float (&x)[B], const float (&dl)[B], float *buf, int n, long long t0, const bool MONO) {
  const double len = (double)n;
  serial0<B, T>(cx, [&] {
    // Fast path: the delay is one value over the lane's B samples, at least B + 1 samples long, and neither the
    // read window [tb, tb+B) nor the write window [floor(tb+d), +B+1) wraps or meets the other.  Then the B reads
    // see the ring as of the start of the run, and the 2B scatter-adds touch B + 1 consecutive slots in a fixed
    // order -- done in registers in exactly that order: independent loads and stores instead of a chain of 4B
    // dependent memory round trips.  Anything else takes the literal loop below.
    {
      bool same = true;
#pragma unroll
      for (int s = 1; s < B; s++) same = same && dl[s] == dl[0];
      double tb0 = (double)t0;
      if (!(tb0 < len)) tb0 = fmod(tb0, len);
      const double dd = (double)dl[0];
      if (same && dd >= (double)(B + 1) && dd <= len - (double)(B + 2) && tb0 + (double)B <= len) {
        double tw0 = tb0 + dd;
        if (tw0 >= len) tw0 -= len;
        const double fr = tw0 - trunc(tw0);
        const long long fl0 = (long long)tw0;
        const int ceo = fr != 0.0 ? 1 : 0;
        if (fl0 + B + ceo <= (long long)n) {
          float *rd = buf + (long long)tb0, *wr = buf + fl0;
          float y[B], w[B + 1];
#pragma unroll
          for (int s = 0; s < B; s++) y[s] = rd[s];
#pragma unroll
          for (int j = 0; j < B + 1; j++) w[j] = (j < B || ceo) ? wr[j] : 0.0f;
#pragma unroll
          for (int s = 0; s < B; s++) {
            w[s] = (float)((double)w[s] + (double)x[s] * (1.0 - fr));
            if (ceo) w[s + 1] = (float)((double)w[s + 1] + (double)x[s] * fr);
            else w[s] = (float)((double)w[s] + (double)x[s] * fr);
          }
#pragma unroll
          for (int j = 0; j < B + 1; j++) if (j < B || ceo) wr[j] = w[j];
#pragma unroll
          for (int s = 0; s < B; s++) { rd[s] = 0.0f; x[s] = y[s]; }
          return;
        }
      }
    }
#pragma unroll
    for (int s = 0; s < B; s++) {
      double tb = (double)(t0 + s);
      if (!(tb < len)) tb = fmod(tb, len);  // (clock + t) % length
      float y = 0.0f;
      if (!MONO) { y = (float)ring_read(buf, len, tb); if (ring_valid(tb, len)) buf[(long long)tb] = 0.0f; }
      double tw = tb + (double)dl[s];  // (tBuffer + delay) % length, exact fast paths for the usual range
      if (tw >= len) { tw = tw < 2.0 * len ? tw - len : fmod(tw, len); }
      else if (!(tw > -len)) tw = fmod(tw, len);
      const double fr = tw - trunc(tw);
      const double fl = floor(tw);
      double ce = ceil(tw);
      if (MONO && !(ce < len && ce > -len)) ce = fmod(ce, len);
      if (ring_valid(fl, len)) buf[(long long)fl] = (float)((double)buf[(long long)fl] + (double)x[s] * (1.0 - fr));
      if (ring_valid(ce, len)) buf[(long long)ce] = (float)((double)buf[(long long)ce] + (double)x[s] * fr);
      if (MONO) { y = (float)ring_read(buf, len, tb); if (ring_valid(tb, len)) buf[(long long)tb] = 0.0f; }
      x[s] = y;
    }
  });
}

// ReadBackDelay.js:23-43
template <int B, int T>
DOPS_DEV void c_readback(const Cx<B, T> &cx, float (&x)[B], const float (&dl)[B], float *buf, int n, double &t) {
  const double len = (double)n;
  serial1<B, T>(cx, t, [&] {
#pragma unroll
    for (int s = 0; s < B; s++) {
      const double wi = fmod(t + len, len);
      if (ring_valid(wi, len)) buf[(long long)wi] = x[s];
      x[s] = (float)ring_read(buf, len, fmod(t - (double)dl[s] + len, len));
      t += 1.0;
    }
  });
}

// CircleBufferWriter.js:12-25 / CircleBufferReader.js:12-25 with a signal-rate OFFSET (admitted only
// for a ring with a single tap unit): sequential.  off: offsets in; for a reader, output out.
template <int B, int T>
DOPS_DEV void c_cbtap(const Cx<B, T> &cx, float (&off)[B], const float (&x)[B], float *buf, int n, double tu, long long t0,
                      const bool WRITER, const bool WIPE, const bool HAS_IN) {
  const double len = (double)n;
  serial0<B, T>(cx, [&] {
#pragma unroll
    for (int s = 0; s < B; s++) {
      const double tt = WRITER ? tu + (double)(t0 + s) + SRD * (double)off[s] : tu + (double)(t0 + s) - SRD * (double)off[s];
      const double i = cb_index(tt, len);
      float y = NAN;
      if (ring_valid(i, len)) {
        if (WRITER) {
          float v = buf[(long long)i];
          if (WIPE) v = 0.0f;                                      // preWipe: write(c, tWrite, 0)
          if (HAS_IN) v = (float)((double)v + (double)x[s]);       // mix: +=
          buf[(long long)i] = v;
        } else {
          y = buf[(long long)i];
          if (WIPE) buf[(long long)i] = 0.0f;                      // postWipe
        }
      }
      off[s] = y;
    }
  });
}

// Same units with a constant OFFSET: the ring index advances by one per sample; its start value is
// derived on the host (CircleBuffer.js:17-20 applied to the first access).  idx < 0: never valid.
// x: writer input in / reader output out.
template <int B, int T>
DOPS_DEV void c_cbtapc(const Cx<B, T> &cx, float (&x)[B], float *buf, int n, double &idx, const bool WRITER, const bool WIPE,
                       const bool HAS_IN) {
  auto run = [&](int i) {
    if (i + B <= n) {  // no wrap inside the run: plain pointer walk
      float *q = buf + i;
      if (!WRITER) {
#pragma unroll
        for (int s = 0; s < B; s++) x[s] = q[s];
        if (WIPE) {
#pragma unroll
          for (int s = 0; s < B; s++) q[s] = 0.0f;
        }
      } else if (WIPE) {                                           // preWipe: write(0) then mix -> 0 + in
#pragma unroll
        for (int s = 0; s < B; s++) q[s] = HAS_IN ? __fadd_rn(0.0f, x[s]) : 0.0f;
      } else if (HAS_IN) {                                         // mix: `+=` (one f32 add == f64 add + rounding)
        float v[B];
#pragma unroll
        for (int s = 0; s < B; s++) v[s] = q[s];
#pragma unroll
        for (int s = 0; s < B; s++) q[s] = __fadd_rn(v[s], x[s]);
      }
      i += B;
      return i == n ? 0 : i;
    }
#pragma unroll
    for (int s = 0; s < B; s++) {  // the run wraps: sample by sample (x[] stays in registers: constant indices)
      if (WRITER) {
        float v = WIPE ? 0.0f : buf[i];
        if (HAS_IN) v = __fadd_rn(v, x[s]);
        if (HAS_IN || WIPE) buf[i] = v;
      } else {
        x[s] = buf[i];
        if (WIPE) buf[i] = 0.0f;
      }
      if (++i == n) i = 0;
    }
    return i;
  };
  if (idx < 0.0) {
    if (!WRITER) {
#pragma unroll
      for (int s = 0; s < B; s++) x[s] = NAN;
    }
  } else if (T == 32 && (int)idx + T * B <= n && ring_base_aligned(buf)) {
    // the warp's window does not wrap: aligned 16-byte accesses, neighbours' elements by shuffle (see ring_load_shifted)
    const int i0 = (int)idx;
    if (!WRITER) {
      ring_load_shifted<B>(x, buf, i0, cx.tl);
      if (WIPE) {
        float z[B];
#pragma unroll
        for (int s = 0; s < B; s++) z[s] = 0.0f;
        __syncwarp();                                              // every lane has read before any lane wipes
        ring_store_shifted<B>(z, buf, i0, cx.tl);
      }
    } else if (WIPE) {                                             // preWipe: write(0) then mix -> 0 + in
      float v[B];
#pragma unroll
      for (int s = 0; s < B; s++) v[s] = HAS_IN ? __fadd_rn(0.0f, x[s]) : 0.0f;
      ring_store_shifted<B>(v, buf, i0, cx.tl);
    } else if (HAS_IN) {                                           // mix: `+=`
      float v[B];
      ring_load_shifted<B>(v, buf, i0, cx.tl);
#pragma unroll
      for (int s = 0; s < B; s++) v[s] = __fadd_rn(v[s], x[s]);
      __syncwarp();
      ring_store_shifted<B>(v, buf, i0, cx.tl);
    }
    const int e = i0 + T * B;
    idx = (double)(e == n ? 0 : e);
    __syncwarp();
  } else if (T == 1 || n >= T * B) {  // distinct positions within the tile: time-lanes are independent
    int i = (int)idx + cx.tl * B;
    if (i >= n) i %= n;
    double e = (double)run(i);
    cx.last(e);
    idx = e;
    if (T > 1) __syncwarp();
  } else {
    serial1<B, T>(cx, idx, [&] { idx = (double)run((int)idx); });
  }
}

// A plain CircleBufferReader with a constant OFFSET, split in two so that a kernel can issue the ring
// loads early (before unrelated work of the same step) and consume them where the unit ticks:
// c_cbpeek_issue starts the loads of what c_cbtapc<reader> would read (raw aligned runs on the shifted path),
// c_cbpeek_finish turns them into the lane's B samples, c_cbadv advances the ring index like the unit does.
// Only valid when no other tap stores into those positions during the step (compile.cc: aux[2]).
template <int B, int T>
DOPS_DEV bool cbpeek_shifted(const float *buf, int n, double idx) { return T == 32 && (int)idx + T * B <= n && ring_base_aligned(buf); }
template <int B, int T>
DOPS_DEV void c_cbpeek_issue(const Cx<B, T> &cx, float (&f)[B + 3], const float *buf, int n, double idx) {
  if (idx < 0.0) return;
  if (cbpeek_shifted<B, T>(buf, n, idx)) { ring_load_issue<B>(f, buf, (int)idx, cx.tl); return; }
  int i = (int)idx + cx.tl * B;
  if (i >= n) i %= n;
  if (i + B <= n) {
#pragma unroll
    for (int s = 0; s < B; s++) f[s] = buf[i + s];
  } else {
#pragma unroll
    for (int s = 0; s < B; s++) { f[s] = buf[i]; if (++i == n) i = 0; }
  }
}
template <int B, int T>
DOPS_DEV void c_cbpeek_finish(const Cx<B, T> &cx, float (&x)[B], float (&f)[B + 3], const float *buf, int n, double idx) {
  if (idx < 0.0) {
#pragma unroll
    for (int s = 0; s < B; s++) x[s] = NAN;
    return;
  }
  if (cbpeek_shifted<B, T>(buf, n, idx)) { ring_load_finish<B>(x, f, (int)idx, cx.tl); return; }
#pragma unroll
  for (int s = 0; s < B; s++) x[s] = f[s];
}
// pulls the ring lines the NEXT step's c_cbpeek will read from HBM into L2 (no register cost)
template <int B, int T>
DOPS_DEV void c_cbprefetch(const Cx<B, T> &cx, const float *buf, int n, double idx) {
  if (idx < 0.0) return;
  const int i = (int)(((long long)idx + T * B + cx.tl * B) % n);
  asm volatile("prefetch.global.L2 [%0];" ::"l"(buf + i));
}
template <int B, int T>
DOPS_DEV void c_cbadv(int n, double &idx) {
  if (idx >= 0.0) {
    int i = (int)idx + T * B;
    if (i >= n) i = T * B <= n ? i - n : i % n;
    idx = (double)i;
  }
}

}  // namespace dops
