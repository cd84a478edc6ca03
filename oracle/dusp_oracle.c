/*
 * dusp_oracle.c -- CPU restatement of the Dusp reference render path.
 *
 * TEST INFRASTRUCTURE ONLY.  This file is the parity oracle (and the timed CPU
 * baseline "port").  Nothing in the product (dusp_b200/, libdusp_b200.so) may
 * include, link or call it.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs load liboracle.
 *
 * PARITY STATUS: "parity unpinned by the reference's own tests" -- the
 * reference (plain JavaScript, /root/reference) ships no golden vectors, no
 * assertions and cannot run here (no node / JS engine in the image).  The
 * oracle is pinned instead against closed-form known answers derived from the
 * reference source (tests/test_oracle_kat.py).
 *
 * Every function cites the reference file:line it restates.  Rounding rules
 * (SURVEY 8c): all arithmetic in double, a (float) cast exactly where the JS
 * stores into a Float32Array, reads of f32 buffers widen to double, JS `%` is
 * fmod, compile with -ffp-contract=off.  The scheduler (Circuit.add /
 * computeProcessIndex / computeOrders) is restated verbatim so that cyclic
 * graphs get the same one-chunk-stale edges as the reference.
 *
 * Input is the neutral "DUSPGRAPH 1" text blob written by
 * dusp_b200/graphblob.py (format documented in DESIGN.md section 3).
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define SR 44100            /* src/config.js:7 */
#define CHUNK 256           /* src/config.js:6 */
#define MAXCH 16
#define MAXPORTS 24

/* ---------------------------------------------------------------- chunks */
/* src/SignalChunk.js:1-11 -- channelData: Float32Array(chunkSize)[nChannels] */
typedef struct Chunk {
  int nch;
  long size;   /* Piglet.js:10 chunkSize: 256 unless the port says otherwise (spectral units) */
  float *ch[MAXCH];
} Chunk;

static Chunk *chunk_new_sized(int nch, long size) {
  Chunk *c = (Chunk *)calloc(1, sizeof(Chunk));
  c->nch = nch; c->size = size;
  for (int i = 0; i < nch; i++) c->ch[i] = (float *)calloc(size, sizeof(float));
  return c;
}
static Chunk *chunk_new(int nch) { return chunk_new_sized(nch, CHUNK); }
static float *chunk_ensure(Chunk *c, int i) { /* `out[c] = out[c] || new Float32Array(chunkSize)` */
  if (i >= MAXCH) { fprintf(stderr, "oracle: too many channels\n"); abort(); }
  while (c->nch <= i) { c->ch[c->nch] = (float *)calloc(c->size, sizeof(float)); c->nch++; }
  return c->ch[i];
}
static void chunk_free(Chunk *c) {
  if (!c) return;
  for (int i = 0; i < c->nch; i++) {
    /* Delay.js:23 may alias one Float32Array under two indices */
    int dup = 0;
    for (int j = 0; j < i; j++) if (c->ch[j] == c->ch[i]) dup = 1;
    if (!dup) free(c->ch[i]);
  }
  free(c);
}

/* src/CircleBuffer.js:3-35 */
typedef struct CircleBuffer {
  int nch;
  long len;
  float **ch;
} CircleBuffer;

/* JS typed-array element access with a double index: non-integer, negative or
 * out-of-range reads give `undefined` (NaN once it meets arithmetic or a
 * Float32Array store); such writes are silently dropped. */
static int ta_valid(double idx, long len) { return idx == floor(idx) && idx >= 0 && idx < (double)len; }
static double ta_read(const float *a, long len, double idx) { return ta_valid(idx, len) ? (double)a[(long)idx] : NAN; }
static void ta_write(float *a, long len, double idx, double v) { if (ta_valid(idx, len)) a[(long)idx] = (float)v; }

static double cb_index(const CircleBuffer *b, double t) { /* CircleBuffer.js:17-20 */
  t = floor(fmod(t, (double)b->len));
  while (t < 0) t += (double)b->len;
  return t;
}
static double cb_read(const CircleBuffer *b, int c, double t) { return ta_read(b->ch[c], b->len, cb_index(b, t)); }
static void cb_write(CircleBuffer *b, int c, double t, double y) { ta_write(b->ch[c], b->len, cb_index(b, t), y); }
static void cb_mix(CircleBuffer *b, int c, double t, double y) { /* CircleBuffer.js:29-34 `+=` */
  double i = cb_index(b, t);
  if (ta_valid(i, b->len)) b->ch[c][(long)i] = (float)((double)b->ch[c][(long)i] + y);
}

/* ---------------------------------------------------------------- graph */
enum Kind {
  K_OSC, K_MCOSC, K_RAMP, K_SHAPE, K_AHD, K_TIMER, K_NOISE, K_FILTER,
  K_MULTIPLY, K_SUM, K_SUBTRACT, K_DIVIDE, K_GAIN, K_POLARITYINVERT, K_ABS, K_CLIP,
  K_HARDCLIPABOVE, K_HARDCLIPBELOW, K_POW, K_RESCALE, K_CROSSFADER, K_PAN, K_REPEATER,
  K_FIXEDMULTIPLY, K_SEMITONETORATIO, K_MIDITOFREQUENCY, K_DECIBELTOSCALER, K_SECONDSTOSAMPLES,
  K_CONCATCHANNELS, K_PICKCHANNEL, K_VECTORMAGNITUDE,
  K_DELAY, K_MONODELAY, K_READBACKDELAY, K_FIXEDDELAY, K_COMBFILTER, K_ALLPASS,
  K_CBWRITER, K_CBREADER, K_RETRIGGERER, K_SPORADICRETRIGGERER, K_SAMPLERATEREDUX, K_MONITOR,
  K_HOPPER, K_WINDOWER, K_FFT, K_IFFT, K_UNHOPPER, K_RECHUNK, K_HARDLOWPASS, K_HARDHIGHPASS, K_SPECTRALGATE, K_BINSHIFT,
  K_AUGMENT, K_SPECTRALSUM, K_NKINDS
};
static const char *KIND_NAMES[K_NKINDS] = {
  "Osc", "MultiChannelOsc", "Ramp", "Shape", "AHD", "Timer", "Noise", "Filter",
  "Multiply", "Sum", "Subtract", "Divide", "Gain", "PolarityInvert", "Abs", "Clip",
  "HardClipAbove", "HardClipBelow", "Pow", "Rescale", "CrossFader", "Pan", "Repeater",
  "FixedMultiply", "SemitoneToRatio", "MidiToFrequency", "DecibelToScaler", "SecondsToSamples",
  "ConcatChannels", "PickChannel", "VectorMagnitude",
  "Delay", "MonoDelay", "ReadBackDelay", "FixedDelay", "CombFilter", "AllPass",
  "CircleBufferWriter", "CircleBufferReader", "Retriggerer", "SporadicRetriggerer", "SampleRateRedux", "Monitor",
  "Hopper", "Windower", "FFT", "IFFT", "UnHopper", "ReChunk", "HardLowPass", "HardHighPass", "SpectralGate", "BinShift",
  "Augment", "SpectralSum"};

struct Unit;
typedef struct Port {
  char name[32];
  int mono;          /* Piglet.js:13-17 exposeAsMono */
  int nodata;        /* Unit.js:88-98 chain ports */
  int connected;
  int src_unit, src_outlet; /* inlets only */
  Chunk *chunk;      /* owned by outlets and unconnected inlets; aliased by connected inlets (Inlet.js:53) */
  int owns;
  int nconn, conn_unit[64]; /* outlets: connections in push order (Outlet.js / Inlet.js:52) */
} Port;

typedef struct Unit {
  int kind;
  char label[48];
  int nin, nout;
  Port in[MAXPORTS], out[MAXPORTS];
  int process_index; /* -1 = undefined (Unit.js:171-209) */
  int in_circuit;
  /* generic numeric / string properties */
  int nprops;
  char pname[24][24];
  double pval[24];
  int nsprops;
  char sname[4][24];
  char sval[4][32];
  /* run-time state */
  double phase, t, y, y0, y1, duration, left_edge, right_edge, sf, comp_db;
  int playing, state, finished, right_edge_is_shape, left_edge_is_shape;
  uint64_t seed, draw;
  const float *table;        /* Osc wave table / Shape table */
  double mc_phase[MAXCH]; int mc_phase_set[MAXCH];
  /* Filter.js:17-20,27-51 */
  int fkind; int lastf_defined; double lastf, a0, a1, a2, b1, b2;
  double x1[MAXCH], x2[MAXCH], fy1[MAXCH], fy2[MAXCH]; int sdef[MAXCH][4];
  /* delays */
  long maxdelay; float *bufs[MAXCH]; long buflen; double tbuffer;
  CircleBuffer *cb; int prewipe, postwipe;
  /* SampleRateRedux.js:9-10 */
  double srr_val[MAXCH]; int srr_nval; double srr_ts;
  /* frame-rate units (components/spectral): Circuit.js:34-37 ticks a unit when clock % tickInterval == 0 */
  long tick_interval, hop, win, frame, in_iv, out_iv;
  double *dbuf;            /* untyped JS Arrays: Hopper / UnHopper / ReChunk `buffer[0]`, IFFT `complexOut` */
  float *env;              /* Windower.envelopeBuffer (Float32Array) */
  struct FFTPlan *fft;
  int invert, nmap; double map_ratio[8], map_gain[8];
} Unit;

/* An Event (Event.js:3-30) whose function only assigns numeric unit state, e.g. the trigger() queued by
 * scheduleTrigger / the `at` operator (UnitOrPatch.js:25-33): runs before the tick of chunk `chunk`
 * (Circuit.js:23,57-65: events with t < clock + 256). */
typedef struct OEvent { long chunk; int unit; char prop[24]; double value; } OEvent;

/* One assignment made by `target.trigger()` of a (Sporadic)Retriggerer (Retriggerer.js:17-18): `chunk` holds the
 * retriggering unit's index instead of a time. */
typedef struct Oracle {
  struct Oracle *next; long next_chunk; /* graph edit: the circuit as it stands from chunk next_chunk on */
  int next_fin_unit;                    /* >= 0: ... made by this unit finishing by itself inside a tick, whenever that happens */
  int ntriggers; OEvent *triggers;
  int nevents; OEvent *events;
  int nunits;
  Unit *units;
  int ncb;
  CircleBuffer *cbs;
  int norder, *order;
  int has_order;
  int root_unit, root_outlet;
  double clock;
  char err[256];
} Oracle;

/* ------------------------------------------------------------ wave tables */
/* src/components/Osc/waveTables.js:5-40 -- 44101-entry float32 tables. */
static float *T_SIN, *T_SAW, *T_TRI, *T_SQUARE, *T_8BIT;
/* src/components/Shape/shapeTables.js:3-38 */
static float *S_DECAY, *S_ATTACK, *S_SEMISINE, *S_DECAYSQ;
static int tables_ready = 0;

static void build_tables(void) {
  if (tables_ready) return;
  const int N = SR + 1;
  const double PHI = 2 * M_PI;
  T_SIN = (float *)calloc(N, 4); T_SAW = (float *)calloc(N, 4); T_TRI = (float *)calloc(N, 4);
  T_SQUARE = (float *)calloc(N, 4); T_8BIT = (float *)calloc(N, 4);
  for (int t = 0; t < N; t++) T_SIN[t] = (float)sin(PHI * t / N);          /* waveTables.js:6-8 */
  for (int t = 0; t < SR; t++) T_SAW[t] = (float)(-1 + t * 2.0 / N);       /* :10-12 (last entry stays 0) */
  { int q = SR / 4;                                                        /* :14-22 */
    for (int t = 0; t < q; t++) {
      T_TRI[t] = (float)((double)t / SR * 4);
      T_TRI[t + q] = (float)(1 - (double)T_TRI[t]);
      T_TRI[t + q * 2] = (float)(-(double)T_TRI[t]);
      T_TRI[t + q * 3] = (float)(-1 + (double)T_TRI[t]);
    }
    T_TRI[SR] = 0; }
  for (int t = 0; t < N; t++) T_SQUARE[t] = t < SR / 2 ? 1.f : -1.f;        /* :24-26 */
  for (int t = 0; t < N; t++) {                                            /* :28-31 Math.round = floor(x+0.5) */
    double s = (double)T_SIN[t] * 128.0;
    T_8BIT[t] = (float)(floor(s + 0.5) / 128.0);
  }
  S_DECAY = (float *)calloc(N, 4); S_ATTACK = (float *)calloc(N, 4);
  S_SEMISINE = (float *)calloc(N, 4); S_DECAYSQ = (float *)calloc(N, 4);
  for (int x = 0; x < N; x++) {                                            /* shapeTables.js:3-9,20-37 */
    double u = (double)x / SR;
    S_DECAY[x] = (float)(1 - u);
    S_ATTACK[x] = (float)u;
    S_SEMISINE[x] = (float)sin(M_PI * u);
    S_DECAYSQ[x] = (float)((1 - u) * (1 - u));
  }
  tables_ready = 1;
}

const float *oracle_wave_table(const char *name) {
  build_tables();
  if (!strcmp(name, "sin") || !strcmp(name, "sine")) return T_SIN;
  if (!strcmp(name, "saw")) return T_SAW;
  if (!strcmp(name, "square")) return T_SQUARE;
  if (!strcmp(name, "triangle")) return T_TRI;
  if (!strcmp(name, "8bit")) return T_8BIT;
  return NULL;
}
const float *oracle_shape_table(const char *name) {
  build_tables();
  if (!strcmp(name, "decay")) return S_DECAY;
  if (!strcmp(name, "attack")) return S_ATTACK;
  if (!strcmp(name, "semiSine")) return S_SEMISINE;
  if (!strcmp(name, "decaySquared")) return S_DECAYSQ;
  return NULL;
}

/* Injected PRNG replacing Math.random inside a tick (Noise.js:10,23, SporadicRetrigger.js:13): counter-based,
 * keyed by (unit seed, draw index), 32 bits of resolution (a legal Math.random: the result is only ever rounded to
 * a Float32Array sample or compared with a probability).  Draw k of stream `seed`:
 *   x = low32(k) + low32(seed);  x ^= x >> 16;  x *= 0x7feb352d;  x ^= x >> 15;  x *= 0x846ca68b;  x ^= x >> 16;
 *   u = x ^ high32(seed);        rand = u / 2^32
 * (the 32-bit integer finaliser "lowbias32"; the stream's low word offsets the counter, its high word the output). */
double oracle_rand(uint64_t seed, uint64_t k) {
  uint32_t x = (uint32_t)k + (uint32_t)seed;
  x ^= x >> 16; x *= 0x7feb352dU;
  x ^= x >> 15; x *= 0x846ca68bU;
  x ^= x >> 16;
  x ^= (uint32_t)(seed >> 32);
  return (double)x * (1.0 / 4294967296.0);
}

/* ----------------------------------------------------------- blob parser */
static double *prop(Unit *u, const char *name) {
  for (int i = 0; i < u->nprops; i++) if (!strcmp(u->pname[i], name)) return &u->pval[i];
  return NULL;
}
static double propd(Unit *u, const char *name, double dflt) { double *p = prop(u, name); return p ? *p : dflt; }
static const char *sprop(Unit *u, const char *name, const char *dflt) {
  for (int i = 0; i < u->nsprops; i++) if (!strcmp(u->sname[i], name)) return u->sval[i];
  return dflt;
}
static Port *find_in(Unit *u, const char *name) {
  for (int i = 0; i < u->nin; i++) if (!strcmp(u->in[i].name, name)) return &u->in[i];
  return NULL;
}
static Port *find_out(Unit *u, const char *name) {
  for (int i = 0; i < u->nout; i++) if (!strcmp(u->out[i].name, name)) return &u->out[i];
  return NULL;
}

static int kind_from_name(const char *s) {
  for (int i = 0; i < K_NKINDS; i++) if (!strcmp(KIND_NAMES[i], s)) return i;
  return -1;
}

void oracle_free(Oracle *o);

struct FFTPlan;
static struct FFTPlan *fft_plan(long size);
static void fft_plan_free(struct FFTPlan *f);
static int fail(Oracle *o, const char *msg) { snprintf(o->err, sizeof o->err, "%s", msg); return -1; }

static int init_unit_state(Oracle *o, Unit *u) {
  u->phase = propd(u, "phase", 0);
  u->t = propd(u, "t", 0);
  u->playing = (int)propd(u, "playing", 0);
  switch (u->kind) {
  case K_OSC: case K_MCOSC:
    u->table = oracle_wave_table(sprop(u, "waveform", "sin"));
    if (!u->table) return fail(o, "waveform doesn't exist"); /* Osc.js:60-61 */
    break;
  case K_RAMP: /* Ramp.js:3-14 */
    u->duration = propd(u, "duration", SR); u->y0 = propd(u, "y0", 1); u->y1 = propd(u, "y1", 0);
    break;
  case K_SHAPE: { /* Shape/index.js:7-24 */
    u->table = oracle_shape_table(sprop(u, "shape", "decay"));
    if (!u->table) return fail(o, "invalid shape function");
    const char *le = sprop(u, "leftEdge", NULL), *re = sprop(u, "rightEdge", NULL);
    u->left_edge_is_shape = le && !strcmp(le, "shape");
    u->right_edge_is_shape = re ? !strcmp(re, "shape") : (prop(u, "rightEdge") == NULL);
    u->left_edge = propd(u, "leftEdge", 0);
    u->right_edge = propd(u, "rightEdge", 0);
    u->finished = (int)propd(u, "finished", 0);
    break; }
  case K_AHD: u->state = (int)propd(u, "state", 0); break; /* AHD.js:19-21 */
  case K_NOISE: /* Noise.js:3-11 */
    u->seed = (uint64_t)propd(u, "seed_lo", 0) | ((uint64_t)propd(u, "seed_hi", 0) << 32);
    u->draw = (uint64_t)propd(u, "draw", 1);
    u->y = propd(u, "y", 0);
    break;
  case K_SPORADICRETRIGGERER: /* Math.random -> the unit's own injected stream */
    u->seed = (uint64_t)propd(u, "seed_lo", 0) | ((uint64_t)propd(u, "seed_hi", 0) << 32);
    u->draw = (uint64_t)propd(u, "draw", 0);
    break;
  case K_FILTER: { /* Filter.js:56-64 */
    const char *k = sprop(u, "kind", "LP");
    if (!strcmp(k, "LP")) u->fkind = 0; else if (!strcmp(k, "HP")) u->fkind = 1;
    else return fail(o, "invalid filter type (only LP/HP are callable in the reference)");
    u->a0 = u->a1 = u->a2 = u->b1 = u->b2 = NAN; /* calculateCoefficients(undefined) */
    u->lastf_defined = 0;
    break; }
  case K_PAN: u->comp_db = propd(u, "compensationDB", 1.5); break; /* Pan.js:12 */
  case K_SAMPLERATEREDUX: u->srr_val[0] = 0; u->srr_nval = 1; u->srr_ts = propd(u, "timeSinceLastUpdate", INFINITY); break; /* SampleRateRedux.js:9-10 */
  case K_FIXEDMULTIPLY: u->sf = propd(u, "sf", NAN); break;
  case K_DELAY: u->maxdelay = (long)propd(u, "maxDelay", SR * 5); break;            /* Delay.js:13-14 */
  case K_MONODELAY: u->maxdelay = SR * 5; u->bufs[0] = (float *)calloc(u->maxdelay, 4); break; /* MonoDelay.js:9-10 */
  case K_READBACKDELAY: u->buflen = (long)propd(u, "bufferLength", SR); u->tbuffer = propd(u, "tBuffer", 0); break;
  case K_FIXEDDELAY: case K_COMBFILTER: case K_ALLPASS: /* FixedDelay.js:22-28 */
    u->buflen = (long)propd(u, "delayTimeInSamples", 0);
    if (u->buflen < 1) return fail(o, "Cannot have fixed delay of length 0 samples");
    u->bufs[0] = (float *)calloc(u->buflen, 4);
    u->tbuffer = propd(u, "tBuffer", 0);
    break;
  case K_CBWRITER: case K_CBREADER: { /* CircleBufferNode.js:7-31 */
    int bi = (int)propd(u, "buffer", -1);
    if (bi < 0 || bi >= o->ncb) return fail(o, "CircleBuffer node without buffer");
    u->cb = &o->cbs[bi];
    u->prewipe = (int)propd(u, "preWipe", 0);
    u->postwipe = (int)propd(u, "postWipe", 0);
    if (u->kind == K_CBREADER) /* set buffer(): grow out to buffer.numberOfChannels */
      for (int c = 0; c < u->cb->nch; c++) chunk_ensure(u->out[0].chunk, c);
    break; }
  case K_HOPPER: /* Hopper.js:5-17 */
    u->hop = (long)propd(u, "hopSize", 0); u->frame = (long)propd(u, "frameSize", 0);
    if (u->hop < 1 || u->frame < 1) return fail(o, "Hopper needs hopSize and frameSize");
    u->dbuf = (double *)calloc(u->frame, sizeof(double));
    break;
  case K_WINDOWER: { /* Windower.js:4-15,28-47 */
    u->win = (long)propd(u, "windowSize", 0);
    if (u->win < 1) return fail(o, "Windower constructor expects a windowSize");
    if (strcmp(sprop(u, "windowKind", "hamming"), "hamming")) return fail(o, "Window type is not defined.");
    u->env = (float *)calloc(u->win, sizeof(float));
    for (long n = 0; n < u->win; n++) u->env[n] = (float)pow(sin((M_PI * (double)n) / (double)(u->win - 1)), 2);
    break; }
  case K_FFT: case K_IFFT: /* FFT.js:5-17, IFFT.js:5-19 */
    u->win = (long)propd(u, "windowSize", 0); u->frame = u->win * 2;
    if (u->win < 2 || (u->win & (u->win - 1))) return fail(o, "FFT size must be a power of two and bigger than 1");
    u->fft = fft_plan(u->win);
    u->dbuf = (double *)calloc(u->frame * 3, sizeof(double));
    break;
  case K_UNHOPPER: /* UnHopper.js:4-17 */
    u->hop = (long)propd(u, "hopSize", 0); u->win = (long)propd(u, "windowSize", 0);
    if (u->hop < 1 || u->win < 1) return fail(o, "UnHopper needs hopSize and windowSize");
    u->dbuf = (double *)calloc(u->win, sizeof(double));
    break;
  case K_RECHUNK: /* ReChunk.js:6-24 */
    u->in_iv = (long)propd(u, "inputInterval", 0); u->out_iv = (long)propd(u, "outputInterval", 0);
    u->frame = (long)propd(u, "bufferSize", 0);
    if (u->in_iv < 1 || u->out_iv < 1 || u->frame < 1) return fail(o, "ReChunk expects 2 numeric contructor arguments");
    u->dbuf = (double *)calloc(u->frame, sizeof(double));
    break;
  case K_HARDLOWPASS: case K_HARDHIGHPASS: case K_SPECTRALGATE: case K_BINSHIFT: case K_AUGMENT: case K_SPECTRALSUM:
    u->win = (long)propd(u, "windowSize", 4096); u->frame = u->win * 2; /* SpectralUnit.js:12-16 */
    u->invert = (int)propd(u, "invert", 1);
    u->nmap = (int)propd(u, "nmap", 0);
    for (int m = 0; m < u->nmap && m < 8; m++) {
      char nm[24];
      snprintf(nm, sizeof nm, "ratio%d", m); u->map_ratio[m] = propd(u, nm, 1);
      snprintf(nm, sizeof nm, "gain%d", m); u->map_gain[m] = propd(u, nm, 1);
    }
    break;
  default: break;
  }
  u->tick_interval = (long)propd(u, "tickInterval", CHUNK);
  if (u->tick_interval < 1) return fail(o, "unit with a tick interval below one sample");
  return 0;
}

/* chunkSize of a unit's outlet (the `chunkSize` option its constructor passes to addOutlet) */
static long outlet_chunk_size(Unit *u) {
  switch (u->kind) {
  case K_HOPPER: return (long)propd(u, "frameSize", CHUNK);
  case K_WINDOWER: case K_IFFT: return (long)propd(u, "windowSize", CHUNK);
  case K_FFT: return 2 * (long)propd(u, "windowSize", CHUNK);
  case K_UNHOPPER: return (long)propd(u, "hopSize", CHUNK);
  case K_RECHUNK: return (long)propd(u, "outputInterval", CHUNK);
  case K_HARDLOWPASS: case K_HARDHIGHPASS: case K_SPECTRALGATE: case K_BINSHIFT: case K_AUGMENT: case K_SPECTRALSUM:
    return 2 * (long)propd(u, "windowSize", 4096);
  default: return CHUNK;
  }
}

Oracle *oracle_load(const char *blob, char *err, int errlen) {
  build_tables();
  Oracle *o = (Oracle *)calloc(1, sizeof(Oracle));
  int cap = 16;
  o->units = (Unit *)calloc(cap, sizeof(Unit));
  o->cbs = (CircleBuffer *)calloc(64, sizeof(CircleBuffer));
  o->root_unit = -1;
  const char *p = blob;
  Unit *u = NULL;
  int bad = 0;
  char line[8192];
  while (*p && !bad) {
    const char *e = strchr(p, '\n');
    size_t n = e ? (size_t)(e - p) : strlen(p);
    if (n >= sizeof line) { bad = 1; break; }
    memcpy(line, p, n); line[n] = 0;
    p = e ? e + 1 : p + n;
    char *save = NULL;
    char *tok = strtok_r(line, " \t\r", &save);
    if (!tok || tok[0] == '#') continue;
    if (!strcmp(tok, "DUSPGRAPH")) continue;
    if (!strcmp(tok, "END")) break;
    if (!strcmp(tok, "CB")) {
      CircleBuffer *b = &o->cbs[o->ncb++];
      b->nch = atoi(strtok_r(NULL, " ", &save));
      b->len = atol(strtok_r(NULL, " ", &save));
      b->ch = (float **)calloc(b->nch, sizeof(float *));
      for (int c = 0; c < b->nch; c++) b->ch[c] = (float *)calloc(b->len, 4);
    } else if (!strcmp(tok, "U")) {
      if (o->nunits == cap) { cap *= 2; o->units = (Unit *)realloc(o->units, cap * sizeof(Unit)); memset(o->units + o->nunits, 0, (cap - o->nunits) * sizeof(Unit)); }
      u = &o->units[o->nunits++];
      char *k = strtok_r(NULL, " ", &save), *l = strtok_r(NULL, " ", &save);
      u->kind = k ? kind_from_name(k) : -1;
      if (u->kind < 0) { snprintf(o->err, sizeof o->err, "unknown unit kind %s", k ? k : "?"); bad = 1; break; }
      snprintf(u->label, sizeof u->label, "%s", l ? l : "");
      u->process_index = -1;
    } else if (!u) { bad = 1;
    } else if (!strcmp(tok, "P")) {
      char *nm = strtok_r(NULL, " ", &save), *v = strtok_r(NULL, " ", &save);
      snprintf(u->pname[u->nprops], 24, "%s", nm); u->pval[u->nprops++] = strtod(v, NULL);
    } else if (!strcmp(tok, "S")) {
      char *nm = strtok_r(NULL, " ", &save), *v = strtok_r(NULL, " ", &save);
      snprintf(u->sname[u->nsprops], 24, "%s", nm); snprintf(u->sval[u->nsprops++], 32, "%s", v);
    } else if (!strcmp(tok, "I")) {
      Port *pt = &u->in[u->nin++];
      snprintf(pt->name, 32, "%s", strtok_r(NULL, " ", &save));
      pt->mono = atoi(strtok_r(NULL, " ", &save));
      char *mode = strtok_r(NULL, " ", &save);
      if (mode[0] == 'C') { /* Inlet.setConstant, Inlet.js:76-93: chunk channels pre-filled */
        int nc = atoi(strtok_r(NULL, " ", &save));
        pt->chunk = chunk_new(nc); pt->owns = 1;
        for (int c = 0; c < nc; c++) {
          float v = (float)strtod(strtok_r(NULL, " ", &save), NULL);
          for (int t = 0; t < CHUNK; t++) pt->chunk->ch[c][t] = v;
        }
      } else {
        pt->connected = 1; pt->nodata = (mode[0] == 'N');
        pt->src_unit = atoi(strtok_r(NULL, " ", &save));
        pt->src_outlet = atoi(strtok_r(NULL, " ", &save));
      }
    } else if (!strcmp(tok, "O")) {
      Port *pt = &u->out[u->nout++];
      snprintf(pt->name, 32, "%s", strtok_r(NULL, " ", &save));
      pt->mono = atoi(strtok_r(NULL, " ", &save));
      pt->chunk = chunk_new_sized(atoi(strtok_r(NULL, " ", &save)), outlet_chunk_size(u)); pt->owns = 1;
      char *cv; /* outlet.connections in connect order (Inlet.js:52) */
      while ((cv = strtok_r(NULL, " ", &save)) && pt->nconn < 64) pt->conn_unit[pt->nconn++] = atoi(cv);
    } else if (!strcmp(tok, "EV")) {
      o->events = (OEvent *)realloc(o->events, (o->nevents + 1) * sizeof(OEvent));
      OEvent *ev = &o->events[o->nevents++];
      ev->chunk = atol(strtok_r(NULL, " ", &save));
      ev->unit = atoi(strtok_r(NULL, " ", &save));
      snprintf(ev->prop, sizeof ev->prop, "%s", strtok_r(NULL, " ", &save));
      ev->value = strtod(strtok_r(NULL, " ", &save), NULL);
    } else if (!strcmp(tok, "KEEP")) { /* device hint: a graph edit follows (unit objects keep everything here anyway) */
    } else if (!strcmp(tok, "TG")) {
      o->triggers = (OEvent *)realloc(o->triggers, (o->ntriggers + 1) * sizeof(OEvent));
      OEvent *ev = &o->triggers[o->ntriggers++];
      ev->chunk = atol(strtok_r(NULL, " ", &save));
      ev->unit = atoi(strtok_r(NULL, " ", &save));
      snprintf(ev->prop, sizeof ev->prop, "%s", strtok_r(NULL, " ", &save));
      ev->value = strtod(strtok_r(NULL, " ", &save), NULL);
    } else if (!strcmp(tok, "ORDER")) {
      o->order = (int *)calloc(o->nunits + 1, sizeof(int));
      char *v;
      while ((v = strtok_r(NULL, " ", &save))) o->order[o->norder++] = atoi(v);
      o->has_order = 1;
    } else if (!strcmp(tok, "ROOT")) {
      o->root_unit = atoi(strtok_r(NULL, " ", &save));
      o->root_outlet = atoi(strtok_r(NULL, " ", &save));
    } else { snprintf(o->err, sizeof o->err, "bad blob line tag %s", tok); bad = 1; }
  }
  /* resolve connections: Inlet.connect (Inlet.js:40-56) aliases the outlet's chunk */
  for (int i = 0; i < o->nunits && !bad; i++) {
    Unit *w = &o->units[i];
    for (int j = 0; j < w->nin; j++) {
      Port *pt = &w->in[j];
      if (!pt->connected) continue;
      if (pt->src_unit < 0 || pt->src_unit >= o->nunits || pt->src_outlet >= o->units[pt->src_unit].nout) { bad = 1; break; }
      Port *src = &o->units[pt->src_unit].out[pt->src_outlet];
      pt->chunk = src->chunk; pt->owns = 0;
    }
  }
  for (int i = 0; i < o->nunits && !bad; i++) if (init_unit_state(o, &o->units[i])) bad = 1;
  if (!bad && (o->root_unit < 0 || o->root_unit >= o->nunits)) { fail(o, "renderAudioBuffer expects an outlet"); bad = 1; }
  /* "EPOCH <chunk>" + another graph: a run-time graph edit made by an event (`then`, an inlet assignment):
   * the host mirror's snapshot of the circuit after the edit; units are matched by label (adopt_state) */
  while (!bad && (*p == '\n' || *p == ' ' || *p == '\r')) p++;
  if (!bad && !strncmp(p, "EPOCH ", 6)) {
    char *rest = NULL;
    o->next_chunk = strtol(p + 6, &rest, 10);
    /* "EPOCH <chunk> FIN <unit>": the edit is made by unit <unit> of THIS section when it finishes by itself inside a
     * tick (Shape/index.js:40-42 -> UnitOrPatch.js:77-84); the chunk number is the host's prediction and is not used
     * here: the switch happens when this oracle's own Shape runs out (oracle_render) */
    o->next_fin_unit = -1;
    while (rest && *rest == ' ') rest++;
    if (rest && !strncmp(rest, "FIN ", 4)) o->next_fin_unit = atoi(rest + 4);
    const char *e = strchr(p, '\n');
    o->next = e ? oracle_load(e + 1, o->err, sizeof o->err) : NULL;
    if (!o->next) { if (!o->err[0]) fail(o, "EPOCH record without a graph"); bad = 1; }
  }
  if (bad) {
    if (err) snprintf(err, errlen, "%s", o->err[0] ? o->err : "malformed graph blob");
    oracle_free(o);
    return NULL;
  }
  return o;
}

void oracle_free(Oracle *o) {
  if (!o) return;
  oracle_free(o->next);
  free(o->events);
  free(o->triggers);
  for (int i = 0; i < o->nunits; i++) {
    Unit *u = &o->units[i];
    for (int j = 0; j < u->nin; j++) if (u->in[j].owns) chunk_free(u->in[j].chunk);
    for (int j = 0; j < u->nout; j++) if (u->out[j].owns) chunk_free(u->out[j].chunk);
    for (int c = 0; c < MAXCH; c++) free(u->bufs[c]);
    free(u->dbuf); free(u->env); fft_plan_free(u->fft);
  }
  for (int i = 0; i < o->ncb; i++) { for (int c = 0; c < o->cbs[i].nch; c++) free(o->cbs[i].ch[c]); free(o->cbs[i].ch); }
  free(o->cbs); free(o->units); free(o->order); free(o);
}

/* ------------------------------------------------------------- scheduler */
/* Unit.js:121-132 inputUnits: connected inlets in declaration order, dedup. */
static int input_units(Oracle *o, int ui, int *list) {
  Unit *u = &o->units[ui]; int n = 0;
  for (int j = 0; j < u->nin; j++) {
    if (!u->in[j].connected) continue;
    int s = u->in[j].src_unit, dup = 0;
    for (int k = 0; k < n; k++) if (list[k] == s) dup = 1;
    if (!dup) list[n++] = s;
  }
  return n;
}
/* Unit.js:133-143 outputUnits: per outlet, per connection, dedup. */
static int output_units(Oracle *o, int ui, int *list) {
  Unit *u = &o->units[ui]; int n = 0;
  for (int j = 0; j < u->nout; j++)
    for (int k = 0; k < u->out[j].nconn; k++) {
      int s = u->out[j].conn_unit[k], dup = 0;
      for (int m = 0; m < n; m++) if (list[m] == s) dup = 1;
      if (!dup) list[n++] = s;
    }
  return n;
}
static int in_hist(const int *h, int nh, int x) { for (int i = 0; i < nh; i++) if (h[i] == x) return 1; return 0; }

/* Unit.js:171-209 computeProcessIndex(history) */
static int compute_process_index(Oracle *o, int ui, const int *hist, int nh) {
  int *h = (int *)malloc((nh + 1) * sizeof(int));
  memcpy(h, hist, nh * sizeof(int)); h[nh] = ui;
  int *tmp = (int *)malloc(sizeof(int) * (o->nunits + 1));
  int n = input_units(o, ui, tmp), max = -1;
  for (int i = 0; i < n; i++) {
    if (in_hist(h, nh + 1, tmp[i])) continue;
    if (o->units[tmp[i]].process_index < 0) compute_process_index(o, tmp[i], h, nh + 1);
    if (o->units[tmp[i]].process_index > max) max = o->units[tmp[i]].process_index;
  }
  o->units[ui].process_index = max + 1;
  n = output_units(o, ui, tmp);
  for (int i = 0; i < n; i++) {
    if (in_hist(h, nh + 1, tmp[i])) continue;
    int pi = o->units[tmp[i]].process_index;
    if (pi < 0 || pi <= o->units[ui].process_index) compute_process_index(o, tmp[i], h, nh + 1);
  }
  free(tmp); free(h);
  return o->units[ui].process_index;
}
/* Circuit.js:125-131 computeOrders: `units.sort((a,b) => a.processIndex - b.processIndex)`.
 * Restated as V8's Array.prototype.sort (TimSort, third_party/v8/builtins/array-sort.tq;
 * Node >= 11) for arrays below the 64-element min-run threshold: one CountAndMakeRun
 * followed by a binary insertion sort.  The comparator returns NaN while a unit's
 * processIndex is still undefined; `order < 0` and `order >= 0` are then both false,
 * which this code reproduces.  With every key defined any stable sort agrees, so
 * larger circuits take the same path. */
static double cmp_pi(Oracle *o, int a, int b) {
  int pa = o->units[a].process_index, pb = o->units[b].process_index;
  if (pa < 0 || pb < 0) return NAN;
  return (double)(pa - pb);
}
static void compute_orders(Oracle *o) {
  int n = o->norder, *w = o->order;
  if (n < 2) return;
  /* CountAndMakeRun(0, n) */
  int run = 2;
  double order = cmp_pi(o, w[1], w[0]);
  int desc = order < 0;
  int prev = w[1];
  for (int idx = 2; idx < n; idx++) {
    order = cmp_pi(o, w[idx], prev);
    if (desc) { if (order >= 0) break; } else { if (order < 0) break; }
    prev = w[idx]; run++;
  }
  if (desc) for (int i = 0, j = run - 1; i < j; i++, j--) { int t = w[i]; w[i] = w[j]; w[j] = t; }
  /* BinaryInsertionSort(0, run, n) */
  for (int start = run; start < n; start++) {
    int left = 0, right = start, pivot = w[start];
    while (left < right) {
      int mid = left + ((right - left) >> 1);
      if (cmp_pi(o, pivot, w[mid]) < 0) right = mid; else left = mid + 1;
    }
    for (int p = start; p > left; p--) w[p] = w[p - 1];
    w[left] = pivot;
  }
}
/* Circuit.js:67-107 add(unit) */
static void circuit_add(Oracle *o, int ui) {
  if (o->units[ui].in_circuit) return;
  o->order[o->norder++] = ui;
  o->units[ui].in_circuit = 1;
  int *tmp = (int *)malloc(sizeof(int) * (o->nunits + 1));
  int n = input_units(o, ui, tmp);
  for (int i = 0; i < n; i++) circuit_add(o, tmp[i]);
  n = output_units(o, ui, tmp);
  for (int i = 0; i < n; i++) circuit_add(o, tmp[i]);
  free(tmp);
  compute_process_index(o, ui, NULL, 0);
  compute_orders(o);
}
/* `new Circuit(outlet.unit)` (renderChannelData.js:20).  Returns #units ticked. */
int oracle_schedule(Oracle *o, int *order_out) {
  free(o->order);
  o->order = (int *)calloc(o->nunits + 1, sizeof(int)); o->norder = 0;
  for (int i = 0; i < o->nunits; i++) { o->units[i].in_circuit = 0; o->units[i].process_index = -1; }
  circuit_add(o, o->root_unit);
  o->has_order = 1;
  if (order_out) memcpy(order_out, o->order, o->norder * sizeof(int));
  return o->norder;
}

/* ----------------------------------------------------------------- ticks */
#define INCH(u, i) ((u)->in[i].chunk)
#define OUTCH(u, i) ((u)->out[i].chunk)
static const float ZERO_CHUNK[CHUNK];
static double js_or0(double x) { return (x != x || x == 0) ? 0.0 : x; } /* `x || 0` */

/* Osc.js:35-47 */
static void tick_osc(Unit *u) {
  const float *f = INCH(u, 0)->ch[0]; float *out = OUTCH(u, 0)->ch[0];
  for (int t = 0; t < CHUNK; t++) {
    u->phase += (double)f[t];
    u->phase = fmod(u->phase, (double)SR);
    if (u->phase < 0) u->phase += SR;
    double fr = fmod(u->phase, 1.0);
    out[t] = (float)(ta_read(u->table, SR + 1, floor(u->phase)) * (1 - fr) + ta_read(u->table, SR + 1, ceil(u->phase)) * fr);
  }
}
/* MultiChannelOsc.js:21-38 (no negative-phase fix) */
static void tick_mcosc(Unit *u) {
  Chunk *fc = INCH(u, 0), *oc = OUTCH(u, 0);
  for (int c = 0; c < fc->nch; c++) {
    if (!u->mc_phase_set[c] || u->mc_phase[c] == 0 || u->mc_phase[c] != u->mc_phase[c]) { u->mc_phase[c] = 0; u->mc_phase_set[c] = 1; } /* `phase[c] || 0` */
    float *out = chunk_ensure(oc, c); const float *f = fc->ch[c];
    for (int t = 0; t < CHUNK; t++) {
      u->mc_phase[c] += (double)f[t];
      u->mc_phase[c] = fmod(u->mc_phase[c], (double)SR);
      double ph = u->mc_phase[c], fr = fmod(ph, 1.0);
      out[t] = (float)(ta_read(u->table, SR + 1, floor(ph)) * (1 - fr) + ta_read(u->table, SR + 1, ceil(ph)) * fr);
    }
  }
}
/* Ramp.js:25-40 */
static void tick_ramp(Unit *u) {
  float *out = OUTCH(u, 0)->ch[0];
  for (int i = 0; i < CHUNK; i++) {
    if (u->playing) {
      u->t++;
      if (u->t > u->duration) { u->playing = 0; u->t = u->duration; }
      if (u->t < 0) { u->playing = 0; u->t = 0; }
    }
    out[i] = (float)(u->y0 + (u->t / u->duration) * (u->y1 - u->y0));
  }
}
/* Shape/index.js:28-59 */
static void tick_shape(Unit *u) {
  const float *dur = INCH(u, 0)->ch[0], *mn = INCH(u, 1)->ch[0], *mx = INCH(u, 2)->ch[0];
  float *out = OUTCH(u, 0)->ch[0];
  for (int t = 0; t < CHUNK; t++) {
    if (u->playing) u->t += 1 / (double)dur[t];
    double lo = mn[t], hi = mx[t];
    if (u->t <= 0) {
      if (u->left_edge_is_shape) out[t] = (float)((double)u->table[0] * (hi - lo) + lo);
      else out[t] = (float)(u->left_edge * (hi - lo) + lo);
    } else if (u->t > SR) {
      if (!u->finished) u->finished = 1; /* finish(): UnitOrPatch.js:77-84 (onFinish graph edits out of scope) */
      if (u->right_edge_is_shape) out[t] = (float)((double)u->table[SR] * (hi - lo) + lo);
      else out[t] = (float)(u->right_edge * (hi - lo) + lo);
    } else {
      double fr = fmod(u->t, 1.0);
      out[t] = (float)(lo + (hi - lo) * (ta_read(u->table, SR + 1, ceil(u->t)) * fr + ta_read(u->table, SR + 1, floor(u->t)) * (1 - fr)));
    }
    /* NaN t (duration NaN): all three comparisons false -> third branch; table[NaN] undefined -> NaN */
  }
}
/* AHD.js:35-76 */
static void tick_ahd(Unit *u) {
  const float *a = INCH(u, 0)->ch[0], *h = INCH(u, 1)->ch[0], *d = INCH(u, 2)->ch[0];
  float *out = OUTCH(u, 0)->ch[0];
  const double sp = 1.0 / SR;
  for (int t = 0; t < CHUNK; t++) {
    switch (u->state) {
    case 1: out[t] = (float)u->t;
      if (u->playing) { u->t += sp / (double)a[t]; if (u->t >= 1) { u->state++; u->t--; } } break;
    case 2: out[t] = 1;
      if (u->playing) { u->t += sp / (double)h[t]; if (u->t >= 1) { u->state++; u->t--; } } break;
    case 3: out[t] = (float)(1 - u->t);
      if (u->playing) { u->t += sp / (double)d[t]; if (u->t >= 1) { u->state = 0; u->playing = 0; } } break;
    case 0: out[t] = 0; break;
    }
  }
}
/* SampleRateRedux.js:20-37 */
static void tick_sampleratereux(Unit *u) {
  Chunk *in = INCH(u, 0), *o = OUTCH(u, 0); const float *amt = INCH(u, 1)->ch[0];
  for (int c = 0; c < in->nch; c++) chunk_ensure(o, c);          /* while(out.length < in.length) push */
  for (int t = 0; t < CHUNK; t++) {
    u->srr_ts++;
    if (u->srr_ts > (double)amt[t]) {
      u->srr_nval = in->nch;
      for (int c = 0; c < in->nch; c++) u->srr_val[c] = (double)in->ch[c][t];
      u->srr_ts = 0;
    }
    for (int c = 0; c < u->srr_nval; c++) o->ch[c][t] = (float)u->srr_val[c];
  }
}
/* Timer.js:37-42 */
static void tick_timer(Unit *u) {
  float *out = OUTCH(u, 0)->ch[0]; const double sp = 1.0 / SR;
  for (int t = 0; t < CHUNK; t++) { u->t += sp; out[t] = (float)u->t; }
}
/* Noise.js:16-28 (Math.random -> injected oracle_rand) */
static void tick_noise(Unit *u) {
  Chunk *oc = OUTCH(u, 0); const float *f0 = INCH(u, 0)->ch[0];
  for (int c = 0; c < oc->nch; c++) {
    float *out = oc->ch[c];
    for (int t = 0; t < CHUNK; t++) {
      u->phase += (double)f0[t];
      if (u->phase >= SR) { u->phase = 0; u->y = 2 * oracle_rand(u->seed, u->draw++) - 1; }
      out[t] = (float)u->y;
    }
  }
}
/* Filter.js:66-84 */
static void filter_coeffs(Unit *u, double f) {
  if (u->fkind == 0) {
    double lamda = 1 / tan(M_PI * f / SR), l2 = lamda * lamda;
    u->a0 = 1 / (1 + 2 * lamda + l2); u->a1 = 2 * u->a0; u->a2 = u->a0;
    u->b1 = 2 * u->a0 * (1 - l2); u->b2 = u->a0 * (1 - 2 * lamda + l2);
  } else {
    double lamda = tan(M_PI * f / SR), l2 = lamda * lamda;
    u->a0 = 1 / (1 + 2 * lamda + l2); u->a1 = 0; u->a2 = -u->a0;
    u->b1 = 2 * u->a0 * (l2 - 1); u->b2 = u->a0 * (1 - 2 * lamda + l2);
  }
}
/* Filter.js:27-51 */
static void tick_filter(Unit *u) {
  Chunk *ic = INCH(u, 0), *oc = OUTCH(u, 0); const float *f = INCH(u, 1)->ch[0];
  int nc = ic->nch;
  for (int c = 0; c < nc; c++) chunk_ensure(oc, c);
  for (int t = 0; t < CHUNK; t++) {
    double ft = f[t];
    if (!u->lastf_defined || ft != u->lastf) { u->lastf_defined = 1; u->lastf = ft; filter_coeffs(u, ft); }
    for (int c = 0; c < nc; c++) {
      double x = ic->ch[c][t];
      double y = u->a0 * x + u->a1 * js_or0(u->x1[c]) + u->a2 * js_or0(u->x2[c])
               - u->b1 * js_or0(u->fy1[c]) - u->b2 * js_or0(u->fy2[c]);
      oc->ch[c][t] = (float)y;
      u->fy2[c] = js_or0(u->fy1[c]);
      u->fy1[c] = (double)oc->ch[c][t];
      u->x2[c] = js_or0(u->x1[c]);
      u->x1[c] = x;
    }
  }
}

#define MAXI(a, b) ((a) > (b) ? (a) : (b))
/* Multiply.js:23-34 / Divide.js:16-27 / Pow.js:19-30 (modulo channel broadcast) */
static void tick_binary_mod(Unit *u, int op) {
  Chunk *a = INCH(u, 0), *b = INCH(u, 1), *o = OUTCH(u, 0);
  int n = MAXI(a->nch, b->nch);
  for (int c = 0; c < n; c++) {
    const float *ac = a->ch[c % a->nch], *bc = b->ch[c % b->nch]; float *oc = chunk_ensure(o, c);
    for (int t = 0; t < CHUNK; t++) {
      double x = ac[t], y = bc[t];
      oc[t] = (float)(op == 0 ? x * y : op == 1 ? x / y : pow(x, y));
    }
  }
}
/* Sum.js:33-44 (a[c % a.length] is always defined, zeroChannel unreachable) */
static void tick_sum(Unit *u) {
  Chunk *a = INCH(u, 0), *b = INCH(u, 1), *o = OUTCH(u, 0);
  int n = MAXI(a->nch, b->nch);
  for (int c = 0; c < n; c++) {
    const float *ac = a->ch[c % a->nch], *bc = b->ch[c % b->nch]; float *oc = chunk_ensure(o, c);
    for (int t = 0; t < CHUNK; t++) oc[t] = (float)((double)ac[t] + (double)bc[t]);
  }
}
/* Subtract.js:19-29 (zero-fill, not modulo) */
static void tick_subtract(Unit *u) {
  Chunk *a = INCH(u, 0), *b = INCH(u, 1), *o = OUTCH(u, 0);
  int n = MAXI(a->nch, b->nch);
  for (int c = 0; c < n; c++) {
    float *oc = chunk_ensure(o, c);
    const float *ac = c < a->nch ? a->ch[c] : ZERO_CHUNK, *bc = c < b->nch ? b->ch[c] : ZERO_CHUNK;
    for (int t = 0; t < CHUNK; t++) oc[t] = (float)((double)ac[t] - (double)bc[t]);
  }
}
static double dB(double db) { return pow(10, db / 20); } /* Gain.js:26-28, Pan.js:26-28 */
/* Gain.js:17-24 */
static void tick_gain(Unit *u) {
  Chunk *in = INCH(u, 0), *o = OUTCH(u, 0); const float *g = INCH(u, 1)->ch[0];
  for (int c = 0; c < in->nch; c++) {
    float *oc = chunk_ensure(o, c);
    for (int t = 0; t < CHUNK; t++) oc[t] = (float)(dB(g[t]) * (double)in->ch[c][t]);
  }
}
/* PolarityInvert.js:13-20, Abs.js:17-24, Repeater.js:23-30, SecondsToSamples.js:13-20,
 * SemitoneToRatio.js:14-22, DecibelToScaler.js:13-19 */
static void tick_unary(Unit *u, int op) {
  Chunk *in = INCH(u, 0), *o = OUTCH(u, 0);
  for (int c = 0; c < in->nch; c++) {
    float *oc = chunk_ensure(o, c);
    for (int t = 0; t < CHUNK; t++) {
      double x = in->ch[c][t], y;
      switch (op) {
      case 0: y = -x; break;
      case 1: y = fabs(x); break;
      case 2: y = x; break;
      case 3: y = x * SR; break;
      case 4: y = pow(2, x / 12); break;
      default: y = pow(10, x / 20); break;
      }
      oc[t] = (float)y;
    }
  }
}
/* MidiToFrequency.js:14-21: `var fOut = this.frequency[c] || new Float32Array` is
 * never stored back, so only already-existing outlet channels receive data. */
static void tick_miditof(Unit *u) {
  Chunk *in = INCH(u, 0), *o = OUTCH(u, 0);
  for (int c = 0; c < in->nch && c < o->nch; c++)
    for (int t = 0; t < CHUNK; t++) o->ch[c][t] = (float)(pow(2, (((double)in->ch[c][t] - 69) / 12)) * 440);
}
/* Clip.js:12-21, HardClipAbove.js:14-24, HardClipBelow.js:14-24 */
static void tick_clip(Unit *u, int op) {
  Chunk *in = INCH(u, 0), *th = INCH(u, 1), *o = OUTCH(u, 0);
  for (int c = 0; c < in->nch; c++) {
    float *oc = chunk_ensure(o, c); const float *ic = in->ch[c], *tc = th->ch[c % th->nch];
    for (int t = 0; t < CHUNK; t++) {
      if (op == 0) oc[t] = fabs((double)ic[t]) > fabs((double)tc[t]) ? tc[t] : ic[t];
      else if (op == 1) oc[t] = ic[t] > tc[t] ? tc[t] : ic[t];
      else oc[t] = ic[t] < tc[t] ? tc[t] : ic[t];
    }
  }
}
/* Rescale.js:24-37 */
static void tick_rescale(Unit *u) {
  Chunk *in = INCH(u, 0), *il = INCH(u, 1), *iu = INCH(u, 2), *ol = INCH(u, 3), *ou = INCH(u, 4), *o = OUTCH(u, 0);
  for (int c = 0; c < in->nch; c++) {
    float *oc = chunk_ensure(o, c);
    const float *x = in->ch[c], *a = il->ch[c % il->nch], *b = iu->ch[c % iu->nch], *p = ol->ch[c % ol->nch], *q = ou->ch[c % ou->nch];
    for (int t = 0; t < CHUNK; t++)
      oc[t] = (float)(((double)x[t] - (double)a[t]) / ((double)b[t] - (double)a[t]) * ((double)q[t] - (double)p[t]) + (double)p[t]);
  }
}
/* CrossFader.js:20-29 (zero-fill) */
static void tick_crossfader(Unit *u) {
  Chunk *a = INCH(u, 0), *b = INCH(u, 1), *o = OUTCH(u, 0); const float *dial = INCH(u, 2)->ch[0];
  int n = MAXI(a->nch, b->nch);
  for (int c = 0; c < n; c++) {
    const float *ac = c < a->nch ? a->ch[c] : ZERO_CHUNK, *bc = c < b->nch ? b->ch[c] : ZERO_CHUNK;
    float *oc = chunk_ensure(o, c);
    for (int t = 0; t < CHUNK; t++) oc[t] = (float)((1 - (double)dial[t]) * (double)ac[t] + (double)dial[t] * (double)bc[t]);
  }
}
/* Pan.js:18-24 */
static void tick_pan(Unit *u) {
  const float *in = INCH(u, 0)->ch[0], *pan = INCH(u, 1)->ch[0];
  float *o0 = chunk_ensure(OUTCH(u, 0), 0), *o1 = chunk_ensure(OUTCH(u, 0), 1);
  for (int t = 0; t < CHUNK; t++) {
    double p = pan[t], x = in[t];
    double comp = dB((1 - fabs(p)) * u->comp_db);
    o0[t] = (float)(x * (1 - p) / 2 * comp);
    o1[t] = (float)(x * (1 + p) / 2 * comp);
  }
}
/* FixedMultiply.js:19-22 */
static void tick_fixedmultiply(Unit *u) {
  const float *in = INCH(u, 0)->ch[0]; float *o = OUTCH(u, 0)->ch[0];
  for (int t = 0; t < CHUNK; t++) o[t] = (float)((double)in[t] * u->sf);
}
/* ConcatChannels.js:16-30 */
static void tick_concat(Unit *u) {
  Chunk *a = INCH(u, 0), *b = INCH(u, 1), *o = OUTCH(u, 0);
  int na = a->nch, n = a->nch + b->nch;
  for (int c = 0; c < na; c++) memcpy(chunk_ensure(o, c), a->ch[c], CHUNK * 4);
  for (int c = na; c < n; c++) memcpy(chunk_ensure(o, c), b->ch[c - na], CHUNK * 4);
}
/* PickChannel.js:16-21 */
static void tick_pickchannel(Unit *u) {
  Chunk *in = INCH(u, 0); const float *cc = INCH(u, 1)->ch[0]; float *o = OUTCH(u, 0)->ch[0];
  for (int t = 0; t < CHUNK; t++) {
    double idx = fmod((double)cc[t], (double)in->nch);
    o[t] = (idx == floor(idx) && idx >= 0 && idx < in->nch) ? in->ch[(int)idx][t] : NAN;
  }
}
/* VectorMagnitude.js:16-28 */
static void tick_vectormagnitude(Unit *u) {
  Chunk *in = INCH(u, 0); float *o = OUTCH(u, 0)->ch[0];
  for (int t = 0; t < CHUNK; t++) {
    double s = 0;
    for (int c = 0; c < in->nch; c++) { double x = in->ch[c][t]; s += x * x; }
    o[t] = (float)sqrt(s);
  }
}
/* Delay.js:20-41 */
static void tick_delay(Unit *u, double clock) {
  Chunk *in = INCH(u, 0), *dl = INCH(u, 1), *o = OUTCH(u, 0);
  int nin0 = in->nch; /* loop bound re-evaluates this.in.length each iteration; :23 may grow it */
  for (int c = 0; c < in->nch || c < dl->nch; c++) {
    float *oc = chunk_ensure(o, c);
    if (c >= in->nch) { /* this.in[c] = this.in[c % this.in.length]  (aliases, :23) */
      if (c >= MAXCH) abort();
      in->ch[c] = in->ch[c % in->nch]; in->nch = c + 1;
    }
    if (!u->bufs[c]) u->bufs[c] = (float *)calloc(u->maxdelay, 4);
    float *buf = u->bufs[c]; long L = u->maxdelay;
    const float *dc = dl->ch[c % dl->nch], *ic = in->ch[c];
    for (int t = 0; t < CHUNK; t++) {
      double tb = fmod(clock + t, (double)L);
      oc[t] = (float)ta_read(buf, L, tb);
      ta_write(buf, L, tb, 0);
      double tw = fmod(tb + (double)dc[t], (double)L), fr = fmod(tw, 1.0);
      double fl = floor(tw), ce = ceil(tw);
      if (ta_valid(fl, L)) buf[(long)fl] = (float)((double)buf[(long)fl] + (double)ic[t] * (1 - fr));
      if (ta_valid(ce, L)) buf[(long)ce] = (float)((double)buf[(long)ce] + (double)ic[t] * fr);
    }
  }
  (void)nin0;
}
/* MonoDelay.js:19-30 */
static void tick_monodelay(Unit *u, double clock) {
  const float *in = INCH(u, 0)->ch[0], *dl = INCH(u, 1)->ch[0]; float *o = OUTCH(u, 0)->ch[0];
  float *buf = u->bufs[0]; long L = u->maxdelay;
  for (int t = 0; t < CHUNK; t++) {
    double tb = fmod(clock + t, (double)L);
    double tw = fmod(tb + (double)dl[t], (double)L), fr = fmod(tw, 1.0);
    double fl = floor(tw), ce = fmod(ceil(tw), (double)L);
    if (ta_valid(fl, L)) buf[(long)fl] = (float)((double)buf[(long)fl] + (double)in[t] * (1 - fr));
    if (ta_valid(ce, L)) buf[(long)ce] = (float)((double)buf[(long)ce] + (double)in[t] * fr);
    o[t] = (float)ta_read(buf, L, tb);
    ta_write(buf, L, tb, 0);
  }
}
/* ReadBackDelay.js:23-43 */
static int tick_readbackdelay(Oracle *o_, Unit *u) {
  Chunk *in = INCH(u, 0), *dl = INCH(u, 1), *o = OUTCH(u, 0);
  double t0 = u->tbuffer, t1 = t0 + CHUNK; long L = u->buflen;
  int n = MAXI(in->nch, dl->nch);
  for (int c = 0; c < n; c++) {
    const float *ic = in->ch[c % in->nch], *dc = dl->ch[c % dl->nch]; float *oc = chunk_ensure(o, c);
    if (!u->bufs[c]) u->bufs[c] = (float *)calloc(L, 4);
    float *buf = u->bufs[c];
    int i = 0;
    for (double t = t0; t < t1; t++) {
      if ((double)dc[i] > (double)L) return fail(o_, "delay may not exceed buffer length");
      ta_write(buf, L, fmod(t + L, (double)L), ic[i]);
      oc[i] = (float)ta_read(buf, L, fmod(t - (double)dc[i] + L, (double)L));
      i++;
    }
  }
  u->tbuffer = t1;
  return 0;
}
/* FixedDelay.js:14-20, CombFilter.js:11-17, AllPass.js:8-15 */
static void tick_fixeddelay(Unit *u) {
  const float *in = INCH(u, 0)->ch[0]; float *o = OUTCH(u, 0)->ch[0]; float *buf = u->bufs[0]; long L = u->buflen;
  const float *g = u->kind == K_FIXEDDELAY ? NULL : INCH(u, 1)->ch[0];
  for (int t = 0; t < CHUNK; t++) {
    u->tbuffer = fmod(u->tbuffer + 1, (double)L);
    long p = (long)u->tbuffer;
    if (u->kind == K_FIXEDDELAY) { o[t] = buf[p]; buf[p] = in[t]; }
    else if (u->kind == K_COMBFILTER) { o[t] = buf[p]; buf[p] = (float)((double)in[t] + (double)o[t] * (double)g[t]); }
    else { double d = buf[p]; buf[p] = (float)((double)in[t] + d * (double)g[t]); o[t] = (float)(d - (double)in[t] * (double)g[t]); }
  }
}
/* CircleBufferWriter.js:12-25 */
static void tick_cbwriter(Unit *u) {
  Chunk *off = INCH(u, 0), *in = INCH(u, 1);
  for (int c = 0; c < u->cb->nch; c++) {
    const float *oc = off->ch[c % off->nch];
    for (int t = 0; t < CHUNK; t++) {
      double tw = u->t + t + (double)SR * (double)oc[t];
      if (u->prewipe) cb_write(u->cb, c, tw, 0);
      if (c < in->nch) cb_mix(u->cb, c, tw, (double)in->ch[c][t]);
    }
  }
  u->t += CHUNK;
}
/* CircleBufferReader.js:12-25 */
static void tick_cbreader(Unit *u) {
  Chunk *off = INCH(u, 0), *o = OUTCH(u, 0);
  for (int c = 0; c < u->cb->nch; c++) {
    const float *oc = off->ch[c % off->nch]; float *out = chunk_ensure(o, c);
    for (int t = 0; t < CHUNK; t++) {
      double tr = u->t + t - (double)SR * (double)oc[t];
      out[t] = (float)cb_read(u->cb, c, tr);
      if (u->postwipe) cb_write(u->cb, c, tr, 0);
    }
  }
  u->t += CHUNK;
}

static int apply_event(Oracle *o, const OEvent *ev);
/* this._target.trigger(): the assignments recorded for unit `u` (TG records) */
static int fire_trigger(Oracle *o, Unit *u) {
  const long me = (long)(u - o->units);
  for (int i = 0; i < o->ntriggers; i++)
    if (o->triggers[i].chunk == me && apply_event(o, &o->triggers[i])) return -1;
  return 0;
}
/* Retriggerer.js:13-22 */
static int tick_retriggerer(Oracle *o, Unit *u) {
  const float *rate = INCH(u, 0)->ch[0];
  for (int t = 0; t < CHUNK; t++) {
    u->t += (double)rate[t];
    if (u->t >= SR) {
      if (fire_trigger(o, u)) return -1;
      u->t -= SR;
    }
  }
  return 0;
}
/* SporadicRetrigger.js:12-16 (tickInterval = 256) */
static int tick_sporadic_retriggerer(Oracle *o, Unit *u) {
  const float *rate = INCH(u, 0)->ch[0];
  if (oracle_rand(u->seed, u->draw++) < (double)rate[0] * CHUNK / SR) return fire_trigger(o, u);
  return 0;
}


/* ------------------------------------------------------- spectral units */
/* THIRD-PARTY ALGORITHM, NOT IN /root/reference: FFT.js:2,16 / IFFT.js:2,12 call `fft.js` (package.json pins
 * "^4.0.3"; node_modules is not vendored).  What follows restates that package's published radix-4 algorithm
 * (constructor tables, _transform4, _realTransform4, _singleTransform2/4, _singleRealTransform2/4,
 * completeSpectrum, inverseTransform) from its documented structure; it cannot be diffed against the package here.
 * It is pinned numerically instead: tests/test_oracle_kat.py checks it against numpy's DFT for every size
 * 4 .. 8192.  Storage rounding follows the call sites: FFT.js:21 transforms into the outlet's Float32Array
 * (every store rounds to binary32, `r32` below), IFFT.js:23 into an untyped Array (doubles). */
typedef struct FFTPlan { long size, csize; int width; double *table; long *bitrev; } FFTPlan;

static struct FFTPlan *fft_plan(long size) {
  FFTPlan *f = (FFTPlan *)calloc(1, sizeof(FFTPlan));
  f->size = size; f->csize = size << 1;
  f->table = (double *)malloc(sizeof(double) * size * 2);
  for (long i = 0; i < size * 2; i += 2) {
    const double angle = M_PI * (double)i / (double)size;
    f->table[i] = cos(angle); f->table[i + 1] = -sin(angle);
  }
  int power = 0;
  for (long t = 1; size > t; t <<= 1) power++;
  f->width = power % 2 == 0 ? power - 1 : power;
  long nb = 1L << f->width;
  f->bitrev = (long *)calloc(nb, sizeof(long));
  for (long j = 0; j < nb; j++)
    for (int shift = 0; shift < f->width; shift += 2) {
      int rev = f->width - shift - 2;
      /* (JS `<< -1` only ever meets a zero digit for the entries the transforms use: j < size / 4) */
      if (rev >= 0) f->bitrev[j] |= ((j >> shift) & 3) << rev;
    }
  return f;
}
static void fft_plan_free(struct FFTPlan *f) { if (f) { free(f->table); free(f->bitrev); free(f); } }

#define FST(i, v) out[i] = r32 ? (double)(float)(v) : (v)
/* the butterflies after the first pass, shared by the complex and the real transform's "full" part */
static void fft_transform4(const FFTPlan *f, double *out, const double *data, int inverse, int r32) {
  const long size = f->csize; long step = 1L << f->width; long len = (size / step) << 1;
  const double inv = inverse ? -1 : 1;
  long outOff, t;
  if (len == 4) {
    for (outOff = 0, t = 0; outOff < size; outOff += len, t++) { /* _singleTransform2 */
      const long off = f->bitrev[t];
      const double evenR = data[off], evenI = data[off + 1], oddR = data[off + step], oddI = data[off + step + 1];
      FST(outOff, evenR + oddR); FST(outOff + 1, evenI + oddI); FST(outOff + 2, evenR - oddR); FST(outOff + 3, evenI - oddI);
    }
  } else {
    for (outOff = 0, t = 0; outOff < size; outOff += len, t++) { /* _singleTransform4 */
      const long off = f->bitrev[t], step2 = step * 2, step3 = step * 3;
      const double Ar = data[off], Ai = data[off + 1], Br = data[off + step], Bi = data[off + step + 1];
      const double Cr = data[off + step2], Ci = data[off + step2 + 1], Dr = data[off + step3], Di = data[off + step3 + 1];
      const double T0r = Ar + Cr, T0i = Ai + Ci, T1r = Ar - Cr, T1i = Ai - Ci;
      const double T2r = Br + Dr, T2i = Bi + Di, T3r = inv * (Br - Dr), T3i = inv * (Bi - Di);
      FST(outOff, T0r + T2r); FST(outOff + 1, T0i + T2i); FST(outOff + 2, T1r + T3i); FST(outOff + 3, T1i - T3r);
      FST(outOff + 4, T0r - T2r); FST(outOff + 5, T0i - T2i); FST(outOff + 6, T1r - T3i); FST(outOff + 7, T1i + T3r);
    }
  }
  const double *table = f->table;
  for (step >>= 2; step >= 2; step >>= 2) {
    len = (size / step) << 1;
    const long quarterLen = len >> 2;
    for (outOff = 0; outOff < size; outOff += len) {
      const long limit = outOff + quarterLen;
      for (long i = outOff, k = 0; i < limit; i += 2, k += step) {
        const long A = i, B = A + quarterLen, C = B + quarterLen, D = C + quarterLen;
        const double Ar = out[A], Ai = out[A + 1], Br = out[B], Bi = out[B + 1], Cr = out[C], Ci = out[C + 1], Dr = out[D], Di = out[D + 1];
        const double MAr = Ar, MAi = Ai;
        const double tableBr = table[k], tableBi = inv * table[k + 1];
        const double MBr = Br * tableBr - Bi * tableBi, MBi = Br * tableBi + Bi * tableBr;
        const double tableCr = table[2 * k], tableCi = inv * table[2 * k + 1];
        const double MCr = Cr * tableCr - Ci * tableCi, MCi = Cr * tableCi + Ci * tableCr;
        const double tableDr = table[3 * k], tableDi = inv * table[3 * k + 1];
        const double MDr = Dr * tableDr - Di * tableDi, MDi = Dr * tableDi + Di * tableDr;
        const double T0r = MAr + MCr, T0i = MAi + MCi, T1r = MAr - MCr, T1i = MAi - MCi;
        const double T2r = MBr + MDr, T2i = MBi + MDi, T3r = inv * (MBr - MDr), T3i = inv * (MBi - MDi);
        FST(A, T0r + T2r); FST(A + 1, T0i + T2i); FST(B, T1r + T3i); FST(B + 1, T1i - T3r);
        FST(C, T0r - T2r); FST(C + 1, T0i - T2i); FST(D, T1r - T3i); FST(D + 1, T1i + T3r);
      }
    }
  }
}
/* realTransform: `data` holds `size` reals; only the first half (+1 bin) of `out` is meaningful afterwards */
static void fft_real_transform4(const FFTPlan *f, double *out, const double *data, int r32) {
  const long size = f->csize; long step = 1L << f->width; long len = (size / step) << 1;
  const double inv = 1;
  long outOff, t;
  if (len == 4) {
    for (outOff = 0, t = 0; outOff < size; outOff += len, t++) { /* _singleRealTransform2(outOff, off >>> 1, step >>> 1) */
      const long off = f->bitrev[t] >> 1, st = step >> 1;
      const double evenR = data[off], oddR = data[off + st];
      FST(outOff, evenR + oddR); FST(outOff + 1, 0.0); FST(outOff + 2, evenR - oddR); FST(outOff + 3, 0.0);
    }
  } else {
    for (outOff = 0, t = 0; outOff < size; outOff += len, t++) { /* _singleRealTransform4 */
      const long off = f->bitrev[t] >> 1, st = step >> 1;
      const double Ar = data[off], Br = data[off + st], Cr = data[off + 2 * st], Dr = data[off + 3 * st];
      const double T0r = Ar + Cr, T1r = Ar - Cr, T2r = Br + Dr, T3r = inv * (Br - Dr);
      FST(outOff, T0r + T2r); FST(outOff + 1, 0.0); FST(outOff + 2, T1r); FST(outOff + 3, -T3r);
      FST(outOff + 4, T0r - T2r); FST(outOff + 5, 0.0); FST(outOff + 6, T1r); FST(outOff + 7, T3r);
    }
  }
  const double *table = f->table;
  for (step >>= 2; step >= 2; step >>= 2) {
    len = (size / step) << 1;
    const long halfLen = len >> 1, quarterLen = halfLen >> 1, hquarterLen = quarterLen >> 1;
    for (outOff = 0; outOff < size; outOff += len) {
      for (long i = 0, k = 0; i <= hquarterLen; i += 2, k += step) {
        const long A = outOff + i, B = A + quarterLen, C = B + quarterLen, D = C + quarterLen;
        const double Ar = out[A], Ai = out[A + 1], Br = out[B], Bi = out[B + 1], Cr = out[C], Ci = out[C + 1], Dr = out[D], Di = out[D + 1];
        const double MAr = Ar, MAi = Ai;
        const double tableBr = table[k], tableBi = inv * table[k + 1];
        const double MBr = Br * tableBr - Bi * tableBi, MBi = Br * tableBi + Bi * tableBr;
        const double tableCr = table[2 * k], tableCi = inv * table[2 * k + 1];
        const double MCr = Cr * tableCr - Ci * tableCi, MCi = Cr * tableCi + Ci * tableCr;
        const double tableDr = table[3 * k], tableDi = inv * table[3 * k + 1];
        const double MDr = Dr * tableDr - Di * tableDi, MDi = Dr * tableDi + Di * tableDr;
        const double T0r = MAr + MCr, T0i = MAi + MCi, T1r = MAr - MCr, T1i = MAi - MCi;
        const double T2r = MBr + MDr, T2i = MBi + MDi, T3r = inv * (MBr - MDr), T3i = inv * (MBi - MDi);
        FST(A, T0r + T2r); FST(A + 1, T0i + T2i); FST(B, T1r + T3i); FST(B + 1, T1i - T3r);
        if (i == 0) { FST(C, T0r - T2r); FST(C + 1, T0i - T2i); continue; }   /* output final middle point */
        if (i == hquarterLen) continue;                                         /* do not overwrite ourselves */
        /* the mirrored butterfly, by symmetry of a real signal's spectrum */
        const double ST0r = T1r, ST0i = -T1i, ST1r = T0r, ST1i = -T0i;
        const double ST2r = -inv * T3i, ST2i = -inv * T3r, ST3r = -inv * T2i, ST3i = -inv * T2r;
        const long SA = outOff + quarterLen - i, SB = outOff + halfLen - i;
        FST(SA, ST0r + ST2r); FST(SA + 1, ST0i + ST2i); FST(SB, ST1r + ST3i); FST(SB + 1, ST1i - ST3r);
      }
    }
  }
}
static void fft_complete_spectrum(const FFTPlan *f, double *sp, int r32) {
  const long size = f->csize, half = size >> 1;
  double *out = sp;
  for (long i = 2; i < half; i += 2) { FST(size - i, sp[i]); FST(size - i + 1, -sp[i + 1]); }
}
#undef FST
/* test hooks (tests/test_oracle_kat.py): one transform on caller arrays */
int oracle_fft(long size, int mode, const double *in, double *out, int r32) {
  if (size < 2 || (size & (size - 1))) return -1;
  FFTPlan *f = fft_plan(size);
  if (mode == 0) { fft_real_transform4(f, out, in, r32); fft_complete_spectrum(f, out, r32); }
  else if (mode == 1) fft_transform4(f, out, in, 0, r32);
  else { fft_transform4(f, out, in, 1, r32); for (long i = 0; i < f->csize; i++) out[i] = r32 ? (double)(float)(out[i] / (double)size) : out[i] / (double)size; }
  fft_plan_free(f);
  return 0;
}

/* Hopper.js:19-37 (buffer: untyped Array; in / out channel 0: the ports' Float32Arrays) */
static void tick_hopper(Unit *u) {
  const float *in = INCH(u, 0)->ch[0]; const long inlen = INCH(u, 0)->size; float *out = OUTCH(u, 0)->ch[0];
  for (long t = 0; t < u->tick_interval; t++)
    u->dbuf[(long)fmod(u->t + t, (double)u->frame)] = (double)in[(long)fmod(u->t + t, (double)inlen)];
  u->t += (double)u->tick_interval;
  if (fmod(u->t, (double)u->hop) == 0)
    for (long t = 0; t < u->frame; t++) out[t] = (float)u->dbuf[(long)fmod((double)t + u->t, (double)u->frame)];
}
/* Windower.js:17-23; envelope :28-47 (Math.pow(Math.sin(pi n / (N-1)), 2) stored in a Float32Array) */
static void tick_windower(Unit *u) {
  const float *in = INCH(u, 0)->ch[0]; float *out = OUTCH(u, 0)->ch[0];
  for (long t = 0; t < u->win; t++) out[t] = (float)((double)in[t] * (double)u->env[t]);
}
/* FFT.js:19-25 */
static void tick_fft(Unit *u) {
  const float *in = INCH(u, 0)->ch[0]; float *out = OUTCH(u, 0)->ch[0];
  double *d = u->dbuf, *w = u->dbuf + u->win; /* scratch: input widened, output image of the Float32Array */
  for (long t = 0; t < u->win; t++) d[t] = (double)in[t];
  for (long t = 0; t < u->frame; t++) w[t] = (double)out[t];
  fft_real_transform4(u->fft, w, d, 1);
  fft_complete_spectrum(u->fft, w, 1);
  for (long t = 0; t < u->frame; t++) out[t] = (float)w[t];
}
/* IFFT.js:21-33 */
static void tick_ifft(Unit *u) {
  const float *in = INCH(u, 0)->ch[0]; float *out = OUTCH(u, 0)->ch[0];
  double *d = u->dbuf, *cplx = u->dbuf + u->frame; /* cplx = this.complexOut (persists, untyped Array) */
  for (long t = 0; t < u->frame; t++) d[t] = (double)in[t];
  fft_transform4(u->fft, cplx, d, 1, 0);
  for (long t = 0; t < u->frame; t++) cplx[t] /= (double)u->win; /* inverseTransform: out[i] /= this.size */
  for (long t = 0; t < u->win; t++) out[t] = (float)cplx[t * 2];
}
/* UnHopper.js:19-44 */
static void tick_unhopper(Unit *u) {
  const float *in = INCH(u, 0)->ch[0]; float *out = OUTCH(u, 0)->ch[0];
  for (long t = 0; t < u->win; t++) u->dbuf[(long)fmod((double)t + u->t, (double)u->win)] += (double)in[t];
  u->t += (double)u->hop;
  if (u->t > (double)u->hop) {
    const double t0 = u->t - (double)u->hop;
    for (long t = 0; t < u->hop; t++) {
      const long tb = (long)fmod(t0 + (double)t, (double)u->win);
      out[t] = (float)u->dbuf[tb];
      u->dbuf[tb] = 0;
    }
  }
}
/* ReChunk.js:26-49 */
static void tick_rechunk(Unit *u) {
  const float *in = INCH(u, 0)->ch[0]; float *out = OUTCH(u, 0)->ch[0];
  if (fmod(u->t, (double)u->in_iv) == 0)
    for (long t = 0; t < u->in_iv; t++) u->dbuf[(long)fmod(u->t + (double)t, (double)u->frame)] = (double)in[t];
  u->t += (double)u->tick_interval;
  if (fmod(u->t, (double)u->out_iv) == 0) {
    const double t0 = u->t - (double)u->out_iv;
    for (long t = 0; t < u->out_iv; t++) out[t] = (float)u->dbuf[(long)fmod(t0 + (double)t, (double)u->frame)];
  }
}
static double js_round(double x) { return floor(x + 0.5); } /* Math.round */
/* HardLowPass.js:20-31 / HardHighPass.js:20-31 */
static void tick_hardpass(Unit *u, int high) {
  const float *in = INCH(u, 0)->ch[0]; float *out = OUTCH(u, 0)->ch[0];
  const double f0 = (double)INCH(u, 1)->ch[0][0];
  const double cut = js_round(f0 / ((double)SR / (double)u->win)) * 2;
  if (cut != cut) return; /* NaN cutoff: neither loop runs, `out` keeps the previous frame */
  for (long i = 0; i < u->frame; i++) {
    const int below = (double)i < cut; /* NaN cutoff: no index is below it */
    out[i] = (below != high) ? in[i] : 0.0f;
  }
}
/* SpectralGate.js:15-32 */
static void tick_spectralgate(Unit *u) {
  const float *in = INCH(u, 0)->ch[0]; float *out = OUTCH(u, 0)->ch[0];
  const double th = (double)INCH(u, 1)->ch[0][0];
  for (long b = 0; b < u->frame; b += 2) {
    const double re = in[b], im = in[b + 1], mag = sqrt(re * re + im * im);
    if (u->invert ? mag < th : mag > th) { out[b] = (float)re; out[b + 1] = (float)im; } else { out[b] = 0; out[b + 1] = 0; }
  }
}
/* BinShift.js:14-26 */
static void tick_binshift(Unit *u) {
  const float *in = INCH(u, 0)->ch[0]; float *out = OUTCH(u, 0)->ch[0];
  const double shift = js_round((double)INCH(u, 1)->ch[0][0]) * 2;
  for (long i = 0; i < u->frame; i++) out[i] = 0;
  /* `bin + shift < frameSize` is false for a NaN shift: nothing moves */
  for (long bin = 1; bin < u->frame && (double)bin + shift < (double)u->frame; bin += 2) {
    if ((double)bin + shift < 0) continue;
    const long d = (long)((double)bin + shift);
    out[d] = in[bin]; out[d - 1] = in[bin - 1];
  }
}
/* Augment.js:14-27 (`out` is the port's Float32Array: every `+=` rounds to binary32) */
static void tick_augment(Unit *u) {
  const float *in = INCH(u, 0)->ch[0]; float *out = OUTCH(u, 0)->ch[0];
  for (long i = 0; i < u->frame; i++) out[i] = 0;
  for (long bin = 0; bin < u->win; bin++)
    for (int m = 0; m < u->nmap; m++) {
      const double b2 = js_round((double)bin * u->map_ratio[m]) * 2;
      if (!(b2 >= 0 && b2 < (double)u->frame)) continue; /* (a NaN index is a no-op store on a typed array) */
      const long d = (long)b2;
      out[d] = (float)((double)out[d] + (double)in[bin * 2] * u->map_gain[m]);
      out[d + 1] = (float)((double)out[d + 1] + (double)in[bin * 2 + 1] * u->map_gain[m]);
    }
}
/* SpectralSum.js:15-24 */
static void tick_spectralsum(Unit *u) {
  const float *a = INCH(u, 0)->ch[0], *b = INCH(u, 1)->ch[0]; float *out = OUTCH(u, 0)->ch[0];
  for (long i = 0; i < u->frame; i++) out[i] = (float)((double)a[i] + (double)b[i]);
}
/* Channels beyond the first of these units are untyped Arrays (`out[c] = out[c] || new Array(n)`: doubles, no
 * binary32 rounding) -- not modelled; and every frame inlet must see a chunk at least as long as it reads. */
static int spectral_check(Oracle *o, Unit *u) {
  for (int i = 0; i < u->nin; i++) if (!u->in[i].nodata && u->in[i].chunk->nch > 1 && i == 0) return fail(o, "multi-channel spectral chains are not modelled (extra channels are untyped Arrays)");
  long need = 0;
  switch (u->kind) {
  case K_WINDOWER: case K_FFT: case K_UNHOPPER: need = u->win; break;
  case K_RECHUNK: need = u->in_iv; break;
  case K_IFFT: case K_HARDLOWPASS: case K_HARDHIGHPASS: case K_SPECTRALGATE: case K_BINSHIFT: case K_AUGMENT: case K_SPECTRALSUM: need = u->frame; break;
  default: break;
  }
  if (need && INCH(u, 0)->size < need) return fail(o, "frame inlet fed by a shorter chunk (reads undefined)");
  if (u->kind == K_SPECTRALSUM && INCH(u, 1)->size < need) return fail(o, "frame inlet fed by a shorter chunk (reads undefined)");
  return 0;
}

/* Unit.js:111-119 tick -> _tick */
static int unit_tick(Oracle *o, Unit *u, double clock) {
  switch (u->kind) {
  case K_OSC: tick_osc(u); break;
  case K_MCOSC: tick_mcosc(u); break;
  case K_RAMP: tick_ramp(u); break;
  case K_SHAPE: tick_shape(u); break;
  case K_AHD: tick_ahd(u); break;
  case K_TIMER: tick_timer(u); break;
  case K_SAMPLERATEREDUX: tick_sampleratereux(u); break;
  case K_MONITOR: break; /* Monitor.js:14-16: console.log only */
  case K_RETRIGGERER: return tick_retriggerer(o, u);
  case K_SPORADICRETRIGGERER: return tick_sporadic_retriggerer(o, u);
  case K_NOISE: tick_noise(u); break;
  case K_FILTER: tick_filter(u); break;
  case K_MULTIPLY: tick_binary_mod(u, 0); break;
  case K_DIVIDE: tick_binary_mod(u, 1); break;
  case K_POW: tick_binary_mod(u, 2); break;
  case K_SUM: tick_sum(u); break;
  case K_SUBTRACT: tick_subtract(u); break;
  case K_GAIN: tick_gain(u); break;
  case K_POLARITYINVERT: tick_unary(u, 0); break;
  case K_ABS: tick_unary(u, 1); break;
  case K_REPEATER: tick_unary(u, 2); break;
  case K_SECONDSTOSAMPLES: tick_unary(u, 3); break;
  case K_SEMITONETORATIO: tick_unary(u, 4); break;
  case K_DECIBELTOSCALER: tick_unary(u, 5); break;
  case K_MIDITOFREQUENCY: tick_miditof(u); break;
  case K_CLIP: tick_clip(u, 0); break;
  case K_HARDCLIPABOVE: tick_clip(u, 1); break;
  case K_HARDCLIPBELOW: tick_clip(u, 2); break;
  case K_RESCALE: tick_rescale(u); break;
  case K_CROSSFADER: tick_crossfader(u); break;
  case K_PAN: tick_pan(u); break;
  case K_FIXEDMULTIPLY: tick_fixedmultiply(u); break;
  case K_CONCATCHANNELS: tick_concat(u); break;
  case K_PICKCHANNEL: tick_pickchannel(u); break;
  case K_VECTORMAGNITUDE: tick_vectormagnitude(u); break;
  case K_DELAY: tick_delay(u, clock); break;
  case K_MONODELAY: tick_monodelay(u, clock); break;
  case K_READBACKDELAY: return tick_readbackdelay(o, u);
  case K_FIXEDDELAY: case K_COMBFILTER: case K_ALLPASS: tick_fixeddelay(u); break;
  case K_CBWRITER: tick_cbwriter(u); break;
  case K_CBREADER: tick_cbreader(u); break;
  case K_HOPPER: if (spectral_check(o, u)) return -1; tick_hopper(u); break;
  case K_WINDOWER: if (spectral_check(o, u)) return -1; tick_windower(u); break;
  case K_FFT: if (spectral_check(o, u)) return -1; tick_fft(u); break;
  case K_IFFT: if (spectral_check(o, u)) return -1; tick_ifft(u); break;
  case K_UNHOPPER: if (spectral_check(o, u)) return -1; tick_unhopper(u); break;
  case K_RECHUNK: if (spectral_check(o, u)) return -1; tick_rechunk(u); break;
  case K_HARDLOWPASS: if (spectral_check(o, u)) return -1; tick_hardpass(u, 0); break;
  case K_HARDHIGHPASS: if (spectral_check(o, u)) return -1; tick_hardpass(u, 1); break;
  case K_SPECTRALGATE: if (spectral_check(o, u)) return -1; tick_spectralgate(u); break;
  case K_BINSHIFT: if (spectral_check(o, u)) return -1; tick_binshift(u); break;
  case K_AUGMENT: if (spectral_check(o, u)) return -1; tick_augment(u); break;
  case K_SPECTRALSUM: if (spectral_check(o, u)) return -1; tick_spectralsum(u); break;
  }
  return 0;
}

/* ------------------------------------------------------------------ sink */
/* the assignments an event function makes (Ramp.js:19-23 trigger, Shape/index.js:92-99, AHD.js:23-33, Timer.js) */
static int apply_event(Oracle *o, const OEvent *ev) {
  if (ev->unit < 0 || ev->unit >= o->nunits) return fail(o, "event names a missing unit");
  Unit *u = &o->units[ev->unit];
  const char *p = ev->prop; double v = ev->value;
  if (!strcmp(p, "t")) u->t = v;
  else if (!strcmp(p, "playing")) u->playing = v != 0;
  else if (!strcmp(p, "state")) u->state = (int)v;
  else if (!strcmp(p, "phase")) u->phase = v;
  else if (!strcmp(p, "phase[*]")) { for (int c = 0; c < MAXCH; c++) if (u->mc_phase_set[c]) u->mc_phase[c] = v; } /* MultiChannelOsc.js:52-55 */
  else if (!strcmp(p, "y0")) u->y0 = v;
  else if (!strcmp(p, "y1")) u->y1 = v;
  else if (!strcmp(p, "duration")) u->duration = v;
  else if (!strcmp(p, "y")) u->y = v;
  else if (!strcmp(p, "sf")) u->sf = v;
  else if (!strcmp(p, "finished")) u->finished = v != 0;
  else return fail(o, "event assigns a property the oracle does not model");
  return 0;
}

static int chain_fail(Oracle *first, Oracle *o) { if (o != first) snprintf(first->err, sizeof first->err, "%s", o->err); return -1; }

/* renderChannelData.js:5-49.  out[c] is malloc'd here (caller frees with
 * oracle_free_output); returns 0 or -1 (message via oracle_error). */
static int g_raw_copy = 0;
/* 1: copy the root outlet's chunk as it is (what RenderStream reads, RenderStream.js:38), no `x || 0` */
void oracle_set_raw(int raw) { g_raw_copy = raw; }

/* A graph edit leaves every unit object as it is (Inlet.connect only re-aliases chunks and adds units,
 * Inlet.js:40-74, Circuit.js:67-107): units of the new snapshot that existed before keep their run-time state,
 * their outlets' chunk contents (a feedback edge reads them one chunk late) and their delay lines. */
static void adopt_state(Oracle *n, Oracle *o) {
  n->clock = o->clock;
  for (int j = 0; j < n->nunits; j++) {
    Unit *a = &n->units[j], *b = NULL;
    for (int i = 0; i < o->nunits && !b; i++)
      if (o->units[i].kind == a->kind && !strcmp(o->units[i].label, a->label)) b = &o->units[i];
    if (!b) continue;
    a->phase = b->phase; a->t = b->t; a->y = b->y; a->y0 = b->y0; a->y1 = b->y1; a->duration = b->duration;
    a->left_edge = b->left_edge; a->right_edge = b->right_edge; a->sf = b->sf; a->comp_db = b->comp_db;
    a->playing = b->playing; a->state = b->state; a->finished = b->finished; a->seed = b->seed; a->draw = b->draw;
    memcpy(a->mc_phase, b->mc_phase, sizeof a->mc_phase); memcpy(a->mc_phase_set, b->mc_phase_set, sizeof a->mc_phase_set);
    a->lastf_defined = b->lastf_defined; a->lastf = b->lastf;
    a->a0 = b->a0; a->a1 = b->a1; a->a2 = b->a2; a->b1 = b->b1; a->b2 = b->b2;
    memcpy(a->x1, b->x1, sizeof a->x1); memcpy(a->x2, b->x2, sizeof a->x2);
    memcpy(a->fy1, b->fy1, sizeof a->fy1); memcpy(a->fy2, b->fy2, sizeof a->fy2); memcpy(a->sdef, b->sdef, sizeof a->sdef);
    a->tbuffer = b->tbuffer;
    memcpy(a->srr_val, b->srr_val, sizeof a->srr_val); a->srr_nval = b->srr_nval; a->srr_ts = b->srr_ts;
    if (a->maxdelay == b->maxdelay && a->buflen == b->buflen)
      for (int c = 0; c < MAXCH; c++) if (b->bufs[c]) { free(a->bufs[c]); a->bufs[c] = b->bufs[c]; b->bufs[c] = NULL; }
    if (a->cb && b->cb && a->cb->nch == b->cb->nch && a->cb->len == b->cb->len)
      for (int c = 0; c < a->cb->nch; c++) memcpy(a->cb->ch[c], b->cb->ch[c], (size_t)a->cb->len * 4);
    for (int k = 0; k < a->nout; k++)
      for (int m = 0; m < b->nout; m++) {
        if (strcmp(a->out[k].name, b->out[m].name) || !a->out[k].chunk || !b->out[m].chunk) continue;
        for (int c = 0; c < b->out[m].chunk->nch; c++)
          if (a->out[k].chunk->size == b->out[m].chunk->size)
            memcpy(chunk_ensure(a->out[k].chunk, c), b->out[m].chunk->ch[c], (size_t)b->out[m].chunk->size * sizeof(float));
      }
  }
}

int oracle_render(Oracle *first, double duration, float **out, int max_ch, int *nch_out, long *nsamples_out) {
  Oracle *o = first;
  if (!o->has_order) oracle_schedule(o, NULL);
  double length = duration * SR;                  /* :24 */
  long n = length > 0 ? (long)length : 0;         /* new Float32Array(x): ToIndex truncates */
  Chunk *root = o->units[o->root_unit].out[o->root_outlet].chunk;
  int nch = 0;
  /* Circuit.js:125-143: the circuit advances by the gcd of its units' tick intervals; only multiples of the
   * standard chunk are modelled (every frame-rate unit of the reference ticks at a multiple of 256 by default) */
  for (Oracle *q = first; q; q = q->next)
    for (int i = 0; i < q->nunits; i++)
      if (q->units[i].tick_interval % CHUNK) return fail(first, "tick interval that is not a multiple of 256 samples");
  long csize = root->size;                        /* :25 outlet.chunkSize */
  for (double t0 = 0; t0 < length; t0 += (double)csize) { /* :29 */
    /* circuit.tickUntil(t1): Circuit.js:19-47 -- one tick per 256 samples; a unit ticks when clock % tickInterval == 0 */
    while (o->clock < t0 + (double)csize) {
      while (o->next && o->next_fin_unit < 0 && (long)(o->clock / CHUNK) >= o->next_chunk) { /* the event that edits the graph runs first */
        adopt_state(o->next, o);
        o = o->next;
        if (!o->has_order) oracle_schedule(o, NULL);
        root = o->units[o->root_unit].out[o->root_outlet].chunk;
      }
      for (int e = 0; e < o->nevents; e++)          /* Circuit.js:23 runEvents(clock + 256), in queue order */
        if (o->events[e].chunk == (long)(o->clock / CHUNK) && apply_event(o, &o->events[e])) return chain_fail(first, o);
      for (int i = 0; i < o->norder; i++) {
        Unit *w = &o->units[o->order[i]];
        if (fmod(o->clock, (double)w->tick_interval) != 0) continue; /* Circuit.js:35 */
        const int was_finished = w->finished;
        if (unit_tick(o, w, o->clock)) return chain_fail(first, o);
        if (o->next && o->next_fin_unit == o->order[i] && w->finished && !was_finished) {
          /* finish() from inside this tick ran onFinish, which edited the graph: Inlet.connect re-aliased chunks and
           * Circuit.add re-sorted `this.units` (Circuit.js:67-107) while Circuit.tick's `for i` loop is running
           * (Circuit.js:34-37).  The loop simply goes on with index i + 1 over the NEW list. */
          Oracle *nx = o->next;
          adopt_state(nx, o);
          if (!nx->has_order) oracle_schedule(nx, NULL);
          for (int e = 0; e < nx->nevents; e++)    /* state onFinish assigned, at this moment */
            if (nx->events[e].chunk == (long)(nx->clock / CHUNK) && apply_event(nx, &nx->events[e])) return chain_fail(first, nx);
          o = nx;
          root = o->units[o->root_unit].out[o->root_outlet].chunk;
        }
      }
      o->clock += CHUNK;
    }
    while (root->nch > nch) {                     /* :38-39 */
      if (nch >= max_ch) return fail(first, "too many output channels");
      out[nch++] = (float *)calloc(n > 0 ? n : 1, sizeof(float));
    }
    for (int c = 0; c < root->nch; c++)           /* :42-44: `x || 0`, out-of-range stores dropped */
      for (long t = 0; t < csize; t++) {
        long i = (long)t0 + t;
        if (i < n) { float x = root->ch[c][t]; out[c][i] = g_raw_copy ? x : (x != x || x == 0) ? 0.f : x; }
      }
  }
  *nch_out = nch; *nsamples_out = n;
  return 0;
}
void oracle_free_output(float **out, int nch) { for (int c = 0; c < nch; c++) free(out[c]); }
const char *oracle_error(Oracle *o) { return o->err; }
int oracle_num_units(Oracle *o) { return o->nunits; }
