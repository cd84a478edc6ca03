// compile.cc -- the Circuit scheduler's tick list -> flat op program.
//
// What the reference does per 256-sample chunk (Circuit.js:34-37): call every
// unit's _tick in `circuit.units` order, each reading its inlets' SignalChunks
// and writing its outlets'.  Here that list is flattened once:
//   * every (outlet, channel) becomes a VALUE; a connected inlet is an alias of
//     the producer's value (Inlet.js:53), an unconnected inlet is a per-voice
//     constant word (Inlet.js:76-93);
//   * multi-channel units are expanded to one instruction per channel with the
//     reference's channel-count growth rules evaluated statically;
//   * a value consumed by a unit that ticks BEFORE its producer (cyclic graphs,
//     Unit.js:171-181 breaks the cycle there) is read one chunk stale through a
//     256-sample per-voice buffer -- exactly what the aliased SignalChunk holds
//     in the reference at that moment;
//   * values get shared-memory slots by linear-scan liveness.
// The device interprets the program sub-block by sub-block (B samples at a time)
// instead of chunk by chunk; that reordering is exact for everything except
// CircleBuffer taps that touch each other's samples inside one chunk, which are
// rejected below (DUSP_EUNSUPPORTED) rather than approximated.
#include "compile.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <set>
#include <tuple>

namespace dusp {

namespace {

const char *OP_NAMES[OP_COUNT] = {
    "nop", "copy", "neg", "abs", "s2s", "semi", "db2s", "midi2f", "fixmul", "mul", "div", "pow", "add", "sub",
    "gain", "clip", "hca", "hcb", "rescale", "xfade", "pan", "vecmag", "osc", "mcosc", "ramp", "shape", "ahd",
    "timer", "noise", "filter", "fixdelay", "comb", "allpass", "delay", "monodelay", "readback", "cbwrite",
    "cbread", "ldstale", "ststale", "out", "filterc", "panc", "gainc", "noise1", "cbwritec", "cbreadc", "add0",
    "retrig", "sretrig", "pokeif", "srredux", "hopper", "window", "fft", "ifft", "unhop", "rechunk", "specpass",
    "specgate", "binshift", "augment", "specsum", "ldframe", "stframe", "outring", "pick"};

struct Opnd {
  int kind = 0;  // 0 none (reads 0), 1 value, 2 const word, 3 immediate (frame ops: ring ids / sizes travel in `in`)
  int id = -1;
};

struct IR {
  int op = OP_NOP;
  Opnd in[5];
  int outv[2] = {-1, -1};
  int p = -1;
  int aux[3] = {0, 0, 0};
  int unit = -1, chan = 0;
  bool root = false;  // has effects beyond its output values
  int pos = 0;        // index of the owning unit in the tick order
};

struct Val {
  int unit, outlet, chan;
  int def = -1;  // producing IR index
};

struct Builder {
  const Graph &g;
  Program prog;
  std::vector<IR> ir;
  std::vector<Val> vals;
  std::map<std::tuple<int, int, int>, int> valmap;
  std::vector<std::vector<int>> nch;  // steady-state channel count per [unit][outlet]
  std::vector<int> pos_of_unit;

  explicit Builder(const Graph &gr) : g(gr) {}

  int val(int u, int o, int c) {
    auto key = std::make_tuple(u, o, c);
    auto it = valmap.find(key);
    if (it != valmap.end()) return it->second;
    vals.push_back({u, o, c, -1});
    return valmap[key] = (int)vals.size() - 1;
  }
  int word(const std::string &name, double dflt, bool f32) {
    prog.words.push_back({name, dflt, f32});
    return (int)prog.words.size() - 1;
  }
  int ring(long long len, const std::string &name) {
    if (len < 1) throw Unsupported("delay line of length < 1 sample");
    prog.ring_names.push_back(name);
    prog.rings.push_back({prog.ring_floats, len});
    prog.ring_floats += (len + 3) & ~3LL;  // keep every ring 16-byte aligned
    return (int)prog.rings.size() - 1;
  }
  static std::string uname(int u) { return "u" + std::to_string(u); }

  int in_nch(const std::vector<std::vector<int>> &n, int u, int i) const {
    const GPort &p = g.units[u].in[i];
    if (p.connected) return n[p.src_unit][p.src_outlet];
    return (int)p.consts.size();
  }
  // operand for channel `c` of inlet `i` (caller has already applied the unit's
  // own modulo / zero-fill rule, so c < channel count or the result is "none")
  Opnd in_ref(int u, int i, int c) {
    const GPort &p = g.units[u].in[i];
    Opnd o;
    if (p.connected) {
      if (p.src_unit == u) throw Unsupported("unit " + g.units[u].label + " feeds its own inlet (sample-level self feedback)");
      if (c >= nch[p.src_unit][p.src_outlet]) return o;
      o.kind = 1;
      o.id = val(p.src_unit, p.src_outlet, c);
    } else {
      if (c >= (int)p.consts.size()) return o;
      o.kind = 2;
      o.id = word(uname(u) + "." + p.name + "[" + std::to_string(c) + "]", p.consts[c], true);
    }
    return o;
  }
  IR &emit(int op, int u, int c) {
    ir.emplace_back();
    IR &r = ir.back();
    r.op = op; r.unit = u; r.chan = c; r.pos = pos_of_unit[u];
    return r;
  }
  void def(IR &r, int k, int u, int o, int c) {
    int v = val(u, o, c);
    r.outv[k] = v;
    vals[v].def = (int)(&r - &ir[0]);
  }
};

int find_inlet(const GUnit &u, const char *name) {
  for (size_t i = 0; i < u.in.size(); i++) if (u.in[i].name == name) return (int)i;
  throw std::runtime_error("unit " + u.kind + " lacks inlet " + name);
}

int table_id(const std::string &s, bool shape) {
  if (!shape) {
    if (s == "sin" || s == "sine") return TAB_SIN;
    if (s == "saw") return TAB_SAW;
    if (s == "square") return TAB_SQUARE;
    if (s == "triangle") return TAB_TRIANGLE;
    if (s == "8bit") return TAB_8BIT;
    throw std::runtime_error("waveform doesn't exist: " + s);  // Osc.js:60-61
  }
  if (s == "decay") return TAB_DECAY;
  if (s == "attack") return TAB_ATTACK;
  if (s == "semiSine") return TAB_SEMISINE;
  if (s == "decaySquared") return TAB_DECAYSQ;
  throw std::runtime_error("invalid shape function: " + s);
}

// ---- frame-rate units (components/spectral) ----
bool is_spectral_unit(const std::string &k) {
  return k == "HardLowPass" || k == "HardHighPass" || k == "SpectralGate" || k == "BinShift" || k == "Augment" || k == "SpectralSum";
}
bool is_frame_kind(const std::string &k) {
  return k == "Hopper" || k == "Windower" || k == "FFT" || k == "IFFT" || k == "UnHopper" || k == "ReChunk" || is_spectral_unit(k);
}
// chunkSize of a frame-family unit's outlet (the option its constructor passes to addOutlet)
long long frame_out_size(const GUnit &u) {
  const std::string &k = u.kind;
  if (k == "Hopper") return (long long)u.num("frameSize", 0);
  if (k == "Windower" || k == "IFFT") return (long long)u.num("windowSize", 0);
  if (k == "FFT") return 2 * (long long)u.num("windowSize", 0);
  if (k == "UnHopper") return (long long)u.num("hopSize", 0);
  if (k == "ReChunk") return (long long)u.num("outputInterval", 0);
  return 2 * (long long)u.num("windowSize", 4096);
}
// does inlet `i` of a frame-family unit read a whole frame (as opposed to a standard 256-sample chunk)?
bool is_frame_inlet(const GUnit &u, int i) {
  if (!is_frame_kind(u.kind) || u.kind == "Hopper") return false;
  if (u.kind == "SpectralSum") return i < 2;
  return i == 0;
}

// How many channels of outlet 0 a unit's _tick writes / grows to, given the
// channel counts it sees (the `out[c] = out[c] || new Float32Array` lines).
int written_channels(const Graph &g, const std::vector<std::vector<int>> &n, int ui, const Builder &b) {
  const GUnit &u = g.units[ui];
  const std::string &k = u.kind;
  auto in = [&](int i) { return b.in_nch(n, ui, i); };
  if (k == "MultiChannelOsc") return in(0);
  if (k == "Filter" || k == "Gain" || k == "PolarityInvert" || k == "Abs" || k == "Repeater" ||
      k == "SecondsToSamples" || k == "SemitoneToRatio" || k == "DecibelToScaler" || k == "Clip" ||
      k == "HardClipAbove" || k == "HardClipBelow" || k == "Rescale" || k == "SampleRateRedux")
    return in(0);
  if (k == "Multiply" || k == "Divide" || k == "Pow" || k == "Sum" || k == "Subtract" || k == "CrossFader" ||
      k == "Delay" || k == "ReadBackDelay")
    return std::max(in(0), in(1));
  if (k == "ConcatChannels") return in(0) + in(1);
  if (k == "Pan") return 2;
  if (k == "CircleBufferReader") return g.cbs.at((size_t)u.num("buffer", -1)).nch;
  return 0;  // mono-outlet units and MidiToFrequency never grow their outlet
}

}  // namespace

const char *op_name(int op) { return op >= 0 && op < OP_COUNT ? OP_NAMES[op] : "?"; }

Program compile(const Graph &g) {
  Builder b(g);
  const int NU = (int)g.units.size();
  b.pos_of_unit.assign(NU, -1);
  for (size_t i = 0; i < g.order.size(); i++) {
    if (b.pos_of_unit[g.order[i]] >= 0) throw std::runtime_error("graph blob: unit listed twice in ORDER");
    b.pos_of_unit[g.order[i]] = (int)i;
  }
  for (int u = 0; u < NU; u++)
    if (b.pos_of_unit[u] < 0) b.pos_of_unit[u] = 1 << 28;  // not ticked: never produces
  if (b.pos_of_unit[g.root_unit] >= (1 << 28)) throw std::runtime_error("rendered unit is not in the circuit");
  b.prog.n_units = (int)g.order.size();
  for (auto &u : g.units) b.prog.unit_labels.push_back(u.kind + ":" + u.label);

  // ---- channel counts: steady state, then check tick 0 already sees it ----
  std::vector<std::vector<int>> init(NU);
  for (int u = 0; u < NU; u++) for (auto &o : g.units[u].out) init[u].push_back(std::max(1, o.nch));
  auto simulate = [&](std::vector<std::vector<int>> n, std::vector<std::vector<int>> *seen) {
    for (int ui : g.order) {
      const GUnit &u = g.units[ui];
      if (seen) { (*seen)[ui].clear(); for (size_t i = 0; i < u.in.size(); i++) (*seen)[ui].push_back(b.in_nch(n, ui, (int)i)); }
      if (!u.out.empty()) n[ui][0] = std::max(n[ui][0], written_channels(g, n, ui, b));
      if (u.kind == "Delay" && u.in[0].connected && b.in_nch(n, ui, 0) < b.in_nch(n, ui, 1))
        throw Unsupported("Delay with more DELAY channels than IN channels aliases channels into its producer (Delay.js:23)");
    }
    return n;
  };
  std::vector<std::vector<int>> seen0(NU), seenS(NU);
  auto n1 = simulate(init, &seen0);
  auto ns = n1;
  for (int it = 0; it < 16; it++) { auto nn = simulate(ns, nullptr); if (nn == ns) break; ns = nn; }
  simulate(ns, &seenS);
  // Warm-up ticks: a unit that the reference's order ticks BEFORE its producer sees the producer's outlet with the
  // channel count of the previous tick, so a multi-channel signal takes a few ticks to spread (`out[c] = out[c] ||
  // new Float32Array`).  Tick k is compiled from the counts its units see, n_k = simulate^k(init); from tick S on
  // the program is the steady one.  The host renders chunk k < S with the program of a `WARMUP k` record and hands
  // the state over like a graph edit (render.SessionRun).
  bool dynamic_nch = false;
  {
    auto nk = init;
    int S = 0;
    for (; S < 64; S++) {
      std::vector<std::vector<int>> seenk(NU);
      auto nn = simulate(nk, &seenk);
      if (seenk == seenS && nn[g.root_unit][g.root_outlet] == ns[g.root_unit][g.root_outlet]) break;
      if (S == g.warmup_tick) { b.nch = nk; dynamic_nch = true; }
      nk = nn;
    }
    if (S >= 64) throw Unsupported("channel counts do not settle within 64 ticks");
    b.prog.warmup_ticks = S;
    if (S > 0 && g.warmup_tick < 0)
      throw Unsupported("channel counts keep growing after the first chunk (multi-channel signal through a feedback edge): "
                        "needs " + std::to_string(S) + " warm-up program(s)");
  }
  if (!dynamic_nch) b.nch = ns;
  for (auto &v : b.nch) for (int c : v) if (c > 16) throw Unsupported("more than 16 channels on one outlet");

  // ---- emit one instruction per unit-channel, in tick order ----
  std::map<std::pair<int, int>, int> cb_ring;  // (buffer, channel) -> ring id
  for (size_t bi = 0; bi < g.cbs.size(); bi++)
    for (int c = 0; c < g.cbs[bi].nch; c++) {
      int owner = 0;  // named after the first unit that holds the buffer (units are matched by label across graph edits)
      for (int u = NU - 1; u >= 0; u--) if (g.units[u].has("buffer") && (size_t)g.units[u].num("buffer", -1) == bi) owner = u;
      cb_ring[{(int)bi, c}] = b.ring(g.cbs[bi].len, Builder::uname(owner) + ".cb[" + std::to_string(c) + "]");
    }
  // ---- frame-rate units: every frame outlet is a per-voice ring that keeps its content between the unit's ticks
  std::map<int, int> frame_ring;  // unit -> ring of its outlet 0
  bool has_frames = false;
  for (int ui : g.order) {
    const GUnit &u = g.units[ui];
    if (!is_frame_kind(u.kind)) continue;
    has_frames = true;
    const long long n = frame_out_size(u);
    if (n < 1 || n > (1 << 20)) throw Unsupported("unit " + u.kind + " with a frame of " + std::to_string(n) + " samples");
    const long long ti = (long long)u.num("tickInterval", DUSP_CHUNK);
    if (ti < DUSP_CHUNK || ti % DUSP_CHUNK)
      throw Unsupported("unit " + u.kind + " ticks every " + std::to_string(ti) + " samples: only multiples of the 256-sample chunk run on the device");
    frame_ring[ui] = b.ring(n, Builder::uname(ui) + ".frame");
  }
  if (has_frames) {
    b.prog.chunk_tiles = true;
    // a standard inlet (or a frame unit's control inlet) wired to a frame outlet reads its first 256 samples
    for (int ui : g.order) {
      const GUnit &u = g.units[ui];
      for (size_t i = 0; i < u.in.size(); i++) {
        const GPort &p = u.in[i];
        if (!p.connected || p.nodata || !frame_ring.count(p.src_unit) || is_frame_inlet(u, (int)i)) continue;
        if (frame_out_size(g.units[p.src_unit]) != DUSP_CHUNK)
          throw Unsupported("standard inlet " + u.label + "." + p.name + " wired to a " + std::to_string(frame_out_size(g.units[p.src_unit])) +
                            "-sample frame outlet (chunk size mismatch; put a ReChunk in between)");
      }
    }
  }
  auto ctab_put = [&](const std::vector<double> &v) {
    const int off = (int)b.prog.ctab.size();
    b.prog.ctab.insert(b.prog.ctab.end(), v.begin(), v.end());
    return off;
  };
  std::map<long long, int> twiddle_at, envelope_at;
  auto twiddles = [&](long long size) {  // fft.js constructor: table[i] = cos(pi i / size), table[i+1] = -sin(pi i / size)
    if (twiddle_at.count(size)) return twiddle_at[size];
    std::vector<double> t((size_t)size * 2);
    for (long long i = 0; i < size * 2; i += 2) { const double a = M_PI * (double)i / (double)size; t[i] = std::cos(a); t[i + 1] = -std::sin(a); }
    return twiddle_at[size] = ctab_put(t);
  };
  auto envelope = [&](long long size) {  // Windower.js:28-47 ("hamming": sin^2), values as the Float32Array holds them
    if (envelope_at.count(size)) return envelope_at[size];
    std::vector<double> t((size_t)size);
    for (long long n = 0; n < size; n++) t[n] = (double)(float)std::pow(std::sin((M_PI * (double)n) / (double)(size - 1)), 2);
    return envelope_at[size] = ctab_put(t);
  };

  struct Tap { int unit; bool writer; Opnd off; int ir; };
  std::map<int, std::vector<Tap>> taps;  // ring -> taps (for the reordering check)

  struct TriggerFixup { int ir; std::string word; };
  std::vector<TriggerFixup> trigger_fixups;  // OP_POKEIF targets: the retriggered unit's words come later in the order
  for (int ui : g.order) {
    const GUnit &u = g.units[ui];
    const std::string &k = u.kind;
    const std::string un = Builder::uname(ui);
    auto inn = [&](int i) { return b.in_nch(b.nch, ui, i); };
    auto need_in = [&](size_t n) { if (u.in.size() < n || u.out.empty()) throw std::runtime_error("unit " + k + " has too few ports in graph blob"); };
    auto unary = [&](int op) { need_in(1); for (int c = 0; c < inn(0); c++) { IR &r = b.emit(op, ui, c); r.in[0] = b.in_ref(ui, 0, c); b.def(r, 0, ui, 0, c); } };
    auto binary_mod = [&](int op) {
      need_in(2);
      int n = std::max(inn(0), inn(1));
      for (int c = 0; c < n; c++) { IR &r = b.emit(op, ui, c); r.in[0] = b.in_ref(ui, 0, c % inn(0)); r.in[1] = b.in_ref(ui, 1, c % inn(1)); b.def(r, 0, ui, 0, c); }
    };
    if (k == "Osc") {
      need_in(1);
      IR &r = b.emit(OP_OSC, ui, 0); r.in[0] = b.in_ref(ui, 0, 0);
      r.p = b.word(un + ".phase", u.num("phase", 0), false);
      r.aux[0] = table_id(u.str("waveform", "sin"), false);
      b.def(r, 0, ui, 0, 0);
    } else if (k == "MultiChannelOsc") {
      need_in(1);
      for (int c = 0; c < inn(0); c++) {
        IR &r = b.emit(OP_MCOSC, ui, c); r.in[0] = b.in_ref(ui, 0, c);
        r.p = b.word(un + ".phase[" + std::to_string(c) + "]", 0, false);
        r.aux[0] = table_id(u.str("waveform", "sin"), false);
        b.def(r, 0, ui, 0, c);
      }
    } else if (k == "Ramp") {
      IR &r = b.emit(OP_RAMP, ui, 0);
      r.p = b.word(un + ".t", u.num("t", 0), false);
      b.word(un + ".playing", u.num("playing", 0), false);
      b.word(un + ".duration", u.num("duration", DUSP_SR), false);
      b.word(un + ".y0", u.num("y0", 1), false);
      b.word(un + ".y1", u.num("y1", 0), false);
      b.def(r, 0, ui, 0, 0);
    } else if (k == "Shape") {
      need_in(3);
      IR &r = b.emit(OP_SHAPE, ui, 0);
      for (int i = 0; i < 3; i++) r.in[i] = b.in_ref(ui, i, 0);
      r.p = b.word(un + ".t", u.num("t", 0), false);
      b.word(un + ".playing", u.num("playing", 0), false);
      b.word(un + ".leftEdge", u.num("leftEdge", 0), false);
      b.word(un + ".rightEdge", u.num("rightEdge", 0), false);
      r.aux[0] = table_id(u.str("shape", "decay"), true);
      bool le_shape = u.str("leftEdge", "") == "shape";
      bool re_shape = u.S.count("rightEdge") ? u.str("rightEdge", "") == "shape" : !u.has("rightEdge");
      if ((u.S.count("leftEdge") && !le_shape) || (u.S.count("rightEdge") && !re_shape))
        throw Unsupported("Shape edge must be \"shape\" or a number");
      r.aux[1] = (le_shape ? 1 : 0) | (re_shape ? 2 : 0);
      b.def(r, 0, ui, 0, 0);
    } else if (k == "AHD") {
      need_in(3);
      IR &r = b.emit(OP_AHD, ui, 0);
      for (int i = 0; i < 3; i++) r.in[i] = b.in_ref(ui, i, 0);
      r.p = b.word(un + ".t", u.num("t", 0), false);
      b.word(un + ".state", u.num("state", 0), false);
      b.word(un + ".playing", u.num("playing", 0), false);
      b.def(r, 0, ui, 0, 0);
    } else if (k == "SampleRateRedux") {
      need_in(2);
      for (int c = 0; c < inn(0); c++) {
        IR &r = b.emit(OP_SRREDUX, ui, c); r.in[0] = b.in_ref(ui, 0, c); r.in[1] = b.in_ref(ui, 1, 0);
        r.p = b.word(un + ".timeSinceLastUpdate[" + std::to_string(c) + "]", u.num("timeSinceLastUpdate", INFINITY), false);
        b.word(un + ".val[" + std::to_string(c) + "]", 0.0, false);
        b.def(r, 0, ui, 0, c);
      }
    } else if (k == "Monitor") {
      // Monitor.js:14-16: console.log only
    } else if (k == "Timer") {
      IR &r = b.emit(OP_TIMER, ui, 0);
      r.p = b.word(un + ".t", u.num("t", 0), false);
      b.def(r, 0, ui, 0, 0);
    } else if (k == "Retriggerer" || k == "SporadicRetriggerer") {
      // control-rate units: no outlet; they decide once per chunk whether to run target.trigger(),
      // whose assignments travel as TG records and become conditional word stores placed right after the unit
      if (u.in.empty()) throw std::runtime_error("unit " + k + " has too few ports in graph blob");
      const bool sporadic = k == "SporadicRetriggerer";
      int flag;
      {
        IR &r = b.emit(sporadic ? OP_SRETRIG : OP_RETRIG, ui, 0); r.in[0] = b.in_ref(ui, 0, 0); r.root = true;
        if (sporadic) {
          r.p = b.word(un + ".draw", u.num("draw", 0), false);
          b.word(un + ".seed_lo", u.num("seed_lo", 0), false);
          b.word(un + ".seed_hi", u.num("seed_hi", 0), false);
          flag = b.word(un + ".~fired", 0, false);
        } else {
          r.p = b.word(un + ".t", u.num("t", 0), false);
          flag = b.word(un + ".~fired", 0, false);
        }
      }
      int k_tg = 0;
      for (const GTrigger &tg : g.triggers) {
        if (tg.unit != ui) continue;
        // (a unit that ticks BEFORE this one -- an input of the chained target, reached through the default
        //  Unit.trigger, Unit.js:293-297 -- sees the assignment from its next tick on, here as in the reference)
        IR &q = b.emit(OP_POKEIF, ui, k_tg); q.root = true;
        q.aux[0] = flag;
        q.aux[1] = b.word(un + ".~tg" + std::to_string(k_tg++), tg.value, false);
        trigger_fixups.push_back({(int)b.ir.size() - 1, Builder::uname(tg.target) + "." + tg.prop});
      }
      b.prog.chunk_tiles = true;
    } else if (k == "Noise") {
      need_in(1);
      // A multi-channel Noise fills one channel after the other with ONE phase / held value / random stream
      // (Noise.js:17-27: `for c in out: for t: phase += f[0][t] ...`): one instruction per channel on the same state
      // words, in whole-chunk tiles so that channel c's 256 samples come before channel c + 1's.
      const int noise_nch = b.nch[ui][0];
      if (noise_nch > 1) b.prog.chunk_tiles = true;
      int pw = -1;
      for (int c = 0; c < noise_nch; c++) {
        IR &r = b.emit(OP_NOISE, ui, c); r.in[0] = b.in_ref(ui, 0, 0);
        // constant F >= sampleRate with phase 0: `phase += f` crosses the threshold on every sample (Noise.js:21-24)
        if (r.in[0].kind == 2 && (float)b.prog.words[r.in[0].id].dflt >= (float)DUSP_SR && u.num("phase", 0) == 0) {
          r.op = OP_NOISE1;
          if (c == 0) b.prog.locked.push_back({r.in[0].id, (double)DUSP_SR, INFINITY});
        }
        if (c == 0) {
          pw = b.word(un + ".phase", u.num("phase", 0), false);
          b.word(un + ".y", u.num("y", 0), false);
          b.word(un + ".draw", u.num("draw", 1), false);
          b.word(un + ".seed_lo", u.num("seed_lo", 0), false);
          b.word(un + ".seed_hi", u.num("seed_hi", 0), false);
        }
        b.ir.back().p = pw;
        b.def(b.ir.back(), 0, ui, 0, c);
      }
    } else if (k == "Filter") {
      need_in(2);
      std::string kind = u.str("kind", "LP");
      if (kind != "LP" && kind != "HP") throw Unsupported("Filter kind " + kind + " (the reference's BP/BR coefficient functions are never passed a bandwidth)");
      for (int c = 0; c < inn(0); c++) {
        IR &r = b.emit(OP_FILTER, ui, c); r.in[0] = b.in_ref(ui, 0, c); r.in[1] = b.in_ref(ui, 1, 0);
        std::string s = "[" + std::to_string(c) + "]";
        if (r.in[1].kind == 2) {  // constant F: coefficients are a pure function of f (Filter.js:66-84)
          r.op = OP_FILTERC;
          const char *names[9] = {"~a0", "~a1", "~a2", "~b1", "~b2", "x1", "x2", "y1", "y2"};
          for (int w = 0; w < 9; w++) {
            int id = b.word(un + "." + names[w] + s, 0.0, false);
            if (w == 0) r.p = id;
          }
          b.prog.derived.push_back({kind == "HP" ? DV_FILTER_HP : DV_FILTER_LP, r.p, r.in[1].id, -1, 0, 0});
        } else {
          const char *names[11] = {"a0", "a1", "a2", "b1", "b2", "lastF", "lastFdefined", "x1", "x2", "y1", "y2"};
          for (int w = 0; w < 11; w++) {
            int id = b.word(un + "." + names[w] + s, w < 5 ? NAN : 0.0, false);
            if (w == 0) r.p = id;
          }
        }
        r.aux[0] = kind == "HP";
        b.def(r, 0, ui, 0, c);
      }
    } else if (k == "Multiply") binary_mod(OP_MUL);
    else if (k == "Divide") binary_mod(OP_DIV);
    else if (k == "Pow") binary_mod(OP_POW);
    else if (k == "Sum") binary_mod(OP_ADD);
    else if (k == "Subtract" || k == "CrossFader") {
      need_in(k == "Subtract" ? 2 : 3);
      int n = std::max(inn(0), inn(1));
      for (int c = 0; c < n; c++) {
        IR &r = b.emit(k == "Subtract" ? OP_SUB : OP_XFADE, ui, c);
        r.in[0] = b.in_ref(ui, 0, c); r.in[1] = b.in_ref(ui, 1, c);  // zero-fill, not modulo (Subtract.js:23-24)
        if (k == "CrossFader") r.in[2] = b.in_ref(ui, 2, 0);
        b.def(r, 0, ui, 0, c);
      }
    } else if (k == "Gain") {
      need_in(2);
      for (int c = 0; c < inn(0); c++) {
        IR &r = b.emit(OP_GAIN, ui, c); r.in[0] = b.in_ref(ui, 0, c); r.in[1] = b.in_ref(ui, 1, 0);
        if (r.in[1].kind == 2) {  // constant GAIN: 10^(gain/20) once per voice
          r.op = OP_GAINC;
          r.p = b.word(un + ".~scale[" + std::to_string(c) + "]", 1.0, false);
          b.prog.derived.push_back({DV_GAIN, r.p, r.in[1].id, -1, 0, 0});
        }
        b.def(r, 0, ui, 0, c);
      }
    } else if (k == "PolarityInvert") unary(OP_NEG);
    else if (k == "Abs") unary(OP_ABS);
    else if (k == "Repeater") unary(OP_COPY);
    else if (k == "SecondsToSamples") unary(OP_S2S);
    else if (k == "SemitoneToRatio") unary(OP_SEMI);
    else if (k == "DecibelToScaler") unary(OP_DB2S);
    else if (k == "MidiToFrequency") {
      need_in(1);
      for (int c = 0; c < std::min(inn(0), b.nch[ui][0]); c++) { IR &r = b.emit(OP_MIDI2F, ui, c); r.in[0] = b.in_ref(ui, 0, c); b.def(r, 0, ui, 0, c); }
    } else if (k == "Clip" || k == "HardClipAbove" || k == "HardClipBelow") {
      need_in(2);
      int op = k == "Clip" ? OP_CLIP : k == "HardClipAbove" ? OP_HCA : OP_HCB;
      for (int c = 0; c < inn(0); c++) { IR &r = b.emit(op, ui, c); r.in[0] = b.in_ref(ui, 0, c); r.in[1] = b.in_ref(ui, 1, c % inn(1)); b.def(r, 0, ui, 0, c); }
    } else if (k == "Rescale") {
      need_in(5);
      for (int c = 0; c < inn(0); c++) {
        IR &r = b.emit(OP_RESCALE, ui, c); r.in[0] = b.in_ref(ui, 0, c);
        for (int i = 1; i < 5; i++) r.in[i] = b.in_ref(ui, i, c % inn(i));
        b.def(r, 0, ui, 0, c);
      }
    } else if (k == "Pan") {
      need_in(2);
      IR &r = b.emit(OP_PAN, ui, 0); r.in[0] = b.in_ref(ui, 0, 0); r.in[1] = b.in_ref(ui, 1, 0);
      if (r.in[1].kind == 2) {  // constant PAN: (1-pan), (1+pan) and the dB compensation once per voice
        r.op = OP_PANC;
        r.p = b.word(un + ".~left", 0, false);
        b.word(un + ".~right", 0, false);
        b.word(un + ".~compensation", 0, false);
        int cdb = b.word(un + ".compensationDB", u.num("compensationDB", 1.5), false);
        b.prog.derived.push_back({DV_PAN, r.p, r.in[1].id, cdb, 0, 0});
      } else {
        r.p = b.word(un + ".compensationDB", u.num("compensationDB", 1.5), false);
      }
      b.def(r, 0, ui, 0, 0); b.def(r, 1, ui, 0, 1);
    } else if (k == "FixedMultiply") {
      need_in(1);
      IR &r = b.emit(OP_FIXMUL, ui, 0); r.in[0] = b.in_ref(ui, 0, 0);
      r.p = b.word(un + ".sf", u.num("sf", NAN), false);
      b.def(r, 0, ui, 0, 0);
    } else if (k == "ConcatChannels") {
      need_in(2);
      int na = inn(0), n = na + inn(1);
      for (int c = 0; c < n; c++) { IR &r = b.emit(OP_COPY, ui, c); r.in[0] = c < na ? b.in_ref(ui, 0, c) : b.in_ref(ui, 1, c - na); b.def(r, 0, ui, 0, c); }
    } else if (k == "PickChannel") {
      need_in(2);
      const GPort &cp = u.in[1];
      if (cp.connected) {  // the index is a signal: a per-sample select over the input's channels (PickChannel.js:19)
        if (inn(0) > 4) throw Unsupported("PickChannel with a signal-rate index over more than 4 channels");
        IR &r = b.emit(OP_PICK, ui, 0);
        r.in[0] = b.in_ref(ui, 1, 0);
        for (int c = 0; c < inn(0); c++) r.in[1 + c] = b.in_ref(ui, 0, c);
        r.aux[0] = inn(0);
        b.def(r, 0, ui, 0, 0);
      } else {
        double idx = std::fmod((double)(float)cp.consts[0], (double)inn(0));  // PickChannel.js:19
        IR &r = b.emit(OP_COPY, ui, 0);
        if (idx == std::floor(idx) && idx >= 0 && idx < inn(0)) r.in[0] = b.in_ref(ui, 0, (int)idx);
        else throw Unsupported("PickChannel index selects no channel (the reference throws a TypeError)");
        b.def(r, 0, ui, 0, 0);
      }
    } else if (k == "VectorMagnitude") {
      need_in(1);
      if (inn(0) > 5) throw Unsupported("VectorMagnitude over more than 5 channels");
      IR &r = b.emit(OP_VECMAG, ui, 0);
      for (int c = 0; c < inn(0); c++) r.in[c] = b.in_ref(ui, 0, c);
      r.aux[0] = inn(0);
      b.def(r, 0, ui, 0, 0);
    } else if (k == "Delay" || k == "ReadBackDelay") {
      need_in(2);
      int n = std::max(inn(0), inn(1));
      for (int c = 0; c < n; c++) {
        IR &r = b.emit(k == "Delay" ? OP_DELAY : OP_READBACK, ui, c);
        // Delay.js:23 aliases in[c % length] while the array grows: every added channel is channel c % (original count)
        r.in[0] = b.in_ref(ui, 0, c % inn(0)); r.in[1] = b.in_ref(ui, 1, c % inn(1));
        const std::string rn = un + ".ring[" + std::to_string(c) + "]";
        if (k == "Delay") r.aux[0] = b.ring((long long)u.num("maxDelay", DUSP_SR * 5), rn);
        else {
          r.aux[0] = b.ring((long long)u.num("bufferLength", DUSP_SR), rn);
          r.p = b.word(un + ".tBuffer[" + std::to_string(c) + "]", u.num("tBuffer", 0), false);
        }
        b.prog.ring_bytes_per_sample += k == "Delay" ? 24 : 8;
        b.def(r, 0, ui, 0, c);
      }
    } else if (k == "MonoDelay") {
      need_in(2);
      IR &r = b.emit(OP_MONODELAY, ui, 0); r.in[0] = b.in_ref(ui, 0, 0); r.in[1] = b.in_ref(ui, 1, 0);
      r.aux[0] = b.ring(DUSP_SR * 5, un + ".ring[0]");
      b.prog.ring_bytes_per_sample += 24;
      b.def(r, 0, ui, 0, 0);
    } else if (k == "FixedDelay" || k == "CombFilter" || k == "AllPass") {
      need_in(k == "FixedDelay" ? 1 : 2);
      IR &r = b.emit(k == "FixedDelay" ? OP_FIXDELAY : k == "CombFilter" ? OP_COMB : OP_ALLPASS, ui, 0);
      r.in[0] = b.in_ref(ui, 0, 0);
      if (k != "FixedDelay") r.in[1] = b.in_ref(ui, 1, 0);
      r.aux[0] = b.ring((long long)u.num("delayTimeInSamples", 0), un + ".ring[0]");
      r.p = b.word(un + ".tBuffer", u.num("tBuffer", 0), false);
      b.prog.ring_bytes_per_sample += 8;
      b.def(r, 0, ui, 0, 0);
    } else if (k == "CircleBufferWriter" || k == "CircleBufferReader") {
      bool writer = k == "CircleBufferWriter";
      int bi = (int)u.num("buffer", -1);
      if (bi < 0 || bi >= (int)g.cbs.size()) throw std::runtime_error("CircleBuffer node without a buffer");
      int io = find_inlet(u, "offset");
      int ii = writer ? find_inlet(u, "in") : -1;
      for (int c = 0; c < g.cbs[bi].nch; c++) {
        IR &r = b.emit(writer ? OP_CBWRITE : OP_CBREAD, ui, c);
        r.in[0] = b.in_ref(ui, io, c % inn(io));
        if (writer) r.in[1] = b.in_ref(ui, ii, c);  // `if(this.in[c])`: absent channel -> no mix
        r.p = b.word(un + ".t[" + std::to_string(c) + "]", u.num("t", 0), false);
        r.aux[0] = cb_ring[{bi, c}];
        if (r.in[0].kind == 2) {  // constant OFFSET: the ring index just advances by one per sample
          r.op = writer ? OP_CBWRITEC : OP_CBREADC;
          int idx = b.word(un + ".~index[" + std::to_string(c) + "]", 0, false);
          b.prog.derived.push_back({DV_CBIDX, idx, r.p, r.in[0].id, writer ? 1.0 : -1.0, (double)g.cbs[bi].len});
        }
        r.aux[1] = (int)u.num(writer ? "preWipe" : "postWipe", 0);
        r.root = writer || r.aux[1];
        if (!writer) b.def(r, 0, ui, 0, c);
        taps[r.aux[0]].push_back({ui, writer || r.aux[1] != 0, r.in[0], (int)b.ir.size() - 1});
        b.prog.ring_bytes_per_sample += writer ? (r.aux[1] ? 4 : 8) : (r.aux[1] ? 8 : 4);
      }
    } else if (is_frame_kind(k)) {
      // SURVEY 8f-3.  Channel 0 of every port is the Float32Array its SignalChunk was built with; further channels
      // would be untyped Arrays (`out[c] = out[c] || new Array(n)`: doubles) -- refused, not approximated.
      const int gate = (int)u.num("tickInterval", DUSP_CHUNK);
      // ring holding the frame that frame inlet `i` reads (`need` samples)
      auto frame_in = [&](int i, long long need) {
        if ((int)u.in.size() <= i) throw std::runtime_error("unit " + k + " has too few ports in graph blob");
        const GPort &p = u.in[i];
        if (!p.connected) throw Unsupported("frame inlet " + u.label + "." + p.name + " is not connected");
        if (b.nch[p.src_unit][p.src_outlet] != 1) throw Unsupported("multi-channel signal into " + k + " (extra channels of the spectral units are untyped Arrays)");
        if (frame_ring.count(p.src_unit)) {
          if (frame_out_size(g.units[p.src_unit]) < need)
            throw Unsupported("frame inlet " + u.label + "." + p.name + " is fed by a shorter chunk (the reference reads undefined)");
          return frame_ring[p.src_unit];
        }
        if (need > DUSP_CHUNK) throw Unsupported("frame inlet " + u.label + "." + p.name + " is fed by a standard 256-sample outlet (the reference reads undefined)");
        const int r = b.ring(DUSP_CHUNK, un + ".in" + std::to_string(i));
        IR &st = b.emit(OP_STFRAME, ui, 0); st.in[0] = b.in_ref(ui, i, 0); st.aux[0] = r; st.root = true;
        return r;
      };
      auto gated = [&](int op) -> IR & { IR &r = b.emit(op, ui, 0); r.root = true; r.in[4].kind = 3; r.in[4].id = gate; return r; };
      auto pow2 = [](long long n) { return n > 1 && (n & (n - 1)) == 0; };
      if (k == "Hopper") {
        need_in(1);
        if (inn(0) != 1) throw Unsupported("multi-channel signal into Hopper (extra channels of the spectral units are untyped Arrays)");
        if (gate != DUSP_CHUNK) throw Unsupported("Hopper whose hop is not a multiple of 256 samples");
        const long long frame = (long long)u.num("frameSize", 0), hop = (long long)u.num("hopSize", 0);
        const int buf = b.ring(frame, un + ".buffer");
        IR &r = gated(OP_HOPPER);
        r.in[0] = b.in_ref(ui, 0, 0);
        r.aux[0] = buf; r.aux[1] = frame_ring[ui]; r.aux[2] = (int)hop;
        r.p = b.word(un + ".t", u.num("t", 0), false);
      } else if (k == "Windower") {
        const long long win = (long long)u.num("windowSize", 0);
        if (u.str("windowKind", "hamming") != "hamming") throw std::runtime_error("Window type '" + u.str("windowKind", "") + "' is not defined.");
        const int in = frame_in(0, win);
        IR &r = gated(OP_WINDOW);
        r.aux[0] = in; r.aux[1] = frame_ring[ui]; r.aux[2] = envelope(win);
      } else if (k == "FFT") {
        const long long win = (long long)u.num("windowSize", 0);
        if (!pow2(win)) throw std::runtime_error("FFT size must be a power of two and bigger than 1");
        const int in = frame_in(0, win);
        IR &r = gated(OP_FFT);
        r.aux[0] = in; r.aux[1] = frame_ring[ui]; r.aux[2] = twiddles(win);
      } else if (k == "IFFT") {
        const long long win = (long long)u.num("windowSize", 0);
        if (!pow2(win)) throw std::runtime_error("FFT size must be a power of two and bigger than 1");
        const int in = frame_in(0, 2 * win);
        const int cplx = b.ring(4 * win, un + ".complexOut");  // 2*win doubles
        IR &r = gated(OP_IFFT);
        r.aux[0] = in; r.aux[1] = frame_ring[ui]; r.aux[2] = twiddles(win);
        r.in[0].kind = 3; r.in[0].id = cplx;
      } else if (k == "UnHopper") {
        const long long win = (long long)u.num("windowSize", 0), hop = (long long)u.num("hopSize", 0);
        if (hop > win) throw Unsupported("UnHopper with a hop longer than its window");
        const int in = frame_in(0, win);
        const int acc = b.ring(2 * win, un + ".buffer");  // win doubles
        IR &r = gated(OP_UNHOP);
        r.aux[0] = in; r.aux[1] = frame_ring[ui]; r.aux[2] = acc;
        r.p = b.word(un + ".t", u.num("t", 0), false);
      } else if (k == "ReChunk") {
        const long long ii = (long long)u.num("inputInterval", 0), oi = (long long)u.num("outputInterval", 0);
        const long long bs = (long long)u.num("bufferSize", 0);
        if (ii < 1 || oi < 1 || bs < 1) throw std::runtime_error("ReChunk expects 2 numeric contructor arguments");
        const int in = frame_in(0, ii);
        const int buf = b.ring(bs, un + ".buffer");
        IR &r = gated(OP_RECHUNK);
        r.aux[0] = in; r.aux[1] = frame_ring[ui]; r.aux[2] = buf;
        r.in[0].kind = 3; r.in[0].id = (int)ii; r.in[1].kind = 3; r.in[1].id = (int)oi;
        r.p = b.word(un + ".t", u.num("t", 0), false);
      } else {
        const long long frame = 2 * (long long)u.num("windowSize", 4096);
        const int in = frame_in(0, frame);
        if (k == "SpectralSum") {
          const int in_b = frame_in(1, frame);
          IR &r = gated(OP_SPECSUM);
          r.aux[0] = in; r.aux[1] = frame_ring[ui]; r.aux[2] = in_b;
        } else if (k == "Augment") {
          IR &r = gated(OP_AUGMENT);
          const int n = (int)u.num("nmap", 0);
          r.aux[0] = in; r.aux[1] = frame_ring[ui]; r.aux[2] = n;
          for (int m = 0; m < n; m++) {
            const int w = b.word(un + ".ratio" + std::to_string(m), u.num("ratio" + std::to_string(m), 1), false);
            b.word(un + ".gain" + std::to_string(m), u.num("gain" + std::to_string(m), 1), false);
            if (m == 0) r.p = w;
          }
        } else {
          need_in(2);
          IR &r = gated(k == "SpectralGate" ? OP_SPECGATE : k == "BinShift" ? OP_BINSHIFT : OP_SPECPASS);
          r.in[0] = b.in_ref(ui, 1, 0);
          r.aux[0] = in; r.aux[1] = frame_ring[ui];
          r.aux[2] = k == "SpectralGate" ? (int)u.num("invert", 1) : k == "HardHighPass" ? 1 : 0;
        }
      }
      // a 256-sample frame outlet can feed standard units: their slots are loaded from the ring every chunk
      if (frame_out_size(u) == DUSP_CHUNK) {
        IR &ld = b.emit(OP_LDFRAME, ui, 0); ld.aux[0] = frame_ring[ui];
        b.def(ld, 0, ui, 0, 0);
      }
    } else {
      throw Unsupported("unit kind " + k + " is outside the accelerated render path");
    }
    // warm-up tick: units later in the order see this outlet with the channels it has just grown
    if (dynamic_nch && !u.out.empty()) b.nch[ui][0] = std::max(b.nch[ui][0], written_channels(g, b.nch, ui, b));
  }

  // ---- retrigger targets: the state words target.trigger() assigns ----
  for (auto &f : trigger_fixups) {
    int w = -1;
    for (size_t i = 0; i < b.prog.words.size(); i++) if (b.prog.words[i].name == f.word) { w = (int)i; break; }
    if (w < 0) throw Unsupported("trigger() assigns " + f.word + ", which is not unit state on the device");
    for (const Derived &d : b.prog.derived)
      if (d.src0 == w || d.src1 == w) throw Unsupported("trigger() assigns " + f.word + ", which was specialised at compile time");
    for (auto &l : b.prog.locked) if (l.word == w) throw Unsupported("trigger() assigns " + f.word + ", which is locked by a specialisation");
    b.ir[f.ir].p = w;
  }

  // ---- CircleBuffer taps: sub-block interleaving must not reorder accesses --
  // Two different units touching one ring see the same samples in the same
  // order as the reference (unit after unit per 256-chunk) iff their constant
  // time shifts are equal or at least one chunk apart on the ring.
  // Graphs outside that condition -- taps of different units closer than a chunk, a shared ring shorter than two
  // chunks, a signal-rate OFFSET on a shared ring (MultiTapDelay.js:28-44 with short or modulated taps) -- need the
  // reference's own order, unit after unit over a WHOLE chunk (CircleBufferWriter.js:12-25, CircleBufferReader.js:
  // 12-25): the program is rendered in whole-chunk tiles (interpreter, B * T = 256: every op covers the chunk before
  // the next one starts, signal-rate taps and rings shorter than the tile run their samples in turn).
  auto tap_sign = [&](const Tap &t) { return g.units[t.unit].kind == "CircleBufferWriter" ? 1 : -1; };
  for (auto &kv : taps) {
    auto &tv = kv.second;
    const long long len = b.prog.rings[kv.first].length;
    if (len < 2 * DUSP_CHUNK && tv.size() > 1) b.prog.chunk_tiles = true;
    for (size_t i = 0; i < tv.size(); i++)
      for (size_t j = i + 1; j < tv.size(); j++) {
        if (tv[i].unit == tv[j].unit || (!tv[i].writer && !tv[j].writer)) continue;
        if (tv[i].off.kind != 2 || tv[j].off.kind != 2) { b.prog.chunk_tiles = true; continue; }
        const double sa = tap_sign(tv[i]) * (double)DUSP_SR * (double)(float)b.prog.words[tv[i].off.id].dflt;
        const double sb = tap_sign(tv[j]) * (double)DUSP_SR * (double)(float)b.prog.words[tv[j].off.id].dflt;
        double d = std::fabs(sa - sb);
        if (!(d == d) || std::isinf(d)) continue;
        d = std::fmod(d, (double)len);
        d = std::min(d, (double)len - d);
        if (d != 0 && d < DUSP_CHUNK) b.prog.chunk_tiles = true;
      }
  }
  std::set<int> not_hoistable;
  for (auto &kv : taps) {
    if (b.prog.chunk_tiles) break;  // whole-chunk tiles keep the reference's order whatever the offsets are
    auto &tv = kv.second;
    long long len = b.prog.rings[kv.first].length;
    for (size_t i = 0; i < tv.size(); i++)
      for (size_t j = i + 1; j < tv.size(); j++) {
        if (tv[i].unit == tv[j].unit) continue;
        if (!tv[i].writer && !tv[j].writer) continue;
        if (tv[i].off.kind != 2 || tv[j].off.kind != 2)
          throw Unsupported("CircleBuffer taps with signal-rate OFFSET sharing a buffer");
        auto sign = [&](const Tap &t) { return g.units[t.unit].kind == "CircleBufferWriter" ? 1 : -1; };
        TapPair tp{tv[i].off.id, tv[j].off.id, sign(tv[i]), sign(tv[j]), len};
        if (!tv[i].writer || !tv[j].writer) {  // plain reader vs writer: strict when apart by default
          double sa = tp.sign_a * (double)DUSP_SR * (double)(float)b.prog.words[tp.word_a].dflt;
          double sb = tp.sign_b * (double)DUSP_SR * (double)(float)b.prog.words[tp.word_b].dflt;
          double d = std::fmod(std::fabs(sa - sb), (double)len);
          d = std::min(d, (double)len - d);
          tp.strict = d >= DUSP_CHUNK;
          if (!tp.strict) not_hoistable.insert(!tv[i].writer ? tv[i].ir : tv[j].ir);
        }
        b.prog.tap_pairs.push_back(tp);
      }
    // a plain constant-offset reader whose every writer stays a chunk or more away: its ring loads do not
    // depend on anything the same step stores, so a kernel may issue them early (aux[2] = 1)
    for (auto &t : tv)
      if (!t.writer && t.off.kind == 2 && !not_hoistable.count(t.ir) && b.ir[t.ir].op == OP_CBREADC) b.ir[t.ir].aux[2] = 1;
  }
  {
    std::vector<double> dv;
    for (auto &w : b.prog.words) dv.push_back(w.dflt);
    std::string msg = b.prog.check_taps(dv.data());
    if (!msg.empty()) throw Unsupported(msg);
  }

  // ---- peephole: preWipe writer + mixing writer on the same ring position -> one writer ----
  // MultiTapDelay.addFeedback (MultiTapDelay.js:35-43) mixes the feedback into the position the input
  // writer has just wiped and written.  With no other tap on that ring ticking in between, the pair
  // leaves (0 + a) + b there: one store instead of store, load, add, store.  Needs the second input
  // to be available where the first writer ticks (same fresh / stale status at both positions).
  if (!getenv("DUSP_NO_WRITER_MERGE")) {
    std::map<int, std::vector<int>> ring_ops;
    for (size_t i = 0; i < b.ir.size(); i++)
      if (b.ir[i].op == OP_CBWRITEC || b.ir[i].op == OP_CBREADC || b.ir[i].op == OP_CBWRITE || b.ir[i].op == OP_CBREAD)
        ring_ops[b.ir[i].aux[0]].push_back((int)i);
    for (auto &kv : ring_ops)
      for (size_t k = 0; k + 1 < kv.second.size(); k++) {
        const int i = kv.second[k], j = kv.second[k + 1];
        IR &wi = b.ir[i], &wj = b.ir[j];
        if (wi.op != OP_CBWRITEC || wj.op != OP_CBWRITEC || wi.aux[1] != 1 || wj.aux[1] != 0) continue;
        if (wi.in[1].kind == 0 || wj.in[1].kind == 0) continue;
        const Param &oi = b.prog.words[wi.in[0].id], &oj = b.prog.words[wj.in[0].id];
        if ((float)oi.dflt != (float)oj.dflt || b.prog.words[wi.p].dflt != b.prog.words[wj.p].dflt) continue;
        const Opnd bo = wj.in[1];
        if (bo.kind == 1) { const int d = b.vals[bo.id].def; if (!(d >= 0 && d < i) && !(d >= j)) continue; }
        b.prog.equal_words.push_back({wi.in[0].id, wj.in[0].id});
        b.prog.equal_words.push_back({wi.p, wj.p});
        b.vals.push_back({wi.unit, -1, wi.chan, i});
        const int tmp = (int)b.vals.size() - 1;
        const Opnd a = wi.in[1];
        wi = IR();
        wi.op = OP_ADD0; wi.in[0] = a; wi.in[1] = bo; wi.outv[0] = tmp; wi.unit = wj.unit; wi.chan = wj.chan; wi.pos = wj.pos;
        wj.in[1].kind = 1; wj.in[1].id = tmp;
        wj.aux[1] = 1;
      }
  }

  // ---- rendered outlet -> OUT instructions ----
  const int out_nch = b.nch[g.root_unit][g.root_outlet];
  b.prog.out_nch = out_nch;
  std::vector<IR> outs;
  for (int c = 0; c < out_nch; c++) {
    IR r; r.op = OP_OUT; r.unit = g.root_unit; r.chan = c; r.aux[0] = c; r.root = true;
    if (frame_ring.count(g.root_unit) && frame_out_size(g.units[g.root_unit]) != DUSP_CHUNK) {
      // renderChannelData.js:25-45 copies outlet.chunkSize samples after every tickUntil(t0 + chunkSize).  A unit that
      // ticks once per such span (at its start) leaves the whole span in its outlet, so each 256-sample chunk can be
      // stored as soon as it is reached.
      const GUnit &ru = g.units[g.root_unit];
      const long long cs = frame_out_size(ru), ti = (long long)ru.num("tickInterval", DUSP_CHUNK);
      if (cs % DUSP_CHUNK || ti % cs) throw Unsupported("rendered outlet has chunks of " + std::to_string(cs) + " samples but its unit ticks every " + std::to_string(ti));
      r.op = OP_OUTRING; r.aux[1] = frame_ring[g.root_unit]; r.aux[2] = (int)cs;
      r.pos = 1 << 29;
      outs.push_back(r);
      continue;
    }
    auto it = b.valmap.find(std::make_tuple(g.root_unit, g.root_outlet, c));
    if (it != b.valmap.end() && b.vals[it->second].def >= 0) { r.in[0].kind = 1; r.in[0].id = it->second; }
    r.pos = 1 << 29;
    outs.push_back(r);
  }

  // ---- dead-code elimination (keep OUT, CircleBuffer side effects, and their inputs) ----
  std::vector<IR> &ir = b.ir;
  for (auto &o : outs) ir.push_back(o);
  // re-index defs (emit() may have reallocated; defs were recorded by index so they are valid)
  std::vector<char> live(ir.size(), 0);
  std::vector<int> stack;
  // (a graph edit ahead: a later program may read any of this one's values one chunk late, see KEEP below)
  for (size_t i = 0; i < ir.size(); i++) if (ir[i].root || g.keep_all) { live[i] = 1; stack.push_back((int)i); }
  while (!stack.empty()) {
    int i = stack.back(); stack.pop_back();
    for (auto &o : ir[i].in)
      if (o.kind == 1) { int d = b.vals[o.id].def; if (d >= 0 && !live[d]) { live[d] = 1; stack.push_back(d); } }
  }
  // values without a producer read as silence (a channel the unit never writes stays zero-filled)
  for (auto &r : ir) for (auto &o : r.in) if (o.kind == 1 && b.vals[o.id].def < 0) o = Opnd();

  // ---- order + stale edges ----
  struct Item { IR r; int orig; };
  std::vector<IR> seq;
  std::vector<int> newidx(ir.size(), -1);
  for (size_t i = 0; i < ir.size(); i++) if (live[i]) { newidx[i] = (int)seq.size(); seq.push_back(ir[i]); }
  for (auto &v : b.vals) v.def = v.def >= 0 ? newidx[v.def] : -1;
  // place each OUT right after its producer (keeps the value's live range short)
  {
    std::vector<IR> body, tail;
    for (auto &r : seq) (r.op == OP_OUT ? tail : body).push_back(r);
    std::vector<IR> res;
    std::map<int, std::vector<IR>> after;  // producer body index -> outs
    std::vector<IR> front;
    // recompute def index within `body`
    std::map<int, int> defidx;
    for (size_t i = 0; i < body.size(); i++) for (int k = 0; k < 2; k++) if (body[i].outv[k] >= 0) defidx[body[i].outv[k]] = (int)i;
    for (auto &o : tail) {
      if (o.in[0].kind == 1) after[defidx.at(o.in[0].id)].push_back(o);
      else front.push_back(o);
    }
    for (auto &o : front) res.push_back(o);
    for (size_t i = 0; i < body.size(); i++) { res.push_back(body[i]); for (auto &o : after[(int)i]) res.push_back(o); }
    seq.swap(res);
  }
  auto recompute_defs = [&]() {
    for (auto &v : b.vals) v.def = -1;
    for (size_t i = 0; i < seq.size(); i++) for (int k = 0; k < 2; k++) if (seq[i].outv[k] >= 0) b.vals[seq[i].outv[k]].def = (int)i;
  };
  recompute_defs();
  // stale reads: consumer scheduled at or before its producer
  {
    std::map<int, int> stale_id;   // value -> stale buffer
    std::map<int, int> stale_tmp;  // value -> temp value holding the loaded previous chunk
    std::vector<IR> res;
    auto keep = [&](int v) {
      stale_id[v] = b.prog.n_stale++;
      b.prog.stale_names.push_back(Builder::uname(b.vals[v].unit) + ".stale[" + std::to_string(b.vals[v].outlet) + "][" +
                                   std::to_string(b.vals[v].chan) + "]");
    };
    // KEEP: the circuit is edited at a later chunk.  The edit may re-level units so that an
    // existing outlet is then read one chunk late (a new feedback edge): every outlet value leaves its last
    // chunk in a stale buffer, where dusp_session_adopt finds it by name.
    b.prog.keeps_all = g.keep_all;
    if (g.keep_all)
      for (size_t v = 0; v < b.vals.size(); v++) if (b.vals[v].def >= 0 && b.vals[v].outlet >= 0) keep((int)v);
    for (size_t i = 0; i < seq.size(); i++) {
      IR r = seq[i];
      for (auto &o : r.in) {
        if (o.kind != 1) continue;
        int d = b.vals[o.id].def;
        if (d >= 0 && d >= (int)i) {
          int v = o.id;
          if (!stale_tmp.count(v)) {
            if (!stale_id.count(v)) keep(v);
            b.vals.push_back({b.vals[v].unit, b.vals[v].outlet, b.vals[v].chan, -1});
            stale_tmp[v] = (int)b.vals.size() - 1;
            IR ld; ld.op = OP_LDSTALE; ld.unit = b.vals[v].unit; ld.chan = b.vals[v].chan;
            ld.aux[0] = stale_id[v]; ld.outv[0] = stale_tmp[v];
            res.push_back(ld);
          }
          o.id = stale_tmp[v];
        }
      }
      res.push_back(r);
      for (int k = 0; k < 2; k++)
        if (r.outv[k] >= 0 && stale_id.count(r.outv[k])) {
          IR st; st.op = OP_STSTALE; st.unit = r.unit; st.chan = r.chan; st.aux[0] = stale_id[r.outv[k]];
          st.in[0].kind = 1; st.in[0].id = r.outv[k];
          res.push_back(st);
        }
    }
    // a producer may precede the discovery of its stale consumers only if the consumer comes later,
    // which by definition is not stale; so every STSTALE was emitted above.
    seq.swap(res);
  }
  recompute_defs();

  // ---- evaluation order ----
  // The reference ticks a feed-forward circuit level by level (Circuit.js:67-107: every leaf first), so all of a
  // tree's leaves are live at once: 28 slots for a 64-unit W4 circuit, 15 KB of shared memory per warp and 14
  // circuits per SM in a heterogeneous batch.  Where the order of two ops cannot be observed -- no value is read one
  // chunk late, every op owns its state words and its ring, nothing pokes another unit's state -- the ops are
  // re-sequenced depth first from the sinks, deeper operand first (Sethi-Ullman): same values, a handful of slots.
  {
    auto own_state_only = [](int op) {
      switch (op) {
        case OP_COPY: case OP_NEG: case OP_ABS: case OP_S2S: case OP_SEMI: case OP_DB2S: case OP_MIDI2F: case OP_FIXMUL:
        case OP_MUL: case OP_DIV: case OP_POW: case OP_ADD: case OP_SUB: case OP_GAIN: case OP_CLIP: case OP_HCA: case OP_HCB:
        case OP_RESCALE: case OP_XFADE: case OP_PAN: case OP_VECMAG: case OP_OSC: case OP_MCOSC: case OP_RAMP: case OP_SHAPE:
        case OP_AHD: case OP_TIMER: case OP_NOISE: case OP_FILTER: case OP_FIXDELAY: case OP_COMB: case OP_ALLPASS:
        case OP_OUT: case OP_FILTERC: case OP_PANC: case OP_GAINC: case OP_NOISE1: case OP_ADD0: case OP_PICK:
          return true;
        default:
          return false;
      }
    };
    auto has_ring = [](int op) { return op == OP_FIXDELAY || op == OP_COMB || op == OP_ALLPASS; };
    bool ok = b.prog.n_stale == 0 && !g.keep_all && !b.prog.chunk_tiles && !getenv("DUSP_KEEP_TICK_ORDER");
    std::set<int> words_seen, rings_seen;
    const int n = (int)seq.size();
    for (int i = 0; i < n && ok; i++) {
      const IR &r = seq[i];
      if (!own_state_only(r.op)) ok = false;
      if (r.p >= 0 && !words_seen.insert(r.p).second) ok = false;
      if (has_ring(r.op) && !rings_seen.insert(r.aux[0]).second) ok = false;
      for (auto &o : r.in) if (o.kind == 1 && !(b.vals[o.id].def >= 0 && b.vals[o.id].def < i)) ok = false;
    }
    if (ok && n > 2) {
      std::vector<std::vector<int>> kids(n);
      std::vector<char> is_read(n, 0);
      std::vector<int> need(n, 1);
      for (int i = 0; i < n; i++) {
        for (auto &o : seq[i].in)
          if (o.kind == 1) {
            const int d = b.vals[o.id].def;
            is_read[d] = 1;
            if (std::find(kids[i].begin(), kids[i].end(), d) == kids[i].end()) kids[i].push_back(d);
          }
        std::stable_sort(kids[i].begin(), kids[i].end(), [&](int a, int c) { return need[a] > need[c]; });
        for (size_t k = 0; k < kids[i].size(); k++) need[i] = std::max(need[i], need[kids[i][k]] + (int)k);
      }
      std::vector<char> done(n, 0);
      std::vector<IR> res;
      res.reserve(n);
      std::vector<std::pair<int, size_t>> stack;
      for (int root = 0; root < n; root++) {
        if (is_read[root] || done[root]) continue;
        stack.push_back({root, 0});
        done[root] = 1;
        while (!stack.empty()) {
          auto &top = stack.back();
          if (top.second < kids[top.first].size()) {
            const int c = kids[top.first][top.second++];
            if (!done[c]) { done[c] = 1; stack.push_back({c, 0}); }
          } else {
            res.push_back(seq[top.first]);
            stack.pop_back();
          }
        }
      }
      if ((int)res.size() == n) { seq.swap(res); recompute_defs(); }
    }
  }

  // ---- slot allocation: linear scan, in-place reuse allowed (ops load all inputs before storing) ----
  std::vector<int> last_use(b.vals.size(), -1);
  for (size_t i = 0; i < seq.size(); i++) for (auto &o : seq[i].in) if (o.kind == 1) last_use[o.id] = (int)i;
  std::vector<int> slot_of(b.vals.size(), -1);
  std::vector<int> free_slots;
  int n_slots = 0;
  for (size_t i = 0; i < seq.size(); i++) {
    IR &r = seq[i];
    std::set<int> freed;
    for (auto &o : r.in)
      if (o.kind == 1 && last_use[o.id] == (int)i && !freed.count(o.id)) { freed.insert(o.id); free_slots.push_back(slot_of[o.id]); }
    for (int k = 0; k < 2; k++) {
      int v = r.outv[k];
      if (v < 0) continue;
      if (last_use[v] < 0) { r.outv[k] = -1; continue; }  // never read: op writes nowhere
      int s;
      if (!free_slots.empty()) { std::sort(free_slots.begin(), free_slots.end(), std::greater<int>()); s = free_slots.back(); free_slots.pop_back(); }
      else s = n_slots++;
      slot_of[v] = s;
    }
  }
  b.prog.n_slots = std::max(n_slots, 1);

  // ---- encode ----
  for (auto &r : seq) {
    DuspInstr di;
    std::memset(&di, 0, sizeof di);
    di.op = r.op;
    for (int k = 0; k < 5; k++)
      di.in[k] = r.in[k].kind == 1 ? slot_of[r.in[k].id] : r.in[k].kind == 2 ? ref_const(r.in[k].id) : r.in[k].kind == 3 ? r.in[k].id : REF_NONE;
    for (int k = 0; k < 2; k++) di.out[k] = r.outv[k] >= 0 ? slot_of[r.outv[k]] : REF_NONE;
    di.p = r.p;
    for (int k = 0; k < 3; k++) di.aux[k] = r.aux[k];
    di.unit = r.unit; di.chan = r.chan;
    b.prog.code.push_back(di);
  }
  return b.prog;
}

std::string Program::check_taps(const double *wv) const {
  for (auto &t : tap_pairs) {
    double sa = t.sign_a * (double)DUSP_SR * (double)(float)wv[t.word_a];
    double sb = t.sign_b * (double)DUSP_SR * (double)(float)wv[t.word_b];
    double d = std::fabs(sa - sb);
    if (!(d == d) || std::isinf(d)) return "CircleBuffer tap OFFSET is not finite";
    d = std::fmod(d, (double)t.ring_len);
    d = std::min(d, (double)t.ring_len - d);
    if (d != 0 && d < DUSP_CHUNK)
      return "CircleBuffer taps closer than one 256-sample chunk (intra-chunk ordering differs from the reference)";
    if (d == 0 && t.strict)
      return "CircleBuffer reader OFFSET coincides with a writer's: the program was compiled with these taps apart "
             "(compile it with the coinciding offsets as the inlet constants instead)";
  }
  return "";
}

// Host evaluation of the per-voice words that depend only on other per-voice constants.  Same
// formulas and the same libm calls as the reference's Math.* (and as the oracle), evaluated in double.
void derive_words(const Program &p, double *w, long long stride) {
  auto W = [&](int i) -> double & { return w[(long long)i * stride]; };
  for (const Derived &d : p.derived) {
    switch (d.kind) {
      case DV_FILTER_LP: {  // Filter.js:67-75
        const double f = (double)(float)W(d.src0);
        const double lamda = 1 / std::tan(M_PI * f / DUSP_SR), l2 = lamda * lamda;
        const double a0 = 1 / (1 + 2 * lamda + l2);
        W(d.dst) = a0; W(d.dst + 1) = 2 * a0; W(d.dst + 2) = a0;
        W(d.dst + 3) = 2 * a0 * (1 - l2); W(d.dst + 4) = a0 * (1 - 2 * lamda + l2);
        break;
      }
      case DV_FILTER_HP: {  // Filter.js:76-84
        const double f = (double)(float)W(d.src0);
        const double lamda = std::tan(M_PI * f / DUSP_SR), l2 = lamda * lamda;
        const double a0 = 1 / (1 + 2 * lamda + l2);
        W(d.dst) = a0; W(d.dst + 1) = 0; W(d.dst + 2) = -a0;
        W(d.dst + 3) = 2 * a0 * (l2 - 1); W(d.dst + 4) = a0 * (1 - 2 * lamda + l2);
        break;
      }
      case DV_PAN: {  // Pan.js:20-22
        const double pan = (double)(float)W(d.src0), cdb = W(d.src1);
        W(d.dst) = 1 - pan; W(d.dst + 1) = 1 + pan;
        W(d.dst + 2) = std::pow(10, ((1 - std::fabs(pan)) * cdb) / 20);
        break;
      }
      case DV_GAIN:  // Gain.js:26-28
        W(d.dst) = std::pow(10, (double)(float)W(d.src0) / 20);
        break;
      case DV_CBIDX: {  // CircleBuffer.js:17-20 applied to the unit's first access (sample 0)
        const double t = W(d.src0) + d.aux0 * ((double)DUSP_SR * (double)(float)W(d.src1));
        double i = std::floor(std::fmod(t, d.aux1));
        if (i < 0) i += d.aux1;
        W(d.dst) = (i == i && i >= 0 && i < d.aux1) ? i : -1.0;  // -1: never a valid index (NaN / Infinity offset)
        break;
      }
    }
  }
}

std::string Program::check_locked(const double *wv) const {
  for (auto &e : equal_words)
    if ((float)wv[e.word_a] != (float)wv[e.word_b])
      return "parameters " + words[e.word_a].name + " and " + words[e.word_b].name +
             " were merged at compile time (two writers on one CircleBuffer position) and must stay equal";
  for (auto &l : locked) {
    const float v = (float)wv[l.word];
    if (!(v >= (float)l.lo && v <= (float)l.hi))
      return "parameter " + words[l.word].name + " was specialised at compile time and must stay in [" +
             std::to_string(l.lo) + ", " + std::to_string(l.hi) + "]";
  }
  return "";
}

std::string Program::listing() const {
  std::string s;
  char buf[256];
  snprintf(buf, sizeof buf, "; %zu instr, %d slots, %d stale, %zu rings (%lld floats/voice), %zu words, %d ch out\n",
           code.size(), n_slots, n_stale, rings.size(), ring_floats, words.size(), out_nch);
  s += buf;
  for (size_t i = 0; i < code.size(); i++) {
    const DuspInstr &d = code[i];
    snprintf(buf, sizeof buf, "%3zu %-9s u%d.%d ", i, op_name(d.op), d.unit, d.chan);
    s += buf;
    for (int k = 0; k < 5; k++) {
      if (d.in[k] == REF_NONE) continue;
      if (d.in[k] >= 0) snprintf(buf, sizeof buf, "s%d ", d.in[k]);
      else snprintf(buf, sizeof buf, "$%s ", words[REF_CONST_BASE - d.in[k]].name.c_str());
      s += buf;
    }
    s += "-> ";
    for (int k = 0; k < 2; k++) if (d.out[k] >= 0) { snprintf(buf, sizeof buf, "s%d ", d.out[k]); s += buf; }
    snprintf(buf, sizeof buf, "[p=%d aux=%d,%d,%d]\n", d.p, d.aux[0], d.aux[1], d.aux[2]);
    s += buf;
  }
  return s;
}

}  // namespace dusp
