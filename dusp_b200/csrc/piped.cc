// piped.cc -- the pipelined render kernel: one CTA renders VC voices, 256 (or 128) samples per step.
//
// Why: inside a 256-sample chunk no unit depends on a later unit (every feedback path of the
// reference has >= one chunk of latency), so a circuit's op program splits into
//   stage 0  time-PARALLEL ops that feed the sequential ones       (e.g. Noise)
//   stage 1  time-SEQUENTIAL ops: recursive filters, envelopes ... (e.g. the biquad recurrence)
//   stage 2  time-parallel ops that consume them                   (rings, Multiply/Sum, Pan, copy-out)
// and the stages can run on DIFFERENT steps at the same moment (software pipeline):
//   * warps 0..3 ("P-warps", VC/4 voices each) run stage 0 of step k+1 and stage 2 of step k-1 with
//     LANES ALONG TIME: lane l owns samples [8l, 8l+8) of the step, so a warp touches 1 KB contiguous
//     of a ring / stale buffer / output row per access (fully coalesced), phase accumulators
//     are warp-shuffle scans (ops.cuh cores with T = 32);
//   * warp 4 (the "S-warp") runs stage 1 of step k with ONE VOICE PER LANE: VC independent
//     recurrences in exactly the reference's sequential order (ops.cuh cores with T = 1), unit
//     state in registers for the whole launch;
//   * signals crossing a stage boundary live in shared memory, [rotation 3][buffer][voice][256]
//     floats; the float4 column is XOR-swizzled with the voice so that both access patterns (a row
//     per warp / a column per warp) are bank-conflict free;
//   * one __syncthreads per step.  The recurrences' latency (the bound of this path: VC chains per
//     S-warp, ~60 cycles per sample) is hidden behind the P-warps' work instead of serialising it.
// The arithmetic is the same ops.cuh cores the tile kernel uses: results are identical.
#include <algorithm>
#include <cstdlib>
#include <map>
#include <set>
#include <sstream>
#include <vector>

#include "fused.h"
#include "fusedgen.h"

namespace dusp {

namespace {

// Samples per pipeline step: 256 (a chunk; lanes own 8 samples) when the exchange buffers leave room for
// three CTAs per SM, else 128 (lanes own 4).  Rings shorter than a chunk never run time-parallel.
constexpr int MAX_STEP = 256;

// ops with no time-parallel form (or none worth having): they run in the lane-per-voice warp
bool sequential_only(const Program &p, const DuspInstr &I) {
  switch (I.op) {
    case OP_MCOSC: case OP_SHAPE: case OP_AHD: case OP_TIMER: case OP_SRREDUX: case OP_NOISE: case OP_FILTER: case OP_FILTERC:
    case OP_DELAY: case OP_MONODELAY: case OP_READBACK: case OP_CBWRITE: case OP_CBREAD:
      return true;
    case OP_FIXDELAY: case OP_COMB: case OP_ALLPASS: case OP_CBWRITEC: case OP_CBREADC:
      return p.rings[I.aux[0]].length < MAX_STEP;  // the step's samples would collide inside the ring
    default:
      return false;
  }
}

// ops that communicate through per-voice memory must stay in one stage (and keep their order)
int coupling_key(const DuspInstr &I) {
  switch (I.op) {
    case OP_FIXDELAY: case OP_COMB: case OP_ALLPASS: case OP_DELAY: case OP_MONODELAY: case OP_READBACK:
    case OP_CBWRITE: case OP_CBREAD: case OP_CBWRITEC: case OP_CBREADC:
      return I.aux[0];
    case OP_LDSTALE: case OP_STSTALE:
      return 1000000 + I.aux[0];
    default:
      return -1;
  }
}

struct Cross {       // a value that crosses a stage boundary
  int def, k;        // defining instruction, output index
  int slot;
  int from, to;      // producing stage, last consuming stage
  bool used_in[3] = {false, false, false};
  int buf = -1;
};

struct Plan {
  std::vector<int> stage;
  std::vector<Cross> cross;
  std::map<std::pair<int, int>, int> cross_of;  // (def, k) -> index
  int n_buffers = 0;
};

bool make_plan(const Program &p, Plan &pl) {
  const int n = (int)p.code.size();
  std::vector<std::vector<std::pair<int, int>>> prod(n);  // per instruction: (def instr, output index) of each slot operand
  std::map<int, std::pair<int, int>> def_of_slot;
  for (int i = 0; i < n; i++) {
    const DuspInstr &I = p.code[i];
    for (int k = 0; k < 5; k++)
      if (I.in[k] >= 0) {
        auto it = def_of_slot.find(I.in[k]);
        if (it != def_of_slot.end()) prod[i].push_back(it->second);
      }
    for (int k = 0; k < 2; k++) if (I.out[k] >= 0) def_of_slot[I.out[k]] = {i, k};
  }
  std::vector<char> seq(n), anc(n, 0), desc(n, 0);
  for (int i = 0; i < n; i++) seq[i] = sequential_only(p, p.code[i]);
  // desc: depends on a sequential op; anc: feeds one.  Ops coupled through a ring or a stale buffer act as one
  // node (a feedback loop closes through its stale buffer: the whole loop then sits between sequential ops and
  // runs in the lane-per-voice warp), so the two properties are propagated to a fixed point.
  for (bool changed = true; changed;) {
    changed = false;
    auto set = [&](std::vector<char> &v, int i) { if (!v[i]) { v[i] = 1; changed = true; } };
    for (int i = 0; i < n; i++)
      for (auto &d : prod[i]) {
        if (seq[d.first] || desc[d.first]) set(desc, i);
        if (seq[i] || anc[i]) set(anc, d.first);
      }
    std::map<int, std::pair<bool, bool>> grp;
    for (int i = 0; i < n; i++) { int key = coupling_key(p.code[i]); if (key >= 0) { grp[key].first |= anc[i] != 0; grp[key].second |= (desc[i] || seq[i]); } }
    for (int i = 0; i < n; i++) {
      int key = coupling_key(p.code[i]);
      if (key < 0) continue;
      if (grp[key].first) set(anc, i);
      if (grp[key].second && !seq[i]) set(desc, i);
    }
  }
  pl.stage.assign(n, 2);
  for (int i = 0; i < n; i++) pl.stage[i] = seq[i] ? 1 : (anc[i] && desc[i]) ? 1 : anc[i] ? 0 : 2;
  // fixed point: consumers never run before producers; memory-coupled ops share a stage
  for (int round = 0; round < 64; round++) {
    bool changed = false;
    std::map<int, int> group_max;
    for (int i = 0; i < n; i++) { int key = coupling_key(p.code[i]); if (key >= 0) group_max[key] = std::max(group_max[key], pl.stage[i]); }
    for (int i = 0; i < n; i++) {
      int s = pl.stage[i];
      int key = coupling_key(p.code[i]);
      if (key >= 0) s = std::max(s, group_max[key]);
      for (auto &d : prod[i]) s = std::max(s, pl.stage[d.first]);
      if (s != pl.stage[i]) { pl.stage[i] = s; changed = true; }
    }
    if (!changed) break;
  }
  for (int i = 0; i < n; i++) if (seq[i] && pl.stage[i] != 1) return false;  // a sequential op behind stage-2 data: needs > 3 stages
  // crossing values
  for (int i = 0; i < n; i++)
    for (auto &d : prod[i]) {
      if (pl.stage[i] == pl.stage[d.first]) continue;
      auto key = std::make_pair(d.first, d.second);
      if (!pl.cross_of.count(key)) {
        Cross c;
        c.def = d.first; c.k = d.second; c.slot = p.code[d.first].out[d.second]; c.from = pl.stage[d.first]; c.to = pl.stage[i];
        pl.cross_of[key] = (int)pl.cross.size();
        pl.cross.push_back(c);
      }
      Cross &c = pl.cross[pl.cross_of[key]];
      c.to = std::max(c.to, pl.stage[i]);
      c.used_in[pl.stage[i]] = true;
    }
  // buffers: a stage-1 output may take over the buffer of a stage-0 value that dies in stage 1
  // (the S-warp loads a 4-sample group before it stores into it)
  std::vector<int> reusable;
  for (auto &c : pl.cross) if (c.from == 0) { c.buf = pl.n_buffers++; if (c.to == 1) reusable.push_back(c.buf); }
  for (auto &c : pl.cross) if (c.from == 1) {
    if (!reusable.empty()) { c.buf = reusable.back(); reusable.pop_back(); }
    else c.buf = pl.n_buffers++;
  }
  return true;
}

}  // namespace

namespace {
std::string generate(const Program &p, int VC, int STEP, PipedInfo *info, bool want_ffp = false);
std::string generate_ffq(const Program &p, int VC, PipedInfo *info);
}

bool program_has_sequential(const Program &p) {
  for (const DuspInstr &I : p.code) if (sequential_only(p, I)) return true;
  return false;
}

std::string generate_piped_source(const Program &p, int VC, PipedInfo *info, int out_mode) {
  if (VC < 8 || VC > 32 || VC % 8) return "";
  std::string src;
  const char *fe = getenv("DUSP_PIPED_FFP");
  if (fe && atoi(fe) == 2) src = generate_ffq(p, VC, info);
  else if (fe && atoi(fe) != 0) src = generate(p, VC, 128, info, true);  // "" when the circuit has no such form: the usual kernel then
  if (!src.empty()) return fused_mode_prefix(out_mode) + src;
  if (const char *se = getenv("DUSP_PIPED_STEP")) src = generate(p, VC, atoi(se) == 128 ? 128 : 256, info);
  else {
    PipedInfo pi;
    src = generate(p, VC, 256, &pi);
    if (src.empty() || pi.smem_bytes > 76400) src = generate(p, VC, 128, &pi);
    if (info) *info = pi;
  }
  return src.empty() ? src : fused_mode_prefix(out_mode) + src;
}

namespace {

std::string generate(const Program &p, int VC, int STEP, PipedInfo *info, bool want_ffp) {
  const int PB = STEP / 32;  // samples per lane in the P-warps and per S-warp group
  int NP = 4;                // P-warps per CTA (VC / NP voices each); DUSP_PIPED_PWARPS overrides (experiments)
  if (const char *e = getenv("DUSP_PIPED_PWARPS")) { int v = atoi(e); if (v >= 1 && v <= 8 && VC % v == 0) NP = v; }
  Plan pl;
  if (!make_plan(p, pl)) return "";
  const int n = (int)p.code.size();
  const int NB = pl.n_buffers;
  // Feed-forward split (DUSP_PIPED_SPLIT=1 turns it on): when the ONE sequential op of the circuit is a biquad with
  // host-derived coefficients, its time-parallel half (a0 x + a1 x1 + a2 x2, Filter.js:40-42) leaves the S-warp for a
  // second lane-per-voice warp (the "F-warp"), which hands the f64 bracket on through a small shared-memory ring, one
  // 8-sample group at a time, with named-barrier handshakes.  The S-warp then issues nothing but the recurrence -- its
  // in-order instruction stream is the critical path of the launch (profiles/r1_v10_piped_barrier_attribution.txt).
  // The exchange buffers rotate over TWO steps instead of three (a P-warp reads a voice's stage-1 results of step
  // k-2 into registers before it writes the stage-0 values of step k into the same row), which pays for the ring.
  int fc = -1;
  {
    int n1 = 0;
    for (int i = 0; i < n; i++) if (pl.stage[i] == 1) { n1++; fc = i; }
    const char *se = getenv("DUSP_PIPED_SPLIT");
    if (n1 != 1 || p.code[fc].op != OP_FILTERC || !se || atoi(se) == 0 || STEP != 256) fc = -1;  // opt-in: measured slower
  }
  const bool split = fc >= 0;
  // Feed-forward half in the P-warps (DUSP_PIPED_FFP=1): same circuits, no extra warp.  Stage 0 of the P-warps, lanes
  // along time, turns the filter's input straight into the f64 brackets c = a0 x + a1 x1 + a2 x2 (the two preceding
  // inputs of a lane's first samples come from its neighbour by shuffle, the voice's x1 / x2 words feed lane 0) and
  // stores THOSE in the exchange row; the S-warp reads eight brackets, runs nothing but the recurrence, and overwrites
  // the row in place with its f32 results (y of samples [8g, 8g+8) lands on the bytes of brackets that are already in
  // registers).  Rows of doubles: 128-sample steps, two-step rotation, so three CTAs per SM still fit.
  int ffc = -1, ff_in = -1;
  if (want_ffp) {
    int n1 = 0;
    for (int i = 0; i < n; i++) if (pl.stage[i] == 1) { n1++; ffc = i; }
    if (n1 != 1 || p.code[ffc].op != OP_FILTERC || STEP != 128 || NB != 1 || split) return "";
    const DuspInstr &F = p.code[ffc];
    bool in_ok = false, out_ok = false;
    for (auto &c : pl.cross) {
      if (c.from == 0 && c.slot == F.in[0] && c.used_in[1] && !c.used_in[2]) { in_ok = true; ff_in = c.def; }
      if (c.def == ffc && c.to == 2) out_ok = true;
    }
    if (!in_ok || !out_ok || pl.cross.size() != 2) return "";
  }
  const bool ffp = ffc >= 0;
  const bool tworot = split || ffp;
  const int NROT = tworot ? 2 : 3;
  int RING_D = 4;  // groups of B samples in flight between the F-warp and the S-warp (DUSP_PIPED_RING overrides)
  if (const char *e = getenv("DUSP_PIPED_RING")) RING_D = std::max(2, std::min(16, atoi(e)));

  // ---- the three stage bodies
  Gen g0(p, PB, 32), g1(p, PB, 1), g2(p, PB, 32);
  Gen *gens[3] = {&g0, &g1, &g2};
  std::ostringstream s_decl, s_load, s_store;  // S-warp: crossing loads first, stores last (in-place buffer reuse)
  // stage-2 loads that depend on nothing this step computes are issued before stage 0 of the next step,
  // so their HBM / L2 latency hides behind that work
  std::ostringstream pre_decl, pre_code, prex_code;
  if (!getenv("DUSP_PIPED_NOPREFETCH"))
    for (int i = 0; i < n; i++)
      if (pl.stage[i] == 2 && Gen::can_prefetch(p, p.code[i], STEP) && p.code[i].out[0] >= 0) {
        pre_decl << "        float pre" << i << "[B" << (p.code[i].op == OP_CBREADC ? " + 3" : "") << "];\n";
        g2.prefetch(i, p.code[i], pre_code);
      }
  for (int st = 0; st < 3; st++) {
    Gen &g = *gens[st];
    for (int i = 0; i < n; i++) {
      const DuspInstr &I = p.code[i];
      if (pl.stage[i] == st) {
        g.instr(i, I);
        for (int k = 0; k < 2; k++) {
          auto it = pl.cross_of.find({i, k});
          if (it == pl.cross_of.end()) continue;
          const Cross &c = pl.cross[it->second];
          if (st == 1) {
            s_decl << "        float xo" << i << "_" << k << "[B];\n";
            g.o << "      CP(xo" << i << "_" << k << ", s" << c.slot << ")\n";
            s_store << "        STX(" << c.buf << ", xo" << i << "_" << k << ")\n";
          } else if (ffp && st == 0 && i == ff_in) {
            const int P = p.code[ffc].p;  // words: a0 a1 a2 b1 b2 (derived), x1 x2 y1 y2
            g.o << "      {  // feed-forward half of instruction " << ffc << " (Filter.js:40-42), lanes along time\n"
                << "        const float xp1_ = __shfl_up_sync(0xffffffffu, s" << c.slot << "[B - 1], 1), xp2_ = __shfl_up_sync(0xffffffffu, s" << c.slot << "[B - 2], 1);\n"
                << "        double o1_ = lane ? (double)js_or0f(xp1_) : js_or0(" << g.w(P + 5) << "), o2_ = lane ? (double)js_or0f(xp2_) : js_or0(" << g.w(P + 6) << ");\n"
                << "        double c_[B];\n"
                << "        filterc_ff<B>(s" << c.slot << ", c_, " << g.w(P) << ", " << g.w(P + 1) << ", " << g.w(P + 2) << ", o1_, o2_);\n"
                << "        " << g.w(P + 5) << " = __shfl_sync(0xffffffffu, (double)s" << c.slot << "[B - 1], 31);  // x1: the raw last input\n"
                << "        " << g.w(P + 6) << " = __shfl_sync(0xffffffffu, o2_, 31);                       // x2: normalised\n"
                << "        double2 *const row_ = reinterpret_cast<double2 *>(xb + ((size_t)rot0 * VC + vl) * ROW);\n"
                << "        UNROLL for (int q_ = 0; q_ < B / 2; q_++) row_[((B / 2) * lane + q_) ^ (vl & 7)] = make_double2(c_[2 * q_], c_[2 * q_ + 1]);\n"
                << "      }\n";
            g.mut(P + 5); g.mut(P + 6);
          } else {
            g.o << "      STX(" << c.buf << ", s" << c.slot << ")\n";
          }
        }
      } else if (pl.stage[i] < st) {
        for (int k = 0; k < 2; k++) {
          auto it = pl.cross_of.find({i, k});
          if (it == pl.cross_of.end() || !pl.cross[it->second].used_in[st]) continue;
          const Cross &c = pl.cross[it->second];
          if (st == 1) {
            s_load << "        float xi" << i << "_" << k << "[B]; LDX(xi" << i << "_" << k << ", " << c.buf << ")\n";
            g.o << "      CP(s" << c.slot << ", xi" << i << "_" << k << ")  // value of instruction " << i << " (stage " << c.from << ")\n";
          } else if (tworot) {  // read before stage 0 of this iteration reuses the row (two-step rotation)
            pre_decl << "        float prex" << i << "_" << k << "[B];\n";
            prex_code << "          LDX(prex" << i << "_" << k << ", " << c.buf << ")  // value of instruction " << i << " (stage " << c.from << ")\n";
            g.o << "      CP(s" << c.slot << ", prex" << i << "_" << k << ")\n";
          } else {
            g.o << "      LDX(s" << c.slot << ", " << c.buf << ")  // value of instruction " << i << " (stage " << c.from << ")\n";
          }
        }
      }
    }
  }
  std::set<int> p_words, p_mut, s_words, s_mut;
  for (int st : {0, 2}) { p_words.insert(gens[st]->used_words.begin(), gens[st]->used_words.end()); p_mut.insert(gens[st]->mutable_words.begin(), gens[st]->mutable_words.end()); }
  s_words = g1.used_words; s_mut = g1.mutable_words;
  std::map<int, int> pj;  // word -> row of the shared-memory word table
  for (int w : p_words) { int j = (int)pj.size(); pj[w] = j; }
  const int NPW = (int)pj.size();
  const size_t xb_bytes = (size_t)NROT * NB * VC * STEP * (ffp ? sizeof(double) : sizeof(float));
  const size_t ring_bytes = split ? (size_t)RING_D * PB * 32 * sizeof(double) + 16 : 0;  // + the two hand-over counters
  // optional dynamic voice queue for the P-warps (measured slightly slower than the static split: off)
  const bool dyn = getenv("DUSP_PIPED_DYN") != nullptr;
  int s_unroll = PB == 4 ? 2 : 1;  // S-warp groups per loop body; DUSP_PIPED_SUNROLL overrides (experiments)
  if (const char *e = getenv("DUSP_PIPED_SUNROLL")) s_unroll = std::max(1, std::min(8, atoi(e)));
  const char *srot_env = getenv("DUSP_PIPED_SROT");  // "rot" or a fixed warp index 0..NP (experiments)
  const bool srot = srot_env != nullptr;
  const int srot_fixed = srot && srot_env[0] >= '0' && srot_env[0] <= '9' ? std::min(atoi(srot_env), NP) : -1;
  const bool slot = getenv("DUSP_PIPED_SLOT") != nullptr && NP == 4;  // S-warp role by per-SM ticket (experiments)
  const bool slot_hw = slot && atoi(getenv("DUSP_PIPED_SLOT")) == 2;   // ... matched against the hardware warp slot
  const size_t smem = xb_bytes + (size_t)NPW * VC * sizeof(double) + 16 + ring_bytes;  // (+16: qctr[3] queue heads, qctr[3] = S-warp ticket)
  const int NWARPS = NP + (split ? 2 : 1);
  if (smem > 200 * 1024) return "";
  if (info) {
    info->voices_per_cta = VC; info->threads = NWARPS * 32; info->smem_bytes = (int)smem; info->n_buffers = NB;
    info->n_ops = n; info->n_ops_sequential_stage = 0;
    for (int i = 0; i < n; i++) info->n_ops_sequential_stage += pl.stage[i] == 1;
  }
  int min_blocks = (int)std::min<size_t>(4, (227 * 1024) / (smem + 1024));
  if (min_blocks < 1) min_blocks = 1;
  if (const char *e = getenv("DUSP_PIPED_MINB")) min_blocks = std::max(1, std::min(min_blocks, atoi(e)));  // experiments

  bool stage_has[3] = {false, false, false};
  for (int i = 0; i < n; i++) stage_has[pl.stage[i]] = true;

  std::ostringstream s;
  s << "// generated by libdusp_b200 (piped.cc) -- pipelined render kernel for one circuit shape\n"
    << "// stages:";
  for (int i = 0; i < n; i++) s << " " << i << ":" << op_name(p.code[i].op) << "=" << pl.stage[i];
  s << "\n" << fused_prelude()
    // exchange buffers: [rotation][buffer][voice][ROW float4]; samples [B*l + 4q, +4) of a step sit at float4
    // q*32 + (l ^ (voice & 7)): a warp reading a row (P) or a column (S) touches every bank group once
    << "#define LDX(d, b) { UNROLL for (int q_ = 0; q_ < B / 4; q_++) { const float4 t_ = xs[(b) * (VC * ROW) + q_ * 32];\\\n"
    << "    d[4*q_] = t_.x; d[4*q_+1] = t_.y; d[4*q_+2] = t_.z; d[4*q_+3] = t_.w; } }\n"
    << "#define STX(b, v) { UNROLL for (int q_ = 0; q_ < B / 4; q_++) xs[(b) * (VC * ROW) + q_ * 32] = make_float4(v[4*q_], v[4*q_+1], v[4*q_+2], v[4*q_+3]); }\n"
    // renderChannelData.js:42-44 for the P-warps: lanes along time, a warp stores 512 contiguous bytes
    << "#define OUT_CHANNEL_P(a, CH) { if (!KP_OUT_RAW) { UNROLL for (int i_ = 0; i_ < B; i_++) a[i_] = js_or0f(a[i_]); }\\\n"
    << "  if (KP_OUT_MODE == 0) { if (active) { float *dst = out_row + (long long)(CH) * kp.out_len + tl0;\\\n"
    << "      if (tl0 + B <= kp.n_samples && ((kp.out_len & 3) | (reinterpret_cast<unsigned long long>(kp.out) & 15ull)) == 0) { UNROLL for (int q_ = 0; q_ < B / 4; q_++)\\\n"
    << "        __stcs(reinterpret_cast<float4 *>(dst) + q_, make_float4(a[4*q_], a[4*q_+1], a[4*q_+2], a[4*q_+3])); }\\\n"
    << "      else { UNROLL for (int i_ = 0; i_ < B; i_++) if (tl0 + i_ < kp.n_samples) dst[i_] = a[i_]; } } }\\\n"
    << "  else if ((CH) < 2 && active) { UNROLL for (int i_ = 0; i_ < B; i_++) macc[CH][i_] += a[i_]; } }\n"
    // same for the S-warp (one voice per lane)
    << "#define OUT_CHANNEL_S(a, CH) { if (!KP_OUT_RAW) { UNROLL for (int i_ = 0; i_ < B; i_++) a[i_] = js_or0f(a[i_]); }\\\n"
    << "  if (KP_OUT_MODE == 0) { if (active) { float *dst = out_row + (long long)(CH) * kp.out_len + tl0;\\\n"
    << "      UNROLL for (int i_ = 0; i_ < B; i_++) if (tl0 + i_ < kp.n_samples) dst[i_] = a[i_]; } }\\\n"
    << "  else if ((CH) < 2 && active) { mix_add<B>(kp.out + (long long)(CH) * kp.out_len + tl0, a, kp.n_samples - tl0);\\\n"
    << "      if (kp.prog_nch == 1) mix_add<B>(kp.out + kp.out_len + tl0, a, kp.n_samples - tl0); } }\n"
    // F-warp -> S-warp hand-over of 8-sample groups of brackets: a ring in shared memory with two monotonic counters
    // (flags[0] groups produced, flags[1] groups taken).  Taking group k: wait until it is produced, load it into
    // registers, then publish the slot as free -- BEFORE the recurrence runs on it.  (Named barriers here cost the
    // S-warp ~210 cycles per group, profiles/r2_c5_split_roles.txt.)
    << "#define RING_TAKE(d, k) { const unsigned k_ = (k); while (flags[0] < k_ + 1u) {} __threadfence_block();\\\n"
    << "    UNROLL for (int i_ = 0; i_ < B; i_++) d[i_] = cring[((int)(k_ % RING_D) * B + i_) * 32 + lane];\\\n"
    << "    __syncwarp(); if (lane == 0) flags[1] = k_ + 1u; }\n"
    << "extern \"C\" __global__ void " << (getenv("DUSP_PIPED_MAXNREG") ? "__maxnreg__(" + std::string(getenv("DUSP_PIPED_MAXNREG")) + ")"
                                             : "__launch_bounds__(" + std::to_string(NWARPS * 32) + ", " + std::to_string(min_blocks) + ")")
    << " dusp_piped(const __grid_constant__ FusedParams kp) {\n"
    << "  constexpr int B = " << PB << ", VC = " << VC << ", NB = " << NB << ", STEP = " << STEP << ", ROW = STEP / " << (ffp ? 2 : 4) << ", NP = " << NP << ", VW = VC / NP, NROT = " << NROT << ", RING_D = " << RING_D << ";\n"
    << "  extern __shared__ float4 smem4[];\n"
    << "  float4 *const xb = smem4;                                                    // [NROT][NB][VC][ROW] float4, swizzled\n"
    << "  double *const pw = reinterpret_cast<double *>(smem4 + NROT * NB * VC * ROW);  // [" << NPW << "][VC] words of the P-warps' units\n"
    << "  int *const qctr = reinterpret_cast<int *>(pw + " << NPW << " * VC);                 // [3] per-step voice queue heads\n"
    << "  double *const cring = reinterpret_cast<double *>(qctr + 4); (void)cring;            // [RING_D][B][32] brackets on their way to the S-warp\n"
    << "  volatile unsigned *const flags = reinterpret_cast<volatile unsigned *>(cring + RING_D * B * 32 * " << (split ? 1 : 0) << "); (void)flags;\n"
    << (split ? "  if (threadIdx.x < 2) flags[threadIdx.x] = 0u;\n" : "")
    << "  if (threadIdx.x < 3) qctr[threadIdx.x] = 0;\n"
    << "  const int lane = threadIdx.x & 31, wraw = threadIdx.x >> 5;\n"
    // Which warp of the CTA is the lane-per-voice warp.  Warps go to the SM's four schedulers by warp index, and the
    // CTAs resident on one SM are blockIdx, blockIdx + #SMs, ...: rotating the role by blockIdx / #SMs puts the
    // (latency-critical) S-warps of co-resident CTAs on different schedulers.  wib: P-warps 0..NP-1, S-warp NP.
    << (slot ? std::string(
                   // co-resident CTAs take consecutive tickets of their SM's counter: their lane-per-voice warps get different
                   // warp indices 0..3, i.e. different SM sub-partitions, whatever the block scheduler's placement was
                   "  if (threadIdx.x == 0) { unsigned smid_; asm volatile(\"mov.u32 %0, %%smid;\" : \"=r\"(smid_));\n"
                   "    qctr[3] = (int)(atomicAdd(kp.sm_slots + (smid_ & 255u), 1u) & 3u); }\n"
                   "  __syncthreads();\n"
                   + std::string(slot_hw ?
                   // by hardware warp slot: the first warp of the CTA whose %warpid % 4 equals the ticket
                   "  __shared__ int hw_of_[8];\n"
                   "  if (lane == 0) { unsigned wid_; asm volatile(\"mov.u32 %0, %%warpid;\" : \"=r\"(wid_)); hw_of_[wraw] = (int)(wid_ & 3u); }\n"
                   "  __syncthreads();\n"
                   "  int s_raw = NP;\n"
                   "  for (int w_ = NP; w_ >= 0; w_--) if (hw_of_[w_] == qctr[3]) s_raw = w_;\n"
                   : "  const int s_raw = qctr[3];\n") +
                   "  const int wib = wraw == s_raw ? NP : wraw - (wraw > s_raw ? 1 : 0);\n")
        : srot ? std::string("  const int s_raw = ") + (srot_fixed >= 0 ? std::to_string(srot_fixed) : std::string("(blockIdx.x / 148) % (NP + 1)")) + ";\n"
               "  const int wib = wraw == s_raw ? NP : wraw - (wraw > s_raw ? 1 : 0);\n"
             : std::string("  const int wib = wraw;\n"))
    << "  const long long v0 = (long long)blockIdx.x * VC;  // allocations are padded to a multiple of 96 voices\n"
    << "  const size_t nv = (size_t)kp.nv_stride;\n"
    << "  const float *const tables = kp.tables; (void)tables; (void)xb; (void)pw; (void)qctr;\n"
    << "  const long long n_proc = (kp.n_samples + DUSP_CHUNK - 1) / DUSP_CHUNK * DUSP_CHUNK;  // whole chunks, like Circuit.tick\n"
    << "  const int n_steps = (int)(n_proc / STEP);\n";
  for (auto &kv : pj)
    s << "  if (wib == " << kv.second % (NP + 1) << " && lane < VC) pw[" << kv.second << " * VC + lane] = kp.words[" << kv.first << " * nv + v0 + lane];  // "
      << p.words[kv.first].name << "\n";
  s << "  __syncthreads();\n"
    << "  if (wib < NP) {\n"
    << "    // ---------------- P-warps: lanes along time, VW voices each\n"
    << "    constexpr int T = 32;\n"
    << "    Cx<B, T> cx; cx.tl = lane; cx.lane = lane;\n"
    << "    _Pragma(\"unroll 1\")\n"
    << "    for (int it = 0; it < n_steps + 2; it++) {\n"
    << "      const bool do0 = it < n_steps, do2 = it >= 2;\n"
    << (tworot ? "      const int rot0 = it & 1, rot2 = it & 1;  // two-step rotation: stage 2 of step it-2 frees the row stage 0 of step it fills\n"
              : "      const int rot0 = it % 3, rot2 = (it + 1) % 3;  // (it - 2) mod 3\n")
    << "      float macc[2][B]; UNROLL for (int i_ = 0; i_ < B; i_++) { macc[0][i_] = 0.0f; macc[1][i_] = 0.0f; }\n"
    << (dyn ? "      if (threadIdx.x == 0) qctr[(it + 1) % 3] = 0;  // next step's queue (last used two steps ago)\n" : "")
    << "      _Pragma(\"unroll 1\")\n"
    << (dyn ? "      for (;;) {  // the CTA's VC voices are a queue: whichever P-warp is free takes the next one\n"
              "        int vl = 0;\n"
              "        if (lane == 0) vl = atomicAdd(&qctr[it % 3], 1);\n"
              "        vl = __shfl_sync(0xffffffffu, vl, 0);\n"
              "        if (vl >= VC) break;\n"
            : "      for (int j = 0; j < VW; j++) {\n"
              "        const int vl = wib * VW + j;\n")
    << "        const long long v = v0 + vl;\n"
    << "        const bool active = v < kp.n_voices;\n"
    << "        float *const rings = kp.rings + v * kp.ring_stride;\n"
    << "        float *const stale = kp.stale + v * (long long)kp.n_stale * DUSP_CHUNK;\n"
    << "        float *const out_row = kp.out + (long long)(kp.voice0 + v) * kp.out_nch * kp.out_len;\n"
    << "        (void)rings; (void)stale; (void)out_row; (void)active;\n";
  for (auto &kv : pj) s << "        double w" << kv.first << " = pw[" << kv.second << " * VC + vl];\n";
  if (!pre_code.str().empty() || !prex_code.str().empty())
    s << pre_decl.str()
      << "        if (do2) {  // ---- stage 2, early loads\n"
      << "          const int cpos = (int)(((long long)(it - 2) * STEP + lane * B) & (DUSP_CHUNK - 1)); (void)cpos;\n"
      << "          float4 *const xs = xb + ((size_t)rot2 * NB * VC + vl) * ROW + (lane ^ (vl & 7)); (void)xs;\n"
      << pre_code.str() << prex_code.str() << "        }\n";
  s << "#define OUT_CHANNEL OUT_CHANNEL_P\n";
  const char *rot[3] = {"rot0", "", "rot2"};
  const char *cond[3] = {"do0", "", "do2"};
  const char *when[3] = {"it", "", "(it - 2)"};
  for (int st : {0, 2}) {
    if (!stage_has[st]) continue;
    s << "        if (" << cond[st] << ") {  // ---- stage " << st << "\n"
      << "          const long long tl0 = (long long)" << when[st] << " * STEP + lane * B;\n"
      << "          const int cpos = (int)(tl0 & (DUSP_CHUNK - 1)); (void)cpos;\n"
      << "          float4 *const xs = xb + ((size_t)" << rot[st] << " * NB * VC + vl) * ROW + (lane ^ (vl & 7)); (void)xs;\n";
    for (int k = 0; k < p.n_slots; k++) s << "          float s" << k << "[B];\n";
    s << gens[st]->o.str() << "        }\n";
  }
  s << "#undef OUT_CHANNEL\n";
  if (!p_mut.empty()) {
    s << "        __syncwarp();\n        if (lane == 0) {\n";
    for (int w : p_mut) s << "          pw[" << pj[w] << " * VC + vl] = w" << w << ";\n";
    s << "        }\n        __syncwarp();\n";
  }
  s << "      }\n"
    << "      if (do2 && KP_OUT_MODE != 0) {  // stereo mixdown of this warp's voices: one atomic per sample and channel\n"
    << "        const long long t2 = (long long)(it - 2) * STEP + lane * B;\n"
    << "        mix_add<B>(kp.out + t2, macc[0], kp.n_samples - t2);\n"
    << "        if (kp.prog_nch == 1) mix_add<B>(kp.out + kp.out_len + t2, macc[0], kp.n_samples - t2);\n"
    << "        else mix_add<B>(kp.out + kp.out_len + t2, macc[1], kp.n_samples - t2);\n"
    << "      }\n"
    << "      __syncthreads();\n"
    << "    }\n";
  if (ffp) {
    const int P = p.code[ffc].p;
    auto W = [&](int k) { return "kp.words[" + std::to_string(P + k) + " * nv + v]"; };
    const char *magic = getenv("DUSP_PIPED_MAGIC");
    const bool use_magic = magic && atoi(magic) != 0;
    s << "  } else {\n"
      << "    // ---------------- S-warp: one voice per lane, nothing but the recurrence  y = c - b1 y1 - b2 y2  (Filter.js:40-48)\n"
      << "    const int vl = lane < VC ? lane : VC - 1;\n"
      << "    const long long v = v0 + vl;\n"
      << "    const double b1 = " << W(3) << ", b2 = " << W(4) << ";\n"
      << "    double q1 = js_or0(" << W(7) << "), q2 = js_or0(" << W(8) << ");\n"
      << "    float ylast = (float)" << W(7) << ";\n"
      << "    _Pragma(\"unroll 1\")\n"
      << "    for (int it = 0; it < n_steps + 2; it++) {\n"
      << "      if (it >= 1 && it <= n_steps && lane < VC) {\n"
      << "        float4 *const row = xb + ((size_t)((it - 1) & 1) * VC + vl) * ROW;\n"
      << "        const double2 *const rowd = reinterpret_cast<const double2 *>(row);\n"
      << "        const int k_ = vl & 7;\n"
      << "        _Pragma(\"unroll 1\")\n"
      << "        for (int g = 0; g < STEP / 8; g++) {\n"
      << "          double c[8];\n"
      << "          UNROLL for (int q_ = 0; q_ < 4; q_++) { const double2 t_ = rowd[(4 * g + q_) ^ k_]; c[2 * q_] = t_.x; c[2 * q_ + 1] = t_.y; }\n"
      << "          float y[8];\n"
      << "          filterc_rec<8, " << (use_magic ? "true" : "false") << ">(c, y, b1, b2, q1, q2);\n"
      << "          ylast = y[7];\n"
      << "          row[(2 * g) ^ k_] = make_float4(y[0], y[1], y[2], y[3]);\n"
      << "          row[(2 * g + 1) ^ k_] = make_float4(y[4], y[5], y[6], y[7]);\n"
      << "        }\n"
      << "      }\n"
      << "      __syncthreads();\n"
      << "    }\n"
      << "    if (lane < VC) { " << W(7) << " = (double)ylast; " << W(8) << " = q2; }\n"
      << "  }\n";
  } else if (split) {
    // ---- S-warp: the recurrence only; F-warp: the bracket.  Both one voice per lane, state in registers for the launch.
    const DuspInstr &F = p.code[fc];
    const int P = F.p;  // words: a0 a1 a2 b1 b2 (derived), x1 x2 y1 y2
    int in_buf = -1, out_buf = -1;
    for (auto &c : pl.cross) {
      if (c.to >= 1 && c.from == 0 && c.slot == F.in[0] && c.used_in[1]) in_buf = c.buf;
      if (c.def == fc) out_buf = c.buf;
    }
    const char *magic = getenv("DUSP_PIPED_MAGIC");
    const bool use_magic = !magic || atoi(magic) != 0;
    auto W = [&](int k) { return "kp.words[" + std::to_string(P + k) + " * nv + v]"; };
    s << "  } else if (wib == NP) {\n"
      << "    // ---------------- S-warp: one voice per lane, nothing but the recurrence  y = c - b1 y1 - b2 y2  (Filter.js:40-48)\n"
      << "    const int vl = lane < VC ? lane : VC - 1;\n"
      << "    const long long v = v0 + vl;\n"
      << "    const double b1 = " << W(3) << ", b2 = " << W(4) << ";\n"
      << "    double q1 = js_or0(" << W(7) << "), q2 = js_or0(" << W(8) << ");\n"
      << "    float ylast = (float)" << W(7) << ";\n"
      << "    unsigned G = 0;  // groups consumed so far: ring slot G % RING_D\n"
      << "    _Pragma(\"unroll 1\")\n"
      << "    for (int it = 0; it < n_steps + 2; it++) {\n"
      << "      if (it >= 1 && it <= n_steps) {\n"
      << "        const int rot1 = (it - 1) & 1;\n"
      << "        double c[B], cn[B];\n"
      << "        RING_TAKE(c, G)  // the step's first group\n"
      << "        _Pragma(\"unroll 2\")\n"
      << "        for (int g = 0; g < STEP / B; g++, G++) {\n"
      << "          if (g + 1 < STEP / B) RING_TAKE(cn, G + 1)  // the next group's brackets: in flight while this group's chain runs\n"
      << "          float y[B];\n"
      << "          filterc_rec<B, " << (use_magic ? "true" : "false") << ">(c, y, b1, b2, q1, q2);\n"
      << "          UNROLL for (int i_ = 0; i_ < B; i_++) c[i_] = cn[i_];\n"
      << "          ylast = y[B - 1];\n";
    if (out_buf >= 0)
      s << "          if (lane < VC) { float4 *const xs = xb + ((size_t)rot1 * NB * VC + vl) * ROW + (g ^ (vl & 7)); STX(" << out_buf << ", y) }\n";
    s << "        }\n"
      << "      }\n"
      << "      __syncthreads();\n"
      << "    }\n"
      << "    if (lane < VC) { " << W(7) << " = (double)ylast; " << W(8) << " = q2; }\n"
      << "  } else {\n"
      << "    // ---------------- F-warp: one voice per lane, the bracket  c = a0 x + a1 x1 + a2 x2  (Filter.js:40-42)\n"
      << "    const int vl = lane < VC ? lane : VC - 1;\n"
      << "    const long long v = v0 + vl;\n"
      << "    const double a0 = " << W(0) << ", a1 = " << W(1) << ", a2 = " << W(2) << ";\n"
      << "    double o1 = js_or0(" << W(5) << "), o2 = js_or0(" << W(6) << ");\n"
      << "    double xlast = " << W(5) << ";\n";
    if (in_buf < 0 && F.in[0] != REF_NONE && F.in[0] < 0)
      s << "    const float xconst = (float)kp.words[" << (REF_CONST_BASE - F.in[0]) << " * nv + v];\n";
    s << "    unsigned G = 0;\n"
      << "    _Pragma(\"unroll 1\")\n"
      << "    for (int it = 0; it < n_steps + 2; it++) {\n"
      << "      if (it >= 1 && it <= n_steps) {\n"
      << "        const int rot1 = (it - 1) & 1; (void)rot1;\n"
      << "        _Pragma(\"unroll 1\")\n"
      << "        for (int g = 0; g < STEP / B; g++, G++) {\n"
      << "          const int slot = (int)(G % RING_D);\n"
      << "          float x[B];\n";
    if (in_buf >= 0)
      s << "          { float4 *const xs = xb + ((size_t)rot1 * NB * VC + vl) * ROW + (g ^ (vl & 7)); LDX(x, " << in_buf << ") }\n";
    else if (F.in[0] == REF_NONE)
      s << "          FILL(x, 0.0f)\n";
    else
      s << "          FILL(x, xconst)\n";
    s << "          double c[B];\n"
      << "          xlast = (double)x[B - 1];\n"
      << "          filterc_ff<B>(x, c, a0, a1, a2, o1, o2);\n"
      << "          if (G >= RING_D) { while (flags[1] < G - RING_D + 1) {} }  // the S-warp has taken what the slot held\n"
      << "          UNROLL for (int i_ = 0; i_ < B; i_++) cring[(slot * B + i_) * 32 + lane] = c[i_];\n"
      << "          __syncwarp();\n"
      << "          if (lane == 0) { __threadfence_block(); flags[0] = G + 1; }\n"
      << "        }\n"
      << "      }\n"
      << "      __syncthreads();\n"
      << "    }\n"
      << "    if (lane < VC) { " << W(5) << " = xlast; " << W(6) << " = o2; }\n"
      << "  }\n";
  } else {
  s << "  } else {\n"
    << "    // ---------------- S-warp: one voice per lane, the reference's sequential order\n"
    << "    constexpr int T = 1;\n"
    << "    Cx<B, T> cx; cx.tl = 0; cx.lane = lane;\n"
    << "    const int vl = lane < VC ? lane : VC - 1;\n"
    << "    const long long v = v0 + vl;\n"
    << "    const bool active = lane < VC && v < kp.n_voices;\n"
    << "    float *const rings = kp.rings + v * kp.ring_stride;\n"
    << "    float *const stale = kp.stale + v * (long long)kp.n_stale * DUSP_CHUNK;\n"
    << "    float *const out_row = kp.out + (long long)(kp.voice0 + v) * kp.out_nch * kp.out_len;\n"
    << "    (void)rings; (void)stale; (void)out_row; (void)active;\n";
  for (int w : s_words) s << "    double w" << w << " = kp.words[" << w << " * nv + v];  // " << p.words[w].name << "\n";
  s << "#define OUT_CHANNEL OUT_CHANNEL_S\n"
    << "    _Pragma(\"unroll 1\")\n"
    << "    for (int it = 0; it < n_steps + 2; it++) {\n";
  if (stage_has[1]) {
    s << "      if (it >= 1 && it <= n_steps && lane < VC) {\n"
      << "        const int rot1 = (it + 2) % 3;  // (it - 1) mod 3\n"
      << "        _Pragma(\"unroll " << s_unroll << "\")\n"
      << "        for (int g = 0; g < STEP / B; g++) {\n"
      << "          const long long tl0 = (long long)(it - 1) * STEP + g * B;\n"
      << "          const int cpos = (int)(tl0 & (DUSP_CHUNK - 1)); (void)cpos;\n"
      << "          float4 *const xs = xb + ((size_t)rot1 * NB * VC + vl) * ROW + (g ^ (vl & 7)); (void)xs;\n";
    for (int k = 0; k < p.n_slots; k++) s << "          float s" << k << "[B];\n";
    s << s_decl.str() << s_load.str() << g1.o.str() << s_store.str() << "        }\n      }\n";
  }
  s << "      __syncthreads();\n"
    << "    }\n"
    << "#undef OUT_CHANNEL\n";
  if (!s_mut.empty()) {
    s << "    if (lane < VC) {\n";
    for (int w : s_mut) s << "      kp.words[" << w << " * nv + v] = w" << w << ";\n";
    s << "    }\n";
  }
  s << "  }\n";
  }  // !split
  if (!p_mut.empty()) {
    s << "  __syncthreads();\n";
    for (int w : p_mut)
      s << "  if (wib == " << pj[w] % (NP + 1) << " && lane < VC) kp.words[" << w << " * nv + v0 + lane] = pw[" << pj[w] << " * VC + lane];\n";
  }
  s << "}\n";
  return s.str();
}


// Quarter-step form (DUSP_PIPED_FFP=2; same circuits as the ffp form above).  The feed-forward half of the biquad runs in
// the P-warps and the S-warp issues nothing but the recurrence, as in ffp -- but only stage 0 pays for the f64 brackets'
// size: it runs on 64-sample quarters (two voices per pass, 16 lanes along time each, 4 samples per lane) into two
// quarter rows of brackets per voice, the S-warp follows one quarter behind and writes its f32 results into the y rows
// of the step (the layout stage 2 reads), and stage 2 keeps whole 256-sample steps with 8 samples per lane -- where the
// P-warps' per-step costs are (ring index arithmetic, shifted ring access, stale buffers, output).  Stage 2 of a step
// is spread over the four quarters after its last recurrence, one of a warp's four voices per quarter (VW = 4, so
// VC = 24 means six P-warps).  Shared memory: y rows [2 steps][VC][1 KB] + bracket rows [2 quarters][VC][512 B] + words:
// 76.6 KB at VC = 24, three CTAs per SM like the default kernel.
std::string generate_ffq(const Program &p, int VC, PipedInfo *info) {
  Plan pl;
  if (!make_plan(p, pl)) return "";
  const int n = (int)p.code.size();
  if (pl.n_buffers != 1 || VC % 4 != 0 || VC / 4 > 7) return "";
  const int NP = VC / 4;
  int ffc = -1, ff_in = -1, n1 = 0;
  for (int i = 0; i < n; i++) if (pl.stage[i] == 1) { n1++; ffc = i; }
  if (n1 != 1 || p.code[ffc].op != OP_FILTERC) return "";
  {
    const DuspInstr &F = p.code[ffc];
    bool in_ok = false, out_ok = false;
    for (auto &c : pl.cross) {
      if (c.from == 0 && c.slot == F.in[0] && c.used_in[1] && !c.used_in[2]) { in_ok = true; ff_in = c.def; }
      if (c.def == ffc && c.to == 2) out_ok = true;
    }
    if (!in_ok || !out_ok || pl.cross.size() != 2) return "";
  }
  const int P = p.code[ffc].p;  // words: a0 a1 a2 b1 b2 (derived), x1 x2 y1 y2
  Gen g0(p, 4, 16), g2(p, 8, 32);
  std::ostringstream pre_decl, pre_code;
  if (!getenv("DUSP_PIPED_NOPREFETCH"))
    for (int i = 0; i < n; i++)
      if (pl.stage[i] == 2 && Gen::can_prefetch(p, p.code[i], 256) && p.code[i].out[0] >= 0) {
        pre_decl << "        float pre" << i << "[B" << (p.code[i].op == OP_CBREADC ? " + 3" : "") << "];\n";
        g2.prefetch(i, p.code[i], pre_code);
      }
  for (int i = 0; i < n; i++) {
    const DuspInstr &I = p.code[i];
    if (pl.stage[i] == 0) {
      g0.instr(i, I);
      if (i == ff_in) {
        int slot = -1;
        for (int k = 0; k < 2; k++) { auto it = pl.cross_of.find({i, k}); if (it != pl.cross_of.end()) slot = pl.cross[it->second].slot; }
        g0.o << "      {  // feed-forward half of instruction " << ffc << " (Filter.js:40-42), 16 lanes along time per voice\n"
             << "        const float xp1_ = __shfl_up_sync(0xffffffffu, s" << slot << "[B - 1], 1, 16), xp2_ = __shfl_up_sync(0xffffffffu, s" << slot << "[B - 2], 1, 16);\n"
             << "        double o1_ = cx.tl ? (double)js_or0f(xp1_) : js_or0(" << g0.w(P + 5) << "), o2_ = cx.tl ? (double)js_or0f(xp2_) : js_or0(" << g0.w(P + 6) << ");\n"
             << "        double c_[B];\n"
             << "        filterc_ff<B>(s" << slot << ", c_, " << g0.w(P) << ", " << g0.w(P + 1) << ", " << g0.w(P + 2) << ", o1_, o2_);\n"
             << "        " << g0.w(P + 5) << " = __shfl_sync(0xffffffffu, (double)s" << slot << "[B - 1], (lane & 16) | 15);  // x1: the raw last input\n"
             << "        " << g0.w(P + 6) << " = __shfl_sync(0xffffffffu, o2_, (lane & 16) | 15);                         // x2: normalised\n"
             << "        double2 *const row_ = reinterpret_cast<double2 *>(bk + ((size_t)(Q & 1) * VC + vl) * 32);\n"
             << "        UNROLL for (int q_ = 0; q_ < B / 2; q_++) row_[((B / 2) * cx.tl + q_) ^ (vl & 7)] = make_double2(c_[2 * q_], c_[2 * q_ + 1]);\n"
             << "      }\n";
        g0.mut(P + 5); g0.mut(P + 6);
      }
    } else if (pl.stage[i] == 2) {
      g2.instr(i, I);
    }
    if (pl.stage[i] == 1) {  // its value reaches stage 2 through the y rows
      for (int k = 0; k < 2; k++) {
        auto it = pl.cross_of.find({i, k});
        if (it == pl.cross_of.end()) continue;
        const Cross &c = pl.cross[it->second];
        g2.o << "      LDX(s" << c.slot << ", " << c.buf << ")  // value of instruction " << i << " (stage 1)\n";
      }
    }
  }
  // (stage-2 code must see the crossing load BEFORE its consumers: instructions are emitted in program order and the
  //  stage-1 op precedes every stage-2 consumer of its value, so the LDX above lands in front of them)
  std::set<int> words0 = g0.used_words, words2 = g2.used_words, mut0 = g0.mutable_words, mut2 = g2.mutable_words, all_words, all_mut;
  for (int w : words0) if (words2.count(w)) return "";  // a unit on both sides of the recurrence: not this form
  all_words.insert(words0.begin(), words0.end()); all_words.insert(words2.begin(), words2.end());
  all_mut.insert(mut0.begin(), mut0.end()); all_mut.insert(mut2.begin(), mut2.end());
  std::map<int, int> pj;
  for (int w : all_words) { int j = (int)pj.size(); pj[w] = j; }
  const int NPW = (int)pj.size();
  const size_t smem = (size_t)2 * VC * 1024 + (size_t)2 * VC * 512 + (size_t)NPW * VC * sizeof(double);
  if (smem > 200 * 1024) return "";
  const int NWARPS = NP + 1;
  if (info) {
    info->voices_per_cta = VC; info->threads = NWARPS * 32; info->smem_bytes = (int)smem; info->n_buffers = 1;
    info->n_ops = n; info->n_ops_sequential_stage = 1;
  }
  int min_blocks = (int)std::min<size_t>(4, (228 * 1024) / (smem + 1024));
  if (min_blocks < 1) min_blocks = 1;
  if (const char *e = getenv("DUSP_PIPED_MINB")) min_blocks = std::max(1, std::min(min_blocks, atoi(e)));
  auto W = [&](int k) { return "kp.words[" + std::to_string(P + k) + " * nv + v]"; };

  std::ostringstream s;
  s << "// generated by libdusp_b200 (piped.cc, quarter-step form) -- pipelined render kernel for one circuit shape\n// stages:";
  for (int i = 0; i < n; i++) s << " " << i << ":" << op_name(p.code[i].op) << "=" << pl.stage[i];
  s << "\n" << fused_prelude()
    << "#define LDX(d, b) { UNROLL for (int q_ = 0; q_ < B / 4; q_++) { const float4 t_ = xs[q_ * 32];\\\n"
    << "    d[4*q_] = t_.x; d[4*q_+1] = t_.y; d[4*q_+2] = t_.z; d[4*q_+3] = t_.w; } }\n"
    << "#define OUT_CHANNEL(a, CH) { if (!KP_OUT_RAW) { UNROLL for (int i_ = 0; i_ < B; i_++) a[i_] = js_or0f(a[i_]); }\\\n"
    << "  if (KP_OUT_MODE == 0) { if (active) { float *dst = out_row + (long long)(CH) * kp.out_len + tl0;\\\n"
    << "      if (tl0 + B <= kp.n_samples && ((kp.out_len & 3) | (reinterpret_cast<unsigned long long>(kp.out) & 15ull)) == 0) { UNROLL for (int q_ = 0; q_ < B / 4; q_++)\\\n"
    << "        __stcs(reinterpret_cast<float4 *>(dst) + q_, make_float4(a[4*q_], a[4*q_+1], a[4*q_+2], a[4*q_+3])); }\\\n"
    << "      else { UNROLL for (int i_ = 0; i_ < B; i_++) if (tl0 + i_ < kp.n_samples) dst[i_] = a[i_]; } } }\\\n"
    << "  else if ((CH) < 2 && active) { UNROLL for (int i_ = 0; i_ < B; i_++) macc[CH][i_] += a[i_]; } }\n"
    << "extern \"C\" __global__ void __launch_bounds__(" << NWARPS * 32 << ", " << min_blocks << ") dusp_piped(const __grid_constant__ FusedParams kp) {\n"
    << "  constexpr int VC = " << VC << ", NP = " << NP << ", VW = 4, STEP = 256, ROW = 64;\n"
    << "  extern __shared__ float4 smem4[];\n"
    << "  float4 *const yb = smem4;                    // [2 steps][VC][64] float4: the recurrences' results, stage 2's layout (swizzled)\n"
    << "  float4 *const bk = yb + 2 * VC * ROW;        // [2 quarters][VC][32] float4: f64 brackets of 64 samples (swizzled)\n"
    << "  double *const pw = reinterpret_cast<double *>(bk + 2 * VC * 32);  // [" << NPW << "][VC] words of the P-warps' units\n"
    << "  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;\n"
    << "  const long long v0 = (long long)blockIdx.x * VC;  // allocations are padded to a multiple of 96 voices\n"
    << "  const size_t nv = (size_t)kp.nv_stride;\n"
    << "  const float *const tables = kp.tables; (void)tables;\n"
    << "  const long long n_proc = (kp.n_samples + DUSP_CHUNK - 1) / DUSP_CHUNK * DUSP_CHUNK;  // whole chunks, like Circuit.tick\n"
    << "  const int NQ = (int)(n_proc / 64);  // quarters; the loop runs 5 more: recurrence one behind, stage 2 five behind\n";
  for (auto &kv : pj)
    s << "  if (wib == " << kv.second % (NP + 1) << " && lane < VC) pw[" << kv.second << " * VC + lane] = kp.words[" << kv.first << " * nv + v0 + lane];  // "
      << p.words[kv.first].name << "\n";
  s << "  __syncthreads();\n"
    << "  if (wib < NP) {\n"
    << "    // ---------------- P-warps\n"
    << "    float macc[2][8]; UNROLL for (int i_ = 0; i_ < 8; i_++) { macc[0][i_] = 0.0f; macc[1][i_] = 0.0f; }\n"
    << "    constexpr int B = 8, T = 32;\n"
    << pre_decl.str()   // stage 2's early loads are issued ONE QUARTER AHEAD (end of the previous pass) and live across the barrier
    << "    _Pragma(\"unroll 1\")\n"
    << "    for (int Q = 0; Q < NQ + 5; Q++) {\n"
    << "      const bool do0 = Q < NQ, do2 = Q >= 5;\n"
    << "      const int stp2_ = (Q - 5) >> 2;  // the step stage 2 works on, one voice of this warp per quarter\n"
    << "      {\n"
    << "        Cx<B, T> cx; cx.tl = lane; cx.lane = lane;\n"
    << "        const int vl = wib * VW + ((Q - 5) & 3);\n"
    << "        const long long v = v0 + vl;\n"
    << "        const bool active = v < kp.n_voices;\n"
    << "        float *const rings = kp.rings + v * kp.ring_stride;\n"
    << "        float *const stale = kp.stale + v * (long long)kp.n_stale * DUSP_CHUNK;\n"
    << "        float *const out_row = kp.out + (long long)(kp.voice0 + v) * kp.out_nch * kp.out_len;\n"
    << "        (void)rings; (void)stale; (void)out_row; (void)active; (void)cx;\n"
    << "        const long long tl0 = (long long)stp2_ * STEP + lane * B;\n"
    << "        const int cpos = (int)(tl0 & (DUSP_CHUNK - 1)); (void)cpos;\n"
    << "        float4 *const xs = yb + ((size_t)(stp2_ & 1) * VC + vl) * ROW + (lane ^ (vl & 7)); (void)xs;\n";
  for (int w : words2) s << "        double w" << w << " = pw[" << pj[w] << " * VC + vl];\n";
  s << "        if (do2) {  // ---- stage 2 of step stp2_, voice vl (its early loads were issued in the previous pass)\n";
  for (int k = 0; k < p.n_slots; k++) s << "          float s" << k << "[B];\n";
  s << g2.o.str();
  if (!mut2.empty()) {
    s << "          __syncwarp();\n          if (lane == 0) {\n";
    for (int w : mut2) s << "            pw[" << pj[w] << " * VC + vl] = w" << w << ";\n";
    s << "          }\n          __syncwarp();\n";
  }
  s << "          if (KP_OUT_MODE != 0 && ((Q - 5) & 3) == 3) {  // stereo mixdown of this warp's four voices of step stp2_\n"
    << "            const long long t2 = (long long)stp2_ * STEP + lane * B;\n"
    << "            mix_add<B>(kp.out + t2, macc[0], kp.n_samples - t2);\n"
    << "            if (kp.prog_nch == 1) mix_add<B>(kp.out + kp.out_len + t2, macc[0], kp.n_samples - t2);\n"
    << "            else mix_add<B>(kp.out + kp.out_len + t2, macc[1], kp.n_samples - t2);\n"
    << "            UNROLL for (int i_ = 0; i_ < B; i_++) { macc[0][i_] = 0.0f; macc[1][i_] = 0.0f; }\n"
    << "          }\n"
    << "        }\n"
    << "        if (Q >= 4 && Q < NQ + 4) {  // ---- early loads of the NEXT pass's stage 2 (voice (Q - 4) & 3, step (Q - 4) >> 2):\n"
    << "          // its stale buffers and ring positions were last written four passes ago\n"
    << "          const int vl = wib * VW + ((Q - 4) & 3);\n"
    << "          const long long v = v0 + vl;\n"
    << "          float *const rings = kp.rings + v * kp.ring_stride;\n"
    << "          float *const stale = kp.stale + v * (long long)kp.n_stale * DUSP_CHUNK;\n"
    << "          const int cpos = (int)(((long long)((Q - 4) >> 2) * STEP + lane * B) & (DUSP_CHUNK - 1));\n"
    << "          (void)rings; (void)stale; (void)cpos;\n";
  for (int w : words2) s << "          const double w" << w << " = pw[" << pj[w] << " * VC + vl]; (void)w" << w << ";\n";
  s << pre_code.str()
    << "        }\n"
    << "        if (do0) {  // ---- stage 0 of quarter Q: this warp's four voices, two per pass\n"
    << "          _Pragma(\"unroll 1\")\n"
    << "          for (int ps_ = 0; ps_ < 2; ps_++) {\n"
    << "            constexpr int B = 4, T = 16;\n"
    << "            Cx<B, T> cx; cx.tl = lane & 15; cx.lane = lane;\n"
    << "            const int vl = wib * VW + ps_ * 2 + (lane >> 4);\n"
    << "            const long long v = v0 + vl;\n"
    << "            const bool active = v < kp.n_voices;\n"
    << "            float *const rings = kp.rings + v * kp.ring_stride;\n"
    << "            float *const stale = kp.stale + v * (long long)kp.n_stale * DUSP_CHUNK;\n"
    << "            (void)rings; (void)stale; (void)active;\n"
    << "            const long long tl0 = (long long)Q * 64 + cx.tl * B;\n"
    << "            const int cpos = (int)(tl0 & (DUSP_CHUNK - 1)); (void)cpos;\n";
  for (int w : words0) s << "            double w" << w << " = pw[" << pj[w] << " * VC + vl];\n";
  for (int k = 0; k < p.n_slots; k++) s << "            float s" << k << "[B];\n";
  s << g0.o.str();
  if (!mut0.empty()) {
    s << "            __syncwarp();\n            if (cx.tl == 0) {\n";
    for (int w : mut0) s << "              pw[" << pj[w] << " * VC + vl] = w" << w << ";\n";
    s << "            }\n            __syncwarp();\n";
  }
  s << "          }\n"
    << "        }\n"
;
  s     << "      }\n"
    << "      __syncthreads();\n"
    << "    }\n"
    << "  } else {\n"
    << "    // ---------------- S-warp: one voice per lane, nothing but the recurrence  y = c - b1 y1 - b2 y2  (Filter.js:40-48)\n"
    << "    const int vl = lane < VC ? lane : VC - 1;\n"
    << "    const long long v = v0 + vl;\n"
    << "    const double b1 = " << W(3) << ", b2 = " << W(4) << ";\n"
    << "    double q1 = js_or0(" << W(7) << "), q2 = js_or0(" << W(8) << ");\n"
    << "    float ylast = (float)" << W(7) << ";\n"
    << "    const int k_ = vl & 7;\n"
    << "    _Pragma(\"unroll 1\")\n"
    << "    for (int Q = 0; Q < NQ + 5; Q++) {\n"
    << "      if (Q >= 1 && Q <= NQ && lane < VC) {\n"
    << "        const int qs = Q - 1;  // the quarter the P-warps filled in the previous pass\n"
    << "        const double2 *const rowd = reinterpret_cast<const double2 *>(bk + ((size_t)(qs & 1) * VC + vl) * 32);\n"
    << "        float4 *const yrow = yb + ((size_t)((qs >> 2) & 1) * VC + vl) * ROW;\n"
    << "        _Pragma(\"unroll 1\")\n"
    << "        for (int g = 0; g < 8; g++) {\n"
    << "          double c[8];\n"
    << "          UNROLL for (int q_ = 0; q_ < 4; q_++) { const double2 t_ = rowd[(4 * g + q_) ^ k_]; c[2 * q_] = t_.x; c[2 * q_ + 1] = t_.y; }\n"
    << "          float y[8];\n"
    << "          filterc_rec<8, false>(c, y, b1, b2, q1, q2);\n"
    << "          ylast = y[7];\n"
    << "          const int l_ = (qs & 3) * 8 + g;  // the stage-2 lane that owns these eight samples\n"
    << "          yrow[l_ ^ k_] = make_float4(y[0], y[1], y[2], y[3]);\n"
    << "          yrow[32 + (l_ ^ k_)] = make_float4(y[4], y[5], y[6], y[7]);\n"
    << "        }\n"
    << "      }\n"
    << "      __syncthreads();\n"
    << "    }\n"
    << "    if (lane < VC) { " << W(7) << " = (double)ylast; " << W(8) << " = q2; }\n"
    << "  }\n";
  if (!all_mut.empty()) {
    s << "  __syncthreads();\n";
    for (int w : all_mut)
      s << "  if (wib == " << pj[w] % (NP + 1) << " && lane < VC) kp.words[" << w << " * nv + v0 + lane] = pw[" << pj[w] << " * VC + lane];\n";
  }
  s << "}\n";
  return s.str();
}

}  // namespace

size_t piped_compile_only(const Program &p, int VC, PipedInfo *info, std::string &err, std::string *src_out) {
  const std::string src = generate_piped_source(p, VC, info);
  if (src_out) *src_out = src;
  if (src.empty()) { err = "the circuit has no 3-stage split"; return 0; }
  return fused_build_cubin(src, err);
}

FusedKernel *piped_get(const Program &p, int VC, PipedInfo *info, std::string &err, int out_mode) {
  const std::string src = generate_piped_source(p, VC, info, out_mode);
  if (src.empty()) { err = "the circuit has no 3-stage split"; return nullptr; }
  return fused_build(src, "dusp_piped", err);
}

}  // namespace dusp
