// capi.cc -- the C ABI of libdusp_b200.so (include/dusp_b200.h): programs,
// device-resident sessions and the render call.  No CPU rendering path exists
// in this library: without a CUDA device every session call fails.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <algorithm>
#include <string>
#include <vector>

#include "../../include/dusp_b200.h"
#include "compile.h"
#include "fused.h"
#include "graph.h"
#include "launch.h"

using namespace dusp;

static thread_local std::string g_err;
static int fail(int code, const std::string &msg) { g_err = msg; return code; }
namespace dusp { int fail_public(int code, const std::string &msg) { return fail(code, msg); } }  // mixpush.cu
#define CUDA_TRY(x)                                                                                   \
  do {                                                                                                \
    cudaError_t e_ = (x);                                                                             \
    if (e_ != cudaSuccess) return fail(DUSP_ECUDA, std::string(#x) + ": " + cudaGetErrorString(e_)); \
  } while (0)

struct dusp_program {
  Program prog;
  std::string listing;
};

namespace {

// waveTables.js:5-40 and shapeTables.js:3-38: 44101-entry float32 tables, built
// with the host libm exactly as the reference builds them with Math.sin.
const std::vector<float> &host_tables() {
  static std::vector<float> t;
  static std::once_flag once;
  std::call_once(once, [] {
    const int N = DUSP_TABLE_LEN, SR = DUSP_SR;
    t.assign((size_t)TAB_COUNT * N, 0.f);
    float *sine = &t[(size_t)TAB_SIN * N], *saw = &t[(size_t)TAB_SAW * N], *sq = &t[(size_t)TAB_SQUARE * N];
    float *tri = &t[(size_t)TAB_TRIANGLE * N], *bit8 = &t[(size_t)TAB_8BIT * N];
    const double PHI = 2 * M_PI;
    for (int i = 0; i < N; i++) sine[i] = (float)std::sin(PHI * i / N);
    for (int i = 0; i < SR; i++) saw[i] = (float)(-1 + i * 2.0 / N);
    const int q = SR / 4;
    for (int i = 0; i < q; i++) {
      tri[i] = (float)((double)i / SR * 4);
      tri[i + q] = (float)(1 - (double)tri[i]);
      tri[i + 2 * q] = (float)(-(double)tri[i]);
      tri[i + 3 * q] = (float)(-1 + (double)tri[i]);
    }
    tri[SR] = 0;
    for (int i = 0; i < N; i++) sq[i] = i < SR / 2 ? 1.f : -1.f;
    for (int i = 0; i < N; i++) bit8[i] = (float)(std::floor((double)sine[i] * 128.0 + 0.5) / 128.0);
    float *dec = &t[(size_t)TAB_DECAY * N], *att = &t[(size_t)TAB_ATTACK * N];
    float *semi = &t[(size_t)TAB_SEMISINE * N], *dsq = &t[(size_t)TAB_DECAYSQ * N];
    for (int i = 0; i < N; i++) {
      const double x = (double)i / SR;
      dec[i] = (float)(1 - x);
      att[i] = (float)x;
      semi[i] = (float)std::sin(M_PI * x);
      dsq[i] = (float)((1 - x) * (1 - x));
    }
  });
  return t;
}

struct Group {
  Program prog;
  int n_voices = 0, voice0 = 0;
  long long nv_pad = 0;
  std::vector<double> h_words;  // [n_words][nv_pad] initial values (defaults + overrides)
  std::vector<DuspDevInstr> dev_code;
  DuspDevInstr *d_code = nullptr;
  int sub_block = 16, time_lanes = 1, warps_per_cta = 1;
  FusedKernel *fused = nullptr;  // specialised straight-line kernel (fused.cc); null -> interpreter.  Built for DUSP_OUT_VOICES;
  FusedKernel *fused_mode[3] = {nullptr, nullptr, nullptr};  // ... the other output modes are built on first use (enqueue_block)
  int piped_vc = 0, fused_wps = 0;
  bool piped = false;            // `fused` is the pipelined kernel (piped.cc)
  PipedInfo pinfo;
  double *d_words = nullptr;
  float *d_rings = nullptr;
  DuspRing *d_ringtab = nullptr;
  float *d_stale = nullptr;
  double *d_ctab = nullptr;  // program constants (FFT twiddles, window envelopes); owned by the session
  bool dirty = true;  // device words/rings need (re)initialisation
  size_t off_words = 0, off_rings = 0, off_stale = 0;  // this group's segments of the session pools (elements)
  bool in_multi = false;  // rendered by the session's one multi-program interpreter launch
  // segmented rendering (ops.cuh "segments of one voice's timeline on different warps"): small batches of circuits whose
  // stateful units can all be put at an arbitrary sample -- a launch's timeline is cut into up to seg_max pieces per voice
  int seg_max = 0, seg_nosc = 0, seg_wps = 0;
  FusedKernel *seg_scan = nullptr;
  FusedKernel *seg_render[3] = {nullptr, nullptr, nullptr};  // per output mode, built on first use
  double *d_seg_totals = nullptr;  // [n_voices][seg_max][max(seg_nosc, 1)]
  int *d_seg_flags = nullptr;      // [n_voices]
};

}  // namespace

struct dusp_session {
  int device = 0;
  std::vector<Group> groups;
  int n_voices = 0, out_nch = 0;
  long long unit_count = 0;
  long long clock = 0;
  float *d_tables = nullptr;
  unsigned *d_sm_slots = nullptr;  // [256] per-SM CTA arrival counters of the pipelined kernel (never reset: only differences matter)
  // every group's arrays are segments of five session-wide allocations (a heterogeneous batch of 10 000
  // circuits is 10 000 groups: one upload / two memsets re-initialise them all)
  double *d_words_pool = nullptr; float *d_rings_pool = nullptr; float *d_stale_pool = nullptr;
  DuspRing *d_ringtab_pool = nullptr; DuspDevInstr *d_code_pool = nullptr;
  size_t words_total = 0, rings_total = 0, stale_total = 0;
  std::vector<double> h_all;  // staging image of the words pool
  long long *d_dbg = nullptr;   // DUSP_MULTI_DEBUG_FILE: per-warp tick counts of the last multi-program launch
  DuspWarpDesc *d_multi = nullptr;  // one descriptor per voice of the interpreter groups (multi-program launch)
  int n_multi = 0, multi_slots = 1;
  float *d_staging[2] = {nullptr, nullptr};  // host-output path: double-buffered device blocks
  size_t staging_floats = 0;
  cudaStream_t copy_stream = nullptr;
  // independent program groups of a block render concurrently: group k > 0 launches on side stream k-1 between a
  // fork event on the caller's stream and a join event per side stream (the latency-bound pipelined kernel of one
  // group leaves issue slots and CTA slots that another group's kernel fills)
  std::vector<cudaStream_t> side_streams;
  std::vector<cudaEvent_t> side_done;
  cudaEvent_t ev_fork = nullptr;
  cudaEvent_t ev_rendered[2] = {nullptr, nullptr}, ev_copied[2] = {nullptr, nullptr};
  long long launches = 0;
  std::string warning;
  int last_grid = 0, last_block = 0, last_smem = 0;
  bool last_fused = false;
};

extern "C" {

int dusp_abi_version(void) { return DUSP_ABI_VERSION; }
const char *dusp_last_error(void) { return g_err.c_str(); }

int dusp_program_create(const char *blob, size_t len, dusp_program **out) {
  if (!blob || !out) return fail(DUSP_EINVAL, "renderAudioBuffer expects an outlet");
  try {
    Graph g = parse_graph(blob, len);
    auto p = std::make_unique<dusp_program>();
    p->prog = compile(g);
    p->listing = p->prog.listing();
    *out = p.release();
    return DUSP_OK;
  } catch (const Unsupported &e) {
    return fail(DUSP_EUNSUPPORTED, e.what());
  } catch (const std::exception &e) {
    return fail(DUSP_EINVAL, e.what());
  }
}
void dusp_program_destroy(dusp_program *p) { delete p; }
int dusp_program_num_channels(const dusp_program *p) { return p->prog.out_nch; }
int dusp_program_num_units(const dusp_program *p) { return p->prog.n_units; }
int dusp_program_warmup_ticks(const dusp_program *p) { return p->prog.warmup_ticks; }
int dusp_program_num_instructions(const dusp_program *p) { return (int)p->prog.code.size(); }
int dusp_program_num_slots(const dusp_program *p) { return p->prog.n_slots; }
double dusp_program_bytes_per_voice_sample(const dusp_program *p, int out_mode) {
  return p->prog.ring_bytes_per_sample + (out_mode != DUSP_OUT_MIX ? 4.0 * p->prog.out_nch : 0.0);
}
int dusp_program_num_params(const dusp_program *p) { return (int)p->prog.words.size(); }
const char *dusp_program_param_name(const dusp_program *p, int i) {
  return i >= 0 && i < (int)p->prog.words.size() ? p->prog.words[i].name.c_str() : nullptr;
}
int dusp_program_param_index(const dusp_program *p, const char *name) {
  for (size_t i = 0; i < p->prog.words.size(); i++) if (p->prog.words[i].name == name) return (int)i;
  return -1;
}
double dusp_program_param_default(const dusp_program *p, int i) {
  return i >= 0 && i < (int)p->prog.words.size() ? p->prog.words[i].dflt : NAN;
}
size_t dusp_program_disassemble(const dusp_program *p, char *buf, size_t cap) {
  if (buf && cap) { size_t n = std::min(cap - 1, p->listing.size()); memcpy(buf, p->listing.data(), n); buf[n] = 0; }
  return p->listing.size() + 1;
}


void dusp_session_destroy(dusp_session *s) {
  if (!s) return;
  cudaSetDevice(s->device);
  cudaFree(s->d_words_pool); cudaFree(s->d_rings_pool); cudaFree(s->d_stale_pool); cudaFree(s->d_ringtab_pool);
  cudaFree(s->d_code_pool); cudaFree(s->d_multi); cudaFree(s->d_dbg);
  cudaFree(s->d_tables); cudaFree(s->d_sm_slots);
  for (auto &g : s->groups) { cudaFree(g.d_ctab); cudaFree(g.d_seg_totals); cudaFree(g.d_seg_flags); }
  for (int i = 0; i < 2; i++) {
    cudaFree(s->d_staging[i]);
    if (s->ev_rendered[i]) cudaEventDestroy(s->ev_rendered[i]);
    if (s->ev_copied[i]) cudaEventDestroy(s->ev_copied[i]);
  }
  if (s->copy_stream) cudaStreamDestroy(s->copy_stream);
  for (cudaStream_t t : s->side_streams) cudaStreamDestroy(t);
  for (cudaEvent_t e : s->side_done) cudaEventDestroy(e);
  if (s->ev_fork) cudaEventDestroy(s->ev_fork);
  delete s;
}

int dusp_session_create(const dusp_program *const *programs, const int *voices, int n_programs, int device,
                        dusp_session **out) {
  if (!programs || !voices || n_programs < 1 || !out) return fail(DUSP_EINVAL, "dusp_session_create: bad arguments");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1)
    return fail(DUSP_ECUDA, "no CUDA device: libdusp_b200 has no CPU rendering path");
  if (device < 0 || device >= ndev) return fail(DUSP_EINVAL, "dusp_session_create: no such CUDA device");
  CUDA_TRY(cudaSetDevice(device));
  std::unique_ptr<dusp_session, void (*)(dusp_session *)> s(new dusp_session, dusp_session_destroy);
  s->device = device;
  const auto &tabs = host_tables();
  CUDA_TRY(cudaMalloc(&s->d_tables, tabs.size() * sizeof(float)));
  CUDA_TRY(cudaMemcpy(s->d_tables, tabs.data(), tabs.size() * sizeof(float), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMalloc(&s->d_sm_slots, 256 * sizeof(unsigned)));
  CUDA_TRY(cudaMemset(s->d_sm_slots, 0, 256 * sizeof(unsigned)));
  int voice0 = 0;
  for (int i = 0; i < n_programs; i++) {
    if (!programs[i] || voices[i] < 1) return fail(DUSP_EINVAL, "dusp_session_create: empty program group");
    s->groups.emplace_back();
    Group &g = s->groups.back();
    g.prog = programs[i]->prog;
    g.n_voices = voices[i];
    g.voice0 = voice0;
    // batches: whole warps / whole CTAs of 16, 24 or 32 voices; small groups render one voice per warp
    const bool forced = getenv("DUSP_TILE") || getenv("DUSP_FUSED") || getenv("DUSP_PIPED");  // experiments / tests
    g.nv_pad = voices[i] >= 64 || forced ? ((long long)voices[i] + 95) / 96 * 96 : ((long long)voices[i] + 7) / 8 * 8;
    voice0 += voices[i];
    s->out_nch = std::max(s->out_nch, g.prog.out_nch);
    s->unit_count += (long long)g.prog.n_units * voices[i];
    const size_t nw = g.prog.words.size();
    g.h_words.resize(std::max<size_t>(nw, 1) * g.nv_pad);
    for (size_t w = 0; w < nw; w++) {
      const Param &p = g.prog.words[w];
      const double v = p.f32 ? (double)(float)p.dflt : p.dflt;
      for (long long k = 0; k < g.nv_pad; k++) g.h_words[w * g.nv_pad + k] = v;
    }
    for (long long k = 0; k < g.nv_pad; k++) derive_words(g.prog, g.h_words.data() + k, g.nv_pad);
    for (const DuspInstr &in : g.prog.code) {
      DuspDevInstr d;
      memset(&d, 0, sizeof d);
      d.op = in.op;
      for (int k = 0; k < 5; k++) d.in[k] = in.in[k];
      for (int k = 0; k < 2; k++) d.out[k] = in.out[k];
      d.p = in.p; d.aux0 = in.aux[0]; d.aux1 = in.aux[1]; d.aux2 = in.aux[2];
      g.dev_code.push_back(d);
    }
    {
      // Tile shape: T time-lanes per voice so that the batch still fills the GPU (>= 8 warps per SM);
      // B samples per lane.  DUSP_TILE="B,T" overrides (experiments).
      // (measured on B200, profiles/r1_tile_sweep.txt: for 10k voices B=8,T=4 wins; B=8 keeps the kernel
      // at 128 registers so 16 warps/SM are resident and the whole batch runs as one wave)
      const int choices[4] = {1, 4, 8, 32};
      int T = 32;
      for (int t : choices) if (((long long)voices[i] + (32 / t) - 1) / (32 / t) >= 148 * 8) { T = t; break; }
      int B = T == 1 ? 16 : 8;
      if (const char *env = getenv("DUSP_TILE")) { int b = 0, t = 0; if (sscanf(env, "%d,%d", &b, &t) == 2) { B = b; T = t; } }
      // Circuit specialisation (fused.cc): batches get a straight-line kernel compiled for this circuit
      // shape; single circuits keep the interpreter (no compile latency).  DUSP_FUSED=0/1 forces either.
      const char *fe = getenv("DUSP_FUSED");
      // per-chunk control ops (Retriggerer ...) need whole-chunk tiles: the interpreter at B = 8, T = 32
      const bool chunky = g.prog.chunk_tiles;
      if (chunky) { B = 8; T = 32; }
      const bool want = !chunky && (fe ? atoi(fe) != 0 : voices[i] >= 64);
      const bool pow2 = T >= 1 && T <= 32 && (T & (T - 1)) == 0 && (B == 4 || B == 8 || B == 16);
      // Pipelined kernel (piped.cc): lanes along time for the parallel units, a lane-per-voice warp for the
      // recurrences.  Default for batches; circuits without a 3-stage split use the tile kernel.  DUSP_PIPED=0/1 forces.
      const char *pe = getenv("DUSP_PIPED");
      const bool has_seq = program_has_sequential(g.prog);
      if ((!has_seq && !getenv("DUSP_TILE")) || chunky) { T = 32; B = 8; }  // measured: W2 10k voices 4.6 ms at (8,32) vs 9.6 ms at (8,4)
      if (want && (pe ? atoi(pe) != 0 : voices[i] >= 64 && has_seq)) {
        int vc = 0;
        if (const char *ve = getenv("DUSP_PIPED_VC")) vc = atoi(ve);
        if (!vc) {  // fewest voices on the fullest SM
          long long best = 1LL << 60;
          for (int c : {32, 24, 16, 8}) {  // ties go to the larger CTA (fuller lane-per-voice warp)
            const long long ctas = ((long long)voices[i] + c - 1) / c, per_sm = (ctas + 147) / 148 * c;
            if (per_sm < best) { best = per_sm; vc = c; }
          }
        }
        std::string err;
        // the pipeline pays while the lane-per-voice warp carries a small part of the circuit (W3: 1 op of 14);
        // a feedback loop that runs there almost entirely (W3b: 7 of 13) is faster on the tile kernel (measured)
        PipedInfo probe;
        const bool worth = pe || (!generate_piped_source(g.prog, vc, &probe).empty() && probe.n_ops_sequential_stage * 3 <= probe.n_ops);
        if (worth) g.fused = piped_get(g.prog, vc, &g.pinfo, err);
        if (g.fused) { g.piped = true; g.piped_vc = vc; g.sub_block = 4; g.time_lanes = 32; g.warps_per_cta = 5; }
        else if (pe) s->warning = "pipelined kernel unavailable: " + err;
      }
      if (want && pow2 && !g.fused) {
        // one-warp CTAs (warps never talk to each other); the register budget is set so that the whole
        // batch is resident at once (no tail wave): W warps per SM, W in [8, 24].  DUSP_FUSED_WPS overrides.
        const long long n_warps = ((long long)voices[i] + (32 / T) - 1) / (32 / T);
        int wps = (int)std::min<long long>(24, std::max<long long>(8, (n_warps + 147) / 148));
        if (const char *we = getenv("DUSP_FUSED_WPS")) wps = std::max(1, atoi(we));
        std::string err;
        g.fused = fused_get(g.prog, B, T, wps, err);
        if (!g.fused && fe) return fail(DUSP_ECUDA, "circuit specialisation failed: " + err);
        if (!g.fused) s->warning = "circuit specialisation unavailable, using the interpreter: " + err;
        if (g.fused) { g.fused_wps = wps; g.sub_block = B; g.time_lanes = T; g.warps_per_cta = 1; }
        // a batch too small to fill the GPU with one warp per voice: several warps per voice, along time
        if (g.fused && B == 8 && T == 32 && !getenv("DUSP_NO_SEGMENTS")) {
          // about 48 warps per SM in total: two waves of the 24 resident ones (measured on W2, 1 000 voices x 10 s: 2.85 ms at
          // 8 segments, 3.09 at 3, 3.32 in one piece; 100 voices: 0.33 ms at 35 segments against 2.93)
          int smax = (int)std::min<long long>(64, (148 * 48) / std::max<long long>(n_warps, 1));
          if (const char *se = getenv("DUSP_SEGMENTS")) smax = std::max(1, std::min(64, atoi(se)));
          std::string scan_src, render_src;
          if (smax >= 2 && generate_segmented_sources(g.prog, 1, 0, scan_src, render_src, &g.seg_nosc)) {
            g.seg_wps = (int)std::min<long long>(24, std::max<long long>(8, (n_warps * smax + 147) / 148));
            generate_segmented_sources(g.prog, g.seg_wps, 0, scan_src, render_src, &g.seg_nosc);
            std::string e2;
            g.seg_scan = fused_build(scan_src, "dusp_fused_scan", e2);
            if (g.seg_scan) g.seg_render[0] = fused_build(render_src, "dusp_fused", e2);
            if (g.seg_scan && g.seg_render[0]) g.seg_max = smax;
            else s->warning = "segmented kernels unavailable: " + e2;
          }
        }
      }
      if (!g.fused) {
        if ((size_t)g.prog.n_slots * B * 128 > 200 * 1024 && T == 1 && !chunky) B = 8;
        if (!dusp_launch_supported(B, T)) return fail(DUSP_EINVAL, "unsupported tile shape (DUSP_TILE)");
        if ((size_t)g.prog.n_slots * B * 128 > 220 * 1024)
          return fail(DUSP_EUNSUPPORTED, "circuit needs more live signal slots than shared memory holds");
        const long long n_warps = ((long long)voices[i] + (32 / T) - 1) / (32 / T);
        int wpc = n_warps >= 148 * 16 ? 4 : 1;
        while (wpc > 1 && (size_t)g.prog.n_slots * B * 128 * wpc > 100 * 1024) wpc >>= 1;
        g.sub_block = B; g.time_lanes = T; g.warps_per_cta = wpc;
      }
    }
  }
  // ---- session pools: every group's words / rings / stale / ring table / code are 256-byte aligned segments
  {
    auto up = [](size_t n, size_t elem) { const size_t a = 256 / elem; return (n + a - 1) / a * a; };
    size_t n_words = 0, n_rings = 0, n_stale = 0, n_rt = 0, n_code = 0;
    std::vector<size_t> off_rt, off_code;
    for (auto &g : s->groups) {
      g.off_words = n_words; n_words += up(g.h_words.size(), sizeof(double));
      g.off_rings = n_rings; n_rings += up((size_t)std::max<long long>(g.prog.ring_floats, 4) * g.nv_pad, sizeof(float));
      g.off_stale = n_stale; n_stale += up((size_t)std::max(g.prog.n_stale, 1) * DUSP_CHUNK * g.nv_pad, sizeof(float));
      off_rt.push_back(n_rt); n_rt += up(std::max<size_t>(g.prog.rings.size(), 1), sizeof(DuspRing));
      off_code.push_back(n_code); n_code += up(std::max<size_t>(g.dev_code.size(), 1), sizeof(DuspDevInstr));
    }
    s->words_total = n_words; s->rings_total = n_rings; s->stale_total = n_stale;
    CUDA_TRY(cudaMalloc(&s->d_words_pool, n_words * sizeof(double)));
    if (cudaMalloc(&s->d_rings_pool, n_rings * sizeof(float)) != cudaSuccess) {
      cudaGetLastError();
      char m[160];
      snprintf(m, sizeof m, "delay lines need %.1f GB of device memory for %d voices", n_rings * 4 / 1e9, voice0);
      return fail(DUSP_ENOMEM, m);
    }
    CUDA_TRY(cudaMalloc(&s->d_stale_pool, n_stale * sizeof(float)));
    CUDA_TRY(cudaMalloc(&s->d_ringtab_pool, n_rt * sizeof(DuspRing)));
    CUDA_TRY(cudaMalloc(&s->d_code_pool, n_code * sizeof(DuspDevInstr)));
    std::vector<DuspRing> h_rt(n_rt);
    std::vector<DuspDevInstr> h_code(n_code);
    memset(h_code.data(), 0, n_code * sizeof(DuspDevInstr));
    for (size_t i = 0; i < s->groups.size(); i++) {
      Group &g = s->groups[i];
      g.d_words = s->d_words_pool + g.off_words; g.d_rings = s->d_rings_pool + g.off_rings;
      g.d_stale = s->d_stale_pool + g.off_stale; g.d_ringtab = s->d_ringtab_pool + off_rt[i];
      g.d_code = s->d_code_pool + off_code[i];
      std::copy(g.prog.rings.begin(), g.prog.rings.end(), h_rt.begin() + off_rt[i]);
      std::copy(g.dev_code.begin(), g.dev_code.end(), h_code.begin() + off_code[i]);
    }
    for (auto &g : s->groups)
      if (g.seg_max >= 2) {
        CUDA_TRY(cudaMalloc(&g.d_seg_totals, (size_t)g.n_voices * g.seg_max * std::max(g.seg_nosc, 1) * sizeof(double)));
        CUDA_TRY(cudaMalloc(&g.d_seg_flags, (size_t)g.n_voices * sizeof(int)));
      }
    for (auto &g : s->groups)
      if (!g.prog.ctab.empty()) {
        CUDA_TRY(cudaMalloc(&g.d_ctab, g.prog.ctab.size() * sizeof(double)));
        CUDA_TRY(cudaMemcpy(g.d_ctab, g.prog.ctab.data(), g.prog.ctab.size() * sizeof(double), cudaMemcpyHostToDevice));
      }
    CUDA_TRY(cudaMemcpy(s->d_ringtab_pool, h_rt.data(), n_rt * sizeof(DuspRing), cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(s->d_code_pool, h_code.data(), n_code * sizeof(DuspDevInstr), cudaMemcpyHostToDevice));
  }
  // ---- heterogeneous batches: all interpreter groups with the one-voice-per-warp tile (B = 8, T = 32)
  //      render in ONE launch, every warp with its own program (vm.h DuspWarpDesc)
  {
    std::vector<DuspWarpDesc> descs;
    int n_cand = 0;
    for (auto &g : s->groups) if (!g.fused && g.sub_block == 8 && g.time_lanes == 32 && !g.prog.chunk_tiles) n_cand++;
    if (n_cand >= 2 && !getenv("DUSP_NO_MULTI"))
      for (auto &g : s->groups) {
        if (g.fused || g.sub_block != 8 || g.time_lanes != 32 || g.prog.chunk_tiles) continue;
        g.in_multi = true;
        s->multi_slots = std::max(s->multi_slots, g.prog.n_slots);
        const long long rs = std::max<long long>(g.prog.ring_floats, 4);
        for (int v = 0; v < g.n_voices; v++) {
          DuspWarpDesc d;
          memset(&d, 0, sizeof d);
          d.code = g.d_code; d.words = g.d_words + v; d.rings = g.d_rings + (long long)v * rs; d.ring_tab = g.d_ringtab;
          d.stale = g.d_stale + (long long)v * g.prog.n_stale * DUSP_CHUNK; d.nv_stride = g.nv_pad; d.out_row = g.voice0 + v;
          d.n_instr = (int)g.dev_code.size(); d.n_stale = g.prog.n_stale; d.prog_nch = g.prog.out_nch;
          descs.push_back(d);
        }
      }
    // longest circuits first: the launch is one CTA per circuit handed out in index order, so its tail is the circuit
    // that starts last; the estimate weighs the ops that run as 32 serial turns of the warp
    if (descs.size() > 1 && !getenv("DUSP_MULTI_UNSORTED")) {
      std::vector<std::pair<long long, size_t>> key(descs.size());
      size_t k = 0;
      for (auto &g : s->groups) {
        if (!g.in_multi) continue;
        long long cost = 0;
        for (auto &I : g.dev_code)
          cost += (I.op == OP_FILTER || I.op == OP_FILTERC || I.op == OP_SHAPE || I.op == OP_AHD || I.op == OP_COMB || I.op == OP_ALLPASS) ? 12
                  : (I.op == OP_OSC || I.op == OP_MCOSC || I.op == OP_NOISE || I.op == OP_RAMP) ? 3 : 1;
        for (int v = 0; v < g.n_voices; v++, k++) key[k] = {-cost, k};
      }
      std::stable_sort(key.begin(), key.end());
      std::vector<DuspWarpDesc> sorted(descs.size());
      for (size_t i = 0; i < key.size(); i++) sorted[i] = descs[key[i].second];
      descs.swap(sorted);
    }
    s->n_multi = (int)descs.size();
    if (s->n_multi) {
      CUDA_TRY(cudaMalloc(&s->d_multi, descs.size() * sizeof(DuspWarpDesc)));
      CUDA_TRY(cudaMemcpy(s->d_multi, descs.data(), descs.size() * sizeof(DuspWarpDesc), cudaMemcpyHostToDevice));
    }
  }
  s->n_voices = voice0;
  *out = s.release();
  return DUSP_OK;
}

int dusp_session_num_voices(const dusp_session *s) { return s->n_voices; }
int dusp_session_num_channels(const dusp_session *s) { return s->out_nch; }
long long dusp_session_unit_count(const dusp_session *s) { return s->unit_count; }
long long dusp_session_launch_count(const dusp_session *s) { return s->launches; }
int dusp_session_launch_info(const dusp_session *s, int *grid, int *block, int *smem) {
  if (grid) *grid = s->last_grid;
  if (block) *block = s->last_block;
  if (smem) *smem = s->last_smem;
  return DUSP_OK;
}

int dusp_session_set_params(dusp_session *s, int gi, const int *ids, int n_ids, const double *values) {
  if (!s || gi < 0 || gi >= (int)s->groups.size() || !ids || !values || n_ids < 0)
    return fail(DUSP_EINVAL, "dusp_session_set_params: bad arguments");
  // parameters (re)initialise a group at clock 0; in a session that has rendered, the other groups would keep
  // their advanced oscillator phases / filter memory / rings while clock-indexed units restart: refuse instead
  if (s->clock != 0) return fail(DUSP_EINVAL, "dusp_session_set_params: the session has rendered; call dusp_session_reset first");
  Group &g = s->groups[gi];
  const int nw = (int)g.prog.words.size();
  for (int k = 0; k < n_ids; k++)
    if (ids[k] < 0 || ids[k] >= nw) return fail(DUSP_EINVAL, "dusp_session_set_params: no such parameter");
  for (int v = 0; v < g.n_voices; v++)
    for (int k = 0; k < n_ids; k++) {
      const Param &p = g.prog.words[ids[k]];
      const double x = values[(size_t)v * n_ids + k];
      g.h_words[(size_t)ids[k] * g.nv_pad + v] = p.f32 ? (double)(float)x : x;
    }
  if (!g.prog.tap_pairs.empty() || !g.prog.locked.empty() || !g.prog.equal_words.empty()) {
    // per-voice values must keep what the compiler proved / specialised on (compile.cc)
    std::vector<double> wv(nw);
    for (int v = 0; v < g.n_voices; v++) {
      for (int w = 0; w < nw; w++) wv[w] = g.h_words[(size_t)w * g.nv_pad + v];
      std::string msg = g.prog.check_taps(wv.data());
      if (msg.empty()) msg = g.prog.check_locked(wv.data());
      if (!msg.empty()) return fail(DUSP_EUNSUPPORTED, "voice " + std::to_string(v) + ": " + msg);
    }
  }
  for (int v = 0; v < g.n_voices; v++) derive_words(g.prog, g.h_words.data() + v, g.nv_pad);
  g.dirty = true;
  return DUSP_OK;
}

int dusp_session_poke(dusp_session *s, int gi, int voice, int n_ids, const int *ids, const double *values, void *cuda_stream) {
  if (!s || gi < 0 || gi >= (int)s->groups.size() || n_ids < 0 || (n_ids && (!ids || !values)))
    return fail(DUSP_EINVAL, "dusp_session_poke: bad arguments");
  Group &g = s->groups[gi];
  if (voice < -1 || voice >= g.n_voices) return fail(DUSP_EINVAL, "dusp_session_poke: no such voice");
  const int nw = (int)g.prog.words.size();
  for (int k = 0; k < n_ids; k++) {
    if (ids[k] < 0 || ids[k] >= nw) return fail(DUSP_EINVAL, "dusp_session_poke: no such parameter");
    for (const Derived &d : g.prog.derived)
      if (d.src0 == ids[k] || d.src1 == ids[k])
        return fail(DUSP_EUNSUPPORTED, "parameter " + g.prog.words[ids[k]].name + " was specialised at compile time and cannot change during a render");
    for (auto &l : g.prog.locked) if (l.word == ids[k]) return fail(DUSP_EUNSUPPORTED, "parameter " + g.prog.words[ids[k]].name + " is locked by a compile-time specialisation");
    for (auto &e : g.prog.equal_words) if (e.word_a == ids[k] || e.word_b == ids[k]) return fail(DUSP_EUNSUPPORTED, "parameter " + g.prog.words[ids[k]].name + " belongs to merged units");
    for (auto &t : g.prog.tap_pairs) if (t.word_a == ids[k] || t.word_b == ids[k]) return fail(DUSP_EUNSUPPORTED, "CircleBuffer tap offsets cannot change during a render");
  }
  CUDA_TRY(cudaSetDevice(s->device));
  cudaStream_t st = (cudaStream_t)cuda_stream;
  // pending (re)initialisation must land first, or it would overwrite the poke
  if (g.dirty) {
    CUDA_TRY(cudaMemcpyAsync(g.d_words, g.h_words.data(), g.h_words.size() * sizeof(double), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemsetAsync(g.d_rings, 0, (size_t)std::max<long long>(g.prog.ring_floats, 4) * g.nv_pad * sizeof(float), st));
    CUDA_TRY(cudaMemsetAsync(g.d_stale, 0, (size_t)std::max(g.prog.n_stale, 1) * DUSP_CHUNK * g.nv_pad * sizeof(float), st));
    g.dirty = false;
  }
  std::vector<double> row;
  for (int k = 0; k < n_ids; k++) {  // in order: a later assignment of the same word wins, like the event's own statements
    const Param &p = g.prog.words[ids[k]];
    const double v = p.f32 ? (double)(float)values[k] : values[k];
    double *dst = g.d_words + (size_t)ids[k] * g.nv_pad;
    if (voice >= 0) {
      CUDA_TRY(cudaMemcpyAsync(dst + voice, &v, sizeof(double), cudaMemcpyHostToDevice, st));
    } else {
      row.assign((size_t)g.nv_pad, v);
      CUDA_TRY(cudaMemcpyAsync(dst, row.data(), row.size() * sizeof(double), cudaMemcpyHostToDevice, st));
    }
    CUDA_TRY(cudaStreamSynchronize(st));  // the staging values live on this stack frame
  }
  return DUSP_OK;
}

// Graph edits (`then`, events that assign inlets).  `dst` holds the program compiled from the circuit
// AFTER the edit; group `dgi` takes over the device state of group `sgi` of `src`: units are matched by kind and
// label, and everything the units own travels -- state words, per-voice constants (unless the edit changed the
// constant's value), delay lines, CircleBuffers and the previous-chunk buffers of feedback edges.  New units keep
// the construction state the blob gave them.  The sessions must be on the same device with equal voice counts.
int dusp_session_adopt(dusp_session *dst, int dgi, dusp_session *src, int sgi, void *cuda_stream) {
  if (!dst || !src || dst == src || dgi < 0 || dgi >= (int)dst->groups.size() || sgi < 0 || sgi >= (int)src->groups.size())
    return fail(DUSP_EINVAL, "dusp_session_adopt: bad arguments");
  Group &d = dst->groups[dgi], &o = src->groups[sgi];
  if (dst->device != src->device || d.n_voices != o.n_voices || d.nv_pad != o.nv_pad)
    return fail(DUSP_EINVAL, "dusp_session_adopt: sessions differ in device or voice count");
  CUDA_TRY(cudaSetDevice(dst->device));
  cudaStream_t st = (cudaStream_t)cuda_stream;
  for (Group *g : {&d, &o})
    if (g->dirty) {  // pending (re)initialisation lands first
      CUDA_TRY(cudaMemcpyAsync(g->d_words, g->h_words.data(), g->h_words.size() * sizeof(double), cudaMemcpyHostToDevice, st));
      CUDA_TRY(cudaMemsetAsync(g->d_rings, 0, (size_t)std::max<long long>(g->prog.ring_floats, 4) * g->nv_pad * sizeof(float), st));
      CUDA_TRY(cudaMemsetAsync(g->d_stale, 0, (size_t)std::max(g->prog.n_stale, 1) * DUSP_CHUNK * g->nv_pad * sizeof(float), st));
      g->dirty = false;
    }
  // "u<j>.<rest>" of the new program -> "u<i>.<rest>" of the old one
  std::map<std::string, int> old_unit;
  for (size_t i = 0; i < o.prog.unit_labels.size(); i++) old_unit[o.prog.unit_labels[i]] = (int)i;
  auto translate = [&](const std::string &name, bool *known_unit) -> std::string {
    *known_unit = false;
    const size_t dot = name.find('.');
    if (name.size() < 2 || name[0] != 'u' || dot == std::string::npos) return "";
    const int j = atoi(name.c_str() + 1);
    if (j < 0 || j >= (int)d.prog.unit_labels.size()) return "";
    auto it = old_unit.find(d.prog.unit_labels[j]);
    if (it == old_unit.end()) return "";
    *known_unit = true;
    return "u" + std::to_string(it->second) + name.substr(dot);
  };
  std::map<std::string, int> ow, orr, os;
  for (size_t i = 0; i < o.prog.words.size(); i++) ow[o.prog.words[i].name] = (int)i;
  for (size_t i = 0; i < o.prog.ring_names.size(); i++) orr[o.prog.ring_names[i]] = (int)i;
  for (size_t i = 0; i < o.prog.stale_names.size(); i++) os[o.prog.stale_names[i]] = (int)i;
  // ---- words
  const size_t nw = d.prog.words.size();
  std::vector<int> from(nw, -1);
  for (size_t w = 0; w < nw; w++) {
    bool known;
    const std::string on = translate(d.prog.words[w].name, &known);
    auto it = ow.find(on);
    if (on.empty() || it == ow.end()) continue;
    const Param &pn = d.prog.words[w], &po = o.prog.words[it->second];
    if (pn.f32 && !(pn.dflt == po.dflt || (pn.dflt != pn.dflt && po.dflt != po.dflt))) continue;  // the edit set this constant
    from[w] = it->second;
  }
  // words derived from a constant that the edit changed keep their fresh values
  auto derived_count = [](int kind) { return kind == DV_FILTER_LP || kind == DV_FILTER_HP ? 5 : kind == DV_PAN ? 3 : 1; };
  for (const Derived &dv : d.prog.derived) {
    bool fresh = false;
    for (int sw : {dv.src0, dv.src1}) if (sw >= 0 && d.prog.words[sw].f32 && from[sw] < 0) fresh = true;
    if (fresh) for (int k = 0; k < derived_count(dv.kind); k++) from[dv.dst + k] = -1;
  }
  for (size_t w = 0; w < nw; w++)
    if (from[w] >= 0)
      CUDA_TRY(cudaMemcpyAsync(d.d_words + w * d.nv_pad, o.d_words + (size_t)from[w] * o.nv_pad, (size_t)d.nv_pad * sizeof(double),
                               cudaMemcpyDeviceToDevice, st));
  // ---- delay lines and CircleBuffers: [voice][ring_floats] arenas with different strides
  const size_t dpitch = (size_t)std::max<long long>(d.prog.ring_floats, 4) * sizeof(float);
  const size_t opitch = (size_t)std::max<long long>(o.prog.ring_floats, 4) * sizeof(float);
  for (size_t r = 0; r < d.prog.rings.size(); r++) {
    bool known;
    auto it = orr.find(translate(d.prog.ring_names[r], &known));
    if (it == orr.end()) continue;
    const DuspRing &rn = d.prog.rings[r], &ro = o.prog.rings[it->second];
    if (rn.length != ro.length) return fail(DUSP_EUNSUPPORTED, "graph edit changes the length of delay line " + d.prog.ring_names[r]);
    CUDA_TRY(cudaMemcpy2DAsync(d.d_rings + rn.offset, dpitch, o.d_rings + ro.offset, opitch, (size_t)rn.length * sizeof(float),
                               (size_t)d.n_voices, cudaMemcpyDeviceToDevice, st));
  }
  // ---- previous-chunk buffers of feedback edges: [voice][n_stale][256]
  for (size_t k = 0; k < d.prog.stale_names.size(); k++) {
    bool known;
    auto it = os.find(translate(d.prog.stale_names[k], &known));
    if (it == os.end()) {
      // a predecessor that kept every value it produced never had this one: a channel (or outlet) that is new in
      // this program, zero-filled like the reference's `new Float32Array` until its producer first writes it
      if (known && !o.prog.keeps_all)
        return fail(DUSP_EUNSUPPORTED, "graph edit makes " + d.prog.stale_names[k] + " a feedback edge: its previous chunk was not kept");
      continue;
    }
    CUDA_TRY(cudaMemcpy2DAsync(d.d_stale + k * DUSP_CHUNK, (size_t)d.prog.n_stale * DUSP_CHUNK * sizeof(float),
                               o.d_stale + (size_t)it->second * DUSP_CHUNK, (size_t)o.prog.n_stale * DUSP_CHUNK * sizeof(float),
                               DUSP_CHUNK * sizeof(float), (size_t)d.n_voices, cudaMemcpyDeviceToDevice, st));
  }
  CUDA_TRY(cudaStreamSynchronize(st));
  dst->clock = src->clock;
  return DUSP_OK;
}

int dusp_session_reset(dusp_session *s) {
  if (!s) return fail(DUSP_EINVAL, "dusp_session_reset: null session");
  for (auto &g : s->groups) g.dirty = true;
  s->clock = 0;
  return DUSP_OK;
}

// Groups that were (re)created, reset or re-parameterised since their last render: initial words up, rings and stale
// chunks cleared -- on the stream, right before the first kernel that needs them.
static int upload_initial_state(dusp_session *s, cudaStream_t st) {
  bool all_dirty = true, any_dirty = false;
  for (auto &g : s->groups) { all_dirty = all_dirty && g.dirty; any_dirty = any_dirty || g.dirty; }
  if (all_dirty && s->groups.size() > 1) {  // one upload + two memsets for the whole session
    s->h_all.resize(s->words_total);
    for (auto &g : s->groups) { std::copy(g.h_words.begin(), g.h_words.end(), s->h_all.begin() + g.off_words); g.dirty = false; }
    CUDA_TRY(cudaMemcpyAsync(s->d_words_pool, s->h_all.data(), s->words_total * sizeof(double), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemsetAsync(s->d_rings_pool, 0, s->rings_total * sizeof(float), st));
    CUDA_TRY(cudaMemsetAsync(s->d_stale_pool, 0, s->stale_total * sizeof(float), st));
  } else if (any_dirty) {
    for (auto &g : s->groups) {
      if (!g.dirty) continue;
      CUDA_TRY(cudaMemcpyAsync(g.d_words, g.h_words.data(), g.h_words.size() * sizeof(double), cudaMemcpyHostToDevice, st));
      CUDA_TRY(cudaMemsetAsync(g.d_rings, 0, (size_t)std::max<long long>(g.prog.ring_floats, 4) * g.nv_pad * sizeof(float), st));
      CUDA_TRY(cudaMemsetAsync(g.d_stale, 0, (size_t)std::max(g.prog.n_stale, 1) * DUSP_CHUNK * g.nv_pad * sizeof(float), st));
      g.dirty = false;
    }
  }
  return DUSP_OK;
}

// ---- state block (include/dusp_b200.h): header, then per group words | rings | stale
namespace {
struct StateHeader { char magic[8]; long long clock, n_groups; };
struct StateShape { long long words, rings, stale, voices; };  // element counts of one group (padded) and its voices
StateShape state_shape(const Group &g) {
  return {(long long)g.h_words.size(), std::max<long long>(g.prog.ring_floats, 4) * g.nv_pad,
          (long long)std::max(g.prog.n_stale, 1) * DUSP_CHUNK * g.nv_pad, (long long)g.n_voices};
}
size_t state_bytes(const dusp_session *s) {
  size_t n = sizeof(StateHeader) + s->groups.size() * sizeof(StateShape);
  for (auto &g : s->groups) { const StateShape h = state_shape(g); n += (size_t)h.words * sizeof(double) + (size_t)(h.rings + h.stale) * sizeof(float); }
  return n;
}
}  // namespace

int dusp_session_state_size(const dusp_session *s, size_t *bytes) {
  if (!s || !bytes) return fail(DUSP_EINVAL, "dusp_session_state_size: null argument");
  *bytes = state_bytes(s);
  return DUSP_OK;
}

int dusp_session_state_export(dusp_session *s, void *host_block, size_t bytes, void *cuda_stream) {
  if (!s || !host_block) return fail(DUSP_EINVAL, "dusp_session_state_export: null argument");
  if (bytes < state_bytes(s)) return fail(DUSP_EINVAL, "dusp_session_state_export: the buffer is smaller than dusp_session_state_size says");
  cudaStream_t st = (cudaStream_t)cuda_stream;
  CUDA_TRY(cudaSetDevice(s->device));
  if (int e = upload_initial_state(s, st)) return e;  // a session that has not rendered yet exports its initial state
  char *p = static_cast<char *>(host_block);
  StateHeader h;
  memcpy(h.magic, "DUSPST01", 8);
  h.clock = s->clock; h.n_groups = (long long)s->groups.size();
  memcpy(p, &h, sizeof h); p += sizeof h;
  for (auto &g : s->groups) { const StateShape sh = state_shape(g); memcpy(p, &sh, sizeof sh); p += sizeof sh; }
  for (auto &g : s->groups) {
    const StateShape sh = state_shape(g);
    CUDA_TRY(cudaMemcpyAsync(p, g.d_words, (size_t)sh.words * sizeof(double), cudaMemcpyDeviceToHost, st)); p += (size_t)sh.words * sizeof(double);
    CUDA_TRY(cudaMemcpyAsync(p, g.d_rings, (size_t)sh.rings * sizeof(float), cudaMemcpyDeviceToHost, st)); p += (size_t)sh.rings * sizeof(float);
    CUDA_TRY(cudaMemcpyAsync(p, g.d_stale, (size_t)sh.stale * sizeof(float), cudaMemcpyDeviceToHost, st)); p += (size_t)sh.stale * sizeof(float);
  }
  CUDA_TRY(cudaStreamSynchronize(st));
  return DUSP_OK;
}

int dusp_session_state_import(dusp_session *s, const void *host_block, size_t bytes, void *cuda_stream) {
  if (!s || !host_block) return fail(DUSP_EINVAL, "dusp_session_state_import: null argument");
  if (bytes < state_bytes(s)) return fail(DUSP_EINVAL, "dusp_session_state_import: the block is smaller than this session's state");
  const char *p = static_cast<const char *>(host_block);
  StateHeader h;
  memcpy(&h, p, sizeof h); p += sizeof h;
  if (memcmp(h.magic, "DUSPST01", 8) != 0) return fail(DUSP_EINVAL, "dusp_session_state_import: not a state block");
  if (h.n_groups != (long long)s->groups.size() || h.clock < 0 || h.clock % DUSP_CHUNK != 0)
    return fail(DUSP_EINVAL, "dusp_session_state_import: the block comes from a session of another shape");
  for (auto &g : s->groups) {
    StateShape sh, want = state_shape(g);
    memcpy(&sh, p, sizeof sh); p += sizeof sh;
    if (sh.words != want.words || sh.rings != want.rings || sh.stale != want.stale || sh.voices != want.voices)
      return fail(DUSP_EINVAL, "dusp_session_state_import: the block comes from a session of another shape (programs or voice counts differ)");
  }
  cudaStream_t st = (cudaStream_t)cuda_stream;
  CUDA_TRY(cudaSetDevice(s->device));
  for (auto &g : s->groups) {
    const StateShape sh = state_shape(g);
    CUDA_TRY(cudaMemcpyAsync(g.d_words, p, (size_t)sh.words * sizeof(double), cudaMemcpyHostToDevice, st)); p += (size_t)sh.words * sizeof(double);
    CUDA_TRY(cudaMemcpyAsync(g.d_rings, p, (size_t)sh.rings * sizeof(float), cudaMemcpyHostToDevice, st)); p += (size_t)sh.rings * sizeof(float);
    CUDA_TRY(cudaMemcpyAsync(g.d_stale, p, (size_t)sh.stale * sizeof(float), cudaMemcpyHostToDevice, st)); p += (size_t)sh.stale * sizeof(float);
    g.dirty = false;
  }
  CUDA_TRY(cudaStreamSynchronize(st));  // the caller's buffer is pageable in general: done with it on return
  s->clock = h.clock;
  return DUSP_OK;
}

// Enqueue the kernels that render the next `n` samples of every group into d_out on stream `st`.  Rows of d_out
// are `pitch` floats apart (>= n); per-voice output has `nch` channel rows per voice (>= the session's own count:
// a run whose later epochs have more channels, run.cc).
static int enqueue_block(dusp_session *s, long long n, int out_mode, float *d_out, cudaStream_t st, long long pitch, int nch) {
  const int raw = out_mode == DUSP_OUT_VOICES_RAW;
  if (raw) out_mode = DUSP_OUT_VOICES;
  const size_t out_rows = out_mode == DUSP_OUT_VOICES ? (size_t)s->n_voices * nch : 2;
  // channels a group does not produce (and the mix accumulator) start at zero
  bool need_zero = out_mode == DUSP_OUT_MIX;
  for (auto &g : s->groups) if (g.prog.out_nch != nch) need_zero = true;
  if (need_zero) {
    if (pitch == n) CUDA_TRY(cudaMemsetAsync(d_out, 0, out_rows * (size_t)n * sizeof(float), st));
    else CUDA_TRY(cudaMemset2DAsync(d_out, (size_t)pitch * sizeof(float), 0, (size_t)n * sizeof(float), out_rows, st));
  }
  if (int e = upload_initial_state(s, st)) return e;
  if (s->n_multi) {  // every one-voice-per-warp interpreter group in one launch
    DuspKernelParams kp;
    memset(&kp, 0, sizeof kp);
    kp.multi = s->d_multi; kp.n_voices = s->n_multi; kp.n_slots = s->multi_slots; kp.nv_stride = 1;
    kp.tables = s->d_tables; kp.out = d_out; kp.out_nch = nch; kp.prog_nch = 1; kp.out_mode = out_mode; kp.out_raw = raw;
    kp.out_len = pitch; kp.clock0 = s->clock; kp.n_samples = n;
    // samples per lane / warps per CTA: 4 / 1 keeps a warp's slots at 15 KB for the W4 circuits, so ~14 circuits are
    // resident per SM (measured, 4000 circuits x 4 s: 689 ms against 1119 ms at 8 / 2); DUSP_MULTI_B / _WPC override
    int mb = 4;
    if (const char *e = getenv("DUSP_MULTI_B")) mb = atoi(e) == 8 ? 8 : 4;
    int wpc = 1;
    if (const char *e = getenv("DUSP_MULTI_WPC")) wpc = std::max(1, std::min(4, atoi(e)));
    while (wpc > 1 && (size_t)s->multi_slots * mb * 128 * wpc > 100 * 1024) wpc >>= 1;
    kp.density = 8;  // 64 registers per thread, 32 circuits per SM now that the slots no longer bound it (W4: 1134 ms against
                     // 1200 at 80 registers and 1302 uncapped); DUSP_MULTI_DENSITY=0 selects the uncapped kernel
    if (const char *e = getenv("DUSP_MULTI_DENSITY")) kp.density = atoi(e);
    // diagnostics: DUSP_MULTI_DEBUG_FILE=<path> appends "launch warp output_row instructions ticks" for every circuit
    const char *dbg_file = getenv("DUSP_MULTI_DEBUG_FILE");
    if (dbg_file && !s->d_dbg) CUDA_TRY(cudaMalloc(&s->d_dbg, (size_t)s->n_multi * sizeof(long long)));
    kp.dbg_cycles = dbg_file ? s->d_dbg : nullptr;
    int grid = 0, smem = 0;
    int e = dusp_launch_render(&kp, mb, 32, wpc, st, &grid, &smem);
    if (e != 0) return fail(DUSP_ECUDA, std::string("render launch: ") + cudaGetErrorString((cudaError_t)e));
    if (dbg_file) {
      std::vector<long long> t(s->n_multi);
      std::vector<DuspWarpDesc> d(s->n_multi);
      CUDA_TRY(cudaStreamSynchronize(st));
      CUDA_TRY(cudaMemcpy(t.data(), s->d_dbg, t.size() * sizeof(long long), cudaMemcpyDeviceToHost));
      CUDA_TRY(cudaMemcpy(d.data(), s->d_multi, d.size() * sizeof(DuspWarpDesc), cudaMemcpyDeviceToHost));
      if (FILE *f = fopen(dbg_file, "a")) {
        for (int i = 0; i < s->n_multi; i++) fprintf(f, "%d %d %lld %d %lld\n", s->launches, i, (long long)d[i].out_row, d[i].n_instr, t[i]);
        fclose(f);
      }
    }
    s->launches++;
    s->last_grid = grid; s->last_block = wpc * 32; s->last_smem = smem; s->last_fused = false;
    for (auto &g : s->groups) if (g.in_multi) g.sub_block = mb;
  }
  // fork (opt-in, DUSP_CONCURRENT_GROUPS=1): groups after the first launch on side streams.  Measured on W5 (6 250 FM
  // voices + 6 250 filter/delay voices): 42.1 ms per step against 38.2 ms in sequence -- the latency-bound pipelined
  // kernel slows down more under the co-running kernel than the overlap wins, so the default is one stream.
  int n_own = 0;
  for (auto &g : s->groups) if (!g.in_multi) n_own++;
  const bool fork = n_own >= 2 && n_own <= 8 && getenv("DUSP_CONCURRENT_GROUPS");
  if (fork) {
    if (!s->ev_fork) CUDA_TRY(cudaEventCreateWithFlags(&s->ev_fork, cudaEventDisableTiming));
    while ((int)s->side_streams.size() < n_own - 1) {
      cudaStream_t t; cudaEvent_t e;
      CUDA_TRY(cudaStreamCreateWithFlags(&t, cudaStreamNonBlocking));
      CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
      s->side_streams.push_back(t); s->side_done.push_back(e);
    }
    CUDA_TRY(cudaEventRecord(s->ev_fork, st));
  }
  const cudaStream_t st_main = st;
  int k_own = 0;
  for (auto &g : s->groups) {
    if (g.in_multi) continue;
    const int k_side = fork ? k_own - 1 : -1;
    k_own++;
    if (k_side >= 0) { st = s->side_streams[k_side]; CUDA_TRY(cudaStreamWaitEvent(st, s->ev_fork, 0)); }
    else st = st_main;
    DuspKernelParams kp;
    memset(&kp, 0, sizeof kp);
    kp.prog_global = g.d_code; kp.n_instr = (int)g.dev_code.size(); kp.n_slots = g.prog.n_slots;
    kp.use_inline = g.dev_code.size() <= DUSP_INLINE_INSTR;
    if (kp.use_inline) memcpy(kp.prog_inline, g.dev_code.data(), g.dev_code.size() * sizeof(DuspDevInstr));
    kp.n_voices = g.n_voices; kp.voice0 = g.voice0; kp.nv_stride = g.nv_pad;
    kp.words = g.d_words; kp.rings = g.d_rings; kp.ring_stride = std::max<long long>(g.prog.ring_floats, 4);
    kp.ring_tab = g.d_ringtab; kp.stale = g.d_stale; kp.n_stale = g.prog.n_stale; kp.tables = s->d_tables;
    kp.ctab = g.d_ctab;
    kp.out = d_out; kp.out_nch = nch; kp.prog_nch = g.prog.out_nch; kp.out_mode = out_mode; kp.out_raw = raw;
    kp.out_len = pitch; kp.clock0 = s->clock; kp.n_samples = n;
    const int wpc = g.warps_per_cta;
    int grid = 0, smem = 0;
    FusedKernel *kernel = nullptr;
    if (g.fused) {  // the kernel compiled for this output mode (NVRTC on first use of a mode, cached per process)
      const int m = raw ? 2 : out_mode == DUSP_OUT_MIX ? 1 : 0;
      if (!g.fused_mode[0]) g.fused_mode[0] = g.fused;
      if (!g.fused_mode[m]) {
        std::string err;
        PipedInfo pi;
        g.fused_mode[m] = g.piped ? piped_get(g.prog, g.piped_vc, &pi, err, m) : fused_get(g.prog, g.sub_block, g.time_lanes, g.fused_wps, err, m);
        if (!g.fused_mode[m]) return fail(DUSP_ECUDA, "circuit specialisation for this output mode failed: " + err);
      }
      kernel = g.fused_mode[m];
    }
    if (g.fused) {
      FusedParams fp;
      memset(&fp, 0, sizeof fp);
      fp.n_voices = kp.n_voices; fp.voice0 = kp.voice0; fp.nv_stride = kp.nv_stride; fp.words = kp.words;
      fp.rings = kp.rings; fp.ring_stride = kp.ring_stride; fp.stale = kp.stale; fp.tables = kp.tables; fp.out = kp.out;
      fp.n_stale = kp.n_stale; fp.out_nch = kp.out_nch; fp.prog_nch = kp.prog_nch; fp.out_mode = kp.out_mode; fp.out_raw = kp.out_raw;
      fp.out_len = kp.out_len; fp.clock0 = kp.clock0; fp.n_samples = kp.n_samples; fp.sm_slots = s->d_sm_slots;
      std::string err;
      if (g.piped) {
        grid = (g.n_voices + g.pinfo.voices_per_cta - 1) / g.pinfo.voices_per_cta;
        smem = g.pinfo.smem_bytes;
        if (fused_launch(kernel, fp, grid, g.pinfo.threads, smem, st, err) != 0) return fail(DUSP_ECUDA, err);
      } else {
        const int V = 32 / g.time_lanes;
        const int n_warps = (g.n_voices + V - 1) / V;
        const long long n_tiles = (n + DUSP_CHUNK - 1) / DUSP_CHUNK;
        long long seg_tiles = g.seg_max >= 2 ? (n_tiles + g.seg_max - 1) / g.seg_max : n_tiles;
        const int seg_n = (int)((n_tiles + seg_tiles - 1) / std::max<long long>(seg_tiles, 1));
        if (g.seg_max >= 2 && seg_n >= 2) {  // a voice's timeline of this launch on seg_n warps
          const int m = raw ? 2 : out_mode == DUSP_OUT_MIX ? 1 : 0;
          if (!g.seg_render[m]) {
            std::string scan_src, render_src;
            generate_segmented_sources(g.prog, g.seg_wps, m, scan_src, render_src, &g.seg_nosc);
            g.seg_render[m] = fused_build(render_src, "dusp_fused", err);
            if (!g.seg_render[m]) return fail(DUSP_ECUDA, "segmented kernel for this output mode failed: " + err);
          }
          fp.seg_totals = g.d_seg_totals; fp.seg_flags = g.d_seg_flags; fp.seg_tiles = seg_tiles; fp.seg_n = seg_n;
          CUDA_TRY(cudaMemsetAsync(g.d_seg_flags, 0, (size_t)g.n_voices * sizeof(int), st));
          grid = g.n_voices * seg_n;
          if (fused_launch(g.seg_scan, fp, grid, 32, 0, st, err) != 0) return fail(DUSP_ECUDA, err);
          if (fused_launch(g.seg_render[m], fp, grid, 32, 0, st, err) != 0) return fail(DUSP_ECUDA, err);
          s->launches++;  // (the scan; the render is counted below)
        } else {
          grid = (n_warps + wpc - 1) / wpc;
          if (fused_launch(kernel, fp, grid, wpc * 32, 0, st, err) != 0) return fail(DUSP_ECUDA, err);
        }
      }
    } else {
      int e = dusp_launch_render(&kp, g.sub_block, g.time_lanes, wpc, st, &grid, &smem);
      if (e != 0) return fail(DUSP_ECUDA, std::string("render launch: ") + cudaGetErrorString((cudaError_t)e));
    }
    s->last_fused = g.fused != nullptr;
    s->launches++;
    s->last_grid = grid; s->last_block = g.piped ? g.pinfo.threads : wpc * 32; s->last_smem = smem;
    if (k_side >= 0) CUDA_TRY(cudaEventRecord(s->side_done[k_side], st));
  }
  st = st_main;
  if (fork) for (int k = 0; k < n_own - 1; k++) CUDA_TRY(cudaStreamWaitEvent(st, s->side_done[k], 0));  // join
  s->clock += (n + DUSP_CHUNK - 1) / DUSP_CHUNK * DUSP_CHUNK;
  return DUSP_OK;
}

// dusp_session_render with the destination's row pitch (floats) and per-voice channel stride made explicit: a run
// (run.cc) renders chunk-aligned segments side by side into one caller buffer.
int dusp_render_impl(dusp_session *s, long long n_samples, int out_mode, float *out_host, float *out_device, void *cuda_stream,
                     long long pitch, int nch) {
  if (!s || n_samples < 0 || (out_mode != DUSP_OUT_VOICES && out_mode != DUSP_OUT_MIX && out_mode != DUSP_OUT_VOICES_RAW) || (!out_host == !out_device))
    return fail(DUSP_EINVAL, "dusp_session_render: bad arguments");
  if (pitch < n_samples || nch < s->out_nch) return fail(DUSP_EINVAL, "dusp_session_render: destination too narrow");
  CUDA_TRY(cudaSetDevice(s->device));
  cudaStream_t st = (cudaStream_t)cuda_stream;
  if (n_samples == 0) return DUSP_OK;
  if (out_device) return enqueue_block(s, n_samples, out_mode, out_device, st, pitch, nch);

  // Host output: render time blocks into two device staging buffers and drain each
  // over PCIe on a second stream while the next block renders.
  const size_t rows = out_mode != DUSP_OUT_MIX ? (size_t)s->n_voices * nch : 2;
  size_t staging_mb = 1024;  // per staging buffer (measured on B200: 1 GB blocks drain at 50 GB/s, 256 MB at 46); DUSP_STAGING_MB overrides (longer rows per 2-D copy)
  if (const char *e = getenv("DUSP_STAGING_MB")) staging_mb = (size_t)std::max(16, atoi(e));
  long long blk = (long long)((staging_mb << 20) / (rows * sizeof(float)));
  blk = std::max<long long>(DUSP_CHUNK, blk / DUSP_CHUNK * DUSP_CHUNK);
  blk = std::min<long long>(blk, (n_samples + DUSP_CHUNK - 1) / DUSP_CHUNK * DUSP_CHUNK);
  const size_t need = rows * (size_t)blk;
  if (s->staging_floats < need) {
    for (int i = 0; i < 2; i++) { cudaFree(s->d_staging[i]); s->d_staging[i] = nullptr; }
    s->staging_floats = 0;
    for (int i = 0; i < 2; i++)
      if (cudaMalloc(&s->d_staging[i], need * sizeof(float)) != cudaSuccess) {
        cudaGetLastError();
        return fail(DUSP_ENOMEM, "not enough device memory for the output staging blocks");
      }
    s->staging_floats = need;
  }
  if (!s->copy_stream) {
    CUDA_TRY(cudaStreamCreateWithFlags(&s->copy_stream, cudaStreamNonBlocking));
    for (int i = 0; i < 2; i++) {
      CUDA_TRY(cudaEventCreateWithFlags(&s->ev_rendered[i], cudaEventDisableTiming));
      CUDA_TRY(cudaEventCreateWithFlags(&s->ev_copied[i], cudaEventDisableTiming));
    }
  }
  int i = 0;
  for (long long t = 0; t < n_samples; t += blk, i ^= 1) {
    const long long n = std::min(blk, n_samples - t);
    CUDA_TRY(cudaStreamWaitEvent(st, s->ev_copied[i], 0));  // buffer i drained (no-op the first time)
    int rc = enqueue_block(s, n, out_mode, s->d_staging[i], st, n, nch);
    if (rc != DUSP_OK) return rc;
    CUDA_TRY(cudaEventRecord(s->ev_rendered[i], st));
    CUDA_TRY(cudaStreamWaitEvent(s->copy_stream, s->ev_rendered[i], 0));
    CUDA_TRY(cudaMemcpy2DAsync(out_host + t, (size_t)pitch * sizeof(float), s->d_staging[i], (size_t)n * sizeof(float),
                               (size_t)n * sizeof(float), rows, cudaMemcpyDeviceToHost, s->copy_stream));
    CUDA_TRY(cudaEventRecord(s->ev_copied[i], s->copy_stream));
  }
  CUDA_TRY(cudaStreamSynchronize(s->copy_stream));
  CUDA_TRY(cudaStreamSynchronize(st));
  return DUSP_OK;
}

int dusp_session_render(dusp_session *s, long long n_samples, int out_mode, float *out_host, float *out_device,
                        void *cuda_stream) {
  if (!s) return fail(DUSP_EINVAL, "dusp_session_render: bad arguments");
  return dusp_render_impl(s, n_samples, out_mode, out_host, out_device, cuda_stream, n_samples, s->out_nch);
}

long long dusp_session_clock(const dusp_session *s) { return s ? s->clock : 0; }

int dusp_session_tile_info(const dusp_session *s, int *sub_block, int *time_lanes, int *specialised) {
  if (!s || s->groups.empty()) return fail(DUSP_EINVAL, "dusp_session_tile_info: empty session");
  const Group &g = s->groups[0];
  if (sub_block) *sub_block = g.sub_block;
  if (time_lanes) *time_lanes = g.time_lanes;
  if (specialised) *specialised = g.fused != nullptr;
  return DUSP_OK;
}
const char *dusp_session_warning(const dusp_session *s) { return s ? s->warning.c_str() : ""; }

size_t dusp_program_specialise(const dusp_program *p, int sub_block, int time_lanes, char *src_buf, size_t cap) {
  std::string err, src;
  int wps = 8;
  if (const char *we = getenv("DUSP_FUSED_WPS")) wps = std::max(1, atoi(we));
  size_t n;
  const char *pe = getenv("DUSP_PIPED");
  if (pe && atoi(pe) != 0) {  // the pipelined kernel instead (time_lanes is then the voices per CTA)
    PipedInfo info;
    n = piped_compile_only(p->prog, time_lanes, &info, err, &src);
  } else if (getenv("DUSP_SPECIALISE_SEGMENTED")) {  // the two segmented kernels (build checks): scan, then render
    std::string scan_src, render_src;
    int nosc = 0;
    if (!generate_segmented_sources(p->prog, wps, 0, scan_src, render_src, &nosc)) { fail(DUSP_EUNSUPPORTED, "the circuit has no segmented form"); return 0; }
    n = fused_build_cubin(scan_src, err);
    if (n) n = fused_build_cubin(render_src, err);
    src = scan_src + "\n// ======== render ========\n" + render_src;
  } else {
    n = fused_compile_only(p->prog, sub_block, time_lanes, wps, err, &src);
  }
  if (src_buf && cap) { size_t m = std::min(cap - 1, src.size()); memcpy(src_buf, src.data(), m); src_buf[m] = 0; }
  if (!n) fail(DUSP_ECUDA, err);
  return n;
}

double dusp_rand(uint64_t seed, uint64_t k) {
  uint32_t x = (uint32_t)k + (uint32_t)seed;
  x ^= x >> 16; x *= 0x7feb352dU;
  x ^= x >> 15; x *= 0x846ca68bU;
  x ^= x >> 16;
  x ^= (uint32_t)(seed >> 32);
  return (double)x * (1.0 / 4294967296.0);
}

}  // extern "C"
