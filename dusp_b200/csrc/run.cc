// run.cc -- a render driven through its circuits' timelines, behind the C ABI (include/dusp_b200.h, dusp_run_*).
//
// The reference runs an event's function before the tick of the chunk that contains its time (Circuit.js:23,57-65,
// Event.js:18-30); events trigger envelopes, assign unit state, or edit the graph (`then`, Unit.js:59-67).  The host
// shim (bindings/node/index.js, dusp_b200/render.py) runs the circuit's event queue ONCE on the reference's own
// objects and writes what happened into the graph blob:
//   EV <chunk> <unit> <prop> <value>   numeric unit state assigned by an event, applied before that chunk's tick
//   EPOCH <chunk> + DUSPGRAPH section  the circuit after an event edited the graph, valid from that chunk on
//   KEEP                               an edit follows: the program keeps every outlet's last chunk for its successor
//   WARMUP <k>                         compile the circuit as its k-th tick sees it (channel counts still growing)
// Everything that follows from those records lives here, once, for every host language: splitting the blob into
// epochs, warm-up programs, chunk-aligned render segments, pokes between segments (dusp_session_poke), and a new
// session that adopts the old one's device state at every epoch (dusp_session_adopt).
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <set>
#include <sstream>
#include <string>
#include <vector>

#include "../../include/dusp_b200.h"

namespace dusp { int fail_public(int code, const std::string &msg); }
extern "C" int dusp_render_impl(dusp_session *s, long long n_samples, int out_mode, float *out_host, float *out_device,
                                void *cuda_stream, long long pitch, int nch);

namespace {

struct Ev { long long chunk; int unit; std::string prop; double value; };

struct Epoch {
  long long k0 = 0;        // first chunk this section is valid for
  std::string text;        // one DUSPGRAPH section
  std::vector<Ev> evs;     // its EV records
  dusp_program *prog = nullptr;
};

std::vector<std::string> lines_of(const std::string &s) {
  std::vector<std::string> out;
  std::istringstream in(s);
  std::string l;
  while (std::getline(in, l)) out.push_back(l);
  return out;
}

// [(start chunk, section text)]: one entry unless events edit the graph
std::vector<Epoch> split_epochs(const std::string &blob) {
  std::vector<Epoch> out;
  Epoch cur;
  for (const std::string &l : lines_of(blob)) {
    if (l.compare(0, 6, "EPOCH ") == 0) {
      out.push_back(cur);
      cur = Epoch();
      cur.k0 = atoll(l.c_str() + 6);
    } else {
      cur.text += l;
      cur.text += '\n';
    }
  }
  out.push_back(cur);
  return out;
}

std::vector<Ev> events_of(const std::string &text) {
  std::vector<Ev> ev;
  for (const std::string &l : lines_of(text)) {
    if (l.compare(0, 3, "EV ") != 0) continue;
    std::istringstream ss(l);
    std::string tag, prop, val;
    Ev e;
    if (!(ss >> tag >> e.chunk >> e.unit >> prop >> val)) continue;
    e.prop = prop;
    e.value = strtod(val.c_str(), nullptr);
    ev.push_back(e);
  }
  return ev;
}

bool has_record(const std::string &text, const char *rec) {
  for (const std::string &l : lines_of(text)) if (l == rec) return true;
  return false;
}

std::string with_records(const std::string &text, const std::vector<std::string> &records) {
  const size_t at = text.rfind("END\n");
  std::string head = at == std::string::npos ? text : text.substr(0, at);
  for (const std::string &r : records) { head += r; head += '\n'; }
  return head + "END\n";
}

}  // namespace

struct dusp_run {
  int device = 0;
  std::vector<int> voices;
  std::vector<std::vector<Epoch>> tl;  // per program group: its epochs
  std::vector<int> epoch;              // current epoch per group
  dusp_session *session = nullptr;
  long long pos = 0;                   // samples rendered so far
  int nch = 0;                         // channels of the widest program of any epoch (row layout of every render)
  int nch_seen = 0;                    // ... of the epochs entered so far (what the reference's result would hold)
  int n_voices = 0;
};

namespace {

int make_session(dusp_run *r, dusp_session **out) {
  std::vector<const dusp_program *> progs;
  for (size_t i = 0; i < r->tl.size(); i++) progs.push_back(r->tl[i][r->epoch[i]].prog);
  return dusp_session_create(progs.data(), r->voices.data(), (int)progs.size(), r->device, out);
}

// What the reference does in runEvents before the tick of the chunk at `pos`: graph edits, then state assignments.
int enter(dusp_run *r, long long pos, void *stream) {
  const long long k = pos / DUSP_CHUNK;
  bool sw = false;
  for (size_t i = 0; i < r->tl.size(); i++)
    while (r->epoch[i] + 1 < (int)r->tl[i].size() && r->tl[i][r->epoch[i] + 1].k0 <= k) { r->epoch[i]++; sw = true; }
  if (sw) {
    dusp_session *next = nullptr;
    int rc = make_session(r, &next);
    if (rc != DUSP_OK) return rc;
    for (size_t i = 0; i < r->tl.size(); i++) {
      rc = dusp_session_adopt(next, (int)i, r->session, (int)i, stream);
      if (rc != DUSP_OK) { dusp_session_destroy(next); return rc; }
    }
    dusp_session_destroy(r->session);
    r->session = next;
  }
  for (size_t i = 0; i < r->tl.size(); i++) r->nch_seen = std::max(r->nch_seen, dusp_program_num_channels(r->tl[i][r->epoch[i]].prog));
  if (pos > 0)
    for (size_t i = 0; i < r->tl.size(); i++) {
      const Epoch &e = r->tl[i][r->epoch[i]];
      std::vector<int> ids;
      std::vector<double> vals;
      for (const Ev &ev : e.evs) {
        if (ev.chunk != k) continue;
        const std::string name = "u" + std::to_string(ev.unit) + "." + ev.prop;
        if (name.size() > 3 && name.compare(name.size() - 3, 3, "[*]") == 0) {  // every channel the program holds
          const std::string prefix = name.substr(0, name.size() - 2);
          const int np = dusp_program_num_params(e.prog);
          for (int j = 0; j < np; j++)
            if (strncmp(dusp_program_param_name(e.prog, j), prefix.c_str(), prefix.size()) == 0) { ids.push_back(j); vals.push_back(ev.value); }
          continue;
        }
        const int id = dusp_program_param_index(e.prog, name.c_str());
        if (id < 0) return dusp::fail_public(DUSP_EUNSUPPORTED, "event assigns state the compiled program does not hold: no such parameter: " + name);
        ids.push_back(id);
        vals.push_back(ev.value);
      }
      if (!ids.empty()) {
        int rc = dusp_session_poke(r->session, (int)i, -1, (int)ids.size(), ids.data(), vals.data(), stream);
        if (rc != DUSP_OK) return rc;
      }
    }
  return DUSP_OK;
}

}  // namespace

extern "C" {

void dusp_run_destroy(dusp_run *r) {
  if (!r) return;
  if (r->session) dusp_session_destroy(r->session);
  for (auto &g : r->tl) for (auto &e : g) if (e.prog) dusp_program_destroy(e.prog);
  delete r;
}

int dusp_run_create(const char *const *blobs, const size_t *lens, const int *voices, int n_programs, int device, dusp_run **out) {
  if (!blobs || !lens || !voices || n_programs < 1 || !out) return dusp::fail_public(DUSP_EINVAL, "dusp_run_create: bad arguments");
  dusp_run *r = new dusp_run;
  r->device = device;
  auto bail = [&](int rc) { dusp_run_destroy(r); return rc; };
  for (int i = 0; i < n_programs; i++) {
    if (!blobs[i] || voices[i] < 1) return bail(dusp::fail_public(DUSP_EINVAL, "dusp_run_create: empty program group"));
    r->voices.push_back(voices[i]);
    r->n_voices += voices[i];
    r->tl.emplace_back();
    r->epoch.push_back(0);
    std::vector<Epoch> &tl = r->tl.back();  // (owned by the run from here on: a failure below frees what was compiled)
    tl = split_epochs(std::string(blobs[i], lens[i]));
    // Channel growth at the start of a render: the reference's tick order is not topological (Unit.js:171-209), so a
    // multi-channel signal can take S ticks to reach every unit.  Tick k < S is compiled from a `WARMUP k` record;
    // those programs become one-chunk epochs in front of the steady one, each keeping every outlet's last chunk.
    {
      dusp_program *p = nullptr;
      int rc = dusp_program_create(tl[0].text.data(), tl[0].text.size(), &p);
      if (rc == DUSP_OK) {
        tl[0].prog = p;
      } else if (rc == DUSP_EUNSUPPORTED && strstr(dusp_last_error(), "warm-up")) {
        const std::string text = tl[0].text;
        const std::string probe = with_records(text, {"WARMUP 0"});
        rc = dusp_program_create(probe.data(), probe.size(), &p);
        if (rc != DUSP_OK) return bail(rc);
        const int S = dusp_program_warmup_ticks(p);
        dusp_program_destroy(p);
        if (tl.size() > 1 && tl[1].k0 <= S) return bail(dusp::fail_public(DUSP_EUNSUPPORTED, "graph edit while channel counts are still growing"));
        const bool keeps = has_record(text, "KEEP");
        std::vector<Epoch> warm;
        for (int k = 0; k <= S; k++) {
          Epoch e;
          e.k0 = k;
          std::vector<std::string> recs = {"WARMUP " + std::to_string(k)};
          if (k < S && !keeps) recs.push_back("KEEP");
          e.text = with_records(text, recs);
          warm.push_back(e);
        }
        warm.insert(warm.end(), tl.begin() + 1, tl.end());
        tl.swap(warm);
      } else {
        return bail(rc);
      }
    }
    for (Epoch &e : tl) {
      if (!e.prog) {
        int rc = dusp_program_create(e.text.data(), e.text.size(), &e.prog);
        if (rc != DUSP_OK) return bail(rc);
      }
      e.evs = events_of(e.text);
      r->nch = std::max(r->nch, dusp_program_num_channels(e.prog));
    }
  }
  *out = r;  // (the session -- the first thing that needs a CUDA device -- is created on first use)
  return DUSP_OK;
}

static int ensure_session(dusp_run *r) { return r->session ? DUSP_OK : make_session(r, &r->session); }

int dusp_run_num_voices(const dusp_run *r) { return r ? r->n_voices : 0; }
int dusp_run_num_channels(const dusp_run *r) { return r ? r->nch : 0; }
int dusp_run_channels_rendered(const dusp_run *r) { return r ? r->nch_seen : 0; }
long long dusp_run_position(const dusp_run *r) { return r ? r->pos : 0; }
dusp_session *dusp_run_session(dusp_run *r) { return r && ensure_session(r) == DUSP_OK ? r->session : nullptr; }
int dusp_run_num_epochs(const dusp_run *r, int program_index) {
  return r && program_index >= 0 && program_index < (int)r->tl.size() ? (int)r->tl[program_index].size() : 0;
}
const dusp_program *dusp_run_epoch_program(const dusp_run *r, int program_index, int epoch, long long *first_chunk) {
  if (epoch < 0 || epoch >= dusp_run_num_epochs(r, program_index)) return nullptr;
  if (first_chunk) *first_chunk = r->tl[program_index][epoch].k0;
  return r->tl[program_index][epoch].prog;
}
const dusp_program *dusp_run_program(const dusp_run *r, int program_index) {
  if (!r || program_index < 0 || program_index >= (int)r->tl.size()) return nullptr;
  return r->tl[program_index][r->epoch[program_index]].prog;
}

int dusp_run_set_params(dusp_run *r, int program_index, const char *const *names, int n_names, const double *values) {
  if (!r || program_index < 0 || program_index >= (int)r->tl.size() || n_names < 0 || (n_names && (!names || !values)))
    return dusp::fail_public(DUSP_EINVAL, "dusp_run_set_params: bad arguments");
  if (r->pos != 0) return dusp::fail_public(DUSP_EINVAL, "dusp_run_set_params: the run has rendered");
  int rc0 = ensure_session(r);
  if (rc0 != DUSP_OK) return rc0;
  const dusp_program *p = dusp_run_program(r, program_index);
  std::vector<int> ids;
  for (int k = 0; k < n_names; k++) {
    const int id = names[k] ? dusp_program_param_index(p, names[k]) : -1;
    if (id < 0) return dusp::fail_public(DUSP_EINVAL, std::string("no such parameter: ") + (names[k] ? names[k] : "(null)"));
    ids.push_back(id);
  }
  return dusp_session_set_params(r->session, program_index, ids.data(), n_names, values);
}

int dusp_run_render(dusp_run *r, long long n_samples, int out_mode, float *out_host, float *out_device, void *cuda_stream) {
  if (!r || n_samples < 0 || (!out_host == !out_device)) return dusp::fail_public(DUSP_EINVAL, "dusp_run_render: bad arguments");
  if (r->pos % DUSP_CHUNK) return dusp::fail_public(DUSP_EINVAL, "a render that stopped inside a chunk cannot be continued");
  if (n_samples == 0) return DUSP_OK;
  int rc0 = ensure_session(r);
  if (rc0 != DUSP_OK) return rc0;
  // segment boundaries: every later epoch's first chunk and every event's chunk inside (pos, pos + n)
  std::set<long long> cuts;
  for (auto &g : r->tl)
    for (size_t e = 0; e < g.size(); e++) {
      if (e > 0) cuts.insert(g[e].k0 * DUSP_CHUNK);
      for (const Ev &ev : g[e].evs) cuts.insert(ev.chunk * DUSP_CHUNK);
    }
  std::vector<long long> ends;
  for (long long c : cuts) if (c > r->pos && c < r->pos + n_samples) ends.push_back(c);
  ends.push_back(r->pos + n_samples);
  const long long pos0 = r->pos;
  // (a segment whose programs have fewer channels than the widest epoch zero-fills the rest: enqueue_block)
  for (long long b : ends) {
    int rc = enter(r, r->pos, cuda_stream);
    if (rc != DUSP_OK) return rc;
    const long long off = r->pos - pos0;
    rc = dusp_render_impl(r->session, b - r->pos, out_mode, out_host ? out_host + off : nullptr, out_device ? out_device + off : nullptr,
                          cuda_stream, n_samples, r->nch);
    if (rc != DUSP_OK) return rc;
    r->pos = b;
  }
  return DUSP_OK;
}

}  // extern "C"
