// compile.h -- circuit description -> flat op program (host side).
#pragma once
#include <stdexcept>
#include <string>
#include <vector>

#include "graph.h"
#include "vm.h"

namespace dusp {

struct Unsupported : std::runtime_error {
  using std::runtime_error::runtime_error;
};

struct Param {
  std::string name;  // "u<unit>.<inlet>[<ch>]" (inlet constant, f32-valued) or "u<unit>.<prop>"
  double dflt;
  bool f32;          // value is what a Float32Array chunk holds: rounded to binary32 on upload
};

// Two CircleBuffer taps of different units on one ring: their per-voice OFFSET
// constants must keep them either on the same sample or >= one chunk apart
// (see compile.cc); re-checked whenever a voice overrides an offset.
struct TapPair {
  int word_a, word_b;    // OFFSET constant words
  int sign_a, sign_b;    // +1 writer (t + SR*off), -1 reader (t - SR*off)
  long long ring_len;
  bool strict = false;   // a plain reader against a writer, a chunk or more apart by default: must STAY apart
                         // (not coincide), so the reader's loads may be issued ahead of the writer's stores
};

// A word whose initial value is a function of other (user-settable) words of the same voice.
enum DerivedKind { DV_FILTER_LP, DV_FILTER_HP, DV_PAN, DV_GAIN, DV_CBIDX };
struct Derived {
  int kind;
  int dst;          // first destination word
  int src0, src1;   // source words (-1 = unused)
  double aux0;      // DV_CBIDX: +1 writer / -1 reader;  aux1: ring length
  double aux1;
};

struct Program {
  std::vector<DuspInstr> code;
  std::vector<Param> words;     // per-voice words: parameters and unit state, in instruction order
  std::vector<DuspRing> rings;  // per-voice delay lines / CircleBuffer channels / frames of the spectral units
  std::vector<double> ctab;     // program constants evaluated on the host with libm (FFT twiddles, window envelopes)
  // identities across graph edits (dusp_session_adopt): "<Kind>:<label>" per blob unit; rings and stale buffers
  // are named "u<unit>.<what>" like the words
  std::vector<std::string> unit_labels, ring_names, stale_names;
  long long ring_floats = 0;
  int n_slots = 0;
  int n_stale = 0;
  int out_nch = 0;
  int n_units = 0;              // units the reference ticks per chunk (metric: unit-samples)
  bool keeps_all = false;       // compiled from a KEEP record: every outlet value it produces leaves its last chunk behind
  int warmup_ticks = 0;         // ticks before the channel counts every unit sees are the steady ones (compile.cc)
  bool chunk_tiles = false;     // has per-chunk control ops (OP_RETRIG ...): must render in whole-chunk tiles (B*T = 256)
  double ring_bytes_per_sample = 0;  // SURVEY 8d accounting
  std::vector<TapPair> tap_pairs;
  std::vector<Derived> derived;
  struct Locked { int word; double lo, hi; };
  struct EqualPair { int word_a, word_b; };
  std::vector<EqualPair> equal_words;  // words a merged op relies on being equal in every voice
  std::vector<Locked> locked;  // constants a specialised op relies on: per-voice overrides must stay in range
  std::string check_locked(const double *word_values) const;
  // returns "" if the offsets (one value per word, as the voice will see them) are admissible
  std::string check_taps(const double *word_values) const;
  std::string listing() const;
};

Program compile(const Graph &g);
void derive_words(const Program &p, double *words, long long stride);  // words[i*stride] = word i of one voice
const char *op_name(int op);

}  // namespace dusp
