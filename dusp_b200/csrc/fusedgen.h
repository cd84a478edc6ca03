// fusedgen.h -- internal: the per-instruction code emitter shared by the two kernel generators
// (fused.cc: one warp = V voices x T time-lanes;  piped.cc: pipelined CTA with a lane-per-voice warp).
#pragma once
#include <set>
#include <sstream>
#include <string>

#include "compile.h"

namespace dusp {

struct Gen {
  const Program &p;
  int B, T;
  std::ostringstream o;
  std::set<int> mutable_words;
  mutable std::set<int> used_words;
  std::set<size_t> hoisted;  // instructions whose loads were issued early into `pre<pc>[B]` (prefetch())
  Gen(const Program &pr, int b, int t) : p(pr), B(b), T(t) {}

  std::string w(int i) const { used_words.insert(i); return "w" + std::to_string(i); }
  void mut(int i) { mutable_words.insert(i); }
  // materialise operand `ref` as a fresh register array named `name`
  void load(const std::string &name, int ref) {
    o << "      float " << name << "[B]; ";
    if (ref >= 0) o << "CP(" << name << ", s" << ref << ")\n";
    else if (ref == REF_NONE) o << "FILL(" << name << ", 0.0f)\n";
    else o << "FILL(" << name << ", (float)" << w(REF_CONST_BASE - ref) << ")\n";
  }
  void store(const std::string &name, int ref) {
    if (ref >= 0) o << "      CP(s" << ref << ", " << name << ")\n";
  }
  std::string ring_ptr(int r) const { return "(rings + " + std::to_string((long long)p.rings[r].offset) + "LL)"; }
  std::string ring_len(int r) const { return std::to_string((long long)p.rings[r].length); }
  std::string table(int t) const { return "(tables + " + std::to_string((long long)t * DUSP_TABLE_LEN) + ")"; }

  void map1(const DuspInstr &I, const char *expr) {
    load("a", I.in[0]);
    o << "      UNROLL for (int i_ = 0; i_ < B; i_++) { const float x = a[i_]; a[i_] = " << expr << "; }\n";
    store("a", I.out[0]);
  }
  void map2(const DuspInstr &I, const char *expr) {
    load("a", I.in[0]); load("b", I.in[1]);
    o << "      UNROLL for (int i_ = 0; i_ < B; i_++) { const float x = a[i_], y = b[i_]; a[i_] = " << expr << "; }\n";
    store("a", I.out[0]);
  }
  void instr(size_t pc, const DuspInstr &I);  // fused.cc
  static bool can_prefetch(const Program &p, const DuspInstr &I, int step);
  void prefetch(size_t pc, const DuspInstr &I, std::ostream &out);  // emits the early loads, marks pc hoisted
  std::string source(int min_blocks);          // fused.cc: the V x T tile kernel
};

// shared pieces of the generated text / the NVRTC + module cache (fused.cc)
std::string fused_prelude();  // experiment defines, #include "ops.cuh", FusedParams, CP/FILL/SEED macros
struct FusedKernel;
FusedKernel *fused_build(const std::string &src, const char *entry, std::string &err);  // NVRTC + load on the current device
size_t fused_build_cubin(const std::string &src, std::string &err);                      // NVRTC only

}  // namespace dusp
