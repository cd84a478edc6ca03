// fused.h -- circuit specialisation: one straight-line sm_100a kernel per circuit shape,
// generated from the flat op program and built from the hand-written cores in ops.cuh.
#pragma once
#include <string>

#include "compile.h"

namespace dusp {

// Kernel parameter block of a specialised kernel (mirrored textually in the generated source).
struct FusedParams {
  int n_voices, voice0;
  long long nv_stride;
  double *words;
  float *rings;
  long long ring_stride;
  float *stale;
  const float *tables;
  float *out;
  int n_stale, out_nch, prog_nch, out_mode;
  int out_raw, pad_;
  long long out_len, clock0, n_samples;
  unsigned *sm_slots;  // [256] per-SM arrival counters (pipelined kernel: which warp of a CTA takes the lane-per-voice role)
  // segmented kernels (a voice's timeline of this launch on `seg_n` warps): wrapped totals of every oscillator's
  // increments per segment [voice][seg_n][n_osc], one flag per voice (non-zero: render it in one piece)
  double *seg_totals;
  int *seg_flags;
  long long seg_tiles;  // 256-sample tiles per segment
  int seg_n, seg_pad_;
};

// Segmented rendering (see ops.cuh "segments of one voice's timeline"): "" when the program does not qualify.
// `scan`: the kernel `dusp_fused_scan` that leaves the totals and flags; `render`: `dusp_fused` with the segment logic.
bool generate_segmented_sources(const Program &p, int min_blocks, int out_mode, std::string &scan, std::string &render, int *n_osc);
struct FusedKernel;
FusedKernel *fused_build(const std::string &src, const char *entry, std::string &err);
size_t fused_build_cubin(const std::string &src, std::string &err);  // NVRTC only (no device needed)

// CUDA C++ source of the specialised kernel `dusp_fused` for tile shape (B, T).
std::string generate_fused_source(const Program &p, int B, int T, int min_blocks, int out_mode = 0);
std::string fused_mode_prefix(int out_mode);  // "#define KP_OUT_MODE ...": the output mode as a compile-time constant (< 0: read from the parameters)

struct FusedKernel;  // a loaded module + function (cached per source text, per device)

// Compiles (NVRTC, sm_100a, -fmad=false) and loads the kernel on the current device.  Returns
// nullptr and fills `err` on failure.  Results are cached for the life of the process.
FusedKernel *fused_get(const Program &p, int B, int T, int min_blocks, std::string &err, int out_mode = 0);
// Enqueue; returns 0 or a CUresult/cudaError-like code with `err` filled.
int fused_launch(FusedKernel *k, const FusedParams &kp, int grid, int block, int smem_bytes, void *stream, std::string &err);
// NVRTC only (no device needed): returns the cubin size or 0 with `err` filled.  Used by build checks/tests.
size_t fused_compile_only(const Program &p, int B, int T, int min_blocks, std::string &err, std::string *src_out);

// Pipelined kernel (piped.cc): a CTA renders VC voices; four warps run the time-parallel units with lanes along
// time (coalesced rings / output), a fifth warp runs the time-sequential units (recursive filters ...) one
// voice per lane, one 128-sample step behind, exchanging signals through swizzled shared memory.
struct PipedInfo {
  int voices_per_cta = 0, threads = 0, smem_bytes = 0, n_buffers = 0;
  int n_ops = 0, n_ops_sequential_stage = 0;  // how much of the circuit lands on the one lane-per-voice warp
};
// "" when the program has no admissible 3-stage split (the caller falls back to the tile kernel).
std::string generate_piped_source(const Program &p, int voices_per_cta, PipedInfo *info, int out_mode = 0);
// true when some op has no time-parallel form (recursive filter, Shape, scatter delay ...): only then does the
// pipeline pay; purely time-parallel circuits run fastest on the tile kernel with all 32 lanes along time
bool program_has_sequential(const Program &p);
FusedKernel *piped_get(const Program &p, int voices_per_cta, PipedInfo *info, std::string &err, int out_mode = 0);
size_t piped_compile_only(const Program &p, int voices_per_cta, PipedInfo *info, std::string &err, std::string *src_out);

}  // namespace dusp
