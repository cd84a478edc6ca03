/*
 * dusp_b200.h -- C ABI of the B200-native Dusp circuit renderer (libdusp_b200.so).
 *
 * This is the drop-in boundary for ONE path of the reference: rendering a
 * circuit (or a batch of independent circuits / voices) to channel data,
 *     /root/reference/src/renderChannelData.js:5-49
 *       -> Circuit.tickUntil / tick            src/Circuit.js:19-47
 *       -> every unit's _tick                  src/components/ *.js  
 * The reference has no FFI for this path (it is plain JavaScript); the entry
 * points below are what a Node.js N-API addon binds (bindings/node/, see
 * INTEGRATION.md) and what the Python host mirror (dusp_b200/capi.py, ctypes)
 * calls.  Plain pointers and sizes only; no C++/torch types cross this line.
 *
 * Error convention: every int-returning call returns 0 on success and a
 * negative DUSP_E* code on failure; dusp_last_error() then returns the message
 * (thread-local).  The reference throws *strings* (renderChannelData.js:13,17);
 * host shims re-throw dusp_last_error() verbatim.
 *
 * There is no CPU fallback: without a CUDA device every session call fails
 * with DUSP_ECUDA.
 */
#ifndef DUSP_B200_H
#define DUSP_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DUSP_ABI_VERSION 1
#define DUSP_SAMPLE_RATE 44100 /* src/config.js:7 */
#define DUSP_CHUNK 256         /* src/config.js:6 */

enum {
  DUSP_OK = 0,
  DUSP_EINVAL = -1,      /* bad argument / malformed graph blob */
  DUSP_EUNSUPPORTED = -2,/* graph uses a unit or configuration outside the accelerated path */
  DUSP_ECUDA = -3,       /* CUDA runtime error, or no device */
  DUSP_ENOMEM = -4
};

/* Where dusp_session_render puts the audio. */
enum {
  DUSP_OUT_VOICES = 0, /* per-voice channel data, layout [voice][channel][n_samples] f32
                          == what renderChannelData returns for each circuit (renderChannelData.js:26-45) */
  DUSP_OUT_MIX = 1,    /* stereo mixdown only, layout [2][n_samples] f32:
                          mix[c] = sum over voices of out_v[c % nch_v]  (Sum's modulo broadcast, Sum.js:33-44) */
  DUSP_OUT_VOICES_RAW = 2 /* as DUSP_OUT_VOICES but exactly what outlet.signalChunk.channelData holds (NaN and -0
                          kept): the view of the streaming sink, which throws on NaN (RenderStream.js:38-52) */
};

typedef struct dusp_program dusp_program; /* a circuit compiled to a flat op program */
typedef struct dusp_session dusp_session; /* n voices of one or more programs, resident on one GPU */

int dusp_abi_version(void);
const char *dusp_last_error(void);

/* ---- compile -------------------------------------------------------------
 * Replaces: `new Circuit(outlet.unit)` + the per-chunk unit dispatch loop
 * (renderChannelData.js:20, Circuit.js:34-37, Unit.js:111-119).  `graph_blob`
 * is the "DUSPGRAPH 1" text description of the circuit: units in tick order
 * (circuit.units), their inlets' constants / connections, unit state, shared
 * CircleBuffers and the rendered outlet (format: dusp_b200/graphblob.py).
 * The tick list becomes a flat op program; SignalChunk buffers become slots. */
int dusp_program_create(const char *graph_blob, size_t len, dusp_program **out);
void dusp_program_destroy(dusp_program *p);

int dusp_program_num_channels(const dusp_program *p); /* channels of the rendered outlet */
int dusp_program_num_units(const dusp_program *p);    /* units the reference would tick per chunk */
int dusp_program_num_instructions(const dusp_program *p);
/* Replaces: outlet channel arrays growing while the circuit runs (`out[c] = out[c] || new Float32Array`, e.g.
 * Repeater.js:25, Multiply.js:26).  A unit the reference's order ticks before its producer (Unit.js:171-209 is not
 * a topological sort) sees the producer's outlet one tick late, so a multi-channel signal needs a few ticks to
 * reach every unit.  A blob holding a `WARMUP k` record compiles the circuit as its k-th tick runs; this returns
 * S, the number of such ticks before the program is the steady one (0 for most circuits).  A blob that needs
 * warm-up programs and has no WARMUP record fails with DUSP_EUNSUPPORTED; the host renders chunk k < S with
 * program k and hands state over with dusp_session_adopt (dusp_b200/render.py: SessionRun). */
int dusp_program_warmup_ticks(const dusp_program *p);
int dusp_program_num_slots(const dusp_program *p);
/* bytes of compulsory HBM traffic per voice-sample in DUSP_OUT_VOICES mode
 * (SURVEY.md 8d accounting: 4*C_out + ring traffic) and the ring part alone */
double dusp_program_bytes_per_voice_sample(const dusp_program *p, int out_mode);

/* Per-voice parameters: every inlet constant ("u<unit>.<inlet>[<ch>]") and every
 * numeric unit property / initial state ("u<unit>.<prop>") can be overridden per
 * voice, so one program renders many differently-tuned voices in one launch. */
int dusp_program_num_params(const dusp_program *p);
const char *dusp_program_param_name(const dusp_program *p, int index);
int dusp_program_param_index(const dusp_program *p, const char *name); /* -1 if absent */
double dusp_program_param_default(const dusp_program *p, int index);
/* human-readable listing of the op program (debugging); returns bytes needed */
size_t dusp_program_disassemble(const dusp_program *p, char *buf, size_t cap);

/* ---- sessions ------------------------------------------------------------
 * Replaces: the circuit objects' run-time state (SignalChunks, unit fields,
 * CircleBuffer / delay rings) -- here device-resident arrays for n voices.
 * `programs[i]` is instantiated `voices[i]` times; voices are numbered in that
 * order.  `device` is the CUDA ordinal.  */
int dusp_session_create(const dusp_program *const *programs, const int *voices, int n_programs,
                        int device, dusp_session **out);
void dusp_session_destroy(dusp_session *s);
int dusp_session_num_voices(const dusp_session *s);
int dusp_session_num_channels(const dusp_session *s); /* max over programs */
long long dusp_session_unit_count(const dusp_session *s); /* sum over voices of units per voice */

/* values[v * n_ids + k] overrides parameter param_ids[k] of voice v of program
 * `program_index` (v counts within that program).  Call before the first render
 * or after dusp_session_reset: a session that has rendered (clock != 0) is refused
 * with DUSP_EINVAL, because the other program groups would keep their advanced state. */
int dusp_session_set_params(dusp_session *s, int program_index, const int *param_ids, int n_ids,
                            const double *values);
int dusp_session_reset(dusp_session *s); /* back to clock 0 / initial unit state */

/* Replaces: an Event's function running before a tick (Event.js:18-30, Circuit.js:23,57-65), e.g. the
 * `trigger()` that scheduleTrigger / the `at` operator queue (UnitOrPatch.js:25-33,
 * constructOperation.js:63-69): between two renders (the session clock is chunk-aligned there, like the
 * reference's event granularity) assign unit-state words of voice `voice` of program `program_index`
 * (-1 = every voice of the group) on the device.  Unlike dusp_session_set_params this keeps the clock and
 * all other state.  Words that compile-time specialisations derive from (a constant filter cutoff, a
 * constant pan ...) cannot be poked: DUSP_EUNSUPPORTED.  Enqueued on `cuda_stream`. */
int dusp_session_poke(dusp_session *s, int program_index, int voice, int n_ids, const int *param_ids,
                      const double *values, void *cuda_stream);

/* Replaces: a graph edit made by an event while the circuit runs -- `then` handing an inlet a new source from
 * onFinish (constructOperation.js:45-60, UnitOrPatch.js:77-84), an event assigning an inlet (Unit.js:59-67):
 * Inlet.connect re-aliases a chunk and adds the new units to the live circuit (Inlet.js:40-74,
 * Circuit.js:67-107) while every existing unit object keeps its state.  Here the host compiles the circuit as it
 * stands after the edit into a new program (a new session `dst`) and group `dst_program` takes over the device
 * state of group `src_program` of `src` between two renders: units are matched by kind and label; their state
 * words, per-voice constants (unless the edit changed that constant), delay lines, CircleBuffers and the
 * previous-chunk buffers of feedback edges are copied device to device, and `dst` continues at `src`'s clock.
 * Same device and voice counts required; DUSP_EUNSUPPORTED if the edit turns an existing value into a feedback
 * edge (its previous chunk was never kept) or changes a delay-line length.  Enqueued on `cuda_stream`. */
int dusp_session_adopt(dusp_session *dst, int dst_program, dusp_session *src, int src_program, void *cuda_stream);

/* Replaces: the `for t0 ... await circuit.tickUntil(t1)` loop and the copy-out
 * (renderChannelData.js:29-45), for all voices at once.  Renders the next
 * `n_samples` samples (the session clock advances by whole chunks, like
 * Circuit.tick).  Exactly one of `out_host` / `out_device` is non-NULL:
 *   out_host   -- caller-owned host buffer; the call returns when it is filled
 *   out_device -- caller-owned device buffer on the session's GPU; the kernels
 *                 are enqueued on `cuda_stream` (a cudaStream_t, may be NULL)
 *                 and the call returns without synchronising
 * Buffer size: DUSP_OUT_VOICES n_voices*n_channels*n_samples floats,
 *              DUSP_OUT_MIX 2*n_samples floats.  NaN and -0 samples are stored
 *              as +0 (`x || 0`, renderChannelData.js:44).
 * Alignment:   any 4-byte aligned `out_device` is accepted.  The kernels store 16 bytes at a time when the
 *              buffer is 16-byte aligned and n_samples is a multiple of 4 (every row then is); otherwise they
 *              fall back to 4-byte stores (slower, same result). */
int dusp_session_render(dusp_session *s, long long n_samples, int out_mode, float *out_host,
                        float *out_device, void *cuda_stream);

long long dusp_session_clock(const dusp_session *s); /* samples ticked so far (whole chunks, like Circuit.clock) */

/* ---- state block: checkpoint / resume ----------------------------------------
 * Replaces: nothing the reference offers as a feature -- its state is the live objects themselves: `circuit.clock`
 * (Circuit.js:8,39) and every unit's own fields (Osc.phase, Filter.x1..y2, the CircleBuffers' channelData, ...), which
 * is why a render can be continued only on the same objects ("checkpoint / resume").  Here that
 * state is device arrays, so it can be taken out and put back: the clock, every voice's words (parameters + unit
 * state), rings and stale chunks of every program group.  A block can be imported into any session created from the
 * same programs with the same voice counts (the header carries the shapes and the call checks them); the render then
 * continues bit for bit.  Runs (dusp_run_*) keep their timeline position themselves and are not covered.
 *   dusp_session_state_size    -- bytes a block of this session takes
 *   dusp_session_state_export  -- block -> caller's HOST buffer of at least that size (synchronises `cuda_stream`)
 *   dusp_session_state_import  -- HOST buffer -> session (enqueued on `cuda_stream`, the buffer may be reused on return) */
int dusp_session_state_size(const dusp_session *s, size_t *bytes);
int dusp_session_state_export(dusp_session *s, void *host_block, size_t bytes, void *cuda_stream);
int dusp_session_state_import(dusp_session *s, const void *host_block, size_t bytes, void *cuda_stream);

/* ---- runs: a render driven through the circuits' timelines -----------------
 * Replaces: the whole `for t0 ... await circuit.tickUntil(t1)` loop of renderChannelData.js:29-45 INCLUDING what
 * Circuit.tick does before each chunk -- runEvents (Circuit.js:23,57-65; Event.js:18-30): triggers / state
 * assignments (UnitOrPatch.js:25-33, constructOperation.js:63-69) and graph edits (`for` / `then`,
 * constructOperation.js:36-60, Inlet.js:40-93).  The host shim runs the circuit's event queue once on the
 * reference's own objects and records what happened in the blob (format: dusp_b200/graphblob.py):
 *   EV <chunk> <unit> <prop> <value>    unit state assigned by an event, applied before that chunk's tick
 *   EPOCH <chunk> + DUSPGRAPH section   the circuit after an event edited the graph, from that chunk on
 *   KEEP / WARMUP <k>                   compile hints (keep every outlet's last chunk / channel counts still growing)
 * A run owns one program per section and the session that renders the current one.  dusp_run_render renders the
 * next n_samples in chunk-aligned segments: before the segment that starts at chunk k it switches every group whose
 * next EPOCH has begun to a new session that adopts the old one's device state (dusp_session_adopt), then pokes
 * chunk k's EV records into every voice of their group (dusp_session_poke).  Circuits whose channel counts grow over
 * their first ticks (dusp_program_warmup_ticks) get their warm-up programs as one-chunk epochs automatically.
 * `blobs[i]` is instantiated `voices[i]` times, as in dusp_session_create. */
typedef struct dusp_run dusp_run;
int dusp_run_create(const char *const *blobs, const size_t *lens, const int *voices, int n_programs, int device,
                    dusp_run **out);
void dusp_run_destroy(dusp_run *r);
int dusp_run_num_voices(const dusp_run *r);
int dusp_run_num_channels(const dusp_run *r);      /* widest program of ANY epoch: the channel rows per voice of every render */
int dusp_run_channels_rendered(const dusp_run *r); /* widest program of the epochs entered so far: what the reference's
                                                      result holds (renderChannelData.js:35-45 grows channels as they appear) */
long long dusp_run_position(const dusp_run *r);    /* samples rendered so far */
/* the session / program rendering group `program_index` right now (borrowed; replaced at the next epoch).  The
 * session is created on first use (first set_params / render / this call): NULL + dusp_last_error() without a GPU. */
dusp_session *dusp_run_session(dusp_run *r);
const dusp_program *dusp_run_program(const dusp_run *r, int program_index);
/* the timeline of group `program_index`: its programs (warm-up ticks, then one per EPOCH) and their first chunks */
int dusp_run_num_epochs(const dusp_run *r, int program_index);
const dusp_program *dusp_run_epoch_program(const dusp_run *r, int program_index, int epoch, long long *first_chunk);
/* per-voice parameters by name (dusp_session_set_params), before the first render */
int dusp_run_set_params(dusp_run *r, int program_index, const char *const *names, int n_names, const double *values);
/* as dusp_session_render; rows are [voice][dusp_run_num_channels][n_samples] (channels a segment's programs do not
 * produce read as silence) or [2][n_samples] (DUSP_OUT_MIX).  A render that ended inside a chunk cannot be continued. */
int dusp_run_render(dusp_run *r, long long n_samples, int out_mode, float *out_host, float *out_device, void *cuda_stream);

/* number of kernel launches issued by this session so far (bench accounting) */
long long dusp_session_launch_count(const dusp_session *s);
/* last launch geometry, for reporting: grid, block, dynamic smem bytes */
int dusp_session_launch_info(const dusp_session *s, int *grid, int *block, int *smem_bytes);

/* tile shape of the session's first program group (B samples per lane, T time-lanes per voice) and
 * whether it renders with a kernel specialised for the circuit shape (1) or the interpreter (0) */
int dusp_session_tile_info(const dusp_session *s, int *sub_block, int *time_lanes, int *specialised);
const char *dusp_session_warning(const dusp_session *s); /* "" or why specialisation was skipped */
/* Generates the specialised sm_100a kernel for tile shape (B, T) and compiles it with NVRTC (no GPU
 * needed).  Returns the cubin size in bytes (0 on failure, see dusp_last_error); when src_buf is
 * non-NULL the generated CUDA source is copied into it. */
size_t dusp_program_specialise(const dusp_program *p, int sub_block, int time_lanes, char *src_buf, size_t cap);

/* ---- multi-GPU mixdown: one-sided sum over NVLink peer memory ----------------
 * The reference sums all voices into one stereo signal (Sum.many, src/components/Sum.js:18-29, then
 * renderChannelData.js:42-44).  With voices sharded over the GPUs of a box, every process adds its DUSP_OUT_MIX
 * block into an accumulator that lives on the root GPU (csrc/mixpush.cu):
 *   root:  dusp_mix_acc_create -> device buffer of n_floats f32 + n_counters u32 arrival counters, and a 64-byte
 *          CUDA IPC handle to hand to the other processes (any byte transport);
 *   peers: dusp_mix_acc_open maps it (NVLink peer access);  dusp_mix_acc_close(ptr, owner) unmaps / frees;
 *   all:   dusp_mix_push(local, remote, n, remote_counter, ticket, stream) adds local[0..n) into remote[0..n) with
 *          system-scope float reductions and then increments *remote_counter once; `ticket` is one zeroed u32 of
 *          scratch on the pushing GPU;
 *   root:  dusp_mix_wait(counter, target, timed_out, stream) holds `stream` until *counter >= target (all ranks
 *          arrived) or ~30 s have passed, in which case *timed_out (device memory, may be NULL) is incremented;
 *          dusp_mix_acc_clear zeroes floats and counters for the next render. */
int dusp_mix_acc_create(int device, size_t n_floats, size_t n_counters, float **acc, unsigned **counters,
                        unsigned char handle[64]);
int dusp_mix_acc_open(int device, const unsigned char handle[64], size_t n_floats, float **acc, unsigned **counters);
int dusp_mix_acc_close(float *acc, int owner);
int dusp_mix_acc_clear(float *acc, size_t n_floats, size_t n_counters, void *cuda_stream);
int dusp_mix_push(const float *local, float *remote, size_t n_floats, unsigned *remote_counter, unsigned *ticket,
                  void *cuda_stream);
int dusp_mix_wait(const unsigned *counter, unsigned target, unsigned *timed_out, void *cuda_stream);

/* dusp_mix_push for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same adds and the same arrival-counter increment as dusp_mix_push, but the one word of scratch that call takes from
 * the caller (`ticket`) lives in a static device array inside the library: 256 bytes that every GPU loading the
 * library holds until exit.  The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine): it
 * selects `device`, looks the array's device address up once per device (cudaGetSymbolAddress; do the first call per
 * device outside a capture), clears the ticket on `cuda_stream` and launches the kernel there.  `local`, `remote` and
 * `remote_counter` are the caller's, as in dusp_mix_push.  Limits: n_floats <= DUSP_MIX_PUSH_STATIC_MAX_FLOATS per call
 * (more: DUSP_EINVAL with a message; split the block or use dusp_mix_push), device ordinals 0..63.  Every call on a
 * device shares the one ticket: calls that overlap on the GPU (different streams, no ordering between them) RACE --
 * keep them on one stream or order them with events. */
#define DUSP_MIX_PUSH_STATIC_MAX_FLOATS (1 << 24)
int dusp_mix_push_static(int device, const float *local, float *remote, size_t n_floats, unsigned *remote_counter,
                         void *cuda_stream);

/* The injected PRNG (replaces Math.random inside a tick: Noise.js:10,23, SporadicRetrigger.js:13): draw k of stream
 * `seed`, uniform in [0,1) with 32 bits of resolution:
 *   x = low32(k) + low32(seed); x ^= x >> 16; x *= 0x7feb352d; x ^= x >> 15; x *= 0x846ca68b; x ^= x >> 16;
 *   rand = (x ^ high32(seed)) / 2^32
 * Counter-based, so the draws of a chunk are independent of each other (lanes along time) and a Noise sample
 * `2 rand - 1` rounded to f32 is one integer-to-float conversion: float(int32(u ^ 0x80000000)) * 2^-31. */
double dusp_rand(uint64_t seed, uint64_t k);

#ifdef __cplusplus
}
#endif
#endif /* DUSP_B200_H */
