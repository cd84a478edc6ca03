// vm.h -- the flat op program that a circuit's tick list is compiled into, shared
// by the host compiler (compile.cc) and the device interpreter (render_sm100.cu).
//
// One instruction = one reference unit `_tick` for ONE channel (multi-channel
// units are expanded per channel at compile time; the reference's `c % length`
// channel broadcasting is resolved there).  Operands name "slots": per-voice
// buffers that replace SignalChunk channels (src/SignalChunk.js:1-11).
#pragma once
#include <stdint.h>

#define DUSP_SR 44100
#define DUSP_CHUNK 256
#define DUSP_TABLE_LEN 44101  // waveTables.js:5 / shapeTables.js:4 (sampleRate + 1)

enum DuspOp : int32_t {
  OP_NOP = 0,
  // --- stateless maps (Multiply.js, Sum.js, ... SURVEY 8a row a8)
  OP_COPY,      // Repeater.js:23-30 / ConcatChannels.js:16-30      out = in
  OP_NEG,       // PolarityInvert.js:13-20
  OP_ABS,       // Abs.js:17-24
  OP_S2S,       // SecondsToSamples.js:13-20                         in * 44100
  OP_SEMI,      // SemitoneToRatio.js:14-22                          2^(in/12)
  OP_DB2S,      // DecibelToScaler.js:13-19                          10^(in/20)
  OP_MIDI2F,    // MidiToFrequency.js:14-21                          2^((in-69)/12)*440
  OP_FIXMUL,    // FixedMultiply.js:19-22                            in * sf          [p: sf]
  OP_MUL,       // Multiply.js:23-34
  OP_DIV,       // Divide.js:16-27
  OP_POW,       // Pow.js:19-30
  OP_ADD,       // Sum.js:33-44
  OP_SUB,       // Subtract.js:19-29
  OP_GAIN,      // Gain.js:17-24                                     10^(gain/20) * in   in0=in in1=gain
  OP_CLIP,      // Clip.js:12-21
  OP_HCA,       // HardClipAbove.js:14-24
  OP_HCB,       // HardClipBelow.js:14-24
  OP_RESCALE,   // Rescale.js:24-37                                  in0..in4
  OP_XFADE,     // CrossFader.js:20-29                               in0=a in1=b in2=dial
  OP_PAN,       // Pan.js:18-24                                      in0=in in1=pan out0,out1 [p: compensationDB]
  OP_VECMAG,    // VectorMagnitude.js:16-28                          in0..in(n-1), aux0 = n (<=5)
  // --- stateful (sequential in time per voice)
  OP_OSC,       // Osc.js:35-47            in0=f [p: phase] aux0=table
  OP_MCOSC,     // MultiChannelOsc.js:21-38 (no negative-phase fix)
  OP_RAMP,      // Ramp.js:25-40           [p: t, playing, duration, y0, y1]
  OP_SHAPE,     // Shape/index.js:28-59    in0=duration in1=min in2=max [p: t, playing, leftEdge, rightEdge] aux0=table aux1=edge flags
  OP_AHD,       // AHD.js:35-76            in0..2 [p: t, state, playing]
  OP_TIMER,     // Timer.js:37-42          [p: t]
  OP_NOISE,     // Noise.js:16-28          in0=f [p: phase, y, draw, seed_lo, seed_hi]
  OP_FILTER,    // Filter.js:27-51         in0=in in1=f [p: a0,a1,a2,b1,b2,lastF,x1,x2,y1,y2] aux0=kind
  // --- rings (delay lines; CircleBuffer.js, Delay.js ...)
  OP_FIXDELAY,  // FixedDelay.js:14-20     in0=in [p: tBuffer] aux0=ring
  OP_COMB,      // CombFilter.js:11-17     in0=in in1=feedbackGain
  OP_ALLPASS,   // AllPass.js:8-15
  OP_DELAY,     // Delay.js:20-41          in0=in in1=delay aux0=ring
  OP_MONODELAY, // MonoDelay.js:19-30
  OP_READBACK,  // ReadBackDelay.js:23-43  in0=in in1=delay [p: tBuffer] aux0=ring
  OP_CBWRITE,   // CircleBufferWriter.js:12-25  in0=offset in1=in(or none) [p: t0] aux0=ring aux1=preWipe
  OP_CBREAD,    // CircleBufferReader.js:12-25  in0=offset [p: t0] aux0=ring aux1=postWipe
  // --- plumbing
  OP_LDSTALE,   // slot <- previous chunk of a value produced later in the tick order (cyclic graphs)
  OP_STSTALE,   // stale buffer <- slot
  OP_OUT,       // renderChannelData.js:42-44: out[channel aux0][t] = slot || 0
  // --- specialisations chosen by the compiler when an inlet is a per-voice constant: the
  //     transcendental / index arithmetic that depends only on that constant is evaluated once
  //     per voice on the host ("derived words") with the same libm formulas the reference uses
  OP_FILTERC,   // Filter, constant F        [p: a0,a1,a2,b1,b2 (derived), x1,x2,y1,y2]
  OP_PANC,      // Pan, constant PAN         [p: 1-pan, 1+pan, compensation (derived)]
  OP_GAINC,     // Gain, constant GAIN       [p: 10^(gain/20) (derived)]
  OP_NOISE1,    // Noise, constant F >= sampleRate: a new draw every sample   [p as OP_NOISE]
  OP_CBWRITEC,  // CircleBufferWriter, constant OFFSET  [p: t0, running ring index (derived start)]
  OP_CBREADC,   // CircleBufferReader, constant OFFSET
  OP_ADD0,      // (0 + in0) + in1: what a preWipe writer followed by a mixing writer leave on one ring position
                //  (CircleBufferWriter.js:17-22 twice); feeds the single merged writer (compile.cc peephole)
  // --- control-rate retriggering: decided once per 256-sample chunk, before the target ticks
  OP_RETRIG,    // Retriggerer.js:13-22        in0=rate [p: t, fired]: fired = 1 if t crossed sampleRate in this chunk
  OP_SRETRIG,   // SporadicRetrigger.js:12-16  in0=rate [p: draw, seed_lo, seed_hi, fired]: fired = rand() < rate[0]*256/SR
  OP_POKEIF,    // the assignments of target.trigger(): if word aux0 != 0 then word p = word aux1
  OP_SRREDUX,   // SampleRateRedux.js:20-37    in0=in in1=ammount [p: timeSinceLastUpdate, val]  (per channel; each keeps
                //                              its own copy of the shared counter, which only depends on `ammount`)
  // --- frame-rate units (components/spectral).  Frames live in per-voice rings (persistent, like the
  //     SignalChunk a slow unit leaves behind between its ticks); whole-chunk tiles only; in[4] = tick interval in
  //     samples (Circuit.js:35: the op runs when clock % interval == 0)
  OP_HOPPER,    // Hopper.js:19-37        in0=in  aux0=buffer ring aux1=out ring aux2=hopSize [p: t]
  OP_WINDOW,    // Windower.js:17-23      aux0=in ring aux1=out ring aux2=ctab offset of the envelope
  OP_FFT,       // FFT.js:19-25           aux0=in ring aux1=out ring aux2=ctab offset of the twiddle table
  OP_IFFT,      // IFFT.js:21-33          aux0=in ring aux1=out ring aux2=ctab offset, in0=complexOut ring (f64)
  OP_UNHOP,     // UnHopper.js:19-44      aux0=in ring aux1=out ring aux2=accumulator ring (f64) [p: t]
  OP_RECHUNK,   // ReChunk.js:26-49       aux0=in ring aux1=out ring aux2=buffer ring, in0=inputInterval in1=outputInterval [p: t]
  OP_SPECPASS,  // HardLowPass.js:20-31 / HardHighPass.js:20-31   in0=f aux0=in ring aux1=out ring aux2=high
  OP_SPECGATE,  // SpectralGate.js:15-32  in0=threshold aux0=in ring aux1=out ring aux2=invert
  OP_BINSHIFT,  // BinShift.js:14-26      in0=shift aux0=in ring aux1=out ring
  OP_AUGMENT,   // Augment.js:14-27       aux0=in ring aux1=out ring aux2=entries [p: ratio0, gain0, ratio1, ...]
  OP_SPECSUM,   // SpectralSum.js:15-24   aux0=a ring aux1=out ring aux2=b ring
  OP_LDFRAME,   // slot <- first 256 samples of a frame ring (a standard unit reading a 256-sample frame outlet)
  OP_STFRAME,   // frame ring <- slot (a frame inlet fed by a standard outlet)
  OP_OUTRING,   // renderChannelData.js:25-45 with outlet.chunkSize = aux2 > 256: out[aux0][t] = ring aux1 [t % aux2] || 0
  OP_PICK,      // PickChannel.js:16-21 with a signal-rate index: out = in[c % n], in0 = c, in1..in4 = the channels, aux0 = n (<= 4)
  OP_COUNT
};

// Operand reference: >= 0 slot index; REF_NONE = absent (reads as 0);
// <= REF_CONST_BASE: per-voice constant, word index = REF_CONST_BASE - ref.
#define REF_NONE (-1)
#define REF_CONST_BASE (-2)
static inline int32_t ref_const(int32_t word) { return REF_CONST_BASE - word; }

struct DuspInstr {
  int32_t op;
  int32_t in[5];
  int32_t out[2];
  int32_t p;       // first per-voice word (parameters + state) of this instruction
  int32_t aux[3];
  int32_t unit;    // source unit index (diagnostics)
  int32_t chan;
  int32_t pad[2];
};
static_assert(sizeof(DuspInstr) == 64, "DuspInstr layout");

struct DuspRing {
  int64_t offset;  // float offset inside a voice's ring arena
  int64_t length;  // samples
};

// table ids (waveTables.js:33-40, shapeTables.js:20-38)
enum { TAB_SIN = 0, TAB_SAW, TAB_SQUARE, TAB_TRIANGLE, TAB_8BIT, TAB_DECAY, TAB_ATTACK, TAB_SEMISINE, TAB_DECAYSQ, TAB_COUNT };

// Device encoding of an instruction (48 bytes, all fields 32-bit: no byte permutes to decode).
// Programs of up to DUSP_INLINE_INSTR instructions travel inside the kernel parameter block
// (constant bank: fetched with uniform LDCs, no global load).
struct __attribute__((aligned(16))) DuspDevInstr {
  int32_t op;
  int32_t in[5];
  int32_t out[2];
  int32_t p;
  int32_t aux0, aux1, aux2;
};
static_assert(sizeof(DuspDevInstr) == 48, "DuspDevInstr layout");
#define DUSP_INLINE_INSTR 80

// Heterogeneous batches (many different circuits, one voice-warp each): per-warp descriptor of a
// multi-program launch of the interpreter (tile shape B = 8, T = 32: a warp renders one voice).
struct DuspWarpDesc {
  const DuspDevInstr *code;
  double *words;           // this voice's word 0; word w at words[w * nv_stride]
  float *rings;            // this voice's ring arena
  const DuspRing *ring_tab;
  float *stale;            // this voice's stale buffers
  int64_t nv_stride;
  int64_t out_row;         // global voice index (output row)
  int32_t n_instr, n_stale, prog_nch, pad;
};

// One launch renders `n_samples` for every voice of one program group (or, with `multi`, one voice per
// warp of many program groups).
struct DuspKernelParams {
  int32_t n_instr;
  int32_t n_slots;
  int32_t n_voices;      // voices in this group
  int32_t voice0;        // global index of the group's first voice (output row)
  int64_t nv_stride;     // padded voice count = stride between words
  double *words;         // [n_words][nv_stride] parameters + unit state (f64)
  float *rings;          // [voice][ring_stride]
  int64_t ring_stride;
  const DuspRing *ring_tab;
  float *stale;          // [voice][n_stale][256]
  int32_t n_stale;
  int32_t use_inline;    // program is in prog_inline (else prog_global)
  const float *tables;   // [TAB_COUNT][DUSP_TABLE_LEN]
  float *out;            // DUSP_OUT_VOICES: [voice][out_nch][out_len]; DUSP_OUT_MIX: [2][out_len]
  int32_t out_nch;       // channel stride of the output (session-wide)
  int32_t prog_nch;      // channels this program produces
  int32_t out_mode;
  int32_t out_raw;       // 1: store samples as the outlet's chunk holds them (no `x || 0`): the streaming sink's view
  int64_t out_len;       // samples per channel row in `out`
  int64_t clock0;        // absolute sample index of the first sample of this launch (multiple of 256)
  int64_t n_samples;     // samples to store (<= processed); processing covers whole chunks
  const double *ctab;    // program constants built on the host with libm: FFT twiddles, window envelopes (f64)
  const DuspDevInstr *prog_global;
  const DuspWarpDesc *multi;  // non-null: n_voices descriptors, one per warp (requires T = 32)
  long long *dbg_cycles; // diagnostics (DUSP_MULTI_DEBUG_FILE): clock64() ticks every warp of a multi launch spent, else null
  int64_t density;       // multi only: > 0 selects the register-capped instantiation (64 registers per thread, 32 circuits per SM)
  DuspDevInstr prog_inline[DUSP_INLINE_INSTR];
};
