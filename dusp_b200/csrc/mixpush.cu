// mixpush.cu -- the one collective of the path as one-sided peer-memory traffic.
//
// In DUSP_OUT_MIX mode every GPU accumulates its voices into a per-block stereo mix; the reference's result is the sum
// over ALL voices (a `Sum.many` chain, src/components/Sum.js:18-29, ahead of renderChannelData.js:42-44).  Instead of a
// rendezvous collective per block (ncclReduce: every rank's kernel spins on SMs until the slowest rank arrives, and
// holds CTA slots the next block's single-wave render kernels were sized for), each rank ADDS its block straight into
// the root GPU's accumulator through NVLink peer memory and bumps an arrival counter there:
//
//   root:   acc = cudaMalloc'd [n_blocks][2][block] f32 + [n_blocks] u32 arrival counters, exported with cudaIpc
//   rank r: dusp_mix_push  -- grid-stride kernel on the render stream: red.add.f32 (system scope) of the local block
//           into acc[b] over NVLink, then __threadfence_system and one counter increment by the last CTA
//   root:   dusp_mix_wait  -- a one-thread kernel on the consumer stream polls counter[b] until all ranks arrived
//           (bounded: after ~30 s it records the miss in a flag the host checks)
//
// Nothing waits for a peer except the root's consumer, so a late rank delays only the block it is late for.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <string>

#include "../../include/dusp_b200.h"

namespace dusp { int fail_public(int code, const std::string &msg); }

namespace {

__global__ void mix_push_kernel(const float *__restrict__ local, float *remote, size_t n, unsigned *remote_counter,
                                unsigned *done_ctas) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float v = local[i];
    if (v != 0.0f) atomicAdd_system(remote + i, v);  // RED.ADD.F32 .SYS: travels as an NVLink atomic when remote is a peer
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned prev = atomicAdd(done_ctas, 1u);
    if (prev == gridDim.x - 1) {  // every CTA's adds are fenced: publish the arrival, re-arm the local ticket
      *done_ctas = 0;
      __threadfence_system();
      atomicAdd_system(remote_counter, 1u);
    }
  }
}

// Bounded: a rank that died must not leave the root's stream spinning for ever.  ~30 s of polling, then the miss is
// recorded in *timed_out (checked by the host after the render) and the stream moves on.
__global__ void mix_wait_kernel(const volatile unsigned *counter, unsigned target, unsigned *timed_out) {
  for (long long spins = 0; *counter < target; spins++) {
    if (spins > 30000000LL) { if (timed_out) atomicAdd(timed_out, 1u); break; }
    __nanosleep(1000);
  }
  __threadfence_system();
}

}  // namespace

extern "C" {

int dusp_mix_acc_create(int device, size_t n_floats, size_t n_counters, float **acc, unsigned **counters,
                        unsigned char handle[64]) {
  if (!acc || !counters || !handle) return dusp::fail_public(DUSP_EINVAL, "dusp_mix_acc_create: null argument");
  if (cudaSetDevice(device) != cudaSuccess) return dusp::fail_public(DUSP_ECUDA, "dusp_mix_acc_create: no such CUDA device");
  const size_t bytes = (n_floats + 3) / 4 * 16 + (n_counters + 64) * sizeof(unsigned);
  void *p = nullptr;
  cudaError_t e = cudaMalloc(&p, bytes);
  if (e == cudaSuccess) e = cudaMemset(p, 0, bytes);
  cudaIpcMemHandle_t h;
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) { if (p) cudaFree(p); return dusp::fail_public(DUSP_ECUDA, std::string("dusp_mix_acc_create: ") + cudaGetErrorString(e)); }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handles are 64 bytes");
  memcpy(handle, &h, 64);
  *acc = (float *)p;
  *counters = (unsigned *)((char *)p + (n_floats + 3) / 4 * 16);
  return DUSP_OK;
}

int dusp_mix_acc_open(int device, const unsigned char handle[64], size_t n_floats, float **acc, unsigned **counters) {
  if (!acc || !counters || !handle) return dusp::fail_public(DUSP_EINVAL, "dusp_mix_acc_open: null argument");
  if (cudaSetDevice(device) != cudaSuccess) return dusp::fail_public(DUSP_ECUDA, "dusp_mix_acc_open: no such CUDA device");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  void *p = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) return dusp::fail_public(DUSP_ECUDA, std::string("dusp_mix_acc_open: ") + cudaGetErrorString(e));
  *acc = (float *)p;
  *counters = (unsigned *)((char *)p + (n_floats + 3) / 4 * 16);
  return DUSP_OK;
}

int dusp_mix_acc_close(float *acc, int owner) {
  if (!acc) return DUSP_OK;
  cudaError_t e = owner ? cudaFree(acc) : cudaIpcCloseMemHandle(acc);
  return e == cudaSuccess ? DUSP_OK : dusp::fail_public(DUSP_ECUDA, std::string("dusp_mix_acc_close: ") + cudaGetErrorString(e));
}

int dusp_mix_acc_clear(float *acc, size_t n_floats, size_t n_counters, void *cuda_stream) {
  const size_t bytes = (n_floats + 3) / 4 * 16 + (n_counters + 1) * sizeof(unsigned);  // + the timed-out word behind the counters
  cudaError_t e = cudaMemsetAsync(acc, 0, bytes, (cudaStream_t)cuda_stream);
  return e == cudaSuccess ? DUSP_OK : dusp::fail_public(DUSP_ECUDA, std::string("dusp_mix_acc_clear: ") + cudaGetErrorString(e));
}

int dusp_mix_push(const float *local, float *remote, size_t n_floats, unsigned *remote_counter, unsigned *ticket,
                  void *cuda_stream) {
  if (!local || !remote || !remote_counter || !ticket) return dusp::fail_public(DUSP_EINVAL, "dusp_mix_push: null argument");
  const int block = 256;
  int grid = (int)((n_floats + block * 4 - 1) / (block * 4));
  if (grid < 1) grid = 1;
  if (grid > 148) grid = 148;
  mix_push_kernel<<<grid, block, 0, (cudaStream_t)cuda_stream>>>(local, remote, n_floats, remote_counter, ticket);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? DUSP_OK : dusp::fail_public(DUSP_ECUDA, std::string("dusp_mix_push: ") + cudaGetErrorString(e));
}

// ---- the push without any allocation on the caller's side (no counterpart in the reference; operators' request) ----
// dusp_mix_push needs one zeroed u32 of device scratch (the CTA-arrival ticket) that the caller has to allocate.
// Callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc) use this variant:
// the ticket lives in a file-scope device array that every GPU loading the library holds until exit (256 bytes).
// In HOST code `g_mix_static_scratch` is only a stand-in for the symbol, not a device address: its address is looked up
// once per device, with that device current, by cudaGetSymbolAddress and that pointer is what the kernel gets.
#define DUSP_MIX_STATIC_WORDS 64
static __device__ __align__(256) unsigned g_mix_static_scratch[DUSP_MIX_STATIC_WORDS];  // word 0: arrival ticket; rest reserved
static unsigned *g_mix_static_ptr[64];  // per device ordinal: device address of the array (0 = not looked up yet)

int dusp_mix_push_static(int device, const float *local, float *remote, size_t n_floats, unsigned *remote_counter,
                         void *cuda_stream) {
  if (!local || !remote || !remote_counter) return dusp::fail_public(DUSP_EINVAL, "dusp_mix_push_static: null argument");
  if (device < 0 || device >= 64) return dusp::fail_public(DUSP_EINVAL, "dusp_mix_push_static: device ordinal out of range (0..63)");
  if (n_floats > (size_t)DUSP_MIX_PUSH_STATIC_MAX_FLOATS) {
    char m[200];
    snprintf(m, sizeof m, "dusp_mix_push_static: %zu floats exceed the maximum of %d per call (DUSP_MIX_PUSH_STATIC_MAX_FLOATS); "
             "split the block or use dusp_mix_push with caller-owned scratch", n_floats, DUSP_MIX_PUSH_STATIC_MAX_FLOATS);
    return dusp::fail_public(DUSP_EINVAL, m);
  }
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return dusp::fail_public(DUSP_ECUDA, std::string("dusp_mix_push_static: cudaSetDevice: ") + cudaGetErrorString(e));
  if (!g_mix_static_ptr[device]) {  // set-up, once per device, with that device current
    unsigned *p = nullptr;
    e = cudaGetSymbolAddress((void **)&p, g_mix_static_scratch);
    if (e != cudaSuccess || !p)
      return dusp::fail_public(DUSP_ECUDA, std::string("dusp_mix_push_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
    g_mix_static_ptr[device] = p;
  }
  unsigned *ticket = g_mix_static_ptr[device];
  cudaStream_t st = (cudaStream_t)cuda_stream;
  // every call shares the array and it keeps the last call's contents: the ticket must start at zero
  e = cudaMemsetAsync(ticket, 0, sizeof(unsigned), st);
  if (e != cudaSuccess) return dusp::fail_public(DUSP_ECUDA, std::string("dusp_mix_push_static: clearing the ticket: ") + cudaGetErrorString(e));
  const int block = 256;
  int grid = (int)((n_floats + block * 4 - 1) / (block * 4));
  if (grid < 1) grid = 1;
  if (grid > 148) grid = 148;
  mix_push_kernel<<<grid, block, 0, st>>>(local, remote, n_floats, remote_counter, ticket);
  e = cudaGetLastError();
  return e == cudaSuccess ? DUSP_OK : dusp::fail_public(DUSP_ECUDA, std::string("dusp_mix_push_static: ") + cudaGetErrorString(e));
}

int dusp_mix_wait(const unsigned *counter, unsigned target, unsigned *timed_out, void *cuda_stream) {
  if (!counter) return dusp::fail_public(DUSP_EINVAL, "dusp_mix_wait: null argument");
  mix_wait_kernel<<<1, 1, 0, (cudaStream_t)cuda_stream>>>(counter, target, timed_out);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? DUSP_OK : dusp::fail_public(DUSP_ECUDA, std::string("dusp_mix_wait: ") + cudaGetErrorString(e));
}

}  // extern "C"
