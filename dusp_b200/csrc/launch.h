// launch.h -- host entry point of the render kernel (render_sm100.cu).
#pragma once
#include "vm.h"

#define DUSP_MAX_CTA_THREADS 128

#ifdef __cplusplus
extern "C" {
#endif
// Enqueues one render launch for a program group on `stream`; returns a cudaError_t value.
// `kp` must already hold the program (inline or global).  sub_block B in {8,16}, time_lanes T in
// {1,4,16,32}: a warp renders 32/T voices, T lanes per voice each walking B consecutive samples.
int dusp_launch_render(const DuspKernelParams *kp, int sub_block, int time_lanes, int warps_per_cta, void *stream,
                       int *grid_out, int *smem_out);
// 1 if the (B, T) pair is compiled in
int dusp_launch_supported(int sub_block, int time_lanes);
#ifdef __cplusplus
}
#endif
