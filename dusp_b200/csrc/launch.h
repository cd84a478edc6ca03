// launch.h -- host entry point of the render kernel (render_sm100.cu).
#pragma once
#include "vm.h"

#define DUSP_MAX_CTA_THREADS 256

#ifdef __cplusplus
extern "C" {
#endif
// Enqueues one render launch for a program group on `stream`; returns a cudaError_t value.
int dusp_launch_render(const DuspKernelParams *kp, int sub_block, int warps_per_cta, void *stream,
                       int *grid_out, int *smem_out);
#ifdef __cplusplus
}
#endif
