// cx_pipeline.h — internal interface of the wavefront pipeline (lunar_pipeline.cu) for the C-ABI
// translation unit (abi_lunar.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <functional>

#include "../../include/cotix_b200.h"

// called on the stream of range r (envs [off, off + n)) before its first / after its last kernel
typedef std::function<int(cudaStream_t stream, int r, int64_t off, int64_t n)> CxRangeHook;

extern "C" int64_t cx_lunar_pipeline_workspace_bytes(int64_t B);
extern "C" int cx_launch_lunar_pipeline(cudaStream_t, int64_t, const float*, float*, int64_t, const float*,
                                        float*, uint8_t*, float*, const CotixLunarParams*, void*, int64_t);
extern "C" int cx_launch_lunar_pipeline_steps(cudaStream_t st, int64_t B, const float* sin0, float* sout0,
                                              int64_t stride, int64_t ld, const float* actions,
                                              int64_t action_stride, int n_actions, int n_steps, float* reward,
                                              uint8_t* done, float* obs, const CotixLunarParams* p,
                                              void* workspace, int64_t workspace_bytes);
// one in-place step cut into `split` ranges (0 = the configured split) with hooks around every
// range's chain: what the host-facing env handle uses to put H2D(action) / D2H(obs, reward, done) of
// a range on the same stream as its kernels
int cx_launch_lunar_pipeline_hooked(cudaStream_t st, int64_t B, float* state, int64_t ld, const float* action,
                                    float* reward, uint8_t* done, float* obs, const CotixLunarParams* p,
                                    void* workspace, int64_t workspace_bytes, int split, const CxRangeHook& pre,
                                    const CxRangeHook& post);
extern "C" int cx_lunar_pipeline_config(int split, int pdl, int priorities, int64_t item_capacity);
extern "C" int cx_lunar_pipeline_counters(cudaStream_t st, int64_t B, const void* workspace, int64_t* out);
extern "C" int cx_lunar_pipeline_profile(int enable);
extern "C" int cx_lunar_pipeline_profiling();
extern "C" int64_t cx_lunar_pipeline_stage_ms(double* out);
