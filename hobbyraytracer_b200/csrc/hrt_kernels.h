// hrt_kernels.h — launcher interface between the C ABI (hrt_api.cu) and the kernels (hrt_kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "hrt_types.h"

namespace hrt {

struct Hit;

// Device-resident working set of one wavefront batch of `capacity` paths (path slot = s_local * npix + pixel).
//   rays[2]   : ping-pong ray queues, 2 x float4 per ray: (o.xyz, slot) (d.xyz, -)            32 B / ray
//   hits      : one 16-byte Hit per queue entry, written by K2, read by K4
//   atten     : per path slot, rayColour's `attenuation` (main.cpp:40)                         16 B / path
//   radiance  : per path slot, rayColour's `result`                                            16 B / path
//   counts[b] : queue length at bounce b (counts[b+1] is appended to by K4 of bounce b)
//   retrace   : queue entries whose answer depends on the visiting order (hrt_hit.cuh), n_retrace[b] of them
struct RenderJob {
    int width, height, npix;
    int max_depth;
    float t_min;
    uint64_t seed;
    int rng_mode;
    float4* rays[2];
    Hit* hits;
    float4* atten;
    float4* radiance;
    unsigned int* counts;
    unsigned int* retrace;
    unsigned int* n_retrace;
    unsigned long long* totals;     // [0] rays traced, [1] rays re-traced in reference order
};

void launch_trace_batch(const DScene& s, const float4* d_rays, long long n, Hit* d_hits, uint64_t seed,
                        unsigned int* d_retrace, unsigned int* d_n_retrace, int sm_count, cudaStream_t st, uint64_t* launches);
void launch_split_hits(const Hit* d_hits, long long n, uint32_t* d_prim, float* d_t, int sm_count, cudaStream_t st, uint64_t* launches);
void launch_eval_batch(const DScene& s, const float4* d_rays, const Hit* d_hits, long long n, float* d_normal, float* d_uv, uint8_t* d_front,
                       int sm_count, cudaStream_t st, uint64_t* launches);
// one batch of s_count samples per pixel: raygen, max_depth x (trace, retrace, shade), film add into d_accum.
// ev_trace (2 per bounce) / ev_shade (1 per bounce) are optional CUDA events recorded around the kernels.
void launch_render_batch(const DScene& s, const RenderJob& j, const DCamera& cam, int s_first, int s_count, float* d_accum,
                         bool mega, int sm_count, cudaStream_t st, uint64_t* launches, cudaEvent_t* ev_trace, cudaEvent_t* ev_shade);
void launch_film_resolve(const float* d_accum, int npix, int spp, float* d_mean, uint8_t* d_rgb8, int sm_count, cudaStream_t st, uint64_t* launches);

} // namespace hrt
