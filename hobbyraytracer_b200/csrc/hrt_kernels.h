// hrt_kernels.h — launcher interface between the C ABI (hrt_api.cu) and the kernels (hrt_kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "hrt_types.h"

namespace hrt {

struct Hit;

// Film layout and arithmetic: hrt_shade.cuh (FILM_LANES, film_quantize, film_channel).

// Device-resident control block of one persistent-wavefront render call.
struct WaveCtl {
    unsigned int n_front[2];            // queue q: survivors, compacted from the front
    unsigned long long base[2];         // queue q: value of `next_path` when the shade pass that fills q started
    unsigned long long next_path;       // next camera path index to hand out (may overshoot total_paths)
    unsigned long long total_paths;
    unsigned long long rays;            // closest-hit queries issued
    unsigned long long resolved;        // rays answered by pass 2 + replay (hrt_trace.cuh)
    unsigned long long fallbacks;       // rays answered by trace_ref
    unsigned long long clamped;         // finite sample channels beyond the film's +-2^21 range (clamped, hrt_shade.cuh)
    unsigned int sec_next[2];           // fused-primary pipeline: hand-out counter of the secondary entries of queue q (per-lane refill)
    unsigned int n_prim[2];             // fused-primary pipeline: camera-path hits appended behind the n_front[q] secondary entries of queue q
    unsigned int pend_count;            // wide pipeline: entries of the current queue whose hit record is pending (WaveJob::pend), reset by the shade pass
    unsigned int pad_;
    float dbg[16][8];                   // first rays that took the trace_ref fallback (o, pixel, d, state), for diagnosis
};

// Working set of a render call.  Queue entry e of queue q (capacity entries each):
//   rays[q][2e]   = (o.xyz, pixel index)            rays[q][2e+1] = (d.xyz, has_result << 31 | sample_local << 8 | bounce)
//   thr[q][e]     = (rayColour's currentAttenuation, -)           res[q][e] = (rayColour's `result`, -) when has_result
// Survivors of a shade pass are appended from the front [0, n_front); regenerated camera paths take slots from the back
// (path p of that pass sits at capacity - 1 - (p - base)), so incoherent secondary rays and coherent primary rays stay in
// separate warps.  hits[e] is written by the trace pass for the entry order of that pass.
struct WaveJob {
    int width, height, npix;
    int tiles_x, tiles_y;               // 8x4 pixel tiles; a path index is (sample_local * tiles + tile) * 32 + lane
    int max_depth;
    float t_min;
    uint64_t seed;
    int rng_mode;
    int rr;                             // HRT_RENDER_RR: Russian roulette in the shade pass (hrt_shade.cuh)
    int s_first, s_count;
    unsigned int capacity;
    float4* rays[2];
    float4* thr[2];
    float4* res[2];
    Hit* hits;
    unsigned int* pend;                 // capacity entry indices: rays pass 1 could not decide (hrt_wide.cuh), settled by the resolve pass
    WaveCtl* ctl;
    long long* film;                    // npix * FILM_LANES
};

// d_pend_idx (n entries) / d_pend_count (one word): scratch for the rays that need pass 2 (hrt_wide.cuh)
void launch_trace_batch(const DScene& s, const float4* d_rays, long long n, Hit* d_hits, uint64_t seed, unsigned long long* d_counter,
                        unsigned int* d_pend_idx, unsigned int* d_pend_count, int sm_count, cudaStream_t st, uint64_t* launches);
// The same batch with its scratch in a file-scope device array of hrt_kernels.cu (no allocation; at most HRT_STATIC_BATCH_MAX rays;
// `device` must be current).  Returns the CUDA status of the set-up (cudaGetSymbolAddress, once per device) and of the launches.
#define HRT_STATIC_BATCH_MAX (1ll << 20)
#define HRT_STATIC_MAX_DEVICES 64
cudaError_t launch_trace_batch_static(const DScene& s, int device, const float4* d_rays, long long n, uint32_t* d_prim, float* d_t, uint64_t seed,
                                      int sm_count, cudaStream_t st, uint64_t* launches);
void launch_camera_rays(const DCamera& cam, const float* d_st, long long n, float* d_out, int sm_count, cudaStream_t st);
void launch_texture_eval(const DScene& s, int tex, const float* d_in, long long n, float* d_out, int sm_count, cudaStream_t st, uint64_t* launches);
void launch_scatter_batch(const DScene& s, int material, const float* d_in, const uint8_t* d_front, long long n, uint64_t seed, int rng_mode, float* d_out,
                          int sm_count, cudaStream_t st, uint64_t* launches);
void launch_split_hits(const Hit* d_hits, long long n, uint32_t* d_prim, float* d_t, int sm_count, cudaStream_t st, uint64_t* launches);
void launch_eval_batch(const DScene& s, const float4* d_rays, const Hit* d_hits, long long n, float* d_normal, float* d_uv, uint8_t* d_front,
                       int sm_count, cudaStream_t st, uint64_t* launches);

// persistent wavefront: init fills queue 0, then iteration `it` is trace(queue it&1) + shade(queue it&1 -> queue (it+1)&1)
void launch_wave_init(const WaveJob& j, const DCamera& cam, int sm_count, cudaStream_t st, uint64_t* launches);
void launch_wave_trace(const DScene& s, const WaveJob& j, int it, int sm_count, cudaStream_t st, uint64_t* launches);
void launch_wave_shade(const DScene& s, const WaveJob& j, const DCamera& cam, int it, int sm_count, cudaStream_t st, uint64_t* launches);
// fused-primary wavefront (default): iteration `it` is trace2(secondary rays of queue it&1 + camera paths until the queue is full)
// + shade2(queue it&1 -> survivors to queue (it+1)&1); no init pass, the control block starts zeroed with total_paths set
unsigned long long wave2_total_claims(const WaveJob& j);   // value of WaveCtl::total_paths for this pipeline (it hands out claims, not paths)
void launch_wave2_trace(const DScene& s, const WaveJob& j, const DCamera& cam, int it, bool secondary_heavy, int sm_count, cudaStream_t st, uint64_t* launches);
// the same pipeline on scenes with the 8-wide trees (DScene::wide): secondary entries by the phase kernel with per-lane refill where
// that pays (hrt_wide.cuh), camera paths by the fused-primary kernel, a small third kernel settles the pending entries
void launch_wave3_trace(const DScene& s, const WaveJob& j, const DCamera& cam, int it, bool secondary_heavy, int sm_count, cudaStream_t st, uint64_t* launches);
void launch_wave2_shade(const DScene& s, const WaveJob& j, int it, int sm_count, cudaStream_t st, uint64_t* launches);
// one thread per path, whole rayColour loop (validation of the wavefront pipeline); same film
void launch_render_mega(const DScene& s, const WaveJob& j, const DCamera& cam, int sm_count, cudaStream_t st, uint64_t* launches);

// film -> per-pixel float sums (may be null) / mean (may be null) / 8-bit tonemapped image (may be null)
// the same over the local film + up to 15 peer films mapped through CUDA IPC (NVLink loads inside the kernel)
void launch_film_resolve_peers(long long* d_film, const long long* const* peers, int n_peers, int npix, int spp, int write_back,
                                float* d_sum, float* d_mean, uint8_t* d_rgb8, int sm_count, cudaStream_t st, uint64_t* launches);
// the reduction spread over the ranks: pixels [pix_first, pix_first + pix_count) of up to 16 films (own + mapped peers) -> outputs (may be peer memory)
void launch_film_resolve_slice(const long long* const* films, int n_films, int pix_first, int pix_count, int spp,
                               float* d_sum, float* d_mean, uint8_t* d_rgb8, int sm_count, cudaStream_t st, uint64_t* launches);
void launch_film_resolve(const long long* d_film, int npix, int spp, float* d_sum, float* d_mean, uint8_t* d_rgb8, int sm_count, cudaStream_t st, uint64_t* launches);

} // namespace hrt
