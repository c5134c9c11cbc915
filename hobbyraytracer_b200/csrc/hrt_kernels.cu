// hrt_kernels.cu — the sm_100a kernels of the wavefront path tracer and their launchers.
//
//   K1 k_raygen          camera rays, Philox jitter                    (main.cpp:115-123, camera.h:32-39)
//   K2 k_trace           closest hit per queued ray                    (world->hit, main.cpp:45)
//      k_trace_retrace   rays whose answer depends on visiting order, in the reference's order (hrt_hit.cuh)
//   K4 k_shade           hit record -> emitted + scatter / background; K3 = warp-ballot + one atomic per warp
//                        appends the surviving rays to the next queue  (main.cpp:47-76, material.h, texture.cpp)
//   K5 k_film_add / k_film_resolve   per-pixel sums in sample order, mean -> ACES -> sqrt -> 8 bit (film.cpp:25-52)
//   K7 k_trace_batch / k_eval_batch  hit() on a caller-supplied ray batch (parity / micro-benchmark entry)
//      k_render_mega     one thread per path, whole rayColour loop (validation of the wavefront pipeline)
//
// All kernels are persistent-style grid-stride loops over device-resident counts: the host never waits for a queue
// length, a render is a fixed sequence of launches.  Compiled with -fmad=false (numerics contract, hrt_math.cuh).
#include <cuda_runtime.h>

#include "hrt_kernels.h"
#include "hrt_shade.cuh"

namespace hrt {

namespace {

constexpr int BLOCK = 128;

__device__ __forceinline__ RngKey make_key(uint64_t seed, uint32_t index, uint32_t sample, int mode)
{
    RngKey k; k.k0 = (uint32_t)seed; k.k1 = (uint32_t)(seed >> 32); k.index = index; k.sample = sample; k.mode = mode;
    return k;
}

// ---- K7 ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(BLOCK) k_trace_batch(DScene s, const float4* __restrict__ rays, long long n, Hit* __restrict__ hits,
                                                       uint64_t seed, unsigned int* __restrict__ retrace, unsigned int* __restrict__ n_retrace)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        float4 a = __ldg(rays + 2 * i), b = __ldg(rays + 2 * i + 1);
        Ray r; r.o = v3(a.x, a.y, a.z); r.d = v3(b.x, b.y, b.z);
        RngKey key = make_key(seed, (uint32_t)i, (uint32_t)(i >> 32), 0);
        Hit h;
        if (s.fast) {
            bool need_ref = false;
            h = trace_fast_try<false>(s, r, a.w, b.w, key, 0, nullptr, need_ref);
            if (need_ref) { retrace[atomicAdd(n_retrace, 1u)] = (unsigned int)i; continue; }
        } else h = trace_ref(s, r, a.w, b.w, key, 0);
        hits[i] = h;
    }
}
__global__ void __launch_bounds__(BLOCK) k_trace_batch_retrace(DScene s, const float4* __restrict__ rays, Hit* __restrict__ hits, uint64_t seed,
                                                               const unsigned int* __restrict__ retrace, const unsigned int* __restrict__ n_retrace)
{
    const unsigned int n = *n_retrace;
    for (unsigned int j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
        long long i = retrace[j];
        float4 a = __ldg(rays + 2 * i), b = __ldg(rays + 2 * i + 1);
        Ray r; r.o = v3(a.x, a.y, a.z); r.d = v3(b.x, b.y, b.z);
        RngKey key = make_key(seed, (uint32_t)i, (uint32_t)(i >> 32), 0);
        hits[i] = trace_ref(s, r, a.w, b.w, key, 0);
    }
}
__global__ void __launch_bounds__(BLOCK) k_split_hits(const Hit* __restrict__ hits, long long n, uint32_t* __restrict__ prim, float* __restrict__ t)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        Hit h = hits[i];
        prim[i] = h.prim; t[i] = h.prim == HRT_MISS ? 0.0f : h.t;
    }
}
__global__ void __launch_bounds__(BLOCK) k_eval_batch(DScene s, const float4* __restrict__ rays, const Hit* __restrict__ hits, long long n,
                                                      float* __restrict__ normal, float* __restrict__ uv, uint8_t* __restrict__ front)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        Hit h = hits[i];
        Rec rec; rec.n = v3(0.0f); rec.u = 0; rec.v = 0; rec.front = 0;
        if (h.prim != HRT_MISS) {
            float4 a = __ldg(rays + 2 * i), b = __ldg(rays + 2 * i + 1);
            Ray r; r.o = v3(a.x, a.y, a.z); r.d = v3(b.x, b.y, b.z);
            eval_hit(s, r, h, rec);
        }
        if (normal) { normal[i * 3] = rec.n.x; normal[i * 3 + 1] = rec.n.y; normal[i * 3 + 2] = rec.n.z; }
        if (uv) { uv[i * 2] = rec.u; uv[i * 2 + 1] = rec.v; }
        if (front) front[i] = rec.front ? 1 : 0;
    }
}

// ---- K1 ---------------------------------------------------------------------------------------------------------------
// path slot = s_local * npix + pixel; the slot is carried in ray.o.w
__global__ void __launch_bounds__(BLOCK) k_raygen(RenderJob j, DCamera cam, int s_first, int s_count)
{
    const long long n = (long long)j.npix * s_count;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long slot = (long long)blockIdx.x * blockDim.x + threadIdx.x; slot < n; slot += stride) {
        int pix = (int)(slot % j.npix), sl = (int)(slot / j.npix);
        RngKey key = make_key(j.seed, (uint32_t)pix, (uint32_t)(s_first + sl), j.rng_mode);
        Ray r = primary_ray(cam, j.width, j.height, pix, key);
        j.rays[0][2 * slot] = make_float4(r.o.x, r.o.y, r.o.z, __int_as_float((int)slot));
        j.rays[0][2 * slot + 1] = make_float4(r.d.x, r.d.y, r.d.z, 0.0f);
        j.atten[slot] = make_float4(1.0f, 1.0f, 1.0f, 0.0f);
        j.radiance[slot] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) j.counts[0] = (unsigned int)n;
}

// ---- K2 ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(BLOCK) k_trace(DScene s, RenderJob j, int bounce, int s_first)
{
    const unsigned int n = j.counts[bounce];
    const float4* __restrict__ rays = j.rays[bounce & 1];
    const unsigned int stride = gridDim.x * blockDim.x;
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        float4 a = rays[2 * (size_t)i], b = rays[2 * (size_t)i + 1];
        Ray r; r.o = v3(a.x, a.y, a.z); r.d = v3(b.x, b.y, b.z);
        int slot = __float_as_int(a.w);
        RngKey key = make_key(j.seed, (uint32_t)(slot % j.npix), (uint32_t)(s_first + slot / j.npix), j.rng_mode);
        Hit h;
        if (s.fast) {
            bool need_ref = false;
            h = trace_fast_try<false>(s, r, j.t_min, hrt_inf(), key, (uint32_t)bounce, nullptr, need_ref);
            if (need_ref) { j.retrace[atomicAdd(&j.n_retrace[bounce], 1u)] = i; continue; }
        } else h = trace_ref(s, r, j.t_min, hrt_inf(), key, (uint32_t)bounce);
        j.hits[i] = h;
    }
}
__global__ void __launch_bounds__(BLOCK) k_trace_retrace(DScene s, RenderJob j, int bounce, int s_first)
{
    const unsigned int n = j.n_retrace[bounce];
    const float4* __restrict__ rays = j.rays[bounce & 1];
    for (unsigned int q = blockIdx.x * blockDim.x + threadIdx.x; q < n; q += gridDim.x * blockDim.x) {
        unsigned int i = j.retrace[q];
        float4 a = rays[2 * (size_t)i], b = rays[2 * (size_t)i + 1];
        Ray r; r.o = v3(a.x, a.y, a.z); r.d = v3(b.x, b.y, b.z);
        int slot = __float_as_int(a.w);
        RngKey key = make_key(j.seed, (uint32_t)(slot % j.npix), (uint32_t)(s_first + slot / j.npix), j.rng_mode);
        j.hits[i] = trace_ref(s, r, j.t_min, hrt_inf(), key, (uint32_t)bounce);
    }
}

// ---- K4 + K3 ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(BLOCK) k_shade(DScene s, RenderJob j, int bounce, int s_first, int last)
{
    const unsigned int n = j.counts[bounce];
    const float4* __restrict__ rays = j.rays[bounce & 1];
    float4* __restrict__ out = j.rays[(bounce + 1) & 1];
    const unsigned int lane = threadIdx.x & 31u;
    const unsigned int stride = gridDim.x * blockDim.x;
    for (unsigned int base = blockIdx.x * blockDim.x + threadIdx.x - lane; base < n; base += stride) {
        const unsigned int i = base + lane;
        bool go = false; Ray r; int slot = 0;
        if (i < n) {
            float4 a = rays[2 * (size_t)i], b = rays[2 * (size_t)i + 1];
            r.o = v3(a.x, a.y, a.z); r.d = v3(b.x, b.y, b.z);
            slot = __float_as_int(a.w);
            Hit h = j.hits[i];
            RngKey key = make_key(j.seed, (uint32_t)(slot % j.npix), (uint32_t)(s_first + slot / j.npix), j.rng_mode);
            float4 at4 = j.atten[slot];
            V3 atten = v3(at4.x, at4.y, at4.z), rad;
            go = shade_segment(s, h, r, atten, rad, key, (uint32_t)bounce);
            if (rad.x != 0.0f || rad.y != 0.0f || rad.z != 0.0f) {               // result += attenuation * emitted / background
                float4 acc = j.radiance[slot];
                V3 sum = v3(acc.x, acc.y, acc.z) + rad;
                j.radiance[slot] = make_float4(sum.x, sum.y, sum.z, 0.0f);
            }
            go = go && !last;                                                   // MAX_DEPTH reached: the partial sum stands (main.cpp:43,78)
            if (go) j.atten[slot] = make_float4(atten.x, atten.y, atten.z, 0.0f);
        }
        const unsigned int m = __ballot_sync(0xffffffffu, go);
        if (m) {
            unsigned int pos = 0;
            if (lane == 0) pos = atomicAdd(&j.counts[bounce + 1], (unsigned int)__popc(m));
            pos = __shfl_sync(0xffffffffu, pos, 0) + __popc(m & ((1u << lane) - 1u));
            if (go) {
                out[2 * (size_t)pos] = make_float4(r.o.x, r.o.y, r.o.z, __int_as_float(slot));
                out[2 * (size_t)pos + 1] = make_float4(r.d.x, r.d.y, r.d.z, 0.0f);
            }
        }
    }
}

// ---- K5 ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(BLOCK) k_film_add(RenderJob j, int s_count, float* __restrict__ accum)
{
    for (int pix = blockIdx.x * blockDim.x + threadIdx.x; pix < j.npix; pix += gridDim.x * blockDim.x) {
        V3 sum = v3(accum[3 * (size_t)pix], accum[3 * (size_t)pix + 1], accum[3 * (size_t)pix + 2]);
        for (int sl = 0; sl < s_count; ++sl) {                                  // fixed order: px += rayColour(...)  main.cpp:123
            float4 c = j.radiance[(size_t)sl * j.npix + pix];
            sum = sum + v3(c.x, c.y, c.z);
        }
        accum[3 * (size_t)pix] = sum.x; accum[3 * (size_t)pix + 1] = sum.y; accum[3 * (size_t)pix + 2] = sum.z;
    }
}
__global__ void __launch_bounds__(BLOCK) k_film_resolve(const float* __restrict__ accum, int npix, int spp, float* __restrict__ mean, uint8_t* __restrict__ rgb8)
{
    for (int pix = blockIdx.x * blockDim.x + threadIdx.x; pix < npix; pix += gridDim.x * blockDim.x) {
        V3 c = v3(accum[3 * (size_t)pix], accum[3 * (size_t)pix + 1], accum[3 * (size_t)pix + 2]) / (float)spp;   // main.cpp:126
        if (mean) { mean[3 * (size_t)pix] = c.x; mean[3 * (size_t)pix + 1] = c.y; mean[3 * (size_t)pix + 2] = c.z; }
        if (rgb8) tonemap_write(c, rgb8 + 3 * (size_t)pix);
    }
}

// ---- megakernel ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(BLOCK) k_render_mega(DScene s, RenderJob j, DCamera cam, int s_first, int s_count)
{
    const long long n = (long long)j.npix * s_count;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long slot = (long long)blockIdx.x * blockDim.x + threadIdx.x; slot < n; slot += stride) {
        int pix = (int)(slot % j.npix), sl = (int)(slot / j.npix);
        RngKey key = make_key(j.seed, (uint32_t)pix, (uint32_t)(s_first + sl), j.rng_mode);
        Ray r = primary_ray(cam, j.width, j.height, pix, key);
        V3 atten = v3(1.0f), result = v3(0.0f);
        for (int b = 0; b < j.max_depth; ++b) {
            Hit h = s.fast ? trace_fast<false>(s, r, j.t_min, hrt_inf(), key, (uint32_t)b, nullptr) : trace_ref(s, r, j.t_min, hrt_inf(), key, (uint32_t)b);
            V3 rad; bool go = shade_segment(s, h, r, atten, rad, key, (uint32_t)b);
            result = result + rad;
            if (!go) break;
        }
        j.radiance[slot] = make_float4(result.x, result.y, result.z, 0.0f);
    }
}

__global__ void k_count_rays(const unsigned int* __restrict__ counts, const unsigned int* __restrict__ n_retrace, int max_depth, unsigned long long* __restrict__ totals)
{
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        unsigned long long rays = 0, re = 0;
        for (int b = 0; b < max_depth; ++b) { rays += counts[b]; re += n_retrace[b]; }
        totals[0] += rays; totals[1] += re;
    }
}

int grid_for(long long n, int sm_count, int per_sm)
{
    long long want = (n + BLOCK - 1) / BLOCK;
    long long cap = (long long)sm_count * per_sm;
    if (want < 1) want = 1;
    return (int)(want < cap ? want : cap);
}

} // namespace

// ---- launchers -------------------------------------------------------------------------------------------------------------
void launch_trace_batch(const DScene& s, const float4* d_rays, long long n, Hit* d_hits, uint64_t seed,
                        unsigned int* d_retrace, unsigned int* d_n_retrace, int sm_count, cudaStream_t st, uint64_t* launches)
{
    cudaMemsetAsync(d_n_retrace, 0, sizeof(unsigned int), st);
    k_trace_batch<<<grid_for(n, sm_count, 16), BLOCK, 0, st>>>(s, d_rays, n, d_hits, seed, d_retrace, d_n_retrace);
    *launches += 1;
    if (s.fast) { k_trace_batch_retrace<<<sm_count * 4, BLOCK, 0, st>>>(s, d_rays, d_hits, seed, d_retrace, d_n_retrace); *launches += 1; }
}
void launch_split_hits(const Hit* d_hits, long long n, uint32_t* d_prim, float* d_t, int sm_count, cudaStream_t st, uint64_t* launches)
{
    k_split_hits<<<grid_for(n, sm_count, 16), BLOCK, 0, st>>>(d_hits, n, d_prim, d_t); *launches += 1;
}
void launch_eval_batch(const DScene& s, const float4* d_rays, const Hit* d_hits, long long n, float* d_normal, float* d_uv, uint8_t* d_front,
                       int sm_count, cudaStream_t st, uint64_t* launches)
{
    k_eval_batch<<<grid_for(n, sm_count, 16), BLOCK, 0, st>>>(s, d_rays, d_hits, n, d_normal, d_uv, d_front); *launches += 1;
}

void launch_render_batch(const DScene& s, const RenderJob& j, const DCamera& cam, int s_first, int s_count, float* d_accum,
                         bool mega, int sm_count, cudaStream_t st, uint64_t* launches, cudaEvent_t* ev_trace, cudaEvent_t* ev_shade)
{
    const long long n = (long long)j.npix * s_count;
    const int g = grid_for(n, sm_count, 16);
    if (mega) {
        k_render_mega<<<g, BLOCK, 0, st>>>(s, j, cam, s_first, s_count); *launches += 1;
    } else {
        cudaMemsetAsync(j.counts, 0, sizeof(unsigned int) * (size_t)(j.max_depth + 1), st);
        cudaMemsetAsync(j.n_retrace, 0, sizeof(unsigned int) * (size_t)(j.max_depth + 1), st);
        k_raygen<<<g, BLOCK, 0, st>>>(j, cam, s_first, s_count); *launches += 1;
        for (int b = 0; b < j.max_depth; ++b) {
            if (ev_trace) cudaEventRecord(ev_trace[2 * b], st);
            k_trace<<<g, BLOCK, 0, st>>>(s, j, b, s_first); *launches += 1;
            if (s.fast) { k_trace_retrace<<<sm_count * 4, BLOCK, 0, st>>>(s, j, b, s_first); *launches += 1; }
            if (ev_trace) cudaEventRecord(ev_trace[2 * b + 1], st);
            k_shade<<<g, BLOCK, 0, st>>>(s, j, b, s_first, b == j.max_depth - 1 ? 1 : 0); *launches += 1;
            if (ev_shade) cudaEventRecord(ev_shade[b], st);
        }
        k_count_rays<<<1, 32, 0, st>>>(j.counts, j.n_retrace, j.max_depth, j.totals); *launches += 1;
    }
    k_film_add<<<grid_for(j.npix, sm_count, 16), BLOCK, 0, st>>>(j, s_count, d_accum); *launches += 1;
}

void launch_film_resolve(const float* d_accum, int npix, int spp, float* d_mean, uint8_t* d_rgb8, int sm_count, cudaStream_t st, uint64_t* launches)
{
    k_film_resolve<<<grid_for(npix, sm_count, 16), BLOCK, 0, st>>>(d_accum, npix, spp, d_mean, d_rgb8); *launches += 1;
}

} // namespace hrt
