// hrt_kernels.cu — the sm_100a kernels of the persistent-wavefront path tracer and their launchers.
//
//   k_wave_init    K1: fills the first queue with camera paths (Philox jitter)      (main.cpp:115-123, camera.h:32-39)
//   k_wave_trace   K2: closest hit per queued ray                                    (world->hit, main.cpp:45; hrt_trace.cuh)
//   k_wave_shade   K4: hit record -> emitted + scatter / background                  (main.cpp:47-76, material.h, texture.cpp)
//                  K3: warp-ballot + one atomic per warp appends the surviving rays to the front of the next queue;
//                  K1: every finished path is replaced by a fresh camera path taken from a global counter (appended
//                      from the back of the next queue), so the queue stays full until the image runs out of samples;
//                  K5: a finished path adds its radiance to the fixed-point film (integer atomics: order independent)
//   k_film_resolve K5: film -> float sums / mean -> ACES -> sqrt -> 8 bit            (main.cpp:126-131, film.cpp:25-52)
//   k_trace_batch / k_eval_batch   K7: hit() on a caller-supplied ray batch (parity / micro-benchmark entry)
//   k_render_mega  one thread per path, whole rayColour loop (validation of the wavefront pipeline)
//
// All kernels are grid-stride loops over device-resident counts: the host never waits for a queue length.
// Compiled with -fmad=false (numerics contract, hrt_math.cuh).
#include <cuda_runtime.h>

#include <cstdlib>
#include <mutex>

#include "hrt_kernels.h"
#include "hrt_shade.cuh"
#include "hrt_wide.cuh"

namespace hrt {

namespace {

constexpr int BLOCK = 128;

__device__ __forceinline__ RngKey make_key(uint64_t seed, uint32_t index, uint32_t sample, int mode)
{
    RngKey k; k.k0 = (uint32_t)seed; k.k1 = (uint32_t)(seed >> 32); k.index = index; k.sample = sample; k.mode = mode;
    return k;
}

// ---- K7 ---------------------------------------------------------------------------------------------------------------
#ifndef HRT_BATCH_MIN_BLOCKS
#define HRT_BATCH_MIN_BLOCKS 5      // 96 registers: ray batch 1355 (4 blocks) / 1493 (5) / 1453 (6) / 1078 (8) Mrays/s on B200 (profiles/README.md r2f)
#endif
// Pending rays (pass 1 could not decide: hrt_wide.cuh) are listed here and settled by k_resolve_* afterwards.
struct PendList { unsigned int* idx; unsigned int* count; unsigned int cap; };

// ---- bounds-checked build (-DHRT_BOUNDS_CHECK, tools/build_variant.sh bounds "-DHRT_BOUNDS_CHECK") ------------------------------------
// compute-sanitizer is closed on the GPU pool this was developed on, so the indexed stores of the queue machinery carry their own
// checks in a variant build: an index outside its array is counted per site and the store is skipped; tools/sanitize_case.py runs
// every kernel family on small cases against that build and expects all counters at zero (site 7 is the self test: it must count).
// In the product build HRT_IN_BOUNDS is the constant true.
#if defined(HRT_BOUNDS_CHECK)
__device__ unsigned int g_bounds[8];
__device__ __forceinline__ bool bounds_ok(unsigned long long i, unsigned long long n, int site) { if (i < n) return true; atomicAdd(&g_bounds[site], 1u); return false; }
#define HRT_IN_BOUNDS(i, n, site) bounds_ok((unsigned long long)(i), (unsigned long long)(n), site)
__global__ void k_bounds_selftest() { if (HRT_IN_BOUNDS(5, 3, 7)) g_bounds[6] = 0xdeadu; }
#else
#define HRT_IN_BOUNDS(i, n, site) true
#endif
enum { BS_PEND = 0, BS_CAMERA_APPEND = 1, BS_SHADE_APPEND = 2, BS_FILM = 3, BS_HITS = 4, BS_BATCH_HITS = 5 };

__device__ __forceinline__ void pend_push(const PendList& p, unsigned int i)
{
    const unsigned int k = atomicAdd(p.count, 1u);
    if (HRT_IN_BOUNDS(k, p.cap, BS_PEND)) p.idx[k] = i;
}

// Persistent warps over the 8-wide trees: a lane whose walk is over takes the next ray from a global counter while the other
// lanes keep theirs, and the warp advances in phases (hrt_wide.cuh).  Scenes without wide trees: one ray per thread (_plain).
__global__ void __launch_bounds__(BLOCK) k_trace_batch_plain(DScene s, const float4* __restrict__ rays, long long n, Hit* __restrict__ hits, uint64_t seed)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        float4 a = __ldg(rays + 2 * i), b = __ldg(rays + 2 * i + 1);
        Ray r; r.o = v3(a.x, a.y, a.z); r.d = v3(b.x, b.y, b.z);
        const RngKey key = make_key(seed, (uint32_t)i, (uint32_t)(i >> 32), 0);
        hits[i] = s.fast ? trace_fast<false>(s, r, a.w, b.w, key, 0, nullptr) : trace_ref(s, r, a.w, b.w, key, 0);
    }
}
template <bool SLOW>
__global__ void __launch_bounds__(BLOCK, HRT_BATCH_MIN_BLOCKS) k_trace_batch(DScene s, const float4* __restrict__ rays, long long n, Hit* __restrict__ hits, uint64_t seed,
                                                                            unsigned long long* __restrict__ counter, PendList pend)
{
    __shared__ int stack_mem[WIDE_STACK * BLOCK];
    WideSmemStack<BLOCK> st; st.base = stack_mem + threadIdx.x; st.sp = 0;
    const unsigned int lane = threadIdx.x & 31u;
    WideLane L; L.live = false;
    long long idx = -1; bool exhausted = false;
    for (;;) {
        const int walking = __popc(__ballot_sync(0xffffffffu, L.live));
        if (walking < (exhausted ? 1 : s.wide_refill)) {
            if (idx >= 0 && !L.live) {                                            // a finished walk: its record, or its place on the pending list
                const Hit h = wide_result(L);
                if (HRT_IN_BOUNDS(idx, n, BS_BATCH_HITS)) hits[idx] = h;
                if (h.flags & HIT_PENDING) pend_push(pend, (unsigned int)idx);
                idx = -1;
            }
            if (!exhausted) {
                const unsigned int m = __ballot_sync(0xffffffffu, idx < 0);
                unsigned long long base = 0;
                if (lane == (unsigned int)(__ffs(m) - 1)) base = atomicAdd(counter, (unsigned long long)__popc(m));
                base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
                const long long my = (long long)base + __popc(m & ((1u << lane) - 1u));
                if (idx < 0 && my < n) {
                    float4 a = __ldg(rays + 2 * my), b = __ldg(rays + 2 * my + 1);
                    Ray r; r.o = v3(a.x, a.y, a.z); r.d = v3(b.x, b.y, b.z);
                    Hit early;
                    st.sp = 0;
                    if (wide_begin(L, s, r, a.w, b.w, early)) idx = my;
                    else {                                                        // nothing to walk: answered on the spot, or trace_ref's business
                        if (L.b.bad) { early = wide_result(L); pend_push(pend, (unsigned int)my); }
                        hits[my] = early;
                    }
                }
                if ((long long)base + __popc(m) >= n) exhausted = true;
            }
        }
        const RngKey key = make_key(seed, (uint32_t)idx, (uint32_t)((unsigned long long)idx >> 32), 0);
        if (!wide_round<WideSmemStack<BLOCK>, SLOW>(L, s, st, key, 0) && exhausted) {
            if (!__any_sync(0xffffffffu, idx >= 0)) break;
        }
    }
}
// pass 2 of the rays k_trace_batch left pending
__global__ void __launch_bounds__(BLOCK) k_resolve_batch(DScene s, const float4* __restrict__ rays, Hit* __restrict__ hits, uint64_t seed, PendList pend)
{
    // entry k goes to warp k mod (number of warps): a short list gives every ray a warp of its own - the ordered pass 2 of 32 unrelated rays
    // in one warp runs every lane's loop to the longest one with both sides of every branch
    const unsigned int n = *pend.count;
    const unsigned int nw = gridDim.x * (blockDim.x >> 5), wid = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31u;
    for (unsigned int k = wid + nw * lane; k < n; k += nw * 32u) {
        const unsigned int i = pend.idx[k];
        float4 a = __ldg(rays + 2 * (size_t)i), b = __ldg(rays + 2 * (size_t)i + 1);
        Ray r; r.o = v3(a.x, a.y, a.z); r.d = v3(b.x, b.y, b.z);
        hits[i] = wide_resolve(s, r, a.w, b.w, make_key(seed, i, 0u, 0), 0, hits[i]);
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

// ---- Texture::colourValue / Material::scatter + emitted on caller-supplied batches (parity entries, like K7 for hit()) -----------------
// Camera::getRay(s, t) on a batch (camera.h:32-39; parity entry hrt_camera_rays): out = origin xyz, direction xyz (not normalised)
__global__ void __launch_bounds__(BLOCK) k_camera_rays(DCamera cam, const float* __restrict__ st, long long n, float* __restrict__ out)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const Ray r = camera_ray<false>(cam, st[2 * i], st[2 * i + 1]);
        float* o = out + 6 * i;
        o[0] = r.o.x; o[1] = r.o.y; o[2] = r.o.z; o[3] = r.d.x; o[4] = r.d.y; o[5] = r.d.z;
    }
}
__global__ void __launch_bounds__(BLOCK) k_texture_eval(DScene s, int tex, const float* __restrict__ in, long long n, float* __restrict__ out)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const V3 c = tex_value(s, tex, in[i * 5], in[i * 5 + 1], v3(in[i * 5 + 2], in[i * 5 + 3], in[i * 5 + 4]));
        out[i * 3] = c.x; out[i * 3 + 1] = c.y; out[i * 3 + 2] = c.z;
    }
}
// in: r_in.o, r_in.d, rec.p, rec.normal, rec.u, rec.v (14 floats) + frontFace; out: scattered?, attenuation, scattered.dir, emitted (10 floats).
// Random numbers: the Philox stream of (seed; index i, sample 0), bounce 0 - what a path at its first segment would draw.
__global__ void __launch_bounds__(BLOCK) k_scatter_batch(DScene s, int material, const float* __restrict__ in, const uint8_t* __restrict__ front, long long n,
                                                         uint64_t seed, int rng_mode, float* __restrict__ out)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float* a = in + i * 14;
        Ray rin; rin.o = v3(a[0], a[1], a[2]); rin.d = v3(a[3], a[4], a[5]);
        Rec rec; rec.p = v3(a[6], a[7], a[8]); rec.n = v3(a[9], a[10], a[11]); rec.u = a[12]; rec.v = a[13];
        rec.front = front[i] != 0; rec.t = 1.0f; rec.mat = material; rec.prim = 0;
        V3 att = v3(0.0f); Ray sc; sc.o = v3(0.0f); sc.d = v3(0.0f);
        const RngKey key = make_key(seed, (uint32_t)i, (uint32_t)((unsigned long long)i >> 32), rng_mode);
        const V3 em = mat_emitted(s, material, rec.u, rec.v, rec.p);
        const bool ok = mat_scatter(s, material, rin, rec, att, sc, key, 0);
        float* o = out + i * 10;
        o[0] = ok ? 1.0f : 0.0f; o[1] = att.x; o[2] = att.y; o[3] = att.z; o[4] = sc.d.x; o[5] = sc.d.y; o[6] = sc.d.z; o[7] = em.x; o[8] = em.y; o[9] = em.z;
    }
}

// ---- film ---------------------------------------------------------------------------------------------------------------
// quantisation / flag fields / reconstruction live in hrt_shade.cuh (shared with the CPU simulation harness)
// a finite sample beyond the film's +-2^FILM_MAG_BITS range is clamped (hrt_shade.cuh) - and counted, so that callers can tell
__device__ __forceinline__ void count_clamped(unsigned long long* counter, float v)
{
    if (fabsf(v) > (float)(1 << FILM_MAG_BITS) && fabsf(v) <= 3.402823466e+38f) atomicAdd(counter, 1ull);
}
__device__ __forceinline__ void film_add(long long* film, int pix, V3 c, unsigned long long* clamped, int npix)
{
    if (!HRT_IN_BOUNDS(pix, npix, BS_FILM)) return;
    long long* px = film + (size_t)pix * FILM_LANES;
    const float v[3] = {c.x, c.y, c.z};
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
        if (v[ch] == 0.0f) continue;
        count_clamped(clamped, v[ch]);
        long long q; unsigned long long flag;
        film_quantize(v[ch], ch, q, flag);
        if (flag) atomicOr((unsigned long long*)(px + 3), flag);
        else if (q != 0) atomicAdd((unsigned long long*)(px + ch), (unsigned long long)q);
    }
}
__global__ void __launch_bounds__(BLOCK) k_film_resolve(const long long* __restrict__ film, int npix, int spp, float* __restrict__ sum,
                                                        float* __restrict__ mean, uint8_t* __restrict__ rgb8)
{
    for (int pix = blockIdx.x * blockDim.x + threadIdx.x; pix < npix; pix += gridDim.x * blockDim.x) {
        const long long* px = film + (size_t)pix * FILM_LANES;
        V3 c = v3(film_channel(px, 0), film_channel(px, 1), film_channel(px, 2));
        if (sum) { sum[3 * (size_t)pix] = c.x; sum[3 * (size_t)pix + 1] = c.y; sum[3 * (size_t)pix + 2] = c.z; }
        c = c / (float)spp;                                                     // main.cpp:126
        if (mean) { mean[3 * (size_t)pix] = c.x; mean[3 * (size_t)pix + 1] = c.y; mean[3 * (size_t)pix + 2] = c.z; }
        if (rgb8) tonemap_write(c, rgb8 + 3 * (size_t)pix);
    }
}

// Reduce + resolve in one pass over NVLink peer memory: the films of the other ranks are mapped into this process
// (CUDA IPC) and every pixel's four int64 lanes are pulled from each peer with two 128-bit loads, summed with the local
// lanes and resolved on the spot — no staging buffer, no second pass over the 32 B/pixel film.  Integer sums: the image
// is the same bit for bit as after an ncclSum reduce, whatever the number of ranks.
struct PeerFilms { const long long* p[15]; int n; };
__global__ void __launch_bounds__(BLOCK) k_film_resolve_peers(long long* __restrict__ film, PeerFilms peers, int npix, int spp, int write_back,
                                                              float* __restrict__ sum, float* __restrict__ mean, uint8_t* __restrict__ rgb8)
{
    for (int pix = blockIdx.x * blockDim.x + threadIdx.x; pix < npix; pix += gridDim.x * blockDim.x) {
        const longlong2* mine = (const longlong2*)(film + (size_t)pix * FILM_LANES);
        longlong2 a = mine[0], b = mine[1];
        for (int k = 0; k < peers.n; ++k) {
            const longlong2* q = (const longlong2*)(peers.p[k] + (size_t)pix * FILM_LANES);
            const longlong2 c = q[0], d = q[1];                                // NVLink reads, 16 B each, coalesced across the warp
            a.x += c.x; a.y += c.y; b.x += d.x; b.y += d.y;                     // lane 3 holds <= 7-bit counters per field: sums of <= 16 ranks cannot carry
        }
        long long px[4] = {a.x, a.y, b.x, b.y};
        if (write_back) { longlong2* o = (longlong2*)(film + (size_t)pix * FILM_LANES); o[0] = a; o[1] = b; }
        V3 c = v3(film_channel(px, 0), film_channel(px, 1), film_channel(px, 2));
        if (sum) { sum[3 * (size_t)pix] = c.x; sum[3 * (size_t)pix + 1] = c.y; sum[3 * (size_t)pix + 2] = c.z; }
        c = c / (float)spp;                                                     // main.cpp:126
        if (mean) { mean[3 * (size_t)pix] = c.x; mean[3 * (size_t)pix + 1] = c.y; mean[3 * (size_t)pix + 2] = c.z; }
        if (rgb8) tonemap_write(c, rgb8 + 3 * (size_t)pix);
    }
}

// The same reduction spread over the ranks: every rank sums and resolves ITS slice of the pixels from all films (its own + the peers')
// and stores the results where the caller says - the root's buffers, mapped into this process - so no GPU gathers 32 B x pixels x
// ranks alone while the others idle (r1: 1.8 ms of a 37 ms step at 8 GPUs went into that serial gather and its two host barriers).
struct FilmSet { const long long* p[16]; int n; };
__global__ void __launch_bounds__(BLOCK) k_film_resolve_slice(FilmSet films, int pix_first, int pix_count, int spp,
                                                              float* __restrict__ sum, float* __restrict__ mean, uint8_t* __restrict__ rgb8)
{
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < pix_count; k += gridDim.x * blockDim.x) {
        const int pix = pix_first + k;
        longlong2 a = make_longlong2(0, 0), b = make_longlong2(0, 0);
        for (int f = 0; f < films.n; ++f) {
            const longlong2* q = (const longlong2*)(films.p[f] + (size_t)pix * FILM_LANES);
            const longlong2 c = q[0], d = q[1];                                // local or NVLink reads, 16 B each, coalesced across the warp
            a.x += c.x; a.y += c.y; b.x += d.x; b.y += d.y;
        }
        long long px[4] = {a.x, a.y, b.x, b.y};
        V3 c = v3(film_channel(px, 0), film_channel(px, 1), film_channel(px, 2));
        if (sum) { sum[3 * (size_t)pix] = c.x; sum[3 * (size_t)pix + 1] = c.y; sum[3 * (size_t)pix + 2] = c.z; }
        c = c / (float)spp;                                                     // main.cpp:126
        if (mean) { mean[3 * (size_t)pix] = c.x; mean[3 * (size_t)pix + 1] = c.y; mean[3 * (size_t)pix + 2] = c.z; }
        if (rgb8) tonemap_write(c, rgb8 + 3 * (size_t)pix);
    }
}

// ---- camera paths ---------------------------------------------------------------------------------------------------------
// Path index p -> (local sample, pixel).  When the film is a whole number of 8x4 pixel tiles, 32 consecutive paths are
// one tile of one sample, so a warp of fresh camera rays is a small screen-space block; otherwise pixels go in row order.
__device__ __forceinline__ void path_pixel(const WaveJob& j, unsigned long long p, int& pix, int& sl)
{
    sl = (int)(p / (unsigned long long)j.npix);
    const unsigned int k = (unsigned int)(p - (unsigned long long)sl * (unsigned long long)j.npix);
    if (j.tiles_x > 0) {
        const unsigned int lane = k & 31u, tile = k >> 5;
        const int x = (int)(tile % (unsigned int)j.tiles_x) * 8 + (int)(lane & 7u);
        const int y = (int)(tile / (unsigned int)j.tiles_x) * 4 + (int)(lane >> 3);
        pix = y * j.width + x;
    } else pix = (int)k;
}
__device__ __forceinline__ unsigned int pack_state(int sl, int bounce, bool has_result)
{ return ((unsigned int)sl << 8) | (unsigned int)bounce | (has_result ? 0x80000000u : 0u); }

// writes the camera ray of path p into entry e of queue q
__device__ __forceinline__ void write_camera_path(const WaveJob& j, const DCamera& cam, int q, unsigned int e, unsigned long long p)
{
    int pix, sl;
    path_pixel(j, p, pix, sl);
    RngKey key = make_key(j.seed, (uint32_t)pix, (uint32_t)(j.s_first + sl), j.rng_mode);
    Ray r = primary_ray(cam, j.width, j.height, pix, key);
    j.rays[q][2 * (size_t)e] = make_float4(r.o.x, r.o.y, r.o.z, __int_as_float(pix));
    j.rays[q][2 * (size_t)e + 1] = make_float4(r.d.x, r.d.y, r.d.z, __uint_as_float(pack_state(sl, 0, false)));
    j.thr[q][e] = make_float4(1.0f, 1.0f, 1.0f, 0.0f);
}

__device__ __forceinline__ unsigned long long min_ull(unsigned long long a, unsigned long long b) { return a < b ? a : b; }

constexpr int SHADE_BLOCK = 256;

#if defined(HRT_WITH_V1_PIPELINE)      // the r1b / r1c pipeline (every ray queued, regeneration in the shade pass): A/B builds only
// ---- K1 (first fill) -------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(BLOCK) k_wave_init(WaveJob j, DCamera cam)
{
    const unsigned long long n = min_ull(j.ctl->total_paths, (unsigned long long)j.capacity);
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long p = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += stride)
        write_camera_path(j, cam, 0, j.capacity - 1u - (unsigned int)p, p);
}
// runs after k_wave_init in stream order: publishes the first queue
__global__ void k_wave_publish(WaveJob j)
{
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        WaveCtl* c = j.ctl;
        c->n_front[0] = 0; c->n_front[1] = 0; c->base[0] = 0; c->base[1] = 0;
        c->next_path = min_ull(c->total_paths, (unsigned long long)j.capacity);
    }
}

// entry index of thread slot i in a queue with n_front survivors: survivors first, then the regenerated paths
__device__ __forceinline__ unsigned int entry_of(unsigned int i, unsigned int n_front, unsigned int capacity)
{ return i < n_front ? i : capacity - 1u - (i - n_front); }

// ---- K2 ---------------------------------------------------------------------------------------------------------------
#ifndef HRT_TRACE_MIN_BLOCKS
#define HRT_TRACE_MIN_BLOCKS 8      // 64 registers: +10 % over 128 registers / 4 blocks on shiny_teapot (A/B on B200, profiles/README.md)
#endif
__global__ void __launch_bounds__(BLOCK, HRT_TRACE_MIN_BLOCKS) k_wave_trace(DScene s, WaveJob j, int q)
{
    WaveCtl* c = j.ctl;
    const unsigned int n_front = c->n_front[q];
    const unsigned long long handed = min_ull(c->next_path, c->total_paths);
    const unsigned int n = n_front + (unsigned int)(handed - c->base[q]);
    if (blockIdx.x == 0 && threadIdx.x == 0) {     // nobody reads these in this kernel: prepare the queue the shade pass will fill
        c->rays += n;
        c->n_front[q ^ 1] = 0;
        c->base[q ^ 1] = handed;
    }
    const float4* __restrict__ rays = j.rays[q];
    const unsigned int stride = gridDim.x * blockDim.x;
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const unsigned int e = entry_of(i, n_front, j.capacity);
        float4 a = rays[2 * (size_t)e], b = rays[2 * (size_t)e + 1];
        Ray r; r.o = v3(a.x, a.y, a.z); r.d = v3(b.x, b.y, b.z);
        const unsigned int st = __float_as_uint(b.w);
        RngKey key = make_key(j.seed, (uint32_t)__float_as_int(a.w), (uint32_t)(j.s_first + (int)((st >> 8) & 0x7fffffu)), j.rng_mode);
        const Hit h = s.fast ? trace_fast<false>(s, r, j.t_min, hrt_inf(), key, st & 0xffu, nullptr) : trace_ref(s, r, j.t_min, hrt_inf(), key, st & 0xffu);
        if (h.flags & (HIT_RESOLVED | HIT_FALLBACK)) {
            const unsigned long long k = atomicAdd((h.flags & HIT_FALLBACK) ? &c->fallbacks : &c->resolved, 1ull);
            if ((h.flags & HIT_FALLBACK) && k < 16) { float* g = c->dbg[k]; g[0] = a.x; g[1] = a.y; g[2] = a.z; g[3] = a.w; g[4] = b.x; g[5] = b.y; g[6] = b.z; g[7] = b.w; }
        }
        j.hits[i] = h;
    }
}

// ---- K4 + K3 + K1 + K5 ------------------------------------------------------------------------------------------------------
// One queue-slot claim per BLOCK and pass of the loop (two returning atomics on two hot addresses): a per-warp claim
// made every warp wait ~1.5 us for the round trip to L2 and serialised 260 k same-address atomics per pass (profiles/ r1b).
__global__ void __launch_bounds__(SHADE_BLOCK) k_wave_shade(DScene s, WaveJob j, DCamera cam, int q)
{
    __shared__ unsigned int w_go[SHADE_BLOCK / 32], w_fresh[SHADE_BLOCK / 32];
    __shared__ unsigned int s_pos; __shared__ unsigned long long s_p0;
    WaveCtl* c = j.ctl;
    const int o = q ^ 1;
    const unsigned int n_front = c->n_front[q];
    const unsigned long long base_out = c->base[o];                             // = paths handed out before this pass (set by k_wave_trace)
    const unsigned int n = n_front + (unsigned int)(base_out - c->base[q]);
    const unsigned long long total = c->total_paths;
    const float4* __restrict__ rays = j.rays[q];
    const unsigned int lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const unsigned int stride = gridDim.x * SHADE_BLOCK;
    for (unsigned int base = blockIdx.x * SHADE_BLOCK; base < n; base += stride) {
        const unsigned int i = base + threadIdx.x;
        bool go = false, fresh = false; Ray r; V3 atten, result; int pix = 0; unsigned int st = 0; bool has_result = false;
        r.o = v3(0.0f); r.d = v3(0.0f); atten = v3(0.0f); result = v3(0.0f);
        if (i < n) {
            const unsigned int e = entry_of(i, n_front, j.capacity);
            float4 a = rays[2 * (size_t)e], b = rays[2 * (size_t)e + 1];
            r.o = v3(a.x, a.y, a.z); r.d = v3(b.x, b.y, b.z);
            pix = __float_as_int(a.w); st = __float_as_uint(b.w);
            const Hit h = j.hits[i];
            const unsigned int bounce = st & 0xffu; const int sl = (int)((st >> 8) & 0x7fffffu);
            has_result = (st >> 31) != 0;
            RngKey key = make_key(j.seed, (uint32_t)pix, (uint32_t)(j.s_first + sl), j.rng_mode);
            float4 at4 = j.thr[q][e];
            atten = v3(at4.x, at4.y, at4.z);
            if (has_result) { float4 r4 = j.res[q][e]; result = v3(r4.x, r4.y, r4.z); }
            V3 rad;
            go = shade_segment(s, h, r, atten, rad, key, bounce);
            if (rad.x != 0.0f || rad.y != 0.0f || rad.z != 0.0f) { result = result + rad; has_result = true; }   // result += attenuation * emitted / background
            go = go && (int)bounce + 1 < j.max_depth;                           // MAX_DEPTH reached: the partial sum stands (main.cpp:43,78)
            if (go) st = pack_state(sl, (int)bounce + 1, has_result);
            else {
                if (has_result) film_add(j.film, pix, result, &c->clamped, j.npix);                  // px += rayColour(...)  main.cpp:123
                fresh = true;
            }
        }
        const unsigned int m_go = __ballot_sync(0xffffffffu, go), m_fresh = __ballot_sync(0xffffffffu, fresh);
        if (lane == 0) { w_go[warp] = (unsigned int)__popc(m_go); w_fresh[warp] = (unsigned int)__popc(m_fresh); }
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned int t_go = 0, t_fresh = 0;
#pragma unroll
            for (int w = 0; w < SHADE_BLOCK / 32; ++w) { unsigned int g = w_go[w], f = w_fresh[w]; w_go[w] = t_go; w_fresh[w] = t_fresh; t_go += g; t_fresh += f; }
            s_pos = t_go ? atomicAdd(&c->n_front[o], t_go) : 0u;
            s_p0 = t_fresh ? atomicAdd(&c->next_path, (unsigned long long)t_fresh) : 0ull;
        }
        __syncthreads();
        const unsigned int below = (1u << lane) - 1u;
        if (go) {
            const unsigned int e2 = s_pos + w_go[warp] + (unsigned int)__popc(m_go & below);
            j.rays[o][2 * (size_t)e2] = make_float4(r.o.x, r.o.y, r.o.z, __int_as_float(pix));
            j.rays[o][2 * (size_t)e2 + 1] = make_float4(r.d.x, r.d.y, r.d.z, __uint_as_float(st));
            j.thr[o][e2] = make_float4(atten.x, atten.y, atten.z, 0.0f);
            if (has_result) j.res[o][e2] = make_float4(result.x, result.y, result.z, 0.0f);
        } else if (fresh) {
            const unsigned long long p = s_p0 + (unsigned long long)(w_fresh[warp] + (unsigned int)__popc(m_fresh & below));
            if (p < total) write_camera_path(j, cam, o, j.capacity - 1u - (unsigned int)(p - base_out), p);
        }
        __syncthreads();                                                        // the next pass of the loop rewrites the shared counters
    }
}

#endif // HRT_WITH_V1_PIPELINE

// ---- fused-primary wavefront (the default pipeline) ---------------------------------------------------------------------------
// profiles/ r1c: on shiny_teapot 90 % of all rays are camera rays and 85 % of those leave the scene at once, yet every
// one of them was written to a queue (64 B), read back twice, given a 16 B hit record and shaded by a second kernel whose
// time went to waiting for those streams (k_wave_shade: 3.3 TB/s of DRAM traffic, issue slots 42 % busy).  Here a camera
// path never touches a queue unless it HITS something:
//   k_wave2_trace, one persistent grid: (1) secondary rays of queue q -> hits[]; (2) every warp then claims CLAIM camera
//       paths at a time from the global counter, generates them (K1), traces them (K2) and
//         miss -> background lookup + film add on the spot (K4/K5: the path is finished),
//         hit  -> appended behind the secondary entries (warp ballot + one atomic per warp, K3) for the shade pass,
//       until the image is out of paths or the queue is (conservatively) full;
//   k_wave2_shade: shades the n_sec + n_prim entries, compacts the survivors to the front of queue q^1.
// The film is integer, so the image is bit-identical to the other pipelines whatever the interleaving.
// A claim is one warp-sized pixel unit (an 8x4 tile, or 32 consecutive pixels when the film is not whole tiles) times
// CLAIM_SPP consecutive samples: the lane keeps its pixel, loops over the samples and sums the finished (missed) ones as
// fixed-point integers in registers — integer sums associate, so one atomic per channel and claim leaves the same film
// as one atomic per sample.  WaveCtl::next_path / total_paths count CLAIMS in this pipeline.
#ifndef HRT_CLAIM_SPP
#define HRT_CLAIM_SPP 4      // A/B on B200 (profiles/README.md r1d): 4 beats 2 and 8 on all four scenes
#endif
constexpr int CLAIM_SPP = HRT_CLAIM_SPP;
constexpr int CLAIM = 32 * CLAIM_SPP;           // camera paths per claim (upper bound of what a warp appends after a room check)

__device__ __forceinline__ unsigned int ld_relaxed(const unsigned int* p) { return *(const volatile unsigned int*)p; }
__device__ __forceinline__ unsigned long long ld_relaxed(const unsigned long long* p) { return *(const volatile unsigned long long*)p; }

__host__ __device__ inline unsigned int wave2_units(const WaveJob& j) { return j.tiles_x > 0 ? (unsigned int)(j.tiles_x * j.tiles_y) : ((unsigned int)j.npix + 31u) / 32u; }
__host__ __device__ inline unsigned int wave2_sblocks(const WaveJob& j) { return ((unsigned int)j.s_count + CLAIM_SPP - 1u) / CLAIM_SPP; }

struct FilmAcc { long long q[3]; unsigned long long flags; };
__device__ __forceinline__ void acc_add(FilmAcc& a, V3 c, unsigned long long* clamped)
{
    const float v[3] = {c.x, c.y, c.z};
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
        count_clamped(clamped, v[ch]);
        long long q; unsigned long long flag;
        film_quantize(v[ch], ch, q, flag);
        a.q[ch] += q; a.flags |= flag;
    }
}
__device__ __forceinline__ void acc_flush(FilmAcc& a, long long* film, int pix, int npix)
{
    if (!HRT_IN_BOUNDS(pix, npix, BS_FILM)) return;
    long long* px = film + (size_t)pix * FILM_LANES;
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) if (a.q[ch] != 0) { atomicAdd((unsigned long long*)(px + ch), (unsigned long long)a.q[ch]); a.q[ch] = 0; }
    if (a.flags) { atomicOr((unsigned long long*)(px + 3), a.flags); a.flags = 0; }
}

// Two builds of the same kernel: 128 registers / 4 blocks per SM when camera paths dominate (shiny_teapot: +8 %), 64 registers /
// 8 blocks per SM when incoherent secondary rays dominate (cornell, scattered balls, teapot_scene: +7-11 %; latency hiding
// beats spilling there).  The host picks per chunk from the secondary-rays-per-camera-path ratio it reads in the control block.
// Camera rays start outside most scenes and the larger part of a frame never comes near an object: such a ray is settled by
// one slab test against the union of the world-list entries' padded bounds (the boxes the top-level tree is built from),
// before trace_fast spends three IEEE reciprocals and the entry set-up of a walk on it.  Approximate reciprocals, so the
// interval is widened by far more than their error; anything unusual about the ray (tiny, huge or non-finite components)
// answers "cannot tell" and takes the full trace.
__device__ __forceinline__ bool camera_ray_clears_scene(const DScene& s, const Ray& r, float t_min)
{
    const float lo = 1.0e-18f, hi = 1.0e18f;
    const float dx = fabsf(r.d.x), dy = fabsf(r.d.y), dz = fabsf(r.d.z);
    if (!(dx > lo && dy > lo && dz > lo && dx < hi && dy < hi && dz < hi)) return false;
    if (!(fabsf(r.o.x) < hi && fabsf(r.o.y) < hi && fabsf(r.o.z) < hi)) return false;
    const float ix = __fdividef(1.0f, r.d.x), iy = __fdividef(1.0f, r.d.y), iz = __fdividef(1.0f, r.d.z);
    const float ax = (s.bounds_lo[0] - r.o.x) * ix, bx = (s.bounds_hi[0] - r.o.x) * ix;
    const float ay = (s.bounds_lo[1] - r.o.y) * iy, by = (s.bounds_hi[1] - r.o.y) * iy;
    const float az = (s.bounds_lo[2] - r.o.z) * iz, bz = (s.bounds_hi[2] - r.o.z) * iz;
    float tn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz));
    float tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
    tn -= fabsf(tn) * 1.0e-5f; tf += fabsf(tf) * 1.0e-5f;
    return tn > tf || tf < fminf(t_min, 0.0f);
}

// WIDE: the walks of a step of the warp (32 entries, or the 32 camera rays of one sample of the claim) go through the 8-wide trees
// together, in phases (hrt_wide.cuh, no refill: the claim structure stays); used on scenes with large meshes, where the
// two-child walk is a chain of ~50 dependent fetches per ray.
template <int MINB, bool LENS, bool WIDE, bool SLOW = true>
__global__ void __launch_bounds__(BLOCK, MINB) k_wave2_trace(DScene s, WaveJob j, DCamera cam, int q, int skip_secondary)
{
    __shared__ int stack_mem[WIDE ? WIDE_STACK * BLOCK : 1];
    WaveCtl* c = j.ctl;
    const unsigned int n_sec = c->n_front[q];
    const unsigned long long total_claims = c->total_paths;
    if (blockIdx.x == 0 && threadIdx.x == 0) { c->rays += n_sec; c->n_front[q ^ 1] = 0; }   // nobody reads these in this kernel
    const unsigned int lane = threadIdx.x & 31u;
    const unsigned int warps_total = gridDim.x * (BLOCK / 32);
    const unsigned int stride = warps_total * 32u;
    const unsigned int room = j.capacity - n_sec;                                 // slots left for camera-path hits
    const unsigned int margin = warps_total * (unsigned int)CLAIM;                // what the other warps may still append after a check
    const unsigned int n_units = wave2_units(j);
    float4* __restrict__ rays = j.rays[q];
    unsigned int sec_i = skip_secondary ? n_sec + lane : (blockIdx.x * (BLOCK / 32) + (threadIdx.x >> 5)) * 32u + lane;   // skip: k_wide_secondary did them
    int c_left = 0, c_sl = 0, c_pix = -1, c_x = 0, c_y = 0;                         // the warp's claim: pixel of this lane, next sample, samples left
    FilmAcc acc; acc.q[0] = acc.q[1] = acc.q[2] = 0; acc.flags = 0;
    for (;;) {
        Ray r; r.o = v3(0.0f); r.d = v3(0.0f);
        unsigned int st = 0; int pix = 0; bool active, primary;
        if (sec_i - lane < n_sec) {                                                 // warp-uniform: secondary entries first
            primary = false; active = sec_i < n_sec;
            if (active) {
                float4 a = rays[2 * (size_t)sec_i], b = rays[2 * (size_t)sec_i + 1];
                r.o = v3(a.x, a.y, a.z); r.d = v3(b.x, b.y, b.z);
                pix = __float_as_int(a.w); st = __float_as_uint(b.w);
            }
        } else {
            primary = true;
            if (c_left == 0) {
                if (c_pix >= 0) acc_flush(acc, j.film, c_pix, j.npix);
                unsigned long long k0 = 0;
                if (lane == 0 && ld_relaxed(&c->n_prim[q]) + margin <= room && ld_relaxed(&c->next_path) < total_claims)
                    k0 = atomicAdd(&c->next_path, 1ull) + 1ull;
                k0 = __shfl_sync(0xffffffffu, k0, 0);
                if (k0 == 0 || k0 - 1ull >= total_claims) break;
                const unsigned int k = (unsigned int)(k0 - 1ull);
                const unsigned int sb = k / n_units, unit = k - sb * n_units;        // claims of one sample block sweep the film
                c_sl = (int)sb * CLAIM_SPP;
                c_left = j.s_count - c_sl < CLAIM_SPP ? j.s_count - c_sl : CLAIM_SPP;
                if (j.tiles_x > 0) {
                    c_x = (int)(unit % (unsigned int)j.tiles_x) * 8 + (int)(lane & 7u);
                    c_y = (int)(unit / (unsigned int)j.tiles_x) * 4 + (int)(lane >> 3);
                    c_pix = c_y * j.width + c_x;
                } else {
                    const unsigned int pp = unit * 32u + lane;
                    c_pix = pp < (unsigned int)j.npix ? (int)pp : -1;
                    c_x = c_pix >= 0 ? c_pix % j.width : 0; c_y = c_pix >= 0 ? c_pix / j.width : 0;
                }
            }
            active = c_pix >= 0;
            pix = c_pix;
            if (active) {
                RngKey key = make_key(j.seed, (uint32_t)pix, (uint32_t)(j.s_first + c_sl), j.rng_mode);
                r = primary_ray_xy<LENS>(cam, j.width, j.height, c_x, j.height - c_y, key);
                st = pack_state(c_sl, 0, false);
            }
            ++c_sl; --c_left;
        }
        Hit h; h.t = 0.0f; h.prim = HRT_MISS; h.obj = -1; h.flags = 0;
        const bool walk = active && !(primary && s.has_bounds && camera_ray_clears_scene(s, r, j.t_min));
        const RngKey key = make_key(j.seed, (uint32_t)pix, (uint32_t)(j.s_first + (int)((st >> 8) & 0x7fffffu)), j.rng_mode);
        if (WIDE) {
            WideSmemStack<BLOCK> stk; stk.base = stack_mem + threadIdx.x; stk.sp = 0;
            WideLane L; L.live = false; L.b.bad = 0;
            bool began = false;
            if (walk) began = wide_begin(L, s, r, j.t_min, hrt_inf(), h);
            while (wide_round<WideSmemStack<BLOCK>, SLOW>(L, s, stk, key, st & 0xffu)) { }
            if (walk && (began || L.b.bad)) {
                h = wide_result(L);
                if (h.flags & HIT_PENDING) h = wide_resolve(s, r, j.t_min, hrt_inf(), key, st & 0xffu, h);
            }
        } else if (walk) h = s.fast ? trace_fast<false>(s, r, j.t_min, hrt_inf(), key, st & 0xffu, nullptr) : trace_ref(s, r, j.t_min, hrt_inf(), key, st & 0xffu);
        if (h.flags & (HIT_RESOLVED | HIT_FALLBACK)) atomicAdd((h.flags & HIT_FALLBACK) ? &c->fallbacks : &c->resolved, 1ull);
        if (!primary) {
            if (active && HRT_IN_BOUNDS(sec_i, j.capacity, BS_HITS)) j.hits[sec_i] = h;
            sec_i += stride;
            continue;
        }
        const bool hit = active && h.prim != HRT_MISS;
        if (active && !hit) acc_add(acc, background_colour(s, r), &c->clamped);                  // main.cpp:47-59: the path ends on the background
        const unsigned int m_hit = __ballot_sync(0xffffffffu, hit);
        if (m_hit) {
            unsigned int e0 = 0;
            if (lane == 0) e0 = atomicAdd(&c->n_prim[q], (unsigned int)__popc(m_hit));
            e0 = __shfl_sync(0xffffffffu, e0, 0);
            if (hit) {
                const unsigned int e = n_sec + e0 + (unsigned int)__popc(m_hit & ((1u << lane) - 1u));
                if (HRT_IN_BOUNDS(e, j.capacity, BS_CAMERA_APPEND)) {
                    rays[2 * (size_t)e] = make_float4(r.o.x, r.o.y, r.o.z, __int_as_float(pix));
                    rays[2 * (size_t)e + 1] = make_float4(r.d.x, r.d.y, r.d.z, __uint_as_float(st));
                    j.hits[e] = h;
                }
            }
        }
    }
}

template <bool RR>
__global__ void __launch_bounds__(SHADE_BLOCK) k_wave2_shade(DScene s, WaveJob j, int q)
{
    __shared__ unsigned int w_go[SHADE_BLOCK / 32];
    __shared__ unsigned int s_pos;
    WaveCtl* c = j.ctl;
    const int o = q ^ 1;
    const unsigned int n = c->n_front[q] + c->n_prim[q];
    if (blockIdx.x == 0 && threadIdx.x == 0) { c->n_prim[o] = 0; c->sec_next[o] = 0; c->pend_count = 0; }   // the next trace pass works on queue o; nobody reads these here
    const float4* __restrict__ rays = j.rays[q];
    const unsigned int lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const unsigned int stride = gridDim.x * SHADE_BLOCK;
    for (unsigned int base = blockIdx.x * SHADE_BLOCK; base < n; base += stride) {
        const unsigned int i = base + threadIdx.x;
        bool go = false; Ray r; V3 atten, result; int pix = 0; unsigned int st = 0; bool has_result = false;
        r.o = v3(0.0f); r.d = v3(0.0f); atten = v3(1.0f); result = v3(0.0f);
        if (i < n) {
            float4 a = rays[2 * (size_t)i], b = rays[2 * (size_t)i + 1];
            r.o = v3(a.x, a.y, a.z); r.d = v3(b.x, b.y, b.z);
            pix = __float_as_int(a.w); st = __float_as_uint(b.w);
            const Hit h = j.hits[i];
            const unsigned int bounce = st & 0xffu; const int sl = (int)((st >> 8) & 0x7fffffu);
            has_result = (st >> 31) != 0;
            RngKey key = make_key(j.seed, (uint32_t)pix, (uint32_t)(j.s_first + sl), j.rng_mode);
            if (bounce > 0) { float4 at4 = j.thr[q][i]; atten = v3(at4.x, at4.y, at4.z); }      // a camera path starts with attenuation 1 (main.cpp:40)
            if (has_result) { float4 r4 = j.res[q][i]; result = v3(r4.x, r4.y, r4.z); }
            V3 rad;
            go = shade_segment(s, h, r, atten, rad, key, bounce);
            if (rad.x != 0.0f || rad.y != 0.0f || rad.z != 0.0f) { result = result + rad; has_result = true; }
            go = go && (int)bounce + 1 < j.max_depth;                               // MAX_DEPTH reached: the partial sum stands (main.cpp:43,78)
            if (RR) go = go && russian_roulette(atten, key, bounce);                // opt-in, separate instantiation: the default kernel is unchanged
            if (go) st = pack_state(sl, (int)bounce + 1, has_result);
            else if (has_result) film_add(j.film, pix, result, &c->clamped, j.npix);                     // px += rayColour(...)  main.cpp:123
        }
        const unsigned int m_go = __ballot_sync(0xffffffffu, go);
        if (lane == 0) w_go[warp] = (unsigned int)__popc(m_go);
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned int t_go = 0;
#pragma unroll
            for (int w = 0; w < SHADE_BLOCK / 32; ++w) { unsigned int g = w_go[w]; w_go[w] = t_go; t_go += g; }
            s_pos = t_go ? atomicAdd(&c->n_front[o], t_go) : 0u;
        }
        __syncthreads();
        if (go) {
            const unsigned int e2 = s_pos + w_go[warp] + (unsigned int)__popc(m_go & ((1u << lane) - 1u));
            if (HRT_IN_BOUNDS(e2, j.capacity, BS_SHADE_APPEND)) {
                j.rays[o][2 * (size_t)e2] = make_float4(r.o.x, r.o.y, r.o.z, __int_as_float(pix));
                j.rays[o][2 * (size_t)e2 + 1] = make_float4(r.d.x, r.d.y, r.d.z, __uint_as_float(st));
                j.thr[o][e2] = make_float4(atten.x, atten.y, atten.z, 0.0f);
                if (has_result) j.res[o][e2] = make_float4(result.x, result.y, result.z, 0.0f);
            }
        }
        __syncthreads();                                                            // the next pass of the loop rewrites the shared counters
    }
}

// ---- secondary rays of queue q over the 8-wide trees (scenes with DScene::wide) --------------------------------------------------
// Persistent warps, per-lane refill from the queue's hand-out counter, phases (hrt_wide.cuh); records go to hits[entry], rays
// that need pass 2 to WaveJob::pend (k_wave3_resolve).  Runs before k_wave2_trace, which is told to skip these entries.
#ifndef HRT_WSEC_MIN_BLOCKS
#define HRT_WSEC_MIN_BLOCKS 5      // teapot_scene 1501 (4 blocks) / 1611 (5) / 1504 (6) / 1292 (8) Mrays/s (r2f)
#endif
template <bool SLOW>
__global__ void __launch_bounds__(BLOCK, HRT_WSEC_MIN_BLOCKS) k_wide_secondary(DScene s, WaveJob j, int q)
{
    __shared__ int stack_mem[WIDE_STACK * BLOCK];
    WaveCtl* c = j.ctl;
    const unsigned int n_sec = c->n_front[q];
    if (n_sec == 0) return;
    WideSmemStack<BLOCK> st; st.base = stack_mem + threadIdx.x; st.sp = 0;
    const unsigned int lane = threadIdx.x & 31u, below = (1u << lane) - 1u;
    const float4* __restrict__ rays = j.rays[q];
    PendList pend; pend.idx = j.pend; pend.count = &c->pend_count; pend.cap = j.capacity;
    WideLane L; L.live = false;
    bool have = false, exhausted = false;
    unsigned int e_idx = 0, r_st = 0; int r_pix = 0;
    for (;;) {
        const int walking = __popc(__ballot_sync(0xffffffffu, L.live));
        if (walking < (exhausted ? 1 : s.wide_refill)) {
            if (have && !L.live) {
                const Hit h = wide_result(L);
                if (HRT_IN_BOUNDS(e_idx, j.capacity, BS_HITS)) j.hits[e_idx] = h;
                if (h.flags & HIT_PENDING) pend_push(pend, e_idx);
                have = false;
            }
            if (!exhausted) {
                const unsigned int m = __ballot_sync(0xffffffffu, !have);
                unsigned int base = 0;
                if (lane == (unsigned int)(__ffs(m) - 1)) base = atomicAdd(&c->sec_next[q], (unsigned int)__popc(m));
                base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
                const unsigned int my = base + (unsigned int)__popc(m & below);
                if (!have && my < n_sec) {
                    const float4 a = rays[2 * (size_t)my], b = rays[2 * (size_t)my + 1];
                    Ray r; r.o = v3(a.x, a.y, a.z); r.d = v3(b.x, b.y, b.z);
                    r_pix = __float_as_int(a.w); r_st = __float_as_uint(b.w); e_idx = my; have = true;
                    Hit early; st.sp = 0;
                    if (!wide_begin(L, s, r, j.t_min, hrt_inf(), early)) { L.b.t = early.t; L.b.prim = early.prim; L.b.obj = early.obj; }
                }
                if (base + (unsigned int)__popc(m) >= n_sec) exhausted = true;
            }
        }
        const RngKey key = make_key(j.seed, (uint32_t)r_pix, (uint32_t)(j.s_first + (int)((r_st >> 8) & 0x7fffffu)), j.rng_mode);
        if (!wide_round<WideSmemStack<BLOCK>, SLOW>(L, s, st, key, r_st & 0xffu) && exhausted) {
            if (!__any_sync(0xffffffffu, have)) break;
        }
    }
}

// pass 2 of the entries k_wide_secondary left pending (ties, badly conditioned triangles, NaNs: hrt_trace.cuh), before the shade pass
__global__ void __launch_bounds__(BLOCK) k_wave3_resolve(DScene s, WaveJob j, int q)
{
    WaveCtl* c = j.ctl;
    const unsigned int n = c->pend_count;
    const float4* __restrict__ rays = j.rays[q];
    const unsigned int nw = gridDim.x * (blockDim.x >> 5), wid = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31u;
    for (unsigned int k = wid + nw * lane; k < n; k += nw * 32u) {                    // one warp per entry while the list is short (k_resolve_batch)
        const unsigned int e = j.pend[k];
        const float4 a = rays[2 * (size_t)e], b = rays[2 * (size_t)e + 1];
        Ray r; r.o = v3(a.x, a.y, a.z); r.d = v3(b.x, b.y, b.z);
        const unsigned int stt = __float_as_uint(b.w);
        const RngKey key = make_key(j.seed, (uint32_t)__float_as_int(a.w), (uint32_t)(j.s_first + (int)((stt >> 8) & 0x7fffffu)), j.rng_mode);
        const Hit h = wide_resolve(s, r, j.t_min, hrt_inf(), key, stt & 0xffu, j.hits[e]);
        if (h.flags & (HIT_RESOLVED | HIT_FALLBACK)) atomicAdd((h.flags & HIT_FALLBACK) ? &c->fallbacks : &c->resolved, 1ull);
        j.hits[e] = h;
    }
}

// ---- megakernel ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(BLOCK) k_render_mega(DScene s, WaveJob j, DCamera cam)
{
    const unsigned long long n = (unsigned long long)j.npix * (unsigned long long)j.s_count;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    unsigned long long rays = 0;
    for (unsigned long long p = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += stride) {
        int pix, sl; path_pixel(j, p, pix, sl);
        RngKey key = make_key(j.seed, (uint32_t)pix, (uint32_t)(j.s_first + sl), j.rng_mode);
        Ray r = primary_ray(cam, j.width, j.height, pix, key);
        V3 atten = v3(1.0f), result = v3(0.0f); bool has_result = false;
        for (int b = 0; b < j.max_depth; ++b) {
            Hit h = s.fast ? trace_fast<false>(s, r, j.t_min, hrt_inf(), key, (uint32_t)b, nullptr) : trace_ref(s, r, j.t_min, hrt_inf(), key, (uint32_t)b);
            ++rays;
            V3 rad; bool go = shade_segment(s, h, r, atten, rad, key, (uint32_t)b);
            if (rad.x != 0.0f || rad.y != 0.0f || rad.z != 0.0f) { result = result + rad; has_result = true; }
            if (go && j.rr && b + 1 < j.max_depth) go = russian_roulette(atten, key, (uint32_t)b);
            if (!go) break;
        }
        if (has_result) film_add(j.film, pix, result, &j.ctl->clamped, j.npix);
    }
    if (rays) atomicAdd(&j.ctl->rays, rays);
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
void launch_trace_batch(const DScene& s, const float4* d_rays, long long n, Hit* d_hits, uint64_t seed, unsigned long long* d_counter,
                        unsigned int* d_pend_idx, unsigned int* d_pend_count, int sm_count, cudaStream_t st, uint64_t* launches)
{
    cudaMemsetAsync(d_counter, 0, sizeof(unsigned long long), st);
    cudaMemsetAsync(d_pend_count, 0, sizeof(unsigned int), st);
    PendList pend; pend.idx = d_pend_idx; pend.count = d_pend_count; pend.cap = (unsigned int)(n < 0xffffffffll ? n : 0xffffffffll);
    if (!s.wide || getenv("HRT_NO_WIDE")) { k_trace_batch_plain<<<grid_for(n, sm_count, 16), BLOCK, 0, st>>>(s, d_rays, n, d_hits, seed); *launches += 1; return; }
    if (s.wide_has_slow) k_trace_batch<true><<<grid_for(n, sm_count, HRT_BATCH_MIN_BLOCKS), BLOCK, 0, st>>>(s, d_rays, n, d_hits, seed, d_counter, pend);
    else k_trace_batch<false><<<grid_for(n, sm_count, HRT_BATCH_MIN_BLOCKS), BLOCK, 0, st>>>(s, d_rays, n, d_hits, seed, d_counter, pend);
    k_resolve_batch<<<sm_count * 8, BLOCK, 0, st>>>(s, d_rays, d_hits, seed, pend); *launches += 2;
}
// ---- hrt_trace_batch_static: the ray batch without any allocation -------------------------------------------------------------------
// The scratch of k_trace_batch / k_resolve_batch / k_split_hits (hit records, hand-out counter, pending list) lives in one file-scope
// device array instead of cudaMalloc memory, for callers that must not allocate (CUDA graph capture, serving loops).  Every GPU that
// loads the library holds the array until the process exits; all calls share it (calls that overlap on one GPU race).
//   [0, 256)                      control words: hand-out counter at byte 0, pending count at byte 64 (both zeroed by every call)
//   [256, 256 + 16 n_max)         hit records (16 B per ray; every record of a call is written before it is read)
//   [256 + 16 n_max, + 4 n_max)   pending list
static __device__ __align__(256) unsigned char g_static_scratch[256 + (size_t)HRT_STATIC_BATCH_MAX * (sizeof(Hit) + sizeof(unsigned int))];
// In host code `g_static_scratch` is only a handle for the runtime, not a device address: the address is looked up once per device,
// with that device current, and cached here.
static unsigned char* g_static_scratch_ptr[HRT_STATIC_MAX_DEVICES];
static std::mutex g_static_scratch_mu;

cudaError_t launch_trace_batch_static(const DScene& s, int device, const float4* d_rays, long long n, uint32_t* d_prim, float* d_t, uint64_t seed,
                                      int sm_count, cudaStream_t st, uint64_t* launches)
{
    if (device < 0 || device >= HRT_STATIC_MAX_DEVICES || n < 0 || n > HRT_STATIC_BATCH_MAX) return cudaErrorInvalidValue;
    unsigned char* base;
    {
        std::lock_guard<std::mutex> lk(g_static_scratch_mu);
        base = g_static_scratch_ptr[device];
        if (!base) {                                                              // set-up, once per device (the caller made `device` current)
            void* p = nullptr;
            cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
            if (e != cudaSuccess) return e;
            if (!p) return cudaErrorInvalidSymbol;
            base = g_static_scratch_ptr[device] = (unsigned char*)p;
        }
    }
    unsigned long long* counter = (unsigned long long*)base;
    unsigned int* pend_count = (unsigned int*)(base + 64);
    Hit* hits = (Hit*)(base + 256);
    unsigned int* pend_idx = (unsigned int*)(base + 256 + (size_t)HRT_STATIC_BATCH_MAX * sizeof(Hit));
    launch_trace_batch(s, d_rays, n, hits, seed, counter, pend_idx, pend_count, sm_count, st, launches);
    launch_split_hits(hits, n, d_prim, d_t, sm_count, st, launches);
    return cudaGetLastError();
}
void launch_camera_rays(const DCamera& cam, const float* d_st, long long n, float* d_out, int sm_count, cudaStream_t st)
{
    k_camera_rays<<<grid_for(n, sm_count, 16), BLOCK, 0, st>>>(cam, d_st, n, d_out);
}
void launch_texture_eval(const DScene& s, int tex, const float* d_in, long long n, float* d_out, int sm_count, cudaStream_t st, uint64_t* launches)
{
    k_texture_eval<<<grid_for(n, sm_count, 16), BLOCK, 0, st>>>(s, tex, d_in, n, d_out); *launches += 1;
}
void launch_scatter_batch(const DScene& s, int material, const float* d_in, const uint8_t* d_front, long long n, uint64_t seed, int rng_mode, float* d_out,
                          int sm_count, cudaStream_t st, uint64_t* launches)
{
    k_scatter_batch<<<grid_for(n, sm_count, 16), BLOCK, 0, st>>>(s, material, d_in, d_front, n, seed, rng_mode, d_out); *launches += 1;
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

#if defined(HRT_WITH_V1_PIPELINE)
void launch_wave_init(const WaveJob& j, const DCamera& cam, int sm_count, cudaStream_t st, uint64_t* launches)
{
    k_wave_init<<<grid_for((long long)j.capacity, sm_count, 16), BLOCK, 0, st>>>(j, cam);
    k_wave_publish<<<1, 32, 0, st>>>(j);
    *launches += 2;
}
void launch_wave_trace(const DScene& s, const WaveJob& j, int it, int sm_count, cudaStream_t st, uint64_t* launches)
{
    k_wave_trace<<<grid_for((long long)j.capacity, sm_count, 16), BLOCK, 0, st>>>(s, j, it & 1); *launches += 1;
}
void launch_wave_shade(const DScene& s, const WaveJob& j, const DCamera& cam, int it, int sm_count, cudaStream_t st, uint64_t* launches)
{
    {
    long long want = ((long long)j.capacity + SHADE_BLOCK - 1) / SHADE_BLOCK, cap = (long long)sm_count * 8;
    k_wave_shade<<<(int)(want < cap ? want : cap), SHADE_BLOCK, 0, st>>>(s, j, cam, it & 1); *launches += 1;
    }
}
#endif
int wave2_trace_blocks(unsigned int capacity, int sm_count, int minb)
{
    long long cap = (long long)capacity / (4ll * CLAIM * (BLOCK / 32));            // keeps the append margin below 1/4 of the queue
    long long want = (long long)sm_count * minb;                                    // one resident wave: the warps loop until the work is gone
    if (cap < 1) cap = 1;
    return (int)(want < cap ? want : cap);
}
unsigned long long wave2_total_claims(const WaveJob& j) { return (unsigned long long)wave2_units(j) * (unsigned long long)wave2_sblocks(j); }
void launch_wave2_trace(const DScene& s, const WaveJob& j, const DCamera& cam, int it, bool secondary_heavy, int sm_count, cudaStream_t st, uint64_t* launches)
{
    // scenes without the 8-wide trees (reference-order scenes, unbounded world-list entries, HRT_NO_WIDE): everything in the lock-step kernel.
    // (Round 1's per-lane-refill pass over the two-child trees - a second hand-kept statement of world_walk - is gone: the wide walk of
    // launch_wave3_trace took its place, profiles/README.md r2b.)
    const int refill = 0;
    const bool lens = cam.lens_radius > 0.0f;             // HRT_RENDER_DEFOCUS: the instantiation with the lens sampling compiled in
    if (secondary_heavy) {
        if (lens) k_wave2_trace<8, true, false><<<wave2_trace_blocks(j.capacity, sm_count, 8), BLOCK, 0, st>>>(s, j, cam, it & 1, refill);
        else k_wave2_trace<8, false, false><<<wave2_trace_blocks(j.capacity, sm_count, 8), BLOCK, 0, st>>>(s, j, cam, it & 1, refill);
    } else {
        if (lens) k_wave2_trace<4, true, false><<<wave2_trace_blocks(j.capacity, sm_count, 4), BLOCK, 0, st>>>(s, j, cam, it & 1, refill);
        else k_wave2_trace<4, false, false><<<wave2_trace_blocks(j.capacity, sm_count, 4), BLOCK, 0, st>>>(s, j, cam, it & 1, refill);
    }
    *launches += 1;
}
#ifndef HRT_WTRACE_MIN_BLOCKS
#define HRT_WTRACE_MIN_BLOCKS 4
#endif
void launch_wave3_trace(const DScene& s, const WaveJob& j, const DCamera& cam, int it, bool secondary_heavy, int sm_count, cudaStream_t st, uint64_t* launches)
{
    // Camera paths: the lock-step fused-primary kernel (most of them are settled by the scene-bounds test and the rest is coherent); on
    // scenes with large meshes its walks go through the wide trees in phases (WIDE), elsewhere through the two-child trees.
    // Secondary rays: the phase kernel with per-lane refill where walks are long and uneven - scenes with meshes whose rays are mostly
    // secondary (teapot_scene 1253 -> 1611 Mrays/s) or whose meshes are large (big_mesh); elsewhere k_wave2_trace's own lock-step
    // pass is as good or better (B200 A/B r2f: shiny_teapot 8874 vs 8658, scattered_balls 5300 vs 3276, cornell_smoke 4070 vs 3626)
    const bool force_wsec = getenv("HRT_FORCE_WIDE_SECONDARY") != nullptr, force_wtrace = getenv("HRT_FORCE_WIDE_PRIMARY") != nullptr;   // tests, sanitizer runs
    const int wsec = ((s.refill_secondary && (secondary_heavy || s.big_mesh)) || force_wsec) ? 1 : 0;
    if (wsec) {
        if (s.wide_has_slow) k_wide_secondary<true><<<sm_count * HRT_WSEC_MIN_BLOCKS, BLOCK, 0, st>>>(s, j, it & 1);
        else k_wide_secondary<false><<<sm_count * HRT_WSEC_MIN_BLOCKS, BLOCK, 0, st>>>(s, j, it & 1);
        *launches += 1;
    }
    const bool lens = cam.lens_radius > 0.0f;
    const int q = it & 1;
    static const int tb = getenv("HRT_TRACE_BUILD") ? atoi(getenv("HRT_TRACE_BUILD")) : 0;      // A/B: 4 = 128-register build, 8 = 64-register build
    if (s.big_mesh || force_wtrace) {
        const int blocks = wave2_trace_blocks(j.capacity, sm_count, HRT_WTRACE_MIN_BLOCKS);
        if (lens) k_wave2_trace<HRT_WTRACE_MIN_BLOCKS, true, true><<<blocks, BLOCK, 0, st>>>(s, j, cam, q, wsec);
        else if (s.wide_has_slow) k_wave2_trace<HRT_WTRACE_MIN_BLOCKS, false, true, true><<<blocks, BLOCK, 0, st>>>(s, j, cam, q, wsec);
        else k_wave2_trace<HRT_WTRACE_MIN_BLOCKS, false, true, false><<<blocks, BLOCK, 0, st>>>(s, j, cam, q, wsec);
    } else if (tb ? tb == 8 : secondary_heavy) {
        if (lens) k_wave2_trace<8, true, false><<<wave2_trace_blocks(j.capacity, sm_count, 8), BLOCK, 0, st>>>(s, j, cam, q, wsec);
        else k_wave2_trace<8, false, false><<<wave2_trace_blocks(j.capacity, sm_count, 8), BLOCK, 0, st>>>(s, j, cam, q, wsec);
    } else {
        // (r3n A/B on shiny_teapot: 3 blocks of 160 registers, no spills: 9351 Mrays/s; 4 of 128, 244 B spilled: 9826; 5 of 96, 628 B spilled: 9407)
        if (lens) k_wave2_trace<4, true, false><<<wave2_trace_blocks(j.capacity, sm_count, 4), BLOCK, 0, st>>>(s, j, cam, q, wsec);
        else k_wave2_trace<4, false, false><<<wave2_trace_blocks(j.capacity, sm_count, 4), BLOCK, 0, st>>>(s, j, cam, q, wsec);
    }
    k_wave3_resolve<<<sm_count * 8, BLOCK, 0, st>>>(s, j, q);
    *launches += 2;
}
void launch_wave2_shade(const DScene& s, const WaveJob& j, int it, int sm_count, cudaStream_t st, uint64_t* launches)
{
    long long want = ((long long)j.capacity + SHADE_BLOCK - 1) / SHADE_BLOCK, cap = (long long)sm_count * 8;
    if (j.rr) k_wave2_shade<true><<<(int)(want < cap ? want : cap), SHADE_BLOCK, 0, st>>>(s, j, it & 1);
    else k_wave2_shade<false><<<(int)(want < cap ? want : cap), SHADE_BLOCK, 0, st>>>(s, j, it & 1);
    *launches += 1;
}
void launch_render_mega(const DScene& s, const WaveJob& j, const DCamera& cam, int sm_count, cudaStream_t st, uint64_t* launches)
{
    k_render_mega<<<grid_for((long long)j.npix * j.s_count, sm_count, 16), BLOCK, 0, st>>>(s, j, cam); *launches += 1;
}
void launch_film_resolve_peers(long long* d_film, const long long* const* peers, int n_peers, int npix, int spp, int write_back,
                                float* d_sum, float* d_mean, uint8_t* d_rgb8, int sm_count, cudaStream_t st, uint64_t* launches)
{
    PeerFilms pf; pf.n = n_peers;
    for (int k = 0; k < 15; ++k) pf.p[k] = k < n_peers ? peers[k] : nullptr;
    k_film_resolve_peers<<<grid_for(npix, sm_count, 16), BLOCK, 0, st>>>(d_film, pf, npix, spp, write_back, d_sum, d_mean, d_rgb8); *launches += 1;
}
void launch_film_resolve_slice(const long long* const* films, int n_films, int pix_first, int pix_count, int spp,
                               float* d_sum, float* d_mean, uint8_t* d_rgb8, int sm_count, cudaStream_t st, uint64_t* launches)
{
    FilmSet fs; fs.n = n_films;
    for (int k = 0; k < 16; ++k) fs.p[k] = k < n_films ? films[k] : nullptr;
    k_film_resolve_slice<<<grid_for(pix_count, sm_count, 16), BLOCK, 0, st>>>(fs, pix_first, pix_count, spp, d_sum, d_mean, d_rgb8); *launches += 1;
}
void launch_film_resolve(const long long* d_film, int npix, int spp, float* d_sum, float* d_mean, uint8_t* d_rgb8, int sm_count, cudaStream_t st, uint64_t* launches)
{
    k_film_resolve<<<grid_for(npix, sm_count, 16), BLOCK, 0, st>>>(d_film, npix, spp, d_sum, d_mean, d_rgb8); *launches += 1;
}

} // namespace hrt

#if defined(HRT_BOUNDS_CHECK)
// bounds-checked variant only: runs the self test, copies the per-site violation counters out and clears them
extern "C" int hrt_debug_bounds(unsigned int* out8)
{
    hrt::k_bounds_selftest<<<1, 1>>>();
    if (cudaDeviceSynchronize() != cudaSuccess) return -3;
    if (cudaMemcpyFromSymbol(out8, hrt::g_bounds, 8 * sizeof(unsigned int)) != cudaSuccess) return -3;
    unsigned int zero[8] = {0};
    if (cudaMemcpyToSymbol(hrt::g_bounds, zero, sizeof(zero)) != cudaSuccess) return -3;
    return 0;
}
#endif
