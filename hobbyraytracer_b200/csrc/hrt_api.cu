// hrt_api.cu — the C ABI of include/hrt.h on top of the CUDA kernels.  This is the drop-in boundary: the
// reference's main() would call hrt_render() where it calls render(NUM_THREADS, background, world, camera, film)
// (main.cpp:176), after replaying its loader's constructor calls (scene.cpp:127-374) through hrt_scene_add_*.
// CUDA only: there is no CPU path in this library; without a device every compute entry point returns HRT_ERR_CUDA.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <algorithm>
#include <atomic>
#include <mutex>
#include <unordered_map>
#include <string>
#include <thread>
#include <utility>
#include <vector>

#include "hrt_shade.cuh"
#include "hrt_bvh_gpu.h"
#include "hrt_kernels.h"
#include "hrt_upload.h"

using namespace hrt;

namespace {
thread_local std::string g_err;
int g_device = 0;
std::atomic<bool> g_last_secondary_heavy{false};      // kernel choice of the last render of the process: where a new scene starts (render_into)

#define CK(call)                                                                                          \
    do {                                                                                                  \
        cudaError_t e_ = (call);                                                                          \
        if (e_ != cudaSuccess) {                                                                          \
            g_err = std::string(#call) + ": " + cudaGetErrorString(e_);                                   \
            return HRT_ERR_CUDA;                                                                          \
        }                                                                                                 \
    } while (0)

// Scene arrays of a destroyed scene are kept (by exact size, bounded) for the next commit of this process: a host that rebuilds
// its scene every frame uploads into the same blocks instead of cudaMalloc / cudaFree of every array (hrt_trim frees them).
struct BlockCache {
    struct Block { int device; size_t bytes; void* p; };
    std::mutex m; std::vector<Block> blocks; size_t total = 0;
    static constexpr size_t MAX_BLOCKS = 1024;
    size_t MAX_TOTAL = 0;                                   // a quarter of the device's memory, at most 32 GiB (C5: 2.9 GB of scene + 1 GB of builder scratch per scene in flight)
    size_t limit()
    {
        if (!MAX_TOTAL) { size_t fr = 0, tot = 0; MAX_TOTAL = (cudaMemGetInfo(&fr, &tot) == cudaSuccess && tot) ? std::min(tot / 4, size_t(32) << 30) : (size_t(1) << 30); }
        return MAX_TOTAL;
    }
    void* take(int device, size_t bytes)
    {
        std::lock_guard<std::mutex> lock(m);
        for (size_t i = 0; i < blocks.size(); ++i)
            if (blocks[i].device == device && blocks[i].bytes == bytes) { void* p = blocks[i].p; total -= bytes; blocks[i] = blocks.back(); blocks.pop_back(); return p; }
        return nullptr;
    }
    void give(int device, size_t bytes, void* p)
    {
        {
            std::lock_guard<std::mutex> lock(m);
            if (blocks.size() < MAX_BLOCKS && total + bytes <= limit()) { blocks.push_back(Block{device, bytes, p}); total += bytes; return; }
        }
        cudaFree(p);
    }
    void trim()
    {
        std::lock_guard<std::mutex> lock(m);
        int cur = 0; cudaGetDevice(&cur);
        for (const Block& b : blocks) { cudaSetDevice(b.device); cudaFree(b.p); }
        cudaSetDevice(cur); blocks.clear(); total = 0;
    }
};
BlockCache g_blocks;

} // namespace
namespace hrt {
cudaError_t pool_alloc(int device, size_t bytes, void** p)
{
    *p = g_blocks.take(device, bytes);
    return *p ? cudaSuccess : cudaMalloc(p, bytes);
}
void pool_give(int device, size_t bytes, void* p) { if (p) g_blocks.give(device, bytes, p); }
} // namespace hrt
namespace {

struct DevBuf {                       // scoped device scratch: freed on every exit path
    void* p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t n) { return cudaMalloc(&p, n ? n : 1); }
};

struct DevUp {
    std::vector<void*>& allocs; std::unordered_map<void*, size_t>& sizes; int device;
    bool failed = false; std::string msg; size_t bytes = 0;
    DevUp(std::vector<void*>& a, std::unordered_map<void*, size_t>& z, int dev) : allocs(a), sizes(z), device(dev) {}
    const void* put(const void* p, size_t n)
    {
        if (!n || failed) return nullptr;
        void* d = g_blocks.take(device, n);
        cudaError_t e = d ? cudaSuccess : cudaMalloc(&d, n);
        if (e == cudaSuccess) e = cudaMemcpy(d, p, n, cudaMemcpyHostToDevice);
        if (e != cudaSuccess) { failed = true; msg = cudaGetErrorString(e); if (d) cudaFree(d); return nullptr; }
        allocs.push_back(d); sizes[d] = n; bytes += n;
        return d;
    }
    const void* put_copy(const void* p, size_t n) { return put(p, n); }
    const std::vector<DMesh>* dev_override = nullptr;           // replicas of hrt_render_multi: the device-built meshes copied to their device
    DMesh dev_mesh(size_t i, const HostShape& sh) { return (dev_override && i < dev_override->size()) ? (*dev_override)[i] : sh.dev; }
};
} // namespace

struct hrt_scene {
    std::shared_ptr<HostScene> host_owner = std::make_shared<HostScene>();   // shared with the per-device replicas of hrt_render_multi
    HostScene& host = *host_owner;
    hrt_scene() = default;
    explicit hrt_scene(const std::shared_ptr<HostScene>& h) : host_owner(h) {}
    hrt_scene(const hrt_scene&) = delete;
    DScene d{};
    int device = 0, sm_count = 148;
    cudaStream_t stream = nullptr;
    cudaStream_t tb_copy_in = nullptr, tb_copy_out = nullptr;   // hrt_trace_batch on large host batches: copies overlap the walks
    std::vector<cudaEvent_t> tb_events;
    std::vector<void*> scene_allocs;
    std::unordered_map<void*, size_t> scene_alloc_bytes;        // sizes of the scene's pool blocks (uploads and device-built arrays)
    size_t scene_bytes = 0, h2d_bytes = 0; double ms_build_upload = 0;
    // persistent-wavefront working set
    WaveJob job{};
    unsigned int job_capacity = 0;
    long long* d_film = nullptr; size_t film_pixels = 0;       // film of hrt_render() + 8-bit image + float sums behind it
    float* d_out = nullptr; size_t out_pixels = 0;
    WaveCtl* h_ctl = nullptr;                                   // pinned snapshots of the control block (ring of CTL_RING)
    std::vector<cudaEvent_t> events;
    // ray-batch working set
    float4* tb_rays = nullptr; Hit* tb_hits = nullptr;
    uint32_t* tb_prim = nullptr; float* tb_t = nullptr; long long tb_capacity = 0;
    unsigned int* tb_pend = nullptr;                            // tb_capacity indices + 1 count: rays left to pass 2
    unsigned long long* d_counter = nullptr;                    // ray hand-out counter of k_trace_batch
    hrt_stats stats{};
    bool last_secondary_heavy = false, have_secondary_prior = false;   // kernel choice of the previous render of this scene (render_into)
    double ms_build_gpu = 0; int bvh_fallbacks = 0;             // HRT_BVH_LBVH_GPU: upload + device build time, meshes handed back to the host builder
    std::vector<DMesh> meshes_view;                             // host copy of DScene::meshes (hrt_scene_mesh_array)
    std::vector<std::pair<cudaIpcMemHandle_t, void*>> peers;    // peer films opened through CUDA IPC (closed with the scene)
    std::vector<hrt_scene*> replicas;                           // hrt_render_multi: this scene on the other devices (same host scene, own uploads)
    std::vector<DMesh> dev_meshes;                              // device-built meshes (HRT_BVH_LBVH_GPU) as they live on THIS device, by mesh index
    bool is_replica = false;
};

namespace {

constexpr int CTL_RING = 4;

// Work buffers of a destroyed scene are parked (one set per process) instead of freed: a host that builds a scene per frame
// (bench.py's end-to-end step, any animation loop) otherwise pays cudaMalloc + cudaFree of ~670 MB of queues and film per
// frame.  The contents are scratch: every render writes a buffer before it reads it.  hrt_trim() releases them.
struct Parked {
    int device = -1;
    unsigned int job_capacity = 0; WaveJob job{};
    size_t film_pixels = 0; long long* d_film = nullptr; float* d_out = nullptr;
    WaveCtl* h_ctl = nullptr;
};
std::mutex g_park_mutex;
Parked g_parked;

void cuda_free_job(WaveJob& job)
{
    for (int q = 0; q < 2; ++q) { cudaFree(job.rays[q]); cudaFree(job.thr[q]); cudaFree(job.res[q]); }
    cudaFree(job.hits); cudaFree(job.pend); cudaFree(job.ctl);
    job = WaveJob{};
}
void trim_parked_locked()
{
    if (g_parked.device < 0) return;
    int cur = 0; cudaGetDevice(&cur); cudaSetDevice(g_parked.device);
    cuda_free_job(g_parked.job); cudaFree(g_parked.d_film); cudaFree(g_parked.d_out);
    if (g_parked.h_ctl) cudaFreeHost(g_parked.h_ctl);
    g_parked = Parked(); cudaSetDevice(cur);
}

void free_job(hrt_scene* s)
{
    for (int q = 0; q < 2; ++q) { cudaFree(s->job.rays[q]); cudaFree(s->job.thr[q]); cudaFree(s->job.res[q]); }
    cudaFree(s->job.hits); cudaFree(s->job.pend); cudaFree(s->job.ctl);
    s->job = WaveJob{}; s->job_capacity = 0;
}
void free_tb(hrt_scene* s)
{
    cudaFree(s->tb_rays); cudaFree(s->tb_hits); cudaFree(s->tb_prim); cudaFree(s->tb_t); cudaFree(s->tb_pend);
    s->tb_pend = nullptr; s->tb_rays = nullptr; s->tb_hits = nullptr; s->tb_prim = nullptr; s->tb_t = nullptr; s->tb_capacity = 0;
}
void free_scene_device(hrt_scene* s)
{
    for (void* p : s->scene_allocs) {
        auto it = s->scene_alloc_bytes.find(p);
        if (it != s->scene_alloc_bytes.end()) g_blocks.give(s->device, it->second, p); else cudaFree(p);
    }
    s->scene_allocs.clear(); s->scene_alloc_bytes.clear(); s->scene_bytes = 0;
}

int ensure_job(hrt_scene* s, unsigned int capacity)
{
    if (capacity <= s->job_capacity) return 0;
    free_job(s);
    {
        std::lock_guard<std::mutex> lock(g_park_mutex);
        if (g_parked.device == s->device && g_parked.job_capacity >= capacity) {       // a parked set fits: take it
            s->job = g_parked.job; s->job_capacity = g_parked.job_capacity;
            g_parked.job = WaveJob{}; g_parked.job_capacity = 0;
            if (!s->h_ctl) { s->h_ctl = g_parked.h_ctl; g_parked.h_ctl = nullptr; }
        }
    }
    if (s->job_capacity >= capacity) { if (!s->h_ctl) CK(cudaMallocHost(&s->h_ctl, sizeof(WaveCtl) * CTL_RING)); return 0; }
    for (int q = 0; q < 2; ++q) {
        CK(cudaMalloc(&s->job.rays[q], sizeof(float4) * 2 * (size_t)capacity));
        CK(cudaMalloc(&s->job.thr[q], sizeof(float4) * (size_t)capacity));
        CK(cudaMalloc(&s->job.res[q], sizeof(float4) * (size_t)capacity));
    }
    CK(cudaMalloc(&s->job.hits, sizeof(Hit) * (size_t)capacity));
    CK(cudaMalloc(&s->job.pend, sizeof(unsigned int) * (size_t)capacity));
    CK(cudaMalloc(&s->job.ctl, sizeof(WaveCtl)));
    if (!s->h_ctl) CK(cudaMallocHost(&s->h_ctl, sizeof(WaveCtl) * CTL_RING));
    s->job_capacity = capacity;
    return 0;
}
int ensure_film(hrt_scene* s, size_t npix)
{
    if (s->film_pixels >= npix) return 0;
    cudaFree(s->d_film); s->d_film = nullptr; s->film_pixels = 0;
    cudaFree(s->d_out); s->d_out = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_park_mutex);
        if (g_parked.device == s->device && g_parked.film_pixels >= npix) {
            s->d_film = g_parked.d_film; s->d_out = g_parked.d_out; s->film_pixels = g_parked.film_pixels;
            g_parked.d_film = nullptr; g_parked.d_out = nullptr; g_parked.film_pixels = 0;
            return 0;
        }
    }
    CK(cudaMalloc(&s->d_film, sizeof(long long) * FILM_LANES * npix));
    CK(cudaMalloc(&s->d_out, sizeof(float) * npix * 3 + npix * 3));
    s->film_pixels = npix;
    return 0;
}
int ensure_tb(hrt_scene* s, long long n, bool own_rays)
{
    if (!s->d_counter) CK(cudaMalloc(&s->d_counter, sizeof(unsigned long long)));
    if (n > s->tb_capacity) {
        free_tb(s);
        CK(cudaMalloc(&s->tb_rays, sizeof(float4) * 2 * (size_t)n));
        CK(cudaMalloc(&s->tb_hits, sizeof(Hit) * (size_t)n));
        CK(cudaMalloc(&s->tb_prim, sizeof(uint32_t) * (size_t)n));
        CK(cudaMalloc(&s->tb_t, sizeof(float) * (size_t)n));
        CK(cudaMalloc(&s->tb_pend, sizeof(unsigned int) * ((size_t)n + 1)));
        s->tb_capacity = n;
    }
    (void)own_rays;
    return 0;
}

unsigned int wave_capacity()
{
    const char* e = getenv("HRT_WAVE_CAPACITY");
    long long v = e ? atoll(e) : 0;
    if (v < 1024) v = 32ll << 20;            // entries per queue (148 B each): B200 A/B r2o 4 / 8 / 16 / 32 Mi: C5 1879 / 2236 / 2474 / 2611, C1 1603 / 1715 / 1790 / 1837 Mrays/s
    if (v > (1ll << 30)) v = 1ll << 30;
    return (unsigned int)v;
}

int check_render_args(hrt_scene* s, const hrt_camera* cam, const hrt_render_params* p)
{
    if (!s || !cam || !p) { g_err = "NULL argument"; return HRT_ERR_ARG; }
    if (!s->host.committed) { g_err = "scene not committed"; return HRT_ERR_STATE; }
    if (p->width < 2 || p->height < 2 || p->spp < 1) { g_err = "width/height must be >= 2 and spp >= 1"; return HRT_ERR_ARG; }
    if ((long long)p->width * p->height > 0x3fffffffll) { g_err = "film too large"; return HRT_ERR_ARG; }
    if (p->sample_first < 0) { g_err = "negative sample_first"; return HRT_ERR_ARG; }
    // the wavefront packs (has_result, local sample index, bounce) into one word: 1 + 23 + 8 bits (hrt_kernels.cu, pack_state)
    if (p->max_depth > 255) { g_err = "max_depth above 255"; return HRT_ERR_ARG; }
    if ((p->sample_count > 0 ? p->sample_count : p->spp) > 0x7fffff) { g_err = "more than 2^23 - 1 samples in one call (render in slices: sample_first / sample_count)"; return HRT_ERR_ARG; }
    return 0;
}

// renders samples [first, first+count) of every pixel and adds them into the fixed-point film d_film
int render_into(hrt_scene* s, const hrt_camera* cam, const hrt_render_params* p, long long* d_film, cudaStream_t st)
{
    const int npix = p->width * p->height;
    const int first = p->sample_first, count = p->sample_count > 0 ? p->sample_count : p->spp;
    const int max_depth = p->max_depth > 0 ? p->max_depth : 50;
    const bool mega = (p->flags & HRT_RENDER_MEGAKERNEL) != 0;
#if defined(HRT_WITH_V1_PIPELINE)
    const bool v1 = (p->flags & HRT_RENDER_WAVEFRONT_V1) != 0;       // queue-everything wavefront with regeneration (A/B builds only)
#else
    const bool v1 = false;                                           // not compiled into the product: the flag selects the default pipeline
#endif
    const unsigned long long total = (unsigned long long)npix * (unsigned long long)count;
    unsigned long long cap = wave_capacity();
    if (cap > total) cap = (total + 1023ull) / 1024ull * 1024ull;
    if (!v1 && cap < (1ull << 18)) cap = 1ull << 18;                      // the fused-primary trace pass keeps an append margin (hrt_kernels.cu)
    if (mega) cap = 1024;
    int rc = ensure_job(s, (unsigned int)cap); if (rc < 0) return rc;
    WaveJob j = s->job;
    j.width = p->width; j.height = p->height; j.npix = npix; j.max_depth = max_depth;
    const bool tiled = (p->width % 8 == 0) && (p->height % 4 == 0);
    j.tiles_x = tiled ? p->width / 8 : 0; j.tiles_y = tiled ? p->height / 4 : 0;
    j.t_min = p->t_min > 0 ? p->t_min : 0.001f; j.seed = p->seed; j.rng_mode = p->rng_mode; j.rr = (p->flags & HRT_RENDER_RR) ? 1 : 0;
    j.s_first = first; j.s_count = count; j.capacity = (unsigned int)cap; j.film = d_film;
    DCamera dc = make_camera(*cam);
    if ((p->flags & HRT_RENDER_DEFOCUS) && cam->aperture > 0.0f) dc.lens_radius = cam->aperture / 2.0f;   // camera.h:29

    constexpr int CHUNK = 8;                              // iterations between two looks at the control block
    constexpr int EV_PER_CHUNK = 2 * CHUNK + 2;           // per iteration: before trace, before shade; + chunk end + snapshot
    while (s->events.size() < (size_t)(EV_PER_CHUNK * CTL_RING + 2)) { cudaEvent_t e; CK(cudaEventCreate(&e)); s->events.push_back(e); }
    cudaEvent_t ev_begin = s->events[EV_PER_CHUNK * CTL_RING], ev_end = s->events[EV_PER_CHUNK * CTL_RING + 1];

    const bool wide = s->d.wide && !mega && !v1 && !getenv("HRT_NO_WIDE");     // the pipeline over the 8-wide trees (hrt_wide.cuh)
    WaveCtl init{}; init.total_paths = (mega || v1) ? total : wave2_total_claims(j);
    CK(cudaMemcpyAsync(j.ctl, &init, sizeof(init), cudaMemcpyHostToDevice, st));
    CK(cudaEventRecord(ev_begin, st));
    s->stats.paths += total;
    if (mega) {
        launch_render_mega(s->d, j, dc, s->sm_count, st, &s->stats.launches);
    } else {
#if defined(HRT_WITH_V1_PIPELINE)
        if (v1) launch_wave_init(j, dc, s->sm_count, st, &s->stats.launches);
#endif
        int it = 0, issued = 0, retired = 0; bool done = false;
        // Kernel choice (lock-step pass vs per-lane refill for the secondary rays, register build of the camera-path kernel) follows the
        // measured secondary rays per camera path.  It starts from what the previous render of this scene found; a scene that has not
        // been rendered yet starts from the last finding of the process (a host that rebuilds its scene per frame) and looks at its
        // first TWO iterations before it queues more: r2o launch list of teapot_scene - 24 iterations with the wrong kernel at the
        // start of every new scene were 29 % of the kernel time (3.3 ms per pass instead of 0.9).
        const bool have_prior = s->have_secondary_prior;
        bool secondary_heavy = have_prior ? s->last_secondary_heavy : g_last_secondary_heavy.load();
        int chunk_len[CTL_RING] = {0}, chunk_end_it[CTL_RING] = {0};
        bool camera_done = false;
        auto retire = [&](int chunk) -> int {               // waits for the chunk, adds its kernel times, looks at its snapshot
            cudaEvent_t* ev = s->events.data() + (size_t)EV_PER_CHUNK * (size_t)(chunk % CTL_RING);
            CK(cudaEventSynchronize(ev[2 * CHUNK + 1]));
            float ms = 0;
            static const bool dbg_iter = getenv("HRT_DEBUG_ITER") != nullptr;     // per-pass timeline on stderr (tuning)
            for (int k = 0; k < chunk_len[chunk % CTL_RING]; ++k) {
                float ms2 = 0;
                CK(cudaEventElapsedTime(&ms, ev[2 * k], ev[2 * k + 1])); s->stats.ms_trace += ms;
                CK(cudaEventElapsedTime(&ms2, ev[2 * k + 1], ev[2 * k + 2])); s->stats.ms_shade += ms2;
                if (dbg_iter) fprintf(stderr, "[hrt] pass %3d: trace %8.3f ms  shade %7.3f ms\n", chunk_end_it[chunk % CTL_RING] - chunk_len[chunk % CTL_RING] + k, ms, ms2);
            }
            if (dbg_iter) {
                const WaveCtl& d = s->h_ctl[chunk % CTL_RING];
                fprintf(stderr, "[hrt]   after pass %d: claims %llu / %llu, queue entries %u / %u, secondary rays so far %llu, heavy %d\n", chunk_end_it[chunk % CTL_RING] - 1,
                        d.next_path < d.total_paths ? d.next_path : d.total_paths, d.total_paths, d.n_front[0], d.n_front[1], d.rays, (int)secondary_heavy);
            }
            const WaveCtl& c = s->h_ctl[chunk % CTL_RING];
            const int o = chunk_end_it[chunk % CTL_RING] & 1; // queue the last shade pass of the chunk filled
            const unsigned long long handed = c.next_path < c.total_paths ? c.next_path : c.total_paths;
            if (!v1 && c.next_path > 0) {                    // fused-primary: secondary rays per camera path so far (claims are CLAIM paths)
                const double paths = (double)handed * (double)total / (double)(c.total_paths ? c.total_paths : 1);
                secondary_heavy = (double)c.rays > 0.75 * paths;
                s->last_secondary_heavy = secondary_heavy; s->have_secondary_prior = true;
                g_last_secondary_heavy.store(secondary_heavy);
            }
            if (handed == c.total_paths) camera_done = true;     // the tail: shorter chunks, the host looks more often
            if (v1 ? (c.n_front[o] == 0 && handed == c.base[o] && handed == c.total_paths) : (c.n_front[o] == 0 && handed == c.total_paths)) done = true;
            return 0;
        };
        while (!done) {
            if (issued - retired >= CTL_RING - 1 || (issued == 1 && !have_prior)) { rc = retire(retired++); if (rc < 0) return rc; if (done) break; }
            // chunks that have finished meanwhile are looked at right away (no wait): in the tail of a render, where a pass takes 20-100 us, the
            // host then sees the empty queue one chunk late instead of three (r2t timeline: 16 empty passes = 0.3 ms of a 31 ms step)
            while (!done && retired < issued) {
                if (cudaEventQuery(s->events[(size_t)EV_PER_CHUNK * (size_t)(retired % CTL_RING) + 2 * CHUNK + 1]) != cudaSuccess) { (void)cudaGetLastError(); break; }   // not ready (a real failure is sticky and shows at the next call)
                rc = retire(retired++); if (rc < 0) return rc;
            }
            if (done) break;
            cudaEvent_t* ev = s->events.data() + (size_t)EV_PER_CHUNK * (size_t)(issued % CTL_RING);
            const int len = (issued == 0 && !have_prior) ? 2 : (camera_done ? 4 : CHUNK);
            chunk_len[issued % CTL_RING] = len; chunk_end_it[issued % CTL_RING] = it + len;
            for (int k = 0; k < len; ++k, ++it) {
                CK(cudaEventRecord(ev[2 * k], st));
#if defined(HRT_WITH_V1_PIPELINE)
                if (v1) launch_wave_trace(s->d, j, it, s->sm_count, st, &s->stats.launches); else
#endif
                if (wide) launch_wave3_trace(s->d, j, dc, it, secondary_heavy, s->sm_count, st, &s->stats.launches);
                else launch_wave2_trace(s->d, j, dc, it, secondary_heavy, s->sm_count, st, &s->stats.launches);
                CK(cudaEventRecord(ev[2 * k + 1], st));
#if defined(HRT_WITH_V1_PIPELINE)
                if (v1) launch_wave_shade(s->d, j, dc, it, s->sm_count, st, &s->stats.launches); else
#endif
                launch_wave2_shade(s->d, j, it, s->sm_count, st, &s->stats.launches);
            }
            CK(cudaEventRecord(ev[2 * len], st));
            CK(cudaMemcpyAsync(&s->h_ctl[issued % CTL_RING], j.ctl, sizeof(WaveCtl), cudaMemcpyDeviceToHost, st));
            CK(cudaEventRecord(ev[2 * CHUNK + 1], st));
            CK(cudaGetLastError());
            ++issued;
        }
        while (retired < issued) { rc = retire(retired++); if (rc < 0) return rc; }
    }
    CK(cudaEventRecord(ev_end, st));
    WaveCtl fin{};
    CK(cudaMemcpyAsync(&s->h_ctl[0], j.ctl, sizeof(WaveCtl), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    fin = s->h_ctl[0];
    float ms = 0; CK(cudaEventElapsedTime(&ms, ev_begin, ev_end));
    s->stats.ms_total += ms;
    s->stats.rays += fin.rays + ((mega || v1) ? 0ull : total);       // fused-primary: the control block counts the secondary rays only
    s->stats.retraced += fin.resolved; s->stats.prim_tests += fin.fallbacks; s->stats.film_clamped += fin.clamped;
    if (fin.fallbacks && getenv("HRT_DEBUG_FALLBACK"))
        for (unsigned long long k = 0; k < fin.fallbacks && k < 16; ++k) {
            const float* g = fin.dbg[k]; unsigned int st; memcpy(&st, &g[7], 4); int pix; memcpy(&pix, &g[3], 4);
            fprintf(stderr, "fallback ray: o=(%.9g %.9g %.9g) d=(%.9g %.9g %.9g) pixel %d bounce %u\n", g[0], g[1], g[2], g[4], g[5], g[6], pix, st & 0xffu);
        }
    s->stats.ms_other = s->stats.ms_total - s->stats.ms_trace - s->stats.ms_shade;
    return 0;
}

} // namespace

extern "C" {

int hrt_abi_version(void) { return HRT_ABI_VERSION; }
const char* hrt_last_error(void) { return g_err.c_str(); }
int hrt_device_count(void)
{
    int n = 0; cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) { g_err = cudaGetErrorString(e); return HRT_ERR_CUDA; }
    return n;
}
int hrt_set_device(int device)
{
    CK(cudaSetDevice(device));
    g_device = device; return 0;
}

hrt_scene* hrt_scene_create(void)
{
    hrt_scene* s = new (std::nothrow) hrt_scene();
    if (!s) { g_err = "out of memory"; return nullptr; }
    s->device = g_device;
    return s;
}
void hrt_trim(void)
{
    { std::lock_guard<std::mutex> lock(g_park_mutex); trim_parked_locked(); }
    g_blocks.trim();
}
void hrt_scene_destroy(hrt_scene* s)
{
    if (!s) return;
    for (hrt_scene* r : s->replicas) hrt_scene_destroy(r);
    s->replicas.clear();
    if (!s->peers.empty()) { cudaSetDevice(s->device); for (auto& pr : s->peers) cudaIpcCloseMemHandle(pr.second); s->peers.clear(); }
    if (s->host.committed || s->stream) {
        int cur_dev = 0; cudaGetDevice(&cur_dev);
        cudaSetDevice(s->device);
        cudaDeviceSynchronize();                                   // nothing may still be using what gets parked
        {
            std::lock_guard<std::mutex> lock(g_park_mutex);
            if (g_parked.device != s->device) trim_parked_locked();
            g_parked.device = s->device;
            if (s->job_capacity > g_parked.job_capacity) {          // keep the larger set
                cuda_free_job(g_parked.job);
                g_parked.job = s->job; g_parked.job_capacity = s->job_capacity; s->job = WaveJob{}; s->job_capacity = 0;
            }
            if (s->film_pixels > g_parked.film_pixels) {
                cudaFree(g_parked.d_film); cudaFree(g_parked.d_out);
                g_parked.d_film = s->d_film; g_parked.d_out = s->d_out; g_parked.film_pixels = s->film_pixels;
                s->d_film = nullptr; s->d_out = nullptr; s->film_pixels = 0;
            }
            if (s->h_ctl && !g_parked.h_ctl) { g_parked.h_ctl = s->h_ctl; s->h_ctl = nullptr; }
        }
        free_job(s); free_tb(s); free_scene_device(s);
        cudaFree(s->d_film); cudaFree(s->d_out); cudaFree(s->d_counter);
        if (s->h_ctl) cudaFreeHost(s->h_ctl);
        for (cudaEvent_t e : s->events) cudaEventDestroy(e);
        if (s->tb_copy_in) { cudaStreamDestroy(s->tb_copy_in); cudaStreamDestroy(s->tb_copy_out); }
        for (cudaEvent_t e : s->tb_events) cudaEventDestroy(e);
        if (s->stream) cudaStreamDestroy(s->stream);
        if (s->is_replica) cudaSetDevice(cur_dev);
    }
    delete s;
}

#define NEED(s) do { if (!(s)) { g_err = "scene is NULL"; return HRT_ERR_ARG; } if ((s)->host.committed) { g_err = "scene already committed"; return HRT_ERR_STATE; } } while (0)
int hrt_scene_add_texture_solid(hrt_scene* s, const float rgb[3]) { NEED(s); return b_add_texture_solid(s->host, rgb, g_err); }
int hrt_scene_add_texture_checker(hrt_scene* s, int e, int o) { NEED(s); return b_add_texture_checker(s->host, e, o, g_err); }
int hrt_scene_add_texture_image(hrt_scene* s, const uint8_t* p, int w, int h) { NEED(s); return b_add_texture_image(s->host, p, w, h, g_err); }
int hrt_scene_add_texture_env(hrt_scene* s, const float* p, int w, int h, int c) { NEED(s); return b_add_texture_env(s->host, p, w, h, c, g_err); }
int hrt_scene_add_material(hrt_scene* s, const hrt_material* m) { NEED(s); return b_add_material(s->host, m, g_err); }
int hrt_scene_add_sphere(hrt_scene* s, const float c[3], float r, int m) { NEED(s); return b_add_sphere(s->host, c, r, m, g_err); }
int hrt_scene_add_rect(hrt_scene* s, int axis, float a0, float a1, float b0, float b1, float k, int m) { NEED(s); return b_add_rect(s->host, axis, a0, a1, b0, b1, k, m, g_err); }
int hrt_scene_add_box(hrt_scene* s, const float c[3], const float d[3], int m) { NEED(s); return b_add_box(s->host, c, d, m, g_err); }
int hrt_scene_add_mesh(hrt_scene* s, const float* p, const float* n, const float* uv, int64_t nt, int m) { NEED(s); return b_add_mesh(s->host, p, n, uv, nt, m, g_err); }
int hrt_scene_add_mesh_indexed(hrt_scene* s, const float* v, const float* n, const float* uv, int64_t nv, const uint32_t* idx, int64_t ni,
                               const hrt_submesh* sub, int n_sub, int m, uint32_t flags) { NEED(s); return b_add_mesh_indexed(s->host, v, n, uv, nv, idx, ni, sub, n_sub, m, flags, g_err); }
int hrt_scene_add_object(hrt_scene* s, int shape, const hrt_xform* ops, int n) { NEED(s); return b_add_object(s->host, shape, ops, n, g_err); }
int hrt_scene_add_medium(hrt_scene* s, int shape, const hrt_xform* ops, int n, float density, const float rgb[3], int tex) { NEED(s); return b_add_medium(s->host, shape, ops, n, density, rgb, tex, g_err); }
int hrt_scene_set_background(hrt_scene* s, int tex) { NEED(s); return b_set_background(s->host, tex, g_err); }
int64_t hrt_scene_prim_base(const hrt_scene* s, int shape)
{
    if (!s || shape < 0 || (size_t)shape >= s->host.shapes.size()) { g_err = "unknown shape"; return HRT_ERR_ARG; }
    return s->host.shapes[(size_t)shape].prim_base;
}
int64_t hrt_scene_prim_count(const hrt_scene* s) { if (!s) { g_err = "scene is NULL"; return HRT_ERR_ARG; } return s->host.n_prims; }

int hrt_scene_commit(hrt_scene* s, int bvh_mode, uint64_t seed)
{
    NEED(s);
    if (bvh_mode < HRT_BVH_REFERENCE || bvh_mode > HRT_BVH_LBVH_HOST) { g_err = "unknown bvh mode"; return HRT_ERR_ARG; }
    CK(cudaSetDevice(s->device));
    CK(cudaDeviceGetAttribute(&s->sm_count, cudaDevAttrMultiProcessorCount, s->device));   // (cudaGetDeviceProperties costs ~1 ms per commit)
    std::string e = flatten(s->host, bvh_mode, seed);
    if (!e.empty()) { g_err = e; return HRT_ERR_ARG; }
    CK(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    s->ms_build_gpu = 0; s->bvh_fallbacks = 0; s->h2d_bytes = 0; s->ms_build_upload = 0;
    size_t dev_bytes = 0;
    for (int shape : s->host.mesh_shapes) {                       // HRT_BVH_LBVH_GPU: per-mesh arrays are built on the device
        HostShape& sh = s->host.shapes[(size_t)shape];
        if (!sh.gpu_pending) continue;
        GpuMeshBuild b; std::string berr;
        MeshSource src;
        if (sh.indexed) { src.indexed = true; src.pos = sh.V(); src.nrm = sh.N(); src.uv = sh.T(); src.idx = sh.I(); src.n_vert = sh.n_vert; }
        else { src.pos = sh.pos.data(); src.nrm = sh.nrm.data(); src.uv = sh.uv.data(); }
        if (!build_mesh_gpu(src, sh.n_tri, seed, shape, s->stream, s->scene_allocs, s->scene_alloc_bytes, dev_bytes, b, berr, &s->stats.launches)) {
            g_err = "device BVH build failed: " + berr; free_scene_device(s); return HRT_ERR_CUDA;
        }
        sh.gpu_pending = false;
        if (b.fallback) {                                         // (non-finite positions, degenerate Morton order): the host builder's SAH tree
            int nt = (int)std::thread::hardware_concurrency(); if (nt < 1) nt = 1;
            build_mesh_host(sh, HRT_BVH_SAH, seed, shape, nt, s->host.build_ms_ref, s->host.build_ms_fast);
            ++s->bvh_fallbacks;
            continue;
        }
        sh.has_dev = true; sh.dev = b.mesh;
        sh.n_ref_nodes = b.n_ref_nodes; sh.n_fast_nodes = b.n_fast_nodes; sh.n_leaves = b.n_leaves; sh.n_wide_nodes = b.n_wide_nodes; sh.last_tri_r = b.last_tri_r;
        sh.fast.depth = b.depth;
        s->ms_build_gpu += b.ms_upload + b.ms_build; s->ms_build_upload += b.ms_upload; s->h2d_bytes += b.h2d_bytes;
        if (getenv("HRT_DEBUG_BUILD")) fprintf(stderr, "[hrt] device build of shape %d: %lld triangles, source %.1f MB (%s) uploaded in %.1f ms, built in %.1f ms\n",
                                               shape, (long long)sh.n_tri, b.h2d_bytes / 1e6, sh.indexed ? "indexed" : "soup", b.ms_upload, b.ms_build);
    }
    DevUp up(s->scene_allocs, s->scene_alloc_bytes, s->device);
    s->d = make_dscene(s->host, up);
    if (const char* e = getenv("HRT_WIDE_PHASE_MIN")) { int v = atoi(e); if (v >= 1 && v <= 32) s->d.wide_phase_min = v; }          // A/B overrides
    if (const char* e = getenv("HRT_WIDE_PHASE_MIN_OBJ")) { int v = atoi(e); if (v >= 1 && v <= 32) s->d.wide_phase_min_obj = v; }
    if (const char* e = getenv("HRT_WIDE_REFILL")) { int v = atoi(e); if (v >= 1 && v <= 32) s->d.wide_refill = v; }
    if (const char* e = getenv("HRT_REFILL_SECONDARY")) s->d.refill_secondary = (atoi(e) != 0 && s->d.fast) ? 1 : 0;   // A/B override
    if (up.failed) { g_err = "scene upload failed: " + up.msg; free_scene_device(s); return HRT_ERR_CUDA; }
    s->scene_bytes = up.bytes + dev_bytes; s->h2d_bytes += up.bytes;
    s->meshes_view.resize(s->host.mesh_shapes.size());
    if (!s->meshes_view.empty()) CK(cudaMemcpy(s->meshes_view.data(), s->d.meshes, s->meshes_view.size() * sizeof(DMesh), cudaMemcpyDeviceToHost));
    CK(cudaStreamSynchronize(s->stream));                          // (uploads are synchronous copies, device builds run on s->stream: no device-wide
                                                                   //  wait, so a host thread can commit the next scene while another one renders)
    for (HostShape& sh : s->host.shapes) sh.release_borrowed();   // HRT_MESH_BORROW ends here
    s->host.committed = true;
    return 0;
}

int64_t hrt_scene_mesh_array(hrt_scene* s, int mesh_index, int which, void* out, int64_t capacity_bytes)
{
    if (!s || !s->host.committed) { g_err = "scene not committed"; return HRT_ERR_STATE; }
    if (mesh_index < 0 || (size_t)mesh_index >= s->meshes_view.size() || which < 0 || which > 5) { g_err = "bad argument"; return HRT_ERR_ARG; }
    const HostShape& sh = s->host.shapes[(size_t)s->host.mesh_shapes[(size_t)mesh_index]];
    const DMesh& m = s->meshes_view[(size_t)mesh_index];
    const void* src[6] = {m.ref_nodes, m.fast_nodes, m.fast_leaves, m.tris, m.attrs, m.wide_nodes};
    const size_t bytes[6] = {sh.n_ref_nodes * 32, sh.n_fast_nodes * 64, sh.n_leaves * 128, (size_t)sh.n_tri * 48, (size_t)sh.n_tri * 64, m.wide_nodes ? sh.n_wide_nodes * 96 : 0};
    if (out) {
        if (capacity_bytes < (int64_t)bytes[which]) { g_err = "buffer too small"; return HRT_ERR_ARG; }
        CK(cudaSetDevice(s->device));
        if (bytes[which]) CK(cudaMemcpy(out, src[which], bytes[which], cudaMemcpyDeviceToHost));
    }
    return (int64_t)bytes[which];
}

int hrt_trace_batch_device(hrt_scene* s, const void* d_rays, int64_t n, void* d_prim_id, void* d_t, uint64_t rng_seed, void* stream)
{
    if (!s || (n > 0 && (!d_rays || !d_prim_id || !d_t)) || n < 0) { g_err = "bad argument"; return HRT_ERR_ARG; }
    if (!s->host.committed) { g_err = "scene not committed"; return HRT_ERR_STATE; }
    if (n > 0xffffffffll) { g_err = "at most 2^32-1 rays per call"; return HRT_ERR_ARG; }
    if (n == 0) return 0;
    CK(cudaSetDevice(s->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
    if (n > s->tb_capacity) {                      // only hits / retrace list are needed for a device call; allocate the full set for simplicity
        int rc = ensure_tb(s, n, false); if (rc < 0) return rc;
    }
    launch_trace_batch(s->d, (const float4*)d_rays, n, s->tb_hits, rng_seed, s->d_counter, s->tb_pend, s->tb_pend + n, s->sm_count, st, &s->stats.launches);
    launch_split_hits(s->tb_hits, n, (uint32_t*)d_prim_id, (float*)d_t, s->sm_count, st, &s->stats.launches);
    CK(cudaGetLastError());
    s->stats.rays += (uint64_t)n;
    return 0;
}

// hrt_trace_batch_device without any allocation: the scratch is a file-scope device array of hrt_kernels.cu (include/hrt.h).
// Nothing here calls cudaMalloc / cudaFree or an allocating library routine; calls share the array and must not overlap on one GPU.
int hrt_trace_batch_static(hrt_scene* s, const void* d_rays, int64_t n, void* d_prim_id, void* d_t, uint64_t rng_seed, void* stream)
{
    if (!s || (n > 0 && (!d_rays || !d_prim_id || !d_t)) || n < 0) { g_err = "bad argument"; return HRT_ERR_ARG; }
    if (!s->host.committed) { g_err = "scene not committed"; return HRT_ERR_STATE; }
    static_assert(HRT_TRACE_BATCH_STATIC_MAX == HRT_STATIC_BATCH_MAX, "include/hrt.h and hrt_kernels.h disagree on the static scratch size");
    if (n > HRT_TRACE_BATCH_STATIC_MAX) {
        g_err = "hrt_trace_batch_static: " + std::to_string((long long)n) + " rays exceed the static scratch of HRT_TRACE_BATCH_STATIC_MAX = "
              + std::to_string((long long)HRT_TRACE_BATCH_STATIC_MAX) + " rays per call (split the batch or use hrt_trace_batch_device)";
        return HRT_ERR_ARG;
    }
    if (n == 0) return 0;
    CK(cudaSetDevice(s->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
    cudaError_t e = launch_trace_batch_static(s->d, s->device, (const float4*)d_rays, n, (uint32_t*)d_prim_id, (float*)d_t, rng_seed, s->sm_count, st, &s->stats.launches);
    if (e != cudaSuccess) { g_err = std::string("hrt_trace_batch_static: ") + cudaGetErrorString(e); return HRT_ERR_CUDA; }
    s->stats.rays += (uint64_t)n;
    return 0;
}

int hrt_trace_batch(hrt_scene* s, const float* rays, int64_t n, uint32_t* prim_id, float* t, float* normal, float* uv, uint8_t* front, uint64_t rng_seed)
{
    if (!s || (n > 0 && (!rays || !prim_id || !t)) || n < 0) { g_err = "bad argument"; return HRT_ERR_ARG; }
    if (!s->host.committed) { g_err = "scene not committed"; return HRT_ERR_STATE; }
    if (n == 0) return 0;
    CK(cudaSetDevice(s->device));
    int rc = ensure_tb(s, n, true); if (rc < 0) return rc;
    cudaStream_t st = s->stream;
    // Large batches in chunks on three streams: chunk c + 1 crosses PCIe while chunk c is walked and the answers of chunk c - 1 go back
    // (16 Mi rays from pinned memory: 512 MB in, 128 MB out; B200: 23 ms serial).  Only where chunking cannot change an answer: the Philox
    // key of a ray is its index in the LAUNCH, which matters to ConstantMedium::hit alone (scenes with media take the one-launch path).
    constexpr int64_t CHUNK = int64_t(2) << 20;
    bool has_medium = false; for (const HostObject& o : s->host.objects) has_medium |= o.medium;
    if (n >= 2 * CHUNK && !normal && !uv && !front && !has_medium && !getenv("HRT_TRACE_BATCH_ONE_LAUNCH")) {
        if (!s->tb_copy_in) { CK(cudaStreamCreateWithFlags(&s->tb_copy_in, cudaStreamNonBlocking)); CK(cudaStreamCreateWithFlags(&s->tb_copy_out, cudaStreamNonBlocking)); }
        const int n_chunks = (int)((n + CHUNK - 1) / CHUNK);
        while ((int)s->tb_events.size() < 2 * n_chunks) { cudaEvent_t e; CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); s->tb_events.push_back(e); }
        CK(cudaStreamSynchronize(st));                                         // earlier work on the scene's stream may still use the buffers
        for (int c = 0; c < n_chunks; ++c) {
            const int64_t off = (int64_t)c * CHUNK, m = std::min(CHUNK, n - off);
            CK(cudaMemcpyAsync(s->tb_rays + 2 * off, rays + 8 * off, sizeof(float) * 8 * (size_t)m, cudaMemcpyHostToDevice, s->tb_copy_in));
            CK(cudaEventRecord(s->tb_events[2 * c], s->tb_copy_in));
            CK(cudaStreamWaitEvent(st, s->tb_events[2 * c], 0));
            launch_trace_batch(s->d, s->tb_rays + 2 * off, m, s->tb_hits, rng_seed, s->d_counter, s->tb_pend, s->tb_pend + n, s->sm_count, st, &s->stats.launches);
            launch_split_hits(s->tb_hits, m, s->tb_prim + off, s->tb_t + off, s->sm_count, st, &s->stats.launches);
            CK(cudaEventRecord(s->tb_events[2 * c + 1], st));
            CK(cudaStreamWaitEvent(s->tb_copy_out, s->tb_events[2 * c + 1], 0));
            CK(cudaMemcpyAsync(prim_id + off, s->tb_prim + off, sizeof(uint32_t) * (size_t)m, cudaMemcpyDeviceToHost, s->tb_copy_out));
            CK(cudaMemcpyAsync(t + off, s->tb_t + off, sizeof(float) * (size_t)m, cudaMemcpyDeviceToHost, s->tb_copy_out));
        }
        CK(cudaStreamSynchronize(s->tb_copy_out)); CK(cudaStreamSynchronize(st));
        CK(cudaGetLastError());
        s->stats.rays += (uint64_t)n;
        return 0;
    }
    CK(cudaMemcpyAsync(s->tb_rays, rays, sizeof(float) * 8 * (size_t)n, cudaMemcpyHostToDevice, st));
    rc = hrt_trace_batch_device(s, s->tb_rays, n, s->tb_prim, s->tb_t, rng_seed, st); if (rc < 0) return rc;
    CK(cudaMemcpyAsync(prim_id, s->tb_prim, sizeof(uint32_t) * (size_t)n, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(t, s->tb_t, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost, st));
    if (normal || uv || front) {
        DevBuf bn, bu, bf;                                                      // freed on every exit path
        if (normal) CK(bn.alloc(sizeof(float) * 3 * (size_t)n));
        if (uv) CK(bu.alloc(sizeof(float) * 2 * (size_t)n));
        if (front) CK(bf.alloc((size_t)n));
        float* d_n = (float*)bn.p; float* d_uv = (float*)bu.p; uint8_t* d_f = (uint8_t*)bf.p;
        launch_eval_batch(s->d, s->tb_rays, s->tb_hits, n, d_n, d_uv, d_f, s->sm_count, st, &s->stats.launches);
        if (normal) CK(cudaMemcpyAsync(normal, d_n, sizeof(float) * 3 * (size_t)n, cudaMemcpyDeviceToHost, st));
        if (uv) CK(cudaMemcpyAsync(uv, d_uv, sizeof(float) * 2 * (size_t)n, cudaMemcpyDeviceToHost, st));
        if (front) CK(cudaMemcpyAsync(front, d_f, (size_t)n, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
    }
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    return 0;
}

// ---- parity entry for Camera::getRay (camera.h:9-39): needs no scene, runs on the device selected by hrt_set_device ----------------------
int hrt_camera_rays(const hrt_camera* cam, const float* st, int64_t n, float* out)
{
    if (!cam || n < 0 || (n > 0 && (!st || !out))) { g_err = "bad argument"; return HRT_ERR_ARG; }
    if (n == 0) return 0;
    CK(cudaSetDevice(g_device));
    int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, g_device);
    DevBuf in, res;
    CK(in.alloc(sizeof(float) * 2 * (size_t)n)); CK(res.alloc(sizeof(float) * 6 * (size_t)n));
    CK(cudaMemcpy(in.p, st, sizeof(float) * 2 * (size_t)n, cudaMemcpyHostToDevice));
    launch_camera_rays(make_camera(*cam), (const float*)in.p, n, (float*)res.p, sms, nullptr);
    CK(cudaGetLastError());
    CK(cudaMemcpy(out, res.p, sizeof(float) * 6 * (size_t)n, cudaMemcpyDeviceToHost));
    return 0;
}

// ---- parity entries for the shading half of the path: Texture::colourValue and Material::scatter / emitted on batches ----------------
int hrt_texture_eval(hrt_scene* s, int texture, const float* uvp, int64_t n, float* out_rgb)
{
    if (!s || n < 0 || (n > 0 && (!uvp || !out_rgb))) { g_err = "bad argument"; return HRT_ERR_ARG; }
    if (!s->host.committed) { g_err = "scene not committed"; return HRT_ERR_STATE; }
    if (texture < 0 || texture >= s->d.n_textures) { g_err = "unknown texture"; return HRT_ERR_ARG; }
    if (n == 0) return 0;
    CK(cudaSetDevice(s->device));
    DevBuf in, out;
    CK(in.alloc(sizeof(float) * 5 * (size_t)n)); CK(out.alloc(sizeof(float) * 3 * (size_t)n));
    cudaStream_t st = s->stream;
    CK(cudaMemcpyAsync(in.p, uvp, sizeof(float) * 5 * (size_t)n, cudaMemcpyHostToDevice, st));
    launch_texture_eval(s->d, texture, (const float*)in.p, n, (float*)out.p, s->sm_count, st, &s->stats.launches);
    CK(cudaMemcpyAsync(out_rgb, out.p, sizeof(float) * 3 * (size_t)n, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    return 0;
}
int hrt_scatter_batch(hrt_scene* s, int material, const float* in14, const uint8_t* front, int64_t n, uint64_t seed, int rng_mode, float* out10)
{
    if (!s || n < 0 || (n > 0 && (!in14 || !front || !out10))) { g_err = "bad argument"; return HRT_ERR_ARG; }
    if (!s->host.committed) { g_err = "scene not committed"; return HRT_ERR_STATE; }
    if (material < 0 || material >= s->d.n_materials) { g_err = "unknown material"; return HRT_ERR_ARG; }
    if (n == 0) return 0;
    CK(cudaSetDevice(s->device));
    DevBuf in, fr, out;
    CK(in.alloc(sizeof(float) * 14 * (size_t)n)); CK(fr.alloc((size_t)n)); CK(out.alloc(sizeof(float) * 10 * (size_t)n));
    cudaStream_t st = s->stream;
    CK(cudaMemcpyAsync(in.p, in14, sizeof(float) * 14 * (size_t)n, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(fr.p, front, (size_t)n, cudaMemcpyHostToDevice, st));
    launch_scatter_batch(s->d, material, (const float*)in.p, (const uint8_t*)fr.p, n, seed, rng_mode, (float*)out.p, s->sm_count, st, &s->stats.launches);
    CK(cudaMemcpyAsync(out10, out.p, sizeof(float) * 10 * (size_t)n, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    return 0;
}

int64_t hrt_film_bytes(int width, int height)
{
    if (width < 1 || height < 1 || (long long)width * height > 0x3fffffffll) { g_err = "bad film size"; return HRT_ERR_ARG; }
    return (int64_t)width * height * FILM_LANES * (int64_t)sizeof(long long);
}

int hrt_render_accumulate_device(hrt_scene* s, const hrt_camera* cam, const hrt_render_params* p, void* d_film, void* stream)
{
    int rc = check_render_args(s, cam, p); if (rc < 0) return rc;
    if (!d_film) { g_err = "d_film is NULL"; return HRT_ERR_ARG; }
    CK(cudaSetDevice(s->device));
    if (!stream) CK(cudaStreamSynchronize(cudaStreamLegacy));      // NULL = the scene's own (non-blocking) stream: what the caller queued on the default stream first
    return render_into(s, cam, p, (long long*)d_film, stream ? (cudaStream_t)stream : s->stream);
}

int hrt_film_resolve_device(hrt_scene* s, const void* d_film, int width, int height, int spp, void* d_sum, void* d_mean, void* d_rgb8, void* stream)
{
    if (!s || !d_film || width < 1 || height < 1 || spp < 1) { g_err = "bad argument"; return HRT_ERR_ARG; }
    CK(cudaSetDevice(s->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
    if (!stream) CK(cudaStreamSynchronize(cudaStreamLegacy));
    launch_film_resolve((const long long*)d_film, width * height, spp, (float*)d_sum, (float*)d_mean, (uint8_t*)d_rgb8, s->sm_count, st, &s->stats.launches);
    CK(cudaGetLastError());
    return 0;
}

int hrt_render(hrt_scene* s, const hrt_camera* cam, const hrt_render_params* p, float* out_sum, uint8_t* out_rgb8)
{
    int rc = check_render_args(s, cam, p); if (rc < 0) return rc;
    CK(cudaSetDevice(s->device));
    const size_t npix = (size_t)p->width * p->height;
    rc = ensure_film(s, npix); if (rc < 0) return rc;
    cudaStream_t st = s->stream;
    CK(cudaMemsetAsync(s->d_film, 0, sizeof(long long) * FILM_LANES * npix, st));
    rc = render_into(s, cam, p, s->d_film, st); if (rc < 0) return rc;
    uint8_t* d_rgb8 = (uint8_t*)(s->d_out + npix * 3);
    launch_film_resolve(s->d_film, (int)npix, p->spp, out_sum ? s->d_out : nullptr, nullptr, out_rgb8 ? d_rgb8 : nullptr, s->sm_count, st, &s->stats.launches);
    if (out_sum) CK(cudaMemcpyAsync(out_sum, s->d_out, sizeof(float) * npix * 3, cudaMemcpyDeviceToHost, st));
    if (out_rgb8) CK(cudaMemcpyAsync(out_rgb8, d_rgb8, npix * 3, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    return 0;
}

// ---- checkpoints (hrt_render_progressive) ------------------------------------------------------------------------
namespace {
struct CheckpointHeader {
    char     magic[8];            // "HRTFILM2"
    int32_t  width, height, spp, samples_done, max_depth, rng_mode, flags_reserved;
    float    t_min;
    uint64_t seed, camera_hash, prim_count;
    uint64_t scene_hash;          // HostScene::content_hash: objects, wrappers, materials, textures, geometry, background, BVH mode
};
const char kCheckpointMagic[8] = {'H', 'R', 'T', 'F', 'I', 'L', 'M', '2'};
uint64_t fnv1a(const void* p, size_t n) { const unsigned char* b = (const unsigned char*)p; uint64_t h = 1469598103934665603ull; for (size_t i = 0; i < n; ++i) { h ^= b[i]; h *= 1099511628211ull; } return h; }
CheckpointHeader checkpoint_header_of(const hrt_scene* s, const hrt_camera* cam, const hrt_render_params* p)
{
    CheckpointHeader h; memset(&h, 0, sizeof(h));
    memcpy(h.magic, kCheckpointMagic, 8);
    h.width = p->width; h.height = p->height; h.spp = p->spp; h.samples_done = 0;
    h.max_depth = p->max_depth > 0 ? p->max_depth : 50; h.rng_mode = p->rng_mode; h.flags_reserved = p->flags & (HRT_RENDER_DEFOCUS | HRT_RENDER_RR);
    h.t_min = p->t_min > 0 ? p->t_min : 0.001f;
    h.seed = p->seed; h.camera_hash = fnv1a(cam, sizeof(*cam)); h.prim_count = (uint64_t)s->host.n_prims; h.scene_hash = s->host.content_hash;
    return h;
}
bool same_render(const CheckpointHeader& a, const CheckpointHeader& b)
{
    return memcmp(a.magic, b.magic, 8) == 0 && a.width == b.width && a.height == b.height && a.spp == b.spp && a.max_depth == b.max_depth &&
           a.rng_mode == b.rng_mode && a.flags_reserved == b.flags_reserved && memcmp(&a.t_min, &b.t_min, 4) == 0 && a.seed == b.seed && a.camera_hash == b.camera_hash && a.prim_count == b.prim_count && a.scene_hash == b.scene_hash;
}
int read_header(const char* path, CheckpointHeader& h, FILE** keep_open)
{
    FILE* f = fopen(path, "rb");
    if (!f) return 1;                                                     // no such file: a fresh render
    if (fread(&h, sizeof(h), 1, f) != 1 || memcmp(h.magic, kCheckpointMagic, 8) != 0) { fclose(f); g_err = std::string("not a checkpoint file: ") + path; return HRT_ERR_IO; }
    if (keep_open) *keep_open = f; else fclose(f);
    return 0;
}
int write_checkpoint(const char* path, const CheckpointHeader& h, const void* film, size_t bytes)
{
    const std::string tmp = std::string(path) + ".tmp";
    FILE* f = fopen(tmp.c_str(), "wb");
    if (!f) { g_err = "cannot write checkpoint " + tmp; return HRT_ERR_IO; }
    const bool ok = fwrite(&h, sizeof(h), 1, f) == 1 && fwrite(film, 1, bytes, f) == bytes;
    if (fclose(f) != 0 || !ok) { remove(tmp.c_str()); g_err = "short write on checkpoint " + tmp; return HRT_ERR_IO; }
    if (rename(tmp.c_str(), path) != 0) { remove(tmp.c_str()); g_err = std::string("cannot move the checkpoint to ") + path; return HRT_ERR_IO; }
    return 0;
}
} // namespace

int hrt_checkpoint_samples(const char* path, int* width, int* height, int* spp)
{
    if (!path) { g_err = "path is NULL"; return HRT_ERR_ARG; }
    CheckpointHeader h; int rc = read_header(path, h, nullptr);
    if (rc == 1) { g_err = std::string("no such checkpoint: ") + path; return HRT_ERR_IO; }
    if (rc < 0) return rc;
    if (width) *width = h.width; if (height) *height = h.height; if (spp) *spp = h.spp;
    return h.samples_done;
}

int hrt_render_progressive(hrt_scene* s, const hrt_camera* cam, const hrt_render_params* p, const char* checkpoint_path,
                           int slice_spp, hrt_progress_fn on_slice, void* user, float* out_sum, uint8_t* out_rgb8)
{
    int rc = check_render_args(s, cam, p); if (rc < 0) return rc;
    if (p->sample_first != 0 || (p->sample_count > 0 && p->sample_count != p->spp)) { g_err = "a progressive render covers all samples"; return HRT_ERR_ARG; }
    CK(cudaSetDevice(s->device));
    const size_t npix = (size_t)p->width * p->height;
    const size_t film_bytes = sizeof(long long) * FILM_LANES * npix;
    rc = ensure_film(s, npix); if (rc < 0) return rc;
    cudaStream_t st = s->stream;
    CheckpointHeader head = checkpoint_header_of(s, cam, p);
    std::vector<unsigned char> host_film;
    if (checkpoint_path) host_film.resize(film_bytes);
    bool resumed = false;
    if (checkpoint_path) {
        CheckpointHeader on_disk; FILE* f = nullptr;
        rc = read_header(checkpoint_path, on_disk, &f);
        if (rc < 0) return rc;
        if (rc == 0) {
            const bool ok = same_render(head, on_disk) && on_disk.samples_done >= 0 && on_disk.samples_done <= p->spp;
            const bool full = ok && fread(host_film.data(), 1, film_bytes, f) == film_bytes;
            fclose(f);
            if (!ok) { g_err = std::string("checkpoint belongs to a different render: ") + checkpoint_path; return HRT_ERR_IO; }
            if (!full) { g_err = std::string("checkpoint is truncated: ") + checkpoint_path; return HRT_ERR_IO; }
            head.samples_done = on_disk.samples_done; resumed = true;
            CK(cudaMemcpyAsync(s->d_film, host_film.data(), film_bytes, cudaMemcpyHostToDevice, st));
        }
    }
    if (!resumed) CK(cudaMemsetAsync(s->d_film, 0, film_bytes, st));
    const int slice = slice_spp > 0 ? slice_spp : p->spp;
    while (head.samples_done < p->spp) {
        hrt_render_params q = *p;
        q.sample_first = head.samples_done; q.sample_count = std::min(slice, p->spp - head.samples_done);
        rc = render_into(s, cam, &q, s->d_film, st); if (rc < 0) return rc;
        head.samples_done += q.sample_count;
        if (checkpoint_path) {
            CK(cudaMemcpyAsync(host_film.data(), s->d_film, film_bytes, cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            rc = write_checkpoint(checkpoint_path, head, host_film.data(), film_bytes); if (rc < 0) return rc;
        } else if (on_slice) CK(cudaStreamSynchronize(st));
        if (on_slice && on_slice(head.samples_done, p->spp, user) != 0 && head.samples_done < p->spp) { CK(cudaStreamSynchronize(st)); return HRT_STOPPED; }
    }
    uint8_t* d_rgb8 = (uint8_t*)(s->d_out + npix * 3);
    launch_film_resolve(s->d_film, (int)npix, p->spp, out_sum ? s->d_out : nullptr, nullptr, out_rgb8 ? d_rgb8 : nullptr, s->sm_count, st, &s->stats.launches);
    if (out_sum) CK(cudaMemcpyAsync(out_sum, s->d_out, sizeof(float) * npix * 3, cudaMemcpyDeviceToHost, st));
    if (out_rgb8) CK(cudaMemcpyAsync(out_rgb8, d_rgb8, npix * 3, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    return 0;
}

void* hrt_film_create(int width, int height)
{
    const int64_t bytes = hrt_film_bytes(width, height);
    if (bytes < 0) return nullptr;
    void* p = nullptr;
    cudaError_t e = cudaSetDevice(g_device);
    if (e == cudaSuccess) e = cudaMalloc(&p, (size_t)bytes);
    if (e == cudaSuccess) e = cudaMemset(p, 0, (size_t)bytes);
    if (e != cudaSuccess) { g_err = std::string("hrt_film_create: ") + cudaGetErrorString(e); if (p) cudaFree(p); return nullptr; }
    return p;
}
void hrt_film_destroy(void* d_film) { if (d_film) cudaFree(d_film); }
int hrt_film_zero(void* d_film, int width, int height, void* stream)
{
    const int64_t bytes = hrt_film_bytes(width, height);
    if (!d_film || bytes < 0) { g_err = "bad argument"; return HRT_ERR_ARG; }
    if (stream) CK(cudaMemsetAsync(d_film, 0, (size_t)bytes, (cudaStream_t)stream));
    else CK(cudaMemset(d_film, 0, (size_t)bytes));                  // NULL: done when the call returns, so ordered before anything queued afterwards
    return 0;
}
int hrt_film_ipc_handle(void* d_film, void* out_handle)
{
    if (!d_film || !out_handle) { g_err = "NULL argument"; return HRT_ERR_ARG; }
    static_assert(sizeof(cudaIpcMemHandle_t) == HRT_IPC_HANDLE_BYTES, "CUDA IPC handle size");
    cudaIpcMemHandle_t h;
    CK(cudaIpcGetMemHandle(&h, d_film));
    memcpy(out_handle, &h, sizeof(h));
    return 0;
}
int hrt_film_resolve_peers(hrt_scene* s, void* d_film, const void* peer_handles, int n_peers, int width, int height, int spp, int write_back,
                           void* d_sum, void* d_mean, void* d_rgb8, void* stream)
{
    if (!s || !d_film || width < 1 || height < 1 || spp < 1 || n_peers < 0 || n_peers > HRT_MAX_PEERS || (n_peers > 0 && !peer_handles)) { g_err = "bad argument"; return HRT_ERR_ARG; }
    CK(cudaSetDevice(s->device));
    const long long* ptrs[HRT_MAX_PEERS];
    for (int k = 0; k < n_peers; ++k) {
        cudaIpcMemHandle_t h; memcpy(&h, (const char*)peer_handles + (size_t)k * HRT_IPC_HANDLE_BYTES, sizeof(h));
        void* p = nullptr;
        for (auto& pr : s->peers) if (memcmp(&pr.first, &h, sizeof(h)) == 0) p = pr.second;
        if (!p) {
            CK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
            s->peers.emplace_back(h, p);
        }
        ptrs[k] = (const long long*)p;
    }
    cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
    if (!stream) CK(cudaStreamSynchronize(cudaStreamLegacy));
    launch_film_resolve_peers((long long*)d_film, ptrs, n_peers, width * height, spp, write_back, (float*)d_sum, (float*)d_mean, (uint8_t*)d_rgb8,
                              s->sm_count, st, &s->stats.launches);
    CK(cudaGetLastError());
    return 0;
}

// ---- peer-mappable buffers and the reduction spread over the ranks ------------------------------------------------------------------
void* hrt_peer_alloc(int64_t bytes)
{
    if (bytes <= 0) { g_err = "bad size"; return nullptr; }
    void* p = nullptr;
    cudaError_t e = cudaSetDevice(g_device);
    if (e == cudaSuccess) e = cudaMalloc(&p, (size_t)bytes);
    if (e == cudaSuccess) e = cudaMemset(p, 0, (size_t)bytes);
    if (e != cudaSuccess) { g_err = std::string("hrt_peer_alloc: ") + cudaGetErrorString(e); if (p) cudaFree(p); return nullptr; }
    return p;
}
void hrt_peer_free(void* p) { if (p) cudaFree(p); }
int hrt_peer_handle(void* p, void* out_handle) { return hrt_film_ipc_handle(p, out_handle); }
void* hrt_peer_open(hrt_scene* s, const void* handle)
{
    if (!s || !handle) { g_err = "NULL argument"; return nullptr; }
    cudaIpcMemHandle_t h; memcpy(&h, handle, sizeof(h));
    for (auto& pr : s->peers) if (memcmp(&pr.first, &h, sizeof(h)) == 0) return pr.second;
    void* p = nullptr;
    cudaError_t e = cudaSetDevice(s->device);
    if (e == cudaSuccess) e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) { g_err = std::string("cudaIpcOpenMemHandle: ") + cudaGetErrorString(e); cudaGetLastError(); return nullptr; }
    s->peers.emplace_back(h, p);
    return p;
}
int hrt_film_resolve_slice(hrt_scene* s, const void* const* films, int n_films, int width, int height, int spp, int pix_first, int pix_count,
                           void* d_sum, void* d_mean, void* d_rgb8, void* stream)
{
    if (!s || !films || n_films < 1 || n_films > 16 || width < 1 || height < 1 || spp < 1 || pix_first < 0 || pix_count < 0 ||
        (long long)pix_first + pix_count > (long long)width * height) { g_err = "bad argument"; return HRT_ERR_ARG; }
    for (int k = 0; k < n_films; ++k) if (!films[k]) { g_err = "NULL film"; return HRT_ERR_ARG; }
    if (pix_count == 0) return 0;
    CK(cudaSetDevice(s->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
    if (!stream) CK(cudaStreamSynchronize(cudaStreamLegacy));
    launch_film_resolve_slice((const long long* const*)films, n_films, pix_first, pix_count, spp, (float*)d_sum, (float*)d_mean, (uint8_t*)d_rgb8,
                              s->sm_count, st, &s->stats.launches);
    CK(cudaGetLastError());
    return 0;
}

int hrt_get_stats(hrt_scene* s, hrt_stats* out)
{
    if (!s || !out) { g_err = "NULL argument"; return HRT_ERR_ARG; }
    *out = s->stats; out->scene_bytes = s->scene_bytes; out->h2d_bytes = s->h2d_bytes; out->ms_build_upload = s->ms_build_upload; out->ms_build = s->host.build_ms_ref + s->host.build_ms_fast + s->ms_build_gpu;
    return 0;
}
int hrt_reset_stats(hrt_scene* s)
{
    if (!s) { g_err = "scene is NULL"; return HRT_ERR_ARG; }
    s->stats = hrt_stats{}; return 0;
}

} // extern "C"

// ---- one process, several GPUs (main.cpp:176 stays ONE synchronous call) -------------------------------------------
namespace {
// this scene on another device: the same host scene, its own uploads, streams and work buffers
int make_replica(hrt_scene* s, int device, hrt_scene** out)
{
    hrt_scene* r = new (std::nothrow) hrt_scene(s->host_owner);
    if (!r) { g_err = "out of memory"; return HRT_ERR_ARG; }
    r->device = device; r->is_replica = true;
    auto fail = [&](const std::string& why) { g_err = why; cudaSetDevice(device); free_scene_device(r); if (r->stream) cudaStreamDestroy(r->stream); delete r; cudaSetDevice(s->device); return HRT_ERR_CUDA; };
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&r->sm_count, cudaDevAttrMultiProcessorCount, device);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&r->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) return fail(std::string("replica set-up: ") + cudaGetErrorString(e));
    // meshes that were built on the primary's device (HRT_BVH_LBVH_GPU): their arrays cross NVLink, nothing is rebuilt
    r->dev_meshes.resize(s->host.mesh_shapes.size());
    for (size_t i = 0; i < s->host.mesh_shapes.size(); ++i) {
        const HostShape& sh = s->host.shapes[(size_t)s->host.mesh_shapes[i]];
        if (!sh.has_dev) continue;
        DMesh m = sh.dev;
        const void* src[6] = {sh.dev.ref_nodes, sh.dev.tris, sh.dev.attrs, sh.dev.fast_nodes, sh.dev.fast_leaves, sh.dev.wide_nodes};
        const size_t bytes[6] = {sh.n_ref_nodes * 32, (size_t)sh.n_tri * 48, (size_t)sh.n_tri * 64, sh.n_fast_nodes * 64, sh.n_leaves * 128, sh.dev.wide_nodes ? sh.n_wide_nodes * 96 : 0};
        void* dst[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
        for (int k = 0; k < 6; ++k) {
            if (!bytes[k] || !src[k]) continue;
            e = cudaMalloc(&dst[k], bytes[k]);
            if (e == cudaSuccess) { r->scene_allocs.push_back(dst[k]); r->scene_bytes += bytes[k]; e = cudaMemcpyPeer(dst[k], device, src[k], s->device, bytes[k]); }
            if (e != cudaSuccess) return fail(std::string("replica mesh copy: ") + cudaGetErrorString(e));
        }
        m.ref_nodes = (const float4*)dst[0]; m.tris = (const float4*)dst[1]; m.attrs = (const float4*)dst[2];
        m.fast_nodes = (const float4*)dst[3]; m.fast_leaves = (const float4*)dst[4]; m.wide_nodes = (const WNode*)dst[5];
        r->dev_meshes[i] = m;
    }
    DevUp up(r->scene_allocs, r->scene_alloc_bytes, device);
    up.dev_override = &r->dev_meshes;
    r->d = make_dscene(s->host, up);
    r->d.refill_secondary = s->d.refill_secondary; r->d.wide_has_slow = s->d.wide_has_slow; r->d.wide_phase_min = s->d.wide_phase_min; r->d.wide_phase_min_obj = s->d.wide_phase_min_obj; r->d.wide_refill = s->d.wide_refill;
    if (up.failed) return fail("replica upload failed: " + up.msg);
    r->scene_bytes += up.bytes;
    e = cudaDeviceSynchronize();
    if (e != cudaSuccess) return fail(std::string("replica upload: ") + cudaGetErrorString(e));
    cudaSetDevice(s->device);
    *out = r;
    return 0;
}
} // namespace

extern "C" int hrt_render_multi(hrt_scene* s, const int* devices, int n_devices, const hrt_camera* cam, const hrt_render_params* p, float* out_sum, uint8_t* out_rgb8)
{
    int rc = check_render_args(s, cam, p); if (rc < 0) return rc;
    if (!devices || n_devices < 1 || n_devices > HRT_MAX_PEERS + 1) { g_err = "bad device list"; return HRT_ERR_ARG; }
    if (devices[0] != s->device) { g_err = "devices[0] must be the device the scene was committed on"; return HRT_ERR_ARG; }
    if (p->sample_first != 0 || (p->sample_count > 0 && p->sample_count != p->spp)) { g_err = "hrt_render_multi splits all spp samples itself"; return HRT_ERR_ARG; }
    for (int a = 0; a < n_devices; ++a) for (int b = a + 1; b < n_devices; ++b) if (devices[a] == devices[b]) { g_err = "a device is listed twice"; return HRT_ERR_ARG; }
    if (n_devices == 1) return hrt_render(s, cam, p, out_sum, out_rgb8);
    CK(cudaSetDevice(s->device));
    std::vector<hrt_scene*> part((size_t)n_devices, nullptr);
    part[0] = s;
    for (int k = 1; k < n_devices; ++k) {
        for (hrt_scene* r : s->replicas) if (r->device == devices[k]) part[(size_t)k] = r;
        if (!part[(size_t)k]) { hrt_scene* r = nullptr; rc = make_replica(s, devices[k], &r); if (rc < 0) return rc; s->replicas.push_back(r); part[(size_t)k] = r; }
    }
    const size_t npix = (size_t)p->width * p->height;
    // every device renders its contiguous share of the sample indices into its own integer film (hrt_render_accumulate_device's contract)
    std::vector<int> status((size_t)n_devices, 0); std::vector<std::string> errs((size_t)n_devices);
    auto work = [&](int k) {
        hrt_scene* r = part[(size_t)k];
        const int base = p->spp / n_devices, extra = p->spp % n_devices;
        hrt_render_params q = *p;
        q.sample_first = k * base + std::min(k, extra); q.sample_count = base + (k < extra ? 1 : 0);
        int st = 0;
        if (cudaSetDevice(r->device) != cudaSuccess) st = HRT_ERR_CUDA;
        if (st == 0) st = ensure_film(r, npix);
        if (st == 0 && cudaMemsetAsync(r->d_film, 0, sizeof(long long) * FILM_LANES * npix, r->stream) != cudaSuccess) st = HRT_ERR_CUDA;
        if (st == 0 && q.sample_count > 0) st = render_into(r, cam, &q, r->d_film, r->stream);
        if (st == 0 && cudaStreamSynchronize(r->stream) != cudaSuccess) st = HRT_ERR_CUDA;
        status[(size_t)k] = st; if (st < 0) errs[(size_t)k] = g_err.empty() ? "CUDA failure on a worker" : g_err;
    };
    std::vector<std::thread> th;
    for (int k = 1; k < n_devices; ++k) th.emplace_back(work, k);
    work(0);
    for (auto& t : th) t.join();
    for (int k = 0; k < n_devices; ++k) if (status[(size_t)k] < 0) { g_err = "device " + std::to_string(devices[k]) + ": " + errs[(size_t)k]; return status[(size_t)k]; }
    // reduce + resolve in one kernel on devices[0]: the other films are read in place over NVLink (peer access), or staged when the
    // devices cannot address each other
    CK(cudaSetDevice(s->device));
    const long long* peers[HRT_MAX_PEERS]; std::vector<void*> staged;
    for (int k = 1; k < n_devices; ++k) {
        int can = 0; CK(cudaDeviceCanAccessPeer(&can, s->device, devices[k]));
        if (can) {
            cudaError_t e = cudaDeviceEnablePeerAccess(devices[k], 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
            if (e != cudaSuccess) can = 0;
        }
        if (can) peers[k - 1] = part[(size_t)k]->d_film;
        else {
            void* tmp = nullptr; CK(cudaMalloc(&tmp, sizeof(long long) * FILM_LANES * npix)); staged.push_back(tmp);
            CK(cudaMemcpyPeer(tmp, s->device, part[(size_t)k]->d_film, devices[k], sizeof(long long) * FILM_LANES * npix));
            peers[k - 1] = (const long long*)tmp;
        }
    }
    cudaStream_t st = s->stream;
    uint8_t* d_rgb8 = (uint8_t*)(s->d_out + npix * 3);
    launch_film_resolve_peers(s->d_film, peers, n_devices - 1, (int)npix, p->spp, 0, out_sum ? s->d_out : nullptr, nullptr, out_rgb8 ? d_rgb8 : nullptr, s->sm_count, st, &s->stats.launches);
    if (out_sum) CK(cudaMemcpyAsync(out_sum, s->d_out, sizeof(float) * npix * 3, cudaMemcpyDeviceToHost, st));
    if (out_rgb8) CK(cudaMemcpyAsync(out_rgb8, d_rgb8, npix * 3, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    for (void* t : staged) cudaFree(t);
    CK(cudaGetLastError());
    for (int k = 1; k < n_devices; ++k) {                                         // one set of counters for the caller
        hrt_stats& a = s->stats; hrt_stats& b = part[(size_t)k]->stats;
        a.rays += b.rays; a.paths += b.paths; a.launches += b.launches; a.retraced += b.retraced; a.prim_tests += b.prim_tests; a.film_clamped += b.film_clamped;
        a.ms_trace = std::max(a.ms_trace, b.ms_trace); a.ms_shade = std::max(a.ms_shade, b.ms_shade); a.ms_total = std::max(a.ms_total, b.ms_total);
        b = hrt_stats{};
    }
    return 0;
}
