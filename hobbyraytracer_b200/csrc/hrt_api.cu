// hrt_api.cu — the C ABI of include/hrt.h on top of the CUDA kernels.  This is the drop-in boundary: the
// reference's main() would call hrt_render() where it calls render(NUM_THREADS, background, world, camera, film)
// (main.cpp:176), after replaying its loader's constructor calls (scene.cpp:127-374) through hrt_scene_add_*.
// CUDA only: there is no CPU path in this library; without a device every compute entry point returns HRT_ERR_CUDA.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "hrt_hit.cuh"
#include "hrt_kernels.h"
#include "hrt_upload.h"

using namespace hrt;

namespace {
thread_local std::string g_err;
int g_device = 0;

#define CK(call)                                                                                          \
    do {                                                                                                  \
        cudaError_t e_ = (call);                                                                          \
        if (e_ != cudaSuccess) {                                                                          \
            g_err = std::string(#call) + ": " + cudaGetErrorString(e_);                                   \
            return HRT_ERR_CUDA;                                                                          \
        }                                                                                                 \
    } while (0)

struct DevUp {
    std::vector<void*>& allocs; bool failed = false; std::string msg; size_t bytes = 0;
    explicit DevUp(std::vector<void*>& a) : allocs(a) {}
    const void* put(const void* p, size_t n)
    {
        if (!n || failed) return nullptr;
        void* d = nullptr;
        cudaError_t e = cudaMalloc(&d, n);
        if (e == cudaSuccess) e = cudaMemcpy(d, p, n, cudaMemcpyHostToDevice);
        if (e != cudaSuccess) { failed = true; msg = cudaGetErrorString(e); if (d) cudaFree(d); return nullptr; }
        allocs.push_back(d); bytes += n;
        return d;
    }
    const void* put_copy(const void* p, size_t n) { return put(p, n); }
};
} // namespace

struct hrt_scene {
    HostScene host;
    DScene d{};
    int device = 0, sm_count = 148;
    cudaStream_t stream = nullptr;
    std::vector<void*> scene_allocs;
    size_t scene_bytes = 0;
    // wavefront working set
    RenderJob job{};
    long long job_capacity = 0; int job_depth = 0;
    float* d_accum = nullptr; size_t accum_elems = 0;
    std::vector<cudaEvent_t> events;            // 2 pools x (2 trace + 1 shade + 2 batch) per bounce
    // ray-batch working set
    float4* tb_rays = nullptr; Hit* tb_hits = nullptr; unsigned int* tb_retrace = nullptr; unsigned int* tb_n_retrace = nullptr;
    uint32_t* tb_prim = nullptr; float* tb_t = nullptr; long long tb_capacity = 0;
    hrt_stats stats{};
};

namespace {

void free_job(hrt_scene* s)
{
    cudaFree(s->job.rays[0]); cudaFree(s->job.rays[1]); cudaFree(s->job.hits); cudaFree(s->job.atten); cudaFree(s->job.radiance);
    cudaFree(s->job.counts); cudaFree(s->job.retrace); cudaFree(s->job.n_retrace); cudaFree(s->job.totals);
    s->job = RenderJob{}; s->job_capacity = 0; s->job_depth = 0;
}
void free_tb(hrt_scene* s)
{
    cudaFree(s->tb_rays); cudaFree(s->tb_hits); cudaFree(s->tb_retrace); cudaFree(s->tb_n_retrace); cudaFree(s->tb_prim); cudaFree(s->tb_t);
    s->tb_rays = nullptr; s->tb_hits = nullptr; s->tb_retrace = nullptr; s->tb_n_retrace = nullptr; s->tb_prim = nullptr; s->tb_t = nullptr; s->tb_capacity = 0;
}
void free_scene_device(hrt_scene* s)
{
    for (void* p : s->scene_allocs) cudaFree(p);
    s->scene_allocs.clear(); s->scene_bytes = 0;
}

int ensure_job(hrt_scene* s, long long capacity, int max_depth)
{
    if (capacity <= s->job_capacity && max_depth <= s->job_depth) return 0;
    free_job(s);
    CK(cudaMalloc(&s->job.rays[0], sizeof(float4) * 2 * (size_t)capacity));
    CK(cudaMalloc(&s->job.rays[1], sizeof(float4) * 2 * (size_t)capacity));
    CK(cudaMalloc(&s->job.hits, sizeof(Hit) * (size_t)capacity));
    CK(cudaMalloc(&s->job.atten, sizeof(float4) * (size_t)capacity));
    CK(cudaMalloc(&s->job.radiance, sizeof(float4) * (size_t)capacity));
    CK(cudaMalloc(&s->job.counts, sizeof(unsigned int) * (size_t)(max_depth + 1)));
    CK(cudaMalloc(&s->job.retrace, sizeof(unsigned int) * (size_t)capacity));
    CK(cudaMalloc(&s->job.n_retrace, sizeof(unsigned int) * (size_t)(max_depth + 1)));
    CK(cudaMalloc(&s->job.totals, sizeof(unsigned long long) * 2));
    s->job_capacity = capacity; s->job_depth = max_depth;
    return 0;
}
int ensure_tb(hrt_scene* s, long long n, bool own_rays)
{
    if (n > s->tb_capacity) {
        free_tb(s);
        CK(cudaMalloc(&s->tb_rays, sizeof(float4) * 2 * (size_t)n));
        CK(cudaMalloc(&s->tb_hits, sizeof(Hit) * (size_t)n));
        CK(cudaMalloc(&s->tb_retrace, sizeof(unsigned int) * (size_t)n));
        CK(cudaMalloc(&s->tb_n_retrace, sizeof(unsigned int)));
        CK(cudaMalloc(&s->tb_prim, sizeof(uint32_t) * (size_t)n));
        CK(cudaMalloc(&s->tb_t, sizeof(float) * (size_t)n));
        s->tb_capacity = n;
    }
    (void)own_rays;
    return 0;
}

long long batch_paths()
{
    const char* e = getenv("HRT_BATCH_PATHS");
    long long v = e ? atoll(e) : 0;
    return v > 0 ? v : (4ll << 20);
}

int check_render_args(hrt_scene* s, const hrt_camera* cam, const hrt_render_params* p)
{
    if (!s || !cam || !p) { g_err = "NULL argument"; return HRT_ERR_ARG; }
    if (!s->host.committed) { g_err = "scene not committed"; return HRT_ERR_STATE; }
    if (p->width < 2 || p->height < 2 || p->spp < 1) { g_err = "width/height must be >= 2 and spp >= 1"; return HRT_ERR_ARG; }
    if ((long long)p->width * p->height > 0x3fffffffll) { g_err = "film too large"; return HRT_ERR_ARG; }
    if (p->sample_first < 0) { g_err = "negative sample_first"; return HRT_ERR_ARG; }
    return 0;
}

// renders samples [first, first+count) of every pixel and adds them into d_accum (sum, not mean)
int render_into(hrt_scene* s, const hrt_camera* cam, const hrt_render_params* p, float* d_accum, cudaStream_t st)
{
    const int npix = p->width * p->height;
    const int first = p->sample_first, count = p->sample_count > 0 ? p->sample_count : p->spp;
    const int max_depth = p->max_depth > 0 ? p->max_depth : 50;
    const bool mega = (p->flags & HRT_RENDER_MEGAKERNEL) != 0;
    long long per_batch = batch_paths() / npix; if (per_batch < 1) per_batch = 1; if (per_batch > count) per_batch = count;
    int rc = ensure_job(s, (long long)npix * per_batch, max_depth); if (rc < 0) return rc;
    RenderJob j = s->job;
    j.width = p->width; j.height = p->height; j.npix = npix; j.max_depth = max_depth;
    j.t_min = p->t_min > 0 ? p->t_min : 0.001f; j.seed = p->seed; j.rng_mode = p->rng_mode;
    DCamera dc = make_camera(*cam);
    CK(cudaMemsetAsync(j.totals, 0, sizeof(unsigned long long) * 2, st));
    // events: per pool 3 per bounce + 2 for the whole batch
    const size_t per_pool = (size_t)max_depth * 3 + 2;
    while (s->events.size() < per_pool * 2) { cudaEvent_t e; CK(cudaEventCreate(&e)); s->events.push_back(e); }
    auto collect = [&](int pool) -> int {
        cudaEvent_t* ev = s->events.data() + per_pool * (size_t)pool;
        CK(cudaEventSynchronize(ev[per_pool - 1]));
        float ms = 0;
        if (!mega)
            for (int b = 0; b < max_depth; ++b) {
                CK(cudaEventElapsedTime(&ms, ev[2 * b], ev[2 * b + 1])); s->stats.ms_trace += ms;
                CK(cudaEventElapsedTime(&ms, ev[2 * b + 1], ev[2 * max_depth + b])); s->stats.ms_shade += ms;
            }
        CK(cudaEventElapsedTime(&ms, ev[per_pool - 2], ev[per_pool - 1])); s->stats.ms_total += ms;
        return 0;
    };
    int batch = 0;
    for (int done = 0; done < count; ++batch) {
        int n = (int)std::min<long long>(per_batch, count - done);
        int pool = batch & 1;
        if (batch >= 2) { rc = collect(pool); if (rc < 0) return rc; }
        cudaEvent_t* ev = s->events.data() + per_pool * (size_t)pool;
        CK(cudaEventRecord(ev[per_pool - 2], st));
        launch_render_batch(s->d, j, dc, first + done, n, d_accum, mega, s->sm_count, st, &s->stats.launches, mega ? nullptr : ev, mega ? nullptr : ev + 2 * max_depth);
        CK(cudaEventRecord(ev[per_pool - 1], st));
        CK(cudaGetLastError());
        done += n;
        s->stats.paths += (uint64_t)npix * (uint64_t)n;
    }
    for (int k = std::max(0, batch - 2); k < batch; ++k) { rc = collect(k & 1); if (rc < 0) return rc; }
    unsigned long long totals[2] = {0, 0};
    CK(cudaMemcpyAsync(totals, j.totals, sizeof(totals), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    s->stats.rays += totals[0]; s->stats.retraced += totals[1];
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
void hrt_scene_destroy(hrt_scene* s)
{
    if (!s) return;
    if (s->host.committed || s->stream) {
        cudaSetDevice(s->device);
        free_job(s); free_tb(s); free_scene_device(s);
        cudaFree(s->d_accum);
        for (cudaEvent_t e : s->events) cudaEventDestroy(e);
        if (s->stream) cudaStreamDestroy(s->stream);
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
    if (bvh_mode < HRT_BVH_REFERENCE || bvh_mode > HRT_BVH_LBVH_GPU) { g_err = "unknown bvh mode"; return HRT_ERR_ARG; }
    if (bvh_mode == HRT_BVH_LBVH_GPU) bvh_mode = HRT_BVH_SAH;     // the GPU builder is not implemented yet: host SAH, same node format
    CK(cudaSetDevice(s->device));
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, s->device));
    s->sm_count = prop.multiProcessorCount;
    std::string e = flatten(s->host, bvh_mode, seed);
    if (!e.empty()) { g_err = e; return HRT_ERR_ARG; }
    CK(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    DevUp up(s->scene_allocs);
    s->d = make_dscene(s->host, up);
    if (up.failed) { g_err = "scene upload failed: " + up.msg; free_scene_device(s); return HRT_ERR_CUDA; }
    s->scene_bytes = up.bytes;
    CK(cudaDeviceSynchronize());
    s->host.committed = true;
    return 0;
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
    launch_trace_batch(s->d, (const float4*)d_rays, n, s->tb_hits, rng_seed, s->tb_retrace, s->tb_n_retrace, s->sm_count, st, &s->stats.launches);
    launch_split_hits(s->tb_hits, n, (uint32_t*)d_prim_id, (float*)d_t, s->sm_count, st, &s->stats.launches);
    CK(cudaGetLastError());
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
    CK(cudaMemcpyAsync(s->tb_rays, rays, sizeof(float) * 8 * (size_t)n, cudaMemcpyHostToDevice, st));
    rc = hrt_trace_batch_device(s, s->tb_rays, n, s->tb_prim, s->tb_t, rng_seed, st); if (rc < 0) return rc;
    CK(cudaMemcpyAsync(prim_id, s->tb_prim, sizeof(uint32_t) * (size_t)n, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(t, s->tb_t, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost, st));
    if (normal || uv || front) {
        float* d_n = nullptr; float* d_uv = nullptr; uint8_t* d_f = nullptr;
        if (normal) CK(cudaMalloc(&d_n, sizeof(float) * 3 * (size_t)n));
        if (uv) CK(cudaMalloc(&d_uv, sizeof(float) * 2 * (size_t)n));
        if (front) CK(cudaMalloc(&d_f, (size_t)n));
        launch_eval_batch(s->d, s->tb_rays, s->tb_hits, n, d_n, d_uv, d_f, s->sm_count, st, &s->stats.launches);
        if (normal) CK(cudaMemcpyAsync(normal, d_n, sizeof(float) * 3 * (size_t)n, cudaMemcpyDeviceToHost, st));
        if (uv) CK(cudaMemcpyAsync(uv, d_uv, sizeof(float) * 2 * (size_t)n, cudaMemcpyDeviceToHost, st));
        if (front) CK(cudaMemcpyAsync(front, d_f, (size_t)n, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        cudaFree(d_n); cudaFree(d_uv); cudaFree(d_f);
    }
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    return 0;
}

int hrt_render_accumulate_device(hrt_scene* s, const hrt_camera* cam, const hrt_render_params* p, void* d_accum, void* stream)
{
    int rc = check_render_args(s, cam, p); if (rc < 0) return rc;
    if (!d_accum) { g_err = "d_accum is NULL"; return HRT_ERR_ARG; }
    CK(cudaSetDevice(s->device));
    return render_into(s, cam, p, (float*)d_accum, stream ? (cudaStream_t)stream : s->stream);
}

int hrt_film_resolve_device(hrt_scene* s, const void* d_accum, int width, int height, int spp, void* d_mean, void* d_rgb8, void* stream)
{
    if (!s || !d_accum || width < 1 || height < 1 || spp < 1) { g_err = "bad argument"; return HRT_ERR_ARG; }
    CK(cudaSetDevice(s->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
    launch_film_resolve((const float*)d_accum, width * height, spp, (float*)d_mean, (uint8_t*)d_rgb8, s->sm_count, st, &s->stats.launches);
    CK(cudaGetLastError());
    return 0;
}

int hrt_render(hrt_scene* s, const hrt_camera* cam, const hrt_render_params* p, float* out_sum, uint8_t* out_rgb8)
{
    int rc = check_render_args(s, cam, p); if (rc < 0) return rc;
    CK(cudaSetDevice(s->device));
    const size_t npix = (size_t)p->width * p->height;
    if (s->accum_elems < npix * 3) {
        cudaFree(s->d_accum); s->d_accum = nullptr; s->accum_elems = 0;
        CK(cudaMalloc(&s->d_accum, sizeof(float) * npix * 3 + npix * 3));
        s->accum_elems = npix * 3;
    }
    cudaStream_t st = s->stream;
    CK(cudaMemsetAsync(s->d_accum, 0, sizeof(float) * npix * 3, st));
    rc = render_into(s, cam, p, s->d_accum, st); if (rc < 0) return rc;
    if (out_sum) CK(cudaMemcpyAsync(out_sum, s->d_accum, sizeof(float) * npix * 3, cudaMemcpyDeviceToHost, st));
    if (out_rgb8) {
        uint8_t* d_rgb8 = (uint8_t*)(s->d_accum + npix * 3);
        launch_film_resolve(s->d_accum, (int)npix, p->spp, nullptr, d_rgb8, s->sm_count, st, &s->stats.launches);
        CK(cudaMemcpyAsync(out_rgb8, d_rgb8, npix * 3, cudaMemcpyDeviceToHost, st));
    }
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    return 0;
}

int hrt_get_stats(hrt_scene* s, hrt_stats* out)
{
    if (!s || !out) { g_err = "NULL argument"; return HRT_ERR_ARG; }
    *out = s->stats; out->scene_bytes = s->scene_bytes; out->ms_build = s->host.build_ms_ref + s->host.build_ms_fast;
    return 0;
}
int hrt_reset_stats(hrt_scene* s)
{
    if (!s) { g_err = "scene is NULL"; return HRT_ERR_ARG; }
    s->stats = hrt_stats{}; return 0;
}

} // extern "C"
