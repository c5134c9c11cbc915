// hostsim.cpp — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Compiles the product's device headers (hobbyraytracer_b200/csrc/hrt_hit.cuh, hrt_shade.cuh — the exact source
// the CUDA kernels are built from) and its host flattener with g++ and runs them on the CPU, so that the
// traversal logic (reference order, SAH order + tie re-trace) and the shading switch can be checked against the
// oracle in the `-m "not gpu"` suite, in a container without a GPU.  It exports the scene-builder ABI of
// include/hrt.h with the prefix hsim_.  The product library never links or loads this file; nothing here is a
// fallback for the CUDA path.
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <limits>
#include <string>
#include <thread>
#include <vector>

#include "../../hobbyraytracer_b200/csrc/hrt_shade.cuh"
#include "../../hobbyraytracer_b200/csrc/hrt_wide.cuh"
#include "../../hobbyraytracer_b200/csrc/hrt_upload.h"

using namespace hrt;

namespace {
thread_local std::string g_err;
struct HostUp {
    std::deque<std::vector<char>> keep;
    const void* put(const void* p, size_t n) { return n ? p : nullptr; }
    DMesh dev_mesh(size_t, const HostShape& sh) { return sh.dev; }
    const void* put_copy(const void* p, size_t n) { if (!n) return nullptr; keep.emplace_back((const char*)p, (const char*)p + n); return keep.back().data(); }
};
template <class F> void parallel_for(int64_t n, int64_t chunk, F&& fn)
{
    int nt = (int)std::thread::hardware_concurrency(); if (nt < 1) nt = 1;
    std::atomic<int64_t> next(0);
    auto worker = [&]() { for (;;) { int64_t b = next.fetch_add(chunk); if (b >= n) break; int64_t e = b + chunk < n ? b + chunk : n; for (int64_t i = b; i < e; ++i) fn(i); } };
    std::vector<std::thread> th; for (int i = 1; i < nt; ++i) th.emplace_back(worker);
    worker(); for (auto& t : th) t.join();
}
} // namespace

struct hsim_scene { HostScene host; HostUp up; DScene d{}; TraceCounters cnt{0, 0, 0, 0, 0}; };

#define B(call) do { int rc = (call); return rc; } while (0)

extern "C" {
const char* hsim_last_error(void) { return g_err.c_str(); }
hsim_scene* hsim_scene_create(void) { return new hsim_scene(); }
void hsim_scene_destroy(hsim_scene* s) { delete s; }
int hsim_scene_add_texture_solid(hsim_scene* s, const float rgb[3]) { B(b_add_texture_solid(s->host, rgb, g_err)); }
int hsim_scene_add_texture_checker(hsim_scene* s, int e, int o) { B(b_add_texture_checker(s->host, e, o, g_err)); }
int hsim_scene_add_texture_image(hsim_scene* s, const uint8_t* p, int w, int h) { B(b_add_texture_image(s->host, p, w, h, g_err)); }
int hsim_scene_add_texture_env(hsim_scene* s, const float* p, int w, int h, int c) { B(b_add_texture_env(s->host, p, w, h, c, g_err)); }
int hsim_scene_add_material(hsim_scene* s, const hrt_material* m) { B(b_add_material(s->host, m, g_err)); }
int hsim_scene_add_sphere(hsim_scene* s, const float c[3], float r, int m) { B(b_add_sphere(s->host, c, r, m, g_err)); }
int hsim_scene_add_rect(hsim_scene* s, int axis, float a0, float a1, float b0, float b1, float k, int m) { B(b_add_rect(s->host, axis, a0, a1, b0, b1, k, m, g_err)); }
int hsim_scene_add_box(hsim_scene* s, const float c[3], const float d[3], int m) { B(b_add_box(s->host, c, d, m, g_err)); }
int hsim_scene_add_mesh(hsim_scene* s, const float* p, const float* n, const float* uv, int64_t nt, int m) { B(b_add_mesh(s->host, p, n, uv, nt, m, g_err)); }
int hsim_scene_add_mesh_indexed(hsim_scene* s, const float* v, const float* n, const float* uv, int64_t nv, const uint32_t* idx, int64_t ni,
                                const hrt_submesh* sub, int n_sub, int m, uint32_t flags) { B(b_add_mesh_indexed(s->host, v, n, uv, nv, idx, ni, sub, n_sub, m, flags, g_err)); }
int64_t hsim_scene_prim_base(const hsim_scene* s, int shape) { return s->host.shapes[(size_t)shape].prim_base; }
int64_t hsim_scene_prim_count(const hsim_scene* s) { return s->host.n_prims; }
int hsim_scene_add_object(hsim_scene* s, int shape, const hrt_xform* ops, int n) { B(b_add_object(s->host, shape, ops, n, g_err)); }
int hsim_scene_add_medium(hsim_scene* s, int shape, const hrt_xform* ops, int n, float density, const float rgb[3], int tex) { B(b_add_medium(s->host, shape, ops, n, density, rgb, tex, g_err)); }
int hsim_scene_set_background(hsim_scene* s, int tex) { B(b_set_background(s->host, tex, g_err)); }

int hsim_scene_commit(hsim_scene* s, int bvh_mode, uint64_t seed)
{
    if (bvh_mode == HRT_BVH_LBVH_GPU) bvh_mode = HRT_BVH_LBVH_HOST;     // no device here: the host mirror of the GPU builder
    std::string e = flatten(s->host, bvh_mode, seed);
    if (!e.empty()) { g_err = e; return HRT_ERR_ARG; }
    s->d = make_dscene(s->host, s->up);
    for (HostShape& sh : s->host.shapes) sh.release_borrowed();
    s->host.committed = true;
    return 0;
}

int hsim_trace_batch(hsim_scene* s, const float* rays, int64_t n, uint32_t* prim_id, float* t,
                     float* normal, float* uv, uint8_t* front, uint64_t rng_seed)
{
    if (!s->host.committed) { g_err = "scene not committed"; return HRT_ERR_STATE; }
    std::atomic<unsigned long long> nodes(0), prims(0), fb(0), amb(0), bad(0);
    parallel_for(n, 4096, [&](int64_t i) {
        const float* q = rays + i * 8;
        Ray r; r.o = v3(q[0], q[1], q[2]); r.d = v3(q[4], q[5], q[6]);
        RngKey key; key.k0 = (uint32_t)rng_seed; key.k1 = (uint32_t)(rng_seed >> 32); key.index = (uint32_t)i; key.sample = (uint32_t)(i >> 32); key.mode = 0;
        TraceCounters c{0, 0, 0, 0, 0};
        Hit h = s->d.fast ? trace_fast<true>(s->d, r, q[3], q[7], key, 0, &c) : trace_ref(s->d, r, q[3], q[7], key, 0);
        nodes += c.nodes; prims += c.prims; fb += c.fallbacks; amb += c.amb; bad += c.bad;
        bool hit = h.prim != HRT_MISS;
        prim_id[i] = h.prim; t[i] = hit ? h.t : 0.0f;
        if (normal || uv || front) {
            Rec rec; rec.n = v3(0.0f); rec.u = rec.v = 0; rec.front = 0;
            if (hit) eval_hit(s->d, r, h, rec);
            if (normal) { normal[i * 3] = hit ? rec.n.x : 0; normal[i * 3 + 1] = hit ? rec.n.y : 0; normal[i * 3 + 2] = hit ? rec.n.z : 0; }
            if (uv) { uv[i * 2] = hit ? rec.u : 0; uv[i * 2 + 1] = hit ? rec.v : 0; }
            if (front) front[i] = hit ? (rec.front ? 1 : 0) : 0;
        }
    });
    s->cnt.nodes = nodes; s->cnt.prims = prims; s->cnt.fallbacks = fb; s->cnt.amb = amb; s->cnt.bad = bad;
    return 0;
}
// The same through the 8-wide trees (hrt_wide.cuh): pass 1 on the wide nodes, pending rays through trace_resolve.
// out_steps (may be NULL) receives the number of node / leaf / entry steps; returns HRT_ERR_STATE when the scene has no wide trees.
int hsim_trace_batch_wide(hsim_scene* s, const float* rays, int64_t n, uint32_t* prim_id, float* t, uint64_t rng_seed, uint64_t* out_steps)
{
    if (!s->host.committed) { g_err = "scene not committed"; return HRT_ERR_STATE; }
    if (!s->d.wide) { g_err = "scene has no wide trees"; return HRT_ERR_STATE; }
    std::atomic<unsigned long long> steps(0);
    parallel_for(n, 4096, [&](int64_t i) {
        const float* q = rays + i * 8;
        Ray r; r.o = v3(q[0], q[1], q[2]); r.d = v3(q[4], q[5], q[6]);
        RngKey key; key.k0 = (uint32_t)rng_seed; key.k1 = (uint32_t)(rng_seed >> 32); key.index = (uint32_t)i; key.sample = (uint32_t)(i >> 32); key.mode = 0;
        unsigned long long st = 0;
        Hit h = wide_trace_scalar(s->d, r, q[3], q[7], key, 0, &st);
        steps += st;
        prim_id[i] = h.prim; t[i] = h.prim != HRT_MISS ? h.t : 0.0f;
    });
    if (out_steps) *out_steps = steps;
    return 0;
}
// Visit counts of the wide walk per ray kind of step (node steps, triangle leaves, mesh entries, other world-list entries) and the rays left to
// pass 2: the figures behind SURVEY 8(d)'s supplementary B_logical (tools/b_logical.py).  out5: the four counters + pending rays.
int hsim_wide_visit_counts(hsim_scene* s, const float* rays, int64_t n, uint64_t rng_seed, uint64_t* out5)
{
    if (!s->host.committed) { g_err = "scene not committed"; return HRT_ERR_STATE; }
    if (!s->d.wide) { g_err = "scene has no wide trees"; return HRT_ERR_STATE; }
    std::atomic<unsigned long long> k0(0), k1(0), k2(0), k3(0), pend(0);
    parallel_for(n, 4096, [&](int64_t i) {
        const float* q = rays + i * 8;
        Ray r; r.o = v3(q[0], q[1], q[2]); r.d = v3(q[4], q[5], q[6]);
        RngKey key; key.k0 = (uint32_t)rng_seed; key.k1 = (uint32_t)(rng_seed >> 32); key.index = (uint32_t)i; key.sample = (uint32_t)(i >> 32); key.mode = 0;
        unsigned long long st = 0, kinds[4] = {0, 0, 0, 0};
        Hit h = wide_trace_scalar(s->d, r, q[3], q[7], key, 0, &st, kinds);
        k0 += kinds[0]; k1 += kinds[1]; k2 += kinds[2]; k3 += kinds[3];
        if (h.flags & (HIT_RESOLVED | HIT_FALLBACK)) pend += 1;
    });
    out5[0] = k0; out5[1] = k1; out5[2] = k2; out5[3] = k3; out5[4] = pend;
    return 0;
}
// Texture::colourValue through the device header on the flattened scene (exercises the lossless RGBE repack of env maps)
int hsim_texture_eval(hsim_scene* s, int tex, const float* in, int64_t n, float* out)
{
    if (!s->host.committed || tex < 0 || tex >= s->d.n_textures) { g_err = "bad texture / scene not committed"; return HRT_ERR_ARG; }
    for (int64_t i = 0; i < n; ++i) {
        V3 c = tex_value(s->d, tex, in[i*5], in[i*5+1], v3(in[i*5+2], in[i*5+3], in[i*5+4]));
        out[i*3] = c.x; out[i*3+1] = c.y; out[i*3+2] = c.z;
    }
    return 0;
}
int hsim_texture_kind(hsim_scene* s, int tex) { return (s->host.committed && tex >= 0 && tex < s->d.n_textures) ? s->d.textures[tex].kind : -1; }
int hsim_counters(hsim_scene* s, uint64_t* out3) { out3[0] = s->cnt.nodes; out3[1] = s->cnt.prims; out3[2] = s->cnt.fallbacks; return 0; }
int hsim_counters5(hsim_scene* s, uint64_t* out5) { out5[0] = s->cnt.nodes; out5[1] = s->cnt.prims; out5[2] = s->cnt.fallbacks; out5[3] = s->cnt.amb; out5[4] = s->cnt.bad; return 0; }

// render(): per-pixel sums of rayColour over `spp` samples, one path at a time (what the megakernel does)
int hsim_render(hsim_scene* s, const hrt_camera* cam, const hrt_render_params* p, float* out_sum, uint8_t* out_rgb8)
{
    if (!s->host.committed) { g_err = "scene not committed"; return HRT_ERR_STATE; }
    DCamera c = make_camera(*cam);
    const int w = p->width, h = p->height;
    const int first = p->sample_first, count = p->sample_count > 0 ? p->sample_count : p->spp;
    const int max_depth = p->max_depth > 0 ? p->max_depth : 50;
    const float t_min = p->t_min > 0 ? p->t_min : 0.001f;
    parallel_for((int64_t)w * h, 256, [&](int64_t pidx) {
        V3 sum = v3(0.0f);
        for (int sidx = first; sidx < first + count; ++sidx) {
            RngKey key; key.k0 = (uint32_t)p->seed; key.k1 = (uint32_t)(p->seed >> 32); key.index = (uint32_t)pidx; key.sample = (uint32_t)sidx; key.mode = p->rng_mode;
            Ray r = primary_ray(c, w, h, (int)pidx, key);
            V3 atten = v3(1.0f), result = v3(0.0f);
            for (int b = 0; b < max_depth; ++b) {
                Hit hh = s->d.fast ? trace_fast<false>(s->d, r, t_min, hrt_inf(), key, (uint32_t)b, nullptr) : trace_ref(s->d, r, t_min, hrt_inf(), key, (uint32_t)b);
                if ((hh.flags & HIT_FALLBACK) && getenv("HSIM_DEBUG")) fprintf(stderr, "fallback ray bounce %d o=(%.9g %.9g %.9g) d=(%.9g %.9g %.9g)\n", b, r.o.x, r.o.y, r.o.z, r.d.x, r.d.y, r.d.z);
                V3 rad; bool go = shade_segment(s->d, hh, r, atten, rad, key, (uint32_t)b);
                result = result + rad;
                if (!go) break;
            }
            sum = sum + result;
        }
        if (out_sum) { out_sum[pidx * 3] = sum.x; out_sum[pidx * 3 + 1] = sum.y; out_sum[pidx * 3 + 2] = sum.z; }
        if (out_rgb8) tonemap_write(sum / (float)p->spp, out_rgb8 + pidx * 3);
    });
    return 0;
}

// The film entry points of include/hrt.h on host memory (same fixed-point arithmetic as the kernels, hrt_shade.cuh):
// what the multi-GPU driver (hobbyraytracer_b200/multigpu.py) calls, so its sample split + integer reduce can be run
// on CPU tensors over gloo.
int hsim_film_bytes(int width, int height) { return width * height * FILM_LANES * (int)sizeof(long long); }
int hsim_render_accumulate_device(hsim_scene* s, const hrt_camera* cam, const hrt_render_params* p, void* film_, void*)
{
    if (!s->host.committed) { g_err = "scene not committed"; return HRT_ERR_STATE; }
    long long* film = (long long*)film_;
    DCamera c = make_camera(*cam);
    const int w = p->width, h = p->height;
    const int first = p->sample_first, count = p->sample_count > 0 ? p->sample_count : p->spp;
    const int max_depth = p->max_depth > 0 ? p->max_depth : 50;
    const float t_min = p->t_min > 0 ? p->t_min : 0.001f;
    parallel_for((int64_t)w * h, 256, [&](int64_t pidx) {
        long long* px = film + pidx * FILM_LANES;
        for (int sidx = first; sidx < first + count; ++sidx) {
            RngKey key; key.k0 = (uint32_t)p->seed; key.k1 = (uint32_t)(p->seed >> 32); key.index = (uint32_t)pidx; key.sample = (uint32_t)sidx; key.mode = p->rng_mode;
            Ray r = primary_ray(c, w, h, (int)pidx, key);
            V3 atten = v3(1.0f), result = v3(0.0f);
            for (int b = 0; b < max_depth; ++b) {
                Hit hh = s->d.fast ? trace_fast<false>(s->d, r, t_min, hrt_inf(), key, (uint32_t)b, nullptr) : trace_ref(s->d, r, t_min, hrt_inf(), key, (uint32_t)b);
                V3 rad; bool go = shade_segment(s->d, hh, r, atten, rad, key, (uint32_t)b);
                if (rad.x != 0.0f || rad.y != 0.0f || rad.z != 0.0f) result = result + rad;
                if (!go) break;
            }
            const float v[3] = {result.x, result.y, result.z};
            for (int ch = 0; ch < 3; ++ch) {
                if (v[ch] == 0.0f) continue;
                long long q; unsigned long long flag; film_quantize(v[ch], ch, q, flag);
                if (flag) px[3] = (long long)((unsigned long long)px[3] | flag); else px[ch] += q;
            }
        }
    });
    return 0;
}
int hsim_film_resolve_device(hsim_scene*, const void* film_, int width, int height, int spp, void* sum_, void* mean_, void* rgb8_, void*)
{
    const long long* film = (const long long*)film_;
    float* sum = (float*)sum_; float* mean = (float*)mean_; uint8_t* rgb8 = (uint8_t*)rgb8_;
    for (int64_t pix = 0; pix < (int64_t)width * height; ++pix) {
        const long long* px = film + pix * FILM_LANES;
        V3 c = v3(film_channel(px, 0), film_channel(px, 1), film_channel(px, 2));
        if (sum) { sum[3 * pix] = c.x; sum[3 * pix + 1] = c.y; sum[3 * pix + 2] = c.z; }
        c = c / (float)spp;
        if (mean) { mean[3 * pix] = c.x; mean[3 * pix + 1] = c.y; mean[3 * pix + 2] = c.z; }
        if (rgb8) tonemap_write(c, rgb8 + 3 * pix);
    }
    return 0;
}

// BVH introspection: the host builder's reference-topology nodes as (mn, mx) + (left, right)
int64_t hsim_mesh_bvh_nodes(hsim_scene* s, int shape, float* boxes, int32_t* children)
{
    const HostShape& sh = s->host.shapes[(size_t)shape];
    size_t n = sh.ref.nodes.size() / 2;
    if (boxes && children)
        for (size_t i = 0; i < n; ++i) {
            const float4& a = sh.ref.nodes[i * 2]; const float4& b = sh.ref.nodes[i * 2 + 1];
            boxes[i * 6] = a.x; boxes[i * 6 + 1] = a.y; boxes[i * 6 + 2] = a.z; boxes[i * 6 + 3] = b.x; boxes[i * 6 + 4] = b.y; boxes[i * 6 + 5] = b.z;
            std::memcpy(&children[i * 2], &a.w, 4); std::memcpy(&children[i * 2 + 1], &b.w, 4);
        }
    return (int64_t)n;
}
// the arrays hrt_scene_mesh_array returns, from the host builder (HRT_BVH_LBVH_HOST is the mirror of the device builder)
int64_t hsim_scene_mesh_array(hsim_scene* s, int mesh_index, int which, void* out, int64_t capacity)
{
    if (mesh_index < 0 || (size_t)mesh_index >= s->host.mesh_shapes.size() || which < 0 || which > 5) { g_err = "bad argument"; return HRT_ERR_ARG; }
    const HostShape& sh = s->host.shapes[(size_t)s->host.mesh_shapes[(size_t)mesh_index]];
    const std::vector<float4>* src[5] = {&sh.ref.nodes, &sh.fast.nodes, &sh.leaves, &sh.tris4, &sh.attrs4};
    const void* data = which == 5 ? (const void*)sh.wide.nodes.data() : (const void*)src[which]->data();
    const int64_t bytes = which == 5 ? (int64_t)(sh.wide.nodes.size() * sizeof(WNode)) : (int64_t)(src[which]->size() * sizeof(float4));
    if (out) { if (capacity < bytes) { g_err = "buffer too small"; return HRT_ERR_ARG; } if (bytes) std::memcpy(out, data, (size_t)bytes); }
    return bytes;
}
} // extern "C"
