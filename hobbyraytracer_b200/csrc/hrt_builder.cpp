// hrt_builder.cpp — the scene-builder half of the C ABI (include/hrt.h) on a HostScene: one function per
// construction site of the reference's loader (scene.cpp:184-211 textures, 246-264 materials, 292-353 objects).
// Argument checking lives here so that the CUDA entry points and the CPU simulation harness share it.
#include <cstring>
#include <thread>

#include "hrt_host.h"

namespace hrt {

namespace {
int fail(std::string& err, const char* msg) { err = msg; return HRT_ERR_ARG; }
int push_shape(HostScene& s, HostShape&& d, int64_t count)
{
    d.prim_base = s.n_prims; d.prim_count = count; s.n_prims += count;
    s.shapes.push_back(std::move(d));
    return (int)s.shapes.size() - 1;
}
bool tex_ok(const HostScene& s, int t) { return t >= 0 && (size_t)t < s.textures.size(); }
} // namespace

int b_add_texture_solid(HostScene& s, const float rgb[3], std::string& err)
{
    if (!rgb) return fail(err, "rgb is NULL");
    HostTexture t; t.kind = TEX_SOLID; t.rgb[0] = rgb[0]; t.rgb[1] = rgb[1]; t.rgb[2] = rgb[2];
    s.textures.push_back(std::move(t)); return (int)s.textures.size() - 1;
}
int b_add_texture_checker(HostScene& s, int even_tex, int odd_tex, std::string& err)
{
    if (!tex_ok(s, even_tex) || !tex_ok(s, odd_tex)) return fail(err, "checker refers to an unknown texture");
    HostTexture t; t.kind = TEX_CHECKER; t.even = even_tex; t.odd = odd_tex;
    s.textures.push_back(std::move(t)); return (int)s.textures.size() - 1;
}
int b_add_texture_image(HostScene& s, const uint8_t* rgb8, int w, int h, std::string& err)
{
    if (w < 0 || h < 0) return fail(err, "negative image size");
    HostTexture t; t.kind = TEX_IMAGE; t.w = w; t.h = h; t.ch = 3;
    if (rgb8 && w > 0 && h > 0) t.u8.assign(rgb8, rgb8 + (size_t)w * h * 3);       // empty -> cyan (texture.cpp:55-56)
    s.textures.push_back(std::move(t)); return (int)s.textures.size() - 1;
}
int b_add_texture_env(HostScene& s, const float* data, int w, int h, int channels, std::string& err)
{
    if (w < 0 || h < 0) return fail(err, "negative image size");
    if (data && w > 0 && h > 0 && channels < 3) return fail(err, "environment maps need >= 3 channels");
    HostTexture t; t.kind = TEX_ENV; t.w = w; t.h = h; t.ch = channels;
    if (data && w > 0 && h > 0 && !try_pack_rgbe(data, w, h, channels, t.rgbe))   // lossless 4 B/texel form, else the floats
        t.f32.assign(data, data + (size_t)w * h * channels);
    s.textures.push_back(std::move(t)); return (int)s.textures.size() - 1;
}
int b_add_material(HostScene& s, const hrt_material* m, std::string& err)
{
    if (!m) return fail(err, "material is NULL");
    if (m->kind < HRT_MAT_LAMBERTIAN || m->kind > HRT_MAT_UVTEST) return fail(err, "unknown material kind");
    if ((m->albedo_tex >= 0 && !tex_ok(s, m->albedo_tex)) || (m->p0_tex >= 0 && !tex_ok(s, m->p0_tex)) || (m->p1_tex >= 0 && !tex_ok(s, m->p1_tex)))
        return fail(err, "material refers to an unknown texture");
    s.materials.push_back(*m); return (int)s.materials.size() - 1;
}
static bool mat_ok(const HostScene& s, int m) { return m >= 0 && (size_t)m < s.materials.size(); }

int b_add_sphere(HostScene& s, const float c[3], float radius, int material, std::string& err)
{
    if (!c) return fail(err, "center is NULL");
    if (!mat_ok(s, material)) return fail(err, "unknown material");
    HostShape d; d.kind = SHAPE_SPHERE; d.f[0] = c[0]; d.f[1] = c[1]; d.f[2] = c[2]; d.f[3] = radius; d.material = material;
    return push_shape(s, std::move(d), 1);
}
int b_add_rect(HostScene& s, int axis, float a0, float a1, float b0, float b1, float k, int material, std::string& err)
{
    if (axis < 0 || axis > 2) return fail(err, "rect axis must be 0, 1 or 2");
    if (!mat_ok(s, material)) return fail(err, "unknown material");
    HostShape d; d.kind = SHAPE_RECT; d.axis = axis; d.f[0] = a0; d.f[1] = a1; d.f[2] = b0; d.f[3] = b1; d.f[4] = k; d.material = material;
    return push_shape(s, std::move(d), 1);
}
int b_add_box(HostScene& s, const float c[3], const float dim[3], int material, std::string& err)
{
    if (!c || !dim) return fail(err, "box arguments are NULL");
    if (!mat_ok(s, material)) return fail(err, "unknown material");
    HostShape d; d.kind = SHAPE_BOX; for (int i = 0; i < 3; ++i) { d.f[i] = c[i]; d.f[3 + i] = dim[i]; } d.material = material;
    return push_shape(s, std::move(d), 6);
}
int b_add_mesh(HostScene& s, const float* pos, const float* nrm, const float* uv, int64_t n_tri, int material, std::string& err)
{
    if (n_tri < 0 || (n_tri > 0 && !pos)) return fail(err, "bad mesh arguments");
    if (n_tri > 0x7fffffff / 3) return fail(err, "mesh too large for 32-bit node ids");
    if (!mat_ok(s, material)) return fail(err, "unknown material");
    HostShape d; d.kind = SHAPE_MESH; d.n_tri = n_tri; d.material = material;
    d.pos.assign(pos, pos + n_tri * 9);
    if (nrm) d.nrm.assign(nrm, nrm + n_tri * 9); else d.nrm.assign((size_t)n_tri * 9, 0.0f);   // mesh.cpp:87-99
    if (uv) d.uv.assign(uv, uv + n_tri * 6); else d.uv.assign((size_t)n_tri * 6, 0.0f);
    return push_shape(s, std::move(d), n_tri);
}

// n items in at most hardware_concurrency() contiguous chunks of at least min_chunk, one thread each; chunk c covers
// [begin, end) and chunks are numbered in array order (so per-chunk results can be combined in a fixed order)
static size_t chunk_count(size_t n, size_t min_chunk)
{
    size_t nt = std::thread::hardware_concurrency(); if (nt < 1) nt = 1; if (nt > 32) nt = 32;
    size_t k = min_chunk ? n / min_chunk : nt; if (k < 1) k = 1; if (k > nt) k = nt;
    return k;
}
void parallel_chunks(size_t n, size_t min_chunk, const std::function<void(size_t, size_t, size_t)>& fn, size_t* n_chunks)
{
    const size_t k = chunk_count(n, min_chunk);
    if (n_chunks) *n_chunks = k;
    if (k == 1) { fn(0, 0, n); return; }
    std::vector<std::thread> th;
    for (size_t c = 0; c < k; ++c) th.emplace_back([&, c]() { fn(c, n * c / k, n * (c + 1) / k); });
    for (std::thread& t : th) t.join();
}

int b_add_mesh_indexed(HostScene& s, const float* vtx, const float* nrm, const float* uv, int64_t n_vert, const uint32_t* idx, int64_t n_idx,
                       const hrt_submesh* sub, int n_sub, int material, uint32_t flags, std::string& err)
{
    if (n_vert < 0 || n_idx < 0 || (n_vert > 0 && !vtx) || (n_idx > 0 && !idx)) return fail(err, "bad mesh arguments");
    if (flags & ~HRT_MESH_BORROW) return fail(err, "unknown mesh flags");
    if (n_idx % 3 != 0) return fail(err, "index count is not a multiple of 3");
    if (n_vert > 0xffffffffll) return fail(err, "more vertices than 32-bit indices can address");
    const int64_t n_tri = n_idx / 3;
    if (n_tri > 0x7fffffff / 3) return fail(err, "mesh too large for 32-bit node ids");
    if (n_sub < 0 || (n_sub > 0 && !sub)) return fail(err, "bad sub-mesh table");
    if (!mat_ok(s, material)) return fail(err, "unknown material");
    const bool borrow = (flags & HRT_MESH_BORROW) != 0;
    HostShape d; d.kind = SHAPE_MESH; d.n_tri = n_tri; d.material = material; d.indexed = true; d.n_vert = n_vert;
    if (sub) {                                                                  // the offset mesh.cpp:111-114 leaves out
        d.vidx.assign(idx, idx + n_idx);
        std::vector<uint8_t> seen((size_t)n_idx, 0);
        for (int m = 0; m < n_sub; ++m) {
            const hrt_submesh& q = sub[m];
            if (q.first_index < 0 || q.n_index < 0 || q.first_index > n_idx || q.n_index > n_idx - q.first_index || q.n_index % 3 != 0 || q.first_index % 3 != 0)
                return fail(err, "sub-mesh index range out of bounds");
            if (q.base_vertex < 0 || q.base_vertex > n_vert) return fail(err, "sub-mesh base vertex out of bounds");
            for (int64_t j = q.first_index; j < q.first_index + q.n_index; ++j) {
                if (seen[(size_t)j]) return fail(err, "sub-mesh index ranges overlap");
                seen[(size_t)j] = 1;
                const uint64_t v = (uint64_t)idx[j] + (uint64_t)q.base_vertex;
                if (v >= (uint64_t)n_vert) return fail(err, "vertex index out of range");
                d.vidx[(size_t)j] = (uint32_t)v;
            }
        }
        for (int64_t j = 0; j < n_idx; ++j) if (!seen[(size_t)j]) return fail(err, "sub-mesh table does not cover every index");
    } else {
        std::vector<uint8_t> bad(64, 0);
        if (!borrow) d.vidx.resize((size_t)n_idx);
        uint32_t* copy = borrow ? nullptr : d.vidx.data();
        parallel_chunks((size_t)n_idx, 1 << 20, [&](size_t c, size_t b, size_t e) {      // range check (+ the copy) in one pass
            uint32_t worst = 0;
            if (copy) for (size_t j = b; j < e; ++j) { const uint32_t q = idx[j]; copy[j] = q; worst = q > worst ? q : worst; }
            else for (size_t j = b; j < e; ++j) { const uint32_t q = idx[j]; worst = q > worst ? q : worst; }
            if ((int64_t)worst >= n_vert && e > b) bad[c] = 1;
        });
        for (uint8_t b : bad) if (b) return fail(err, "vertex index out of range");
    }
    if (d.vidx.empty()) d.pi = idx;                                             // borrowed (or no indices at all)
    auto take = [&](std::vector<float>& own, const float*& view, const float* src, size_t count) {
        if (src && borrow) { view = src; return; }
        if (src) {
            own.resize(count);
            float* dst = own.data();
            parallel_chunks(count, 1 << 22, [&](size_t, size_t b, size_t e) { std::memcpy(dst + b, src + b, (e - b) * sizeof(float)); });
        } else own.assign(count, 0.0f);                                         // mesh.cpp:87-99: zeros when the file has none
    };
    take(d.vpos, d.pv, vtx, (size_t)n_vert * 3);
    take(d.vnrm, d.pn, nrm, (size_t)n_vert * 3);
    take(d.vuv, d.pt, uv, (size_t)n_vert * 2);
    return push_shape(s, std::move(d), n_tri);
}

void ensure_soup(HostShape& sh)
{
    if (!sh.indexed || sh.pos.size() == (size_t)sh.n_tri * 9) return;
    const size_t n = (size_t)sh.n_tri;
    sh.pos.resize(n * 9); sh.nrm.resize(n * 9); sh.uv.resize(n * 6);
    const float *V = sh.V(), *N = sh.N(), *T = sh.T(); const uint32_t* I = sh.I();
    parallel_chunks(n * 3, 1 << 18, [&](size_t, size_t b, size_t e) {
        for (size_t c = b; c < e; ++c) {                                        // corner c = triangle c/3, vertex c%3 (mesh.cpp:26-29)
            const size_t v = I[c];
            for (int a = 0; a < 3; ++a) { sh.pos[c * 3 + a] = V[v * 3 + a]; sh.nrm[c * 3 + a] = N[v * 3 + a]; }
            sh.uv[c * 2] = T[v * 2]; sh.uv[c * 2 + 1] = T[v * 2 + 1];
        }
    });
}

static int fill_ops(HostObject& o, const hrt_xform* ops, int n_ops, std::string& err)
{
    if (n_ops < 0 || n_ops > MAX_XF || (n_ops > 0 && !ops)) return fail(err, "bad wrapper chain (at most 8 wrappers)");
    for (int i = 0; i < n_ops; ++i) {
        if (ops[i].kind < HRT_XF_TRANSLATE || ops[i].kind > HRT_XF_ROTATE_Y_DEG) return fail(err, "unknown wrapper kind");
        o.ops.push_back(make_xform(ops[i]));
    }
    return 0;
}
static bool shape_ok(const HostScene& s, int sh) { return sh >= 0 && (size_t)sh < s.shapes.size() && s.shapes[(size_t)sh].kind != SHAPE_MEDIUM; }

int b_add_object(HostScene& s, int shape, const hrt_xform* ops, int n_ops, std::string& err)
{
    if (!shape_ok(s, shape)) return fail(err, "unknown shape");
    HostObject o; o.shape = shape;
    int rc = fill_ops(o, ops, n_ops, err); if (rc < 0) return rc;
    s.objects.push_back(std::move(o)); return (int)s.objects.size() - 1;
}
int b_add_medium(HostScene& s, int shape, const hrt_xform* ops, int n_ops, float density, const float rgb[3], int albedo_tex, std::string& err)
{
    if (!shape_ok(s, shape)) return fail(err, "unknown shape");
    if (albedo_tex >= 0 && !tex_ok(s, albedo_tex)) return fail(err, "unknown texture");
    HostObject o; o.shape = shape;
    int rc = fill_ops(o, ops, n_ops, err); if (rc < 0) return rc;
    o.medium = true; o.neg_inv_density = -1 / density;                        // constantMedium.h:11
    o.phase_tex = albedo_tex;
    for (int a = 0; a < 3; ++a) o.phase_rgb[a] = rgb ? rgb[a] : 0.0f;
    HostShape d; d.kind = SHAPE_MEDIUM; d.material = -1;
    o.pseudo_shape = push_shape(s, std::move(d), 1);
    int id = o.pseudo_shape;
    s.objects.push_back(std::move(o));
    return id;
}
int b_set_background(HostScene& s, int tex, std::string& err)
{
    if (!tex_ok(s, tex)) return fail(err, "unknown texture");
    s.background = tex; return 0;
}

} // namespace hrt
