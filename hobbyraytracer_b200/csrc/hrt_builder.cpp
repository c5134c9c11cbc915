// hrt_builder.cpp — the scene-builder half of the C ABI (include/hrt.h) on a HostScene: one function per
// construction site of the reference's loader (scene.cpp:184-211 textures, 246-264 materials, 292-353 objects).
// Argument checking lives here so that the CUDA entry points and the CPU simulation harness share it.
#include <cstring>

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
