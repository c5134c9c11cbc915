// ref_harness.cpp — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Drives the UNMODIFIED reference (Todegal/HobbyRaytracer, mounted read-only at /root/reference) so that its
// own hit()/rayColour()/render() can serve as the parity oracle and the CPU baseline.  Nothing from the
// reference is copied: this TU textually #includes hobbyraytracer/main.cpp from where it lies (with `main`
// renamed) to reach the file-static rayColour()/render(), and is linked with the reference's other TUs
// compiled in place by oracle/Makefile into oracle/_ref/libhrt_ref.so.
//
// Exposes the same scene-builder ABI as include/hrt.h with the prefix ref_.
// Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may load this library.
#include "hobbyraytracer.h"

#define main hobby_main
#include "main.cpp"   // hobbyraytracer/main.cpp (rayColour 38-79, render 81-140)
#undef main

#include <cstring>
#include <fstream>
#include <sstream>
#include <map>
#include <mutex>
#include <thread>
#include <atomic>

#include "../include/hrt.h"
#include "obj_reader.h"

// ---------------------------------------------------------------------------------------------------------
// rand() interposer.  The reference draws every random number from std::rand() through glm::linearRand
// (dependencies/glm/glm/gtc/random.inl:25).  glibc's rand() is one locked global stream; a definition inside
// this library (bound with -Bsymbolic-functions) gives each thread its own reproducible 31-bit stream:
//   state = state * 6364136223846793005 + 1442695040888963407;  return (state >> 33) & 0x7fffffff
// The oracle restatement (oracle/oracle.cpp) implements the identical stream, which is what makes
// bit-exact comparison of whole renders possible.
// ---------------------------------------------------------------------------------------------------------
static thread_local unsigned long long tl_rand_state = 0x853c49e6748fea9bULL;
static thread_local unsigned long long tl_rand_calls = 0;
extern "C" int rand(void) noexcept
{
    tl_rand_state = tl_rand_state * 6364136223846793005ULL + 1442695040888963407ULL;
    ++tl_rand_calls;
    return (int)((tl_rand_state >> 33) & 0x7fffffffULL);
}
static inline unsigned long long mix64(unsigned long long z)
{   // splitmix64 finaliser
    z += 0x9e3779b97f4a7c15ULL;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
    return z ^ (z >> 31);
}
static inline void seed_stream(unsigned long long seed, unsigned long long stream)
{
    tl_rand_state = mix64(seed ^ mix64(stream));
}

// ---------------------------------------------------------------------------------------------------------
// Mesh shim.  mesh.cpp needs Assimp, whose build requires cmake-generated headers (config.h, revision.h), so
// it is NOT compiled; these definitions give class Mesh (mesh.h, unmodified) the behaviour of mesh.cpp:13-51
// on top of a small OBJ reader that reproduces what Assimp's OBJ importer hands to mesh.cpp:53-120 with
// aiProcess_Triangulate | aiProcess_FlipUVs (unshared vertices per face corner, fan triangulation, v -> 1-v,
// zero normals / uvs when the file has none).
// ---------------------------------------------------------------------------------------------------------
Mesh::Mesh(std::string filepath, std::shared_ptr<Material> m)
{
    HittableList triangleStrip;
    objr::TriSoup soup;
    if (objr::load_obj(filepath, soup)) {
        for (size_t i = 0; i < soup.n_tri(); ++i) {
            const float* P = &soup.pos[i * 9]; const float* N = &soup.nrm[i * 9]; const float* T = &soup.uv[i * 6];
            triangleStrip.add(std::make_shared<ITriangle>(
                std::array<glm::vec3, 3>({ glm::vec3(P[0],P[1],P[2]), glm::vec3(P[3],P[4],P[5]), glm::vec3(P[6],P[7],P[8]) }),
                std::array<glm::vec3, 3>({ glm::vec3(N[0],N[1],N[2]), glm::vec3(N[3],N[4],N[5]), glm::vec3(N[6],N[7],N[8]) }),
                std::array<glm::vec2, 3>({ glm::vec2(T[0],T[1]), glm::vec2(T[2],T[3]), glm::vec2(T[4],T[5]) }), m));
        }
        std::cout << "Loaded mesh: " << filepath << std::endl;
    } else {
        std::cerr << "obj reader error: " << filepath << std::endl;
    }
    matPtr = m;
    tree = std::make_shared<BVHNode>(triangleStrip);
    std::cout << "Indexed file: " << filepath << std::endl;
}
bool Mesh::hit(const ray& r, float t_min, float t_max, hitRecord& rec) const { return tree->hit(r, t_min, t_max, rec); }
bool Mesh::boundingBox(AABB& outputBox) { return tree->boundingBox(outputBox); }

// ---------------------------------------------------------------------------------------------------------
namespace {

thread_local std::string g_err;

// A material that forwards to the real one and carries the primitive id (hitRecord has none, hittable.h:8-25).
struct TagMaterial : public Material {
    uint32_t id; std::shared_ptr<Material> inner;
    TagMaterial(uint32_t i, std::shared_ptr<Material> m) : id(i), inner(m) {}
    bool scatter(const ray& r_in, const hitRecord& rec, glm::vec3& att, ray& sc) const override { return inner->scatter(r_in, rec, att, sc); }
    glm::vec3 emitted(float u, float v, const glm::vec3& p) const override { return inner->emitted(u, v, p); }
};
// Tags whatever material a child reports (used for ConstantMedium, whose Isotropic is private).
struct TaggingWrapper : public Hittable {
    std::shared_ptr<Hittable> child; uint32_t id; mutable std::shared_ptr<TagMaterial> tag;
    TaggingWrapper(std::shared_ptr<Hittable> c, uint32_t i) : child(c), id(i) {}
    bool hit(const ray& r, float t_min, float t_max, hitRecord& rec) const override {
        if (!child->hit(r, t_min, t_max, rec)) return false;
        if (!tag) tag = std::make_shared<TagMaterial>(id, rec.matPtr);
        rec.matPtr = tag; return true;
    }
    bool boundingBox(AABB& b) override { return child->boundingBox(b); }
};
thread_local unsigned long long tl_ray_count = 0;
struct CountingWorld : public Hittable {
    std::shared_ptr<Hittable> w;
    explicit CountingWorld(std::shared_ptr<Hittable> x) : w(x) {}
    bool hit(const ray& r, float t_min, float t_max, hitRecord& rec) const override { ++tl_ray_count; return w->hit(r, t_min, t_max, rec); }
    bool boundingBox(AABB& b) override { return w->boundingBox(b); }
};

struct ShapeDesc {
    int kind;              // 0 sphere 1 rect 2 box 3 mesh 4 medium(pseudo)
    float f[8]; int axis; int material;
    std::vector<float> pos, nrm, uv; int64_t n_tri = 0;
    int64_t prim_base = 0, prim_count = 0;
};
struct ObjectDesc {
    int shape; std::vector<hrt_xform> ops; bool medium = false; float density = 0; float rgb[3] = {0,0,0}; int albedo_tex = -1;
    int pseudo_shape = -1;
};

} // namespace

struct ref_scene {
    bool tagged = true;
    std::vector<std::shared_ptr<Texture>> textures;
    std::vector<hrt_material> mat_descs;
    std::vector<std::shared_ptr<Material>> materials;
    std::vector<ShapeDesc> shapes;
    std::vector<ObjectDesc> objects;
    int64_t n_prims = 0;
    std::shared_ptr<Texture> background;
    std::shared_ptr<HittableList> world;
    bool committed = false;
    // YAML path
    std::unique_ptr<Scene> yaml_scene;
    Camera yaml_camera; std::shared_ptr<Film> yaml_film;
};

namespace {

std::shared_ptr<Material> make_material(ref_scene* s, const hrt_material& m)
{
    auto vec3of = [&](const float* c, int tex) -> MatVec3 { if (tex >= 0) return MatVec3(s->textures[tex]); return MatVec3(glm::vec3(c[0], c[1], c[2])); };
    auto scal = [&](float c, int tex) -> MatScalar { if (tex >= 0) return MatScalar(s->textures[tex]); return MatScalar(c); };
    switch (m.kind) {
    case HRT_MAT_LAMBERTIAN:    return std::make_shared<Lambertian>(vec3of(m.albedo, m.albedo_tex));
    case HRT_MAT_METAL:         return std::make_shared<Metal>(vec3of(m.albedo, m.albedo_tex), scal(m.p0, m.p0_tex));
    case HRT_MAT_DIELECTRIC:    return std::make_shared<Dielectric>(scal(m.p0, m.p0_tex), scal(m.p1, m.p1_tex));
    case HRT_MAT_ISOTROPIC:     if (m.albedo_tex >= 0) return std::make_shared<Isotropic>(s->textures[m.albedo_tex]);
                                return std::make_shared<Isotropic>(glm::vec3(m.albedo[0], m.albedo[1], m.albedo[2]));
    case HRT_MAT_DIFFUSE_LIGHT: return std::make_shared<DiffuseLight>(vec3of(m.albedo, m.albedo_tex), scal(m.p0, m.p0_tex));
    case HRT_MAT_PBR:           if (m.albedo_tex >= 0 && m.p1_tex >= 0) return std::make_shared<PBR>(s->textures[m.albedo_tex], s->textures[m.p1_tex], m.p0);
                                return std::make_shared<PBR>(glm::vec3(m.albedo[0], m.albedo[1], m.albedo[2]), m.p1, m.p0);
    case HRT_MAT_UVTEST:        return std::make_shared<UVTest>();
    }
    return nullptr;
}

std::shared_ptr<Material> mat_for(ref_scene* s, int material, uint32_t prim)
{
    if (s->tagged) return std::make_shared<TagMaterial>(prim, s->materials[material]);
    return s->materials[material];
}

std::shared_ptr<Hittable> build_shape(ref_scene* s, int shape_id, unsigned long long seed)
{
    ShapeDesc& sh = s->shapes[shape_id];
    uint32_t base = (uint32_t)sh.prim_base;
    switch (sh.kind) {
    case 0: return std::make_shared<Sphere>(glm::vec3(sh.f[0], sh.f[1], sh.f[2]), sh.f[3], mat_for(s, sh.material, base));
    case 1:
        if (sh.axis == 0) return std::make_shared<YZRect>(sh.f[0], sh.f[1], sh.f[2], sh.f[3], sh.f[4], mat_for(s, sh.material, base));
        if (sh.axis == 1) return std::make_shared<XZRect>(sh.f[0], sh.f[1], sh.f[2], sh.f[3], sh.f[4], mat_for(s, sh.material, base));
        return std::make_shared<XYRect>(sh.f[0], sh.f[1], sh.f[2], sh.f[3], sh.f[4], mat_for(s, sh.material, base));
    case 2: {
        glm::vec3 c(sh.f[0], sh.f[1], sh.f[2]), d(sh.f[3], sh.f[4], sh.f[5]);
        if (!s->tagged) return std::make_shared<Box>(c, d, s->materials[sh.material]);
        // tagged: the six sides of Box::constructBox (box.h:27-39), same expressions and order, one tag each.
        // Box::hit is `sides.hit(...)` (box.h:46-49), so a HittableList of the same rects is the same hit().
        glm::vec3 mn = c - d / 2.0f, mx = c + d / 2.0f;
        auto l = std::make_shared<HittableList>();
        l->add(std::make_shared<XYRect>(mn.x, mx.x, mn.y, mx.y, mx.z, mat_for(s, sh.material, base + 0)));
        l->add(std::make_shared<XYRect>(mn.x, mx.x, mn.y, mx.y, mn.z, mat_for(s, sh.material, base + 1)));
        l->add(std::make_shared<XZRect>(mn.x, mx.x, mn.z, mx.z, mx.y, mat_for(s, sh.material, base + 2)));
        l->add(std::make_shared<XZRect>(mn.x, mx.x, mn.z, mx.z, mn.y, mat_for(s, sh.material, base + 3)));
        l->add(std::make_shared<YZRect>(mn.y, mx.y, mn.z, mx.z, mx.x, mat_for(s, sh.material, base + 4)));
        l->add(std::make_shared<YZRect>(mn.y, mx.y, mn.z, mx.z, mn.x, mat_for(s, sh.material, base + 5)));
        return l;
    }
    case 3: {
        // mesh.cpp:13-41 without the file: one ITriangle per face in a HittableList, then BVHNode(list).
        HittableList strip;
        for (int64_t i = 0; i < sh.n_tri; ++i) {
            const float* P = &sh.pos[i * 9]; const float* N = &sh.nrm[i * 9]; const float* T = &sh.uv[i * 6];
            strip.add(std::make_shared<ITriangle>(
                std::array<glm::vec3, 3>({ glm::vec3(P[0],P[1],P[2]), glm::vec3(P[3],P[4],P[5]), glm::vec3(P[6],P[7],P[8]) }),
                std::array<glm::vec3, 3>({ glm::vec3(N[0],N[1],N[2]), glm::vec3(N[3],N[4],N[5]), glm::vec3(N[6],N[7],N[8]) }),
                std::array<glm::vec2, 3>({ glm::vec2(T[0],T[1]), glm::vec2(T[2],T[3]), glm::vec2(T[4],T[5]) }),
                mat_for(s, sh.material, base + (uint32_t)i)));
        }
        seed_stream(seed, (unsigned long long)shape_id);   // the split-axis stream of bvh.cpp:10
        return std::make_shared<BVHNode>(strip);
    }
    }
    return nullptr;
}

std::shared_ptr<Hittable> wrap(std::shared_ptr<Hittable> o, const std::vector<hrt_xform>& ops)
{
    for (const hrt_xform& x : ops) {
        switch (x.kind) {
        case HRT_XF_TRANSLATE:   o = std::make_shared<Translate>(o, glm::vec3(x.v[0], x.v[1], x.v[2])); break;
        case HRT_XF_SCALE:       o = std::make_shared<Scale>(o, glm::vec3(x.v[0], x.v[1], x.v[2])); break;
        case HRT_XF_ROTATE_QUAT: o = std::make_shared<RotateQuat>(o, glm::quat(x.v[0], x.v[1], x.v[2], x.v[3])); break;
        case HRT_XF_ROTATE_EULER_DEG: o = std::make_shared<RotateQuat>(o, glm::quat(glm::radians(glm::vec3(x.v[0], x.v[1], x.v[2])))); break; // scene.cpp:340
        case HRT_XF_ROTATE_Y_DEG: o = std::make_shared<RotateY>(o, x.v[0]); break;
        case HRT_XF_ROTATE_Y:    /* sin/cos form is not constructible on the reference; callers use *_DEG */ break;
        }
    }
    return o;
}

struct CoutSilencer {
    std::streambuf* old;
    CoutSilencer() { old = std::cout.rdbuf(nullptr); }
    ~CoutSilencer() { std::cout.rdbuf(old); std::cout.clear(); }
};

Camera make_camera(const hrt_camera* c)
{
    return Camera(glm::vec3(c->position[0], c->position[1], c->position[2]), glm::vec3(c->look_at[0], c->look_at[1], c->look_at[2]),
                  glm::vec3(c->up[0], c->up[1], c->up[2]), c->vfov_deg, c->aspect, c->aperture, c->focal_distance);
}

} // namespace

extern "C" {

const char* ref_last_error(void) { return g_err.c_str(); }

ref_scene* ref_scene_create(void) { return new ref_scene(); }
void ref_scene_destroy(ref_scene* s) { delete s; }
int ref_scene_set_tagged(ref_scene* s, int tagged) { s->tagged = tagged != 0; return 0; }

int ref_scene_add_texture_solid(ref_scene* s, const float rgb[3])
{ s->textures.push_back(std::make_shared<SolidColourTexture>(glm::vec3(rgb[0], rgb[1], rgb[2]))); return (int)s->textures.size() - 1; }
int ref_scene_add_texture_checker(ref_scene* s, int even_tex, int odd_tex)
{ s->textures.push_back(std::make_shared<CheckeredTexture>(s->textures[even_tex], s->textures[odd_tex])); return (int)s->textures.size() - 1; }

// ImageTexture / EnvironmentMap keep their pixels private and only load from files (texture.cpp:30-51,99-115):
// the harness writes the pixels to a temporary file in a format stb decodes losslessly (binary PPM for 8-bit,
// Radiance RGBE is lossy so float data goes through a raw-float PFM... stb has no PFM reader) -- instead the
// caller passes a *file path* for env maps; see ref_scene_add_texture_env_file.
int ref_scene_add_texture_image(ref_scene* s, const uint8_t* rgb8, int w, int h)
{
    char path[] = "/tmp/hrt_ref_img_XXXXXX";
    int fd = mkstemp(path); if (fd < 0) { g_err = "mkstemp failed"; return HRT_ERR_IO; }
    FILE* f = fdopen(fd, "wb");
    fprintf(f, "P6\n%d %d\n255\n", w, h); fwrite(rgb8, 1, (size_t)w * h * 3, f); fclose(f);
    { CoutSilencer q; s->textures.push_back(std::make_shared<ImageTexture>(std::string(path))); }
    remove(path);
    return (int)s->textures.size() - 1;
}
int ref_scene_add_texture_env_file(ref_scene* s, const char* path)
{
    CoutSilencer q; s->textures.push_back(std::make_shared<EnvironmentMap>(std::string(path)));
    return (int)s->textures.size() - 1;
}
int ref_scene_add_texture_image_file(ref_scene* s, const char* path)
{
    CoutSilencer q; s->textures.push_back(std::make_shared<ImageTexture>(std::string(path)));
    return (int)s->textures.size() - 1;
}

int ref_scene_add_material(ref_scene* s, const hrt_material* m)
{
    auto p = make_material(s, *m); if (!p) { g_err = "bad material kind"; return HRT_ERR_ARG; }
    s->mat_descs.push_back(*m); s->materials.push_back(p); return (int)s->materials.size() - 1;
}

static int push_shape(ref_scene* s, ShapeDesc&& d, int64_t count)
{ d.prim_base = s->n_prims; d.prim_count = count; s->n_prims += count; s->shapes.push_back(std::move(d)); return (int)s->shapes.size() - 1; }

int ref_scene_add_sphere(ref_scene* s, const float c[3], float radius, int material)
{ ShapeDesc d; d.kind = 0; d.f[0] = c[0]; d.f[1] = c[1]; d.f[2] = c[2]; d.f[3] = radius; d.material = material; return push_shape(s, std::move(d), 1); }
int ref_scene_add_rect(ref_scene* s, int axis, float a0, float a1, float b0, float b1, float k, int material)
{ ShapeDesc d; d.kind = 1; d.axis = axis; d.f[0] = a0; d.f[1] = a1; d.f[2] = b0; d.f[3] = b1; d.f[4] = k; d.material = material; return push_shape(s, std::move(d), 1); }
int ref_scene_add_box(ref_scene* s, const float c[3], const float dim[3], int material)
{ ShapeDesc d; d.kind = 2; for (int i = 0; i < 3; ++i) { d.f[i] = c[i]; d.f[3 + i] = dim[i]; } d.material = material; return push_shape(s, std::move(d), 6); }
int ref_scene_add_mesh(ref_scene* s, const float* pos, const float* nrm, const float* uv, int64_t n_tri, int material)
{
    ShapeDesc d; d.kind = 3; d.n_tri = n_tri; d.material = material;
    d.pos.assign(pos, pos + n_tri * 9);
    if (nrm) d.nrm.assign(nrm, nrm + n_tri * 9); else d.nrm.assign((size_t)n_tri * 9, 0.0f);
    if (uv) d.uv.assign(uv, uv + n_tri * 6); else d.uv.assign((size_t)n_tri * 6, 0.0f);
    return push_shape(s, std::move(d), n_tri);
}
int64_t ref_scene_prim_base(const ref_scene* s, int shape) { return s->shapes[shape].prim_base; }
int64_t ref_scene_prim_count(const ref_scene* s) { return s->n_prims; }

int ref_scene_add_object(ref_scene* s, int shape, const hrt_xform* ops, int n_ops)
{ ObjectDesc o; o.shape = shape; o.ops.assign(ops, ops + n_ops); s->objects.push_back(o); return (int)s->objects.size() - 1; }
int ref_scene_add_medium(ref_scene* s, int shape, const hrt_xform* ops, int n_ops, float density, const float rgb[3], int albedo_tex)
{
    ObjectDesc o; o.shape = shape; o.ops.assign(ops, ops + n_ops); o.medium = true; o.density = density;
    for (int i = 0; i < 3; ++i) o.rgb[i] = rgb ? rgb[i] : 0.0f; o.albedo_tex = albedo_tex;
    ShapeDesc d; d.kind = 4; d.material = -1; o.pseudo_shape = push_shape(s, std::move(d), 1);
    s->objects.push_back(o); return o.pseudo_shape;
}
int ref_scene_set_background(ref_scene* s, int tex) { s->background = s->textures[tex]; return 0; }

int ref_scene_commit(ref_scene* s, int /*bvh_mode*/, uint64_t seed)
{
    CoutSilencer q;
    s->world = std::make_shared<HittableList>();
    for (ObjectDesc& o : s->objects) {
        std::shared_ptr<Hittable> h = wrap(build_shape(s, o.shape, seed), o.ops);
        if (o.medium) {
            std::shared_ptr<Hittable> m;
            if (o.albedo_tex >= 0) m = std::make_shared<ConstantMedium>(h, o.density, s->textures[o.albedo_tex]);
            else m = std::make_shared<ConstantMedium>(h, o.density, glm::vec3(o.rgb[0], o.rgb[1], o.rgb[2]));
            if (s->tagged) m = std::make_shared<TaggingWrapper>(m, (uint32_t)s->shapes[o.pseudo_shape].prim_base);
            h = m;
        }
        s->world->add(h);
    }
    s->committed = true;
    return 0;
}

// world->hit(r, t_min, t_max, rec) per ray (main.cpp:45).  rng: the calling thread's rand stream is seeded per ray
// with (rng_seed, ray index) so media (constantMedium.cpp:25) are reproducible.
int ref_trace_batch(ref_scene* s, const float* rays, int64_t n, uint32_t* prim_id, float* t,
                    float* normal, float* uv, uint8_t* front, uint64_t rng_seed)
{
    if (!s->committed) { g_err = "scene not committed"; return HRT_ERR_STATE; }
    for (int64_t i = 0; i < n; ++i) {
        const float* r = rays + i * 8;
        seed_stream(rng_seed, (unsigned long long)i);
        hitRecord rec; rec.t = 0; rec.u = rec.v = 0; rec.frontFace = false; rec.normal = glm::vec3(0); rec.p = glm::vec3(0);
        bool h = s->world->hit(ray(glm::vec3(r[0], r[1], r[2]), glm::vec3(r[4], r[5], r[6])), r[3], r[7], rec);
        uint32_t id = HRT_MISS;
        if (h) {
            if (auto* tm = dynamic_cast<TagMaterial*>(rec.matPtr.get())) id = tm->id; else id = 0xFFFFFFFEu;
        }
        prim_id[i] = id; t[i] = h ? rec.t : 0.0f;
        if (normal) { normal[i*3] = h ? rec.normal.x : 0; normal[i*3+1] = h ? rec.normal.y : 0; normal[i*3+2] = h ? rec.normal.z : 0; }
        if (uv) { uv[i*2] = h ? rec.u : 0; uv[i*2+1] = h ? rec.v : 0; }
        if (front) front[i] = h ? (rec.frontFace ? 1 : 0) : 0;
    }
    return 0;
}

// The unmodified render() (main.cpp:81-140) on the calling thread's rand stream seeded with (seed, 0).
int ref_render_u8(ref_scene* s, const hrt_camera* cam, int w, int h, int spp, uint64_t seed, uint8_t* out_rgb8)
{
    if (!s->committed) { g_err = "scene not committed"; return HRT_ERR_STATE; }
    CoutSilencer q;
    Camera c = make_camera(cam);
    std::shared_ptr<Film> film = std::make_shared<Film>(w, h, spp, "unused.png");
    seed_stream(seed, 0);
    render(NUM_THREADS, s->background, s->world, c, film);
    std::copy_n(film->getPixels(), (size_t)w * h * 3, out_rgb8);
    return 0;
}

// The pixel loop of main.cpp:111-126 around the unmodified rayColour(), multi-threaded over rows with a
// per-row rand stream (seed, row+1) so the result does not depend on the thread count.  Writes per-pixel SUMS
// (pre-division, pre-tonemap) in the reference's buffer order (row 0 = top).  sample_first/sample_count select
// a sub-range of samples whose streams are (seed, row+1) advanced past the skipped samples by re-drawing.
int ref_render_f32(ref_scene* s, const hrt_camera* cam, int w, int h, int spp, uint64_t seed, int n_threads,
                   float* out_sum, uint64_t* rays_out)
{
    if (!s->committed) { g_err = "scene not committed"; return HRT_ERR_STATE; }
    Camera c = make_camera(cam);
    std::shared_ptr<Hittable> world = std::make_shared<CountingWorld>(s->world);
    std::shared_ptr<Texture> bg = s->background;
    std::atomic<int> next_row(0); std::atomic<unsigned long long> rays(0);
    if (n_threads <= 0) n_threads = (int)std::thread::hardware_concurrency();
    auto worker = [&]() {
        tl_ray_count = 0;
        for (;;) {
            int row = next_row.fetch_add(1); if (row >= h) break;
            seed_stream(seed, (unsigned long long)row + 1);
            for (int x = 0; x < w; ++x) {
                int pIdx = row * w + x;
                int px = pIdx % w; int py = h - pIdx / w;           // main.cpp:115-116
                glm::vec3 pixelColour(0.0f);
                for (int sidx = 0; sidx < spp; ++sidx) {
                    float u = ((float)px + glm::linearRand(0.0f, 1.0f)) / (w - 1);   // main.cpp:120
                    float v = ((float)py + glm::linearRand(0.0f, 1.0f)) / (h - 1);   // main.cpp:121
                    pixelColour += rayColour(c.getRay(u, v), bg, world);
                }
                out_sum[pIdx * 3 + 0] = pixelColour.r; out_sum[pIdx * 3 + 1] = pixelColour.g; out_sum[pIdx * 3 + 2] = pixelColour.b;
            }
        }
        rays += tl_ray_count;
    };
    std::vector<std::thread> th;
    for (int i = 0; i < n_threads; ++i) th.emplace_back(worker);
    for (auto& x : th) x.join();
    if (rays_out) *rays_out = rays.load();
    return 0;
}

// Film::tonemap + Film::writeColour (film.cpp:25-52) on a buffer of per-pixel sums.
int ref_film_resolve(const float* sum, int w, int h, int spp, uint8_t* out_rgb8)
{
    Film film(w, h, spp, "unused.png");
    std::vector<uint8_t> px((size_t)w * h * 3);
    for (int i = 0; i < w * h; ++i) {
        glm::vec3 c(sum[i * 3], sum[i * 3 + 1], sum[i * 3 + 2]);
        c /= static_cast<float>(spp);
        film.tonemap(c);
        Film::writeColour(c, px.begin() + (size_t)i * 3);
    }
    std::copy(px.begin(), px.end(), out_rgb8);
    return 0;
}

// Camera::getRay (camera.h:32-39): out = n x 6 (origin, direction) for n (s,t) pairs.
int ref_camera_rays(const hrt_camera* cam, const float* st, int64_t n, float* out)
{
    Camera c = make_camera(cam);
    for (int64_t i = 0; i < n; ++i) {
        ray r = c.getRay(st[i * 2], st[i * 2 + 1]);
        out[i*6] = r.o.x; out[i*6+1] = r.o.y; out[i*6+2] = r.o.z; out[i*6+3] = r.dir.x; out[i*6+4] = r.dir.y; out[i*6+5] = r.dir.z;
    }
    return 0;
}

// Raw draws of the reference's RNG helpers on the stream (seed, stream): kind 0 linearRand(0f,1f), 1 linearRand(0,2) int,
// 2 sphericalRand(1f) (3 floats each), 3 ballRand(1f) (3 floats each), 4 linearRand(0.0,1.0) double (returned as 2 floats hi/lo? no: as double array)
int ref_rng_draw(int kind, uint64_t seed, uint64_t stream, int64_t n, double* out)
{
    seed_stream(seed, stream);
    for (int64_t i = 0; i < n; ++i) {
        switch (kind) {
        case 0: out[i] = glm::linearRand(0.0f, 1.0f); break;
        case 1: out[i] = glm::linearRand(0, 2); break;
        case 2: { glm::vec3 v = glm::sphericalRand(1.0f); out[i*3] = v.x; out[i*3+1] = v.y; out[i*3+2] = v.z; } break;
        case 3: { glm::vec3 v = glm::ballRand(1.0f); out[i*3] = v.x; out[i*3+1] = v.y; out[i*3+2] = v.z; } break;
        case 4: out[i] = glm::linearRand(0.0, 1.0); break;
        default: return HRT_ERR_ARG;
        }
    }
    return 0;
}

// Material::scatter / emitted on explicit hit records (for per-material distribution tests).
// in: n x 14 floats: r_in.o(3) r_in.dir(3) p(3) normal(3) u v ; front: n bytes.  out: n x 10: ok, att(3), dir(3), emitted(3)
int ref_scatter_batch(ref_scene* s, int material, const float* in, const uint8_t* front, int64_t n, uint64_t seed, float* out)
{
    std::shared_ptr<Material> m = s->materials[material];
    for (int64_t i = 0; i < n; ++i) {
        const float* a = in + i * 14; seed_stream(seed, (unsigned long long)i);
        ray rin(glm::vec3(a[0], a[1], a[2]), glm::vec3(a[3], a[4], a[5]));
        hitRecord rec; rec.p = glm::vec3(a[6], a[7], a[8]); rec.normal = glm::vec3(a[9], a[10], a[11]); rec.u = a[12]; rec.v = a[13];
        rec.frontFace = front[i] != 0; rec.t = 1.0f;
        glm::vec3 att(0.0f); ray sc(glm::vec3(0.0f), glm::vec3(0.0f));
        glm::vec3 em = m->emitted(rec.u, rec.v, rec.p);
        bool ok = m->scatter(rin, rec, att, sc);
        float* o = out + i * 10;
        o[0] = ok ? 1.0f : 0.0f; o[1] = att.x; o[2] = att.y; o[3] = att.z; o[4] = sc.dir.x; o[5] = sc.dir.y; o[6] = sc.dir.z; o[7] = em.x; o[8] = em.y; o[9] = em.z;
    }
    return 0;
}

// Texture::colourValue on (u, v, p) tuples: in n x 5, out n x 3.
int ref_texture_eval(ref_scene* s, int tex, const float* in, int64_t n, float* out)
{
    for (int64_t i = 0; i < n; ++i) {
        glm::vec3 c = s->textures[tex]->colourValue(in[i*5], in[i*5+1], glm::vec3(in[i*5+2], in[i*5+3], in[i*5+4]));
        out[i*3] = c.x; out[i*3+1] = c.y; out[i*3+2] = c.z;
    }
    return 0;
}

// ---- the reference's own YAML loader (scene.cpp:127-374) ------------------------------------------------
// Loads `path` with Scene::loadScene (meshes through the Mesh shim above); the resulting world/camera/film are
// then used by ref_yaml_render_f32 exactly like main() does (main.cpp:153-162).
int ref_scene_load_yaml(ref_scene* s, const char* path, uint64_t seed, int* w, int* h, int* spp)
{
    CoutSilencer q;
    s->yaml_scene.reset(new Scene());
    seed_stream(seed, 0);
    int rc;
    {
        // relative asset paths in the YAML resolve against the process cwd in the reference; keep that.
        rc = s->yaml_scene->loadScene(std::string(path));
    }
    if (rc < 1) { g_err = "Scene::loadScene failed"; return HRT_ERR_IO; }
    s->yaml_camera = s->yaml_scene->getCamera();
    s->yaml_film = s->yaml_scene->getFilm();
    s->background = s->yaml_scene->getBackground();
    s->world = s->yaml_scene->getScene();
    s->committed = true; s->tagged = false;
    film_desc f = s->yaml_film->getFilm();
    if (w) *w = f.dimensions.x; if (h) *h = f.dimensions.y; if (spp) *spp = f.samples;
    return 0;
}
int ref_yaml_render_f32(ref_scene* s, int w, int h, int spp, uint64_t seed, int n_threads, float* out_sum, uint64_t* rays_out)
{
    if (!s->yaml_scene) { g_err = "no yaml scene"; return HRT_ERR_STATE; }
    Camera c = s->yaml_camera;
    std::shared_ptr<Hittable> world = std::make_shared<CountingWorld>(s->world);
    std::shared_ptr<Texture> bg = s->background;
    std::atomic<int> next_row(0); std::atomic<unsigned long long> rays(0);
    if (n_threads <= 0) n_threads = (int)std::thread::hardware_concurrency();
    auto worker = [&]() {
        tl_ray_count = 0;
        for (;;) {
            int row = next_row.fetch_add(1); if (row >= h) break;
            seed_stream(seed, (unsigned long long)row + 1);
            for (int x = 0; x < w; ++x) {
                int pIdx = row * w + x; int px = pIdx % w; int py = h - pIdx / w;
                glm::vec3 pixelColour(0.0f);
                for (int sidx = 0; sidx < spp; ++sidx) {
                    float u = ((float)px + glm::linearRand(0.0f, 1.0f)) / (w - 1);
                    float v = ((float)py + glm::linearRand(0.0f, 1.0f)) / (h - 1);
                    pixelColour += rayColour(c.getRay(u, v), bg, world);
                }
                out_sum[pIdx * 3 + 0] = pixelColour.r; out_sum[pIdx * 3 + 1] = pixelColour.g; out_sum[pIdx * 3 + 2] = pixelColour.b;
            }
        }
        rays += tl_ray_count;
    };
    std::vector<std::thread> th;
    for (int i = 0; i < n_threads; ++i) th.emplace_back(worker);
    for (auto& x : th) x.join();
    if (rays_out) *rays_out = rays.load();
    return 0;
}
int ref_yaml_trace_batch(ref_scene* s, const float* rays, int64_t n, float* t, float* normal, uint8_t* hit)
{
    for (int64_t i = 0; i < n; ++i) {
        const float* r = rays + i * 8;
        seed_stream(0, (unsigned long long)i);
        hitRecord rec; rec.t = 0; rec.normal = glm::vec3(0);
        bool h = s->world->hit(ray(glm::vec3(r[0], r[1], r[2]), glm::vec3(r[4], r[5], r[6])), r[3], r[7], rec);
        hit[i] = h; t[i] = h ? rec.t : 0.0f;
        if (normal) { normal[i*3] = h ? rec.normal.x : 0; normal[i*3+1] = h ? rec.normal.y : 0; normal[i*3+2] = h ? rec.normal.z : 0; }
    }
    return 0;
}

int ref_hardware_threads(void) { return (int)std::thread::hardware_concurrency(); }

} // extern "C"
