// host/hrt_host.cpp — HobbyRaytracer's own host in front of libhrt_b200.so.
//
// This is the drop-in of INTEGRATION.md as a program: the reference's scene loader (hobbyraytracer/scene.cpp + yaml-cpp),
// its Assimp import (mesh.cpp), stb decoding (texture.cpp), camera parameters and Film::outputFilm (film.cpp) run UNMODIFIED —
// they are compiled from /root/reference where they lie (host/Makefile) — and the one thing that changes is main.cpp:176:
// instead of render(NUM_THREADS, background, world, camera, film) the loaded scene is handed to the CUDA library through the
// C ABI of include/hrt.h and hrt_render() fills the film's pixel vector.
//
// How the scene crosses the boundary.  Every field of the reference's scene classes is private (hittable.h, material.h,
// texture.h ...), so INTEGRATION.md describes the binding as one hrt_scene_add_* call beside each std::make_shared<X>(...)
// of scene.cpp.  This translation unit gets the same information without editing scene.cpp: it includes the reference's
// class definitions with `private` spelled `public` (layout is unchanged) and walks the finished object graph — the world
// list, instance wrappers, meshes (their ITriangles in left-to-right tree order), materials, MatVec3 / MatScalar variants
// and textures with their decoded pixels — replaying each object's constructor arguments into the builder.
//
// SURVEY.md §8(f)1, loader extension in the retained surface: after Scene::loadScene() the same YAML file is read once more
// (yaml-cpp) for what the shipped schema cannot express but the reference's classes can: materials `dielectric` (ior,
// roughness), `isotropic`, `pbr` (roughness, metalness), `uv_test`; objects `box` (center, dimensions); `rotate_y: deg` in
// transform (innermost, like RTIOW); `medium: {density, albedo}` (ConstantMedium around the transformed object).  Files that
// use none of these go through the reference's loader alone.
//
//   hobbyraytracer_b200 scene.yaml [--gpus N] [--seed S] [--bvh sah|lbvh_gpu|reference] [--spp N] [--size W H] [--sum out.f32] [--quiet]
//
// Exit code: the reference's (Film::outputFilm's stb result, 1 = success; -1 on a load / CUDA failure), main.cpp:155,194.
#include <algorithm>
#include <cassert>
#include <cstdio>
#include <cstring>
#include <execution>
#include <map>
#include <numeric>
#include <ranges>
#include <unordered_map>
#include <variant>

#include <glm/glm.hpp>
#include <glm/common.hpp>
#include <glm/geometric.hpp>
#include <glm/vec3.hpp>
#include <glm/mat3x3.hpp>
#include <glm/gtx/rotate_vector.hpp>
#include <glm/gtx/quaternion.hpp>
#include <glm/gtc/random.hpp>
#include <glm/gtc/epsilon.hpp>
#include <glm/gtx/common.hpp>
#include <glm/gtx/string_cast.hpp>
#include <yaml-cpp/yaml.h>

#include <iostream>
#include <iomanip>
#include <vector>
#include <array>
#include <memory>
#include <random>
#include <thread>
#include <future>
#include <atomic>
#include <chrono>
#include <string>
#include <filesystem>

// the reference's classes, members readable
#define private public
#define protected public
#include "hobbyraytracer.h"
#include "ray.h"
#include "aabb.h"
#include "sphere.h"
#include "hittableList.h"
#include "camera.h"
#include "texture.h"
#include "material.h"
#include "bvh.h"
#include "aarect.h"
#include "box.h"                 // defines Box's members: included by exactly one translation unit (the reference: main.cpp:14)
#include "translate.h"
#include "rotateY.h"
#include "rotateQuat.h"
#include "scale.h"
#include "constantMedium.h"
#include "triangle.h"
#include "mesh.h"
#include "film.h"
#include "scene.h"
#undef private
#undef protected

#include "../include/hrt.h"

namespace {

constexpr int MAX_DEPTH = 50;      // main.cpp:32

[[noreturn]] void die(const std::string& what)
{
    std::cout << what << ": " << hrt_last_error() << std::endl;
    std::exit(-1);                 // main.cpp:155-156: a failed load returns -1
}
int ck(int rc, const char* what) { if (rc < 0) die(what); return rc; }

// ---- the object graph, replayed into the builder --------------------------------------------------------------------------------
struct Flattener {
    hrt_scene* gpu;
    std::map<const Texture*, int> tex_ids;
    std::map<const Material*, int> mat_ids;
    std::map<const Hittable*, int> shape_ids;       // a shape listed twice (instanced) is added once
    // constructor arguments the objects no longer hold (Box keeps min / max, ConstantMedium -1 / density): recorded by the
    // extension loader below, the only place that constructs them, so that the builder gets the very same floats
    std::map<const Hittable*, std::array<float, 6>> box_args;
    std::map<const Hittable*, float> medium_density;

    int texture(const std::shared_ptr<Texture>& t)
    {
        auto it = tex_ids.find(t.get());
        if (it != tex_ids.end()) return it->second;
        int id;
        if (auto* s = dynamic_cast<SolidColourTexture*>(t.get())) id = ck(hrt_scene_add_texture_solid(gpu, &s->c.x), "hrt_scene_add_texture_solid");
        else if (auto* c = dynamic_cast<CheckeredTexture*>(t.get())) { const int e = texture(c->e), o = texture(c->o); id = ck(hrt_scene_add_texture_checker(gpu, e, o), "hrt_scene_add_texture_checker"); }
        else if (auto* i = dynamic_cast<ImageTexture*>(t.get())) id = ck(hrt_scene_add_texture_image(gpu, i->data.empty() ? nullptr : i->data.data(), i->data.empty() ? 0 : i->width, i->data.empty() ? 0 : i->height), "hrt_scene_add_texture_image");
        else if (auto* e = dynamic_cast<EnvironmentMap*>(t.get())) id = ck(hrt_scene_add_texture_env(gpu, e->data.empty() ? nullptr : e->data.data(), e->data.empty() ? 0 : e->width, e->data.empty() ? 0 : e->height, e->channels), "hrt_scene_add_texture_env");
        else { std::cout << "unknown Texture subclass" << std::endl; std::exit(-1); }
        tex_ids[t.get()] = id;
        return id;
    }
    void vec3_of(const MatVec3& v, float rgb[3], int32_t& tex)                   // material.h:10-35
    {
        rgb[0] = rgb[1] = rgb[2] = 0.0f; tex = -1;
        if (auto* c = std::get_if<glm::vec3>(&v.variant)) { rgb[0] = c->x; rgb[1] = c->y; rgb[2] = c->z; }
        else tex = texture(std::get<std::shared_ptr<Texture>>(v.variant));
    }
    void scalar_of(const MatScalar& v, float& f, int32_t& tex)                   // material.h:36-58
    {
        f = 0.0f; tex = -1;
        if (auto* c = std::get_if<float>(&v.variant)) f = *c;
        else tex = texture(std::get<std::shared_ptr<Texture>>(v.variant));
    }
    int material(const std::shared_ptr<Material>& m)
    {
        auto it = mat_ids.find(m.get());
        if (it != mat_ids.end()) return it->second;
        hrt_material d; std::memset(&d, 0, sizeof(d)); d.albedo_tex = d.p0_tex = d.p1_tex = -1;
        if (auto* l = dynamic_cast<Lambertian*>(m.get())) { d.kind = HRT_MAT_LAMBERTIAN; vec3_of(l->albedo, d.albedo, d.albedo_tex); }
        else if (auto* x = dynamic_cast<Metal*>(m.get())) { d.kind = HRT_MAT_METAL; vec3_of(x->albedo, d.albedo, d.albedo_tex); scalar_of(x->r, d.p0, d.p0_tex); }
        else if (auto* g = dynamic_cast<Dielectric*>(m.get())) { d.kind = HRT_MAT_DIELECTRIC; scalar_of(g->ir, d.p0, d.p0_tex); scalar_of(g->r, d.p1, d.p1_tex); }
        else if (auto* e = dynamic_cast<DiffuseLight*>(m.get())) { d.kind = HRT_MAT_DIFFUSE_LIGHT; vec3_of(e->emit, d.albedo, d.albedo_tex); scalar_of(e->s, d.p0, d.p0_tex); }
        else if (auto* i = dynamic_cast<Isotropic*>(m.get())) { d.kind = HRT_MAT_ISOTROPIC; d.albedo_tex = texture(i->albedo); }
        else if (auto* p = dynamic_cast<PBR*>(m.get())) {                          // material.cpp:4-28: Metal + Lambertian of one albedo, chosen by the mix texture
            d.kind = HRT_MAT_PBR; vec3_of(p->metal->albedo, d.albedo, d.albedo_tex); scalar_of(p->metal->r, d.p0, d.p0_tex); d.p1_tex = texture(p->mix);
        }
        else if (dynamic_cast<UVTest*>(m.get())) d.kind = HRT_MAT_UVTEST;
        else { std::cout << "unknown Material subclass" << std::endl; std::exit(-1); }
        const int id = ck(hrt_scene_add_material(gpu, &d), "hrt_scene_add_material");
        mat_ids[m.get()] = id;
        return id;
    }

    // Mesh (mesh.cpp:13-41): its ITriangles, collected from the tree left to right (bvh.cpp:6-61 lists a lone triangle as both children)
    void triangles_of(const std::shared_ptr<Hittable>& h, std::vector<float>& pos, std::vector<float>& nrm, std::vector<float>& uv, std::shared_ptr<Material>& mat)
    {
        if (auto* n = dynamic_cast<BVHNode*>(h.get())) {
            if (n->left) triangles_of(n->left, pos, nrm, uv, mat);
            if (n->right && n->right != n->left) triangles_of(n->right, pos, nrm, uv, mat);
        } else if (auto* t = dynamic_cast<ITriangle*>(h.get())) {
            for (int v = 0; v < 3; ++v) {
                pos.insert(pos.end(), {t->vertices[v].x, t->vertices[v].y, t->vertices[v].z});
                nrm.insert(nrm.end(), {t->normals[v].x, t->normals[v].y, t->normals[v].z});
                uv.insert(uv.end(), {t->uvs[v].x, t->uvs[v].y});
            }
            mat = t->matPtr;
        }
    }
    int shape(const std::shared_ptr<Hittable>& h)
    {
        auto it = shape_ids.find(h.get());
        if (it != shape_ids.end()) return it->second;
        int id;
        if (auto* s = dynamic_cast<Sphere*>(h.get())) id = ck(hrt_scene_add_sphere(gpu, &s->center.x, s->radius, material(s->matPtr)), "hrt_scene_add_sphere");
        else if (auto* r = dynamic_cast<YZRect*>(h.get())) id = ck(hrt_scene_add_rect(gpu, 0, r->y0, r->y1, r->z0, r->z1, r->k, material(r->mp)), "hrt_scene_add_rect");
        else if (auto* r = dynamic_cast<XZRect*>(h.get())) id = ck(hrt_scene_add_rect(gpu, 1, r->x0, r->x1, r->z0, r->z1, r->k, material(r->mp)), "hrt_scene_add_rect");
        else if (auto* r = dynamic_cast<XYRect*>(h.get())) id = ck(hrt_scene_add_rect(gpu, 2, r->x0, r->x1, r->y0, r->y1, r->k, material(r->mp)), "hrt_scene_add_rect");
        else if (auto* b = dynamic_cast<Box*>(h.get())) {                          // box.h:41-44: centre and dimensions back from min / max
            glm::vec3 c = (b->boxMin + b->boxMax) / glm::vec3(2.0f), d = b->boxMax - b->boxMin;
            if (auto a = box_args.find(h.get()); a != box_args.end()) { c = glm::vec3(a->second[0], a->second[1], a->second[2]); d = glm::vec3(a->second[3], a->second[4], a->second[5]); }
            auto* side = dynamic_cast<XYRect*>(b->sides.objects.at(0).get());
            id = ck(hrt_scene_add_box(gpu, &c.x, &d.x, material(side->mp)), "hrt_scene_add_box");
        }
        else if (auto* m = dynamic_cast<Mesh*>(h.get())) {
            std::vector<float> pos, nrm, uv; std::shared_ptr<Material> mat = m->matPtr;
            if (m->tree) triangles_of(m->tree, pos, nrm, uv, mat);
            if (!mat) { std::cout << "mesh without a material" << std::endl; std::exit(-1); }
            id = ck(hrt_scene_add_mesh(gpu, pos.data(), nrm.data(), uv.data(), (int64_t)(pos.size() / 9), material(mat)), "hrt_scene_add_mesh");
        }
        else { std::cout << "unknown Hittable subclass in the world list" << std::endl; std::exit(-1); }
        shape_ids[h.get()] = id;
        return id;
    }
    // one world-list entry (scene.cpp:356): wrappers peeled outermost first, handed over innermost first
    void object(std::shared_ptr<Hittable> h)
    {
        const ConstantMedium* medium = dynamic_cast<ConstantMedium*>(h.get());
        if (medium) h = medium->boundary;
        std::vector<hrt_xform> ops;
        for (;;) {
            if (auto* t = dynamic_cast<Translate*>(h.get())) { ops.push_back({HRT_XF_TRANSLATE, {t->offset.x, t->offset.y, t->offset.z, 0}}); h = t->ptr; }
            else if (auto* s = dynamic_cast<Scale*>(h.get())) { ops.push_back({HRT_XF_SCALE, {s->factor.x, s->factor.y, s->factor.z, 0}}); h = s->ptr; }
            else if (auto* q = dynamic_cast<RotateQuat*>(h.get())) { ops.push_back({HRT_XF_ROTATE_QUAT, {q->rotation.w, q->rotation.x, q->rotation.y, q->rotation.z}}); h = q->ptr; }
            else if (auto* y = dynamic_cast<RotateY*>(h.get())) { ops.push_back({HRT_XF_ROTATE_Y, {y->sinTheta, y->cosTheta, 0, 0}}); h = y->ptr; }
            else break;
        }
        std::reverse(ops.begin(), ops.end());
        const int sh = shape(h);
        if (!medium) { ck(hrt_scene_add_object(gpu, sh, ops.data(), (int)ops.size()), "hrt_scene_add_object"); return; }
        auto* iso = dynamic_cast<Isotropic*>(medium->phaseFunction.get());        // constantMedium.h:10-14
        float density = -1.0f / medium->negInvDensity;
        if (auto a = medium_density.find(medium); a != medium_density.end()) density = a->second;
        const float none[3] = {0, 0, 0};
        ck(hrt_scene_add_medium(gpu, sh, ops.data(), (int)ops.size(), density, none, texture(iso->albedo)), "hrt_scene_add_medium");
    }
};

// ---- what the shipped schema cannot say ---------------------------------------------------------------------------
// Runs after Scene::loadScene (scene.cpp:127-374), on the same file.  Materials of the new types are added to the scene's
// material map; when the file uses any extension at all, the world list is rebuilt from the objects node in file order with
// the superset schema (the reference skipped the objects whose material it did not know and stored nullptr for unknown types,
// scene.cpp:279-290,356).  Error convention of the reference: message on stdout, -1.
struct ExtLoader {
    Scene& scene; Flattener& flat;
    static glm::vec3 vec3(const YAML::Node& n, const char* key)
    {
        if (!n[key] || !n[key].IsSequence() || n[key].size() != 3) throw YAML::ParserException(n.Mark(), std::string("Invalid value for vector 3: ") + key);
        return glm::vec3(n[key][0].as<float>(), n[key][1].as<float>(), n[key][2].as<float>());
    }
    static glm::vec2 vec2(const YAML::Node& n, const char* key)
    {
        if (!n[key] || !n[key].IsSequence() || n[key].size() != 2) throw YAML::ParserException(n.Mark(), std::string("Invalid value for vector 2: ") + key);
        return glm::vec2(n[key][0].as<float>(), n[key][1].as<float>());
    }
    template <class T> static T req(const YAML::Node& n, const char* key)
    {
        if (!n[key]) throw YAML::ParserException(n.Mark(), std::string("Could not find required property: ") + key);
        return n[key].as<T>();
    }
    std::shared_ptr<Texture> texture_named(const std::string& name)           // scene.cpp:83-93: a known texture, else an image file
    {
        if (scene.textures.count(name) == 1) return scene.textures[name];
        scene.textures[name] = std::make_shared<ImageTexture>(name);
        return scene.textures[name];
    }
    MatVec3 mat_vec3(const YAML::Node& n, const char* key)                      // scene.cpp:72-98
    {
        if (!n[key]) throw YAML::ParserException(n.Mark(), std::string("Could not find required property: ") + key);
        if (n[key].IsSequence()) return MatVec3(vec3(n, key));
        return MatVec3(texture_named(n[key].as<std::string>()));
    }
    MatScalar mat_scalar(const YAML::Node& n, const char* key, float fallback, bool required)   // scene.cpp:100-125
    {
        if (!n[key]) { if (required) throw YAML::ParserException(n.Mark(), std::string("Could not find required property: ") + key); return MatScalar(fallback); }
        try { return MatScalar(n[key].as<float>()); } catch (const std::exception&) { }
        return MatScalar(texture_named(n[key].as<std::string>()));
    }
    static bool is_ext_material(const std::string& t) { return t == "dielectric" || t == "isotropic" || t == "pbr" || t == "uv_test"; }
    static bool uses_extensions(const YAML::Node& root)
    {
        if (root["materials"]) for (auto m : root["materials"]) if (m["type"] && is_ext_material(m["type"].as<std::string>())) return true;
        if (root["objects"] && root["objects"].IsSequence())
            for (auto o : root["objects"]) {
                if (o["type"] && o["type"].as<std::string>() == "box") return true;
                if (o["medium"]) return true;
                if (o["transform"] && o["transform"]["rotate_y"]) return true;
            }
        return false;
    }
    int run(const std::string& path)
    {
        try {
            YAML::Node root = YAML::LoadFile(path);
            if (!uses_extensions(root)) return 1;
            if (root["materials"])
                for (auto m : root["materials"]) {
                    const std::string name = req<std::string>(m, "name"), type = req<std::string>(m, "type");
                    if (type == "dielectric") scene.materials[name] = std::make_shared<Dielectric>(mat_scalar(m, "ior", 1.5f, true), mat_scalar(m, "roughness", 0.0f, false));   // material.h:202
                    else if (type == "isotropic") {                                                                                                                         // material.h:76-77
                        if (m["albedo"] && m["albedo"].IsSequence()) scene.materials[name] = std::make_shared<Isotropic>(vec3(m, "albedo"));
                        else scene.materials[name] = std::make_shared<Isotropic>(texture_named(req<std::string>(m, "albedo")));
                    }
                    else if (type == "pbr") {                                                                                                                               // material.cpp:4-16
                        const float rough = req<float>(m, "roughness");
                        if (m["albedo"] && m["albedo"].IsSequence()) scene.materials[name] = std::make_shared<PBR>(vec3(m, "albedo"), req<float>(m, "metalness"), rough);
                        else {
                            std::shared_ptr<Texture> metal;
                            try { metal = std::make_shared<SolidColourTexture>(m["metalness"].as<float>()); } catch (const std::exception&) { metal = texture_named(req<std::string>(m, "metalness")); }
                            scene.materials[name] = std::make_shared<PBR>(texture_named(req<std::string>(m, "albedo")), metal, rough);
                        }
                    }
                    else if (type == "uv_test") scene.materials[name] = std::make_shared<UVTest>();
                }
            scene.objects.clear();
            if (root["objects"] && root["objects"].IsSequence())
                for (auto object : root["objects"]) {
                    const std::string key = req<std::string>(object, "material");
                    if (scene.materials.count(key) != 1) { std::cout << "Material " << key << " does not exist!" << std::endl; continue; }    // scene.cpp:281-290
                    std::shared_ptr<Material> m = scene.materials[key];
                    const std::string type = req<std::string>(object, "type");
                    std::shared_ptr<Hittable> o;
                    if (type == "mesh") o = std::make_shared<Mesh>(req<std::string>(object, "path"), m);                                      // scene.cpp:292-332
                    else if (type == "sphere") o = std::make_shared<Sphere>(vec3(object, "center"), req<float>(object, "radius"), m);
                    else if (type == "yz_rect") { glm::vec2 Y = vec2(object, "y"), Z = vec2(object, "z"); o = std::make_shared<YZRect>(Y.x, Y.y, Z.x, Z.y, req<float>(object, "k"), m); }
                    else if (type == "xz_rect") { glm::vec2 X = vec2(object, "x"), Z = vec2(object, "z"); o = std::make_shared<XZRect>(X.x, X.y, Z.x, Z.y, req<float>(object, "k"), m); }
                    else if (type == "xy_rect") { glm::vec2 X = vec2(object, "x"), Y = vec2(object, "y"); o = std::make_shared<XYRect>(X.x, X.y, Y.x, Y.y, req<float>(object, "k"), m); }
                    else if (type == "box") {                                                                                                // box.h:9
                        const glm::vec3 c = vec3(object, "center"), d = vec3(object, "dimensions");
                        o = std::make_shared<Box>(c, d, m);
                        flat.box_args[o.get()] = {c.x, c.y, c.z, d.x, d.y, d.z};
                    }
                    else throw YAML::ParserException(object.Mark(), "Unknown object type: " + type);
                    if (YAML::Node tr = object["transform"]) {                                                                               // scene.cpp:335-353, rotate_y first
                        if (tr["rotate_y"]) o = std::make_shared<RotateY>(o, tr["rotate_y"].as<float>());                                     // rotateY.h:8
                        if (tr["rotate"]) o = std::make_shared<RotateQuat>(o, glm::quat(glm::radians(vec3(tr, "rotate"))));
                        if (tr["scale"]) o = std::make_shared<Scale>(o, vec3(tr, "scale"));
                        if (tr["translate"]) o = std::make_shared<Translate>(o, vec3(tr, "translate"));
                    }
                    if (YAML::Node med = object["medium"]) {                                                                                 // constantMedium.h:13-14
                        const float density = req<float>(med, "density");
                        o = std::make_shared<ConstantMedium>(o, density, vec3(med, "albedo"));
                        flat.medium_density[o.get()] = density;
                    }
                    scene.objects.add(o);
                }
        } catch (const YAML::Exception& ex) {
            std::cout << ex.what() << std::endl;
            return -1;
        }
        return 1;
    }
};

void print_elapsed(const char* what, const std::string& name, std::chrono::high_resolution_clock::time_point start)      // main.cpp:164-171
{
    auto eMS = std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::high_resolution_clock::now() - start);
    unsigned int iH = (unsigned int)std::chrono::duration_cast<std::chrono::hours>(eMS).count();
    unsigned int iM = (unsigned int)std::chrono::duration_cast<std::chrono::minutes>(eMS).count() - (iH * 60);
    long double fS = (eMS.count() / 1000.0) - (double)((iM * 60) + (iH * 3600));
    std::cout << std::endl << std::setprecision(6) << what << name << " (completed in " << iH << ":" << iM << ":" << fS << ")" << std::endl;
}

int progress(int done, int spp, void* user)                                       // the reporter thread's line, main.cpp:95-109
{
    if (!*(bool*)user) std::cout << "\rSamples rendered: " << done << "/" << spp << std::flush;
    return 0;
}

} // namespace

int main(int argc, char** argv)
{
    auto start = std::chrono::high_resolution_clock::now();
    std::filesystem::path file = "teapot_scene.yaml";                            // main.cpp:146
    int gpus = 1, spp_override = 0, size_w = 0, size_h = 0, bvh = HRT_BVH_SAH; uint64_t seed = 1; bool quiet = false;
    std::string sum_path;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        if (a == "--gpus" && i + 1 < argc) gpus = std::atoi(argv[++i]);
        else if (a == "--seed" && i + 1 < argc) seed = std::strtoull(argv[++i], nullptr, 10);
        else if (a == "--spp" && i + 1 < argc) spp_override = std::atoi(argv[++i]);
        else if (a == "--size" && i + 2 < argc) { size_w = std::atoi(argv[++i]); size_h = std::atoi(argv[++i]); }
        else if (a == "--sum" && i + 1 < argc) sum_path = argv[++i];
        else if (a == "--quiet") quiet = true;
        else if (a == "--bvh" && i + 1 < argc) { const std::string m = argv[++i]; bvh = m == "reference" ? HRT_BVH_REFERENCE : (m == "lbvh_gpu" ? HRT_BVH_LBVH_GPU : HRT_BVH_SAH); }
        else file = a;
    }

    Scene scene;
    if (scene.loadScene(file.string()) < 1) return -1;                           // the reference's loader, unmodified (scene.cpp:127-374)
    hrt_scene* gpu = hrt_scene_create();
    if (!gpu) { std::cout << hrt_last_error() << std::endl; return -1; }
    Flattener flat{gpu};
    ExtLoader ext{scene, flat};
    if (ext.run(file.string()) < 1) return -1;
    if (spp_override > 0 || size_w > 0) {                                        // test conveniences: another film of the same output name
        const film_desc d = scene.film->getFilm();
        scene.film = std::make_shared<Film>(size_w > 0 ? size_w : d.dimensions.x, size_h > 0 ? size_h : d.dimensions.y, spp_override > 0 ? spp_override : d.samples, scene.film->outputName);
    }
    std::shared_ptr<Film> film = scene.getFilm();
    const film_desc f = film->getFilm();

    // the world list, the background and the camera cross the C ABI
    for (const std::shared_ptr<Hittable>& o : scene.objects.objects) {
        if (!o) { std::cout << "Object of an unknown type in the world list (scene.cpp:279,356)" << std::endl; return -1; }
        flat.object(o);
    }
    ck(hrt_scene_set_background(gpu, flat.texture(scene.getBackground())), "hrt_scene_set_background");
    hrt_camera cam; std::memset(&cam, 0, sizeof(cam));
    {   // Camera keeps derived vectors only (camera.h:41-47): its constructor arguments are read where the loader read them (scene.cpp:156-167)
        const YAML::Node c = YAML::LoadFile(file.string())["camera"];
        const glm::vec3 p = ExtLoader::vec3(c, "position"), at = ExtLoader::vec3(c, "look_at"), up = ExtLoader::vec3(c, "up");
        for (int a = 0; a < 3; ++a) { cam.position[a] = p[a]; cam.look_at[a] = at[a]; cam.up[a] = up[a]; }
        cam.vfov_deg = c["fov"].as<float>(); cam.aperture = c["aperture"].as<float>(); cam.focal_distance = c["focal_distance"].as<float>();
        cam.aspect = film->getAspectRatio();
    }
    print_elapsed("Loaded scene: ", file.string() + "!", start);                  // main.cpp:164-171

    if (hrt_device_count() < 1) die("no CUDA device");
    ck(hrt_scene_commit(gpu, bvh, seed), "hrt_scene_commit");                    // BVH builds + upload (mesh.cpp:38 -> bvh.cpp:6-61)
    hrt_render_params p; std::memset(&p, 0, sizeof(p));
    p.width = f.dimensions.x; p.height = f.dimensions.y; p.spp = f.samples; p.max_depth = MAX_DEPTH; p.t_min = 0.001f; p.seed = seed;
    std::vector<float> sum;
    if (!sum_path.empty()) sum.resize((size_t)p.width * p.height * 3);
    uint8_t* pixels = film->pixels.data();                                       // what render() writes through film->getPixels() (main.cpp:131)
    // RENDER (main.cpp:176): was render(NUM_THREADS, background, world, camera, film)
    int rc;
    if (gpus > 1) {
        std::vector<int> devices((size_t)gpus); std::iota(devices.begin(), devices.end(), 0);
        rc = hrt_render_multi(gpu, devices.data(), gpus, &cam, &p, sum.empty() ? nullptr : sum.data(), pixels);
    } else {
        const int slice = std::max(1, p.spp / 16);
        rc = hrt_render_progressive(gpu, &cam, &p, nullptr, slice, progress, &quiet, sum.empty() ? nullptr : sum.data(), pixels);
    }
    if (rc < 0) die("render");
    if (!quiet) std::cout << "\n";
    hrt_stats st; hrt_get_stats(gpu, &st);
    std::cout << "GPU: " << st.rays << " rays, " << st.paths << " camera samples, " << std::setprecision(6) << st.ms_total << " ms on the device ("
              << (st.ms_total > 0 ? st.rays / st.ms_total / 1e3 : 0.0) << " Mrays/s), " << gpus << " GPU(s)" << std::endl;
    if (!sum_path.empty()) {
        FILE* o = std::fopen(sum_path.c_str(), "wb");
        if (!o || std::fwrite(sum.data(), sizeof(float), sum.size(), o) != sum.size()) { std::cout << "cannot write " << sum_path << std::endl; return -1; }
        std::fclose(o);
    }
    hrt_scene_destroy(gpu);

    // OUTPUT IMAGE (main.cpp:180): unchanged
    int r = film->outputFilm();
    print_elapsed("Done!", "", start);
    return r;
}
