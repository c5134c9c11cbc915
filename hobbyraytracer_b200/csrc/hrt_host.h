// hrt_host.h — host-side scene builder: collects the constructor calls of scene.cpp:127-374 / mesh.cpp:13-41 and
// flattens them into the arrays of hrt_types.h.  No CUDA in here (plain C++17).
#pragma once
#include <cstdint>
#include <functional>
#include <string>
#include <vector>

#include "hrt_types.h"

namespace hrt {

// Recycled word buffers for packed environment maps: a fresh 32 MiB allocation costs more in page faults (zero fill + first
// touch, ~20 ms on the GPU box) than packing into it; scenes are typically rebuilt with an environment of the same size.
std::vector<uint32_t> take_words(size_t n);
void give_words(std::vector<uint32_t>&& v);

struct HostTexture {
    int kind = TEX_SOLID;
    float rgb[3] = {0, 0, 0};
    int even = -1, odd = -1;
    int w = 0, h = 0, ch = 0;
    std::vector<uint8_t> u8;
    std::vector<float> f32;
    std::vector<uint32_t> rgbe;     // env maps whose every texel is exactly RGBE-representable are kept ONLY in this form (f32 stays empty)
    HostTexture() = default;
    HostTexture(HostTexture&&) = default;
    HostTexture& operator=(HostTexture&&) = default;
    HostTexture(const HostTexture&) = default;
    HostTexture& operator=(const HostTexture&) = default;
    ~HostTexture() { if (!rgbe.empty()) give_words(std::move(rgbe)); }
};

struct RefBvh {                      // bvh.cpp:6-61 in index form
    std::vector<float4> nodes;       // 2 per node
    int root = -1;
};
struct FastBvh {
    std::vector<float4> nodes;       // 4 per node
    std::vector<int> order;          // leaf order -> item index
    int root = -1;
    int depth = 0;
};

struct WideBvh {                     // 8-wide collapse of a FastBvh (hrt_bvh_rules.h (4)); breadth-first numbering, root = node 0
    std::vector<WNode> nodes;
    int root = -1;                   // >= 0 node, < 0 leaf ~item (a tree over one item), -1 with no nodes: empty
    int depth = 0;
    bool ok = true;                  // false: some node box is not representable on the 8-bit grid (unbounded entries)
};

struct HostShape {
    int kind = SHAPE_SPHERE;
    float f[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int axis = 0;
    int material = -1;
    std::vector<float> pos, nrm, uv;     // triangle soup: 9 + 9 + 6 floats per triangle (hrt_scene_add_mesh)
    // hrt_scene_add_mesh_indexed: vertex / index buffers (indices already carry their sub-mesh's base vertex); the soup
    // above stays empty until a host builder needs it (ensure_soup) - the device builder gathers through the indices
    bool indexed = false;
    std::vector<float> vpos, vnrm, vuv;  // 3 + 3 + 2 floats per vertex (owned copies; empty when borrowed)
    std::vector<uint32_t> vidx;          // 3 per triangle
    const float *pv = nullptr, *pn = nullptr, *pt = nullptr;   // HRT_MESH_BORROW: the caller's arrays (valid until the commit
    const uint32_t* pi = nullptr;                              // returns, dropped by release_borrowed), else NULL
    const float* V() const { return pv ? pv : vpos.data(); }   // what the builders read
    const float* N() const { return pn ? pn : vnrm.data(); }
    const float* T() const { return pt ? pt : vuv.data(); }
    const uint32_t* I() const { return pi ? pi : vidx.data(); }
    void release_borrowed() { pv = pn = pt = nullptr; pi = nullptr; }
    int64_t n_vert = 0;
    int64_t n_tri = 0;
    int64_t prim_base = 0, prim_count = 0;
    bool used = false;
    int mesh_index = -1;
    RefBvh ref;
    FastBvh fast;
    WideBvh wide;
    std::vector<float4> tris4, attrs4;   // device layouts of hrt_types.h (DMesh::tris, DMesh::attrs)
    std::vector<float4> leaves;          // DMesh::fast_leaves
    // HRT_BVH_LBVH_GPU: flatten() leaves the mesh to the device builder (gpu_pending); hrt_scene_commit then stores the
    // device-built view here (has_dev) and make_dscene hands it out instead of uploading host arrays
    bool gpu_pending = false, has_dev = false;
    DMesh dev{};
    size_t n_ref_nodes = 0, n_fast_nodes = 0, n_leaves = 0, n_wide_nodes = 0;   // element counts behind the DMesh pointers (either builder)
    int last_tri_r = -1;                 // right triangle of the last bottom node, -1: empty mesh
};

struct HostObject {
    int shape = -1;
    std::vector<DXform> ops;         // innermost first
    bool medium = false;
    float neg_inv_density = 0;
    int phase_tex = -1;
    float phase_rgb[3] = {0, 0, 0};
    int pseudo_shape = -1;
};

struct HostScene {
    std::vector<HostTexture> textures;
    std::vector<hrt_material> materials;
    std::vector<HostShape> shapes;
    std::vector<HostObject> objects;
    int64_t n_prims = 0;
    int background = -1;
    bool committed = false;
    int bvh_mode = HRT_BVH_REFERENCE;
    // results of flatten()
    std::vector<DObject> d_objects;
    std::vector<DXform> d_xforms;
    std::vector<int> mesh_shapes;    // DScene::meshes[i] is built from shapes[mesh_shapes[i]]
    FastBvh tlas;
    WideBvh tlas_wide;
    bool tlas_has_slow = false;      // some top-level leaf of tlas_wide is tagged WIDE_SLOW_BIT (hrt_wide.cuh)
    float tlas_cull_scale = 1.0f;
    float bounds_lo[3] = {0, 0, 0}, bounds_hi[3] = {0, 0, 0};   // union of the world-list entries' padded bounds
    bool has_bounds = false;         // false: some entry is unbounded (or reference-order scene)
    double build_ms_ref = 0, build_ms_fast = 0;
    uint64_t content_hash = 0;       // of everything that decides the image (flatten(): objects, wrappers, materials, textures, geometry; large arrays sampled): checkpoints carry it
};

// Converts the ABI's hrt_xform (incl. the *_DEG conveniences) into the device record, with the host-side
// trigonometry of scene.cpp:340 / rotateY.cpp:7-9.
DXform make_xform(const hrt_xform& x);

// bvh.cpp:6-61: median split of box.min[axis] with axis = glm::linearRand(0,2) on the stream (seed, shape_id).
// stable = false: ranges ordered with std::sort like bvh.cpp:39 (equal keys permute as libstdc++'s introsort does);
// stable = true: equal keys keep their order, the rule of the GPU builder (hrt_bvh_rules.h).
void build_ref_bvh(HostShape& sh, uint64_t seed, int shape_id, int n_threads, bool stable = false);
// Morton LBVH over arbitrary boxes by the rules of hrt_bvh_rules.h: the host mirror of hrt_bvh_gpu.cu.
void build_lbvh_host(const std::vector<float>& bmin, const std::vector<float>& bmax, FastBvh& out);
// Per-mesh part of flatten() on the host: reference topology, device layouts, leaf records, fast tree.
void build_mesh_host(HostShape& sh, int bvh_mode, uint64_t seed, int shape_id, int n_threads, double& ms_ref, double& ms_fast);
// Binned-SAH BVH2 over arbitrary boxes (used for meshes and for the top-level BVH).
void build_fast_bvh(const std::vector<float>& bmin, const std::vector<float>& bmax, FastBvh& out, int n_threads);
// 8-wide collapse of `fast` (hrt_bvh_rules.h (4)): the tree the hot kernels walk.
void build_wide(const FastBvh& fast, WideBvh& out);
void triangle_boxes(const HostShape& sh, std::vector<float>& bmin, std::vector<float>& bmax);   // triangle.cpp:133-151

// Fills d_objects / d_xforms / tlas; builds the per-mesh BVHs.  Returns "" or an error message.
std::string flatten(HostScene& s, int bvh_mode, uint64_t seed);

// Exact-repack test for environment maps: true (and rgbe filled) when every texel equals ldexp(m, e - 136) with
// 8-bit mantissas and one shared exponent, which is what stb_image produces for Radiance .hdr files (texture.cpp:99-115).
bool try_pack_rgbe(const float* f32, int w, int h, int ch, std::vector<uint32_t>& rgbe);

// Camera ctor camera.h:9-30
DCamera make_camera(const hrt_camera& c);

// scene-builder half of the C ABI (hrt_builder.cpp): ids >= 0 or HRT_ERR_* with `err` filled
int b_add_texture_solid(HostScene& s, const float rgb[3], std::string& err);
int b_add_texture_checker(HostScene& s, int even_tex, int odd_tex, std::string& err);
int b_add_texture_image(HostScene& s, const uint8_t* rgb8, int w, int h, std::string& err);
int b_add_texture_env(HostScene& s, const float* data, int w, int h, int channels, std::string& err);
int b_add_material(HostScene& s, const hrt_material* m, std::string& err);
int b_add_sphere(HostScene& s, const float c[3], float radius, int material, std::string& err);
int b_add_rect(HostScene& s, int axis, float a0, float a1, float b0, float b1, float k, int material, std::string& err);
int b_add_box(HostScene& s, const float c[3], const float dim[3], int material, std::string& err);
int b_add_mesh(HostScene& s, const float* pos, const float* nrm, const float* uv, int64_t n_tri, int material, std::string& err);
int b_add_mesh_indexed(HostScene& s, const float* vtx, const float* nrm, const float* uv, int64_t n_vert, const uint32_t* idx, int64_t n_idx,
                       const hrt_submesh* sub, int n_sub, int material, uint32_t flags, std::string& err);
void parallel_chunks(size_t n, size_t min_chunk, const std::function<void(size_t chunk, size_t begin, size_t end)>& fn, size_t* n_chunks = nullptr);
void ensure_soup(HostShape& sh);         // indexed mesh: materialise pos / nrm / uv (mesh.cpp:24-33) for the host builders
int b_add_object(HostScene& s, int shape, const hrt_xform* ops, int n_ops, std::string& err);
int b_add_medium(HostScene& s, int shape, const hrt_xform* ops, int n_ops, float density, const float rgb[3], int albedo_tex, std::string& err);
int b_set_background(HostScene& s, int tex, std::string& err);

} // namespace hrt
