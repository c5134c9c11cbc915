// hrt_types.h — plain-old-data records shared by the host flattener (hrt_host.cpp) and the CUDA kernels
// (hrt_kernels.cu).  The reference's pointer tree (Hittable / Material / Texture, hittable.h:27-32,
// material.h:60-71, texture.h:3-7) becomes these GPU-resident arrays.
#pragma once
#include <stdint.h>
#include <vector_types.h>

#include "../../include/hrt.h"

namespace hrt {

constexpr int MAX_XF = 8;                    // instance wrappers per world-list entry
constexpr int MAT_MEDIUM_BASE = 1 << 24;     // material id >= this: Isotropic phase function of object (id - base)
constexpr int REF_STACK = 64;                // reference topology is balanced: depth <= ceil(log2 n) + 1
constexpr int FAST_STACK = 96;               // SAH tree depth cap (FAST_MAX_DEPTH) + top-level tree + markers
constexpr int FAST_MAX_DEPTH = 60;

enum { SHAPE_SPHERE = 0, SHAPE_RECT = 1, SHAPE_BOX = 2, SHAPE_MESH = 3, SHAPE_MEDIUM = 4 };
enum { TEX_SOLID = 0, TEX_CHECKER = 1, TEX_IMAGE = 2, TEX_ENV = 3, TEX_ENV_RGBE = 4 };

// translate.h / scale.h: (a,b,c) = vector.  rotateQuat.h: (a,b,c,d) = (w,x,y,z).  rotateY.h: a = sin, b = cos.
struct DXform { int kind; float a, b, c, d; };

// One world-list entry (scene.cpp:356 objects.add) = a shape under its wrapper chain, optionally a ConstantMedium.
struct DObject {
    int kind;               // SHAPE_* of the wrapped shape
    int axis;               // rects
    int material;
    uint32_t prim_base;     // primitive id of the first primitive of the shape
    uint32_t medium_prim;   // media: the pseudo primitive id reported for a hit inside the volume
    int mesh;               // index into DScene::meshes
    int xf_first, xf_count; // chain in DScene::xforms, innermost first (scene.cpp:335-353 wrap order)
    int medium;             // 1: ConstantMedium(boundary = this wrapped shape)   constantMedium.h:10-14
    int has_quat;           // chain contains a RotateQuat (its hit() reports t in other units, rotateQuat.cpp:51)
    float f[8];             // sphere: centre, radius | rect: a0 a1 b0 b1 k | box: min xyz, max xyz (box.h:41-44)
    float neg_inv_density;
    int phase_tex;
    float phase_rgb[3];
};

// Mesh (mesh.cpp:13-41): triangles + two acceleration structures.
//  ref_nodes  : the reference's topology (bvh.cpp:6-61) — 2 x float4 per node: (min.xyz, left) (max.xyz, right);
//               child >= 0 is a node, child < 0 is triangle ~child.  Traversed left-then-right (bvh.cpp:69-78).
//  tris       : 3 x float4 per triangle in input order: (p0, -) (p1, -) (p2, -)
//  attrs      : 4 x float4 per triangle in input order: (n0, uv0.x) (n1, uv0.y) (n2, uv1.x) (uv1.y, uv2.x, uv2.y, -)
//  fast_leaves: one 128-byte record per BOTTOM node of the reference topology (a node whose children are triangles),
//               8 x float4: (min.xyz, triL) (max.xyz, triR) (L.p0, triL) (L.p1) (L.p2) (R.p0, triR) (R.p1) (R.p2): the last
//               96 bytes are all a leaf visit reads (three 256-bit loads); the box only when a triangle accepts.  The box is
//               the reference node's box, tested with the reference's own slab arithmetic, so which triangles a
//               ray gets to test below t_min (ITriangle::hit ignores t_min, triangle.cpp:57-131) is unchanged.
//  fast_nodes : binned-SAH BVH2 over those bottom nodes — 4 x float4 per node: (c0.lo.x c0.hi.x c0.lo.y c0.hi.y)
//               (c1.lo.x c1.hi.x c1.lo.y c1.hi.y) (c0.lo.z c0.hi.z c1.lo.z c1.hi.z) (child0, child1, -, -);
//               child >= 0 is a node, child < 0 is leaf record ~child.
// Wide node (hrt_wide.cuh): 8 children with boxes quantised to 8 bits on a per-node grid, 96 bytes = three 32-byte
// sectors read with three 256-bit loads.
//   [ 0] p[3]    grid origin (the node box's min corner)
//   [12] meta    ex | ey << 8 | ez << 16 | valid << 24: biased exponents of the grid steps (step = 2^(e - 127)) and the
//                mask of occupied child slots
//   [16] q[6][8] lo.x hi.x lo.y hi.y lo.z hi.z of the eight slots: child box = [p + lo * step, p + hi * step], a superset
//                of the child's true box
//   [64] child[8] >= 0: wide node, < 0: leaf ~id (mesh: leaf record, top level: world-list entry)
// Slot s holds the child whose centre lies most towards ((s & 1) ? +x : -x, (s & 2) ? +y : -y, (s & 4) ? +z : -z) of the
// node centre, so that slot ^ (sign bits of the ray direction) orders the children roughly front to back (Ylitie et al. 2017).
struct WNode { float p[3]; uint32_t meta; uint8_t q[6][8]; int32_t child[8]; };
static_assert(sizeof(WNode) == 96, "WNode layout");

struct DMesh {
    const float4* ref_nodes;
    const float4* tris;
    const float4* attrs;
    const float4* fast_nodes;
    const float4* fast_leaves;
    const WNode* wide_nodes; // 8-wide collapse of fast_nodes (the tree the hot kernels walk)
    int wide_root;          // >= 0 node, < 0 leaf ~id
    int ref_root;           // -1: empty mesh
    int fast_root;          // >= 0 node, < 0 leaf ~id; only valid when ref_root >= 0
    int64_t n_tri;
    uint32_t prim_base;
    int material;
};

struct DTexture {
    int kind;
    float rgb[3];
    int even, odd;
    int w, h, ch;
    const void* data;       // uint8 rgb (image), float (env) or uint32 rgbe (env, lossless repack)
};

struct DScene {
    const DObject* objects;
    const DXform* xforms;
    const DMesh* meshes;
    const hrt_material* materials;
    const DTexture* textures;
    const float4* tlas_nodes;   // top-level BVH over world-list entries (same node format as fast_nodes), or null
    int tlas_root;              // >= 0 node, < 0 leaf = object ~id
    const WNode* tlas_wide;     // 8-wide collapse of tlas_nodes
    int tlas_wide_root;
    int wide;                   // 1: the wide trees exist for every level (fast scenes whose boxes fit the 8-bit grids): hrt_wide.cuh walks them
    int n_objects;
    int n_materials;
    int n_textures;
    int background;
    int fast;                   // 0: reference visiting order everywhere; 1: SAH BVHs + reference-order re-trace of ambiguous rays
    int big_mesh;               // 1: some mesh has >= 2^17 triangles (its trees leave L1: the wide walk pays for every ray kind)
    int wide_has_slow;          // 1: some top-level leaf carries WIDE_SLOW_BIT (an entry that takes the general object test): the kernels with the fifth phase
    int wide_phase_min;         // wide walk: a phase runs when at least this many lanes wait for it (hrt_wide.cuh; 8 with meshes, 16 without)
    int wide_phase_min_obj;     // the same for world-list entries that are tested on the spot (spheres, rects, boxes, media)
    int wide_refill;            // wide walk: idle lanes take new rays as soon as fewer than this many lanes are walking (24)
    int refill_secondary;       // 1: when secondary rays dominate they are traced by k_wide_secondary (scenes with meshes, or with at most 64 world-list entries)
    float tlas_cull_scale;      // lower bound of (reported t) / (world t * |d|) over RotateQuat objects; 0 if none
    int has_bounds;             // 1: bounds_lo/hi enclose every world-list entry (camera rays that pass outside are misses)
    float bounds_lo[3], bounds_hi[3];
    int nan_obj;                // what world->hit returns for a ray whose origin is NaN in all components: every
    uint32_t nan_prim;          // primitive accepts it (all comparisons false), so the last one in reference order wins; -1: miss
};

struct DCamera { float origin[3], llc[3], horiz[3], vert[3];      // camera.h:9-30
                 float u[3], v[3], lens_radius; };                 // lens basis; lens_radius 0 unless HRT_RENDER_DEFOCUS (camera.h:29,34)

} // namespace hrt
