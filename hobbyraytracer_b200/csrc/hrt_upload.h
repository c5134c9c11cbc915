// hrt_upload.h — turns a flattened HostScene into the DScene view the kernels read.  `Up::put(src, bytes)` returns
// where the bytes live for the consumer: a cudaMalloc'ed copy in the product (hrt_api.cu), the host pointer itself
// in the CPU simulation harness (tests/hostsim).
#pragma once
#include <cstring>
#include <vector>

#include "hrt_host.h"

namespace hrt {

template <class Up>
DScene make_dscene(HostScene& s, Up& up)
{
    DScene d{};
    const bool fast = s.bvh_mode != HRT_BVH_REFERENCE;
    std::vector<DMesh> meshes(s.mesh_shapes.size());
    for (size_t i = 0; i < s.mesh_shapes.size(); ++i) {
        HostShape& sh = s.shapes[(size_t)s.mesh_shapes[i]];
        DMesh m{};
        if (sh.has_dev) { m = up.dev_mesh(i, sh); m.prim_base = (uint32_t)sh.prim_base; m.material = sh.material; meshes[i] = m; continue; }   // built on the device: its arrays as they live on the consumer's device
        m.ref_nodes = (const float4*)up.put(sh.ref.nodes.data(), sh.ref.nodes.size() * sizeof(float4));
        m.tris = (const float4*)up.put(sh.tris4.data(), sh.tris4.size() * sizeof(float4));
        m.attrs = (const float4*)up.put(sh.attrs4.data(), sh.attrs4.size() * sizeof(float4));
        m.fast_nodes = (const float4*)up.put(sh.fast.nodes.data(), sh.fast.nodes.size() * sizeof(float4));
        m.fast_leaves = (const float4*)up.put(sh.leaves.data(), sh.leaves.size() * sizeof(float4));
        m.wide_nodes = (const WNode*)up.put(sh.wide.nodes.data(), sh.wide.nodes.size() * sizeof(WNode));
        m.wide_root = sh.wide.root;
        m.ref_root = sh.ref.root; m.fast_root = sh.fast.root;
        m.n_tri = sh.n_tri; m.prim_base = (uint32_t)sh.prim_base; m.material = sh.material;
        meshes[i] = m;
    }
    std::vector<DTexture> tex(s.textures.size());
    for (size_t i = 0; i < s.textures.size(); ++i) {
        HostTexture& t = s.textures[i];
        DTexture x{};
        x.kind = t.kind; x.rgb[0] = t.rgb[0]; x.rgb[1] = t.rgb[1]; x.rgb[2] = t.rgb[2];
        x.even = t.even; x.odd = t.odd; x.w = t.w; x.h = t.h; x.ch = t.ch; x.data = nullptr;
        if (t.kind == TEX_IMAGE && !t.u8.empty()) x.data = up.put(t.u8.data(), t.u8.size());
        if (t.kind == TEX_ENV) {
            if (!t.rgbe.empty()) { x.kind = TEX_ENV_RGBE; x.data = up.put(t.rgbe.data(), t.rgbe.size() * sizeof(uint32_t)); }
            else if (!t.f32.empty()) x.data = up.put(t.f32.data(), t.f32.size() * sizeof(float));
        }
        tex[i] = x;
    }
    d.objects = (const DObject*)up.put(s.d_objects.data(), s.d_objects.size() * sizeof(DObject));
    d.xforms = (const DXform*)up.put(s.d_xforms.data(), s.d_xforms.size() * sizeof(DXform));
    d.meshes = (const DMesh*)up.put_copy(meshes.data(), meshes.size() * sizeof(DMesh));
    d.materials = (const hrt_material*)up.put(s.materials.data(), s.materials.size() * sizeof(hrt_material));
    d.textures = (const DTexture*)up.put_copy(tex.data(), tex.size() * sizeof(DTexture));
    d.tlas_nodes = (const float4*)up.put(s.tlas.nodes.data(), s.tlas.nodes.size() * sizeof(float4));
    d.tlas_root = s.tlas.root;
    d.tlas_wide = (const WNode*)up.put(s.tlas_wide.nodes.data(), s.tlas_wide.nodes.size() * sizeof(WNode));
    d.tlas_wide_root = s.tlas_wide.root;
    d.n_objects = (int)s.d_objects.size();
    d.n_materials = (int)s.materials.size();
    d.n_textures = (int)s.textures.size();
    d.background = s.background;
    d.fast = fast ? 1 : 0;
    d.wide = (fast && s.tlas_wide.ok) ? 1 : 0;
    for (const DMesh& m : meshes) if (m.ref_root >= 0 && m.wide_root >= 0 && m.wide_nodes == nullptr) d.wide = 0;
    for (size_t i = 0; i < s.mesh_shapes.size(); ++i) if (!s.shapes[(size_t)s.mesh_shapes[i]].wide.ok) d.wide = 0;
    // scenes with meshes (long, uneven walks), and scenes whose whole world list fits two levels of the wide top-level tree (a Cornell box
    // with boxes / media: B200 A/B r2w cornell_smoke 3480 -> 3780 Mrays/s; 485 spheres lose: scattered_balls 5896 -> 3644)
    // r3d A/B (Mrays/s at phase_min 8 / 12 / 16): teapot_scene 2136 / 2138 / 2133, ray batch 1705 / 1529 / 1536, big_mesh 2498 / 2403 / 2398,
    // cornell_smoke 3798 / 4013 / 4342 - without meshes every leaf is a world-list entry and fuller phases pay
    // (r3e, cornell_smoke at phase_min / refill 8/24, 16/24, 16/16, 20/20, 24/24: 3809, 4352, 4705, 4508, 4359; refill 16 costs the mesh scenes 5-10 %)
    d.wide_has_slow = s.tlas_has_slow ? 1 : 0;
    d.wide_phase_min = meshes.empty() ? 16 : 8; d.wide_phase_min_obj = meshes.empty() ? 16 : 12;   /* r3g teapot_scene at 8/8, 8/12, 8/16, 8/24: 2099, 2127, 2115, 2127; mesh entries keep the node threshold (r3i: ray batch 1722 vs 1554 when they wait for 12) */ d.wide_refill = meshes.empty() ? 16 : 24;
    d.refill_secondary = (fast && (!meshes.empty() || d.n_objects <= 64)) ? 1 : 0;
    d.big_mesh = 0;
    for (const DMesh& m : meshes) if (m.n_tri >= (int64_t(1) << 17)) d.big_mesh = 1;
    d.tlas_cull_scale = s.tlas_cull_scale;
    d.has_bounds = (fast && s.has_bounds) ? 1 : 0;
    for (int a = 0; a < 3; ++a) { d.bounds_lo[a] = s.bounds_lo[a]; d.bounds_hi[a] = s.bounds_hi[a]; }
    d.nan_obj = -1; d.nan_prim = HRT_MISS;
    for (int i = (int)s.d_objects.size() - 1; i >= 0 && d.nan_obj < 0; --i) {
        const DObject& ob = s.d_objects[(size_t)i];
        if (ob.medium) { d.nan_obj = i; d.nan_prim = ob.medium_prim; }           // (an empty-mesh boundary never hits, see below)
        else if (ob.kind == SHAPE_BOX) { d.nan_obj = i; d.nan_prim = ob.prim_base + 5u; }
        else if (ob.kind != SHAPE_MESH) { d.nan_obj = i; d.nan_prim = ob.prim_base; }
        if (ob.kind == SHAPE_MESH) {                                             // last triangle of the left-then-right walk (bvh.cpp:69-78)
            const HostShape& sh = s.shapes[(size_t)s.mesh_shapes[(size_t)ob.mesh]];
            if (sh.n_leaves == 0 || sh.last_tri_r < 0) { d.nan_obj = -1; d.nan_prim = HRT_MISS; continue; }
            if (!ob.medium) { d.nan_obj = i; d.nan_prim = ob.prim_base + (uint32_t)sh.last_tri_r; }
        }
    }
    return d;
}

} // namespace hrt
