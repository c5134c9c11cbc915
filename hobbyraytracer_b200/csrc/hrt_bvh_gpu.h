// hrt_bvh_gpu.h — interface of the device-side mesh builder (hrt_bvh_gpu.cu) used by hrt_scene_commit(HRT_BVH_LBVH_GPU).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <unordered_map>
#include <vector>

#include "hrt_types.h"

namespace hrt {

struct GpuMeshBuild {
    DMesh mesh{};                    // device pointers + roots (prim_base / material are filled in by the caller)
    size_t n_ref_nodes = 0, n_fast_nodes = 0, n_leaves = 0, n_wide_nodes = 0;
    int wide_depth = 0;
    int last_tri_r = -1;             // right triangle of the last bottom node (what a NaN-origin ray reports, hrt_upload.h)
    int depth = 0;                   // depth of the fast tree
    bool fallback = false;           // true: this mesh cannot be built here (why), the host builder takes over
    std::string why;
    double ms_upload = 0, ms_build = 0;
    size_t h2d_bytes = 0;            // bytes of mesh source copied host -> device
};

// Host-side view of one mesh: the raw triangle soup (9 + 9 + 6 floats per triangle, the layout of hrt_scene_add_mesh), or
// - indexed - per-vertex arrays (3 + 3 + 2 floats) with 3 indices per triangle (hrt_scene_add_mesh_indexed)
struct MeshSource {
    const float *pos = nullptr, *nrm = nullptr, *uv = nullptr;
    const uint32_t* idx = nullptr;
    int64_t n_vert = 0;
    bool indexed = false;
};

// Size-keyed cache of device blocks shared by the uploads of hrt_scene_commit and the device builder (hrt_api.cu): blocks of a
// destroyed scene and the builder's scratch are kept (bounded; hrt_trim frees them) for the next build of this process, so a host
// that rebuilds its scene every frame neither pays cudaMalloc nor the device-wide synchronisation of cudaFree while another scene
// of the process is rendering.  A block is only given back once no stream uses it any more.
cudaError_t pool_alloc(int device, size_t bytes, void** p);   // a cached block of exactly this size, else cudaMalloc
void pool_give(int device, size_t bytes, void* p);            // keeps the block, or cudaFree when the cache is full

// Uploads the mesh source and builds every per-mesh array on the device (an indexed source is gathered there, so only
// the compact buffers cross PCIe).  Result allocations are appended to `allocs`, their sizes to `alloc_bytes` (owned by the scene; pool blocks).  Returns false
// with `err` set on a CUDA failure.
bool build_mesh_gpu(const MeshSource& src, int64_t n_tri, uint64_t seed, int shape_id,
                    cudaStream_t st, std::vector<void*>& allocs, std::unordered_map<void*, size_t>& alloc_bytes, size_t& bytes, GpuMeshBuild& out,
                    std::string& err, uint64_t* launches);

} // namespace hrt
