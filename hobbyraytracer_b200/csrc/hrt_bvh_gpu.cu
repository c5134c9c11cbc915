// hrt_bvh_gpu.cu — HRT_BVH_LBVH_GPU: the whole per-mesh build of hrt_scene_commit on the device.
//
// Replaces, for one mesh, what the host path does in hrt_host.cpp::flatten (mesh.cpp:24-38 + bvh.cpp:6-61 behind it):
//   k_mesh_pack      raw positions / normals / uvs -> DMesh::tris, DMesh::attrs, per-triangle boxes (triangle.cpp:133-151)
//   k_node_axes      split axis of every node of the reference topology (rand() stream position 4k, hrt_bvh_rules.h)
//   k_level_keys     + one stable radix sort per tree level: each range of level l ordered by box.min[axis] (bvh.cpp:39)
//   k_emit_level     bottom-up: reference-topology nodes (DMesh::ref_nodes) and, for bottom nodes, the 128-byte leaf
//                    records (DMesh::fast_leaves)
//   k_centre_bounds, k_morton + one radix sort, k_karras, k_refit, k_depth:   Morton LBVH over the bottom nodes in the
//                    fast_nodes format (both children's boxes in the node)
//   k_wide_level, one exclusive scan, k_wide_link per level of the 8-wide collapse (DMesh::wide_nodes, the tree the hot
//                    kernels walk): breadth first, numbering and bytes of hrt_host.cpp::build_wide (hrt_bvh_rules.h (4))
// The arrays are bit-identical to the host mirror (HRT_BVH_LBVH_HOST): every rule is shared source (hrt_bvh_rules.h),
// boxes are unions taken in a fixed operand order.  The two sorts are cub::DeviceRadixSort (stable LSD radix sort);
// everything else is hand-written.  The tree shape is known from the triangle count alone, so the level loop runs
// without any device -> host read; the only reads back are the validity flag, the depth and the last leaf's triangle.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cuda_runtime.h>

#include <chrono>
#include <cstring>
#include <string>
#include <vector>

#include "hrt_bvh_gpu.h"
#include "hrt_bvh_rules.h"
#include "hrt_math.cuh"

namespace hrt {
namespace {

constexpr int TB = 256;
inline unsigned int grid_for(long long n) { return (unsigned int)((n + TB - 1) / TB); }

__device__ __forceinline__ int as_i(float f) { return __float_as_int(f); }
__device__ __forceinline__ float as_f(int i) { return __int_as_float(i); }

// flags[0] != 0: some position is not finite (keys would not be ordered; the host path takes over)
// INDEXED: pos / nrm / uv are per-vertex arrays (3 + 3 + 2 floats) and corner v of triangle i is vertex idx[3i + v] -
// Mesh::Mesh's gather (mesh.cpp:24-33) fused into the packing pass
template <bool INDEXED>
__global__ void k_mesh_pack(const float* __restrict__ pos, const float* __restrict__ nrm, const float* __restrict__ uv,
                            const unsigned int* __restrict__ idx, long long n,
                            float4* __restrict__ tris, float4* __restrict__ attrs, float4* __restrict__ tmin, float4* __restrict__ tmax,
                            unsigned int* __restrict__ order, int* __restrict__ flags)
{
    long long i = (long long)blockIdx.x * TB + threadIdx.x;
    if (i >= n) return;
    float P[9], N[9], T[6];
    if (INDEXED) {
        #pragma unroll
        for (int v = 0; v < 3; ++v) {
            const size_t q = __ldg(idx + i * 3 + v);
            #pragma unroll
            for (int a = 0; a < 3; ++a) { P[v * 3 + a] = __ldg(pos + q * 3 + a); N[v * 3 + a] = __ldg(nrm + q * 3 + a); }
            T[v * 2] = __ldg(uv + q * 2); T[v * 2 + 1] = __ldg(uv + q * 2 + 1);
        }
    } else {
        #pragma unroll
        for (int k = 0; k < 9; ++k) { P[k] = __ldg(pos + i * 9 + k); N[k] = __ldg(nrm + i * 9 + k); }
        #pragma unroll
        for (int k = 0; k < 6; ++k) T[k] = __ldg(uv + i * 6 + k);
    }
    bool bad = false;
    #pragma unroll
    for (int k = 0; k < 9; ++k) bad |= !(fabsf(P[k]) <= 3.0e38f);
    if (bad) atomicOr(flags, 1);
    #pragma unroll
    for (int v = 0; v < 3; ++v) tris[i * 3 + v] = make_float4(P[v * 3], P[v * 3 + 1], P[v * 3 + 2], 0.0f);
    attrs[i * 4] = make_float4(N[0], N[1], N[2], T[0]);
    attrs[i * 4 + 1] = make_float4(N[3], N[4], N[5], T[1]);
    attrs[i * 4 + 2] = make_float4(N[6], N[7], N[8], T[2]);
    attrs[i * 4 + 3] = make_float4(T[3], T[4], T[5], 0.0f);
    float mn[3], mx[3];
    #pragma unroll
    for (int a = 0; a < 3; ++a) {                                    // triangle.cpp:133-151
        mn[a] = cmin(cmin(P[a], P[3 + a]), P[6 + a]) - 0.0001f;
        mx[a] = cmax(cmax(P[a], P[3 + a]), P[6 + a]) + 0.0001f;
    }
    tmin[i] = make_float4(mn[0], mn[1], mn[2], 0.0f);
    tmax[i] = make_float4(mx[0], mx[1], mx[2], 0.0f);
    order[i] = (unsigned int)i;
}

__global__ void k_node_axes(uint64_t stream0, long long n_nodes, uint8_t* __restrict__ axis)
{
    long long k = (long long)blockIdx.x * TB + threadIdx.x;
    if (k < n_nodes) axis[k] = (uint8_t)rule_axis_of_node(stream0, (uint64_t)k);
}

__device__ __forceinline__ float axis_of(float4 v, int a) { return a == 0 ? v.x : (a == 1 ? v.y : v.z); }

// key of position p for the ordering of level `level`: (range start >> shift, box.min[axis of the range's node]); positions
// in ranges that are already bottom ranges get the bare prefix and keep their order (the sort is stable).  Every range of a
// level holds at least 2^shift positions, so start >> shift still tells the ranges apart, in order, with ~level bits
// instead of log2(n): the radix sort runs over 33 + level key bits (r1h: 56 bits on every level, 169 passes for 10.5 M)
__global__ void k_level_keys(TreeShape ts, int level, int shift, const unsigned int* __restrict__ order, const float4* __restrict__ tmin,
                             const uint8_t* __restrict__ axis, unsigned long long* __restrict__ keys)
{
    unsigned int p = blockIdx.x * TB + threadIdx.x;
    if (p >= ts.n) return;
    TreeRange r = tree_descend(ts, p, level);
    unsigned long long key = (unsigned long long)(r.start >> shift) << 32;
    if (r.n > 2) key |= rule_float_key((uint32_t)as_i(axis_of(__ldg(tmin + order[p]), axis[r.self])));
    keys[p] = key;
}

// One thread per range of level `level` (range j is reached from the root by the bits of j, most significant first).
__global__ void k_emit_level(TreeShape ts, int level, const unsigned int* __restrict__ order, const float4* __restrict__ tmin,
                             const float4* __restrict__ tmax, const float4* __restrict__ tris, const uint8_t* __restrict__ axis,
                             float4* __restrict__ ref_nodes, float4* __restrict__ leaves)
{
    unsigned int j = blockIdx.x * TB + threadIdx.x;
    if (level < 32 && j >= (1u << level)) return;
    TreeRange r; r.start = 0; r.n = ts.n; r.self = 0; r.bottoms_before = 0; r.level = 0;
    for (int l = 0; l < level; ++l) {
        if (r.n <= 2) return;                                     // that branch ended above this level
        const uint32_t nl = r.n >> 1, k = nl - ts.lo[l + 1];
        if (((j >> (level - 1 - l)) & 1u) == 0) { r.n = nl; r.self += 1; }
        else { r.start += nl; r.n -= nl; r.self += 1 + ts.nodes[l + 1][k]; r.bottoms_before += ts.bottoms[l + 1][k]; }
    }
    int left, right;
    float4 lmn, lmx, rmn, rmx;
    if (r.n <= 2) {
        unsigned int t0 = order[r.start], t1 = t0;
        if (r.n == 2) {                                           // bvh.cpp:25-35: the smaller key goes left, a tie swaps
            t1 = order[r.start + 1];
            const int a = axis[r.self];
            if (!(axis_of(tmin[t0], a) < axis_of(tmin[t1], a))) { unsigned int s = t0; t0 = t1; t1 = s; }
        }
        left = ~(int)t0; right = ~(int)t1;
        lmn = tmin[t0]; lmx = tmax[t0]; rmn = tmin[t1]; rmx = tmax[t1];
    } else {
        const uint32_t nl = r.n >> 1;
        left = (int)r.self + 1; right = (int)(r.self + 1 + ts.nodes[level + 1][nl - ts.lo[level + 1]]);
        lmn = ref_nodes[2 * (size_t)left]; lmx = ref_nodes[2 * (size_t)left + 1];
        rmn = ref_nodes[2 * (size_t)right]; rmx = ref_nodes[2 * (size_t)right + 1];
    }
    float4 a, b;                                                  // surroundingBox aabb.h:41-56
    a.x = cmin(lmn.x, rmn.x); a.y = cmin(lmn.y, rmn.y); a.z = cmin(lmn.z, rmn.z); a.w = as_f(left);
    b.x = cmax(lmx.x, rmx.x); b.y = cmax(lmx.y, rmx.y); b.z = cmax(lmx.z, rmx.z); b.w = as_f(right);
    ref_nodes[2 * (size_t)r.self] = a; ref_nodes[2 * (size_t)r.self + 1] = b;
    if (r.n <= 2) {
        const size_t tl = (size_t)(~left), tr = (size_t)(~right);
        float4* rec = leaves + 8 * (size_t)r.bottoms_before;
        a.w = as_f((int)tl); b.w = as_f((int)tr);
        rec[0] = a; rec[1] = b;
        #pragma unroll
        for (int v = 0; v < 3; ++v) {
            float4 l = tris[tl * 3 + v], r = tris[tr * 3 + v];
            if (v == 0) { l.w = as_f((int)tl); r.w = as_f((int)tr); }     // the ids again, inside the 96 bytes a leaf visit reads
            rec[2 + v] = l; rec[5 + v] = r;
        }
    }
}

// bounds[0..2] = min, bounds[3..5] = max of the leaf box centres, as rule_float_key images
__global__ void k_centre_bounds(const float4* __restrict__ leaves, long long n, unsigned int* __restrict__ bounds)
{
    __shared__ unsigned int s_lo[3], s_hi[3];
    if (threadIdx.x < 3) { s_lo[threadIdx.x] = 0xffffffffu; s_hi[threadIdx.x] = 0u; }
    __syncthreads();
    const long long i = (long long)blockIdx.x * TB + threadIdx.x;
    unsigned int klo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, khi[3] = {0u, 0u, 0u};
    if (i < n) {
        const float4 mn = leaves[8 * i], mx = leaves[8 * i + 1];
        const float c[3] = {rule_centre(mn.x, mx.x), rule_centre(mn.y, mx.y), rule_centre(mn.z, mx.z)};
        #pragma unroll
        for (int a = 0; a < 3; ++a) klo[a] = khi[a] = rule_float_key((uint32_t)as_i(c[a]));
    }
    #pragma unroll
    for (int a = 0; a < 3; ++a) {
        const unsigned int lo = __reduce_min_sync(0xffffffffu, klo[a]), hi = __reduce_max_sync(0xffffffffu, khi[a]);
        if ((threadIdx.x & 31) == 0) { atomicMin(&s_lo[a], lo); atomicMax(&s_hi[a], hi); }
    }
    __syncthreads();
    if (threadIdx.x < 3) { atomicMin(&bounds[threadIdx.x], s_lo[threadIdx.x]); atomicMax(&bounds[3 + threadIdx.x], s_hi[threadIdx.x]); }
}

__device__ __forceinline__ float key_to_float(unsigned int k) { return as_f((int)((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k)); }

__global__ void k_morton(const float4* __restrict__ leaves, long long n, const unsigned int* __restrict__ bounds,
                         unsigned long long* __restrict__ codes, unsigned int* __restrict__ vals)
{
    long long i = (long long)blockIdx.x * TB + threadIdx.x;
    if (i >= n) return;
    const float4 mn = leaves[8 * i], mx = leaves[8 * i + 1];
    const float c[3] = {rule_centre(mn.x, mx.x), rule_centre(mn.y, mx.y), rule_centre(mn.z, mx.z)};
    uint32_t cell[3];
    #pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float lo = key_to_float(bounds[a]), hi = key_to_float(bounds[3 + a]);
        cell[a] = rule_cell(c[a], lo, rule_scale(lo, hi));
    }
    codes[i] = rule_morton(cell[0], cell[1], cell[2]);
    vals[i] = (unsigned int)i;
}

// internal node i: children + parent links (link = parent * 2 + side)
__global__ void k_karras(const unsigned long long* __restrict__ codes, const unsigned int* __restrict__ vals, long long n,
                         float4* __restrict__ nodes, int* __restrict__ node_parent, int* __restrict__ leaf_parent)
{
    long long i = (long long)blockIdx.x * TB + threadIdx.x;
    if (i >= n - 1) return;
    int64_t first, last, split;
    rule_karras((const uint64_t*)codes, n, i, first, last, split);
    int lc, rc;
    if (split == first) { lc = ~(int)vals[split]; leaf_parent[split] = (int)i * 2; } else { lc = (int)split; node_parent[split] = (int)i * 2; }
    if (split + 1 == last) { rc = ~(int)vals[split + 1]; leaf_parent[split + 1] = (int)i * 2 + 1; } else { rc = (int)(split + 1); node_parent[split + 1] = (int)i * 2 + 1; }
    nodes[4 * i + 3] = make_float4(as_f(lc), as_f(rc), 0.0f, 0.0f);
    if (i == 0) node_parent[0] = -1;
}

struct Box6 { float lx, hx, ly, hy, lz, hz; };

__device__ __forceinline__ void store_child_box(float4* node, int side, const Box6& b)
{
    node[side] = make_float4(b.lx, b.hx, b.ly, b.hy);
    float2* z = reinterpret_cast<float2*>(node + 2) + side;
    *z = make_float2(b.lz, b.hz);
}

// bottom-up: the second thread to reach a node unions the two child boxes (always child 0 first) and carries on
__global__ void k_refit(const unsigned int* __restrict__ vals, const float4* __restrict__ leaves, long long n, float4* nodes,
                        const int* __restrict__ node_parent, const int* __restrict__ leaf_parent, unsigned int* __restrict__ arrived)
{
    long long q = (long long)blockIdx.x * TB + threadIdx.x;
    if (q >= n) return;
    const size_t leaf = vals[q];
    const float4 mn = leaves[8 * leaf], mx = leaves[8 * leaf + 1];
    Box6 b = {mn.x, mx.x, mn.y, mx.y, mn.z, mx.z};
    int link = leaf_parent[q];
    while (link >= 0) {
        const int P = link >> 1, side = link & 1;
        float4* node = nodes + 4 * (size_t)P;
        store_child_box(node, side, b);
        __threadfence();
        if (atomicAdd(&arrived[P], 1u) == 0u) return;
        const float4 o = __ldcg(node + (side ^ 1));
        const float2 oz = __ldcg(reinterpret_cast<const float2*>(node + 2) + (side ^ 1));
        Box6 c0 = b, c1 = {o.x, o.y, o.z, o.w, oz.x, oz.y};
        if (side == 1) { Box6 t = c0; c0 = c1; c1 = t; }
        b.lx = cmin(c0.lx, c1.lx); b.hx = cmax(c0.hx, c1.hx);
        b.ly = cmin(c0.ly, c1.ly); b.hy = cmax(c0.hy, c1.hy);
        b.lz = cmin(c0.lz, c1.lz); b.hz = cmax(c0.hz, c1.hz);
        link = node_parent[P];
    }
}

__global__ void k_depth(long long n, const int* __restrict__ node_parent, const int* __restrict__ leaf_parent, int* __restrict__ depth)
{
    const long long q = (long long)blockIdx.x * TB + threadIdx.x;
    int d = 0;
    if (q < n) for (int link = leaf_parent[q]; link >= 0; link = node_parent[link >> 1]) ++d;
    d = __reduce_max_sync(0xffffffffu, d);
    if ((threadIdx.x & 31) == 0) atomicMax(depth, d);
}

// ---- 8-wide collapse, one level: item i (a node of the two-child tree) becomes wide node base + i ------------------------------
// k_wide_level packs the node (boxes, slots) and leaves the two-child references of its inner children in child[] together with
// their count; after an exclusive scan of the counts k_wide_link turns them into the indices of the next level and lists them.
__global__ void k_wide_level(const float4* __restrict__ fast, const int* __restrict__ items, int n_items, int base, WNode* __restrict__ wide,
                             int* __restrict__ inner_count, int* __restrict__ flags)
{
    const int i = blockIdx.x * TB + threadIdx.x;
    if (i >= n_items) return;
    WideItem w; rule_wide_gather(fast, items[i], w);
    WNode node; int slot_of[8];
    if (!rule_wide_pack(w, node, slot_of)) atomicOr(flags + 1, 1);
    int by_slot[8];
#pragma unroll
    for (int s = 0; s < 8; ++s) by_slot[s] = 0x7fffffff;
    for (int c = 0; c < w.count; ++c) {
#pragma unroll
        for (int s = 0; s < 8; ++s) if (slot_of[c] == s) by_slot[s] = w.child[c];
    }
    int k = 0;
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        node.child[s] = by_slot[s] == 0x7fffffff ? -1 : by_slot[s];
        if (by_slot[s] != 0x7fffffff && by_slot[s] >= 0) ++k;
    }
    inner_count[i] = k;
    wide[(size_t)base + (size_t)i] = node;
}
__global__ void k_wide_link(const int* __restrict__ first, int n_items, int base, int next_base, WNode* __restrict__ wide, int* __restrict__ next_items)
{
    const int i = blockIdx.x * TB + threadIdx.x;
    if (i >= n_items) return;
    WNode* node = wide + (size_t)base + (size_t)i;
    const unsigned int valid = node->meta >> 24;
    int k = first[i];
    for (int s = 0; s < 8; ++s) {
        const int ch = node->child[s];
        if (((valid >> s) & 1u) && ch >= 0) { next_items[k] = ch; node->child[s] = next_base + k; ++k; }
    }
}

struct Tmp {                          // scratch of one build: pool blocks, given back once the stream has drained
    std::vector<std::pair<void*, size_t>> ptrs; int device = 0; cudaStream_t st = nullptr;
    ~Tmp() { if (!ptrs.empty()) cudaStreamSynchronize(st); for (auto& b : ptrs) pool_give(device, b.second, b.first); }
    template <class T> bool get(T*& p, size_t count)
    {
        void* v = nullptr; const size_t bytes = (count ? count : 1) * sizeof(T);
        if (pool_alloc(device, bytes, &v) != cudaSuccess) { p = nullptr; return false; }
        ptrs.emplace_back(v, bytes); p = (T*)v; return true;
    }
};

inline double now_ms()
{
    using namespace std::chrono;
    return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}
inline int bits_for(uint64_t v) { int b = 0; while (v) { ++b; v >>= 1; } return b; }

} // namespace

#define GB_CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { err = std::string(#x) + ": " + cudaGetErrorString(e_); return false; } } while (0)

bool build_mesh_gpu(const MeshSource& src, int64_t n_tri, uint64_t seed, int shape_id,
                    cudaStream_t st, std::vector<void*>& allocs, std::unordered_map<void*, size_t>& alloc_bytes, size_t& bytes, GpuMeshBuild& out,
                    std::string& err, uint64_t* launches)
{
    out = GpuMeshBuild();
    if (n_tri <= 0) { out.mesh.ref_root = -1; return true; }
    if (n_tri >= (1ll << 30)) { out.fallback = true; out.why = "more than 2^30 triangles"; return true; }
    const double t_begin = now_ms();
    const size_t n = (size_t)n_tri;
    TreeShape ts; tree_shape_fill(ts, (uint32_t)n);
    const size_t n_nodes = ts.nodes[0][0], n_leaves = ts.bottoms[0][0];

    int device = 0; cudaGetDevice(&device);
    Tmp tmp; tmp.device = device; tmp.st = st;
    const size_t allocs_at_entry = allocs.size(); const size_t bytes_at_entry = bytes;
    auto drop_results = [&]() {
        cudaStreamSynchronize(st);
        while (allocs.size() > allocs_at_entry) { void* p = allocs.back(); pool_give(device, alloc_bytes[p], p); alloc_bytes.erase(p); allocs.pop_back(); }
        bytes = bytes_at_entry;
    };
    float *d_pos, *d_nrm, *d_uv; float4 *d_tmin, *d_tmax; unsigned int *d_order[2]; unsigned long long* d_keys[2];
    uint8_t* d_axis; int* d_flags; unsigned int* d_idx = nullptr;
    const size_t n_in = src.indexed ? (size_t)src.n_vert : n * 3;        // vertices that cross PCIe
    if (src.indexed && !tmp.get(d_idx, n * 3)) { err = "out of device memory (index buffer)"; return false; }
    if (!tmp.get(d_pos, n_in * 3) || !tmp.get(d_nrm, n_in * 3) || !tmp.get(d_uv, n_in * 2) || !tmp.get(d_tmin, n) || !tmp.get(d_tmax, n) ||
        !tmp.get(d_order[0], n) || !tmp.get(d_order[1], n) || !tmp.get(d_keys[0], n) || !tmp.get(d_keys[1], n) ||
        !tmp.get(d_axis, n_nodes) || !tmp.get(d_flags, 16)) { err = "out of device memory (BVH build scratch)"; return false; }
    // results (owned by the scene)
    float4 *d_tris = nullptr, *d_attrs = nullptr, *d_ref = nullptr, *d_leaves = nullptr, *d_fast = nullptr;
    auto keep = [&](float4*& p, size_t count) {
        void* v = nullptr; const size_t nb = (count ? count : 1) * sizeof(float4);
        if (pool_alloc(device, nb, &v) != cudaSuccess) return false;
        allocs.push_back(v); alloc_bytes[v] = nb; bytes += count * sizeof(float4); p = (float4*)v; return true;
    };
    if (!keep(d_tris, n * 3) || !keep(d_attrs, n * 4) || !keep(d_ref, n_nodes * 2) || !keep(d_leaves, n_leaves * 8) || !keep(d_fast, n_leaves > 1 ? (n_leaves - 1) * 4 : 0)) { err = "out of device memory (mesh arrays)"; return false; }

    GB_CK(cudaMemcpyAsync(d_pos, src.pos, n_in * 3 * sizeof(float), cudaMemcpyHostToDevice, st));
    GB_CK(cudaMemcpyAsync(d_nrm, src.nrm, n_in * 3 * sizeof(float), cudaMemcpyHostToDevice, st));
    GB_CK(cudaMemcpyAsync(d_uv, src.uv, n_in * 2 * sizeof(float), cudaMemcpyHostToDevice, st));
    if (src.indexed) GB_CK(cudaMemcpyAsync(d_idx, src.idx, n * 3 * sizeof(unsigned int), cudaMemcpyHostToDevice, st));
    out.h2d_bytes = n_in * 8 * sizeof(float) + (src.indexed ? n * 3 * sizeof(unsigned int) : 0);
    GB_CK(cudaMemsetAsync(d_flags, 0, 16 * sizeof(int), st));
    GB_CK(cudaStreamSynchronize(st));
    out.ms_upload = now_ms() - t_begin;
    const double t_build = now_ms();
    uint64_t nl = 0;

    if (src.indexed) k_mesh_pack<true><<<grid_for((long long)n), TB, 0, st>>>(d_pos, d_nrm, d_uv, d_idx, (long long)n, d_tris, d_attrs, d_tmin, d_tmax, d_order[0], d_flags);
    else k_mesh_pack<false><<<grid_for((long long)n), TB, 0, st>>>(d_pos, d_nrm, d_uv, nullptr, (long long)n, d_tris, d_attrs, d_tmin, d_tmax, d_order[0], d_flags);
    ++nl;
    const uint64_t stream0 = rule_mix64(seed ^ rule_mix64((uint64_t)shape_id));
    k_node_axes<<<grid_for((long long)n_nodes), TB, 0, st>>>(stream0, (long long)n_nodes, d_axis); ++nl;

    // reference topology: one stable sort per level that still has a range of more than two triangles
    size_t sort_bytes = 0;
    {
        cub::DoubleBuffer<unsigned long long> kb(d_keys[0], d_keys[1]); cub::DoubleBuffer<unsigned int> vb(d_order[0], d_order[1]);
        GB_CK(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, kb, vb, (int)n, 0, 64, st));
    }
    unsigned char* d_sort = nullptr; if (!tmp.get(d_sort, sort_bytes)) { err = "out of device memory (sort scratch)"; return false; }
    int cur = 0;
    for (int level = 0; level < ts.levels; ++level) {
        if (((n + ((size_t)1 << level) - 1) >> level) <= 2) break;
        const size_t smallest = n >> level;                               // every range of this level holds >= n >> level positions
        const int shift = smallest > 1 ? bits_for(smallest) - 1 : 0;      // floor(log2(smallest))
        const int key_bits = 32 + bits_for((n - 1) >> shift);
        k_level_keys<<<grid_for((long long)n), TB, 0, st>>>(ts, level, shift, d_order[cur], d_tmin, d_axis, d_keys[0]); ++nl;
        cub::DoubleBuffer<unsigned long long> kb(d_keys[0], d_keys[1]);
        cub::DoubleBuffer<unsigned int> vb(d_order[cur], d_order[cur ^ 1]);
        GB_CK(cub::DeviceRadixSort::SortPairs(d_sort, sort_bytes, kb, vb, (int)n, 0, key_bits, st)); 
        if (vb.Current() != d_order[cur]) cur ^= 1;
    }
    for (int level = ts.levels - 1; level >= 0; --level) {
        const long long ranges = 1ll << level;
        k_emit_level<<<grid_for(ranges), TB, 0, st>>>(ts, level, d_order[cur], d_tmin, d_tmax, d_tris, d_axis, d_ref, d_leaves); ++nl;
    }

    // Morton LBVH over the bottom nodes
    int fast_root = ~0, depth = 0;
    int last_tri_r = -1;
    if (n_leaves > 1) {
        const long long m = (long long)n_leaves;
        unsigned int* d_bounds; int *d_node_parent, *d_leaf_parent, *d_depth; unsigned int* d_arrived;
        if (!tmp.get(d_bounds, 8) || !tmp.get(d_node_parent, n_leaves) || !tmp.get(d_leaf_parent, n_leaves) || !tmp.get(d_arrived, n_leaves) || !tmp.get(d_depth, 4)) { err = "out of device memory (LBVH scratch)"; return false; }
        const unsigned int init_bounds[8] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u, 0u, 0u};
        GB_CK(cudaMemcpyAsync(d_bounds, init_bounds, sizeof(init_bounds), cudaMemcpyHostToDevice, st));
        GB_CK(cudaMemsetAsync(d_arrived, 0, n_leaves * sizeof(unsigned int), st));
        GB_CK(cudaMemsetAsync(d_depth, 0, 4 * sizeof(int), st));
        k_centre_bounds<<<grid_for(m), TB, 0, st>>>(d_leaves, m, d_bounds); ++nl;
        // keys / order buffers of the level sorts are reused (n_leaves <= n)
        k_morton<<<grid_for(m), TB, 0, st>>>(d_leaves, m, d_bounds, d_keys[0], d_order[0]); ++nl;
        cub::DoubleBuffer<unsigned long long> kb(d_keys[0], d_keys[1]); cub::DoubleBuffer<unsigned int> vb(d_order[0], d_order[1]);
        GB_CK(cub::DeviceRadixSort::SortPairs(d_sort, sort_bytes, kb, vb, (int)m, 0, 63, st)); 
        k_karras<<<grid_for(m - 1), TB, 0, st>>>(kb.Current(), vb.Current(), m, d_fast, d_node_parent, d_leaf_parent); ++nl;
        k_refit<<<grid_for(m), TB, 0, st>>>(vb.Current(), d_leaves, m, d_fast, d_node_parent, d_leaf_parent, d_arrived); ++nl;
        k_depth<<<grid_for(m), TB, 0, st>>>(m, d_node_parent, d_leaf_parent, d_depth); ++nl;
        GB_CK(cudaMemcpyAsync(&depth, d_depth, sizeof(int), cudaMemcpyDeviceToHost, st));
        fast_root = 0;
    }
    // 8-wide collapse of the fast tree, breadth first (hrt_bvh_rules.h (4)); one small read back per level (the size of the next one)
    WNode* d_wide = nullptr; size_t n_wide = 0; int wide_depth = 0;
    if (n_leaves > 1) {
        const size_t bound = n_leaves - 1;                                  // a wide node replaces at least one two-child node
        WNode* d_wtmp; int *d_items[2], *d_cnt, *d_first;
        if (!tmp.get(d_wtmp, bound) || !tmp.get(d_items[0], bound) || !tmp.get(d_items[1], bound) || !tmp.get(d_cnt, bound + 1) || !tmp.get(d_first, bound + 1)) { err = "out of device memory (wide collapse scratch)"; return false; }
        size_t scan_bytes = 0;
        GB_CK(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, d_cnt, d_first, (int)(bound + 1), st));
        unsigned char* d_scan = nullptr; if (!tmp.get(d_scan, scan_bytes)) { err = "out of device memory (scan scratch)"; return false; }
        GB_CK(cudaMemsetAsync(d_items[0], 0, sizeof(int), st));              // level 0: the root, two-child node 0
        int n_items = 1, base = 0, cur_items = 0;
        while (n_items > 0) {
            k_wide_level<<<grid_for(n_items), TB, 0, st>>>(d_fast, d_items[cur_items], n_items, base, d_wtmp, d_cnt, d_flags); ++nl;
            GB_CK(cub::DeviceScan::ExclusiveSum(d_scan, scan_bytes, d_cnt, d_first, n_items + 1, st));
            k_wide_link<<<grid_for(n_items), TB, 0, st>>>(d_first, n_items, base, base + n_items, d_wtmp, d_items[cur_items ^ 1]); ++nl;
            int total = 0;
            GB_CK(cudaMemcpyAsync(&total, d_first + n_items, sizeof(int), cudaMemcpyDeviceToHost, st));
            GB_CK(cudaStreamSynchronize(st));
            base += n_items; n_items = total; cur_items ^= 1; ++wide_depth;
        }
        n_wide = (size_t)base;
        float4* d_w4 = nullptr;
        if (!keep(d_w4, n_wide * 6)) { err = "out of device memory (wide nodes)"; return false; }
        d_wide = (WNode*)d_w4;
        GB_CK(cudaMemcpyAsync(d_wide, d_wtmp, n_wide * sizeof(WNode), cudaMemcpyDeviceToDevice, st));
    }
    int flags = 0, wide_bad = 0;
    GB_CK(cudaMemcpyAsync(&wide_bad, d_flags + 1, sizeof(int), cudaMemcpyDeviceToHost, st));
    GB_CK(cudaMemcpyAsync(&flags, d_flags, sizeof(int), cudaMemcpyDeviceToHost, st));
    GB_CK(cudaMemcpyAsync(&last_tri_r, reinterpret_cast<const int*>(d_leaves + 8 * (n_leaves - 1) + 1) + 3, sizeof(int), cudaMemcpyDeviceToHost, st));
    GB_CK(cudaStreamSynchronize(st));
    GB_CK(cudaGetLastError());
    if (launches) *launches += nl;
    out.ms_build = now_ms() - t_build;
    if (flags) { drop_results(); out.fallback = true; out.why = "non-finite vertex positions"; return true; }
    if (depth > FAST_MAX_DEPTH) { drop_results(); out.fallback = true; out.why = "LBVH deeper than the traversal stack allows"; return true; }

    DMesh& m = out.mesh;
    m.ref_nodes = d_ref; m.tris = d_tris; m.attrs = d_attrs; m.fast_nodes = d_fast; m.fast_leaves = d_leaves;
    m.wide_nodes = (wide_bad || n_wide >= (size_t(1) << 24)) ? nullptr : d_wide; m.wide_root = fast_root;      // wide_bad: boxes off the 8-bit grids - the scene keeps to the two-child walk
    m.ref_root = 0; m.fast_root = fast_root; m.n_tri = n_tri;
    out.n_wide_nodes = n_wide; out.wide_depth = wide_depth;
    out.n_ref_nodes = n_nodes; out.n_fast_nodes = n_leaves > 1 ? n_leaves - 1 : 0; out.n_leaves = n_leaves;
    out.last_tri_r = last_tri_r; out.depth = depth;
    return true;
}

} // namespace hrt
