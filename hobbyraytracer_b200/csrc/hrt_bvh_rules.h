// hrt_bvh_rules.h — the rules that make the GPU BVH builder (hrt_bvh_gpu.cu, HRT_BVH_LBVH_GPU) and its host mirror
// (hrt_host.cpp, HRT_BVH_LBVH_HOST) produce the same arrays bit for bit.  Everything here is integer arithmetic or
// single IEEE operations (the library is compiled without FMA contraction on both sides), shared as source.
//
// (1) Reference topology (bvh.cpp:6-61): a range of n > 2 triangles is ordered by box.min[axis] and cut at n/2, so the
//     SHAPE of the tree (range starts, sizes, pre-order node ids) depends on the triangle count alone: the ranges of
//     level l have floor(N/2^l) or ceil(N/2^l) items.  `TreeShape` tabulates, per level, the node count and the number
//     of bottom nodes of a sub-tree of either size; `descend` walks from the root to the range that holds a position.
//     The reference orders with std::sort, whose treatment of equal keys is libstdc++'s introsort; the GPU rule is the
//     algorithm-independent one: equal keys keep their previous order (stable).  Both rules agree when keys differ.
// (2) The split axis of node k is glm::linearRand(0,2) drawn at position 4k of the mesh's rand() stream (one
//     linearRand<int> = four rand() bytes, random.inl:19-27,140-147).
// (3) Fast tree of HRT_BVH_LBVH_*: 63-bit Morton codes (21 bits per axis) of the bottom nodes' box centres, ordered by
//     (code, bottom-node index); hierarchy by the longest-common-prefix rule of Karras 2012 with the position as the
//     tie-breaker; internal node i of the hierarchy is node i, the root is node 0.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define HRT_RULE_HD __host__ __device__ inline
#else
#define HRT_RULE_HD inline
#endif

namespace hrt {

constexpr int TREE_MAX_LEVELS = 34;

struct TreeShape {
    uint32_t n;                               // triangles
    int levels;                               // nodes exist on levels 0 .. levels-1
    uint32_t lo[TREE_MAX_LEVELS];             // smaller range size of level l (the other one is lo + 1)
    uint32_t nodes[TREE_MAX_LEVELS][2];       // nodes of a sub-tree over lo / lo + 1 items
    uint32_t bottoms[TREE_MAX_LEVELS][2];     // bottom nodes (ranges of 1 or 2 items) of such a sub-tree
};

inline void tree_shape_fill(TreeShape& ts, uint32_t n)
{
    ts.n = n; ts.levels = 0;
    int L = 0;
    for (; L < TREE_MAX_LEVELS; ++L) { ts.lo[L] = n >> L; if ((n >> L) <= 1) { ++L; break; } }   // the last level only has ranges of <= 2
    ts.levels = L;
    for (int l = L - 1; l >= 0; --l)
        for (int k = 0; k < 2; ++k) {
            uint32_t m = ts.lo[l] + (uint32_t)k;
            if (m <= 2) { ts.nodes[l][k] = 1; ts.bottoms[l][k] = 1; continue; }
            uint32_t a = m / 2, b = m - m / 2;              // both are sizes of level l + 1
            ts.nodes[l][k] = 1 + ts.nodes[l + 1][a - ts.lo[l + 1]] + ts.nodes[l + 1][b - ts.lo[l + 1]];
            ts.bottoms[l][k] = ts.bottoms[l + 1][a - ts.lo[l + 1]] + ts.bottoms[l + 1][b - ts.lo[l + 1]];
        }
}

struct TreeRange { uint32_t start, n, self, bottoms_before; int level; };

// The range of level `level` that holds position p, or (if the walk reaches a range of <= 2 items earlier) that
// bottom range; r.level tells which.
HRT_RULE_HD TreeRange tree_descend(const TreeShape& ts, uint32_t p, int level)
{
    TreeRange r; r.start = 0; r.n = ts.n; r.self = 0; r.bottoms_before = 0; r.level = 0;
    while (r.level < level && r.n > 2) {
        const int l1 = r.level + 1;
        const uint32_t nl = r.n >> 1, k = nl - ts.lo[l1];
        if (p < r.start + nl) { r.n = nl; r.self += 1; }
        else { r.start += nl; r.n -= nl; r.self += 1 + ts.nodes[l1][k]; r.bottoms_before += ts.bottoms[l1][k]; }
        r.level = l1;
    }
    return r;
}

// ---- rand() stream of the split axes (shared with the test oracle) -------------------------------------------
constexpr uint64_t RULE_LCG_A = 6364136223846793005ULL, RULE_LCG_C = 1442695040888963407ULL;
HRT_RULE_HD uint64_t rule_mix64(uint64_t z)
{
    z += 0x9e3779b97f4a7c15ULL;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
    return z ^ (z >> 31);
}
HRT_RULE_HD int rule_axis_of_node(uint64_t stream0, uint64_t node)
{
    uint64_t k = 4 * node, a = RULE_LCG_A, c = RULE_LCG_C, A = 1, C = 0;
    while (k) {
        if (k & 1) { A = A * a; C = C * a + c; }
        c = (a + 1) * c; a = a * a; k >>= 1;
    }
    uint64_t s = A * stream0 + C;
    uint32_t v = 0;
    for (int i = 0; i < 4; ++i) {
        s = s * RULE_LCG_A + RULE_LCG_C;
        uint32_t byte = (uint32_t)((int)((s >> 33) & 0x7fffffffULL) % 255);
        v |= byte << (8 * i);
    }
    return (int)(v % 3u);
}

// ---- ordering keys ----------------------------------------------------------------------------------------
// unsigned image of a float under which `<` on floats and `<` on the images agree (-0 is taken as +0)
HRT_RULE_HD uint32_t rule_float_key(uint32_t bits)
{
    if ((bits << 1) == 0) bits = 0;
    return (bits & 0x80000000u) ? ~bits : (bits | 0x80000000u);
}
HRT_RULE_HD uint64_t rule_spread21(uint32_t v)
{
    uint64_t x = v & 0x1fffffu;
    x = (x | (x << 32)) & 0x1f00000000ffffULL;
    x = (x | (x << 16)) & 0x1f0000ff0000ffULL;
    x = (x | (x << 8)) & 0x100f00f00f00f00fULL;
    x = (x | (x << 4)) & 0x10c30c30c30c30c3ULL;
    x = (x | (x << 2)) & 0x1249249249249249ULL;
    return x;
}
// centre of a box along one axis, its cell on the 2^21 grid spanned by [lo, lo + extent]
HRT_RULE_HD float rule_centre(float mn, float mx) { return 0.5f * mn + 0.5f * mx; }
HRT_RULE_HD uint32_t rule_cell(float c, float lo, float scale)
{
    float f = (c - lo) * scale;
    if (!(f > 0.0f)) return 0u;
    if (f >= 2097151.0f) return 2097151u;
    return (uint32_t)f;
}
HRT_RULE_HD float rule_scale(float lo, float hi) { return (hi > lo) ? 2097152.0f / (hi - lo) : 0.0f; }
HRT_RULE_HD uint64_t rule_morton(uint32_t x, uint32_t y, uint32_t z) { return (rule_spread21(x) << 2) | (rule_spread21(y) << 1) | rule_spread21(z); }

HRT_RULE_HD int rule_clz64(uint64_t v)
{
#if defined(__CUDA_ARCH__)
    return __clzll((long long)v);
#else
    return v ? __builtin_clzll(v) : 64;
#endif
}
// common-prefix length of the sorted keys at positions i and j (Karras 2012, with the position as tie-breaker)
HRT_RULE_HD int rule_delta(const uint64_t* codes, int64_t n, int64_t i, int64_t j)
{
    if (j < 0 || j >= n) return -1;
    const uint64_t a = codes[i], b = codes[j];
    if (a != b) return rule_clz64(a ^ b);
    return 64 + rule_clz64((uint64_t)i ^ (uint64_t)j);
}

// Range [first, last] of sorted positions covered by internal node i, and its split position (the left child covers
// [first, split], the right one [split + 1, last]).
HRT_RULE_HD void rule_karras(const uint64_t* codes, int64_t n, int64_t i, int64_t& first, int64_t& last, int64_t& split)
{
    const int d = (rule_delta(codes, n, i, i + 1) - rule_delta(codes, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = rule_delta(codes, n, i, i - d);
    int64_t lmax = 2;
    while (rule_delta(codes, n, i, i + lmax * d) > dmin) lmax *= 2;
    int64_t l = 0;
    for (int64_t t = lmax / 2; t >= 1; t /= 2)
        if (rule_delta(codes, n, i, i + (l + t) * d) > dmin) l += t;
    const int64_t j = i + l * d;
    const int dnode = rule_delta(codes, n, i, j);
    int64_t s = 0;
    for (int64_t div = 2, t = (l + div - 1) / div;; div *= 2, t = (l + div - 1) / div) {
        if (rule_delta(codes, n, i, i + (s + t) * d) > dnode) s += t;
        if (t <= 1) break;
    }
    split = i + s * d + (d < 0 ? -1 : 0);
    first = d > 0 ? i : j; last = d > 0 ? j : i;
}

} // namespace hrt
