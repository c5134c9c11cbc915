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
#include <math.h>
#include <stddef.h>
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

// ---- (4) 8-wide collapse of a two-child tree (hrt_types.h, WNode) ------------------------------------------------------
// Shared by the host builder (hrt_host.cpp, build_wide) and the device builder (hrt_bvh_gpu.cu, k_wide_level) so that both
// produce the same bytes.  A wide node is made from one node of the two-child tree (`fast_nodes` format: both children's
// boxes stored in the parent): starting from its two children, the inner child with the largest half area is replaced by
// its own two children until there are eight or no inner child is left (ties: the earlier position).  The children are
// then put into slots by a greedy assignment on the alignment of (child centre - node centre) with the slot's octant
// direction, and their boxes are quantised outwards on the node's 8-bit grid.  Wide nodes are numbered breadth first: the
// inner children of the nodes of one level, taken in node order and slot order, are the nodes of the next level.
namespace hrt {

struct WideItem { int child[8]; float lo[8][3], hi[8][3]; int count; };

HRT_RULE_HD float rule_half_area(const float* lo, const float* hi)
{
    const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    const float a = dx * dy, b = dy * dz, c = dz * dx;
    return (a + b) + c;
}
HRT_RULE_HD int rule_f2i(float f) { union { float f; int i; } u; u.f = f; return u.i; }
template <class F4>
HRT_RULE_HD void rule_node_children(const F4* nodes, int n, int& c0, int& c1, float* lo0, float* hi0, float* lo1, float* hi1)
{
    const F4 n0 = nodes[4 * (size_t)n], n1 = nodes[4 * (size_t)n + 1], n2 = nodes[4 * (size_t)n + 2], n3 = nodes[4 * (size_t)n + 3];
    lo0[0] = n0.x; hi0[0] = n0.y; lo0[1] = n0.z; hi0[1] = n0.w; lo0[2] = n2.x; hi0[2] = n2.y;
    lo1[0] = n1.x; hi1[0] = n1.y; lo1[1] = n1.z; hi1[1] = n1.w; lo1[2] = n2.z; hi1[2] = n2.w;
    c0 = rule_f2i(n3.x); c1 = rule_f2i(n3.y);
}
template <class F4>
HRT_RULE_HD void rule_wide_gather(const F4* nodes, int n, WideItem& w)
{
    rule_node_children(nodes, n, w.child[0], w.child[1], w.lo[0], w.hi[0], w.lo[1], w.hi[1]);
    w.count = 2;
    while (w.count < 8) {
        int best = -1; float best_area = -1.0f;
        for (int i = 0; i < w.count; ++i) {
            if (w.child[i] < 0) continue;
            const float a = rule_half_area(w.lo[i], w.hi[i]);
            if (best < 0 || a > best_area) { best = i; best_area = a; }
        }
        if (best < 0) break;
        const int m = w.child[best];
        rule_node_children(nodes, m, w.child[best], w.child[w.count], w.lo[best], w.hi[best], w.lo[w.count], w.hi[w.count]);
        ++w.count;
    }
}

constexpr int WIDE_EXP_MIN = 27, WIDE_EXP_MAX = 239;      // biased exponents of the grid step: 2^-100 .. 2^112

// Fills p, meta and q of `out` and leaves the slot of every gathered child in slot_of[]; out.child[] is left to the caller
// (it depends on the numbering).  Returns false when the node box cannot be represented (non-finite or > 2^119 wide).
template <class WN>
HRT_RULE_HD bool rule_wide_pack(const WideItem& w, WN& out, int* slot_of)
{
    float lo[3], hi[3];
    for (int a = 0; a < 3; ++a) {
        lo[a] = w.lo[0][a]; hi[a] = w.hi[0][a];
        for (int i = 1; i < w.count; ++i) { if (w.lo[i][a] < lo[a]) lo[a] = w.lo[i][a]; if (w.hi[i][a] > hi[a]) hi[a] = w.hi[i][a]; }
    }
    bool ok = true;
    int eb[3]; double step[3];
    for (int a = 0; a < 3; ++a) {
        const double ext = (double)hi[a] - (double)lo[a];
        if (!(ext >= 0.0) || !(ext < 1.0e36) || !(lo[a] > -1.0e36f) || !(lo[a] < 1.0e36f)) { ok = false; eb[a] = WIDE_EXP_MAX; step[a] = 1.0; continue; }
        int k = WIDE_EXP_MIN - 127;
        if (ext > 0.0) { int e2; (void)frexp(ext / 255.0, &e2); k = e2; }     // 2^k >= ext / 255
        if (k < WIDE_EXP_MIN - 127) k = WIDE_EXP_MIN - 127;
        if (k > WIDE_EXP_MAX - 127) { ok = false; k = WIDE_EXP_MAX - 127; }
        eb[a] = k + 127; step[a] = ldexp(1.0, k);
        out.p[a] = lo[a];
    }
    // slots: greedy on the alignment of the child's centre offset with the slot's octant direction
    float off[8][3];
    for (int i = 0; i < w.count; ++i)
        for (int a = 0; a < 3; ++a) {
            const float cn = 0.5f * lo[a] + 0.5f * hi[a], cc = 0.5f * w.lo[i][a] + 0.5f * w.hi[i][a];
            off[i][a] = cc - cn;
        }
    unsigned used_slots = 0, placed = 0;
    for (int round = 0; round < w.count; ++round) {
        int bi = -1, bs = -1; float bc = 0.0f;
        for (int i = 0; i < w.count; ++i) {
            if (placed & (1u << i)) continue;
            for (int s = 0; s < 8; ++s) {
                if (used_slots & (1u << s)) continue;
                const float x = (s & 1) ? off[i][0] : -off[i][0], y = (s & 2) ? off[i][1] : -off[i][1], z = (s & 4) ? off[i][2] : -off[i][2];
                const float c = (x + y) + z;
                if (bi < 0 || c > bc) { bi = i; bs = s; bc = c; }
            }
        }
        slot_of[bi] = bs; placed |= 1u << bi; used_slots |= 1u << bs;
    }
    out.meta = (uint32_t)eb[0] | ((uint32_t)eb[1] << 8) | ((uint32_t)eb[2] << 16) | (used_slots << 24);
    for (int a = 0; a < 3; ++a) for (int s = 0; s < 8; ++s) { out.q[2 * a][s] = 255; out.q[2 * a + 1][s] = 0; }     // empty slots: inverted box
    for (int i = 0; i < w.count; ++i) {
        const int s = slot_of[i];
        for (int a = 0; a < 3; ++a) {
            const double p = (double)lo[a];
            double fl = floor(((double)w.lo[i][a] - p) / step[a]), fh = ceil(((double)w.hi[i][a] - p) / step[a]);
            if (!(fl >= 0.0)) fl = 0.0;
            if (fl > 255.0) fl = 255.0;
            if (!(fh >= 0.0)) fh = 0.0;
            if (fh > 255.0) fh = 255.0;
            int ql = (int)fl, qh = (int)fh;
            while (ql > 0 && p + (double)ql * step[a] > (double)w.lo[i][a]) --ql;
            while (qh < 255 && p + (double)qh * step[a] < (double)w.hi[i][a]) ++qh;
            if (p + (double)qh * step[a] < (double)w.hi[i][a]) ok = false;
            out.q[2 * a][s] = (uint8_t)ql; out.q[2 * a + 1][s] = (uint8_t)qh;
        }
    }
    return ok;
}

} // namespace hrt
