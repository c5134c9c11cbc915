// hrt_wide.cuh — pass 1 of the ordered closest-hit traversal (hrt_trace.cuh) over the 8-wide quantised trees
// (hrt_types.h, WNode), written for warps of incoherent rays.
//
// What the r2a profiles of the two-child walk said (profiles/README.md): 4.8 of 32 lanes active per instruction in the ray
// batch kernel and 5.6 in the secondary-ray kernel although idle lanes were refilled — inside a warp the lanes need
// different numbers of node steps between two leaves and the while-while loop runs every lane's node loop to the longest
// one; the walk state lived in local memory (7 % of the executed instructions were LDL / STL); a node step was four
// scattered 128-bit loads and a leaf eight (the L1 wavefront pipe 50 % busy), and a ray needed ~50 dependent node fetches.
// Here:
//   * a node has eight children with boxes quantised to 8 bits on a per-node grid: 96 bytes, three 256-bit loads, one
//     dependent fetch per ~3 levels of the two-child tree.  Boxes only grow by quantisation and the slab arithmetic is
//     widened by more than its rounding error, so the walk visits a superset of what the exact boxes admit: everything
//     hrt_trace.cuh says about pass 1 (candidates, ambiguity, pass 2 + replay) holds unchanged;
//   * a leaf visit reads the last 96 bytes of the leaf record (three 256-bit loads); the reference node's box only when a
//     triangle accepted;
//   * the traversal stack is in shared memory (one column per thread), all other state in registers;
//   * a warp advances in PHASES: "node step", "triangle leaf", "world-list entry".  A phase runs when at least
//     WIDE_PHASE_MIN lanes wait for it (or nothing else can run), so no instruction of the hot loop executes for one lane
//     while thirty-one wait; lanes whose walk ended take new rays as soon as fewer than WIDE_REFILL lanes are walking.
// Rays whose answer needs pass 2 (ties, badly conditioned triangles, NaNs, stack overflow) are not resolved here: the
// caller records them and a second, small kernel runs trace_resolve() on that list (hrt_kernels.cu, k_resolve_*).
#pragma once
#include "hrt_trace.cuh"

namespace hrt {

constexpr int WIDE_STACK = 32;                        // stack entries per lane (one per node with children left + one per mesh entered); deeper: pass 2
constexpr float WIDE_WIDEN = 9.5367431640625e-7f;     // 2^-20, relative to |(p - o) / d| + |grid extent / d| (see wide_hits)
constexpr int HIT_PENDING = 8;                        // Hit::flags: pass 1 could not decide; Hit::t carries the cull bound
constexpr int PEND_BAD = 16, PEND_SHEAR = 32;         // Hit::flags of a pending entry: arguments of trace_resolve
constexpr uint32_t WIDE_SLOW_BIT = 1u << 29;          // top-level leaf: an entry that takes the general, out-of-line object test (wrapped shapes, boxes, media):
                                                      // those lanes get a phase of their own, so the large cold function runs for many lanes at a time
constexpr uint32_t WIDE_TAG_BITS = (1u << 30) | (1u << 29);
constexpr uint32_t WIDE_MESH_BIT = 1u << 30;          // top-level leaf ~(entry | WIDE_MESH_BIT): the entry is a mesh that gets ENTERED (no medium),
                                                      // so a warp can tell the lanes that wait for a mesh entry from those that wait for
                                                      // a sphere / rect / box / medium test without a load (set by flatten(), hrt_host.cpp)

struct __align__(32) W32 { uint32_t w[8]; };

HRT_HD W32 wide_ld32(const void* p)
{
    W32 r;
#if defined(__CUDA_ARCH__)
    asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]), "=r"(r.w[4]), "=r"(r.w[5]), "=r"(r.w[6]), "=r"(r.w[7]) : "l"(p));
#else
    __builtin_memcpy(&r, p, 32);
#endif
    return r;
}
HRT_HD void wide_prefetch(const void* p)          // the sector is wanted in L1 a few hundred cycles from now (a node's child ids)
{
#if defined(__CUDA_ARCH__)
    asm volatile("prefetch.global.L1 [%0];" :: "l"(p));
#else
    (void)p;
#endif
}
HRT_HD int wide_ld_child(const WNode* nodes, uint32_t node, uint32_t slot)
{
#if defined(__CUDA_ARCH__)
    return __ldg(&nodes[node].child[slot]);
#else
    return nodes[node].child[slot];
#endif
}
HRT_HD uint32_t wide_perm(uint32_t a, uint32_t b, uint32_t sel)
{
#if defined(__CUDA_ARCH__)
    return __byte_perm(a, b, sel);
#else
    const uint64_t v = ((uint64_t)b << 32) | a; uint32_t r = 0;
    for (int i = 0; i < 4; ++i) r |= (uint32_t)((v >> (8 * ((sel >> (4 * i)) & 7u))) & 0xffu) << (8 * i);
    return r;
#endif
}
HRT_HD float wide_f(uint32_t u) { return int_as_float((int)u); }
HRT_HD float4 wide_f4(uint32_t x, uint32_t y, uint32_t z) { float4 r; r.x = wide_f(x); r.y = wide_f(y); r.z = wide_f(z); r.w = 0.0f; return r; }

// The ray a tree level is walked with: origin, reciprocal direction, sign bits.
struct WRay {
    V3 o, id;
    uint32_t neg;       // bit a: d[a] < 0 (the child's hi plane is the near one)
};
// A ray is only walked when the reciprocals of its direction are finite (wide_begin, wide_mesh_step): no component is zero or NaN, so
// "d[a] < 0" is the sign bit and "d[a] > 0" its complement.
HRT_HD WRay wide_ray(V3 o, V3 d, V3 id)
{
    WRay q; q.o = o; q.id = id;
    q.neg = ((uint32_t)float_as_int(d.x) >> 31) | (((uint32_t)float_as_int(d.y) >> 31) << 1) | (((uint32_t)float_as_int(d.z) >> 31) << 2);
    return q;
}
HRT_HD uint32_t wide_oct(const WRay& q) { return q.neg ^ 7u; }    // bit a: d[a] > 0; slot ^ oct = 7 for the nearest octant, 0 for the farthest

// Slab test of the eight children of a node: bit s of the result is set when the ray MAY pass through slot s's box within
// [t_lo, bnd0].  c0 / c1 are the first 64 bytes of the node.
// Arithmetic.  A child plane on axis a is p + k * step (k = 0..255); its parameter is t = (p - o) * id + k * step * id.
// The byte k is placed in bits 8..15 of the float 1.0 (one PRMT), giving v = 1 + k * 2^-15 exactly, and t = fma(v, A, B) with
// A = step * 2^15 * id and B = (p - o) * id - A: one PRMT + one FMA per plane.  Rounding: (p - o) and its product with id
// (two roundings, relative to |(p - o) * id|), A (one), B (one, relative to |(p - o) id| + |A|), the fma (one, |t| <= the
// same sum): below 2^-22 * (|(p - o) id| + |A|) in all; the reference's own (plane - o) / d rounds twice.  Near planes are
// moved down and far planes up by WIDE_WIDEN (2^-20) times that sum: a quarter of a percent of a grid step plus 2^-20 of
// the distance to the grid origin.  NaNs (overflowing products) drop out of fminf / fmaxf and the final test is "not
// greater", so anything that cannot be computed counts as a hit.
HRT_HD uint32_t wide_hits(const W32& c0, const W32& c1, const WRay& q, float t_lo, float bnd0)
{
    const uint32_t meta = c0.w[3];
    const float Ax = wide_f(((meta & 0xffu) + 15u) << 23) * q.id.x;
    const float Ay = wide_f((((meta >> 8) & 0xffu) + 15u) << 23) * q.id.y;
    const float Az = wide_f((((meta >> 16) & 0xffu) + 15u) << 23) * q.id.z;
    const float ox = (wide_f(c0.w[0]) - q.o.x) * q.id.x, oy = (wide_f(c0.w[1]) - q.o.y) * q.id.y, oz = (wide_f(c0.w[2]) - q.o.z) * q.id.z;
    const float wx = WIDE_WIDEN * (fabsf(ox) + fabsf(Ax)), wy = WIDE_WIDEN * (fabsf(oy) + fabsf(Ay)), wz = WIDE_WIDEN * (fabsf(oz) + fabsf(Az));
    const float Bx = ox - Ax, By = oy - Ay, Bz = oz - Az;
    const float Bnx = Bx - wx, Bfx = Bx + wx, Bny = By - wy, Bfy = By + wy, Bnz = Bz - wz, Bfz = Bz + wz;
    const bool nx = (q.neg & 1u) != 0, ny = (q.neg & 2u) != 0, nz = (q.neg & 4u) != 0;
    // near / far plane bytes of slots 0-3 and 4-7 per axis
    const uint32_t nx0 = nx ? c0.w[6] : c0.w[4], nx1 = nx ? c0.w[7] : c0.w[5], fx0 = nx ? c0.w[4] : c0.w[6], fx1 = nx ? c0.w[5] : c0.w[7];
    const uint32_t ny0 = ny ? c1.w[2] : c1.w[0], ny1 = ny ? c1.w[3] : c1.w[1], fy0 = ny ? c1.w[0] : c1.w[2], fy1 = ny ? c1.w[1] : c1.w[3];
    const uint32_t nz0 = nz ? c1.w[6] : c1.w[4], nz1 = nz ? c1.w[7] : c1.w[5], fz0 = nz ? c1.w[4] : c1.w[6], fz1 = nz ? c1.w[5] : c1.w[7];
    uint32_t mask = 0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int s = 0; s < 8; ++s) {
        const uint32_t sel = 0x7604u | ((uint32_t)(s & 3) << 4);
        const uint32_t one = 0x3F800000u;
        const float tnx = hrt_fma(wide_f(wide_perm(s < 4 ? nx0 : nx1, one, sel)), Ax, Bnx), tfx = hrt_fma(wide_f(wide_perm(s < 4 ? fx0 : fx1, one, sel)), Ax, Bfx);
        const float tny = hrt_fma(wide_f(wide_perm(s < 4 ? ny0 : ny1, one, sel)), Ay, Bny), tfy = hrt_fma(wide_f(wide_perm(s < 4 ? fy0 : fy1, one, sel)), Ay, Bfy);
        const float tnz = hrt_fma(wide_f(wide_perm(s < 4 ? nz0 : nz1, one, sel)), Az, Bnz), tfz = hrt_fma(wide_f(wide_perm(s < 4 ? fz0 : fz1, one, sel)), Az, Bfz);
        const float tn = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, t_lo));
        const float tf = fminf(fminf(tfx, tfy), fminf(tfz, bnd0));
        if (!(tn > tf)) mask |= 1u << s;
    }
    return mask & (meta >> 24);
}

// bit k of the result = bit (k ^ r) of m
HRT_HD uint32_t wide_xor_bits(uint32_t m, uint32_t r)
{
    if (r & 1u) m = ((m & 0xAAu) >> 1) | ((m & 0x55u) << 1);
    if (r & 2u) m = ((m & 0xCCu) >> 2) | ((m & 0x33u) << 2);
    if (r & 4u) m = ((m & 0xF0u) >> 4) | ((m & 0x0Fu) << 4);
    return m;
}
// ---- one lane's walk ------------------------------------------------------------------------------------------------------------
// St is the stack: push(int) -> false when full, top(), set_top(int), drop(), empty().  In the kernels it is a column of shared memory; in the CPU simulation
// (tests/hostsim) a plain array.
struct WideLane {
    Ray r; V3 rid; float t_min;                    // the world ray
    float icull_top;                               // 1 / (top-level cull scale), rounded up (RotateQuat objects, see world_walk)
    WRay q;                                        // the ray of the current level
    const WNode* nodes; const float4* leaves;      // current level: top-level tree, or a mesh's tree + leaf records
    uint32_t prim_base; int obj;                   // obj >= 0: inside the mesh of world-list entry obj
    V3 ld; TriRay tq;                              // mesh level: local direction (exact box test of an accepted triangle), ITriangle's per-ray part
    Best b;
    float bnd0;                                    // max(0, cull bound in the current level's units)
    int cur;                                       // >= 0 wide node, < 0 leaf ~id of the current level
    bool live;
};

HRT_HD void wide_set_bound(WideLane& L)
{
    const float bound = cull_bound(L.b);
    L.bnd0 = fmaxf(0.0f, L.obj >= 0 ? bound : bound * L.icull_top);
}
HRT_HD void wide_enter_top(WideLane& L, const DScene& s)
{
    L.q = wide_ray(L.r.o, L.r.d, L.rid);
    L.nodes = s.tlas_wide; L.leaves = nullptr; L.prim_base = 0; L.obj = -1;
    wide_set_bound(L);
}
// Stack entries are GROUPS: (wide node << 8) | the children of that node still to be walked, as a mask in walking order (bit k
// = slot k ^ oct; the highest set bit is the nearest octant), or WALK_MARK ("leave the mesh").  One entry per node however many
// of its children were hit, so the stack never holds more than the depth of the trees; a child's id is read from its node only
// when it is walked (the sector was prefetched when the node was tested).
HRT_HD int wide_fls(uint32_t m)
{
#if defined(__CUDA_ARCH__)
    return 31 - __clz((int)m);
#else
    return 31 - __builtin_clz(m);
#endif
}
// What every step ends with: the lane's next node or leaf.  `group` is the group a node step just produced (its nearest child is
// taken, the others are stacked) or 0 (the nearest remaining child of the group on top of the stack).  ONE copy of this code per
// kernel, run after the phases of a round for all the lanes that took a step (r2o profile: inlined at the end of each step it
// ran with 4.8 of 32 lanes and made a third of the hot code).
// The group travels in L.cur (dead between a step and the pop): steps leave it there, 0 for "none".
template <class St> HRT_HD void wide_pop(WideLane& L, const DScene& s, St& st)
{
    // (no loop: a marker is never followed by another one - meshes are only entered from the top level - and a loop around
    // wide_enter_top made the compiler shuffle the whole lane state through a second register set, 18 % of teapot_scene's instructions, r2z)
    uint32_t v = (uint32_t)L.cur; L.cur = 0;
    if (v == 0u) {
        if (st.empty() || L.b.bad) { L.live = false; return; }
        v = (uint32_t)st.top(); st.drop();
        if (v == (uint32_t)WALK_MARK) {                                                   // the mesh is done: back to the world ray
            wide_enter_top(L, s);
            if (st.empty()) { L.live = false; return; }
            v = (uint32_t)st.top(); st.drop();
        }
    }
    uint32_t mp = v & 0xffu; const uint32_t node = v >> 8;
    const int k = wide_fls(mp);
    mp &= ~(1u << k);
    if (mp && !st.push((int)((node << 8) | mp))) { L.b.bad = 1; L.live = false; return; }
    L.cur = wide_ld_child(L.nodes, node, (uint32_t)k ^ wide_oct(L.q));
}

// Starts the walk of ray r.  Returns false when pass 1 has nothing to walk (L.live stays false): no scene, an all-NaN origin
// (answered on the spot in `early`, see trace_fast), or a ray the fast walk cannot take (L.b.bad: trace_ref decides).
HRT_HD bool wide_begin(WideLane& L, const DScene& s, const Ray& r, float t_min, float t_max, Hit& early)
{
    early.t = 0.0f; early.prim = HRT_MISS; early.obj = -1; early.flags = 0;
    L.live = false;
    L.b.t = 0.0f; L.b.a = 0.0f; L.b.prim = HRT_MISS; L.b.obj = -1; L.b.amb = 0; L.b.bad = 0; L.b.shear = 0; L.b.t_max = t_max;
    if (s.n_objects <= 0) return false;
    if (r.o.x != r.o.x && r.o.y != r.o.y && r.o.z != r.o.z) {
        if (s.nan_obj >= 0) { early.t = int_as_float(0x7fc00000); early.prim = s.nan_prim; early.obj = s.nan_obj; }
        return false;
    }
    const V3 id = v3(1.0f / r.d.x, 1.0f / r.d.y, 1.0f / r.d.z);
    if (!(recip_ok(id) && hrt_finite(r.o.x) && hrt_finite(r.o.y) && hrt_finite(r.o.z))) { L.b.bad = 1; return false; }
    L.r = r; L.rid = id; L.t_min = t_min;
    const float cull_top = s.tlas_cull_scale > 0.0f ? fminf(1.0f, length(r.d) * s.tlas_cull_scale) * 0.9999f : 1.0f;
    L.icull_top = cull_top < 1.0f ? (1.0f / cull_top) * 1.000001f : 1.0f;
    L.ld = r.d; L.tq.o = r.o; L.tq.sX = L.tq.sY = L.tq.sZ = 0.0f; L.tq.kz = 2;
    wide_enter_top(L, s);
    L.cur = s.tlas_wide_root; L.live = true;
    return true;
}

// ---- node step: tests the eight children of node L.cur; returns the group (node << 8 | children hit, in walking order: the highest
// bit is the nearest octant) for wide_pop in L.cur, 0 when the ray misses them all
HRT_HD void wide_node_step(WideLane& L)
{
    const uint32_t node = (uint32_t)L.cur;
    const char* N = (const char*)(L.nodes + (size_t)node);
    const W32 c0 = wide_ld32(N), c1 = wide_ld32(N + 32);
    wide_prefetch(N + 64);
    const float t_lo = fminf(L.t_min, 0.0f);                                    // ITriangle::hit ignores t_min but only reports t > 0
    const uint32_t mp = wide_xor_bits(wide_hits(c0, c1, L.q, t_lo, L.bnd0), wide_oct(L.q));
    L.cur = mp ? (int)((node << 8) | mp) : 0;
}

// itriangle_t<3> (hrt_trace.cuh) without early exits: the same operations in the same order on every lane, the verdict formed at
// the end with non-short-circuit logic, so the lanes of a leaf phase stay together (r2b profile: the early exits left 5.5 of
// 32 lanes in the second half of the test).  A rejected triangle's t is garbage and never read.
HRT_HD bool wide_itriangle(const float4& P0, const float4& P1, const float4& P2, const TriRay& q, float t_max, float& t)
{
    V3 p0t = permute<3>(xyz(P0) - q.o, q.kz), p1t = permute<3>(xyz(P1) - q.o, q.kz), p2t = permute<3>(xyz(P2) - q.o, q.kz);
    p0t.x += q.sX * p0t.z; p0t.y += q.sY * p0t.z;
    p1t.x += q.sX * p1t.z; p1t.y += q.sY * p1t.z;
    p2t.x += q.sX * p2t.z; p2t.y += q.sY * p2t.z;
    const float e0 = p1t.x * p2t.y - p1t.y * p2t.x;
    const float e1 = p2t.x * p0t.y - p2t.y * p0t.x;
    const float e2 = p0t.x * p1t.y - p0t.y * p1t.x;
    const bool mixed = ((e0 < 0) | (e1 < 0) | (e2 < 0)) & ((e0 > 0) | (e1 > 0) | (e2 > 0));
    const float det = e0 + e1 + e2;
    p0t.z *= q.sZ; p1t.z *= q.sZ; p2t.z *= q.sZ;
    const float tScaled = e0 * p0t.z + e1 * p1t.z + e2 * p2t.z;
    const float lim = t_max * det;
    const bool out_neg = (det < 0) & ((tScaled >= 0) | (tScaled < lim));
    const bool out_pos = (det > 0) & ((tScaled <= 0) | (tScaled > lim));
    const float invDet = 1 / det;
    t = tScaled * invDet;
    return !mixed & (det != 0) & !out_neg & !out_pos;
}

// ---- triangle leaf: the 1-2 triangles of a bottom node of the reference topology, exact ITriangle arithmetic; the reference's own
// slab test of the node box only for a triangle that accepted (world_walk has the reasoning)
HRT_HD void wide_leaf_step(WideLane& L)
{
    const uint32_t leaf = (uint32_t)(~L.cur);
    const char* R = (const char*)(L.leaves + 8 * (size_t)leaf);
    const W32 a = wide_ld32(R + 32), b = wide_ld32(R + 64), c = wide_ld32(R + 96);
    const int triL = (int)a.w[3], triR = (int)b.w[7];
    const float bound = cull_bound(L.b);
    const float4 L0 = wide_f4(a.w[0], a.w[1], a.w[2]), L1 = wide_f4(a.w[4], a.w[5], a.w[6]), L2 = wide_f4(b.w[0], b.w[1], b.w[2]);
    const float4 R0 = wide_f4(b.w[4], b.w[5], b.w[6]), R1 = wide_f4(c.w[0], c.w[1], c.w[2]), R2 = wide_f4(c.w[4], c.w[5], c.w[6]);
    float tL = 0, tR = 0;
    const bool hL = wide_itriangle(L0, L1, L2, L.tq, bound, tL);
    const bool hR = wide_itriangle(R0, R1, R2, L.tq, bound, tR) & (triR != triL);
    if (hL || hR) {
        const W32 box = wide_ld32(R);
        Ray lr; lr.o = L.q.o; lr.d = L.ld;
        float entry;
        if (hit_aabb_entry(v3(wide_f(box.w[0]), wide_f(box.w[1]), wide_f(box.w[2])), v3(wide_f(box.w[4]), wide_f(box.w[5]), wide_f(box.w[6])), lr, L.t_min, bound, entry)) {
            if (hL) consider(L.b, tL, fmaxf(entry, tL), L.prim_base + (uint32_t)triL, L.obj);
            if (hR) consider(L.b, tR, fmaxf(entry, tR), L.prim_base + (uint32_t)triR, L.obj);
            wide_set_bound(L);
        }
    }
    L.cur = 0;
}

// ---- world-list entry that is a mesh (WIDE_MESH_BIT): entered with the local ray, a marker goes on the stack.  Returns true when the
// mesh was NOT entered (empty mesh, or a ray pass 2 has to take: L.b.bad / L.b.shear) and the lane pops instead.
template <class St> HRT_HD bool wide_mesh_step(WideLane& L, const DScene& s, St& st)
{
    const int i = (int)((uint32_t)(~L.cur) & ~WIDE_TAG_BITS);
    const DObject& ob = s.objects[i];
    const DMesh& m = s.meshes[ob.mesh];
    L.cur = 0;
    if (m.ref_root < 0) return true;
    Ray lr = L.r; V3 id = L.rid; bool ok = true;
    if (ob.xf_count > 0) {
        lr = to_local(s, ob, L.r);
        id = v3(1.0f / lr.d.x, 1.0f / lr.d.y, 1.0f / lr.d.z);
        ok = recip_ok(id) && hrt_finite(lr.o.x) && hrt_finite(lr.o.y) && hrt_finite(lr.o.z);
    }
    const int kz = tri_kz(lr.o);
    if (!ok) { L.b.bad = 1; return true; }
    if (!shear_ok(lr, kz)) { L.b.shear = 1; return true; }                      // pass 2 collects every hit along this ray
    if (!st.push(WALK_MARK)) { L.b.bad = 1; return true; }
    L.q = wide_ray(lr.o, lr.d, id);
    L.nodes = m.wide_nodes; L.leaves = m.fast_leaves; L.prim_base = m.prim_base; L.obj = i;
    L.ld = lr.d; L.tq = tri_ray<3>(lr, kz);
    wide_set_bound(L);
    L.cur = m.wide_root;
    return false;
}

// ---- any other world-list entry: tested on the spot.  SLOW = false: the entries without WIDE_SLOW_BIT (plain spheres and rects, in line);
// SLOW = true: everything else through the general object test (out of line)
// (one entry per call: taking a lane's next entries in the same call - a Cornell-box ray holds three in a row - was slower,
// teapot_scene 1612 -> 1496 Mrays/s, and unrolled it outgrew the instruction cache: 931, profiles/README.md r2h / r2i)
template <bool SLOW>
HRT_HD void wide_object_step(WideLane& L, const DScene& s, const RngKey& key, uint32_t bounce)
{
    const int i = (int)((uint32_t)(~L.cur) & ~WIDE_TAG_BITS);
    const DObject& ob = s.objects[i];
    float t = 0.0f; uint32_t prim = ob.prim_base; bool hit;
    const float bound = cull_bound(L.b);
    if (SLOW) {
        const Ray wr = L.r;                                             // a copy: hit_object_ref is out of line
        hit = hit_object_ref(s, i, wr, L.t_min, bound, t, prim, key, bounce);
    } else if (ob.kind == SHAPE_SPHERE) hit = sphere_t(ob.f, L.r, L.t_min, bound, t);
    else { float a, b; hit = rect_t(ob.axis, ob.f[0], ob.f[1], ob.f[2], ob.f[3], ob.f[4], L.r, L.t_min, bound, t, a, b); }
    if (hit) { consider(L.b, t, t, prim, i); wide_set_bound(L); }
    L.cur = 0;
}

// what pass 1 found, as a hit record: final, or pending (the caller queues the ray for trace_resolve)
HRT_HD Hit wide_result(const WideLane& L)
{
    Hit h; h.obj = L.b.obj; h.flags = 0;
    if (!L.b.bad && !L.b.amb && !L.b.shear) { h.t = L.b.prim == HRT_MISS ? 0.0f : L.b.t; h.prim = L.b.prim; return h; }
    h.t = cull_bound(L.b); h.prim = HRT_MISS;
    h.flags = HIT_PENDING | (L.b.bad ? PEND_BAD : 0) | (L.b.shear ? PEND_SHEAR : 0);
    return h;
}
// pass 2 of a pending record (hrt_trace.cuh, trace_resolve)
HRT_HD Hit wide_resolve(const DScene& s, const Ray& r, float t_min, float t_max, const RngKey& key, uint32_t bounce, const Hit& pending)
{
    return trace_resolve<false>(s, r, t_min, t_max, key, bounce, nullptr, (pending.flags & PEND_BAD) ? 1 : 0, (pending.flags & PEND_SHEAR) ? 1 : 0, pending.t);
}

// ---- one ray, start to finish, on one thread (CPU simulation; validation) --------------------------------------------------------------
struct WideArrayStack {
    int v[WIDE_STACK]; int sp = 0;
    HRT_HD bool push(int x) { if (sp >= WIDE_STACK) return false; v[sp++] = x; return true; }
    HRT_HD int top() const { return v[sp - 1]; }
    HRT_HD void set_top(int x) { v[sp - 1] = x; }
    HRT_HD void drop() { --sp; }
    HRT_HD bool empty() const { return sp == 0; }
};
// kinds (may be NULL): 4 counters - node steps, triangle leaves, mesh entries, other world-list entries (the visit counts behind SURVEY 8(d)'s B_logical)
HRT_HDN inline Hit wide_trace_scalar(const DScene& s, const Ray& r, float t_min, float t_max, const RngKey& key, uint32_t bounce, unsigned long long* steps,
                                     unsigned long long* kinds = nullptr)
{
    WideLane L; Hit h;
    if (!wide_begin(L, s, r, t_min, t_max, h)) {
        if (!L.b.bad) return h;
    } else {
        WideArrayStack st;
        while (L.live) {
            if (steps) ++*steps;
            bool pop = true;
            if (L.cur >= 0) { if (kinds) ++kinds[0]; wide_node_step(L); }
            else if (L.obj >= 0) { if (kinds) ++kinds[1]; wide_leaf_step(L); }
            else if ((uint32_t)(~L.cur) & WIDE_MESH_BIT) { if (kinds) ++kinds[2]; pop = wide_mesh_step(L, s, st); }
            else if ((uint32_t)(~L.cur) & WIDE_SLOW_BIT) { if (kinds) ++kinds[3]; wide_object_step<true>(L, s, key, bounce); }
            else { if (kinds) ++kinds[3]; wide_object_step<false>(L, s, key, bounce); }
            if (pop) wide_pop(L, s, st);
        }
    }
    h = wide_result(L);
    if (h.flags & HIT_PENDING) h = wide_resolve(s, r, t_min, t_max, key, bounce, h);
    return h;
}

// ---- a warp of lanes (device only) ------------------------------------------------------------------------------------------------
#if defined(__CUDACC__)
// Thresholds of a warp's rounds, per scene (DScene::wide_phase_min / wide_refill, set at commit in hrt_upload.h; B200 A/B r2c, r3d):
// a phase runs when at least wide_phase_min lanes wait for it (or nothing larger can run); idle lanes take new rays as soon as fewer
// than wide_refill lanes are walking.

// One column of a shared-memory array per thread: entry k of thread t lives at base[k * THREADS], so the lanes of a warp
// hit 32 different banks whatever their stack depths.
template <int THREADS> struct WideSmemStack {
    int* base; int sp;
    __device__ __forceinline__ bool push(int x) { if (sp >= WIDE_STACK) return false; base[sp * THREADS] = x; ++sp; return true; }
    __device__ __forceinline__ int top() const { return base[(sp - 1) * THREADS]; }
    __device__ __forceinline__ void set_top(int x) { base[(sp - 1) * THREADS] = x; }
    __device__ __forceinline__ void drop() { --sp; }
    __device__ __forceinline__ bool empty() const { return sp == 0; }
};

// One round of the warp: every lane that is walking waits for a node step, a triangle leaf, a mesh entry or another world-list
// entry; the phases that enough lanes wait for run, the others wait for the next round; the lanes that took a step then fetch
// their next node together (wide_pop).  Must be called by all 32 lanes; returns false when no lane is walking.
// SLOW = false: the build for scenes whose world list has no entry with WIDE_SLOW_BIT (DScene::wide_has_slow == 0: meshes, plain spheres and
// rects only) - four phases, four ballots; the fifth phase costs such scenes 3 - 5 % (r3v: big_mesh 2602 -> 2483, ray batch 1706 -> 1680).
template <class St, bool SLOW>
__device__ __forceinline__ bool wide_round(WideLane& L, const DScene& s, St& st, const RngKey& key, uint32_t bounce)
{
    if (!SLOW) {
        const bool top = L.live && L.cur < 0 && L.obj < 0;
        const bool mesh = ((uint32_t)(~L.cur) & WIDE_MESH_BIT) != 0u;
        const bool a_in = L.live && L.cur >= 0;
        const bool a_tl = L.live && L.cur < 0 && L.obj >= 0;
        const bool a_me = top && mesh, a_ob = top && !mesh;
        const int c_in = __popc(__ballot_sync(0xffffffffu, a_in)), c_tl = __popc(__ballot_sync(0xffffffffu, a_tl));
        const int c_me = __popc(__ballot_sync(0xffffffffu, a_me)), c_ob = __popc(__ballot_sync(0xffffffffu, a_ob));
        const int top_n = max(max(c_in, c_tl), max(c_me, c_ob));
        if (top_n == 0) return false;
        const int t0 = s.wide_phase_min, t1 = s.wide_phase_min_obj;
        bool go = false;
        if (c_in >= t0 || c_in == top_n) { if (a_in) { wide_node_step(L); go = true; } }
        if (c_tl >= t0 || c_tl == top_n) { if (a_tl) { wide_leaf_step(L); go = true; } }
        if (c_me >= t0 || c_me == top_n) { if (a_me) go = wide_mesh_step(L, s, st); }
        if (c_ob >= t1 || c_ob == top_n) { if (a_ob) { wide_object_step<false>(L, s, key, bounce); go = true; } }
        if (go) wide_pop(L, s, st);
        return true;
    }
    const bool top_leaf = L.live && L.cur < 0 && L.obj < 0;
    const uint32_t tag = (uint32_t)(~L.cur) & WIDE_TAG_BITS;
    const bool w_in = L.live && L.cur >= 0;
    const bool w_tl = L.live && L.cur < 0 && L.obj >= 0;
    const bool w_me = top_leaf && tag == WIDE_MESH_BIT;
    const bool w_ob = top_leaf && tag == 0u;
    const bool w_sl = top_leaf && tag == WIDE_SLOW_BIT;
    const int n_in = __popc(__ballot_sync(0xffffffffu, w_in)), n_tl = __popc(__ballot_sync(0xffffffffu, w_tl));
    int n_me = 0, n_ob = 0, n_sl = 0;
    if (__ballot_sync(0xffffffffu, top_leaf)) {                                // warp-uniform: inside a large mesh most rounds have no lane at the top level
        n_me = __popc(__ballot_sync(0xffffffffu, w_me)); n_ob = __popc(__ballot_sync(0xffffffffu, w_ob)); n_sl = __popc(__ballot_sync(0xffffffffu, w_sl));
    }
    const int most = max(max(max(n_in, n_tl), max(n_me, n_ob)), n_sl);
    if (most == 0) return false;
    // a phase runs when enough lanes wait for it, or when no phase has more (the largest group always runs: progress)
    const int thr = s.wide_phase_min, thr_ob = s.wide_phase_min_obj;
    bool pop = false;
    if (n_in >= thr || n_in == most) { if (w_in) { wide_node_step(L); pop = true; } }
    if (n_tl >= thr || n_tl == most) { if (w_tl) { wide_leaf_step(L); pop = true; } }
    if (n_me >= thr || n_me == most) { if (w_me) pop = wide_mesh_step(L, s, st); }        // (a mesh entry that waits keeps a long walk from starting: r3i)
    if (n_ob >= thr_ob || n_ob == most) { if (w_ob) { wide_object_step<false>(L, s, key, bounce); pop = true; } }
    if (n_sl >= thr_ob || n_sl == most) { if (w_sl) { wide_object_step<true>(L, s, key, bounce); pop = true; } }
    if (pop) wide_pop(L, s, st);
    return true;
}
#endif

} // namespace hrt
