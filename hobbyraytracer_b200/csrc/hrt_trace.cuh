// hrt_trace.cuh — ordered closest-hit traversal that returns exactly what the reference's sequential world->hit
// (trace_ref in hrt_hit.cuh) returns, without walking the scene in the reference's order.
//
// The reference visits primitives in a fixed order (flat world list, per-mesh BVH left-then-right) with one running
// `closest`: primitive i is ACCEPTED when it is visited with closest >= a_i, and then closest = t_i.  For spheres,
// rects, boxes and media a_i = t_i.  A triangle is only reached through its bottom BVH node, whose slab test needs
// closest > entry_i, so a_i = max(entry_i, t_i) (ancestor boxes contain the bottom box exactly, and fp division is
// monotone, so they never reject first).  The answer is the LAST accepted primitive.
//
//   pass 1  SAH BVHs (top level over world-list entries, per mesh over the reference's bottom nodes), nearest child
//           first, culled at 6.25 % beyond the best candidate.  If the candidate m with the smallest t has a_m below
//           every other candidate's t, every visiting order ends with m: done (the overwhelmingly common case).
//   pass 2  otherwise (near-ties, a_m reaching past another candidate) the candidates up to the cull bound are
//           collected with the reference's own arithmetic and the reference's sequential process is replayed over
//           them in the reference's order — a handful of primitive tests instead of a whole-scene walk.
//   shear   ITriangle::hit takes its shear axis from the ray ORIGIN (triangle.cpp:70); when |d[kZ]| is small against
//           the other components its edge functions are garbage and a triangle can report a t far in front of its own
//           box, so distance culling is unsound for that ray.  Such rays collect EVERY hit along the ray (no culling by
//           distance) before the replay.  What a replay needs is only what the reference itself would have accepted
//           at closest = infinity, so it is exact however badly conditioned the triangles are.
//   last resort: NaN t, zero direction components, non-finite origins, more candidates than the collector holds —
//           trace_ref on the spot (vanishing probability for rendered rays; adversarial test batches exercise it).
#pragma once
#include "hrt_hit.cuh"

namespace hrt {

constexpr float NEAR_EPS = 1e-5f;            // relative distance under which two candidate hits count as a tie
constexpr float CULL_MARGIN = 0.0625f;       // pass 1 keeps everything that starts within 6.25 % beyond the best candidate
constexpr float SHEAR_LIMIT = 64.0f;         // max|d| / |d[kZ]| above which ITriangle's t is too inaccurate to cull by
constexpr float SLAB_WIDEN = 4.76837158e-7f; // 2^-21: widening of the SAH-node slab interval (conservative vs aabb.h)
constexpr int HIT_RESOLVED = 2;              // Hit::flags: answer came from pass 2 + replay
constexpr int HIT_FALLBACK = 4;              // Hit::flags: answer came from trace_ref (last resort)
constexpr int KMAX = 16;                     // candidates a replay can hold

struct TraceCounters { unsigned long long nodes, prims, fallbacks, amb, bad; };

struct Best { float t, a; uint32_t prim; int obj; int amb; int bad; int shear; float t_max; };

HRT_HD float cull_bound(const Best& b)      // nothing that starts beyond this can take part in the decision
{ return b.prim == HRT_MISS ? b.t_max : b.a + CULL_MARGIN * fabsf(b.a); }

HRT_HD void consider(Best& b, float t, float a, uint32_t prim, int obj)
{
    if (t != t || a != a) { b.bad = 1; return; }                                // NaN t poisons `closest` in the reference
    if (b.prim == HRT_MISS) { b.t = t; b.a = a; b.prim = prim; b.obj = obj; return; }
    if (t < b.t) {
        if (a + 2.0f * NEAR_EPS * fabsf(a) >= b.t) b.amb = 1;                   // not clear of the old best (or of its ties)
        b.t = t; b.a = a; b.prim = prim; b.obj = obj;
    } else if (t <= b.a + NEAR_EPS * fabsf(b.a)) b.amb = 1;
}

struct Cand { float t; int obj; uint32_t prim; uint32_t key; };   // key: mesh -> leaf * 2 + side (reference DFS order), else 0
struct Collector { Cand c[KMAX]; int n; int overflow; int bad; float beta; };

HRT_HD void collect(Collector& k, float t, int obj, uint32_t prim, uint32_t key)
{
    if (t != t) { k.bad = 1; return; }
    if (k.n >= KMAX) { k.overflow = 1; return; }
    Cand c; c.t = t; c.obj = obj; c.prim = prim; c.key = key;
    k.c[k.n++] = c;
}

// Slab test of the two children stored in an SAH node against (o, 1/d); conservative with respect to aabb.h's
// divisions (one extra rounding from the reciprocal, covered by SLAB_WIDEN).
HRT_HD void slab2(const float4& n0, const float4& n1, const float4& n2, V3 o, V3 id, float& tn0, float& tf0, float& tn1, float& tf1)
{
    float ax = (n0.x - o.x) * id.x, bx = (n0.y - o.x) * id.x, ay = (n0.z - o.y) * id.y, by = (n0.w - o.y) * id.y;
    float az = (n2.x - o.z) * id.z, bz = (n2.y - o.z) * id.z;
    tn0 = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz));
    tf0 = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
    float cx = (n1.x - o.x) * id.x, dx = (n1.y - o.x) * id.x, cy = (n1.z - o.y) * id.y, dy = (n1.w - o.y) * id.y;
    float cz = (n2.z - o.z) * id.z, dz = (n2.w - o.z) * id.z;
    tn1 = fmaxf(fmaxf(fminf(cx, dx), fminf(cy, dy)), fminf(cz, dz));
    tf1 = fminf(fminf(fmaxf(cx, dx), fmaxf(cy, dy)), fmaxf(cz, dz));
    tn0 -= fabsf(tn0) * SLAB_WIDEN; tf0 += fabsf(tf0) * SLAB_WIDEN;
    tn1 -= fabsf(tn1) * SLAB_WIDEN; tf1 += fabsf(tf1) * SLAB_WIDEN;
}

HRT_HD bool recip_ok(V3 id) { return hrt_finite(id.x) && hrt_finite(id.y) && hrt_finite(id.z); }   // id = 1/d computed by the caller
HRT_HD bool dir_ok(V3 d)                    // every component non-zero with a finite reciprocal
{
    return hrt_finite(1.0f / d.x) && hrt_finite(1.0f / d.y) && hrt_finite(1.0f / d.z);
}

// ---- ITriangle::hit with the per-ray part hoisted -------------------------------------------------------------------
// triangle.cpp:70-84 depend on the ray only: the permutation (kZ from the ORIGIN) and the shear constants.  KZ is a
// template parameter so that the permutation costs no selects; the arithmetic per triangle is itriangle()'s, verbatim.
struct TriRay { V3 o; float sX, sY, sZ; int kz; };

// KZ = 0, 1, 2: compile-time permutation (no selects).  KZ = 3: the permutation is taken from q.kz at run time (two
// selects per component) so that a warp whose rays disagree on kZ — secondary rays, whose origins lie all over the
// mesh — walks the tree once instead of once per kZ value.  The arithmetic on the permuted components is the same.
template <int KZ> HRT_HD V3 permute(V3 v, int kz)
{
    if (KZ == 3) return kz == 0 ? v3(v.y, v.z, v.x) : (kz == 1 ? v3(v.z, v.x, v.y) : v);
    return KZ == 0 ? v3(v.y, v.z, v.x) : (KZ == 1 ? v3(v.z, v.x, v.y) : v);   // (kX, kY, kZ) = (KZ+1, KZ+2, KZ) mod 3
}
template <int KZ> HRT_HD TriRay tri_ray(const Ray& r, int kz)
{
    V3 d = permute<KZ>(r.d, kz);
    TriRay q; q.o = r.o; q.kz = kz;
    q.sX = -d.x / d.z; q.sY = -d.y / d.z; q.sZ = 1.0f / d.z;
    return q;
}
HRT_HD int tri_kz(V3 o) { return o.x > o.z ? (o.x > o.y ? 0 : 1) : 2; }        // triangle.cpp:70

template <int KZ>
HRT_HD bool itriangle_t(const float4& P0, const float4& P1, const float4& P2, const TriRay& q, float t_max, float& t)
{
    V3 p0t = permute<KZ>(xyz(P0) - q.o, q.kz), p1t = permute<KZ>(xyz(P1) - q.o, q.kz), p2t = permute<KZ>(xyz(P2) - q.o, q.kz);
    p0t.x += q.sX * p0t.z; p0t.y += q.sY * p0t.z;
    p1t.x += q.sX * p1t.z; p1t.y += q.sY * p1t.z;
    p2t.x += q.sX * p2t.z; p2t.y += q.sY * p2t.z;
    float e0 = p1t.x * p2t.y - p1t.y * p2t.x;
    float e1 = p2t.x * p0t.y - p2t.y * p0t.x;
    float e2 = p0t.x * p1t.y - p0t.y * p1t.x;
    if ((e0 < 0 || e1 < 0 || e2 < 0) && (e0 > 0 || e1 > 0 || e2 > 0)) return false;
    float det = e0 + e1 + e2;
    if (det == 0) return false;
    p0t.z *= q.sZ; p1t.z *= q.sZ; p2t.z *= q.sZ;
    float tScaled = e0 * p0t.z + e1 * p1t.z + e2 * p2t.z;
    if (det < 0 && (tScaled >= 0 || tScaled < t_max * det)) return false;
    else if (det > 0 && (tScaled <= 0 || tScaled > t_max * det)) return false;
    float invDet = 1 / det;
    t = tScaled * invDet;
    return true;
}

// shear conditioning of ITriangle for this (local) ray
HRT_HD bool shear_ok(const Ray& lr, int kZ)
{
    float dmax = fmaxf(fmaxf(fabsf(lr.d.x), fabsf(lr.d.y)), fabsf(lr.d.z));
    return dmax <= SHEAR_LIMIT * fabsf(comp(lr.d, kZ));
}

// ---- one loop over both levels ---------------------------------------------------------------------------------------------
// Top level: SAH tree over world-list entries.  Bottom level: per-mesh SAH tree over the reference's bottom nodes (a leaf
// is one bottom node of bvh.cpp's topology: its 1-2 triangles are tested with exact ITriangle arithmetic, and the
// reference's own slab test of the node box — six IEEE divisions — is only evaluated for a triangle that was accepted:
// it decides whether the reference could have reached the triangle at this bound and yields entry_i).
// Both levels share one stack and ONE node loop: entering a mesh pushes a marker, swaps in the local ray and carries on
// with the mesh's nodes; popping the marker swaps the world ray back.  With a walk function per level (r1c) the lanes of
// a warp reached a mesh in different iterations of the top-level loop and in three kZ instantiations, and the mesh walk
// ran with 1.5-2.3 active lanes of 32 on scenes where most rays reach a mesh (profiles/ r1d, teapot_scene).
// COLLECT = false: pass 1 (track the best candidate, cull by it).  COLLECT = true: pass 2 (fixed bound k.beta, every
// candidate goes to the collector).
constexpr int WALK_MARK = (int)0x80000000;     // stack marker: "leave the mesh"; never a node (>= 0) nor ~index of anything
constexpr int WALK_STACK = 48 + FAST_MAX_DEPTH + 8;

template <bool COLLECT, bool COUNT>
HRT_HDN inline void world_walk(const DScene& s, const Ray& r, const V3 rid, float t_min, Best& b, Collector& k, const RngKey& key, uint32_t bounce, TraceCounters* cnt)
{
    // RotateQuat objects report t in units of their normalised direction (rotateQuat.cpp:51):
    // reported t >= world t * |d| * tlas_cull_scale, so entry distances are scaled down before they cull.
    const float cull_top = s.tlas_cull_scale > 0.0f ? fminf(1.0f, length(r.d) * s.tlas_cull_scale) * 0.9999f : 1.0f;
    const float t_lo = fminf(t_min, 0.0f);            // ITriangle::hit ignores t_min but only reports t > 0
    int stack[WALK_STACK]; int sp = 0;
    // current level: ray the nodes are tested against, node array, cull scale
    V3 co = r.o, cid = rid; float cull = cull_top;
    const float4* nodes = s.tlas_nodes;
    // mesh level state
    int obj = -1; Ray lr = r; TriRay q; q.o = r.o; q.sX = q.sY = q.sZ = 0.0f; q.kz = 2; bool q_ready = false;
    const float4* leaves = nullptr; uint32_t prim_base = 0;
    int node = s.tlas_root;
    for (;;) {
        while (node >= 0) {                                                    // inner nodes of either level
            const float4* N = nodes + 4 * (size_t)node;
            float4 n0 = HRT_LD4(N), n1 = HRT_LD4(N + 1), n2 = HRT_LD4(N + 2), n3 = HRT_LD4(N + 3);
            float tn0, tf0, tn1, tf1;
            slab2(n0, n1, n2, co, cid, tn0, tf0, tn1, tf1);
            if (COUNT) cnt->nodes += 2;
            const float bound = COLLECT ? k.beta : cull_bound(b);
            bool h0 = tn0 <= tf0 && tf0 >= t_lo && (tn0 <= 0.0f || tn0 * cull <= bound);
            bool h1 = tn1 <= tf1 && tf1 >= t_lo && (tn1 <= 0.0f || tn1 * cull <= bound);
            int c0 = float_as_int(n3.x), c1 = float_as_int(n3.y);
            if (h0 && h1) {
                if (tn1 < tn0) { int tmp = c0; c0 = c1; c1 = tmp; }
                if (sp < WALK_STACK - 2) stack[sp++] = c1; else b.bad = 1;
                node = c0;
            } else if (h0) node = c0;
            else if (h1) node = c1;
            else { if (sp == 0) return; node = stack[--sp]; }
        }
        if (node == WALK_MARK) {                                               // the mesh is done: back to the world ray
            co = r.o; cid = rid; cull = cull_top; nodes = s.tlas_nodes; obj = -1;
        } else if (obj >= 0) {                                                 // a leaf of the mesh tree
            const uint32_t leaf = (uint32_t)(~node);
            const float4* L = leaves + 8 * (size_t)leaf;
            const float4 a = HRT_LD4(L), c = HRT_LD4(L + 1);
            const int triL = float_as_int(a.w), triR = float_as_int(c.w);
            if (!q_ready) { q = tri_ray<3>(lr, tri_kz(lr.o)); q_ready = true; }
            const float bound = COLLECT ? k.beta : cull_bound(b);
            float tL = 0, tR = 0;
            if (COUNT) cnt->prims += (triR != triL) ? 2 : 1;
            bool hL = itriangle_t<3>(HRT_LD4(L + 2), HRT_LD4(L + 3), HRT_LD4(L + 4), q, bound, tL);
            bool hR = triR != triL && itriangle_t<3>(HRT_LD4(L + 5), HRT_LD4(L + 6), HRT_LD4(L + 7), q, bound, tR);
            if (hL || hR) {
                float entry;
                if (COUNT) cnt->nodes += 1;
                if (hit_aabb_entry(xyz(a), xyz(c), lr, t_min, bound, entry)) {       // the reference's bottom node box, its arithmetic
                    if (COLLECT) {
                        if (hL) collect(k, tL, obj, prim_base + (uint32_t)triL, leaf * 2u);
                        if (hR) collect(k, tR, obj, prim_base + (uint32_t)triR, leaf * 2u + 1u);
                    } else {
                        if (hL) consider(b, tL, fmaxf(entry, tL), prim_base + (uint32_t)triL, obj);
                        if (hR) consider(b, tR, fmaxf(entry, tR), prim_base + (uint32_t)triR, obj);   // beyond the new bound: no effect
                    }
                }
            }
        } else {                                                               // a world-list entry
            const int i = ~node;
            const DObject& ob = s.objects[i];
            if (ob.kind == SHAPE_MESH && !ob.medium) {
                const DMesh& m = s.meshes[ob.mesh];
                if (m.ref_root >= 0) {
                    // the three IEEE reciprocals of a ray are computed once per coordinate frame: an unwrapped mesh shares the world ray's
                    lr = r; V3 id = rid; bool ok = true;
                    if (ob.xf_count > 0) {
                        lr = to_local(s, ob, r);
                        id = v3(1.0f / lr.d.x, 1.0f / lr.d.y, 1.0f / lr.d.z);
                        ok = recip_ok(id) && hrt_finite(lr.o.x) && hrt_finite(lr.o.y) && hrt_finite(lr.o.z);
                    }
                    if (!ok) b.bad = 1;
                    else if (!COLLECT && !shear_ok(lr, tri_kz(lr.o))) b.shear = 1;      // pass 2 will collect every hit along this ray
                    else {
                        if (sp < WALK_STACK - 2) stack[sp++] = WALK_MARK; else b.bad = 1;
                        co = lr.o; cid = id; cull = 1.0f; nodes = m.fast_nodes; leaves = m.fast_leaves; prim_base = m.prim_base;
                        obj = i; q_ready = false;
                        node = m.fast_root;                                     // >= 0 node, < 0 leaf: handled by the next turn of the loop
                        if (!b.bad) continue;
                    }
                }
            } else {
                float t; uint32_t prim;
                if (COUNT) cnt->prims += 1;
                if (hit_object_ref(s, i, r, t_min, COLLECT ? k.beta : cull_bound(b), t, prim, key, bounce)) {
                    if (COLLECT) collect(k, t, i, prim, 0u); else consider(b, t, t, prim, i);
                }
            }
        }
        if (sp == 0 || b.bad) return;
        node = stack[--sp];
    }
}

// ---- replay: the reference's sequential process over the collected candidates, in the reference's order ----------------------
// Each step re-evaluates the reference's own acceptance test against the running `closest` (the bottom node's slab test
// once per node, then ITriangle::hit; hit_object_ref for everything else), so the outcome is the reference's whatever
// the conditioning.  Returns false when a NaN shows up (the caller falls back to trace_ref).
HRT_HDN inline bool replay(const DScene& s, const Ray& r, float t_min, float t_max, Collector& k, const RngKey& key, uint32_t bounce, Hit& h)
{
    h.t = 0.0f; h.prim = HRT_MISS; h.obj = -1; h.flags = 0;
    float closest = t_max;
    int box_obj = -1; uint32_t box_leaf = 0xFFFFFFFFu; bool box_ok = false;
    Ray lr = r; int lr_obj = -1;
    for (int done = 0; done < k.n; ++done) {
        int best = -1;                                                          // next candidate in (object, leaf, side) order
        for (int j = 0; j < k.n; ++j) {
            const Cand& c = k.c[j];
            if (c.obj < 0) continue;
            if (best < 0 || c.obj < k.c[best].obj || (c.obj == k.c[best].obj && c.key < k.c[best].key)) best = j;
        }
        if (best < 0) break;
        const Cand c = k.c[best]; k.c[best].obj = -1;
        const DObject& ob = s.objects[c.obj];
        if (ob.kind == SHAPE_MESH && !ob.medium) {
            const DMesh& m = s.meshes[ob.mesh];
            if (lr_obj != c.obj) { lr = to_local(s, ob, r); lr_obj = c.obj; }
            const uint32_t leaf = c.key >> 1;
            const float4* L = m.fast_leaves + 8 * (size_t)leaf;
            if (box_obj != c.obj || box_leaf != leaf) {                        // bvh.cpp:71: one slab test per node visit
                box_obj = c.obj; box_leaf = leaf;
                box_ok = hit_aabb(xyz(HRT_LD4(L)), xyz(HRT_LD4(L + 1)), lr, t_min, closest);
            }
            if (!box_ok) continue;
            const float4* T = L + 2 + 3 * (c.key & 1u);
            float t, b0, b1, b2;
            if (itriangle(xyz(HRT_LD4(T)), xyz(HRT_LD4(T + 1)), xyz(HRT_LD4(T + 2)), lr, closest, t, b0, b1, b2)) {
                if (t != t) return false;
                closest = t; h.t = t; h.prim = c.prim; h.obj = c.obj;
            }
        } else {
            float t; uint32_t prim;
            if (hit_object_ref(s, c.obj, r, t_min, closest, t, prim, key, bounce)) {
                if (t != t) return false;
                closest = t; h.t = t; h.prim = prim; h.obj = c.obj;
            }
        }
    }
    return true;
}

// ---- closest hit ---------------------------------------------------------------------------------------------------------------
#if defined(__CUDACC__)
#define HRT_COLD __host__ __device__ __noinline__
#else
#define HRT_COLD __attribute__((noinline))
#endif

// cold path, kept out of line so that the hot kernel body is pass 1 only: pass 2 + replay, or trace_ref as the last resort
template <bool COUNT>
HRT_COLD Hit trace_resolve(const DScene& s, const Ray& r, float t_min, float t_max, const RngKey& key, uint32_t bounce, TraceCounters* cnt,
                           int bad, int shear, float beta)
{
    Hit h; h.t = 0.0f; h.prim = HRT_MISS; h.obj = -1; h.flags = 0;
    if (!bad) {
        if (COUNT) cnt->amb += 1;
        Best b; b.t = 0.0f; b.a = 0.0f; b.prim = HRT_MISS; b.obj = -1; b.amb = 0; b.bad = 0; b.shear = 0; b.t_max = t_max;
        Collector k; k.n = 0; k.overflow = 0; k.bad = 0;
        k.beta = shear ? t_max : beta;
        world_walk<true, COUNT>(s, r, v3(1.0f / r.d.x, 1.0f / r.d.y, 1.0f / r.d.z), t_min, b, k, key, bounce, cnt);
        if (!k.overflow && !k.bad && !b.bad && replay(s, r, t_min, t_max, k, key, bounce, h)) { h.flags |= HIT_RESOLVED; return h; }
    }
    if (COUNT) { cnt->fallbacks += 1; cnt->bad += 1; }
    h = trace_ref(s, r, t_min, t_max, key, bounce);
    h.flags |= HIT_FALLBACK;
    return h;
}

template <bool COUNT>
HRT_HD Hit trace_fast(const DScene& s, const Ray& r, float t_min, float t_max, const RngKey& key, uint32_t bounce, TraceCounters* cnt)
{
    Hit h; h.t = 0.0f; h.prim = HRT_MISS; h.obj = -1; h.flags = 0;
    if (s.n_objects <= 0) return h;
    if (r.o.x != r.o.x && r.o.y != r.o.y && r.o.z != r.o.z) {
        // A path whose hit point became NaN (a NaN-t hit) keeps bouncing until MAX_DEPTH.  With an
        // all-NaN origin every comparison in every hit() is false, so every slab test passes and every primitive accepts
        // with t = NaN; the reference walks the whole scene and ends with the LAST primitive in its order.  Same answer,
        // without the walk (a whole-scene reference walk costs milliseconds of single-thread latency per pass here).
        if (s.nan_obj >= 0) { h.t = int_as_float(0x7fc00000); h.prim = s.nan_prim; h.obj = s.nan_obj; }
        return h;
    }
    const V3 id = v3(1.0f / r.d.x, 1.0f / r.d.y, 1.0f / r.d.z);
    const bool ok = recip_ok(id) && hrt_finite(r.o.x) && hrt_finite(r.o.y) && hrt_finite(r.o.z);
    Best b; b.t = 0.0f; b.a = 0.0f; b.prim = HRT_MISS; b.obj = -1; b.amb = 0; b.bad = ok ? 0 : 1; b.shear = 0; b.t_max = t_max;
    if (ok) {
        Collector* none = nullptr;
        world_walk<false, COUNT>(s, r, id, t_min, b, *none, key, bounce, cnt);       // pass 1 never touches the collector
        if (!b.bad && !b.amb && !b.shear) { h.t = b.prim == HRT_MISS ? 0.0f : b.t; h.prim = b.prim; h.obj = b.obj; return h; }
    }
    return trace_resolve<COUNT>(s, r, t_min, t_max, key, bounce, cnt, b.bad, b.shear, cull_bound(b));
}

} // namespace hrt
