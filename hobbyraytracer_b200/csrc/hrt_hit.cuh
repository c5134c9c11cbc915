// hrt_hit.cuh — closest-hit semantics of the reference (world->hit of main.cpp:45) over the flattened scene.
//
// Two traversal orders produce the SAME answer:
//   trace_ref  : the reference's own visiting order — flat world list (hittableList.cpp:4-21), per-mesh BVH in
//                bvh.cpp's topology walked left-then-right (bvh.cpp:69-78).  Exact by construction.
//   trace_fast : SAH BVHs (top level over world-list entries, per mesh over the reference's bottom nodes), nearest
//                child first.  Closest-hit is order independent except for (near-)ties, NaN hits and rays with a
//                zero direction component; those are detected and re-traced with trace_ref.
// Everything that decides accept/reject or a value of t uses the reference's arithmetic verbatim (fp32, source
// order, no FMA contraction, C++ min/max NaN behaviour) — SURVEY.md §8a "numerics contract".
#pragma once
#include "hrt_math.cuh"
#include "hrt_rng.cuh"
#include "hrt_types.h"

namespace hrt {

#if defined(__CUDA_ARCH__)
#define HRT_LD4(p) __ldg(p)
#else
#define HRT_LD4(p) (*(p))
#endif

struct Hit {                                 // 16-byte hit record written by the traversal kernel
    float t;
    uint32_t prim;                           // HRT_MISS on a miss
    int obj;                                 // world-list entry that reported the hit
    int flags;                               // bit 0: stale frontFace of an unwrapped mesh hit (see trace_ref)
};

struct Rec {                                 // hitRecord (hittable.h:8-25) + ids
    V3 p, n; float t, u, v; int front; int mat; uint32_t prim;
};

HRT_HD V3 xyz(const float4& a) { return v3(a.x, a.y, a.z); }

HRT_HD void set_face_normal(Rec& rec, const Ray& r, V3 outward)                  // hittable.h:21-24
{
    rec.front = dot(r.d, outward) < 0;
    rec.n = rec.front ? outward : -outward;
}

// ---- AABB::hit aabb.h:26-39 ------------------------------------------------------------------------------------
HRT_HD bool hit_aabb(V3 mn, V3 mx, const Ray& r, float t_min, float t_max)
{
    {
        float t0 = cmin((mn.x - r.o.x) / r.d.x, (mx.x - r.o.x) / r.d.x);
        float t1 = cmax((mn.x - r.o.x) / r.d.x, (mx.x - r.o.x) / r.d.x);
        t_min = cmax(t0, t_min); t_max = cmin(t1, t_max);
        if (t_max <= t_min) return false;
    }
    {
        float t0 = cmin((mn.y - r.o.y) / r.d.y, (mx.y - r.o.y) / r.d.y);
        float t1 = cmax((mn.y - r.o.y) / r.d.y, (mx.y - r.o.y) / r.d.y);
        t_min = cmax(t0, t_min); t_max = cmin(t1, t_max);
        if (t_max <= t_min) return false;
    }
    {
        float t0 = cmin((mn.z - r.o.z) / r.d.z, (mx.z - r.o.z) / r.d.z);
        float t1 = cmax((mn.z - r.o.z) / r.d.z, (mx.z - r.o.z) / r.d.z);
        t_min = cmax(t0, t_min); t_max = cmin(t1, t_max);
        if (t_max <= t_min) return false;
    }
    return true;
}

// Same test, also handing back the interval start it ends with (max of t_min and the three slab entries).
HRT_HD bool hit_aabb_entry(V3 mn, V3 mx, const Ray& r, float t_min, float t_max, float& entry)
{
    {
        float t0 = cmin((mn.x - r.o.x) / r.d.x, (mx.x - r.o.x) / r.d.x);
        float t1 = cmax((mn.x - r.o.x) / r.d.x, (mx.x - r.o.x) / r.d.x);
        t_min = cmax(t0, t_min); t_max = cmin(t1, t_max);
        if (t_max <= t_min) return false;
    }
    {
        float t0 = cmin((mn.y - r.o.y) / r.d.y, (mx.y - r.o.y) / r.d.y);
        float t1 = cmax((mn.y - r.o.y) / r.d.y, (mx.y - r.o.y) / r.d.y);
        t_min = cmax(t0, t_min); t_max = cmin(t1, t_max);
        if (t_max <= t_min) return false;
    }
    {
        float t0 = cmin((mn.z - r.o.z) / r.d.z, (mx.z - r.o.z) / r.d.z);
        float t1 = cmax((mn.z - r.o.z) / r.d.z, (mx.z - r.o.z) / r.d.z);
        t_min = cmax(t0, t_min); t_max = cmin(t1, t_max);
        if (t_max <= t_min) return false;
    }
    entry = t_min;
    return true;
}

// ---- Sphere::hit sphere.cpp:20-49 (t only) ---------------------------------------------------------------------
HRT_HD bool sphere_t(const float* f, const Ray& r, float t_min, float t_max, float& t)
{
    V3 center = v3(f[0], f[1], f[2]); float radius = f[3];
    V3 oc = r.o - center;
    float a = length(r.d) * length(r.d);
    float half_b = dot(oc, r.d);
    float c = length(oc) * length(oc) - radius * radius;
    float disc = half_b * half_b - a * c;
    if (disc < 0) return false;
    float sqrtd = sqrtf(disc);
    float root = (-half_b - sqrtd) / a;
    if (root < t_min || root > t_max) {
        root = (-half_b + sqrtd) / a;
        if (root < t_min || root > t_max) return false;
    }
    t = root;
    return true;
}

// ---- XYRect / XZRect / YZRect::hit aarect.h:12-39 (axis 0), 59-86 (axis 1), 106-133 (axis 2) --------------------
HRT_HD bool rect_t(int axis, float a0, float a1, float b0, float b1, float k, const Ray& r, float t_min, float t_max,
                   float& t, float& a, float& b)
{
    int ia = (axis == 0) ? 1 : 0, ib = (axis == 2) ? 1 : 2;
    float tt = (k - comp(r.o, axis)) / comp(r.d, axis);
    if (tt < t_min || tt > t_max) return false;
    a = comp(r.o, ia) + tt * comp(r.d, ia);
    b = comp(r.o, ib) + tt * comp(r.d, ib);
    if (a < a0 || a > a1 || b < b0 || b > b1) return false;
    t = tt;
    return true;
}

// Box (box.h:27-55): six rects in the order +z, -z, +y, -y, +x, -x; f = min xyz, max xyz
HRT_HD void box_face(const float* f, int face, int& axis, float& a0, float& a1, float& b0, float& b1, float& k)
{
    switch (face) {
    case 0: axis = 2; a0 = f[0]; a1 = f[3]; b0 = f[1]; b1 = f[4]; k = f[5]; break;
    case 1: axis = 2; a0 = f[0]; a1 = f[3]; b0 = f[1]; b1 = f[4]; k = f[2]; break;
    case 2: axis = 1; a0 = f[0]; a1 = f[3]; b0 = f[2]; b1 = f[5]; k = f[4]; break;
    case 3: axis = 1; a0 = f[0]; a1 = f[3]; b0 = f[2]; b1 = f[5]; k = f[1]; break;
    case 4: axis = 0; a0 = f[1]; a1 = f[4]; b0 = f[2]; b1 = f[5]; k = f[3]; break;
    default: axis = 0; a0 = f[1]; a1 = f[4]; b0 = f[2]; b1 = f[5]; k = f[0]; break;
    }
}
HRT_HD bool box_t(const float* f, const Ray& r, float t_min, float t_max, float& t, int& face)
{
    bool any = false; float closest = t_max;                                   // hittableList.cpp:4-21
    for (int i = 0; i < 6; ++i) {
        int axis; float a0, a1, b0, b1, k, tt, a, b;
        box_face(f, i, axis, a0, a1, b0, b1, k);
        if (rect_t(axis, a0, a1, b0, b1, k, r, t_min, closest, tt, a, b)) { any = true; closest = tt; t = tt; face = i; }
    }
    return any;
}

// ---- ITriangle::hit triangle.cpp:57-131 -------------------------------------------------------------------------
// t_min is ignored; the projection axis comes from the ray ORIGIN; a zero d[kZ] yields a NaN "hit".
HRT_HD bool itriangle(V3 P0, V3 P1, V3 P2, const Ray& r, float t_max, float& t, float& b0, float& b1, float& b2)
{
    V3 d = r.d, o = r.o;
    V3 p0t = P0 - o, p1t = P1 - o, p2t = P2 - o;
    int kZ = o.x > o.z ? (o.x > o.y ? 0 : 1) : 2;
    int kX = kZ + 1 == 3 ? 0 : kZ + 1;
    int kY = kX + 1 == 3 ? 0 : kX + 1;
    d = v3(comp(d, kX), comp(d, kY), comp(d, kZ));
    p0t = v3(comp(p0t, kX), comp(p0t, kY), comp(p0t, kZ));
    p1t = v3(comp(p1t, kX), comp(p1t, kY), comp(p1t, kZ));
    p2t = v3(comp(p2t, kX), comp(p2t, kY), comp(p2t, kZ));
    float sX = -d.x / d.z;
    float sY = -d.y / d.z;
    float sZ = 1.0f / d.z;
    p0t.x += sX * p0t.z; p0t.y += sY * p0t.z;
    p1t.x += sX * p1t.z; p1t.y += sY * p1t.z;
    p2t.x += sX * p2t.z; p2t.y += sY * p2t.z;
    float e0 = p1t.x * p2t.y - p1t.y * p2t.x;
    float e1 = p2t.x * p0t.y - p2t.y * p0t.x;
    float e2 = p0t.x * p1t.y - p0t.y * p1t.x;
    if ((e0 < 0 || e1 < 0 || e2 < 0) && (e0 > 0 || e1 > 0 || e2 > 0)) return false;
    float det = e0 + e1 + e2;
    if (det == 0) return false;
    p0t.z *= sZ; p1t.z *= sZ; p2t.z *= sZ;
    float tScaled = e0 * p0t.z + e1 * p1t.z + e2 * p2t.z;
    if (det < 0 && (tScaled >= 0 || tScaled < t_max * det)) return false;
    else if (det > 0 && (tScaled <= 0 || tScaled > t_max * det)) return false;
    float invDet = 1 / det;
    b0 = e0 * invDet; b1 = e1 * invDet; b2 = e2 * invDet;
    t = tScaled * invDet;
    return true;
}

// ---- instance wrappers: the ray half of translate.cpp:7-19, scale.cpp:11-27, rotateQuat.cpp:44-66, rotateY.cpp:44-75
HRT_HD Ray xform_ray(const DXform& x, const Ray& r)
{
    Ray m;
    switch (x.kind) {
    case HRT_XF_TRANSLATE: m.o = r.o - v3(x.a, x.b, x.c); m.d = r.d; break;
    case HRT_XF_SCALE: m.o = r.o / v3(x.a, x.b, x.c); m.d = r.d / v3(x.a, x.b, x.c); break;
    case HRT_XF_ROTATE_QUAT: {
        Quat q; q.w = x.a; q.x = x.b; q.y = x.c; q.z = x.d;
        Quat inv = qconj(q);
        m.o = qrot(inv, r.o); m.d = normalize(qrot(inv, r.d));
    } break;
    default: {                                                                  // HRT_XF_ROTATE_Y: a = sin, b = cos
        m.o = r.o; m.d = r.d;
        m.o.x = x.b * r.o.x - x.a * r.o.z;
        m.o.z = x.a * r.o.x + x.b * r.o.z;
        m.d.x = x.b * r.d.x - x.a * r.d.z;
        m.d.z = x.a * r.d.x + x.b * r.d.z;
    } break;
    }
    return m;
}
HRT_HD Ray to_local(const DScene& s, const DObject& ob, Ray r)
{
    for (int l = ob.xf_count - 1; l >= 0; --l) r = xform_ray(s.xforms[ob.xf_first + l], r);
    return r;
}

// ---- BVHNode::hit bvh.cpp:69-78, iteratively.  The recursion hands the right child t_max = rec.t when the left
// child hit and its own t_max otherwise, which is the t of the most recent accepted hit (or the mesh's incoming
// t_max) — one running `closest`, assigned (not min'ed) on every accept.
HRT_HDN inline bool mesh_hit_ref(const DMesh& m, const Ray& r, float t_min, float t_max, float& t, uint32_t& prim)
{
    if (m.ref_root < 0) return false;
    int stack[REF_STACK]; int sp = 0; stack[sp++] = m.ref_root;
    float closest = t_max; bool any = false;
    while (sp > 0) {
        int c = stack[--sp];
        if (c >= 0) {
            float4 a = HRT_LD4(m.ref_nodes + 2 * (size_t)c), b = HRT_LD4(m.ref_nodes + 2 * (size_t)c + 1);
            if (!hit_aabb(xyz(a), xyz(b), r, t_min, closest)) continue;
            int left = float_as_int(a.w), right = float_as_int(b.w);
            if (right != left && sp < REF_STACK - 1) stack[sp++] = right;        // a 1-triangle node lists it twice; the second test cannot change (t, id)
            stack[sp++] = left;
        } else {
            size_t tri = (size_t)(~c);
            float4 p0 = HRT_LD4(m.tris + 3 * tri), p1 = HRT_LD4(m.tris + 3 * tri + 1), p2 = HRT_LD4(m.tris + 3 * tri + 2);
            float tt, b0, b1, b2;
            if (itriangle(xyz(p0), xyz(p1), xyz(p2), r, closest, tt, b0, b1, b2)) { closest = tt; any = true; t = tt; prim = m.prim_base + (uint32_t)tri; }
        }
    }
    return any;
}

// shape under its wrapper chain, reference order; returns (t, prim) only
HRT_HDN inline bool hit_wrapped_ref(const DScene& s, const DObject& ob, const Ray& r, float t_min, float t_max, float& t, uint32_t& prim)
{
    Ray l = to_local(s, ob, r);
    switch (ob.kind) {
    case SHAPE_SPHERE: prim = ob.prim_base; return sphere_t(ob.f, l, t_min, t_max, t);
    case SHAPE_RECT: { float a, b; prim = ob.prim_base; return rect_t(ob.axis, ob.f[0], ob.f[1], ob.f[2], ob.f[3], ob.f[4], l, t_min, t_max, t, a, b); }
    case SHAPE_BOX: { int face = 0; bool h = box_t(ob.f, l, t_min, t_max, t, face); prim = ob.prim_base + (uint32_t)face; return h; }
    case SHAPE_MESH: return mesh_hit_ref(s.meshes[ob.mesh], l, t_min, t_max, t, prim);
    }
    return false;
}

// one world-list entry, including ConstantMedium::hit constantMedium.cpp:4-38
HRT_HDN inline bool hit_object_ref(const DScene& s, int i, const Ray& r, float t_min, float t_max, float& t, uint32_t& prim,
                                   const RngKey& key, uint32_t bounce)
{
    const DObject& ob = s.objects[i];
    if (!ob.medium) return hit_wrapped_ref(s, ob, r, t_min, t_max, t, prim);
    const float inf = hrt_inf();
    float t1, t2; uint32_t p1, p2;
    if (!hit_wrapped_ref(s, ob, r, -inf, inf, t1, p1)) return false;
    if (!hit_wrapped_ref(s, ob, r, t1 + 0.0001f, inf, t2, p2)) return false;
    if (t1 < t_min) t1 = t_min;
    if (t2 > t_max) t2 = t_max;
    if (t1 >= t2) return false;
    if (t1 < 0) t1 = 0;
    const float ray_length = length(r.d);
    const float dist_inside = (t2 - t1) * ray_length;
    float xi = linear_f(rng_draw(key, bounce, SLOT_MEDIUM + (uint32_t)i).x, key.mode, 0.0f, 1.0f);
    const float hit_distance = (float)((double)ob.neg_inv_density * log((double)xi));   // constantMedium.cpp:25 (double log)
    if (hit_distance > dist_inside) return false;
    t = t1 + hit_distance / ray_length;
    prim = ob.medium_prim;
    return true;
}

// ---- full hit record of a known hit: what hit() leaves in `rec` for (object, primitive, t) ------------------------
// Re-evaluates the single primitive on the same local ray, so every value is bit-identical to what a traversal that
// carried the whole record would have produced.
HRT_HDN inline void eval_hit(const DScene& s, const Ray& r, const Hit& h, Rec& rec)
{
    const DObject& ob = s.objects[h.obj];
    rec.prim = h.prim; rec.t = h.t; rec.u = 0; rec.v = 0; rec.front = 0;
    if (ob.medium) {                                                            // constantMedium.cpp:30-36
        rec.p = ray_at(r, h.t); rec.n = v3(1, 0, 0); rec.front = 1; rec.mat = MAT_MEDIUM_BASE + h.obj;
        return;
    }
    Ray lv[MAX_XF + 1];
    lv[ob.xf_count] = r;
    for (int l = ob.xf_count - 1; l >= 0; --l) lv[l] = xform_ray(s.xforms[ob.xf_first + l], lv[l + 1]);
    const Ray& lr = lv[0];
    rec.mat = ob.material;
    switch (ob.kind) {
    case SHAPE_SPHERE: {                                                        // sphere.cpp:37-46, getSphereUV 4-18
        V3 center = v3(ob.f[0], ob.f[1], ob.f[2]);
        rec.p = ray_at(lr, h.t);
        V3 outward = (rec.p - center) / ob.f[3];
        set_face_normal(rec, lr, outward);
        const float pi = 3.14159265358979323846264338327950288f;
        float theta = acosf(-outward.y);
        float phi = atan2f(-outward.z, outward.x) + pi;
        rec.u = phi / (2 * pi);
        rec.v = theta / pi;
    } break;
    case SHAPE_RECT: case SHAPE_BOX: {
        int axis = ob.axis; float a0 = ob.f[0], a1 = ob.f[1], b0 = ob.f[2], b1 = ob.f[3], k = ob.f[4];
        if (ob.kind == SHAPE_BOX) box_face(ob.f, (int)(h.prim - ob.prim_base), axis, a0, a1, b0, b1, k);
        int ia = (axis == 0) ? 1 : 0, ib = (axis == 2) ? 1 : 2;
        float a = comp(lr.o, ia) + h.t * comp(lr.d, ia);
        float b = comp(lr.o, ib) + h.t * comp(lr.d, ib);
        rec.u = (a - a0) / (a1 - a0);
        rec.v = (b - b0) / (b1 - b0);
        set_face_normal(rec, lr, v3(axis == 0 ? 1.0f : 0.0f, axis == 1 ? 1.0f : 0.0f, axis == 2 ? 1.0f : 0.0f));
        rec.p = ray_at(lr, h.t);
    } break;
    case SHAPE_MESH: {                                                          // triangle.cpp:113-129
        const DMesh& m = s.meshes[ob.mesh];
        size_t tri = (size_t)(h.prim - m.prim_base);
        float4 p0 = HRT_LD4(m.tris + 3 * tri), p1 = HRT_LD4(m.tris + 3 * tri + 1), p2 = HRT_LD4(m.tris + 3 * tri + 2);
        float tt = h.t, b0 = 0, b1 = 0, b2 = 0;
        itriangle(xyz(p0), xyz(p1), xyz(p2), lr, hrt_inf(), tt, b0, b1, b2);
        float4 q0 = HRT_LD4(m.attrs + 4 * tri), q1 = HRT_LD4(m.attrs + 4 * tri + 1), q2 = HRT_LD4(m.attrs + 4 * tri + 2), q3 = HRT_LD4(m.attrs + 4 * tri + 3);
        rec.p = ray_at(lr, h.t);
        rec.n = b0 * xyz(q0) + b1 * xyz(q1) + b2 * xyz(q2);
        rec.u = b0 * q0.w + b1 * q2.w + b2 * q3.y;
        rec.v = b0 * q1.w + b1 * q3.x + b2 * q3.z;
        rec.front = h.flags & 1;                                                // ITriangle::hit never writes frontFace
        rec.mat = m.material;
    } break;
    }
    // the way back up: translate.cpp:14-16, scale.cpp:21-24, rotateQuat.cpp:57-63, rotateY.cpp:60-72
    for (int l = 0; l < ob.xf_count; ++l) {
        const DXform& x = s.xforms[ob.xf_first + l];
        V3 n = rec.n;
        switch (x.kind) {
        case HRT_XF_TRANSLATE: rec.p = rec.p + v3(x.a, x.b, x.c); break;
        case HRT_XF_SCALE: rec.p = rec.p * v3(x.a, x.b, x.c); break;
        case HRT_XF_ROTATE_QUAT: { Quat q; q.w = x.a; q.x = x.b; q.y = x.c; q.z = x.d; rec.p = qrot(q, rec.p); n = qrot(q, rec.n); } break;
        default: {
            V3 p = rec.p;
            p.x = x.b * rec.p.x + x.a * rec.p.z;
            p.z = -x.a * rec.p.x + x.b * rec.p.z;
            n.x = x.b * rec.n.x + x.a * rec.n.z;
            n.z = -x.a * rec.n.x + x.b * rec.n.z;
            rec.p = p;
        } break;
        }
        set_face_normal(rec, lv[l], n);
    }
}

HRT_HD bool is_bare_mesh(const DObject& ob) { return ob.kind == SHAPE_MESH && ob.xf_count == 0 && !ob.medium; }

// ---- HittableList::hit hittableList.cpp:4-21 over the world list ---------------------------------------------------
// The list reuses one tempRec, and ITriangle::hit never writes frontFace, so a hit on a mesh that has no wrapper
// reports the frontFace left behind by the latest earlier successful hit of another kind (false if there is none —
// the reference reads an indeterminate value there).  flags bit 0 carries it.
HRT_HDN inline Hit trace_ref(const DScene& s, const Ray& r, float t_min, float t_max, const RngKey& key, uint32_t bounce)
{
    Hit h; h.t = 0.0f; h.prim = HRT_MISS; h.obj = -1; h.flags = 0;
    Hit stale; stale.obj = -1; stale.t = 0; stale.prim = HRT_MISS; stale.flags = 0;
    float closest = t_max;
    for (int i = 0; i < s.n_objects; ++i) {
        float t; uint32_t prim;
        if (hit_object_ref(s, i, r, t_min, closest, t, prim, key, bounce)) {
            closest = t; h.t = t; h.prim = prim; h.obj = i;
            if (!is_bare_mesh(s.objects[i])) stale = h;
        }
    }
    if (h.obj >= 0 && stale.obj >= 0 && is_bare_mesh(s.objects[h.obj])) {
        Rec tmp; eval_hit(s, r, stale, tmp);
        h.flags = tmp.front ? 1 : 0;
    }
    return h;
}

} // namespace hrt
