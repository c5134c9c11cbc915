// oracle.cpp — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// CPU restatement of the path-tracing hot path of Todegal/HobbyRaytracer (camera ray -> closest hit -> scatter ->
// accumulate -> tonemap), written from the reference's behaviour without GLM or any reference source.  Every
// function cites the reference file:line it restates.  It operates on the flattened scene description of
// include/hrt.h (prefix orc_) — the same description the CUDA library consumes — and follows the reference's
// sequential visiting order (flat world list, BVH left-then-right), its arithmetic order (no FMA contraction:
// compile with -ffp-contract=off), its RNG consumption order (glm::linearRand on a 31-bit rand() stream) and its
// quirks.
//
// PARITY PINNED: tests/test_oracle_vs_ref.py checks this file bit-for-bit against the unmodified reference
// (oracle/_ref/libhrt_ref.so, built by oracle/Makefile from /root/reference) — hit ids, t, normals, uvs, scatter
// outputs and whole renders on the same rand() stream — and tests/golden/ holds vectors generated from the
// reference for machines where /root/reference is absent.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this library.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <limits>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "../include/hrt.h"

namespace {

// ------------------------------------------------------------------------------------------------------------
// vector arithmetic in GLM's evaluation order (dependencies/glm/glm/detail/func_geometric.inl)
// ------------------------------------------------------------------------------------------------------------
struct V3 { float x, y, z; };
inline V3 v3(float a, float b, float c) { return V3{a, b, c}; }
inline V3 v3(float a) { return V3{a, a, a}; }
inline V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator-(V3 a) { return V3{-a.x, -a.y, -a.z}; }
inline V3 operator*(V3 a, V3 b) { return V3{a.x * b.x, a.y * b.y, a.z * b.z}; }
inline V3 operator/(V3 a, V3 b) { return V3{a.x / b.x, a.y / b.y, a.z / b.z}; }
inline V3 operator*(V3 a, float s) { return V3{a.x * s, a.y * s, a.z * s}; }
inline V3 operator*(float s, V3 a) { return V3{s * a.x, s * a.y, s * a.z}; }
inline V3 operator/(V3 a, float s) { return V3{a.x / s, a.y / s, a.z / s}; }
inline float comp(const V3& a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }
// func_geometric.inl:48-55: tmp = a*b; tmp.x + tmp.y + tmp.z
inline float dot(V3 a, V3 b) { V3 t = a * b; return t.x + t.y + t.z; }
// func_geometric.inl:68-79
inline V3 cross(V3 x, V3 y) { return V3{x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y}; }
inline float length(V3 a) { return sqrtf(dot(a, a)); }                       // func_geometric.inl:8-14
inline V3 normalize(V3 a) { return a * (1.0f / sqrtf(dot(a, a))); }          // func_geometric.inl:82-90 (v * inversesqrt)
inline V3 reflect(V3 I, V3 N) { return I - N * dot(N, I) * 2.0f; }           // func_geometric.inl:104-110
inline V3 refract(V3 I, V3 N, float eta)                                     // func_geometric.inl:113-123
{
    float d = dot(N, I);
    float k = 1.0f - eta * eta * (1.0f - d * d);
    return (k >= 0.0f) ? (eta * I - (eta * d + std::sqrt(k)) * N) : v3(0.0f);
}
// std::min / std::max / glm::min / glm::max: (b < a) ? b : a  and  (a < b) ? b : a  (aabb.h:29-34 relies on this)
inline float cmin(float a, float b) { return (b < a) ? b : a; }
inline float cmax(float a, float b) { return (a < b) ? b : a; }

struct Quat { float w, x, y, z; };
// type_quat.inl:347-354
inline V3 qrot(Quat q, V3 v)
{
    V3 qv = v3(q.x, q.y, q.z);
    V3 uv = cross(qv, v);
    V3 uuv = cross(qv, uv);
    return v + ((uv * q.w) + uuv) * 2.0f;
}
inline Quat qconj(Quat q) { return Quat{q.w, -q.x, -q.y, -q.z}; }
// qua(vec3 eulerAngle) type_quat.inl:208-217
inline Quat quat_from_euler(V3 e)
{
    V3 h = e * 0.5f;
    V3 c = v3(std::cos(h.x), std::cos(h.y), std::cos(h.z));
    V3 s = v3(std::sin(h.x), std::sin(h.y), std::sin(h.z));
    Quat q;
    q.w = c.x * c.y * c.z + s.x * s.y * s.z;
    q.x = s.x * c.y * c.z - c.x * s.y * s.z;
    q.y = c.x * s.y * c.z + s.x * c.y * s.z;
    q.z = c.x * c.y * s.z - s.x * s.y * c.z;
    return q;
}
inline float radiansf(float deg) { return deg * 0.01745329251994329576923690768489f; }   // glm::radians

// ------------------------------------------------------------------------------------------------------------
// RNG: glm::linearRand & friends (dependencies/glm/glm/gtc/random.inl) on a per-thread 31-bit rand() stream.
// The stream definition is shared with ref_harness.cpp's rand() interposer.
// Byte assembly as compiled by g++ 13 (pinned by tests against the reference build): the FIRST byte drawn is the
// LEAST significant; vector variants draw z, y, x per call.
// ------------------------------------------------------------------------------------------------------------
inline uint64_t mix64(uint64_t z)
{
    z += 0x9e3779b97f4a7c15ULL;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
    return z ^ (z >> 31);
}
struct Rng {
    uint64_t s;
    void seed(uint64_t seed, uint64_t stream) { s = mix64(seed ^ mix64(stream)); }
    int rand31() { s = s * 6364136223846793005ULL + 1442695040888963407ULL; return (int)((s >> 33) & 0x7fffffffULL); }
    uint32_t byte() { return (uint32_t)(rand31() % 255); }                         // random.inl:25
    uint32_t u32() { uint32_t b0 = byte(), b1 = byte(), b2 = byte(), b3 = byte(); return b0 | (b1 << 8) | (b2 << 16) | (b3 << 24); }
    uint64_t u64() { uint64_t lo = u32(); uint64_t hi = u32(); return lo | (hi << 32); }
    float linear(float mn, float mx)                                               // random.inl:176-183 (float)
    { return (float)u32() / static_cast<float>(std::numeric_limits<uint32_t>::max()) * (mx - mn) + mn; }
    double lineard(double mn, double mx)                                           // random.inl:185-192 (double)
    { return (double)u64() / static_cast<double>(std::numeric_limits<uint64_t>::max()) * (mx - mn) + mn; }
    int lineari(int mn, int mx) { return (int)(u32() % (uint32_t)(mx + 1 - mn)) + mn; }   // random.inl:140-147
    V3 linear3(float mn, float mx)
    {
        uint32_t c[3] = {0, 0, 0};
        for (int k = 0; k < 4; ++k) { uint32_t z = byte(), y = byte(), x = byte(); c[0] |= x << (8 * k); c[1] |= y << (8 * k); c[2] |= z << (8 * k); }
        const float d = static_cast<float>(std::numeric_limits<uint32_t>::max());
        return v3((float)c[0] / d * (mx - mn) + mn, (float)c[1] / d * (mx - mn) + mn, (float)c[2] / d * (mx - mn) + mn);
    }
    V3 spherical()                                                                 // random.inl:290-302, Radius = 1
    {
        float theta = linear(0.0f, 6.283185307179586476925286766559f);
        float phi = std::acos(linear(-1.0f, 1.0f));
        float x = std::sin(phi) * std::cos(theta);
        float y = std::sin(phi) * std::sin(theta);
        float z = std::cos(phi);
        return v3(x, y, z) * 1.0f;
    }
    V3 ball()                                                                      // random.inl:261-278, Radius = 1
    {
        V3 r; float len;
        do { r = linear3(-1.0f, 1.0f); len = length(r); } while (len > 1.0f);
        return r;
    }
};

// ------------------------------------------------------------------------------------------------------------
// scene description
// ------------------------------------------------------------------------------------------------------------
struct Texture {
    int kind;            // 0 solid 1 checker 2 image 3 env
    V3 rgb; int even, odd; int w, h, ch;
    std::vector<uint8_t> u8; std::vector<float> f32;
};
struct Shape {
    int kind;            // 0 sphere 1 rect 2 box 3 mesh 4 medium(pseudo)
    float f[8]; int axis; int material;
    std::vector<float> pos, nrm, uv; int64_t n_tri = 0;
    int64_t prim_base = 0, prim_count = 0;
    // mesh BVH in reference topology: child >= 0 -> node index, child < 0 -> ~triangle index
    struct Node { V3 mn, mx; int left, right; };
    std::vector<Node> nodes; int root = -1;
    std::vector<V3> tri_mn, tri_mx;
};
struct Xf { int kind; V3 v; Quat q; float s, c; };
struct Object {
    int shape; std::vector<Xf> ops;     // innermost first
    bool medium = false; float neg_inv_density = 0; int phase_tex = -1; V3 phase_rgb; int pseudo_shape = -1;
};

struct Rec {                            // hitRecord (hittable.h:8-25) + primitive id
    V3 p, n; float t, u, v; bool front; int mat; uint32_t prim;
};
struct Ray { V3 o, d; };
inline V3 at(const Ray& r, float t) { return r.o + (t * r.d); }                   // ray.h:10
inline void set_face_normal(Rec& rec, const Ray& r, V3 outward)                   // hittable.h:21-24
{
    rec.front = dot(r.d, outward) < 0;
    rec.n = rec.front ? outward : -outward;
}

} // namespace

struct orc_scene {
    std::vector<Texture> textures;
    std::vector<hrt_material> materials;
    std::vector<Shape> shapes;
    std::vector<Object> objects;
    int64_t n_prims = 0;
    int background = -1;
    bool committed = false;
    int n_user_materials = 0;
};

namespace {

thread_local std::string g_err;
constexpr int MAT_MEDIUM_BASE = 1 << 24;   // rec.mat >= this: Isotropic phase function of object (mat - base)

// ------------------------------------------------------------------------------------------------------------
// textures (texture.h / texture.cpp)
// ------------------------------------------------------------------------------------------------------------
V3 tex_value(const orc_scene& s, int id, float u, float v, V3 p)
{
    const Texture& t = s.textures[id];
    switch (t.kind) {
    case 0: return t.rgb;                                                         // texture.h:18-21
    case 1: {                                                                     // texture.cpp:17-28
        float sines = std::sin(10 * p.x) * std::sin(10 * p.y) * std::sin(10 * p.z);
        return sines < 0 ? tex_value(s, t.odd, u, v, p) : tex_value(s, t.even, u, v, p);
    }
    case 2: {                                                                     // texture.cpp:53-74
        if (t.u8.empty()) return v3(0, 1, 1);
        u = cmin(cmax(u, 0.0f), 1.0f);                                            // glm::clamp = min(max(x,lo),hi)
        v = 1.0f - cmin(cmax(v, 0.0f), 1.0f);
        int i = static_cast<int>(u * t.w);
        int j = static_cast<int>(v * t.h);
        if (i >= t.w) i = t.w - 1;
        if (j >= t.h) j = t.h - 1;
        const float sc = 1.0f / 255.0f;
        int px = j * (3 * t.w) + i * 3;
        return v3(sc * t.u8[px], sc * t.u8[px + 1], sc * t.u8[px + 2]);
    }
    case 3: {                                                                     // texture.cpp:76-97
        if (t.f32.empty()) return v3(0, 1, 1);
        u = cmin(cmax(u, 0.0f), 1.0f);
        v = cmin(cmax(v, 0.0f), 1.0f);
        int i = static_cast<int>((u * (t.w - 1)) + 0.5f);
        int j = static_cast<int>((v * (t.h - 1)) + 0.5f);
        size_t b = ((size_t)j * t.w + i) * t.ch;
        return v3(t.f32[b], t.f32[b + 1], t.f32[b + 2]);
    }
    }
    return v3(0);
}
V3 matvec3(const orc_scene& s, const float c[3], int tex, float u, float v, V3 p)          // material.h:20-31
{ return tex >= 0 ? tex_value(s, tex, u, v, p) : v3(c[0], c[1], c[2]); }
float matscalar(const orc_scene& s, float c, int tex, float u, float v, V3 p)              // material.h:43-54
{ return tex >= 0 ? length(tex_value(s, tex, u, v, p)) : c; }

// ------------------------------------------------------------------------------------------------------------
// primitives
// ------------------------------------------------------------------------------------------------------------
bool hit_sphere(const float* f, int mat, uint32_t prim, const Ray& r, float t_min, float t_max, Rec& rec)   // sphere.cpp:20-49
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
    rec.t = root;
    rec.p = at(r, rec.t);
    V3 outward = (rec.p - center) / radius;
    set_face_normal(rec, r, outward);
    // getSphereUV sphere.cpp:4-18
    const float pi = 3.14159265358979323846264338327950288f;
    float theta = std::acos(-outward.y);
    float phi = std::atan2(-outward.z, outward.x) + pi;
    rec.u = phi / (2 * pi);
    rec.v = theta / pi;
    rec.mat = mat; rec.prim = prim;
    return true;
}

// aarect.h:12-39 (axis 0, YZ), 59-86 (axis 1, XZ), 106-133 (axis 2, XY): params a0,a1,b0,b1,k
bool hit_rect(int axis, float a0, float a1, float b0, float b1, float k, int mat, uint32_t prim,
              const Ray& r, float t_min, float t_max, Rec& rec)
{
    int ka = axis, ia = (axis == 0) ? 1 : 0, ib = (axis == 2) ? 1 : 2;
    float t = (k - comp(r.o, ka)) / comp(r.d, ka);
    if (t < t_min || t > t_max) return false;
    float a = comp(r.o, ia) + t * comp(r.d, ia);
    float b = comp(r.o, ib) + t * comp(r.d, ib);
    if (a < a0 || a > a1 || b < b0 || b > b1) return false;
    rec.u = (a - a0) / (a1 - a0);
    rec.v = (b - b0) / (b1 - b0);
    rec.t = t;
    V3 outward = v3(axis == 0 ? 1.0f : 0.0f, axis == 1 ? 1.0f : 0.0f, axis == 2 ? 1.0f : 0.0f);
    set_face_normal(rec, r, outward);
    rec.mat = mat; rec.prim = prim;
    rec.p = at(r, t);
    return true;
}

// Box (box.h:27-55): HittableList of 6 rects; f = center(3), dimensions(3)
bool hit_box(const float* f, int mat, uint32_t prim, const Ray& r, float t_min, float t_max, Rec& rec)
{
    V3 c = v3(f[0], f[1], f[2]), d = v3(f[3], f[4], f[5]);
    V3 mn = c - d / 2.0f, mx = c + d / 2.0f;                                       // box.h:41-44
    Rec temp = rec; bool any = false; float closest = t_max;                        // hittableList.cpp:4-21
    if (hit_rect(2, mn.x, mx.x, mn.y, mx.y, mx.z, mat, prim + 0, r, t_min, closest, temp)) { any = true; closest = temp.t; rec = temp; }
    if (hit_rect(2, mn.x, mx.x, mn.y, mx.y, mn.z, mat, prim + 1, r, t_min, closest, temp)) { any = true; closest = temp.t; rec = temp; }
    if (hit_rect(1, mn.x, mx.x, mn.z, mx.z, mx.y, mat, prim + 2, r, t_min, closest, temp)) { any = true; closest = temp.t; rec = temp; }
    if (hit_rect(1, mn.x, mx.x, mn.z, mx.z, mn.y, mat, prim + 3, r, t_min, closest, temp)) { any = true; closest = temp.t; rec = temp; }
    if (hit_rect(0, mn.y, mx.y, mn.z, mx.z, mx.x, mat, prim + 4, r, t_min, closest, temp)) { any = true; closest = temp.t; rec = temp; }
    if (hit_rect(0, mn.y, mx.y, mn.z, mx.z, mn.x, mat, prim + 5, r, t_min, closest, temp)) { any = true; closest = temp.t; rec = temp; }
    return any;
}

// ITriangle::hit triangle.cpp:57-131.  t_min is ignored; the projection axis comes from the ray ORIGIN; frontFace is
// not written; the normal is the unnormalised barycentric blend.
bool hit_itriangle(const float* P, const float* N, const float* T, int mat, uint32_t prim,
                   const Ray& r, float /*t_min*/, float t_max, Rec& rec)
{
    V3 d = r.d, o = r.o;
    V3 p0t = v3(P[0], P[1], P[2]) - o;
    V3 p1t = v3(P[3], P[4], P[5]) - o;
    V3 p2t = v3(P[6], P[7], P[8]) - o;
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
    float b0 = e0 * invDet, b1 = e1 * invDet, b2 = e2 * invDet;
    float t = tScaled * invDet;
    rec.t = t;
    rec.p = at(r, rec.t);
    rec.mat = mat; rec.prim = prim;
    V3 n = b0 * v3(N[0], N[1], N[2]) + b1 * v3(N[3], N[4], N[5]) + b2 * v3(N[6], N[7], N[8]);
    rec.n = n;
    rec.u = b0 * T[0] + b1 * T[2] + b2 * T[4];
    rec.v = b0 * T[1] + b1 * T[3] + b2 * T[5];
    return true;
}

// AABB::hit aabb.h:26-39
bool hit_aabb(V3 mn, V3 mx, const Ray& r, float t_min, float t_max)
{
    for (int a = 0; a < 3; ++a) {
        float lo = comp(mn, a), hi = comp(mx, a), o = comp(r.o, a), d = comp(r.d, a);
        float t0 = cmin((lo - o) / d, (hi - o) / d);
        float t1 = cmax((lo - o) / d, (hi - o) / d);
        t_min = cmax(t0, t_min);
        t_max = cmin(t1, t_max);
        if (t_max <= t_min) return false;
    }
    return true;
}

// BVHNode::hit bvh.cpp:69-78 over the mesh of `sh`; a child < 0 is triangle ~child.
bool hit_bvh(const Shape& sh, int node, const Ray& r, float t_min, float t_max, Rec& rec)
{
    const Shape::Node& n = sh.nodes[node];
    if (!hit_aabb(n.mn, n.mx, r, t_min, t_max)) return false;
    auto child = [&](int c, float tmax) -> bool {
        if (c >= 0) return hit_bvh(sh, c, r, t_min, tmax, rec);
        int64_t i = ~c;
        return hit_itriangle(&sh.pos[i * 9], &sh.nrm[i * 9], &sh.uv[i * 6], sh.material, (uint32_t)(sh.prim_base + i), r, t_min, tmax, rec);
    };
    bool hl = child(n.left, t_max);
    bool hr = child(n.right, hl ? rec.t : t_max);
    return hl || hr;
}

bool hit_shape(const orc_scene& s, const Shape& sh, const Ray& r, float t_min, float t_max, Rec& rec)
{
    switch (sh.kind) {
    case 0: return hit_sphere(sh.f, sh.material, (uint32_t)sh.prim_base, r, t_min, t_max, rec);
    case 1: return hit_rect(sh.axis, sh.f[0], sh.f[1], sh.f[2], sh.f[3], sh.f[4], sh.material, (uint32_t)sh.prim_base, r, t_min, t_max, rec);
    case 2: return hit_box(sh.f, sh.material, (uint32_t)sh.prim_base, r, t_min, t_max, rec);
    case 3: if (sh.root < 0) return false;    // empty mesh: BVHNode over 0 objects has null children in the reference (would crash)
            return hit_bvh(sh, sh.root, r, t_min, t_max, rec);
    }
    return false;
}

// wrapper chain, outermost = ops[level-1]
bool hit_wrapped(const orc_scene& s, const Object& ob, int level, const Ray& r, float t_min, float t_max, Rec& rec)
{
    if (level == 0) return hit_shape(s, s.shapes[ob.shape], r, t_min, t_max, rec);
    const Xf& x = ob.ops[level - 1];
    switch (x.kind) {
    case HRT_XF_TRANSLATE: {                                                       // translate.cpp:7-19
        Ray m{r.o - x.v, r.d};
        if (!hit_wrapped(s, ob, level - 1, m, t_min, t_max, rec)) return false;
        rec.p = rec.p + x.v;
        set_face_normal(rec, m, rec.n);
        return true;
    }
    case HRT_XF_SCALE: {                                                           // scale.cpp:11-27
        Ray m{r.o / x.v, r.d / x.v};
        if (!hit_wrapped(s, ob, level - 1, m, t_min, t_max, rec)) return false;
        rec.p = rec.p * x.v;
        set_face_normal(rec, m, rec.n);
        return true;
    }
    case HRT_XF_ROTATE_QUAT: {                                                     // rotateQuat.cpp:44-66
        Quat inv = qconj(x.q);
        Ray m{qrot(inv, r.o), normalize(qrot(inv, r.d))};
        if (!hit_wrapped(s, ob, level - 1, m, t_min, t_max, rec)) return false;
        rec.p = qrot(x.q, rec.p);
        rec.n = qrot(x.q, rec.n);
        set_face_normal(rec, m, rec.n);
        return true;
    }
    case HRT_XF_ROTATE_Y: {                                                        // rotateY.cpp:44-75
        V3 o = r.o, d = r.d;
        o.x = x.c * r.o.x - x.s * r.o.z;
        o.z = x.s * r.o.x + x.c * r.o.z;
        d.x = x.c * r.d.x - x.s * r.d.z;
        d.z = x.s * r.d.x + x.c * r.d.z;
        Ray m{o, d};
        if (!hit_wrapped(s, ob, level - 1, m, t_min, t_max, rec)) return false;
        V3 p = rec.p, n = rec.n;
        p.x = x.c * rec.p.x + x.s * rec.p.z;
        p.z = -x.s * rec.p.x + x.c * rec.p.z;
        n.x = x.c * rec.n.x + x.s * rec.n.z;
        n.z = -x.s * rec.n.x + x.c * rec.n.z;
        rec.p = p;
        set_face_normal(rec, m, n);
        return true;
    }
    }
    return false;
}

bool hit_object(const orc_scene& s, int obj_index, const Ray& r, float t_min, float t_max, Rec& rec, Rng& rng)
{
    const Object& ob = s.objects[obj_index];
    int nl = (int)ob.ops.size();
    if (!ob.medium) return hit_wrapped(s, ob, nl, r, t_min, t_max, rec);
    // ConstantMedium::hit constantMedium.cpp:4-38
    const float inf = std::numeric_limits<float>::infinity();
    Rec rec1, rec2; rec1 = rec; rec2 = rec;
    if (!hit_wrapped(s, ob, nl, r, -inf, inf, rec1)) return false;
    if (!hit_wrapped(s, ob, nl, r, rec1.t + 0.0001f, inf, rec2)) return false;
    if (rec1.t < t_min) rec1.t = t_min;
    if (rec2.t > t_max) rec2.t = t_max;
    if (rec1.t >= rec2.t) return false;
    if (rec1.t < 0) rec1.t = 0;
    const float ray_length = length(r.d);
    const float dist_inside = (rec2.t - rec1.t) * ray_length;
    const float hit_distance = ob.neg_inv_density * ::log((double)rng.linear(0.0f, 1.0f));   // constantMedium.cpp:25 resolves to the double log
    if (hit_distance > dist_inside) return false;
    rec.t = rec1.t + hit_distance / ray_length;
    rec.p = at(r, rec.t);
    rec.n = v3(1, 0, 0);
    rec.front = true;
    rec.mat = MAT_MEDIUM_BASE + obj_index;
    rec.prim = (uint32_t)s.shapes[ob.pseudo_shape].prim_base;
    return true;
}

// HittableList::hit hittableList.cpp:4-21 over the world list.  tempRec starts with front=false, u=v=0 where the
// reference has indeterminate stack contents (ITriangle never writes frontFace; ConstantMedium never writes u,v).
bool hit_world(const orc_scene& s, const Ray& r, float t_min, float t_max, Rec& rec, Rng& rng)
{
    Rec temp; temp.front = false; temp.u = temp.v = 0; temp.t = 0; temp.p = v3(0); temp.n = v3(0); temp.mat = -1; temp.prim = HRT_MISS;
    bool any = false; float closest = t_max;
    for (int i = 0; i < (int)s.objects.size(); ++i) {
        if (hit_object(s, i, r, t_min, closest, temp, rng)) { any = true; closest = temp.t; rec = temp; }
    }
    return any;
}

// ------------------------------------------------------------------------------------------------------------
// materials (material.h / material.cpp)
// ------------------------------------------------------------------------------------------------------------
inline bool near_zero(V3 e)                                                        // hobbyraytracer.h:34-38
{ const auto s = 1e-8; return (fabs(e.x) < s) && (fabs(e.y) < s) && (fabs(e.z) < s); }

V3 mat_emitted(const orc_scene& s, int mat, float u, float v, V3 p)
{
    if (mat >= MAT_MEDIUM_BASE) return v3(0);
    const hrt_material& m = s.materials[mat];
    if (m.kind != HRT_MAT_DIFFUSE_LIGHT) return v3(0);                             // material.h:67-70
    return matvec3(s, m.albedo, m.albedo_tex, u, v, p) * matscalar(s, m.p0, m.p0_tex, u, v, p);   // material.h:101-104
}

bool scatter_lambertian(const orc_scene& s, const hrt_material& m, const Rec& rec, V3& att, Ray& sc, Rng& rng)   // material.h:137-153
{
    V3 dir = rec.n + rng.spherical();
    if (near_zero(dir)) dir = rec.n;
    sc = Ray{rec.p, dir};
    att = matvec3(s, m.albedo, m.albedo_tex, rec.u, rec.v, rec.p);
    return true;
}
bool scatter_metal(const orc_scene& s, const hrt_material& m, const Ray& rin, const Rec& rec, V3& att, Ray& sc, Rng& rng)   // material.h:166-177
{
    V3 reflected = reflect(normalize(rin.d), normalize(rec.n));
    float rough = std::fabs(matscalar(s, m.p0, m.p0_tex, rec.u, rec.v, rec.p));   // glm::length(float) = abs
    rough = rough < 1 ? rough : 1;
    sc = Ray{rec.p, reflected + rough * rng.spherical() + v3(std::numeric_limits<float>::epsilon())};
    att = matvec3(s, m.albedo, m.albedo_tex, rec.u, rec.v, rec.p);
    return dot(sc.d, normalize(rec.n)) > 0;
}
double reflectance(double cosine, float refIdx)                                    // material.h:236-241
{
    double r0 = (1 - refIdx) / (1 + refIdx);
    r0 = r0 * r0;
    return r0 + (1 - r0) * std::pow((1 - cosine), 5);
}
bool scatter_dielectric(const orc_scene& s, const hrt_material& m, const Ray& rin, const Rec& rec, V3& att, Ray& sc, Rng& rng)   // material.h:204-228
{
    att = v3(1, 1, 1);
    float ratio = rec.front ? (1.0f / matscalar(s, m.p0, m.p0_tex, rec.u, rec.v, rec.p)) : matscalar(s, m.p0, m.p0_tex, rec.u, rec.v, rec.p);
    V3 unit = normalize(rin.d);
    double cosTheta = cmin(dot(-unit, rec.n), 1.0f);
    double sinTheta = std::sqrt(1.0 - cosTheta * cosTheta);
    bool cannot = ratio * sinTheta > 1.0;
    V3 direction;
    double ref = reflectance(cosTheta, ratio);
    if (cannot || ref > rng.lineard(0.0, 1.0)) direction = reflect(unit, rec.n);
    else direction = refract(unit, rec.n, ratio);
    sc = Ray{rec.p, direction + matscalar(s, m.p1, m.p1_tex, rec.u, rec.v, rec.p) * rng.spherical()};
    return true;
}

bool mat_scatter(const orc_scene& s, int mat, const Ray& rin, const Rec& rec, V3& att, Ray& sc, Rng& rng)
{
    if (mat >= MAT_MEDIUM_BASE) {                                                  // Isotropic of a ConstantMedium, material.h:79-85
        const Object& ob = s.objects[mat - MAT_MEDIUM_BASE];
        sc = Ray{rec.p, rng.ball()};
        att = ob.phase_tex >= 0 ? tex_value(s, ob.phase_tex, rec.u, rec.v, rec.p) : ob.phase_rgb;
        return true;
    }
    const hrt_material& m = s.materials[mat];
    switch (m.kind) {
    case HRT_MAT_LAMBERTIAN: return scatter_lambertian(s, m, rec, att, sc, rng);
    case HRT_MAT_METAL:      return scatter_metal(s, m, rin, rec, att, sc, rng);
    case HRT_MAT_DIELECTRIC: return scatter_dielectric(s, m, rin, rec, att, sc, rng);
    case HRT_MAT_ISOTROPIC:                                                        // material.h:79-85
        sc = Ray{rec.p, rng.ball()};
        att = matvec3(s, m.albedo, m.albedo_tex, rec.u, rec.v, rec.p);
        return true;
    case HRT_MAT_DIFFUSE_LIGHT: return false;                                      // material.h:96-99
    case HRT_MAT_PBR: {                                                            // material.cpp:18-28
        V3 mix = m.p1_tex >= 0 ? tex_value(s, m.p1_tex, rec.u, rec.v, rec.p) : v3(m.p1);
        if (length(mix) > 0.5f) return scatter_metal(s, m, rin, rec, att, sc, rng);
        return scatter_lambertian(s, m, rec, att, sc, rng);
    }
    case HRT_MAT_UVTEST: {                                                         // material.h:116-129
        V3 dir = rec.n + rng.spherical();
        if (near_zero(dir)) dir = rec.n;
        sc = Ray{rec.p, dir};
        att = rec.n;
        return true;
    }
    }
    return false;
}

// rayColour main.cpp:38-79
V3 ray_colour(const orc_scene& s, Ray r, int max_depth, float t_min, Rng& rng, uint64_t& n_rays)
{
    V3 atten = v3(1.0f), result = v3(0.0f);
    const float inf = std::numeric_limits<float>::infinity();
    for (int i = 0; i < max_depth; ++i) {
        Rec rec; ++n_rays;
        if (!hit_world(s, r, t_min, inf, rec, rng)) {
            V3 nD = normalize(r.d);
            float phi = ::atan2((double)nD.z, (double)nD.x);                         // main.cpp:51 (double atan2)
            float theta = ::acos((double)nD.y);                                      // main.cpp:52 (double acos)
            const float pi = 3.14159265358979323846264338327950288f;
            float u = phi / (2 * pi) + 0.5;                                        // main.cpp:55 (double add)
            float v = theta / pi;
            result = result + atten * tex_value(s, s.background, u, v, v3(0));
            break;
        }
        Ray sc{v3(0), v3(0)}; V3 att = v3(0);
        V3 emitted = mat_emitted(s, rec.mat, rec.u, rec.v, rec.p);
        bool b = mat_scatter(s, rec.mat, r, rec, att, sc, rng);
        if (!b) { result = result + atten * emitted; break; }
        result = result + atten * emitted;
        atten = atten * att;
        r = sc;
    }
    return result;
}

struct Cam { V3 origin, llc, horiz, vert, u, v; float lens_radius = 0.0f; };
Cam make_camera(const hrt_camera* c)                                               // camera.h:9-30
{
    V3 from = v3(c->position[0], c->position[1], c->position[2]), lookat = v3(c->look_at[0], c->look_at[1], c->look_at[2]);
    V3 up = v3(c->up[0], c->up[1], c->up[2]);
    float theta = radiansf(c->vfov_deg);
    float h = std::tan(theta / 2);
    float vh = 2.0f * h;
    float vw = c->aspect * vh;
    V3 w = normalize(from - lookat);
    V3 u = normalize(cross(up, w));
    V3 v = cross(w, u);
    Cam k;
    k.origin = from;
    k.horiz = c->focal_distance * vw * u;
    k.vert = c->focal_distance * vh * v;
    k.llc = k.origin - k.horiz / 2.0f - k.vert / 2.0f - c->focal_distance * w;
    k.u = u; k.v = v;                                                              // lens_radius stays 0: camera.h:34 zeroes rd
    return k;
}
inline Ray get_ray(const Cam& c, float s, float t)                                 // camera.h:32-39 (lens offset is zero)
{
    V3 off = v3(0.0f);                                                             // u*rd.x + v*rd.y with rd = 0
    return Ray{c.origin + off, c.llc + s * c.horiz + t * c.vert - c.origin - off};
}
// The variant camera.h:34 has commented out ("TODO: Add back in randomness"): rd = glm::circularRand(lensRadius)
// (random.inl:281-287), drawn when getRay is called, i.e. after the two jitter draws of main.cpp:120-121.  Only reached
// through orc_render_f32_defocus (the opt-in HRT_RENDER_DEFOCUS of the product); not reference behaviour.
template <class R> inline Ray get_ray_defocus(const Cam& c, float s, float t, R& rng)
{
    V3 off = v3(0.0f);
    if (c.lens_radius > 0.0f) {
        float a = rng.linear(0.0f, 6.283185307179586476925286766559f);
        float rx = std::cos(a) * c.lens_radius, ry = std::sin(a) * c.lens_radius;
        off = c.u * rx + c.v * ry;
    }
    return Ray{c.origin + off, c.llc + s * c.horiz + t * c.vert - c.origin - off};
}

// Film::tonemap + writeColour film.cpp:25-52
void tonemap_write(V3 c, uint8_t* px)
{
    if (c.x != c.x) c.x = 0.0f;
    if (c.y != c.y) c.y = 0.0f;
    if (c.z != c.z) c.z = 0.0f;
    const float a = 2.51f, b = 0.03f, cc = 2.43f, d = 0.59f, e = 0.14f;
    V3 num = c * (a * c + v3(b));
    V3 den = c * (cc * c + v3(d)) + v3(e);
    V3 q = num / den;
    q = v3(cmin(cmax(q.x, 0.0f), 1.0f), cmin(cmax(q.y, 0.0f), 1.0f), cmin(cmax(q.z, 0.0f), 1.0f));
    q = v3(std::sqrt(q.x), std::sqrt(q.y), std::sqrt(q.z));
    px[0] = static_cast<uint8_t>(256 * cmin(cmax(q.x, 0.0f), 0.9999f));
    px[1] = static_cast<uint8_t>(256 * cmin(cmax(q.y, 0.0f), 0.9999f));
    px[2] = static_cast<uint8_t>(256 * cmin(cmax(q.z, 0.0f), 0.9999f));
}

// ------------------------------------------------------------------------------------------------------------
// BVHNode::BVHNode bvh.cpp:6-61 in index form.  The split axis is glm::linearRand(0,2) on the current rand stream;
// ranges > 2 are std::sort'ed (same libstdc++ introsort as the reference, so equal keys permute identically).
// ------------------------------------------------------------------------------------------------------------
struct BvhBuilder {
    Shape& sh; Rng& rng; std::vector<int> order;
    bool stable = false;      // bvh_mode >= 2 (HRT_BVH_LBVH_*): equal keys keep their order - the algorithm-independent rule of the GPU builder
    void tri_box(int64_t i, V3& mn, V3& mx) { mn = sh.tri_mn[i]; mx = sh.tri_mx[i]; }
    void child_box(int c, V3& mn, V3& mx) { if (c >= 0) { mn = sh.nodes[c].mn; mx = sh.nodes[c].mx; } else tri_box(~c, mn, mx); }
    int build(size_t start, size_t end)
    {
        int axis = rng.lineari(0, 2);
        size_t n = end - start;
        int self = (int)sh.nodes.size(); sh.nodes.push_back(Shape::Node{});
        int left, right;
        auto key = [&](int tri) { return comp(sh.tri_mn[tri], axis); };
        if (n == 1) { left = right = ~order[start]; }
        else if (n == 2) {
            if (key(order[start]) < key(order[start + 1])) { left = ~order[start]; right = ~order[start + 1]; }
            else { left = ~order[start + 1]; right = ~order[start]; }
        } else {
            if (stable) std::stable_sort(order.begin() + start, order.begin() + end, [&](int a, int b) { return key(a) < key(b); });
            else std::sort(order.begin() + start, order.begin() + end, [&](int a, int b) { return key(a) < key(b); });
            size_t mid = start + n / 2;
            left = build(start, mid);
            right = build(mid, end);
        }
        V3 lmn, lmx, rmn, rmx; child_box(left, lmn, lmx); child_box(right, rmn, rmx);
        Shape::Node& nd = sh.nodes[self];
        nd.left = left; nd.right = right;
        nd.mn = v3(cmin(lmn.x, rmn.x), cmin(lmn.y, rmn.y), cmin(lmn.z, rmn.z));   // aabb.h:41-56
        nd.mx = v3(cmax(lmx.x, rmx.x), cmax(lmx.y, rmx.y), cmax(lmx.z, rmx.z));
        return self;
    }
};

void build_mesh_bvh(Shape& sh, uint64_t seed, int shape_id, bool stable)
{
    sh.nodes.clear(); sh.root = -1;
    sh.tri_mn.resize(sh.n_tri); sh.tri_mx.resize(sh.n_tri);
    for (int64_t i = 0; i < sh.n_tri; ++i) {                                       // triangle.cpp:133-151
        const float* P = &sh.pos[i * 9];
        float mnx = cmin(cmin(P[0], P[3]), P[6]), mny = cmin(cmin(P[1], P[4]), P[7]), mnz = cmin(cmin(P[2], P[5]), P[8]);
        float mxx = cmax(cmax(P[0], P[3]), P[6]), mxy = cmax(cmax(P[1], P[4]), P[7]), mxz = cmax(cmax(P[2], P[5]), P[8]);
        sh.tri_mn[i] = v3(mnx - 0.0001f, mny - 0.0001f, mnz - 0.0001f);
        sh.tri_mx[i] = v3(mxx + 0.0001f, mxy + 0.0001f, mxz + 0.0001f);
    }
    if (sh.n_tri == 0) return;
    Rng rng; rng.seed(seed, (uint64_t)shape_id);
    BvhBuilder b{sh, rng, {}, stable};
    b.order.resize(sh.n_tri); for (int64_t i = 0; i < sh.n_tri; ++i) b.order[i] = (int)i;
    sh.nodes.reserve(sh.n_tri * 2);
    sh.root = b.build(0, (size_t)sh.n_tri);
}

Xf make_xf(const hrt_xform& x)
{
    Xf r{}; r.kind = x.kind; r.v = v3(x.v[0], x.v[1], x.v[2]);
    switch (x.kind) {
    case HRT_XF_ROTATE_QUAT: r.q = Quat{x.v[0], x.v[1], x.v[2], x.v[3]}; break;
    case HRT_XF_ROTATE_EULER_DEG:                                                 // scene.cpp:340
        r.kind = HRT_XF_ROTATE_QUAT; r.q = quat_from_euler(v3(radiansf(x.v[0]), radiansf(x.v[1]), radiansf(x.v[2]))); break;
    case HRT_XF_ROTATE_Y: r.s = x.v[0]; r.c = x.v[1]; break;
    case HRT_XF_ROTATE_Y_DEG: { r.kind = HRT_XF_ROTATE_Y; float rad = radiansf(x.v[0]); r.s = std::sin(rad); r.c = std::cos(rad); } break;   // rotateY.cpp:7-9
    }
    return r;
}

int push_shape(orc_scene* s, Shape&& d, int64_t count)
{ d.prim_base = s->n_prims; d.prim_count = count; s->n_prims += count; s->shapes.push_back(std::move(d)); return (int)s->shapes.size() - 1; }

template <class F> void parallel_rows(int h, int n_threads, F&& fn)
{
    if (n_threads <= 0) n_threads = (int)std::thread::hardware_concurrency();
    if (n_threads < 1) n_threads = 1;
    std::atomic<int> next(0);
    auto worker = [&]() { for (;;) { int row = next.fetch_add(1); if (row >= h) break; fn(row); } };
    if (n_threads == 1) { worker(); return; }
    std::vector<std::thread> th;
    for (int i = 0; i < n_threads; ++i) th.emplace_back(worker);
    for (auto& t : th) t.join();
}

} // namespace

extern "C" {

const char* orc_last_error(void) { return g_err.c_str(); }
orc_scene* orc_scene_create(void) { return new orc_scene(); }
void orc_scene_destroy(orc_scene* s) { delete s; }

int orc_scene_add_texture_solid(orc_scene* s, const float rgb[3])
{ Texture t{}; t.kind = 0; t.rgb = v3(rgb[0], rgb[1], rgb[2]); s->textures.push_back(t); return (int)s->textures.size() - 1; }
int orc_scene_add_texture_checker(orc_scene* s, int even_tex, int odd_tex)
{ Texture t{}; t.kind = 1; t.even = even_tex; t.odd = odd_tex; s->textures.push_back(t); return (int)s->textures.size() - 1; }
int orc_scene_add_texture_image(orc_scene* s, const uint8_t* rgb8, int w, int h)
{ Texture t{}; t.kind = 2; t.w = w; t.h = h; t.ch = 3; if (rgb8 && w > 0 && h > 0) t.u8.assign(rgb8, rgb8 + (size_t)w * h * 3); s->textures.push_back(std::move(t)); return (int)s->textures.size() - 1; }
int orc_scene_add_texture_env(orc_scene* s, const float* data, int w, int h, int channels)
{ Texture t{}; t.kind = 3; t.w = w; t.h = h; t.ch = channels; if (data && w > 0 && h > 0) t.f32.assign(data, data + (size_t)w * h * channels); s->textures.push_back(std::move(t)); return (int)s->textures.size() - 1; }
int orc_scene_add_material(orc_scene* s, const hrt_material* m) { s->materials.push_back(*m); return (int)s->materials.size() - 1; }

int orc_scene_add_sphere(orc_scene* s, const float c[3], float radius, int material)
{ Shape d{}; d.kind = 0; d.f[0] = c[0]; d.f[1] = c[1]; d.f[2] = c[2]; d.f[3] = radius; d.material = material; return push_shape(s, std::move(d), 1); }
int orc_scene_add_rect(orc_scene* s, int axis, float a0, float a1, float b0, float b1, float k, int material)
{ Shape d{}; d.kind = 1; d.axis = axis; d.f[0] = a0; d.f[1] = a1; d.f[2] = b0; d.f[3] = b1; d.f[4] = k; d.material = material; return push_shape(s, std::move(d), 1); }
int orc_scene_add_box(orc_scene* s, const float c[3], const float dim[3], int material)
{ Shape d{}; d.kind = 2; for (int i = 0; i < 3; ++i) { d.f[i] = c[i]; d.f[3 + i] = dim[i]; } d.material = material; return push_shape(s, std::move(d), 6); }
int orc_scene_add_mesh(orc_scene* s, const float* pos, const float* nrm, const float* uv, int64_t n_tri, int material)
{
    Shape d{}; d.kind = 3; d.n_tri = n_tri; d.material = material;
    d.pos.assign(pos, pos + n_tri * 9);
    if (nrm) d.nrm.assign(nrm, nrm + n_tri * 9); else d.nrm.assign((size_t)n_tri * 9, 0.0f);
    if (uv) d.uv.assign(uv, uv + n_tri * 6); else d.uv.assign((size_t)n_tri * 6, 0.0f);
    return push_shape(s, std::move(d), n_tri);
}
int64_t orc_scene_prim_base(const orc_scene* s, int shape) { return s->shapes[shape].prim_base; }
int64_t orc_scene_prim_count(const orc_scene* s) { return s->n_prims; }

int orc_scene_add_object(orc_scene* s, int shape, const hrt_xform* ops, int n_ops)
{ Object o; o.shape = shape; for (int i = 0; i < n_ops; ++i) o.ops.push_back(make_xf(ops[i])); s->objects.push_back(std::move(o)); return (int)s->objects.size() - 1; }
int orc_scene_add_medium(orc_scene* s, int shape, const hrt_xform* ops, int n_ops, float density, const float rgb[3], int albedo_tex)
{
    Object o; o.shape = shape; for (int i = 0; i < n_ops; ++i) o.ops.push_back(make_xf(ops[i]));
    o.medium = true; o.neg_inv_density = -1 / density;                              // constantMedium.h:11
    o.phase_tex = albedo_tex; o.phase_rgb = rgb ? v3(rgb[0], rgb[1], rgb[2]) : v3(0);
    Shape d{}; d.kind = 4; d.material = -1; o.pseudo_shape = push_shape(s, std::move(d), 1);
    s->objects.push_back(std::move(o)); return s->objects.back().pseudo_shape;
}
int orc_scene_set_background(orc_scene* s, int tex) { s->background = tex; return 0; }

int orc_scene_commit(orc_scene* s, int bvh_mode, uint64_t seed)
{
    // meshes are built in world-list order, like the reference constructs them while loading (scene.cpp:292-297)
    std::vector<char> done(s->shapes.size(), 0);
    for (Object& o : s->objects) {
        Shape& sh = s->shapes[o.shape];
        if (sh.kind == 3 && !done[o.shape]) { build_mesh_bvh(sh, seed, o.shape, bvh_mode >= 2); done[o.shape] = 1; }
    }
    s->committed = true;
    return 0;
}

int orc_trace_batch(orc_scene* s, const float* rays, int64_t n, uint32_t* prim_id, float* t,
                    float* normal, float* uv, uint8_t* front, uint64_t rng_seed)
{
    if (!s->committed) { g_err = "scene not committed"; return HRT_ERR_STATE; }
    for (int64_t i = 0; i < n; ++i) {
        const float* r = rays + i * 8;
        Rng rng; rng.seed(rng_seed, (uint64_t)i);
        Rec rec; rec.prim = HRT_MISS;
        bool h = hit_world(*s, Ray{v3(r[0], r[1], r[2]), v3(r[4], r[5], r[6])}, r[3], r[7], rec, rng);
        prim_id[i] = h ? rec.prim : HRT_MISS; t[i] = h ? rec.t : 0.0f;
        if (normal) { normal[i*3] = h ? rec.n.x : 0; normal[i*3+1] = h ? rec.n.y : 0; normal[i*3+2] = h ? rec.n.z : 0; }
        if (uv) { uv[i*2] = h ? rec.u : 0; uv[i*2+1] = h ? rec.v : 0; }
        if (front) front[i] = h ? (rec.front ? 1 : 0) : 0;
    }
    return 0;
}

// render() main.cpp:81-140 on one rand stream (seed, 0), pixels in index order: 8-bit output.
int orc_render_u8(orc_scene* s, const hrt_camera* cam, int w, int h, int spp, uint64_t seed, uint8_t* out_rgb8)
{
    if (!s->committed) { g_err = "scene not committed"; return HRT_ERR_STATE; }
    Cam c = make_camera(cam);
    Rng rng; rng.seed(seed, 0); uint64_t nr = 0;
    for (int pIdx = 0; pIdx < w * h; ++pIdx) {
        V3 px = v3(0.0f);
        int x = pIdx % w; int y = h - pIdx / w;                                     // main.cpp:115-116
        for (int sidx = 0; sidx < spp; ++sidx) {
            float u = ((float)x + rng.linear(0.0f, 1.0f)) / (w - 1);
            float v = ((float)y + rng.linear(0.0f, 1.0f)) / (h - 1);
            px = px + ray_colour(*s, get_ray(c, u, v), 50, 0.001f, rng, nr);
        }
        px = px / static_cast<float>(spp);
        tonemap_write(px, out_rgb8 + (size_t)pIdx * 3);
    }
    return 0;
}

// Per-row streams (seed, row+1): per-pixel sums, any thread count gives the same result.
int orc_render_f32(orc_scene* s, const hrt_camera* cam, int w, int h, int spp, uint64_t seed, int n_threads,
                   float* out_sum, uint64_t* rays_out)
{
    if (!s->committed) { g_err = "scene not committed"; return HRT_ERR_STATE; }
    Cam c = make_camera(cam);
    std::atomic<unsigned long long> rays(0);
    parallel_rows(h, n_threads, [&](int row) {
        Rng rng; rng.seed(seed, (uint64_t)row + 1); uint64_t nr = 0;
        for (int xx = 0; xx < w; ++xx) {
            int pIdx = row * w + xx; int x = pIdx % w; int y = h - pIdx / w;
            V3 px = v3(0.0f);
            for (int sidx = 0; sidx < spp; ++sidx) {
                float u = ((float)x + rng.linear(0.0f, 1.0f)) / (w - 1);
                float v = ((float)y + rng.linear(0.0f, 1.0f)) / (h - 1);
                px = px + ray_colour(*s, get_ray(c, u, v), 50, 0.001f, rng, nr);
            }
            out_sum[pIdx * 3] = px.x; out_sum[pIdx * 3 + 1] = px.y; out_sum[pIdx * 3 + 2] = px.z;
        }
        rays += nr;
    });
    if (rays_out) *rays_out = rays.load();
    return 0;
}

// orc_render_f32 with the lens sampling of get_ray_defocus (lensRadius = aperture / 2, camera.h:29)
int orc_render_f32_defocus(orc_scene* s, const hrt_camera* cam, int w, int h, int spp, uint64_t seed, int n_threads,
                           float* out_sum, uint64_t* rays_out)
{
    if (!s->committed) { g_err = "scene not committed"; return HRT_ERR_STATE; }
    Cam c = make_camera(cam);
    if (cam->aperture > 0.0f) c.lens_radius = cam->aperture / 2.0f;
    std::atomic<unsigned long long> rays(0);
    parallel_rows(h, n_threads, [&](int row) {
        Rng rng; rng.seed(seed, (uint64_t)row + 1); uint64_t nr = 0;
        for (int xx = 0; xx < w; ++xx) {
            int pIdx = row * w + xx; int x = pIdx % w; int y = h - pIdx / w;
            V3 px = v3(0.0f);
            for (int sidx = 0; sidx < spp; ++sidx) {
                float u = ((float)x + rng.linear(0.0f, 1.0f)) / (w - 1);
                float v = ((float)y + rng.linear(0.0f, 1.0f)) / (h - 1);
                px = px + ray_colour(*s, get_ray_defocus(c, u, v, rng), 50, 0.001f, rng, nr);
            }
            out_sum[pIdx * 3] = px.x; out_sum[pIdx * 3 + 1] = px.y; out_sum[pIdx * 3 + 2] = px.z;
        }
        rays += nr;
    });
    if (rays_out) *rays_out = rays.load();
    return 0;
}

int orc_film_resolve(const float* sum, int w, int h, int spp, uint8_t* out_rgb8)
{
    for (int i = 0; i < w * h; ++i) {
        V3 c = v3(sum[i * 3], sum[i * 3 + 1], sum[i * 3 + 2]);
        c = c / static_cast<float>(spp);
        tonemap_write(c, out_rgb8 + (size_t)i * 3);
    }
    return 0;
}

int orc_camera_rays(const hrt_camera* cam, const float* st, int64_t n, float* out)
{
    Cam c = make_camera(cam);
    for (int64_t i = 0; i < n; ++i) {
        Ray r = get_ray(c, st[i * 2], st[i * 2 + 1]);
        out[i*6] = r.o.x; out[i*6+1] = r.o.y; out[i*6+2] = r.o.z; out[i*6+3] = r.d.x; out[i*6+4] = r.d.y; out[i*6+5] = r.d.z;
    }
    return 0;
}

int orc_rng_draw(int kind, uint64_t seed, uint64_t stream, int64_t n, double* out)
{
    Rng rng; rng.seed(seed, stream);
    for (int64_t i = 0; i < n; ++i) {
        switch (kind) {
        case 0: out[i] = rng.linear(0.0f, 1.0f); break;
        case 1: out[i] = rng.lineari(0, 2); break;
        case 2: { V3 v = rng.spherical(); out[i*3] = v.x; out[i*3+1] = v.y; out[i*3+2] = v.z; } break;
        case 3: { V3 v = rng.ball(); out[i*3] = v.x; out[i*3+1] = v.y; out[i*3+2] = v.z; } break;
        case 4: out[i] = rng.lineard(0.0, 1.0); break;
        default: return HRT_ERR_ARG;
        }
    }
    return 0;
}

int orc_scatter_batch(orc_scene* s, int material, const float* in, const uint8_t* front, int64_t n, uint64_t seed, float* out)
{
    for (int64_t i = 0; i < n; ++i) {
        const float* a = in + i * 14; Rng rng; rng.seed(seed, (uint64_t)i);
        Ray rin{v3(a[0], a[1], a[2]), v3(a[3], a[4], a[5])};
        Rec rec; rec.p = v3(a[6], a[7], a[8]); rec.n = v3(a[9], a[10], a[11]); rec.u = a[12]; rec.v = a[13];
        rec.front = front[i] != 0; rec.t = 1.0f; rec.mat = material; rec.prim = 0;
        V3 att = v3(0.0f); Ray sc{v3(0.0f), v3(0.0f)};
        V3 em = mat_emitted(*s, material, rec.u, rec.v, rec.p);
        bool ok = mat_scatter(*s, material, rin, rec, att, sc, rng);
        float* o = out + i * 10;
        o[0] = ok ? 1.0f : 0.0f; o[1] = att.x; o[2] = att.y; o[3] = att.z; o[4] = sc.d.x; o[5] = sc.d.y; o[6] = sc.d.z; o[7] = em.x; o[8] = em.y; o[9] = em.z;
    }
    return 0;
}

int orc_texture_eval(orc_scene* s, int tex, const float* in, int64_t n, float* out)
{
    for (int64_t i = 0; i < n; ++i) {
        V3 c = tex_value(*s, tex, in[i*5], in[i*5+1], v3(in[i*5+2], in[i*5+3], in[i*5+4]));
        out[i*3] = c.x; out[i*3+1] = c.y; out[i*3+2] = c.z;
    }
    return 0;
}

// BVH introspection for the topology-parity test of the product's host builder: nodes as (mn, mx, left, right)
int64_t orc_mesh_bvh_nodes(orc_scene* s, int shape, float* boxes /* n*6 or NULL */, int32_t* children /* n*2 or NULL */)
{
    const Shape& sh = s->shapes[shape];
    if (boxes && children)
        for (size_t i = 0; i < sh.nodes.size(); ++i) {
            const Shape::Node& n = sh.nodes[i];
            boxes[i*6] = n.mn.x; boxes[i*6+1] = n.mn.y; boxes[i*6+2] = n.mn.z; boxes[i*6+3] = n.mx.x; boxes[i*6+4] = n.mx.y; boxes[i*6+5] = n.mx.z;
            children[i*2] = n.left; children[i*2+1] = n.right;
        }
    return (int64_t)sh.nodes.size();
}

int orc_hardware_threads(void) { return (int)std::thread::hardware_concurrency(); }

} // extern "C"
