// hrt_host.cpp — host flattener and BVH builders of the B200 path-tracing core (no CUDA in this file).
//
// What it replaces in the reference: the object graph that Scene::loadScene (scene.cpp:127-374) and Mesh::Mesh
// (mesh.cpp:13-41) build out of shared_ptr<Hittable> nodes.  Here each constructor call appends to flat arrays,
// and commit turns them into GPU records.  Load-time arithmetic that the reference does on the host (quaternion
// from Euler angles scene.cpp:340, RotateY's sin/cos rotateY.cpp:7-9, Box::minMax box.h:41-44, triangle bounding
// boxes triangle.cpp:133-151, the camera basis camera.h:9-30) is done here on the host too, in fp32, in the
// reference's evaluation order (build with -ffp-contract=off).
#include "hrt_host.h"
#include "hrt_bvh_rules.h"

#include <algorithm>
#include <array>
#include <chrono>
#include <cmath>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <map>
#include <mutex>
#include <thread>

namespace hrt {

namespace {

struct F3 { float x, y, z; };
inline F3 f3(float a, float b, float c) { return F3{a, b, c}; }
inline F3 operator+(F3 a, F3 b) { return F3{a.x + b.x, a.y + b.y, a.z + b.z}; }
inline F3 operator-(F3 a, F3 b) { return F3{a.x - b.x, a.y - b.y, a.z - b.z}; }
inline F3 operator*(F3 a, float s) { return F3{a.x * s, a.y * s, a.z * s}; }
inline F3 operator*(float s, F3 a) { return F3{s * a.x, s * a.y, s * a.z}; }
inline F3 operator/(F3 a, float s) { return F3{a.x / s, a.y / s, a.z / s}; }
inline float dot3(F3 a, F3 b) { float x = a.x * b.x, y = a.y * b.y, z = a.z * b.z; return x + y + z; }       // func_geometric.inl:48-55
inline F3 cross3(F3 x, F3 y) { return F3{x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y}; }
inline F3 normalize3(F3 a) { return a * (1.0f / sqrtf(dot3(a, a))); }                                      // func_geometric.inl:82-90
inline float radiansf(float deg) { return deg * 0.01745329251994329576923690768489f; }                    // glm::radians
inline float cmin(float a, float b) { return (b < a) ? b : a; }                                           // std::min
inline float cmax(float a, float b) { return (a < b) ? b : a; }                                           // std::max

double now_ms()
{
    using namespace std::chrono;
    return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}

// ---- the rand() stream behind glm::linearRand (random.inl:19-27, 140-147) --------------------------------
// 31-bit LCG shared with the test oracle so that seeded split axes agree; jump() lets sub-trees be built in
// parallel with exactly the draws a sequential pre-order build would hand them.
constexpr uint64_t LCG_A = 6364136223846793005ULL, LCG_C = 1442695040888963407ULL;
inline uint64_t mix64(uint64_t z)
{
    z += 0x9e3779b97f4a7c15ULL;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
    return z ^ (z >> 31);
}
struct Lcg {
    uint64_t s;
    int rand31() { s = s * LCG_A + LCG_C; return (int)((s >> 33) & 0x7fffffffULL); }
    uint32_t byte() { return (uint32_t)(rand31() % 255); }
    uint32_t u32() { uint32_t b0 = byte(), b1 = byte(), b2 = byte(), b3 = byte(); return b0 | (b1 << 8) | (b2 << 16) | (b3 << 24); }
    int lineari(int mn, int mx) { return (int)(u32() % (uint32_t)(mx + 1 - mn)) + mn; }
    void jump(uint64_t k)
    {
        uint64_t a = LCG_A, c = LCG_C, A = 1, C = 0;
        while (k) {
            if (k & 1) { A = A * a; C = C * a + c; }
            c = (a + 1) * c; a = a * a; k >>= 1;
        }
        s = A * s + C;
    }
};

} // namespace

DXform make_xform(const hrt_xform& x)
{
    DXform r{}; r.kind = x.kind; r.a = x.v[0]; r.b = x.v[1]; r.c = x.v[2]; r.d = x.v[3];
    switch (x.kind) {
    case HRT_XF_ROTATE_EULER_DEG: {                          // glm::quat(glm::radians(angles)) type_quat.inl:208-217
        F3 h = f3(radiansf(x.v[0]), radiansf(x.v[1]), radiansf(x.v[2])) * 0.5f;
        F3 c = f3(std::cos(h.x), std::cos(h.y), std::cos(h.z));
        F3 s = f3(std::sin(h.x), std::sin(h.y), std::sin(h.z));
        r.kind = HRT_XF_ROTATE_QUAT;
        r.a = c.x * c.y * c.z + s.x * s.y * s.z;
        r.b = s.x * c.y * c.z - c.x * s.y * s.z;
        r.c = c.x * s.y * c.z + s.x * c.y * s.z;
        r.d = c.x * c.y * s.z - s.x * s.y * c.z;
    } break;
    case HRT_XF_ROTATE_Y_DEG: {                              // rotateY.cpp:7-9
        float rad = radiansf(x.v[0]);
        r.kind = HRT_XF_ROTATE_Y; r.a = std::sin(rad); r.b = std::cos(rad); r.c = r.d = 0;
    } break;
    default: break;
    }
    return r;
}

DCamera make_camera(const hrt_camera& c)                      // camera.h:9-30
{
    F3 from = f3(c.position[0], c.position[1], c.position[2]), at = f3(c.look_at[0], c.look_at[1], c.look_at[2]);
    F3 up = f3(c.up[0], c.up[1], c.up[2]);
    float theta = radiansf(c.vfov_deg);
    float h = std::tan(theta / 2);
    float vh = 2.0f * h;
    float vw = c.aspect * vh;
    F3 w = normalize3(from - at);
    F3 u = normalize3(cross3(up, w));
    F3 v = cross3(w, u);
    F3 horiz = c.focal_distance * vw * u;
    F3 vert = c.focal_distance * vh * v;
    F3 llc = from - horiz / 2.0f - vert / 2.0f - c.focal_distance * w;
    DCamera k;
    k.origin[0] = from.x; k.origin[1] = from.y; k.origin[2] = from.z;
    k.llc[0] = llc.x; k.llc[1] = llc.y; k.llc[2] = llc.z;
    k.horiz[0] = horiz.x; k.horiz[1] = horiz.y; k.horiz[2] = horiz.z;
    k.vert[0] = vert.x; k.vert[1] = vert.y; k.vert[2] = vert.z;
    k.u[0] = u.x; k.u[1] = u.y; k.u[2] = u.z; k.v[0] = v.x; k.v[1] = v.y; k.v[2] = v.z;
    k.lens_radius = 0.0f;                                       // the caller sets aperture / 2 under HRT_RENDER_DEFOCUS
    return k;
}

void triangle_boxes(const HostShape& sh, std::vector<float>& bmin, std::vector<float>& bmax)   // triangle.cpp:133-151
{
    bmin.resize((size_t)sh.n_tri * 3); bmax.resize((size_t)sh.n_tri * 3);
    for (int64_t i = 0; i < sh.n_tri; ++i) {
        const float* P = &sh.pos[(size_t)i * 9];
        for (int a = 0; a < 3; ++a) {
            float mn = cmin(cmin(P[a], P[3 + a]), P[6 + a]);
            float mx = cmax(cmax(P[a], P[3 + a]), P[6 + a]);
            bmin[(size_t)i * 3 + a] = mn - 0.0001f;
            bmax[(size_t)i * 3 + a] = mx + 0.0001f;
        }
    }
}

// ------------------------------------------------------------------------------------------------------------
// Reference-topology BVH (bvh.cpp:6-61).  Node ids are pre-order positions, so they and the rand() stream
// position of every sub-tree are known from the primitive count alone (nodes(n) = 1 for n <= 2, else
// 1 + nodes(n/2) + nodes(n - n/2)); sub-trees are therefore built by independent threads and still come out
// identical to the sequential recursion.  Ranges > 2 are ordered with std::sort like bvh.cpp:39.
// ------------------------------------------------------------------------------------------------------------
namespace {
struct RefBuilder {
    const std::vector<float>& bmin; const std::vector<float>& bmax;
    std::vector<float4>& nodes; std::vector<int>& order;
    std::map<size_t, uint64_t> count;
    bool stable = false;
    uint64_t n_nodes(size_t n)
    {
        if (n <= 2) return 1;
        auto it = count.find(n);
        if (it != count.end()) return it->second;
        uint64_t v = 1 + n_nodes(n / 2) + n_nodes(n - n / 2);
        count[n] = v; return v;
    }
    uint64_t n_nodes_ro(size_t n) const { return n <= 2 ? 1 : count.find(n)->second; }
    void child_box(int c, float* mn, float* mx) const
    {
        if (c >= 0) { const float4& a = nodes[(size_t)c * 2]; const float4& b = nodes[(size_t)c * 2 + 1]; mn[0] = a.x; mn[1] = a.y; mn[2] = a.z; mx[0] = b.x; mx[1] = b.y; mx[2] = b.z; }
        else { size_t t = (size_t)(~c); for (int a = 0; a < 3; ++a) { mn[a] = bmin[t * 3 + a]; mx[a] = bmax[t * 3 + a]; } }
    }
    void build(size_t start, size_t end, int self, Lcg rng, int par_depth)
    {
        int axis = rng.lineari(0, 2);
        size_t n = end - start;
        int left, right;
        auto key = [&](int tri) { return bmin[(size_t)tri * 3 + axis]; };
        if (n == 1) { left = right = ~order[start]; }
        else if (n == 2) {
            if (key(order[start]) < key(order[start + 1])) { left = ~order[start]; right = ~order[start + 1]; }
            else { left = ~order[start + 1]; right = ~order[start]; }
        } else {
            if (stable) std::stable_sort(order.begin() + start, order.begin() + end, [&](int a, int b) { return key(a) < key(b); });
            else std::sort(order.begin() + start, order.begin() + end, [&](int a, int b) { return key(a) < key(b); });
            size_t mid = start + n / 2;
            uint64_t nl = n_nodes_ro(mid - start);
            left = self + 1; right = self + 1 + (int)nl;
            Lcg rl = rng, rr = rng; rr.jump(4 * nl);
            if (par_depth > 0 && n > 32768) {
                std::thread t([&, rl]() { build(start, mid, left, rl, par_depth - 1); });
                build(mid, end, right, rr, par_depth - 1);
                t.join();
            } else {
                build(start, mid, left, rl, 0);
                build(mid, end, right, rr, 0);
            }
        }
        float lmn[3], lmx[3], rmn[3], rmx[3];
        child_box(left, lmn, lmx); child_box(right, rmn, rmx);
        float4 a, b;                                                     // surroundingBox aabb.h:41-56
        a.x = cmin(lmn[0], rmn[0]); a.y = cmin(lmn[1], rmn[1]); a.z = cmin(lmn[2], rmn[2]);
        b.x = cmax(lmx[0], rmx[0]); b.y = cmax(lmx[1], rmx[1]); b.z = cmax(lmx[2], rmx[2]);
        std::memcpy(&a.w, &left, 4); std::memcpy(&b.w, &right, 4);
        nodes[(size_t)self * 2] = a; nodes[(size_t)self * 2 + 1] = b;
    }
};
} // namespace

void build_ref_bvh(HostShape& sh, uint64_t seed, int shape_id, int n_threads, bool stable)
{
    sh.ref.nodes.clear(); sh.ref.root = -1;
    if (sh.n_tri == 0) return;
    std::vector<float> bmin, bmax; triangle_boxes(sh, bmin, bmax);
    std::vector<int> order((size_t)sh.n_tri);
    for (int64_t i = 0; i < sh.n_tri; ++i) order[(size_t)i] = (int)i;
    RefBuilder b{bmin, bmax, sh.ref.nodes, order, {}, stable};
    uint64_t total = b.n_nodes((size_t)sh.n_tri);
    sh.ref.nodes.resize((size_t)total * 2);
    Lcg rng; rng.s = mix64(seed ^ mix64((uint64_t)shape_id));
    int par = 0; while ((1 << par) < n_threads) ++par;
    b.build(0, (size_t)sh.n_tri, 0, rng, par);
    sh.ref.root = 0;
}

// ------------------------------------------------------------------------------------------------------------
// Binned-SAH BVH2 over arbitrary boxes, one item per leaf.  Node ids are pre-order positions (a sub-tree over m
// items owns exactly m - 1 consecutive nodes), so the layout is deterministic and sub-trees build in parallel.
// Each node stores the boxes of its two children (hrt_types.h, fast_nodes).
// ------------------------------------------------------------------------------------------------------------
namespace {
struct Box3 { float mn[3], mx[3]; };
inline Box3 empty_box()
{
    Box3 b; for (int a = 0; a < 3; ++a) { b.mn[a] = std::numeric_limits<float>::infinity(); b.mx[a] = -std::numeric_limits<float>::infinity(); }
    return b;
}
inline void grow(Box3& b, const float* mn, const float* mx) { for (int a = 0; a < 3; ++a) { b.mn[a] = std::min(b.mn[a], mn[a]); b.mx[a] = std::max(b.mx[a], mx[a]); } }
inline void grow(Box3& b, const Box3& o) { grow(b, o.mn, o.mx); }
inline double half_area(const Box3& b)
{
    double dx = (double)b.mx[0] - b.mn[0], dy = (double)b.mx[1] - b.mn[1], dz = (double)b.mx[2] - b.mn[2];
    if (!(dx >= 0) || !(dy >= 0) || !(dz >= 0)) return 0.0;
    return dx * dy + dy * dz + dz * dx;
}

struct FastBuilder {
    const std::vector<float>& bmin; const std::vector<float>& bmax;
    std::vector<float4>& nodes; std::vector<int>& idx;
    static constexpr int BINS = 16;
    static constexpr int SAH_DEPTH = 32;     // deeper than this: object-median splits (bounded depth, FAST_MAX_DEPTH)

    Box3 item_box(int i) const { Box3 b; for (int a = 0; a < 3; ++a) { b.mn[a] = bmin[(size_t)i * 3 + a]; b.mx[a] = bmax[(size_t)i * 3 + a]; } return b; }

    // returns the child reference (node id or ~item) and its box
    int build(size_t start, size_t end, int base, int depth, int par_depth, Box3& out_box)
    {
        size_t n = end - start;
        if (n == 1) { out_box = item_box(idx[start]); return ~idx[start]; }
        Box3 cb = empty_box();
        for (size_t i = start; i < end; ++i) {
            int it = idx[i]; float c[3];
            for (int a = 0; a < 3; ++a) c[a] = 0.5f * bmin[(size_t)it * 3 + a] + 0.5f * bmax[(size_t)it * 3 + a];
            grow(cb, c, c);
        }
        size_t mid = start;
        bool split_done = false;
        if (depth < SAH_DEPTH && n > 2) {
            double best_cost = std::numeric_limits<double>::infinity(); int best_axis = -1, best_bin = -1;
            for (int axis = 0; axis < 3; ++axis) {
                float lo = cb.mn[axis], hi = cb.mx[axis];
                if (!(hi > lo)) continue;
                float scale = (float)BINS / (hi - lo);
                Box3 bb[BINS]; size_t cnt[BINS];
                for (int b = 0; b < BINS; ++b) { bb[b] = empty_box(); cnt[b] = 0; }
                for (size_t i = start; i < end; ++i) {
                    int it = idx[i];
                    float c = 0.5f * bmin[(size_t)it * 3 + axis] + 0.5f * bmax[(size_t)it * 3 + axis];
                    int b = (int)((c - lo) * scale); if (b < 0) b = 0; if (b >= BINS) b = BINS - 1;
                    grow(bb[b], &bmin[(size_t)it * 3], &bmax[(size_t)it * 3]); ++cnt[b];
                }
                double right_area[BINS]; size_t right_cnt[BINS];
                Box3 acc = empty_box(); size_t c = 0;
                for (int b = BINS - 1; b > 0; --b) { grow(acc, bb[b]); c += cnt[b]; right_area[b] = half_area(acc); right_cnt[b] = c; }
                acc = empty_box(); c = 0;
                for (int b = 0; b < BINS - 1; ++b) {
                    grow(acc, bb[b]); c += cnt[b];
                    if (c == 0 || right_cnt[b + 1] == 0) continue;
                    double cost = half_area(acc) * (double)c + right_area[b + 1] * (double)right_cnt[b + 1];
                    if (cost < best_cost) { best_cost = cost; best_axis = axis; best_bin = b; }
                }
            }
            if (best_axis >= 0) {
                float lo = cb.mn[best_axis], hi = cb.mx[best_axis];
                float scale = (float)BINS / (hi - lo);
                auto it = std::partition(idx.begin() + start, idx.begin() + end, [&](int item) {
                    float c = 0.5f * bmin[(size_t)item * 3 + best_axis] + 0.5f * bmax[(size_t)item * 3 + best_axis];
                    int b = (int)((c - lo) * scale); if (b < 0) b = 0; if (b >= BINS) b = BINS - 1;
                    return b <= best_bin;
                });
                mid = (size_t)(it - idx.begin());
                split_done = mid > start && mid < end;
            }
        }
        if (!split_done) {
            int axis = 0; float ext = -1;
            for (int a = 0; a < 3; ++a) { float e = cb.mx[a] - cb.mn[a]; if (e > ext) { ext = e; axis = a; } }
            mid = start + n / 2;
            std::nth_element(idx.begin() + start, idx.begin() + mid, idx.begin() + end, [&](int x, int y) {
                float cx = 0.5f * bmin[(size_t)x * 3 + axis] + 0.5f * bmax[(size_t)x * 3 + axis];
                float cy = 0.5f * bmin[(size_t)y * 3 + axis] + 0.5f * bmax[(size_t)y * 3 + axis];
                return cx < cy || (cx == cy && x < y);
            });
        }
        size_t nl = mid - start;
        int self = base;
        int lbase = base + 1, rbase = base + 1 + (int)(nl - 1);
        Box3 lb, rb; int lc, rc;
        if (par_depth > 0 && n > 65536) {
            std::thread t([&]() { lc = build(start, mid, lbase, depth + 1, par_depth - 1, lb); });
            rc = build(mid, end, rbase, depth + 1, par_depth - 1, rb);
            t.join();
        } else {
            lc = build(start, mid, lbase, depth + 1, 0, lb);
            rc = build(mid, end, rbase, depth + 1, 0, rb);
        }
        float4 n0, n1, n2, n3;
        n0.x = lb.mn[0]; n0.y = lb.mx[0]; n0.z = lb.mn[1]; n0.w = lb.mx[1];
        n1.x = rb.mn[0]; n1.y = rb.mx[0]; n1.z = rb.mn[1]; n1.w = rb.mx[1];
        n2.x = lb.mn[2]; n2.y = lb.mx[2]; n2.z = rb.mn[2]; n2.w = rb.mx[2];
        std::memcpy(&n3.x, &lc, 4); std::memcpy(&n3.y, &rc, 4); n3.z = 0; n3.w = 0;
        nodes[(size_t)self * 4] = n0; nodes[(size_t)self * 4 + 1] = n1; nodes[(size_t)self * 4 + 2] = n2; nodes[(size_t)self * 4 + 3] = n3;
        out_box = lb; grow(out_box, rb);
        return self;
    }
};
} // namespace

void build_fast_bvh(const std::vector<float>& bmin, const std::vector<float>& bmax, FastBvh& out, int n_threads)
{
    size_t n = bmin.size() / 3;
    out.nodes.clear(); out.order.clear(); out.root = -1; out.depth = 0;
    if (n == 0) return;
    std::vector<int> idx(n);
    for (size_t i = 0; i < n; ++i) idx[i] = (int)i;
    out.nodes.resize((n - 1) * 4);
    FastBuilder b{bmin, bmax, out.nodes, idx};
    int par = 0; while ((1 << par) < n_threads) ++par;
    Box3 box;
    out.root = b.build(0, n, 0, 0, par, box);
}

// ------------------------------------------------------------------------------------------------------------
// Host mirror of the device LBVH (hrt_bvh_gpu.cu: k_centre_bounds, k_morton, sort, k_karras, k_refit) — same rules
// (hrt_bvh_rules.h), same node numbering (internal node i of the hierarchy is node i), boxes united child 0 first.
// ------------------------------------------------------------------------------------------------------------
namespace {
inline uint32_t f2u(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
inline float u2f(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }
inline float key_to_float(uint32_t k) { return u2f((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k); }
}

void build_lbvh_host(const std::vector<float>& bmin, const std::vector<float>& bmax, FastBvh& out)
{
    const int64_t n = (int64_t)(bmin.size() / 3);
    out.nodes.clear(); out.order.clear(); out.root = -1; out.depth = 0;
    if (n == 0) return;
    if (n == 1) { out.root = ~0; return; }
    uint32_t lo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, hi[3] = {0u, 0u, 0u};
    for (int64_t i = 0; i < n; ++i)
        for (int a = 0; a < 3; ++a) {
            uint32_t k = rule_float_key(f2u(rule_centre(bmin[(size_t)i * 3 + a], bmax[(size_t)i * 3 + a])));
            lo[a] = std::min(lo[a], k); hi[a] = std::max(hi[a], k);
        }
    std::vector<uint64_t> code((size_t)n);
    for (int64_t i = 0; i < n; ++i) {
        uint32_t cell[3];
        for (int a = 0; a < 3; ++a) {
            const float l = key_to_float(lo[a]), h = key_to_float(hi[a]);
            cell[a] = rule_cell(rule_centre(bmin[(size_t)i * 3 + a], bmax[(size_t)i * 3 + a]), l, rule_scale(l, h));
        }
        code[(size_t)i] = rule_morton(cell[0], cell[1], cell[2]);
    }
    std::vector<int> order((size_t)n);
    for (int64_t i = 0; i < n; ++i) order[(size_t)i] = (int)i;
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return code[(size_t)a] < code[(size_t)b]; });
    std::vector<uint64_t> sorted((size_t)n);
    for (int64_t i = 0; i < n; ++i) sorted[(size_t)i] = code[(size_t)order[(size_t)i]];
    out.nodes.assign((size_t)(n - 1) * 4, float4{0, 0, 0, 0});
    std::vector<int> lc((size_t)n - 1), rc((size_t)n - 1);
    for (int64_t i = 0; i < n - 1; ++i) {
        int64_t first, last, split;
        rule_karras(sorted.data(), n, i, first, last, split);
        lc[(size_t)i] = (split == first) ? ~order[(size_t)split] : (int)split;
        rc[(size_t)i] = (split + 1 == last) ? ~order[(size_t)split + 1] : (int)(split + 1);
    }
    // boxes bottom-up (explicit stack: the hierarchy can be ~100 levels deep on degenerate input)
    std::vector<Box3> nb((size_t)n - 1); std::vector<int> nd((size_t)n - 1, 0);
    std::vector<std::pair<int, int>> stack; stack.push_back({0, 0});
    auto child_box = [&](int c, Box3& b, int& d) {
        if (c < 0) { size_t it = (size_t)(~c); for (int a = 0; a < 3; ++a) { b.mn[a] = bmin[it * 3 + a]; b.mx[a] = bmax[it * 3 + a]; } d = 0; }
        else { b = nb[(size_t)c]; d = nd[(size_t)c]; }
    };
    while (!stack.empty()) {
        auto [node, stage] = stack.back();
        if (stage == 0) {
            stack.back().second = 1;
            if (rc[(size_t)node] >= 0) stack.push_back({rc[(size_t)node], 0});
            if (lc[(size_t)node] >= 0) stack.push_back({lc[(size_t)node], 0});
            continue;
        }
        stack.pop_back();
        Box3 lb, rb; int dl, dr;
        child_box(lc[(size_t)node], lb, dl); child_box(rc[(size_t)node], rb, dr);
        float4 n0, n1, n2, n3;
        n0.x = lb.mn[0]; n0.y = lb.mx[0]; n0.z = lb.mn[1]; n0.w = lb.mx[1];
        n1.x = rb.mn[0]; n1.y = rb.mx[0]; n1.z = rb.mn[1]; n1.w = rb.mx[1];
        n2.x = lb.mn[2]; n2.y = lb.mx[2]; n2.z = rb.mn[2]; n2.w = rb.mx[2];
        std::memcpy(&n3.x, &lc[(size_t)node], 4); std::memcpy(&n3.y, &rc[(size_t)node], 4); n3.z = 0; n3.w = 0;
        float4* N = &out.nodes[(size_t)node * 4]; N[0] = n0; N[1] = n1; N[2] = n2; N[3] = n3;
        Box3 u;
        for (int a = 0; a < 3; ++a) { u.mn[a] = cmin(lb.mn[a], rb.mn[a]); u.mx[a] = cmax(lb.mx[a], rb.mx[a]); }
        nb[(size_t)node] = u; nd[(size_t)node] = std::max(dl, dr) + 1;
    }
    out.root = 0; out.depth = nd[0]; out.order = std::move(order);
}

// ------------------------------------------------------------------------------------------------------------
// 8-wide collapse (hrt_bvh_rules.h (4)), breadth first: the inner children of the nodes of one level, in node order and
// slot order, are the nodes of the next level - the numbering the device builder reproduces with one scan per level.
// ------------------------------------------------------------------------------------------------------------
void build_wide(const FastBvh& fast, WideBvh& out)
{
    out.nodes.clear(); out.root = fast.root; out.depth = 0; out.ok = true;
    if (fast.root < 0) return;                                           // empty, or a single item: the root is that leaf
    std::vector<int> level{fast.root}, next;                             // two-child nodes that become the wide nodes of this level
    size_t base = 0;                                                     // wide index of level[0]
    std::atomic<bool> ok{true};
    while (!level.empty()) {
        ++out.depth;
        const size_t n = level.size();
        out.nodes.resize(base + n);
        std::vector<uint8_t> inner_count(n);
        std::vector<std::array<int, 8>> inner(n);
        parallel_chunks(n, 4096, [&](size_t, size_t lo, size_t hi) {
            for (size_t i = lo; i < hi; ++i) {
                WideItem w; rule_wide_gather(fast.nodes.data(), level[i], w);
                WNode& node = out.nodes[base + i];
                int slot_of[8];
                if (!rule_wide_pack(w, node, slot_of)) ok = false;
                int by_slot[8]; for (int s = 0; s < 8; ++s) by_slot[s] = 0x7fffffff;       // 0x7fffffff: empty
                for (int c = 0; c < w.count; ++c) by_slot[slot_of[c]] = w.child[c];
                int k = 0;
                for (int s = 0; s < 8; ++s) {
                    node.child[s] = by_slot[s] == 0x7fffffff ? -1 : by_slot[s];          // inner children are renumbered below
                    if (by_slot[s] != 0x7fffffff && by_slot[s] >= 0) inner[i][(size_t)k++] = s;
                }
                inner_count[i] = (uint8_t)k;
            }
        });
        const size_t next_base = base + n;
        size_t total = 0;
        std::vector<size_t> first(n);
        for (size_t i = 0; i < n; ++i) { first[i] = total; total += inner_count[i]; }
        next.resize(total);
        parallel_chunks(n, 4096, [&](size_t, size_t lo, size_t hi) {
            for (size_t i = lo; i < hi; ++i) {
                WNode& node = out.nodes[base + i];
                for (int k = 0; k < inner_count[i]; ++k) {
                    const int s = inner[i][(size_t)k];
                    next[first[i] + (size_t)k] = node.child[s];
                    node.child[s] = (int)(next_base + first[i] + (size_t)k);
                }
            }
        });
        base = next_base; level.swap(next);
    }
    out.root = 0; out.ok = ok && out.nodes.size() < (size_t(1) << 24);       // stack entries carry a node index in 24 bits (hrt_wide.cuh)
}

// ------------------------------------------------------------------------------------------------------------
// Exact repack of a decoded Radiance image (texture.cpp:99-115 -> stb_image stbi__hdr_convert: byte * 2^(e-136))
// into one 32-bit word per texel; refuses (returns false) unless every texel round-trips bit for bit.
// ------------------------------------------------------------------------------------------------------------
namespace {
std::mutex g_pool_mutex;
std::vector<uint32_t> g_pool;                      // one spare buffer (the most recently returned one)
}
std::vector<uint32_t> take_words(size_t n)
{
    std::vector<uint32_t> v;
    { std::lock_guard<std::mutex> lock(g_pool_mutex); v.swap(g_pool); }
    v.resize(n);                                   // same size as last time: nothing is filled, the pages are already mapped
    return v;
}
void give_words(std::vector<uint32_t>&& v)
{
    if (v.capacity() > (size_t(1) << 28)) return;   // keep at most 1 GiB around
    std::lock_guard<std::mutex> lock(g_pool_mutex);
    if (v.capacity() >= g_pool.capacity()) g_pool.swap(v);
}

bool try_pack_rgbe(const float* f32, int w, int h, int ch, std::vector<uint32_t>& rgbe)
{
    if (!f32 || ch < 3 || w <= 0 || h <= 0) return false;
    const size_t n = (size_t)w * h;
    rgbe = take_words(n);
    std::atomic<bool> ok{true};
    // Integer form of "p == q * 2^(e-136) with q an integer in [0,255]": with mexp the biased exponent of the largest
    // channel, e = mexp + 2 (= frexp exponent + 128) and a channel 1.mant * 2^(cexp-127) is q = 1.mant * 2^(cexp-mexp+7).
    // everything the loop reads is captured BY VALUE: with reference captures each iteration reloaded `ch` / `f32` from the
    // calling thread's stack (the uint32 stores may alias them), the line the caller's own slice keeps writing its
    // temporaries to - 16 threads ran no faster than one (20 ms for 4096x2048; 2 ms now)
    uint32_t* const out = rgbe.data();
    std::atomic<bool>* const okp = &ok;
    auto work = [f32, ch, out, okp](size_t i0, size_t i1) {
        std::atomic<bool>& ok = *okp;
        uint32_t* const rgbe = out;
        for (size_t i = i0; i < i1; ++i) {
            if ((i & 4095u) == 0 && !ok.load(std::memory_order_relaxed)) return;
            uint32_t b[3]; std::memcpy(b, f32 + i * (size_t)ch, 12);
            const uint32_t m = std::max(b[0], std::max(b[1], b[2]));
            if ((b[0] | b[1] | b[2]) & 0x80000000u) { ok = false; return; }     // negative (or -0): keep the float array
            if (m == 0) { rgbe[i] = 0; continue; }
            const int mexp = (int)(m >> 23);
            if (mexp < 8 || mexp > 253) { ok = false; return; }                 // denormal / tiny / inf / NaN
            uint32_t word = (uint32_t)(mexp + 2) << 24;
            for (int c = 0; c < 3; ++c) {
                if (b[c] == 0) continue;
                const int s = (int)(b[c] >> 23) - mexp + 7;                     // q = 1.mant * 2^s, s <= 7
                if (s < 0) { ok = false; return; }
                const uint32_t mant = (b[c] & 0x7fffffu) | 0x800000u;
                if (mant & ((1u << (23 - s)) - 1u)) { ok = false; return; }     // q is not an integer
                word |= (mant >> (23 - s)) << (8 * c);
            }
            rgbe[i] = word;
        }
    };
    int nt = (int)std::thread::hardware_concurrency(); if (nt < 1) nt = 1; if (nt > 16) nt = 16;
    if (n < (1u << 16)) nt = 1;
    std::vector<std::thread> th;
    for (int t = 1; t < nt; ++t) th.emplace_back(work, n * (size_t)t / (size_t)nt, n * (size_t)(t + 1) / (size_t)nt);
    work(0, n / (size_t)nt);
    for (auto& t : th) t.join();
    if (!ok) { give_words(std::move(rgbe)); rgbe = std::vector<uint32_t>(); return false; }
    return true;
}

// ------------------------------------------------------------------------------------------------------------
// flatten(): world list -> DObject records, per-mesh arrays, top-level tree
// ------------------------------------------------------------------------------------------------------------
namespace {

struct DBox { double mn[3], mx[3]; bool finite = true; };

// world-space bounds of a shape under its wrapper chain: the forward maps of translate.cpp:14, scale.cpp:21,
// rotateQuat.cpp:57, rotateY.cpp:60-66 applied to the eight corners (double precision, then padded)
DBox world_bounds(const HostShape& sh, const HostObject& o)
{
    DBox b; for (int a = 0; a < 3; ++a) { b.mn[a] = 1e300; b.mx[a] = -1e300; }
    switch (sh.kind) {
    case SHAPE_SPHERE: for (int a = 0; a < 3; ++a) { b.mn[a] = sh.f[a] - std::fabs((double)sh.f[3]); b.mx[a] = sh.f[a] + std::fabs((double)sh.f[3]); } break;
    case SHAPE_RECT: {
        int ka = sh.axis, ia = (sh.axis == 0) ? 1 : 0, ib = (sh.axis == 2) ? 1 : 2;
        b.mn[ka] = sh.f[4] - 1e-4; b.mx[ka] = sh.f[4] + 1e-4;
        b.mn[ia] = std::min(sh.f[0], sh.f[1]); b.mx[ia] = std::max(sh.f[0], sh.f[1]);
        b.mn[ib] = std::min(sh.f[2], sh.f[3]); b.mx[ib] = std::max(sh.f[2], sh.f[3]);
    } break;
    case SHAPE_BOX:                                                         // HostShape::f = centre, dimensions (box.h:41-44)
        for (int a = 0; a < 3; ++a) {
            double lo = sh.f[a] - sh.f[3 + a] / 2.0, hi = sh.f[a] + sh.f[3 + a] / 2.0;
            b.mn[a] = std::min(lo, hi); b.mx[a] = std::max(lo, hi);
        } break;
    case SHAPE_MESH:
        {   // every corner, through the index buffer when there is one.  Contiguous chunks, combined in array order: the
            // same first-of-equals result as one sequential pass
            const float* V = sh.indexed ? sh.V() : sh.pos.data(); const uint32_t* I = sh.indexed ? sh.I() : nullptr;
            std::vector<DBox> part(64, b); size_t used = 1;
            parallel_chunks((size_t)sh.n_tri * 3, 1 << 18, [&](size_t c, size_t lo, size_t hi) {
                DBox q = part[c];
                for (size_t i = lo; i < hi; ++i) {
                    const float* P = V + (I ? (size_t)I[i] : i) * 3;
                    for (int a = 0; a < 3; ++a) { double v = P[a]; q.mn[a] = std::min(q.mn[a], v); q.mx[a] = std::max(q.mx[a], v); }
                }
                part[c] = q;
            }, &used);
            for (size_t c = 0; c < used; ++c)
                for (int a = 0; a < 3; ++a) { b.mn[a] = std::min(b.mn[a], part[c].mn[a]); b.mx[a] = std::max(b.mx[a], part[c].mx[a]); }
        }
        if (sh.n_tri == 0) for (int a = 0; a < 3; ++a) { b.mn[a] = 0; b.mx[a] = 0; }
        break;
    default: break;
    }
    for (const DXform& x : o.ops) {
        double c[8][3];
        for (int k = 0; k < 8; ++k) for (int a = 0; a < 3; ++a) c[k][a] = ((k >> a) & 1) ? b.mx[a] : b.mn[a];
        for (int k = 0; k < 8; ++k) {
            double p[3] = {c[k][0], c[k][1], c[k][2]};
            switch (x.kind) {
            case HRT_XF_TRANSLATE: p[0] += x.a; p[1] += x.b; p[2] += x.c; break;
            case HRT_XF_SCALE: p[0] *= x.a; p[1] *= x.b; p[2] *= x.c; break;
            case HRT_XF_ROTATE_QUAT: {
                double w = x.a, qx = x.b, qy = x.c, qz = x.d;
                double nrm = w * w + qx * qx + qy * qy + qz * qz;
                if (!(std::fabs(nrm - 1.0) < 1e-3)) b.finite = false;            // not a rotation: no usable bound
                double ux = qy * p[2] - qz * p[1], uy = qz * p[0] - qx * p[2], uz = qx * p[1] - qy * p[0];
                double vx = qy * uz - qz * uy, vy = qz * ux - qx * uz, vz = qx * uy - qy * ux;
                p[0] += 2 * (w * ux + vx); p[1] += 2 * (w * uy + vy); p[2] += 2 * (w * uz + vz);
            } break;
            default: {                                                      // ROTATE_Y: a = sin, b = cos
                double px = x.b * p[0] + x.a * p[2], pz = -(double)x.a * p[0] + x.b * p[2];
                p[0] = px; p[2] = pz;
            } break;
            }
            for (int a = 0; a < 3; ++a) c[k][a] = p[a];
        }
        for (int a = 0; a < 3; ++a) { b.mn[a] = 1e300; b.mx[a] = -1e300; }
        for (int k = 0; k < 8; ++k) for (int a = 0; a < 3; ++a) { b.mn[a] = std::min(b.mn[a], c[k][a]); b.mx[a] = std::max(b.mx[a], c[k][a]); }
    }
    for (int a = 0; a < 3; ++a) {
        if (!std::isfinite(b.mn[a]) || !std::isfinite(b.mx[a]) || std::fabs(b.mn[a]) > 1e30 || std::fabs(b.mx[a]) > 1e30) b.finite = false;
        double pad = 1e-4 + 1e-5 * (std::fabs(b.mn[a]) + std::fabs(b.mx[a]));
        b.mn[a] -= pad; b.mx[a] += pad;
    }
    return b;
}

} // namespace

void build_mesh_host(HostShape& sh, int bvh_mode, uint64_t seed, int shape_id, int nt, double& ms_ref, double& ms_fast)
{
    const bool fast = bvh_mode != HRT_BVH_REFERENCE;
    const bool lbvh = bvh_mode == HRT_BVH_LBVH_HOST || bvh_mode == HRT_BVH_LBVH_GPU;
    const double t_enter = now_ms();
    ensure_soup(sh);
    double t0 = now_ms();
    build_ref_bvh(sh, seed, shape_id, nt, lbvh);
    ms_ref += now_ms() - t0;
    size_t n = (size_t)sh.n_tri;
    sh.tris4.resize(n * 3); sh.attrs4.resize(n * 4);
    for (size_t i = 0; i < n; ++i) {
        const float* P = &sh.pos[i * 9]; const float* N = &sh.nrm[i * 9]; const float* T = &sh.uv[i * 6];
        for (int v = 0; v < 3; ++v) { float4 q; q.x = P[v * 3]; q.y = P[v * 3 + 1]; q.z = P[v * 3 + 2]; q.w = 0; sh.tris4[i * 3 + v] = q; }
        float4 a0, a1, a2, a3;
        a0.x = N[0]; a0.y = N[1]; a0.z = N[2]; a0.w = T[0];
        a1.x = N[3]; a1.y = N[4]; a1.z = N[5]; a1.w = T[1];
        a2.x = N[6]; a2.y = N[7]; a2.z = N[8]; a2.w = T[2];
        a3.x = T[3]; a3.y = T[4]; a3.z = T[5]; a3.w = 0;
        sh.attrs4[i * 4] = a0; sh.attrs4[i * 4 + 1] = a1; sh.attrs4[i * 4 + 2] = a2; sh.attrs4[i * 4 + 3] = a3;
    }
    sh.leaves.clear(); sh.fast = FastBvh(); sh.wide = WideBvh(); sh.last_tri_r = -1;
    if (fast && sh.ref.root >= 0) {
        t0 = now_ms();
        size_t n_nodes = sh.ref.nodes.size() / 2;
        std::vector<float> bmin, bmax;
        for (size_t k = 0; k < n_nodes; ++k) {
            const float4& a = sh.ref.nodes[k * 2]; const float4& b = sh.ref.nodes[k * 2 + 1];
            int left, right; std::memcpy(&left, &a.w, 4); std::memcpy(&right, &b.w, 4);
            if (left >= 0) continue;                                     // not a bottom node
            bmin.push_back(a.x); bmin.push_back(a.y); bmin.push_back(a.z);
            bmax.push_back(b.x); bmax.push_back(b.y); bmax.push_back(b.z);
            size_t tl = (size_t)(~left), tr = (size_t)(~right);
            float4 rec[8];
            rec[0] = a; rec[1] = b;
            int il = (int)tl, ir = (int)tr; std::memcpy(&rec[0].w, &il, 4); std::memcpy(&rec[1].w, &ir, 4);
            for (int v = 0; v < 3; ++v) { rec[2 + v] = sh.tris4[tl * 3 + v]; rec[5 + v] = sh.tris4[tr * 3 + v]; }
            std::memcpy(&rec[2].w, &il, 4); std::memcpy(&rec[5].w, &ir, 4);   // the ids again, inside the 96 bytes a leaf visit reads
            sh.leaves.insert(sh.leaves.end(), rec, rec + 8);
            sh.last_tri_r = ir;
        }
        if (lbvh) {
            build_lbvh_host(bmin, bmax, sh.fast);
            if (sh.fast.depth > FAST_MAX_DEPTH) build_fast_bvh(bmin, bmax, sh.fast, nt);   // same decision as the device builder's fallback
        } else build_fast_bvh(bmin, bmax, sh.fast, nt);
        build_wide(sh.fast, sh.wide);
        ms_fast += now_ms() - t0;
    }
    if (getenv("HRT_DEBUG_BUILD")) fprintf(stderr, "[hrt] host build of shape %d: %lld triangles, total %.2f ms (reference topology so far %.2f, fast tree so far %.2f)\n",
                                           shape_id, (long long)sh.n_tri, now_ms() - t_enter, ms_ref, ms_fast);
    sh.n_ref_nodes = sh.ref.nodes.size() / 2; sh.n_fast_nodes = sh.fast.nodes.size() / 4; sh.n_leaves = sh.leaves.size() / 8; sh.n_wide_nodes = sh.wide.nodes.size();
}

// FNV-1a over the scene description.  Arrays above 64 KiB contribute their size and 4096 evenly spaced 16-byte blocks: the hash guards a
// checkpoint against being resumed on an EDITED scene (hrt_render_progressive), it is not a checksum of every texel.
namespace {
struct Fnv {
    uint64_t h = 1469598103934665603ull;
    void bytes(const void* p, size_t n) { const unsigned char* b = (const unsigned char*)p; for (size_t i = 0; i < n; ++i) { h ^= b[i]; h *= 1099511628211ull; } }
    template <class T> void pod(const T& v) { bytes(&v, sizeof(T)); }
    void array(const void* p, size_t n)
    {
        pod(n);
        if (!p || n == 0) return;
        if (n <= (size_t(64) << 10)) { bytes(p, n); return; }
        const size_t blocks = 4096, step = (n - 16) / (blocks - 1);
        for (size_t k = 0; k < blocks; ++k) bytes((const unsigned char*)p + k * step, 16);
    }
};
uint64_t scene_content_hash(const HostScene& s)
{
    Fnv f;
    f.pod(s.background); f.pod(s.bvh_mode); f.pod(s.n_prims);
    for (const HostTexture& t : s.textures) {
        f.pod(t.kind); f.bytes(t.rgb, sizeof(t.rgb)); f.pod(t.even); f.pod(t.odd); f.pod(t.w); f.pod(t.h); f.pod(t.ch);
        f.array(t.u8.data(), t.u8.size()); f.array(t.f32.data(), t.f32.size() * 4); f.array(t.rgbe.data(), t.rgbe.size() * 4);
    }
    for (const hrt_material& m : s.materials) { f.pod(m.kind); f.bytes(m.albedo, sizeof(m.albedo)); f.pod(m.albedo_tex); f.pod(m.p0); f.pod(m.p0_tex); f.pod(m.p1); f.pod(m.p1_tex); }
    for (const HostShape& sh : s.shapes) {
        f.pod(sh.kind); f.bytes(sh.f, sizeof(sh.f)); f.pod(sh.axis); f.pod(sh.material); f.pod(sh.n_tri); f.pod(sh.n_vert);
        if (sh.indexed) { f.array(sh.V(), (size_t)sh.n_vert * 12); f.array(sh.N(), (size_t)sh.n_vert * 12); f.array(sh.T(), (size_t)sh.n_vert * 8); f.array(sh.I(), (size_t)sh.n_tri * 12); }
        else { f.array(sh.pos.data(), sh.pos.size() * 4); f.array(sh.nrm.data(), sh.nrm.size() * 4); f.array(sh.uv.data(), sh.uv.size() * 4); }
    }
    for (const HostObject& o : s.objects) {
        f.pod(o.shape); f.pod(o.medium); f.pod(o.neg_inv_density); f.pod(o.phase_tex); f.bytes(o.phase_rgb, sizeof(o.phase_rgb));
        for (const DXform& x : o.ops) f.bytes(&x, sizeof(DXform));
    }
    return f.h;
}
} // namespace

std::string flatten(HostScene& s, int bvh_mode, uint64_t seed)
{
    s.content_hash = 0;
    const bool fast = bvh_mode != HRT_BVH_REFERENCE;
    int nt = (int)std::thread::hardware_concurrency(); if (nt < 1) nt = 1;
    s.bvh_mode = bvh_mode;
    s.content_hash = scene_content_hash(s);      // while every source array is still there (borrowed meshes are dropped after the commit)
    s.d_objects.clear(); s.d_xforms.clear(); s.mesh_shapes.clear();
    s.build_ms_ref = s.build_ms_fast = 0;
    for (HostShape& sh : s.shapes) { sh.used = false; sh.mesh_index = -1; }

    // meshes, in world-list order like the reference constructs them while loading (scene.cpp:292-297)
    for (const HostObject& o : s.objects) {
        if (o.shape < 0 || (size_t)o.shape >= s.shapes.size()) return "object refers to an unknown shape";
        HostShape& sh = s.shapes[(size_t)o.shape];
        if (sh.kind != SHAPE_MESH || sh.used) { sh.used = true; continue; }
        sh.used = true;
        sh.mesh_index = (int)s.mesh_shapes.size(); s.mesh_shapes.push_back(o.shape);
        sh.gpu_pending = sh.has_dev = false;
        if (bvh_mode == HRT_BVH_LBVH_GPU && sh.n_tri > 0) { sh.gpu_pending = true; continue; }   // built on the device by hrt_scene_commit
        build_mesh_host(sh, bvh_mode, seed, o.shape, nt, s.build_ms_ref, s.build_ms_fast);
    }

    // world list
    float cull_scale = 0.0f; bool any_quat = false;
    std::vector<float> obmin, obmax;
    for (size_t i = 0; i < s.objects.size(); ++i) {
        const HostObject& o = s.objects[i];
        const HostShape& sh = s.shapes[(size_t)o.shape];
        if ((int)o.ops.size() > MAX_XF) return "too many instance wrappers on one object";
        DObject d; std::memset(&d, 0, sizeof(d));
        d.kind = sh.kind; d.axis = sh.axis; d.material = sh.material; d.prim_base = (uint32_t)sh.prim_base;
        d.mesh = sh.mesh_index; d.xf_first = (int)s.d_xforms.size(); d.xf_count = (int)o.ops.size();
        d.medium = o.medium ? 1 : 0; d.neg_inv_density = o.neg_inv_density; d.phase_tex = o.phase_tex;
        for (int a = 0; a < 3; ++a) d.phase_rgb[a] = o.phase_rgb[a];
        d.medium_prim = o.medium ? (uint32_t)s.shapes[(size_t)o.pseudo_shape].prim_base : 0u;
        if (sh.kind == SHAPE_BOX) {                                          // Box::Box box.h:41-44
            for (int a = 0; a < 3; ++a) { d.f[a] = sh.f[a] - sh.f[3 + a] / 2.0f; d.f[3 + a] = sh.f[a] + sh.f[3 + a] / 2.0f; }
        } else for (int a = 0; a < 8; ++a) d.f[a] = sh.f[a];
        int innermost_quat = -1;
        for (int l = (int)o.ops.size() - 1; l >= 0; --l) if (o.ops[(size_t)l].kind == HRT_XF_ROTATE_QUAT) innermost_quat = l;
        d.has_quat = innermost_quat >= 0;
        if (d.has_quat) {
            float a = 1.0f;
            for (int l = innermost_quat + 1; l < (int)o.ops.size(); ++l)
                if (o.ops[(size_t)l].kind == HRT_XF_SCALE) {
                    float m = std::max(std::fabs(o.ops[(size_t)l].a), std::max(std::fabs(o.ops[(size_t)l].b), std::fabs(o.ops[(size_t)l].c)));
                    a *= 1.0f / m;
                }
            if (!(a > 0) || !std::isfinite(a)) a = 1e-30f;
            cull_scale = any_quat ? std::min(cull_scale, a) : a; any_quat = true;
        }
        for (const DXform& x : o.ops) s.d_xforms.push_back(x);
        s.d_objects.push_back(d);
        if (fast) {
            DBox b = world_bounds(sh, o);
            for (int a = 0; a < 3; ++a) {
                obmin.push_back(b.finite ? (float)std::max(b.mn[a], -3.0e38) : -3.0e38f);
                obmax.push_back(b.finite ? (float)std::min(b.mx[a], 3.0e38) : 3.0e38f);
            }
        }
    }
    s.tlas_cull_scale = any_quat ? cull_scale : 0.0f;
    s.tlas = FastBvh(); s.tlas_wide = WideBvh(); s.has_bounds = false; s.tlas_has_slow = false;
    if (fast && !s.objects.empty()) {
        // float(double) may round inwards by half an ulp: push the bounds one step outwards
        for (size_t k = 0; k < obmin.size(); ++k) { obmin[k] = std::nextafter(obmin[k], -std::numeric_limits<float>::infinity()); obmax[k] = std::nextafter(obmax[k], std::numeric_limits<float>::infinity()); }
        build_fast_bvh(obmin, obmax, s.tlas, 1);
        build_wide(s.tlas, s.tlas_wide);
        // top-level leaves of meshes that get entered carry WIDE_MESH_BIT (hrt_wide.cuh: the mesh-entry phase is told apart without a load)
        auto mark = [&](int32_t c) -> int32_t {
            if (c >= 0) return c;
            const DObject& ob = s.d_objects[(size_t)(~c)];
            if (ob.kind == SHAPE_MESH && !ob.medium) return (int32_t)~((uint32_t)(~c) | (1u << 30));        // WIDE_MESH_BIT
            const bool plain = !ob.medium && ob.xf_count == 0 && (ob.kind == SHAPE_SPHERE || ob.kind == SHAPE_RECT);
            if (!plain) s.tlas_has_slow = true;
            return plain ? c : (int32_t)~((uint32_t)(~c) | (1u << 29));                                     // WIDE_SLOW_BIT
        };
        for (WNode& n : s.tlas_wide.nodes)
            for (int k = 0; k < 8; ++k) if ((n.meta >> (24 + k)) & 1u) n.child[k] = mark(n.child[k]);
        s.tlas_wide.root = mark(s.tlas_wide.root);
        s.has_bounds = true;
        for (int a = 0; a < 3; ++a) { s.bounds_lo[a] = 3.0e38f; s.bounds_hi[a] = -3.0e38f; }
        for (size_t k = 0; k < obmin.size(); ++k) {
            const int a = (int)(k % 3);
            s.bounds_lo[a] = std::min(s.bounds_lo[a], obmin[k]); s.bounds_hi[a] = std::max(s.bounds_hi[a], obmax[k]);
            if (!(obmin[k] > -1.0e30f) || !(obmax[k] < 1.0e30f)) s.has_bounds = false;
        }
    }
    return "";
}

} // namespace hrt
