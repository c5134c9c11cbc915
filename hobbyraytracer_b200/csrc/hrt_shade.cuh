// hrt_shade.cuh — Texture::colourValue (texture.h / texture.cpp), Material::scatter / emitted (material.h /
// material.cpp), the background lookup of rayColour (main.cpp:47-58), the camera (camera.h:32-39) and the film
// (film.cpp:25-52) as a switch over packed records.  Random numbers come from Philox (hrt_rng.cuh), so parity of
// this file with the reference is statistical; the arithmetic is kept in the reference's order and precision
// (including its double-precision spots) so that the only difference is the random stream and libm.
#pragma once
#include "hrt_trace.cuh"

namespace hrt {

// ---- textures --------------------------------------------------------------------------------------------------------
HRT_HDN inline V3 tex_value(const DScene& s, int id, float u, float v, V3 p)
{
    for (int guard = 0; guard < 16; ++guard) {
        const DTexture& t = s.textures[id];
        switch (t.kind) {
        case TEX_SOLID: return v3(t.rgb[0], t.rgb[1], t.rgb[2]);                 // texture.h:18-21
        case TEX_CHECKER: {                                                     // texture.cpp:17-28
            float sines = sinf(10 * p.x) * sinf(10 * p.y) * sinf(10 * p.z);
            id = sines < 0 ? t.odd : t.even;
            continue;
        }
        case TEX_IMAGE: {                                                       // texture.cpp:53-74
            if (t.data == nullptr) return v3(0, 1, 1);
            u = cmin(cmax(u, 0.0f), 1.0f);
            v = 1.0f - cmin(cmax(v, 0.0f), 1.0f);
            int i = (int)(u * t.w);
            int j = (int)(v * t.h);
            if (i >= t.w) i = t.w - 1;
            if (j >= t.h) j = t.h - 1;
            const float sc = 1.0f / 255.0f;
            const uint8_t* d = (const uint8_t*)t.data + ((size_t)j * (3 * (size_t)t.w) + (size_t)i * 3);
            return v3(sc * d[0], sc * d[1], sc * d[2]);
        }
        case TEX_ENV: case TEX_ENV_RGBE: {                                      // texture.cpp:76-97
            if (t.data == nullptr) return v3(0, 1, 1);
            u = cmin(cmax(u, 0.0f), 1.0f);
            v = cmin(cmax(v, 0.0f), 1.0f);
            int i = (int)((u * (t.w - 1)) + 0.5f);
            int j = (int)((v * (t.h - 1)) + 0.5f);
            size_t px = (size_t)j * t.w + i;
            if (t.kind == TEX_ENV) {
                const float* d = (const float*)t.data + px * t.ch;
                return v3(d[0], d[1], d[2]);
            }
            // lossless repack of a decoded Radiance file: byte * 2^(e - 136), exactly stb_image's stbi__hdr_convert
            uint32_t w = ((const uint32_t*)t.data)[px];
            int e = (int)(w >> 24);
            if (e == 0) return v3(0.0f);
            float f = ldexpf(1.0f, e - 136);
            return v3((float)(w & 255u) * f, (float)((w >> 8) & 255u) * f, (float)((w >> 16) & 255u) * f);
        }
        }
        break;
    }
    return v3(0.0f);
}
HRT_HD V3 matvec3(const DScene& s, const float c[3], int tex, float u, float v, V3 p)         // material.h:20-31
{ return tex >= 0 ? tex_value(s, tex, u, v, p) : v3(c[0], c[1], c[2]); }
HRT_HD float matscalar(const DScene& s, float c, int tex, float u, float v, V3 p)             // material.h:43-54
{ return tex >= 0 ? length(tex_value(s, tex, u, v, p)) : c; }

// ---- materials -------------------------------------------------------------------------------------------------------
HRT_HD bool near_zero(V3 e)                                                      // hobbyraytracer.h:34-38 (double threshold)
{ const double s = 1e-8; return ((double)fabsf(e.x) < s) && ((double)fabsf(e.y) < s) && ((double)fabsf(e.z) < s); }

HRT_HDN inline V3 mat_emitted(const DScene& s, int mat, float u, float v, V3 p)
{
    if (mat >= MAT_MEDIUM_BASE) return v3(0.0f);
    const hrt_material& m = s.materials[mat];
    if (m.kind != HRT_MAT_DIFFUSE_LIGHT) return v3(0.0f);                        // material.h:67-70
    return matvec3(s, m.albedo, m.albedo_tex, u, v, p) * matscalar(s, m.p0, m.p0_tex, u, v, p);   // material.h:101-104
}

HRT_HD bool scatter_lambertian(const DScene& s, const hrt_material& m, const Rec& rec, V3& att, Ray& sc, const U4& w, int mode)   // material.h:137-153
{
    V3 dir = rec.n + spherical_rand(w.x, w.y, mode);
    if (near_zero(dir)) dir = rec.n;
    sc.o = rec.p; sc.d = dir;
    att = matvec3(s, m.albedo, m.albedo_tex, rec.u, rec.v, rec.p);
    return true;
}
HRT_HD bool scatter_metal(const DScene& s, const hrt_material& m, const Ray& rin, const Rec& rec, V3& att, Ray& sc, const U4& w, int mode)   // material.h:166-177
{
    V3 reflected = reflect(normalize(rin.d), normalize(rec.n));
    float rough = fabsf(matscalar(s, m.p0, m.p0_tex, rec.u, rec.v, rec.p));      // glm::length(float) = abs
    rough = rough < 1 ? rough : 1;
    sc.o = rec.p; sc.d = reflected + rough * spherical_rand(w.x, w.y, mode) + v3(1.1920928955078125e-7f);
    att = matvec3(s, m.albedo, m.albedo_tex, rec.u, rec.v, rec.p);
    return dot(sc.d, normalize(rec.n)) > 0;
}
HRT_HD double reflectance(double cosine, float refIdx)                           // material.h:236-241
{
    double r0 = (1 - refIdx) / (1 + refIdx);
    r0 = r0 * r0;
    return r0 + (1 - r0) * pow((1 - cosine), 5.0);
}
HRT_HD bool scatter_dielectric(const DScene& s, const hrt_material& m, const Ray& rin, const Rec& rec, V3& att, Ray& sc, const U4& w, int mode)   // material.h:204-228
{
    att = v3(1.0f, 1.0f, 1.0f);
    float ior = matscalar(s, m.p0, m.p0_tex, rec.u, rec.v, rec.p);
    float ratio = rec.front ? (1.0f / ior) : ior;
    V3 unit = normalize(rin.d);
    double cosTheta = cmin(dot(-unit, rec.n), 1.0f);
    double sinTheta = sqrt(1.0 - cosTheta * cosTheta);
    bool cannot = ratio * sinTheta > 1.0;
    V3 direction;
    if (cannot || reflectance(cosTheta, ratio) > linear_d(w.x, w.y, mode, 0.0, 1.0)) direction = reflect(unit, rec.n);
    else direction = refract(unit, rec.n, ratio);
    sc.o = rec.p; sc.d = direction + matscalar(s, m.p1, m.p1_tex, rec.u, rec.v, rec.p) * spherical_rand(w.z, w.w, mode);
    return true;
}

HRT_HDN inline bool mat_scatter(const DScene& s, int mat, const Ray& rin, const Rec& rec, V3& att, Ray& sc, const RngKey& key, uint32_t bounce)
{
    if (mat >= MAT_MEDIUM_BASE) {                                               // Isotropic of a ConstantMedium, material.h:79-85
        const DObject& ob = s.objects[mat - MAT_MEDIUM_BASE];
        sc.o = rec.p; sc.d = ball_rand(key, bounce);
        att = ob.phase_tex >= 0 ? tex_value(s, ob.phase_tex, rec.u, rec.v, rec.p) : v3(ob.phase_rgb[0], ob.phase_rgb[1], ob.phase_rgb[2]);
        return true;
    }
    const hrt_material& m = s.materials[mat];
    if (m.kind == HRT_MAT_DIFFUSE_LIGHT) return false;                          // material.h:96-99
    if (m.kind == HRT_MAT_ISOTROPIC) {                                          // material.h:79-85
        sc.o = rec.p; sc.d = ball_rand(key, bounce);
        att = matvec3(s, m.albedo, m.albedo_tex, rec.u, rec.v, rec.p);
        return true;
    }
    const U4 w = rng_draw(key, bounce, SLOT_SCATTER);
    switch (m.kind) {
    case HRT_MAT_LAMBERTIAN: return scatter_lambertian(s, m, rec, att, sc, w, key.mode);
    case HRT_MAT_METAL:      return scatter_metal(s, m, rin, rec, att, sc, w, key.mode);
    case HRT_MAT_DIELECTRIC: return scatter_dielectric(s, m, rin, rec, att, sc, w, key.mode);
    case HRT_MAT_PBR: {                                                         // material.cpp:18-28
        V3 mix = m.p1_tex >= 0 ? tex_value(s, m.p1_tex, rec.u, rec.v, rec.p) : v3(m.p1);
        if (length(mix) > 0.5f) return scatter_metal(s, m, rin, rec, att, sc, w, key.mode);
        return scatter_lambertian(s, m, rec, att, sc, w, key.mode);
    }
    case HRT_MAT_UVTEST: {                                                      // material.h:116-129
        V3 dir = rec.n + spherical_rand(w.x, w.y, key.mode);
        if (near_zero(dir)) dir = rec.n;
        sc.o = rec.p; sc.d = dir;
        att = rec.n;
        return true;
    }
    }
    return false;
}

// background lookup of rayColour, main.cpp:47-58.  The reference's unqualified atan2 / acos resolve to the double
// overloads and are rounded back to float; here they are the fp32 functions (<= 2 ulp from that value), which moves a
// nearest-texel lookup only when u or v sits within 2 ulp of a texel boundary.  Transcendentals are outside the bit-exact
// contract; the +0.5 stays a double add like main.cpp:55.
HRT_HDN inline V3 background_colour(const DScene& s, const Ray& r)
{
    if (s.background < 0) return v3(0.0f);
    if (s.textures[s.background].kind == TEX_SOLID) return tex_value(s, s.background, 0.0f, 0.0f, v3(0.0f));
    V3 nD = normalize(r.d);
    float phi = atan2f(nD.z, nD.x);
    float theta = acosf(nD.y);
    const float pi = 3.14159265358979323846264338327950288f;
    float u = (float)((double)(phi / (2 * pi)) + 0.5);
    float v = theta / pi;
    return tex_value(s, s.background, u, v, v3(0.0f));
}

// One iteration of rayColour's loop (main.cpp:43-76) after world->hit: returns true when the path continues with
// `r` and `atten` updated; `radiance` receives this segment's contribution (attenuation * emitted or * background).
HRT_HDN inline bool shade_segment(const DScene& s, const Hit& h, Ray& r, V3& atten, V3& radiance, const RngKey& key, uint32_t bounce)
{
    if (h.prim == HRT_MISS) { radiance = atten * background_colour(s, r); return false; }
    Rec rec; eval_hit(s, r, h, rec);
    Ray sc; sc.o = v3(0.0f); sc.d = v3(0.0f); V3 att = v3(0.0f);
    V3 emitted = mat_emitted(s, rec.mat, rec.u, rec.v, rec.p);
    bool ok = mat_scatter(s, rec.mat, r, rec, att, sc, key, bounce);
    radiance = atten * emitted;
    if (!ok) return false;
    atten = atten * att;
    r = sc;
    return true;
}

// ---- Russian roulette (opt-in, HRT_RENDER_RR): the reference has none (main.cpp:43-76 runs every path to MAX_DEPTH or a miss) ------
// After the scatter of segment `bounce` (0-based), from the third segment on, the path survives with probability
// p = clamp(max(attenuation), 0.05, 1) and its attenuation is divided by p: the estimator is unchanged in expectation, dim paths stop
// early.  The draw is its own Philox slot, so no other random number of the path moves.  Returns false when the path is cut.
constexpr uint32_t RR_FIRST_BOUNCE = 2;
HRT_HD bool russian_roulette(V3& atten, const RngKey& key, uint32_t bounce)
{
    if (bounce < RR_FIRST_BOUNCE) return true;
    float p = fmaxf(atten.x, fmaxf(atten.y, atten.z));
    p = p < 0.05f ? 0.05f : (p > 1.0f ? 1.0f : p);
    if (!(p < 1.0f)) return true;                                             // (also: NaN attenuation goes on like the reference's)
    const float xi = (float)rng_draw(key, bounce, SLOT_RR).x * 2.3283064365386963e-10f;   // [0, 1): plain uniform whatever the rng mode
    if (xi >= p) return false;
    atten = atten / p;
    return true;
}

// ---- camera: getRay camera.h:32-39 (the lens offset is forced to zero there) ---------------------------------------------
// lens_word: a spare word of the camera-jitter draw, used only under HRT_RENDER_DEFOCUS (c.lens_radius > 0) for
// glm::circularRand(lensRadius) (random.inl:281-287): a = linearRand(0, 2 pi), rd = (cos a, sin a) * Radius
// LENS = false compiles the lens code out: the fused trace kernel is instantiated both ways and picked by the host, because
// even the never-taken branch cost the headline 2.8 % (B200 A/B: 8287 -> 8054 Mrays/s, profiles/README.md r1m).
template <bool LENS = true>
HRT_HD Ray camera_ray(const DCamera& c, float s, float t, uint32_t lens_word = 0, int mode = 0)
{
    V3 origin = v3(c.origin[0], c.origin[1], c.origin[2]), llc = v3(c.llc[0], c.llc[1], c.llc[2]);
    V3 horiz = v3(c.horiz[0], c.horiz[1], c.horiz[2]), vert = v3(c.vert[0], c.vert[1], c.vert[2]);
    V3 off = v3(0.0f);
    if (LENS && c.lens_radius > 0.0f) {
        const float a = linear_f(lens_word, mode, 0.0f, 6.283185307179586476925286766559f);
        const float rx = cosf(a) * c.lens_radius, ry = sinf(a) * c.lens_radius;
        off = v3(c.u[0], c.u[1], c.u[2]) * rx + v3(c.v[0], c.v[1], c.v[2]) * ry;    // camera.h:35
    }
    Ray r; r.o = origin + off; r.d = llc + s * horiz + t * vert - origin - off;
    return r;
}
// the primary ray of (pixel, sample): render() main.cpp:115-123 — y = H - row, u-jitter drawn before v-jitter
HRT_HD Ray primary_ray(const DCamera& c, int w, int h, int pidx, const RngKey& key)
{
    int x = pidx % w, y = h - pidx / w;
    U4 j = rng_draw(key, BOUNCE_CAMERA, 0);
    float u = ((float)x + linear_f(j.x, key.mode, 0.0f, 1.0f)) / (w - 1);
    float v = ((float)y + linear_f(j.y, key.mode, 0.0f, 1.0f)) / (h - 1);
    return camera_ray(c, u, v, j.z, key.mode);
}

// same ray from the pixel coordinates the caller already has (x = pIdx % W, y = H - pIdx / W, main.cpp:115-116)
template <bool LENS = true>
HRT_HD Ray primary_ray_xy(const DCamera& c, int w, int h, int x, int y, const RngKey& key)
{
    U4 j = rng_draw(key, BOUNCE_CAMERA, 0);
    float u = ((float)x + linear_f(j.x, key.mode, 0.0f, 1.0f)) / (w - 1);
    float v = ((float)y + linear_f(j.y, key.mode, 0.0f, 1.0f)) / (h - 1);
    return camera_ray<LENS>(c, u, v, j.z, key.mode);
}

// ---- Film::tonemap + writeColour film.cpp:25-52 -----------------------------------------------------------------------------
HRT_HD void tonemap_write(V3 c, uint8_t* px)
{
    if (c.x != c.x) c.x = 0.0f;
    if (c.y != c.y) c.y = 0.0f;
    if (c.z != c.z) c.z = 0.0f;
    const float a = 2.51f, b = 0.03f, cc = 2.43f, d = 0.59f, e = 0.14f;
    V3 num = c * (a * c + v3(b));
    V3 den = c * (cc * c + v3(d)) + v3(e);
    V3 q = num / den;
    q = v3(cmin(cmax(q.x, 0.0f), 1.0f), cmin(cmax(q.y, 0.0f), 1.0f), cmin(cmax(q.z, 0.0f), 1.0f));
    q = v3(sqrtf(q.x), sqrtf(q.y), sqrtf(q.z));
    px[0] = (uint8_t)(256 * cmin(cmax(q.x, 0.0f), 0.9999f));
    px[1] = (uint8_t)(256 * cmin(cmax(q.y, 0.0f), 0.9999f));
    px[2] = (uint8_t)(256 * cmin(cmax(q.z, 0.0f), 0.9999f));
}

// ---- film: order-independent sample accumulation ---------------------------------------------------------------------------
// Per pixel four int64 lanes: lanes 0-2 hold the r, g, b sample SUMS in fixed point (units of 2^-FILM_FRAC_BITS, each
// sample rounded to nearest and clamped to +-2^FILM_MAG_BITS), lane 3 holds nine 7-bit fields "a NaN / +inf / -inf
// sample was added to channel c" (set with an atomic OR locally, so <= 1 per rank; a cross-rank integer SUM of up to
// 127 ranks keeps them non-zero).  Integer addition commutes, so the film does not depend on the order in which paths
// finish, on the wavefront capacity or on how samples are split over GPUs.  The reference sums floats in sample order
// (main.cpp:123); with 26 fractional bits the rounding of one sample is <= 7.5e-9 absolute, below fp32's own rounding
// of a running sum >= 0.25, and unbiased.
constexpr int FILM_LANES = 4;
constexpr int FILM_FRAC_BITS = 26;
constexpr int FILM_MAG_BITS = 21;

HRT_HD void film_quantize(float v, int ch, long long& q, unsigned long long& flag)
{
    q = 0; flag = 0;
    if (v != v) { flag = 1ull << (7 * ch); return; }
    if (fabsf(v) > 3.402823466e+38f) { flag = 1ull << (7 * (3 + ch + (v < 0 ? 3 : 0))); return; }
    const float lim = (float)(1 << FILM_MAG_BITS);
    v = v < -lim ? -lim : (v > lim ? lim : v);
    const float scaled = v * (float)(1 << FILM_FRAC_BITS);                      // power-of-two scale: exact
#if defined(__CUDA_ARCH__)
    q = __float2ll_rn(scaled);
#else
    q = (long long)nearbyintf(scaled);                                          // round-to-nearest-even, like cvt.rni
#endif
}
HRT_HD float film_channel(const long long* px, int ch)
{
    const unsigned long long f = (unsigned long long)px[3];
    const bool nan = ((f >> (7 * ch)) & 0x7f) != 0, pinf = ((f >> (7 * (3 + ch))) & 0x7f) != 0, ninf = ((f >> (7 * (6 + ch))) & 0x7f) != 0;
    if (nan || (pinf && ninf)) return int_as_float(0x7fc00000);
    if (pinf) return hrt_inf();
    if (ninf) return -hrt_inf();
    return (float)((double)px[ch] * (1.0 / (double)(1 << FILM_FRAC_BITS)));
}

} // namespace hrt
