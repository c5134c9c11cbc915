// hrt_rng.cuh — counter-based Philox4x32-10 and the glm::linearRand / sphericalRand / ballRand distributions
// (dependencies/glm/glm/gtc/random.inl) on top of it.
//
// The reference draws every number from the process-global std::rand() (random.inl:25); a GPU needs a stateless
// generator, so the stream is different and parity of everything that consumes random numbers is statistical.
// What IS kept (rng_mode 0) is the reference's distribution: each 32-bit integer is assembled from four
// "bytes" rand() % 255, i.e. four base-255 digits, so xi = float(u32) / float(UINT32_MAX) lives on [0, 0.99608]
//  — sphericalRand therefore never samples a 1.4 deg azimuth wedge and a 7.2 deg polar cap.
// rng_mode 1 uses all 32 bits (plain uniform).
//
// Counter layout: key = seed (lo, hi); counter = (index, sample, bounce, slot).
//   index  : pixel index (render) or ray index (trace_batch)
//   sample : global sample index of the pixel -> the image does not depend on how samples are split over GPUs
//   bounce : path segment, 0xFFFFFFFF for the camera jitter
//   slot   : SLOT_* below; one Philox call yields four 32-bit words
#pragma once
#include "hrt_math.cuh"

namespace hrt {

enum { SLOT_SCATTER = 0, SLOT_RR = 8, SLOT_BALL = 16, SLOT_MEDIUM = 0x10000 };      // SLOT_BALL + rejection round stays below 0x10000
constexpr uint32_t BOUNCE_CAMERA = 0xFFFFFFFFu;

struct U4 { uint32_t x, y, z, w; };

HRT_HD uint32_t mulhi32(uint32_t a, uint32_t b)
{
#if defined(__CUDA_ARCH__)
    return __umulhi(a, b);
#else
    return (uint32_t)(((uint64_t)a * (uint64_t)b) >> 32);
#endif
}

HRT_HD U4 philox4x32_10(uint32_t k0, uint32_t k1, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3)
{
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = mulhi32(M0, c0), lo0 = M0 * c0;
        uint32_t hi1 = mulhi32(M1, c2), lo1 = M1 * c2;
        uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += W0; k1 += W1;
    }
    U4 o; o.x = c0; o.y = c1; o.z = c2; o.w = c3; return o;
}

struct RngKey {
    uint32_t k0, k1;       // seed
    uint32_t index;        // pixel / ray
    uint32_t sample;
    int mode;              // 0 reference lattice, 1 plain
};
HRT_HD U4 rng_draw(const RngKey& k, uint32_t bounce, uint32_t slot)
{ return philox4x32_10(k.k0, k.k1, k.index, k.sample, bounce, slot); }

// A 32-bit word -> the integer glm's detail::compute_rand assembles from four rand() % 255 bytes
// (random.inl:19-27, 65-85): four base-255 digits placed in the four bytes.
HRT_HD uint32_t lattice_u32(uint32_t w, int mode)
{
    if (mode != 0) return w;
    uint32_t v = mulhi32(w, 4228250625u);            // ~uniform on [0, 255^4)
    uint32_t b0 = v % 255u; v /= 255u;
    uint32_t b1 = v % 255u; v /= 255u;
    uint32_t b2 = v % 255u; v /= 255u;
    return b0 | (b1 << 8) | (b2 << 16) | (v << 24);
}
// linearRand(float Min, float Max) random.inl:176-183
HRT_HD float linear_f(uint32_t w, int mode, float mn, float mx)
{ return (float)lattice_u32(w, mode) / 4294967295.0f * (mx - mn) + mn; }
// linearRand(double Min, double Max) random.inl:185-192
HRT_HD double linear_d(uint32_t wlo, uint32_t whi, int mode, double mn, double mx)
{
    uint64_t u = (uint64_t)lattice_u32(wlo, mode) | ((uint64_t)lattice_u32(whi, mode) << 32);
    return (double)u / 18446744073709551615.0 * (mx - mn) + mn;
}
// sphericalRand(1.0f) random.inl:290-302
HRT_HD V3 spherical_rand(uint32_t w0, uint32_t w1, int mode)
{
    float theta = linear_f(w0, mode, 0.0f, 6.283185307179586476925286766559f);
    float phi = acosf(linear_f(w1, mode, -1.0f, 1.0f));
    float sp, cp, st, ct;
#if defined(__CUDA_ARCH__)
    sincosf(phi, &sp, &cp); sincosf(theta, &st, &ct);
#else
    sp = sinf(phi); cp = cosf(phi); st = sinf(theta); ct = cosf(theta);
#endif
    return v3(sp * ct, sp * st, cp) * 1.0f;
}
// ballRand(1.0f) random.inl:261-278: rejection sampling of linearRand(vec3(-1), vec3(1)); one Philox call per round
HRT_HD V3 ball_rand(const RngKey& k, uint32_t bounce)
{
    V3 r = v3(0.0f);
    for (uint32_t round = 0; round < 64; ++round) {
        U4 w = rng_draw(k, bounce, SLOT_BALL + round);
        r = v3(linear_f(w.x, k.mode, -1.0f, 1.0f), linear_f(w.y, k.mode, -1.0f, 1.0f), linear_f(w.z, k.mode, -1.0f, 1.0f));
        if (!(length(r) > 1.0f)) break;
    }
    return r;
}

} // namespace hrt
