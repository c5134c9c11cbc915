// hrt_math.cuh — fp32 vector arithmetic in GLM 0.9.9.9's evaluation order (the arithmetic spec of the reference,
// dependencies/glm/glm/detail/func_geometric.inl, type_quat.inl), usable from device code and — for the CPU
// simulation harness in tests/hostsim, which compiles these same headers with g++ — from host code.
//
// Numerics contract: every translation unit including this header is compiled with
// -fmad=false (nvcc) / -ffp-contract=off (g++), IEEE division and sqrt, no flush-to-zero.  Places that WANT a
// fused multiply-add (the conservative slab test of the SAH BVH) ask for it explicitly with hrt_fma().
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define HRT_HD __host__ __device__ __forceinline__
#define HRT_HDN __host__ __device__
#else
#define HRT_HD inline
#define HRT_HDN
#endif

namespace hrt {

struct V3 { float x, y, z; };
HRT_HD V3 v3(float a, float b, float c) { V3 r; r.x = a; r.y = b; r.z = c; return r; }
HRT_HD V3 v3(float a) { return v3(a, a, a); }
HRT_HD V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
HRT_HD V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
HRT_HD V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
HRT_HD V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
HRT_HD V3 operator/(V3 a, V3 b) { return v3(a.x / b.x, a.y / b.y, a.z / b.z); }
HRT_HD V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
HRT_HD V3 operator*(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
HRT_HD V3 operator/(V3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
HRT_HD float comp(const V3& a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }

// func_geometric.inl:48-55: tmp = a*b; (tmp.x + tmp.y) + tmp.z
HRT_HD float dot(V3 a, V3 b) { V3 t = a * b; return t.x + t.y + t.z; }
// func_geometric.inl:68-79
HRT_HD V3 cross(V3 x, V3 y) { return v3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y); }
HRT_HD float length(V3 a) { return sqrtf(dot(a, a)); }                         // func_geometric.inl:8-14
HRT_HD V3 normalize(V3 a) { return a * (1.0f / sqrtf(dot(a, a))); }            // func_geometric.inl:82-90
HRT_HD V3 reflect(V3 I, V3 N) { return I - N * dot(N, I) * 2.0f; }             // func_geometric.inl:104-110
HRT_HD V3 refract(V3 I, V3 N, float eta)                                       // func_geometric.inl:113-123
{
    float d = dot(N, I);
    float k = 1.0f - eta * eta * (1.0f - d * d);
    return (k >= 0.0f) ? (eta * I - (eta * d + sqrtf(k)) * N) : v3(0.0f);
}
// std::min/max, glm::min/max: (b < a) ? b : a and (a < b) ? b : a — NaN in the first operand propagates, in the
// second it is dropped (aabb.h:29-34 depends on this).  Never fminf/fmaxf in reference-exact code.
HRT_HD float cmin(float a, float b) { return (b < a) ? b : a; }
HRT_HD float cmax(float a, float b) { return (a < b) ? b : a; }

struct Quat { float w, x, y, z; };
HRT_HD V3 qrot(Quat q, V3 v)                                                   // type_quat.inl:347-354
{
    V3 qv = v3(q.x, q.y, q.z);
    V3 uv = cross(qv, v);
    V3 uuv = cross(qv, uv);
    return v + ((uv * q.w) + uuv) * 2.0f;
}
HRT_HD Quat qconj(Quat q) { Quat r; r.w = q.w; r.x = -q.x; r.y = -q.y; r.z = -q.z; return r; }

HRT_HD float hrt_fma(float a, float b, float c)
{
#if defined(__CUDA_ARCH__)
    return __fmaf_rn(a, b, c);
#else
    return fmaf(a, b, c);
#endif
}
HRT_HD float hrt_inf()
{
#if defined(__CUDA_ARCH__)
    return __int_as_float(0x7f800000);
#else
    return INFINITY;
#endif
}
HRT_HD bool hrt_finite(float x) { return fabsf(x) <= 3.402823466e+38f; }        // false for NaN and +-inf
HRT_HD float int_as_float(int i)
{
#if defined(__CUDA_ARCH__)
    return __int_as_float(i);
#else
    float f; __builtin_memcpy(&f, &i, 4); return f;
#endif
}
HRT_HD int float_as_int(float f)
{
#if defined(__CUDA_ARCH__)
    return __float_as_int(f);
#else
    int i; __builtin_memcpy(&i, &f, 4); return i;
#endif
}

struct Ray { V3 o, d; };
HRT_HD V3 ray_at(const Ray& r, float t) { return r.o + (t * r.d); }            // ray.h:10

} // namespace hrt
