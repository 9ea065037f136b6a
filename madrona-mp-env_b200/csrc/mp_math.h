// mp_math.h -- float math for the MadronaMPEnv step, written so that the SAME
// sequence of IEEE-754 single-precision operations runs on the GPU (nvcc
// -fmad=false, default -prec-div/-prec-sqrt) and on the host (g++
// -ffp-contract=off).  No implicit fused multiply-add anywhere; fmaf() only
// where the reference itself calls fmaf (mesh_bvh.inl:451-461, 509).
//
// Vector / quaternion operation ORDER follows the reference's math library
// (reference src/brt/math.inl: Vector3 ops :240-301, :412-475; Quat :686-739,
// :890-898; AABB::contains :1637-1647) because rounding depends on it.
//
// The transcendental functions (mp_sinf, mp_cosf, mp_logf, mp_atan2f,
// mp_asinf) are this project's own single-precision implementations (Cody &
// Waite range reduction + minimax polynomials in the classic Cephes style).
// libm's and CUDA's built-ins each differ from one another by 1-2 ulp, which a
// chaotic simulation amplifies into different integer state after a few
// hundred steps; a shared implementation makes GPU and CPU runs bit-identical.
// The oracle is built twice: once with libm (the reference's own arithmetic)
// and once with these routed in (oracle/det_math_override.h).
#pragma once

#include "mp_defs.h"

namespace mpenv {

// ------------------------------------------------------------------ bit casts
MP_HD uint32_t f2u(float f)
{
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    uint32_t u; __builtin_memcpy(&u, &f, 4); return u;
#endif
}
MP_HD float u2f(uint32_t u)
{
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float f; __builtin_memcpy(&f, &u, 4); return f;
#endif
}

// ------------------------------------------------------------- transcendentals
namespace detail {
constexpr float kFOPI = 1.27323954473516f;          // 4/pi
constexpr float kDP1 = 0.78515625f;                 // pi/4 split in three
constexpr float kDP2 = 2.4187564849853515625e-4f;
constexpr float kDP3 = 3.77489497744594108e-8f;

MP_HD float sinPoly(float x, float z)
{
    float y = -1.9515295891e-4f * z;
    y = y + 8.3321608736e-3f;
    y = y * z;
    y = y - 1.6666654611e-1f;
    y = y * z;
    y = y * x;
    return y + x;
}
MP_HD float cosPoly(float z)
{
    float y = 2.443315711809948e-5f * z;
    y = y - 1.388731625493765e-3f;
    y = y * z;
    y = y + 4.166664568298827e-2f;
    y = y * z;
    y = y * z;
    y = y - 0.5f * z;
    return y + 1.0f;
}
// Reduces |x| to [-pi/4, pi/4]; returns octant-pair index j in {0,2,4,6}/2 form.
MP_HD float reduce(float ax, int32_t *j_out)
{
    int32_t j = (int32_t)(kFOPI * ax);
    float y = (float)j;
    if (j & 1) { j += 1; y += 1.0f; }
    *j_out = j;
    float r = ax - y * kDP1;
    r = r - y * kDP2;
    r = r - y * kDP3;
    return r;
}
}  // namespace detail

MP_HD float mp_sinf(float xx)
{
    float sign = 1.f;
    float x = xx;
    if (x < 0.f) { sign = -1.f; x = -x; }
    if (!(x < 8388608.f)) return 0.f;   // out of the supported range (incl. NaN/inf)
    int32_t j;
    x = detail::reduce(x, &j);
    j &= 7;
    if (j > 3) { sign = -sign; j -= 4; }
    float z = x * x;
    float y = (j == 1 || j == 2) ? detail::cosPoly(z) : detail::sinPoly(x, z);
    return sign < 0.f ? -y : y;
}

MP_HD float mp_cosf(float xx)
{
    float sign = 1.f;
    float x = xx < 0.f ? -xx : xx;
    if (!(x < 8388608.f)) return 1.f;
    int32_t j;
    x = detail::reduce(x, &j);
    j &= 7;
    if (j > 3) { j -= 4; sign = -sign; }
    if (j > 1) { sign = -sign; }
    float z = x * x;
    float y = (j == 1 || j == 2) ? detail::sinPoly(x, z) : detail::cosPoly(z);
    return sign < 0.f ? -y : y;
}

MP_HD float mp_logf(float xx)
{
    if (xx != xx) return xx;
    if (xx < 0.f) return u2f(0x7FC00000u);
    if (xx == 0.f) return u2f(0xFF800000u);         // -inf
    if (xx == u2f(0x7F800000u)) return xx;
    uint32_t bits = f2u(xx);
    int32_t e = 0;
    if ((bits & 0x7F800000u) == 0) {                // subnormal: scale by 2^23
        xx = xx * 8388608.f;
        bits = f2u(xx);
        e = -23;
    }
    // frexp: xx = m * 2^e, m in [0.5, 1)
    e += (int32_t)((bits >> 23) & 0xFF) - 126;
    float x = u2f((bits & 0x807FFFFFu) | 0x3F000000u);
    if (x < 0.707106781186547524f) {
        e -= 1;
        x = x + x;
        x = x - 1.0f;
    } else {
        x = x - 1.0f;
    }
    float z = x * x;
    float y = 7.0376836292e-2f * x;
    y = y - 1.1514610310e-1f; y = y * x;
    y = y + 1.1676998740e-1f; y = y * x;
    y = y - 1.2420140846e-1f; y = y * x;
    y = y + 1.4249322787e-1f; y = y * x;
    y = y - 1.6668057665e-1f; y = y * x;
    y = y + 2.0000714765e-1f; y = y * x;
    y = y - 2.4999993993e-1f; y = y * x;
    y = y + 3.3333331174e-1f; y = y * x;
    y = y * z;
    float fe = (float)e;
    y = y + -2.12194440e-4f * fe;
    y = y + -0.5f * z;
    z = x + y;
    z = z + 0.693359375f * fe;
    return z;
}

MP_HD float mp_atanf(float xx)
{
    float sign = 1.f;
    float x = xx;
    if (x < 0.f) { sign = -1.f; x = -x; }
    float y;
    if (x > 2.414213562373095f) {            // tan(3pi/8)
        y = 1.5707963267948966192f;
        x = -(1.0f / x);
    } else if (x > 0.4142135623730950f) {    // tan(pi/8)
        y = 0.7853981633974483096f;
        x = (x - 1.0f) / (x + 1.0f);
    } else {
        y = 0.f;
    }
    float z = x * x;
    float p = 8.05374449538e-2f * z;
    p = p - 1.38776856032e-1f; p = p * z;
    p = p + 1.99777106478e-1f; p = p * z;
    p = p - 3.33329491539e-1f; p = p * z;
    p = p * x;
    p = p + x;
    y = y + p;
    return sign < 0.f ? -y : y;
}

// atan2f(y, x), C argument order.
MP_HD float mp_atan2f(float y, float x)
{
    constexpr float PIF = 3.141592653589793238f;
    constexpr float PIO2F = 1.5707963267948966192f;
    if (x != x || y != y) return u2f(0x7FC00000u);
    int code = 0;
    if (x < 0.f) code = 2;
    if (y < 0.f) code |= 1;
    if (x == 0.f) {
        if (code & 1) return -PIO2F;
        if (y == 0.f) return 0.f;
        return PIO2F;
    }
    if (y == 0.f) {
        if (code & 2) return PIF;
        return 0.f;
    }
    float w;
    switch (code) {
    default:
    case 0:
    case 1: w = 0.f; break;
    case 2: w = PIF; break;
    case 3: w = -PIF; break;
    }
    return w + mp_atanf(y / x);
}

MP_HD float mp_asinf(float xx)
{
    float sign = 1.f;
    float a = xx;
    if (a < 0.f) { sign = -1.f; a = -a; }
    if (a > 1.0f || a != a) return u2f(0x7FC00000u);
    if (a < 1.0e-4f) return xx;
    float x, z;
    bool flag;
    if (a > 0.5f) {
        z = 0.5f * (1.0f - a);
        x = sqrtf(z);
        flag = true;
    } else {
        x = a;
        z = x * x;
        flag = false;
    }
    float p = 4.2163199048e-2f * z;
    p = p + 2.4181311049e-2f; p = p * z;
    p = p + 4.5470025998e-2f; p = p * z;
    p = p + 7.4953002686e-2f; p = p * z;
    p = p + 1.6666752422e-1f; p = p * z;
    p = p * x;
    p = p + x;
    if (flag) {
        p = p + p;
        p = 1.5707963267948966192f - p;
    }
    return sign < 0.f ? -p : p;
}

// ------------------------------------------------------------------- Vector3
MP_HD Vec3 v3(float x, float y, float z) { Vec3 r; r.x = x; r.y = y; r.z = z; return r; }
MP_HD Vec3 operator+(Vec3 a, Vec3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
MP_HD Vec3 operator-(Vec3 a, Vec3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
MP_HD Vec3 operator-(Vec3 a) { return v3(-a.x, -a.y, -a.z); }
MP_HD Vec3 operator*(Vec3 a, float b) { return v3(a.x * b, a.y * b, a.z * b); }
MP_HD Vec3 operator*(float b, Vec3 a) { return v3(a.x * b, a.y * b, a.z * b); }
// brt/math.inl:421-426: division multiplies by the reciprocal.
MP_HD Vec3 operator/(Vec3 a, float b) { float inv = 1.f / b; return a * inv; }
MP_HD float dot(Vec3 a, Vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
MP_HD Vec3 cross(Vec3 a, Vec3 b)
{
    return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
MP_HD float length2(Vec3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }
MP_HD float length(Vec3 a) { return sqrtf(length2(a)); }
// CPU branch of Vector3::invLength (brt/math.inl:279-287): 1 / sqrt, not rsqrt.
MP_HD float invLength(Vec3 a) { return 1.f / length(a); }
MP_HD Vec3 normalize(Vec3 a) { return a * invLength(a); }
MP_HD float distance(Vec3 a, Vec3 b) { return length(a - b); }
MP_HD float comp(Vec3 v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : v.z); }

MP_HD bool aabbContains(AABB b, Vec3 p)
{
    return b.pMin.x <= p.x && b.pMin.y <= p.y && b.pMin.z <= p.z &&
           b.pMax.x >= p.x && b.pMax.y >= p.y && b.pMax.z >= p.z;
}

// ---------------------------------------------------------------------- Quat
MP_HD Quat quatInv(Quat q) { Quat r; r.w = q.w; r.x = -q.x; r.y = -q.y; r.z = -q.z; return r; }
MP_HD Quat quatNormalize(Quat q)
{
    float inv_length = 1.f / sqrtf(q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z);
    Quat r; r.w = q.w * inv_length; r.x = q.x * inv_length;
    r.y = q.y * inv_length; r.z = q.z * inv_length; return r;
}
MP_HD Quat quatMul(Quat a, Quat b)
{
    Quat r;
    r.w = (a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z);
    r.x = (a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y);
    r.y = (a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x);
    r.z = (a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w);
    return r;
}
MP_HD Vec3 rotateVec(Quat q, Vec3 v)
{
    Vec3 pure = v3(q.x, q.y, q.z);
    float scalar = q.w;
    Vec3 pure_x_v = cross(pure, v);
    Vec3 pure_x_pure_x_v = cross(pure, pure_x_v);
    return v + 2.f * ((pure_x_v * scalar) + pure_x_pure_x_v);
}
MP_HD Quat angleAxis(float angle, Vec3 n)
{
    float coshalf = mp_cosf(angle / 2.f);
    float sinhalf = mp_sinf(angle / 2.f);
    Quat r; r.w = coshalf; r.x = n.x * sinhalf; r.y = n.y * sinhalf; r.z = n.z * sinhalf;
    return r;
}

MP_HD Vec3 kUp() { return v3(0.f, 0.f, 1.f); }
MP_HD Vec3 kFwd() { return v3(0.f, 1.f, 0.f); }
MP_HD Vec3 kRight() { return v3(1.f, 0.f, 0.f); }

// ----------------------------------------------------------------------- RNG
// Threefry2x32-20 key splitting, the JAX scheme (reference src/brt/rand.cpp:24-96).
MP_HD uint32_t rotl32(uint32_t v, uint32_t d) { return (v << d) | (v >> (32 - d)); }

MP_HD RandKey randSplit(RandKey src, uint32_t idx, uint32_t idx_upper = 0)
{
    uint32_t ks0 = src.a, ks1 = src.b, ks2 = 0x1BD11BDAu ^ src.a ^ src.b;
    uint32_t x0 = idx + ks0, x1 = idx_upper + ks1;
#define MP_TF_ROUND(r) { x0 += x1; x1 = rotl32(x1, r); x1 ^= x0; }
    MP_TF_ROUND(13) MP_TF_ROUND(15) MP_TF_ROUND(26) MP_TF_ROUND(6)
    x0 += ks1; x1 += ks2 + 1u;
    MP_TF_ROUND(17) MP_TF_ROUND(29) MP_TF_ROUND(16) MP_TF_ROUND(24)
    x0 += ks2; x1 += ks0 + 2u;
    MP_TF_ROUND(13) MP_TF_ROUND(15) MP_TF_ROUND(26) MP_TF_ROUND(6)
    x0 += ks0; x1 += ks1 + 3u;
    MP_TF_ROUND(17) MP_TF_ROUND(29) MP_TF_ROUND(16) MP_TF_ROUND(24)
    x0 += ks1; x1 += ks2 + 4u;
    MP_TF_ROUND(13) MP_TF_ROUND(15) MP_TF_ROUND(26) MP_TF_ROUND(6)
#undef MP_TF_ROUND
    RandKey out; out.a = x0 + ks2; out.b = x1 + ks0 + 5u;
    return out;
}
MP_HD RandKey randInitKey(uint32_t seed, uint32_t seed_upper = 0)
{
    RandKey k; k.a = seed; k.b = seed_upper;
    return randSplit(k, 0);
}
MP_HD uint32_t randBits32(RandKey k) { return k.a ^ k.b; }
// brt/rand.inl:87-98: 24 random mantissa bits in [0, 1).
MP_HD float randUniformFromKey(RandKey k) { return (float)(randBits32(k) >> 8) * 0x1p-24f; }
// Lemire's unbiased bounded sampler (brt/rand.cpp:98-140); the rejection loop
// re-splits the sample key and never touches the RNG counter.
MP_HD int32_t randI32FromKey(RandKey k, int32_t a, int32_t b)
{
    uint32_t s = (uint32_t)(b - a);
    uint32_t x = randBits32(k);
    uint64_t m = (uint64_t)x * (uint64_t)s;
    uint32_t l = (uint32_t)m, h = (uint32_t)(m >> 32);
    if (l < s) {
        uint32_t t = (0u - s) % s;
        while (l < t) {
            k = randSplit(k, 0);
            x = randBits32(k);
            m = (uint64_t)x * (uint64_t)s;
            l = (uint32_t)m; h = (uint32_t)(m >> 32);
        }
    }
    return (int32_t)h + a;
}
MP_HD RNG rngInit(RandKey k) { RNG r; r.k = k; r.count = 0; return r; }
MP_HD RandKey rngAdvance(RNG &r) { RandKey k = randSplit(r.k, r.count); r.count += 1; return k; }
MP_HD float rngUniform(RNG &r) { return randUniformFromKey(rngAdvance(r)); }
MP_HD int32_t rngI32(RNG &r, int32_t a, int32_t b) { return randI32FromKey(rngAdvance(r), a, b); }

}  // namespace mpenv
