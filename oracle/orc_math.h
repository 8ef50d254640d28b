// TEST INFRASTRUCTURE -- NOT PRODUCT CODE.
// CPU restatement ("oracle") of the reference's arithmetic for the render pass of
// GuidedVolPathVSPGIntegrator.  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load this.  Built with -ffp-contract=off so the
// source-level operation order of the reference is what is evaluated ("Hard parts").
//
// This header: scalar/vector helpers, PCG32, MurmurHash, FastExp, sampling routines.
// Each function cites the reference file:line it follows (paths relative to /root/reference/src).
#pragma once
#include "../include/vspg_libm.h"   // the libm the path is defined with (portable: same bits as the CUDA kernels; oracle/portm_shim.cpp gives it to the reference)
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>

namespace orc {

static constexpr float Pi = 3.14159265358979323846f;
static constexpr float InvPi = 0.31830988618379067154f;
static constexpr float Inv4Pi = 0.07957747154594766788f;
static constexpr float PiOver2 = 1.57079632679489661923f;
static constexpr float PiOver4 = 0.78539816339744830961f;
static constexpr float Infinity = std::numeric_limits<float>::infinity();
static constexpr float MachineEpsilon = std::numeric_limits<float>::epsilon() * 0.5f;  // pbrt/util/float.h:28
static constexpr float OneMinusEpsilon = 0x1.fffffep-1f;                                // pbrt/util/float.h:62
static constexpr float ShadowEpsilon = 0.0001f;                                         // pbrt/util/math.h:42

inline constexpr float gamma_(int n) { return (n * MachineEpsilon) / (1 - n * MachineEpsilon); }  // float.h:195
inline float Sqr(float v) { return v * v; }
inline float SafeSqrt(float x) { return std::sqrt(std::max(0.f, x)); }
inline float SafeASin(float x) { return vspg_asinf(std::min(1.f, std::max(-1.f, x))); }
inline float Lerp(float x, float a, float b) { return (1 - x) * a + x * b; }  // math.h:210
template <typename T, typename U, typename V>
inline T Clamp(T val, U low, V high) {  // math.h:137
    if (val < low) return T(low);
    else if (val > high) return T(high);
    else return val;
}
inline uint32_t FloatToBits(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
inline float BitsToFloat(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }
inline float NextFloatUp(float v) {  // float.h:160
    if (std::isinf(v) && v > 0.f) return v;
    if (v == -0.f) v = 0.f;
    uint32_t ui = FloatToBits(v);
    if (v >= 0) ++ui; else --ui;
    return BitsToFloat(ui);
}
inline float NextFloatDown(float v) {  // float.h:175
    if (std::isinf(v) && v < 0.f) return v;
    if (v == 0.f) v = -0.f;
    uint32_t ui = FloatToBits(v);
    if (v > 0) --ui; else ++ui;
    return BitsToFloat(ui);
}
inline int Exponent(float v) { return (int)(FloatToBits(v) >> 23) - 127; }  // float.h:131
inline float FMA(float a, float b, float c) { return std::fma(a, b, c); }
inline float DifferenceOfProducts(float a, float b, float c, float d) {  // math.h:569
    float cd = c * d;
    float dop = FMA(a, b, -cd);
    float err = FMA(-c, d, cd);
    return dop + err;
}

// pbrt/util/math.h:450-475 (CPU branch: cubic polynomial, NOT expf)
inline float FastExp(float x) {
    float xp = x * 1.442695041f;
    float fxp = std::floor(xp), f = xp - fxp;
    int i = (int)fxp;
    // EvaluatePolynomial(f, 1, c1, c2, c3) = FMA(f, FMA(f, FMA(f, c3, c2), c1), 1)   math.h:330-337
    float twoToF = FMA(f, FMA(f, FMA(f, 0.0781455737f, 0.226173572f), 0.695556856f), 1.f);
    int exponent = Exponent(twoToF) + i;
    if (exponent < -126) return 0;
    if (exponent > 127) return Infinity;
    uint32_t bits = FloatToBits(twoToF);
    bits &= 0b10000000011111111111111111111111u;
    bits |= (uint32_t)(exponent + 127) << 23;
    return BitsToFloat(bits);
}

struct V3 {
    float x = 0, y = 0, z = 0;
    V3() = default;
    V3(float x, float y, float z) : x(x), y(y), z(z) {}
    explicit V3(const float *p) : x(p[0]), y(p[1]), z(p[2]) {}
    float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
    float &operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
};
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }
inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 operator*(float s, V3 a) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 operator/(V3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline bool operator==(V3 a, V3 b) { return a.x == b.x && a.y == b.y && a.z == b.z; }
inline bool operator!=(V3 a, V3 b) { return !(a == b); }
inline float Dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }  // vecmath.h:964
inline float AbsDot(V3 a, V3 b) { return std::abs(Dot(a, b)); }
// Any Dot with a Normal3 operand is a different function in the reference (vecmath.h:1056-1077):
// FMA(n.x, v.x, SumOfProducts(n.y, v.y, n.z, v.z)), with SumOfProducts the error-compensated form of math.h:577-582.
inline float SumOfProducts(float a, float b, float c, float d) {
    float cd = c * d;
    float sop = FMA(a, b, cd);
    float err = FMA(c, d, -cd);
    return sop + err;
}
inline float DotN(V3 a, V3 b) { return FMA(a.x, b.x, SumOfProducts(a.y, b.y, a.z, b.z)); }
inline float AbsDotN(V3 a, V3 b) { return std::abs(DotN(a, b)); }
inline V3 Cross(V3 v, V3 w) {  // vecmath.h:999
    return {DifferenceOfProducts(v.y, w.z, v.z, w.y), DifferenceOfProducts(v.z, w.x, v.x, w.z),
            DifferenceOfProducts(v.x, w.y, v.y, w.x)};
}
inline float LengthSquared(V3 v) { return Sqr(v.x) + Sqr(v.y) + Sqr(v.z); }
inline float Length(V3 v) { return std::sqrt(LengthSquared(v)); }
inline V3 Normalize(V3 v) { return v / Length(v); }
inline V3 Abs(V3 v) { return {std::abs(v.x), std::abs(v.y), std::abs(v.z)}; }
inline float MaxComponentValue(V3 v) { return std::max({v.x, v.y, v.z}); }
inline V3 LerpV(float t, V3 a, V3 b) { return (1 - t) * a + t * b; }  // vecmath.h Lerp(Float, Tuple3, Tuple3)
inline float DistanceSquared(V3 a, V3 b) { return LengthSquared(a - b); }
inline float Distance(V3 a, V3 b) { return Length(a - b); }
inline V3 FaceForward(V3 n, V3 v) { return (Dot(n, v) < 0.f) ? -n : n; }
inline void CoordinateSystem(V3 v1, V3 *v2, V3 *v3) {  // vecmath.h:1007
    float sign = std::copysign(1.f, v1.z);
    float a = -1 / (sign + v1.z);
    float b = v1.x * v1.y * a;
    *v2 = V3(1 + sign * Sqr(v1.x) * a, sign * b, -sign * v1.x);
    *v3 = V3(b, sign + Sqr(v1.y) * a, -v1.y);
}
struct Frame {  // vecmath.h Frame
    V3 x, y, z;
    static Frame FromZ(V3 z) { Frame f; f.z = z; CoordinateSystem(z, &f.x, &f.y); return f; }
    static Frame FromXZ(V3 x, V3 z) { Frame f; f.x = x; f.y = Cross(z, x); f.z = z; return f; }
    V3 ToLocal(V3 v) const { return {Dot(v, x), Dot(v, y), Dot(v, z)}; }
    V3 FromLocal(V3 v) const { return v.x * x + v.y * y + v.z * z; }
};
inline V3 SphericalDirection(float sinTheta, float cosTheta, float phi) {  // vecmath.h
    return {Clamp(sinTheta, -1, 1) * vspg_cosf(phi), Clamp(sinTheta, -1, 1) * vspg_sinf(phi), Clamp(cosTheta, -1, 1)};
}

// RGB "SampledSpectrum" (NSpectrumSamples = 3 in the reference's RGB build, util/spectrum.h:40-44)
struct S3 {
    float v[3] = {0, 0, 0};
    S3() = default;
    explicit S3(float c) { v[0] = v[1] = v[2] = c; }
    S3(float a, float b, float c) { v[0] = a; v[1] = b; v[2] = c; }
    explicit S3(const float *p) { v[0] = p[0]; v[1] = p[1]; v[2] = p[2]; }
    float operator[](int i) const { return v[i]; }
    float &operator[](int i) { return v[i]; }
    explicit operator bool() const { return v[0] != 0 || v[1] != 0 || v[2] != 0; }
    float Average() const { float s = v[0]; s += v[1]; s += v[2]; return s / 3; }  // spectrum.h Average
    float MaxComponentValue() const { return std::max({v[0], v[1], v[2]}); }
    bool HasNaNs() const { return std::isnan(v[0]) || std::isnan(v[1]) || std::isnan(v[2]); }
    bool HasInf() const { return std::isinf(v[0]) || std::isinf(v[1]) || std::isinf(v[2]); }
};
inline S3 operator*(S3 a, S3 b) { return {a[0] * b[0], a[1] * b[1], a[2] * b[2]}; }
inline S3 operator/(S3 a, S3 b) { return {a[0] / b[0], a[1] / b[1], a[2] / b[2]}; }
inline S3 operator+(S3 a, S3 b) { return {a[0] + b[0], a[1] + b[1], a[2] + b[2]}; }
inline S3 operator-(S3 a, S3 b) { return {a[0] - b[0], a[1] - b[1], a[2] - b[2]}; }
inline S3 operator*(S3 a, float s) { return {a[0] * s, a[1] * s, a[2] * s}; }
inline S3 operator*(float s, S3 a) { return {a[0] * s, a[1] * s, a[2] * s}; }
inline S3 operator/(S3 a, float s) { return {a[0] / s, a[1] / s, a[2] / s}; }
inline S3 operator-(S3 a) { return {-a[0], -a[1], -a[2]}; }
inline S3 &operator*=(S3 &a, S3 b) { a = a * b; return a; }
inline S3 &operator*=(S3 &a, float s) { a = a * s; return a; }
inline S3 &operator/=(S3 &a, float s) { a = a / s; return a; }
inline S3 &operator+=(S3 &a, S3 b) { a = a + b; return a; }
inline S3 ClampZero(S3 a) { return {std::max(0.f, a[0]), std::max(0.f, a[1]), std::max(0.f, a[2])}; }
inline S3 FastExp(S3 a) { return {FastExp(a[0]), FastExp(a[1]), FastExp(a[2])}; }
// spectrum.h y(lambda): in the RGB build IsInf(L.y(lambda)) is used as an "any inf" guard; the RGB
// build's y() is the Rec.709-style luminance of the triple -- inf/NaN propagate the same way.
inline bool YIsInf(S3 a) { return std::isinf(0.2126f * a[0] + 0.7152f * a[1] + 0.0722f * a[2]); }

// util/hash.h:19-63
inline uint64_t MurmurHash64A(const unsigned char *key, size_t len, uint64_t seed) {
    const uint64_t m = 0xc6a4a7935bd1e995ull;
    const int r = 47;
    uint64_t h = seed ^ (len * m);
    const unsigned char *end = key + 8 * (len / 8);
    while (key != end) {
        uint64_t k;
        std::memcpy(&k, key, sizeof(uint64_t));
        key += 8;
        k *= m; k ^= k >> r; k *= m;
        h ^= k; h *= m;
    }
    switch (len & 7) {
    case 7: h ^= uint64_t(key[6]) << 48; [[fallthrough]];
    case 6: h ^= uint64_t(key[5]) << 40; [[fallthrough]];
    case 5: h ^= uint64_t(key[4]) << 32; [[fallthrough]];
    case 4: h ^= uint64_t(key[3]) << 24; [[fallthrough]];
    case 3: h ^= uint64_t(key[2]) << 16; [[fallthrough]];
    case 2: h ^= uint64_t(key[1]) << 8; [[fallthrough]];
    case 1: h ^= uint64_t(key[0]); h *= m;
    };
    h ^= h >> r; h *= m; h ^= h >> r;
    return h;
}
inline uint64_t MixBits(uint64_t v) {  // hash.h:69-76
    v ^= (v >> 31); v *= 0x7fb5d329728ea185ull;
    v ^= (v >> 27); v *= 0x81dadef4bc2dd44dull;
    v ^= (v >> 33);
    return v;
}
inline uint64_t HashFloat1(float f) { return MurmurHash64A((const unsigned char *)&f, 4, 0); }        // Hash(Float)
inline uint64_t HashV3(V3 v) { float b[3] = {v.x, v.y, v.z}; return MurmurHash64A((const unsigned char *)b, 12, 0); }
inline uint64_t HashPixelSeed(int px, int py, int seed) {  // Hash(Point2i, int)  samplers.h:456
    int b[3] = {px, py, seed};
    return MurmurHash64A((const unsigned char *)b, 12, 0);
}

// util/rng.h:29-149 PCG32
struct RNG {
    uint64_t state = 0x853c49e6748fea9bULL, inc = 0xda3e39cb94b95bdbULL;
    static constexpr uint64_t MULT = 0x5851f42d4c957f2dULL;
    RNG() = default;
    RNG(uint64_t seq, uint64_t off) { SetSequence(seq, off); }
    void SetSequence(uint64_t seq, uint64_t seed) {
        state = 0u; inc = (seq << 1u) | 1u;
        U32(); state += seed; U32();
    }
    void SetSequence(uint64_t seq) { SetSequence(seq, MixBits(seq)); }
    uint32_t U32() {
        uint64_t old = state;
        state = old * MULT + inc;
        uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u);
        uint32_t rot = (uint32_t)(old >> 59u);
        return (xs >> rot) | (xs << ((~rot + 1u) & 31));
    }
    float Uniform() { return std::min<float>(OneMinusEpsilon, U32() * 0x1p-32f); }
    void Advance(int64_t idelta) {
        uint64_t curMult = MULT, curPlus = inc, accMult = 1u, accPlus = 0u, delta = (uint64_t)idelta;
        while (delta > 0) {
            if (delta & 1) { accMult *= curMult; accPlus = accPlus * curMult + curPlus; }
            curPlus = (curMult + 1) * curPlus;
            curMult *= curMult;
            delta /= 2;
        }
        state = accMult * state + accPlus;
    }
};

// IndependentSampler (samplers.h:442-476)
struct Sampler {
    RNG rng;
    void StartPixelSample(int px, int py, int seed, int sampleIndex, int dim = 0) {
        rng.SetSequence(HashPixelSeed(px, py, seed));
        rng.Advance(sampleIndex * 65536ull + dim);
    }
    float Get1D() { return rng.Uniform(); }
    void Get2D(float *u) { u[0] = rng.Uniform(); u[1] = rng.Uniform(); }
};

inline float SampleExponential(float u, float a) { return -vspg_logf(1 - u) / a; }  // sampling.h:222
inline float HenyeyGreenstein(float cosTheta, float g) {                             // scattering.h:50
    g = Clamp(g, -.99, .99);
    float denom = 1 + Sqr(g) + 2 * g * cosTheta;
    return Inv4Pi * (1 - Sqr(g)) / (denom * SafeSqrt(denom));
}
inline V3 SampleHenyeyGreenstein(V3 wo, float g, const float *u, float *pdf) {  // sampling.cpp:348
    g = Clamp(g, -.99, .99);
    float cosTheta;
    if (std::abs(g) < 1e-3f) cosTheta = 1 - 2 * u[0];
    else cosTheta = -1 / (2 * g) * (1 + Sqr(g) - Sqr((1 - Sqr(g)) / (1 + g - 2 * g * u[0])));
    float sinTheta = SafeSqrt(1 - Sqr(cosTheta));
    float phi = 2 * Pi * u[1];
    Frame wFrame = Frame::FromZ(wo);
    V3 wi = wFrame.FromLocal(SphericalDirection(sinTheta, cosTheta, phi));
    if (pdf) *pdf = HenyeyGreenstein(cosTheta, g);
    return wi;
}
inline void SampleUniformDiskConcentric(const float *u, float *d) {  // sampling.h
    float ox = 2 * u[0] - 1, oy = 2 * u[1] - 1;
    if (ox == 0 && oy == 0) { d[0] = d[1] = 0; return; }
    float theta, r;
    if (std::abs(ox) > std::abs(oy)) { r = ox; theta = PiOver4 * (oy / ox); }
    else { r = oy; theta = PiOver2 - PiOver4 * (ox / oy); }
    d[0] = r * vspg_cosf(theta); d[1] = r * vspg_sinf(theta);
}
inline V3 SampleCosineHemisphere(const float *u) {
    float d[2]; SampleUniformDiskConcentric(u, d);
    float z = SafeSqrt(1 - Sqr(d[0]) - Sqr(d[1]));
    return {d[0], d[1], z};
}
inline float CosineHemispherePDF(float cosTheta) { return cosTheta * InvPi; }
inline V3 SampleUniformSphere(const float *u) {
    float z = 1 - 2 * u[0];
    float r = SafeSqrt(1 - Sqr(z));
    float phi = 2 * Pi * u[1];
    return {r * vspg_cosf(phi), r * vspg_sinf(phi), z};
}
// sampling.h:134-160
inline float BilinearPDF(const float *p, const float *w) {
    if (p[0] < 0 || p[0] > 1 || p[1] < 0 || p[1] > 1) return 0;
    if (w[0] + w[1] + w[2] + w[3] == 0) return 1;
    return 4 * ((1 - p[0]) * (1 - p[1]) * w[0] + p[0] * (1 - p[1]) * w[1] + (1 - p[0]) * p[1] * w[2] + p[0] * p[1] * w[3]) /
           (w[0] + w[1] + w[2] + w[3]);
}
inline float SampleLinear(float u, float a, float b) {  // sampling.h:114
    if (u == 0 && a == 0) return 0;
    float x = u * (a + b) / (a + std::sqrt(Lerp(u, Sqr(a), Sqr(b))));
    return std::min(x, OneMinusEpsilon);
}
inline void SampleBilinear(const float *u, const float *w, float *p) {
    p[1] = SampleLinear(u[1], w[0] + w[1], w[2] + w[3]);
    p[0] = SampleLinear(u[0], Lerp(p[1], w[0], w[2]), Lerp(p[1], w[1], w[3]));
}

}  // namespace orc
