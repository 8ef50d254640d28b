// Device-side scalar / vector arithmetic for the VSPG render pass (sm_100a).
// Everything here is compiled with -fmad=false: the reference is evaluated by a plain SSE4.2 CPU
// build, so an a*b+c in the source is two roundings; fused operations appear only where the reference
// itself calls FMA() (util/math.h).  File:line citations are relative to /root/reference/src/pbrt.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/vspg_rsqrtss.h"   // embree::rsqrt as OpenPGL evaluates it on Intel hardware
#include "../../include/vspg_libm.h"   // log / sin / cos / asin / atan2 / pow: the path's portable libm (bit-identical on host and device)

#define VD __device__ __forceinline__

namespace vspg {

constexpr float kPi = 3.14159265358979323846f;
constexpr float kInvPi = 0.31830988618379067154f;
constexpr float kInv4Pi = 0.07957747154594766788f;
constexpr float kPiOver2 = 1.57079632679489661923f;
constexpr float kPiOver4 = 0.78539816339744830961f;
constexpr float kMachineEpsilon = 5.9604644775390625e-8f;  // util/float.h:28  (FLT_EPSILON / 2)
constexpr float kOneMinusEpsilon = 0x1.fffffep-1f;         // util/float.h:62
constexpr float kShadowEpsilon = 0.0001f;                  // util/math.h:42
constexpr float kFltMax = 3.402823466e+38f;

VD float inff() { return __int_as_float(0x7f800000); }
VD constexpr float gammaN(int n) { return (n * kMachineEpsilon) / (1 - n * kMachineEpsilon); }  // util/float.h:195
VD float sqr(float v) { return v * v; }
VD float safe_sqrt(float x) { return sqrtf(fmaxf(0.f, x)); }
VD float lerpf(float x, float a, float b) { return (1 - x) * a + x * b; }  // util/math.h:210
VD float clampf(float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); }
VD int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
VD float next_up(float v) {  // util/float.h:160
    if (isinf(v) && v > 0.f) return v;
    if (v == -0.f) v = 0.f;
    uint32_t ui = __float_as_uint(v);
    if (v >= 0) ++ui; else --ui;
    return __uint_as_float(ui);
}
VD float next_down(float v) {  // util/float.h:175
    if (isinf(v) && v < 0.f) return v;
    if (v == 0.f) v = -0.f;
    uint32_t ui = __float_as_uint(v);
    if (v > 0) --ui; else ++ui;
    return __uint_as_float(ui);
}
// util/math.h:569  DifferenceOfProducts (the reference's explicit FMA use)
VD float diff_of_products(float a, float b, float c, float d) {
    float cd = c * d;
    float dop = fmaf(a, b, -cd);
    float err = fmaf(-c, d, cd);
    return dop + err;
}
// util/math.h:450-475 FastExp, CPU branch: 2^f by a cubic evaluated with FMA (math.h:330-337), exponent patched in
VD float fast_exp(float x) {
    float xp = x * 1.442695041f;
    float fxp = floorf(xp), f = xp - fxp;
    int i = (int)fxp;
    float twoToF = fmaf(f, fmaf(f, fmaf(f, 0.0781455737f, 0.226173572f), 0.695556856f), 1.f);
    int exponent = (int)(__float_as_uint(twoToF) >> 23) - 127 + i;
    if (exponent < -126) return 0.f;
    if (exponent > 127) return inff();
    uint32_t bits = __float_as_uint(twoToF);
    bits &= 0x807fffffu;
    bits |= (uint32_t)(exponent + 127) << 23;
    return __uint_as_float(bits);
}

struct V3 {
    float x, y, z;
    VD V3() : x(0), y(0), z(0) {}
    VD V3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
    VD float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
    VD void set(int i, float v) { if (i == 0) x = v; else if (i == 1) y = v; else z = v; }
};
VD V3 operator+(V3 a, V3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
VD V3 operator-(V3 a, V3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
VD V3 operator-(V3 a) { return V3(-a.x, -a.y, -a.z); }
VD V3 operator*(V3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
VD V3 operator*(float s, V3 a) { return V3(a.x * s, a.y * s, a.z * s); }
VD V3 operator*(V3 a, V3 b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }
VD V3 operator/(V3 a, V3 b) { return V3(a.x / b.x, a.y / b.y, a.z / b.z); }
VD V3 operator/(V3 a, float s) { return V3(a.x / s, a.y / s, a.z / s); }
VD bool operator==(V3 a, V3 b) { return a.x == b.x && a.y == b.y && a.z == b.z; }
VD float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }  // util/vecmath.h:964
VD float absdot(V3 a, V3 b) { return fabsf(dot(a, b)); }
// A Dot with a Normal3 operand is a different function in the reference (util/vecmath.h:1056-1077):
// FMA(n.x, v.x, SumOfProducts(n.y, v.y, n.z, v.z)), SumOfProducts being the compensated form of util/math.h:577-582.
VD float sum_of_products(float a, float b, float c, float d) {
    float cd = c * d;
    float sop = fmaf(a, b, cd);
    float err = fmaf(c, d, -cd);
    return sop + err;
}
VD float dotn(V3 a, V3 b) { return fmaf(a.x, b.x, sum_of_products(a.y, b.y, a.z, b.z)); }
VD float absdotn(V3 a, V3 b) { return fabsf(dotn(a, b)); }
VD V3 cross(V3 v, V3 w) {  // util/vecmath.h:999
    return V3(diff_of_products(v.y, w.z, v.z, w.y), diff_of_products(v.z, w.x, v.x, w.z),
              diff_of_products(v.x, w.y, v.y, w.x));
}
VD float length_sq(V3 v) { return sqr(v.x) + sqr(v.y) + sqr(v.z); }
VD float length(V3 v) { return sqrtf(length_sq(v)); }
VD V3 normalize(V3 v) { return v / length(v); }
VD V3 vabs(V3 v) { return V3(fabsf(v.x), fabsf(v.y), fabsf(v.z)); }
// Point3fi(p, err) (util/vecmath.h Point3fi, util/math.h Interval::FromValueAndError): every coordinate becomes the
// interval [SubRoundDown(v, e), AddRoundUp(v, e)]; p() is its midpoint and Error() half its width -- both differ from
// (p, err) in the last bits, and ray origins spawned from them (hence the shadow trackers' RNG seeds) follow.
VD void point_interval(V3 p, V3 e, V3 *mid, V3 *err) {
    float lo, hi;
    if (e.x == 0) { lo = hi = p.x; } else { lo = next_down(p.x - e.x); hi = next_up(p.x + e.x); }
    mid->x = (lo + hi) / 2; err->x = (hi - lo) / 2;
    if (e.y == 0) { lo = hi = p.y; } else { lo = next_down(p.y - e.y); hi = next_up(p.y + e.y); }
    mid->y = (lo + hi) / 2; err->y = (hi - lo) / 2;
    if (e.z == 0) { lo = hi = p.z; } else { lo = next_down(p.z - e.z); hi = next_up(p.z + e.z); }
    mid->z = (lo + hi) / 2; err->z = (hi - lo) / 2;
}
VD float max_comp(V3 v) { return fmaxf(v.x, fmaxf(v.y, v.z)); }
VD V3 lerpv(float t, V3 a, V3 b) { return (1 - t) * a + t * b; }
VD float dist_sq(V3 a, V3 b) { return length_sq(a - b); }
VD bool nonzero(V3 a) { return a.x != 0 || a.y != 0 || a.z != 0; }               // SampledSpectrum::operator bool
VD float avg3(V3 a) { float s = a.x; s += a.y; s += a.z; return s / 3; }           // SampledSpectrum::Average
VD bool has_nan(V3 a) { return isnan(a.x) || isnan(a.y) || isnan(a.z); }
VD bool y_is_inf(V3 a) { return isinf(0.2126f * a.x + 0.7152f * a.y + 0.0722f * a.z); }
VD V3 clamp_zero(V3 a) { return V3(fmaxf(0.f, a.x), fmaxf(0.f, a.y), fmaxf(0.f, a.z)); }
VD V3 fast_exp3(V3 a) { return V3(fast_exp(a.x), fast_exp(a.y), fast_exp(a.z)); }
VD V3 splat(float v) { return V3(v, v, v); }

// A medium whose sigma_a and sigma_s are equal across R, G and B ("gray": every BASELINE config except the rgbgrid of
// C5) makes every spectral quantity of the reservoir walk -- sigma_maj, T_maj, trRatioEst, beta_rs, r_u_rs, the
// selected candidate's numerator / denominator -- three copies of one float: each channel performs the identical
// IEEE operation sequence.  G1 carries that one float; the GRAY kernel is bit-identical to the V3 kernel on such
// scenes at a third of the arithmetic (16 -> 6 IEEE divisions, 9 -> 3 FastExp per collision) and ~20 fewer registers.
struct G1 {
    float v;
    VD G1() : v(0) {}
    VD explicit G1(float a) : v(a) {}
    VD float operator[](int) const { return v; }
};
VD G1 operator*(G1 a, G1 b) { return G1(a.v * b.v); }
VD G1 operator*(G1 a, float s) { return G1(a.v * s); }
VD G1 operator*(float s, G1 a) { return G1(a.v * s); }
VD G1 operator/(G1 a, G1 b) { return G1(a.v / b.v); }
VD G1 operator/(G1 a, float s) { return G1(a.v / s); }
VD G1 operator+(G1 a, G1 b) { return G1(a.v + b.v); }
VD G1 operator-(G1 a, G1 b) { return G1(a.v - b.v); }
VD G1 operator-(G1 a) { return G1(-a.v); }
VD G1 clamp_zero(G1 a) { return G1(fmaxf(0.f, a.v)); }
VD G1 fast_exp3(G1 a) { return G1(fast_exp(a.v)); }
template <bool GRAY> struct SpecT { typedef V3 type; };
template <> struct SpecT<true> { typedef G1 type; };
template <bool GRAY> VD typename SpecT<GRAY>::type to_spec(V3 v);
template <> VD V3 to_spec<false>(V3 v) { return v; }
template <> VD G1 to_spec<true>(V3 v) { return G1(v.x); }
VD V3 to_v3(V3 v) { return v; }
VD V3 to_v3(G1 g) { return splat(g.v); }
VD float avg3(G1 a) { float s = a.v; s += a.v; s += a.v; return s / 3; }   // SampledSpectrum::Average of three equal channels
VD float max_comp(G1 a) { return a.v; }
VD bool nonzero(G1 a) { return a.v != 0; }


VD void coordinate_system(V3 v1, V3 *v2, V3 *v3) {  // util/vecmath.h:1007
    float sign = copysignf(1.f, v1.z);
    float a = -1 / (sign + v1.z);
    float b = v1.x * v1.y * a;
    *v2 = V3(1 + sign * sqr(v1.x) * a, sign * b, -sign * v1.x);
    *v3 = V3(b, sign + sqr(v1.y) * a, -v1.y);
}
struct Frame {
    V3 x, y, z;
    VD static Frame from_z(V3 z) { Frame f; f.z = z; coordinate_system(z, &f.x, &f.y); return f; }
    VD static Frame from_xz(V3 x, V3 z) { Frame f; f.x = x; f.y = cross(z, x); f.z = z; return f; }
    VD V3 to_local(V3 v) const { return V3(dot(v, x), dot(v, y), dot(v, z)); }
    VD V3 from_local(V3 v) const { return v.x * x + v.y * y + v.z * z; }
};
// util/vecmath.h SphericalDirection with phi = 2*pi*u.  sin/cos are taken as sincospi(2u): the same angle up to the
// float rounding of the product 2*pi*u (<= 4e-7 rad), without dragging libdevice's large-argument reduction into
// every scatter kernel (it was 470 SASS instructions per call site).
VD V3 spherical_direction_2pi(float sinTheta, float cosTheta, float u) {
    float sp, cp; vspg_sincosf(2 * kPi * u, &sp, &cp);   // phi = 2 * Pi * u[1]; cos(phi), sin(phi)
    return V3(clampf(sinTheta, -1, 1) * cp, clampf(sinTheta, -1, 1) * sp, clampf(cosTheta, -1, 1));
}

// ------------------------------------------------------------------ hashing / RNG (integer: bit-exact)
// util/hash.h:19-63 MurmurHash64A specialised to the key lengths the path uses (4 and 12 bytes, seed 0)
VD uint64_t murmur_tail_finish(uint64_t h) {
    const uint64_t m = 0xc6a4a7935bd1e995ull;
    h ^= h >> 47; h *= m; h ^= h >> 47;
    return h;
}
VD uint64_t hash_u32(uint32_t a) {  // Hash(Float): 4 bytes
    const uint64_t m = 0xc6a4a7935bd1e995ull;
    uint64_t h = 0ull ^ (4ull * m);
    h ^= (uint64_t)a;  // bytes 0..3 of the tail switch
    h *= m;
    return murmur_tail_finish(h);
}
VD uint64_t hash_3u32(uint32_t a, uint32_t b, uint32_t c) {  // Hash(Point3f) / Hash(Point2i,int): 12 bytes
    const uint64_t m = 0xc6a4a7935bd1e995ull;
    uint64_t h = 0ull ^ (12ull * m);
    uint64_t k = (uint64_t)a | ((uint64_t)b << 32);
    k *= m; k ^= k >> 47; k *= m;
    h ^= k; h *= m;
    h ^= (uint64_t)c;
    h *= m;
    return murmur_tail_finish(h);
}
VD uint64_t hash_float(float f) { return hash_u32(__float_as_uint(f)); }
VD uint64_t hash_v3(V3 v) { return hash_3u32(__float_as_uint(v.x), __float_as_uint(v.y), __float_as_uint(v.z)); }
VD uint64_t mix_bits(uint64_t v) {  // util/hash.h:69-76
    v ^= (v >> 31); v *= 0x7fb5d329728ea185ull;
    v ^= (v >> 27); v *= 0x81dadef4bc2dd44dull;
    v ^= (v >> 33);
    return v;
}
// util/rng.h:29-149 PCG32
struct Rng {
    uint64_t state, inc;
    static constexpr uint64_t MULT = 0x5851f42d4c957f2dULL;
    VD uint32_t u32() {
        uint64_t old = state;
        state = old * MULT + inc;
        uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u);
        uint32_t rot = (uint32_t)(old >> 59u);
        return (xs >> rot) | (xs << ((~rot + 1u) & 31));
    }
    VD float uniform() { return fminf(kOneMinusEpsilon, u32() * 0x1p-32f); }
    VD void set_sequence(uint64_t seq, uint64_t seed) {
        state = 0u; inc = (seq << 1u) | 1u;
        u32(); state += seed; u32();
    }
    VD void set_sequence(uint64_t seq) { set_sequence(seq, mix_bits(seq)); }
    VD void advance(uint64_t delta) {
        uint64_t curMult = MULT, curPlus = inc, accMult = 1u, accPlus = 0u;
        while (delta > 0) {
            if (delta & 1) { accMult *= curMult; accPlus = accPlus * curMult + curPlus; }
            curPlus = (curMult + 1) * curPlus;
            curMult *= curMult;
            delta /= 2;
        }
        state = accMult * state + accPlus;
    }
};

// ------------------------------------------------------------------ sampling routines
VD float sample_exponential(float u, float a) { return -vspg_logf(1 - u) / a; }  // util/sampling.h:222
VD float henyey_greenstein(float cosTheta, float g) {                        // util/scattering.h:50
    g = clampf(g, -.99f, .99f);
    float denom = 1 + sqr(g) + 2 * g * cosTheta;
    return kInv4Pi * (1 - sqr(g)) / (denom * safe_sqrt(denom));
}
VD V3 sample_henyey_greenstein(V3 wo, float g, float u0, float u1, float *pdf) {  // util/sampling.cpp:348
    g = clampf(g, -.99f, .99f);
    float cosTheta;
    if (fabsf(g) < 1e-3f) cosTheta = 1 - 2 * u0;
    else cosTheta = -1 / (2 * g) * (1 + sqr(g) - sqr((1 - sqr(g)) / (1 + g - 2 * g * u0)));
    float sinTheta = safe_sqrt(1 - sqr(cosTheta));
    Frame wFrame = Frame::from_z(wo);
    V3 wi = wFrame.from_local(spherical_direction_2pi(sinTheta, cosTheta, u1));
    *pdf = henyey_greenstein(cosTheta, g);
    return wi;
}
VD void sample_uniform_disk_concentric(float u0, float u1, float *dx, float *dy) {
    float ox = 2 * u0 - 1, oy = 2 * u1 - 1;
    if (ox == 0 && oy == 0) { *dx = *dy = 0; return; }
    float theta, r;
    if (fabsf(ox) > fabsf(oy)) { r = ox; theta = kPiOver4 * (oy / ox); }
    else { r = oy; theta = kPiOver2 - kPiOver4 * (ox / oy); }
    *dx = r * vspg_cosf(theta); *dy = r * vspg_sinf(theta);
}
VD V3 sample_cosine_hemisphere(float u0, float u1) {
    float dx, dy; sample_uniform_disk_concentric(u0, u1, &dx, &dy);
    float z = safe_sqrt(1 - sqr(dx) - sqr(dy));
    return V3(dx, dy, z);
}
VD float bilinear_pdf(float p0, float p1, const float *w) {  // util/sampling.h:134
    if (p0 < 0 || p0 > 1 || p1 < 0 || p1 > 1) return 0;
    if (w[0] + w[1] + w[2] + w[3] == 0) return 1;
    return 4 * ((1 - p0) * (1 - p1) * w[0] + p0 * (1 - p1) * w[1] + (1 - p0) * p1 * w[2] + p0 * p1 * w[3]) /
           (w[0] + w[1] + w[2] + w[3]);
}
VD float sample_linear(float u, float a, float b) {  // util/sampling.h:114
    if (u == 0 && a == 0) return 0;
    float x = u * (a + b) / (a + sqrtf(lerpf(u, sqr(a), sqr(b))));
    return fminf(x, kOneMinusEpsilon);
}
VD void sample_bilinear(float u0, float u1, const float *w, float *p0, float *p1) {
    *p1 = sample_linear(u1, w[0] + w[1], w[2] + w[3]);
    *p0 = sample_linear(u0, lerpf(*p1, w[0], w[2]), lerpf(*p1, w[1], w[3]));
}

}  // namespace vspg
