// Small vector / frame / PCG toolbox for the device code, templated on the
// arithmetic type R (double = "exact" mode that mirrors lajolla's Real=double,
// float = "fast" mode).  Operation order follows the reference so that the
// exact mode reproduces its results to rounding:
//   vector ops       /root/reference/src/vector.h:150-300 (normalize: :249-257 returns 0 for l<=0)
//   Frame / ONB      /root/reference/src/frame.h:11-56 (Frisvad)
//   PCG32            /root/reference/src/pcg.h:16-68
//   luminance        /root/reference/src/spectrum.h:31-33
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include <math.h>

#define GD_HD __host__ __device__ __forceinline__
#define GD_D __device__ __forceinline__
// large device functions (BSDF dispatch, vertex construction): one out-of-line copy per
// arithmetic type keeps code size, I-cache pressure and compile time down
#ifndef GD_BIG
#define GD_BIG __device__ __noinline__
#endif

namespace gdpt {

template <typename R> struct Consts;
template <> struct Consts<double> {
    static constexpr double PI = 3.14159265358979323846;
    static constexpr double INV_PI = 1.0 / 3.14159265358979323846;
    static constexpr double TWO_PI = 2.0 * 3.14159265358979323846;
};
template <> struct Consts<float> {
    static constexpr float PI = 3.14159265358979323846f;
    static constexpr float INV_PI = (float)(1.0 / 3.14159265358979323846);
    static constexpr float TWO_PI = (float)(2.0 * 3.14159265358979323846);
};

template <typename R> struct V3 {
    R x, y, z;
    V3() = default;
    GD_HD V3(R a, R b, R c) : x(a), y(b), z(c) {}
    template <typename S> GD_HD explicit V3(const V3<S> &o) : x((R)o.x), y((R)o.y), z((R)o.z) {}
    GD_HD R operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
template <typename R> struct V2 {
    R x, y;
    V2() = default;
    GD_HD V2(R a, R b) : x(a), y(b) {}
};

template <typename R> GD_HD V3<R> operator+(const V3<R> &a, const V3<R> &b) { return V3<R>(a.x + b.x, a.y + b.y, a.z + b.z); }
template <typename R> GD_HD V3<R> operator-(const V3<R> &a, const V3<R> &b) { return V3<R>(a.x - b.x, a.y - b.y, a.z - b.z); }
template <typename R> GD_HD V3<R> operator-(const V3<R> &a) { return V3<R>(-a.x, -a.y, -a.z); }
template <typename R> GD_HD V3<R> operator*(const V3<R> &a, const V3<R> &b) { return V3<R>(a.x * b.x, a.y * b.y, a.z * b.z); }
template <typename R> GD_HD V3<R> operator*(const V3<R> &a, R s) { return V3<R>(a.x * s, a.y * s, a.z * s); }
template <typename R> GD_HD V3<R> operator*(R s, const V3<R> &a) { return V3<R>(s * a.x, s * a.y, s * a.z); }
// vector.h divides by computing the reciprocal once: v * (1/s)
template <typename R> GD_HD V3<R> operator/(const V3<R> &a, R s) { R inv = R(1) / s; return V3<R>(a.x * inv, a.y * inv, a.z * inv); }
template <typename R> GD_HD V3<R> operator/(const V3<R> &a, const V3<R> &b) { return V3<R>(a.x / b.x, a.y / b.y, a.z / b.z); }
template <typename R> GD_HD R dot(const V3<R> &a, const V3<R> &b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
template <typename R> GD_HD V3<R> cross(const V3<R> &a, const V3<R> &b) {
    return V3<R>(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
template <typename R> GD_HD R length_squared(const V3<R> &a) { return dot(a, a); }
template <typename R> GD_HD R length(const V3<R> &a) { return sqrt(dot(a, a)); }
template <typename R> GD_HD V3<R> normalize(const V3<R> &a) {
    R l = length(a);
    if (l <= R(0)) return V3<R>(R(0), R(0), R(0));
    return a / l;
}
template <typename R> GD_HD R distance_squared(const V3<R> &a, const V3<R> &b) { return length_squared(a - b); }
template <typename R> GD_HD R distance(const V3<R> &a, const V3<R> &b) { return length(a - b); }
template <typename R> GD_HD R luminance(const V3<R> &s) { return s.x * R(0.212671) + s.y * R(0.715160) + s.z * R(0.072169); }
template <typename R> GD_HD R max3(const V3<R> &s) { return fmax(fmax(s.x, s.y), s.z); }
template <typename R> GD_HD R clampr(R v, R lo, R hi) { return v < lo ? lo : (hi < v ? hi : v); }  // std::clamp semantics
template <typename R> GD_HD R rmax(R a, R b) { return a > b ? a : b; }   // lajolla.h max(): a > b ? a : b
template <typename R> GD_HD R rmin(R a, R b) { return a < b ? a : b; }

template <typename R> struct Frame {
    V3<R> x, y, n;
    Frame() = default;
    GD_HD Frame(const V3<R> &x_, const V3<R> &y_, const V3<R> &n_) : x(x_), y(y_), n(n_) {}
    GD_HD explicit Frame(const V3<R> &n_) : n(n_) {
        if (n_.z < R(-1 + 1e-6)) { x = V3<R>(0, -1, 0); y = V3<R>(-1, 0, 0); }
        else {
            R a = R(1) / (R(1) + n_.z);
            R b = -n_.x * n_.y * a;
            x = V3<R>(R(1) - n_.x * n_.x * a, b, -n_.x);
            y = V3<R>(b, R(1) - n_.y * n_.y * a, -n_.y);
        }
    }
};
template <typename R> GD_HD Frame<R> operator-(const Frame<R> &f) { return Frame<R>(-f.x, -f.y, -f.n); }
template <typename R> GD_HD V3<R> to_local(const Frame<R> &f, const V3<R> &v) { return V3<R>(dot(v, f.x), dot(v, f.y), dot(v, f.n)); }
template <typename R> GD_HD V3<R> to_world(const Frame<R> &f, const V3<R> &v) { return f.x * v.x + f.y * v.y + f.n * v.z; }
// coordinate_system() as a pair (frame.h:11-22)
template <typename R> GD_HD void coordinate_system(const V3<R> &n, V3<R> &a, V3<R> &b) { Frame<R> f(n); a = f.x; b = f.y; }

// ---- PCG32 (pcg.h) --------------------------------------------------------
struct Pcg { uint64_t state, inc; };
GD_HD uint32_t pcg_next(Pcg &r) {
    uint64_t old = r.state;
    r.state = old * 6364136223846793005ULL + (r.inc | 1);
    uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u);
    uint32_t rot = (uint32_t)(old >> 59u);
    return (xs >> rot) | (xs << ((0u - rot) & 31u));
}
GD_HD Pcg pcg_init(uint64_t stream, uint64_t seed) {
    Pcg r; r.state = 0u; r.inc = (stream << 1u) | 1u;
    pcg_next(r); r.state += seed; pcg_next(r);
    return r;
}
template <typename R> GD_HD R pcg_real(Pcg &r);
template <> GD_HD double pcg_real<double>(Pcg &r) {
    uint64_t u = ((uint64_t)pcg_next(r) << 20) | 0x3ff0000000000000ULL;
#ifdef __CUDA_ARCH__
    return __longlong_as_double((long long)u) - 1.0;
#else
    union { uint64_t u; double d; } x; x.u = u; return x.d - 1.0;
#endif
}
// fast mode: the same 32 random bits, rounded from the double the reference would draw
template <> GD_HD float pcg_real<float>(Pcg &r) {
    float f = (float)pcg_real<double>(r);
    return f < 1.0f ? f : 0.99999994f;
}

} // namespace gdpt
