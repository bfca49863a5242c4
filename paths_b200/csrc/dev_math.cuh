// dev_math.cuh -- device-side restatement of the reference's f64 value types (src/vector.rs, src/colour.rs, src/matrix.rs)
// templated on the arithmetic type R: R = double keeps the reference's operation order bit for bit (compile that TU with
// -fmad=false: rustc never contracts a*b+c); R = float is the production instantiation.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace paths {

template <typename R> struct V3 {
    R x, y, z;
};
template <typename R> __host__ __device__ __forceinline__ V3<R> mk3(R x, R y, R z) { return V3<R>{x, y, z}; }
template <typename R> __device__ __forceinline__ V3<R> operator+(V3<R> a, V3<R> b) { return mk3<R>(a.x + b.x, a.y + b.y, a.z + b.z); }
template <typename R> __device__ __forceinline__ V3<R> operator-(V3<R> a, V3<R> b) { return mk3<R>(a.x - b.x, a.y - b.y, a.z - b.z); }
template <typename R> __device__ __forceinline__ V3<R> operator*(V3<R> a, V3<R> b) { return mk3<R>(a.x * b.x, a.y * b.y, a.z * b.z); }
template <typename R> __device__ __forceinline__ V3<R> operator*(V3<R> a, R s) { return mk3<R>(a.x * s, a.y * s, a.z * s); }
template <typename R> __device__ __forceinline__ V3<R> operator/(V3<R> a, R s) { return mk3<R>(a.x / s, a.y / s, a.z / s); }
template <typename R> __device__ __forceinline__ V3<R> neg(V3<R> a) { return mk3<R>(-a.x, -a.y, -a.z); } // v * -1 (exact)
template <typename R> __device__ __forceinline__ R dot(V3<R> a, V3<R> b) { return a.x * b.x + a.y * b.y + a.z * b.z; } // vector.rs:23-25
// vector.rs:43-49
template <typename R> __device__ __forceinline__ V3<R> cross(V3<R> a, V3<R> b) {
    return mk3<R>(a.y * b.z - a.z * b.y, -(a.x * b.z - a.z * b.x), a.x * b.y - a.y * b.x);
}

__device__ __forceinline__ float sqrt_(float x) { return sqrtf(x); }
__device__ __forceinline__ double sqrt_(double x) { return sqrt(x); }
__device__ __forceinline__ float fmin_(float a, float b) { return fminf(a, b); }
__device__ __forceinline__ double fmin_(double a, double b) { return fmin(a, b); }
__device__ __forceinline__ float fmax_(float a, float b) { return fmaxf(a, b); }
__device__ __forceinline__ double fmax_(double a, double b) { return fmax(a, b); }
__device__ __forceinline__ float pow_(float a, float b) { return powf(a, b); }
__device__ __forceinline__ double pow_(double a, double b) { return pow(a, b); }
__device__ __forceinline__ float acos_(float a) { return acosf(a); }
__device__ __forceinline__ double acos_(double a) { return acos(a); }
__device__ __forceinline__ float atan2_(float a, float b) { return atan2f(a, b); }
__device__ __forceinline__ double atan2_(double a, double b) { return atan2(a, b); }
__device__ __forceinline__ float tan_(float a) { return tanf(a); }
__device__ __forceinline__ double tan_(double a) { return tan(a); }
__device__ __forceinline__ float exp_(float a) { return expf(a); }
__device__ __forceinline__ double exp_(double a) { return exp(a); }
__device__ __forceinline__ void sincos_(float a, float *s, float *c) { sincosf(a, s, c); }
__device__ __forceinline__ void sincos_(double a, double *s, double *c) { *s = sin(a); *c = cos(a); }
__device__ __forceinline__ bool isnan_(float a) { return isnan(a); }
__device__ __forceinline__ bool isnan_(double a) { return isnan(a); }

// vector.rs:39-41: normed() divides by sqrt(magnitude()), magnitude() being the SQUARED length (vector.rs:27-29)
template <typename R> __device__ __forceinline__ V3<R> normed(V3<R> a) { return a / sqrt_(dot(a, a)); }
// production (f32) shortcuts: one reciprocal (square root) and three multiplies instead of three IEEE divisions; the f64
// instantiation keeps the reference's operations
__device__ __forceinline__ V3<float> operator/(V3<float> a, float s) {
    float r = 1.0f / s;
    return mk3<float>(a.x * r, a.y * r, a.z * r);
}
template <> __device__ __forceinline__ V3<float> normed<float>(V3<float> a) {
    float r = rsqrtf(a.x * a.x + a.y * a.y + a.z * a.z);
    return mk3<float>(a.x * r, a.y * r, a.z * r);
}
// x^5 (material.rs:304 powf(5.0)) and sin/cos of 2*pi*v (geom.rs:15, geom.rs:150)
__device__ __forceinline__ double pow5_(double x) { return pow(x, 5.0); }
__device__ __forceinline__ float pow5_(float x) {
    float x2 = x * x;
    return x2 * x2 * x;
}
__device__ __forceinline__ void sincos_2pi(double v, double *s, double *c) {
    double theta = 2.0 * 3.14159265358979323846 * v;
    *s = sin(theta);
    *c = cos(theta);
}
__device__ __forceinline__ void sincos_2pi(float v, float *s, float *c) { sincospif(2.0f * v, s, c); }
template <typename R> __device__ __forceinline__ R vmax3(V3<R> a) { return fmax_(a.x, fmax_(a.y, a.z)); } // vector.rs:31-33
template <typename R> __device__ __forceinline__ R vmin3(V3<R> a) { return fmin_(a.x, fmin_(a.y, a.z)); } // vector.rs:35-37
// colour.rs:38-41 (plain comparisons, not fmax)
template <typename R> __device__ __forceinline__ R cmax3(V3<R> c) {
    R w = c.x > c.y ? c.x : c.y;
    return w > c.z ? w : c.z;
}
template <typename R> __device__ __forceinline__ bool col_bad(V3<R> c) { return c.x < R(0) || c.y < R(0) || c.z < R(0); } // colour.rs:56-60

// vector.rs:51-61 + geom.rs:26-28
template <typename R> __device__ __forceinline__ V3<R> to_world(V3<R> v, V3<R> n) {
    V3<R> j = n, i;
    if (fabs(j.x) == R(0)) i = mk3<R>(R(1), R(0), R(0));
    else i = normed(cross(j, mk3<R>(R(0), R(1), R(0))));
    V3<R> k = cross(i, j);
    return i * v.x + j * v.y + k * v.z;
}

// matrix.rs:94-104, m row-major (stored as double, converted on use)
template <typename R, typename S> __device__ __forceinline__ V3<R> mat_apply(const S *m, V3<R> o) {
    return mk3<R>(R(m[0]) * o.x + R(m[1]) * o.y + R(m[2]) * o.z, R(m[3]) * o.x + R(m[4]) * o.y + R(m[5]) * o.z,
                  R(m[6]) * o.x + R(m[7]) * o.y + R(m[8]) * o.z);
}

template <typename R> struct Vec4T;
template <> struct Vec4T<float> {
    typedef float4 type;
};
template <> struct Vec4T<double> {
    typedef double4 type;
};
__device__ __forceinline__ float4 mk4(float x, float y, float z, float w) { return make_float4(x, y, z, w); }
__device__ __forceinline__ double4 mk4(double x, double y, double z, double w) { return make_double4(x, y, z, w); }
__device__ __forceinline__ float bits_to_real(uint32_t b, float) { return __uint_as_float(b); }
__device__ __forceinline__ double bits_to_real(uint32_t b, double) { return __longlong_as_double((long long)b); }
__device__ __forceinline__ uint32_t real_to_bits(float v) { return __float_as_uint(v); }
__device__ __forceinline__ uint32_t real_to_bits(double v) { return (uint32_t)__double_as_longlong(v); }

template <typename R> struct Consts;
template <> struct Consts<double> {
    static __device__ __forceinline__ double pi() { return 3.14159265358979323846; }
};
template <> struct Consts<float> {
    static __device__ __forceinline__ float pi() { return 3.14159265358979323846f; }
};

} // namespace paths
