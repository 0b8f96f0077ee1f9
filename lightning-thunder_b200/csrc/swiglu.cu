// SwiGLU forward / backward: pure streaming elementwise kernels.
//
// Each thread moves kUnroll independent 16-byte vectors per tensor per trip (all loads issued
// before the first use), grid = a few waves of the SM count with a grid-stride loop.  Math is fp32
// with a storage-dtype rounding at every point where the reference's trace holds a tensor of the
// storage dtype (see include/b200k.h).
#include "common.cuh"

namespace b200k {
namespace {

constexpr int kThreads = 256;
constexpr int kUnroll = 4;

// silu(a) = a / (1 + exp(-a)) in fp32, the expression ATen evaluates.  For bf16 storage every result is rounded
// to 8 significant bits right away, so the 2-ulp hardware exp2 / reciprocal forms are used (a flip of that
// rounding needs the fp32 value within 2^-22 of a rounding boundary, and costs one bf16 ulp, well inside the
// bf16 tolerance); fp16 and fp32 storage keep expf and IEEE division.
template <bool kFast>
__device__ __forceinline__ float exp_neg(float a) {
    if constexpr (kFast) {
        float y;
        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(a * -1.4426950408889634f));
        return y;
    } else {
        return expf(-a);
    }
}
template <bool kFast>
__device__ __forceinline__ float recip(float x) {
    if constexpr (kFast) {
        float y;
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
        return y;
    } else {
        return 1.0f / x;
    }
}
template <bool kFast>
__device__ __forceinline__ float silu_f32(float a) {
    if constexpr (kFast) return a * recip<true>(1.0f + exp_neg<true>(a));
    else return a / (1.0f + expf(-a));
}

template <typename T>
__device__ __forceinline__ void swiglu_fwd_vec(const uint4& av, const uint4& bv, uint4& ov) {
    using V = Vec<T>;
    constexpr int kE = V::kElems;
    constexpr bool kFast = V::kDtype == B200K_BF16;
    float a[kE], b[kE], o[kE];
    V::unpack(av, a);
    V::unpack(bv, b);
#pragma unroll
    for (int e = 0; e < kE; e += 2) {
        float u0 = silu_f32<kFast>(a[e]), u1 = silu_f32<kFast>(a[e + 1]);
        V::rnd2(u0, u1);  // silu result is a tensor of the storage dtype
        o[e] = u0 * b[e];
        o[e + 1] = u1 * b[e + 1];
    }
    ov = V::pack(o);
}

template <typename T>
__device__ __forceinline__ void swiglu_bwd_vec(const uint4& av, const uint4& bv, const uint4& gv, uint4& dav, uint4& dbv) {
    using V = Vec<T>;
    constexpr int kE = V::kElems;
    constexpr bool kFast = V::kDtype == B200K_BF16;
    float a[kE], b[kE], g[kE], da[kE], db[kE];
    V::unpack(av, a);
    V::unpack(bv, b);
    V::unpack(gv, g);
#pragma unroll
    for (int e = 0; e < kE; e += 2) {
        float u0 = silu_f32<kFast>(a[e]), u1 = silu_f32<kFast>(a[e + 1]);
        V::rnd2(u0, u1);
        db[e] = u0 * g[e];
        db[e + 1] = u1 * g[e + 1];
        float gu0 = b[e] * g[e], gu1 = b[e + 1] * g[e + 1];
        V::rnd2(gu0, gu1);
        // silu vjp evaluated op by op in the storage dtype (clang.silu = a * sigmoid(a)):
        float t0 = exp_neg<kFast>(a[e]), t1 = exp_neg<kFast>(a[e + 1]);
        V::rnd2(t0, t1);  // exp(-a)
        t0 = 1.0f + t0, t1 = 1.0f + t1;
        V::rnd2(t0, t1);  // 1 + exp(-a)
        float s0 = recip<kFast>(t0), s1 = recip<kFast>(t1);
        V::rnd2(s0, s1);  // sigmoid
        float p0 = gu0 * s0, p1 = gu1 * s1;
        V::rnd2(p0, p1);  // g * sigmoid
        float q0 = 1.0f - s0, q1 = 1.0f - s1;
        V::rnd2(q0, q1);
        q0 = a[e] * q0, q1 = a[e + 1] * q1;
        V::rnd2(q0, q1);  // a * (1 - sigmoid)
        q0 = 1.0f + q0, q1 = 1.0f + q1;
        V::rnd2(q0, q1);
        da[e] = p0 * q0;
        da[e + 1] = p1 * q1;
    }
    dav = V::pack(da);
    dbv = V::pack(db);
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
swiglu_fwd_kernel(const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ out, int64_t nvec) {
    const int64_t stride = static_cast<int64_t>(gridDim.x) * kThreads * kUnroll;
    for (int64_t base = static_cast<int64_t>(blockIdx.x) * kThreads * kUnroll + threadIdx.x; base < nvec; base += stride) {
        uint4 av[kUnroll], bv[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const int64_t v = base + static_cast<int64_t>(u) * kThreads;
            if (v < nvec) {
                av[u] = ld_stream(reinterpret_cast<const uint4*>(a) + v);
                bv[u] = ld_stream(reinterpret_cast<const uint4*>(b) + v);
            }
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const int64_t v = base + static_cast<int64_t>(u) * kThreads;
            if (v < nvec) {
                uint4 ov;
                swiglu_fwd_vec<T>(av[u], bv[u], ov);
                st_stream(reinterpret_cast<uint4*>(out) + v, ov);
            }
        }
    }
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
swiglu_bwd_kernel(const T* __restrict__ a, const T* __restrict__ b, const T* __restrict__ g, T* __restrict__ da,
                  T* __restrict__ db, int64_t nvec) {
    constexpr int kU = 4;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * kThreads * kU;
    for (int64_t base = static_cast<int64_t>(blockIdx.x) * kThreads * kU + threadIdx.x; base < nvec; base += stride) {
        uint4 av[kU], bv[kU], gv[kU];
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            const int64_t v = base + static_cast<int64_t>(u) * kThreads;
            if (v < nvec) {
                av[u] = ld_stream(reinterpret_cast<const uint4*>(a) + v);
                bv[u] = ld_stream(reinterpret_cast<const uint4*>(b) + v);
                gv[u] = ld_stream(reinterpret_cast<const uint4*>(g) + v);
            }
        }
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            const int64_t v = base + static_cast<int64_t>(u) * kThreads;
            if (v < nvec) {
                uint4 dav, dbv;
                swiglu_bwd_vec<T>(av[u], bv[u], gv[u], dav, dbv);
                st_stream(reinterpret_cast<uint4*>(da) + v, dav);
                st_stream(reinterpret_cast<uint4*>(db) + v, dbv);
            }
        }
    }
}

// Scalar tail (n not a multiple of the vector width) and unaligned pointers.
template <typename T>
__global__ void swiglu_fwd_tail_kernel(const T* a, const T* b, T* out, int64_t begin, int64_t n) {
    using V = Vec<T>;
    for (int64_t i = begin + blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const float u = V::rnd(silu_f32<false>(V::load1(a, i)));
        V::store1(out, i, u * V::load1(b, i));
    }
}

template <typename T>
__global__ void swiglu_bwd_tail_kernel(const T* a, const T* b, const T* g, T* da, T* db, int64_t begin, int64_t n) {
    using V = Vec<T>;
    for (int64_t i = begin + blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const float x = V::load1(a, i), y = V::load1(b, i), gg = V::load1(g, i);
        const float u = V::rnd(silu_f32<false>(x));
        V::store1(db, i, u * gg);
        const float gu = V::rnd(y * gg);
        const float t = V::rnd(1.0f + V::rnd(expf(-x)));
        const float s = V::rnd(1.0f / t);
        const float p = V::rnd(gu * s);
        const float q = V::rnd(1.0f + V::rnd(x * V::rnd(1.0f - s)));
        V::store1(da, i, p * q);
    }
}

int grid_for(int64_t nvec, int per_cta) {
    const int64_t want = (nvec + per_cta - 1) / per_cta;
    const int64_t cap = static_cast<int64_t>(sm_count()) * 16;
    return static_cast<int>(want < cap ? (want > 0 ? want : 1) : cap);
}

template <typename T>
int run_fwd(const void* a, const void* b, void* out, int64_t n, cudaStream_t st) {
    constexpr int kE = Vec<T>::kElems;
    const bool vec_ok = aligned16(a) && aligned16(b) && aligned16(out);
    const int64_t nvec = vec_ok ? n / kE : 0;
    if (nvec > 0) {
        swiglu_fwd_kernel<T><<<grid_for(nvec, kThreads * kUnroll), kThreads, 0, st>>>(
            static_cast<const T*>(a), static_cast<const T*>(b), static_cast<T*>(out), nvec);
        if (int rc = launch_status()) return rc;
    }
    if (nvec * kE < n) {
        const int64_t rest = n - nvec * kE;
        swiglu_fwd_tail_kernel<T><<<grid_for(rest, 256), 256, 0, st>>>(static_cast<const T*>(a), static_cast<const T*>(b),
                                                                      static_cast<T*>(out), nvec * kE, n);
        return launch_status();
    }
    return 0;
}

template <typename T>
int run_bwd(const void* a, const void* b, const void* g, void* da, void* db, int64_t n, cudaStream_t st) {
    constexpr int kE = Vec<T>::kElems;
    const bool vec_ok = aligned16(a) && aligned16(b) && aligned16(g) && aligned16(da) && aligned16(db);
    const int64_t nvec = vec_ok ? n / kE : 0;
    if (nvec > 0) {
        swiglu_bwd_kernel<T><<<grid_for(nvec, kThreads * 4), kThreads, 0, st>>>(
            static_cast<const T*>(a), static_cast<const T*>(b), static_cast<const T*>(g), static_cast<T*>(da), static_cast<T*>(db), nvec);
        if (int rc = launch_status()) return rc;
    }
    if (nvec * kE < n) {
        const int64_t rest = n - nvec * kE;
        swiglu_bwd_tail_kernel<T><<<grid_for(rest, 256), 256, 0, st>>>(static_cast<const T*>(a), static_cast<const T*>(b),
                                                                      static_cast<const T*>(g), static_cast<T*>(da),
                                                                      static_cast<T*>(db), nvec * kE, n);
        return launch_status();
    }
    return 0;
}

}  // namespace
}  // namespace b200k

using namespace b200k;

extern "C" int b200k_swiglu_fwd(const void* a, const void* b, void* out, int64_t n, int dtype, void* stream) {
    if (n < 0) return B200K_EINVAL;
    if (n == 0) return 0;
    if (a == nullptr || b == nullptr || out == nullptr) return B200K_EINVAL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    switch (dtype) {
        case B200K_F32: return run_fwd<float>(a, b, out, n, st);
        case B200K_BF16: return run_fwd<__nv_bfloat16>(a, b, out, n, st);
        case B200K_F16: return run_fwd<__half>(a, b, out, n, st);
        default: return B200K_EDTYPE;
    }
}

extern "C" int b200k_swiglu_bwd(const void* a, const void* b, const void* g, void* da, void* db, int64_t n, int dtype,
                                void* stream) {
    if (n < 0) return B200K_EINVAL;
    if (n == 0) return 0;
    if (a == nullptr || b == nullptr || g == nullptr || da == nullptr || db == nullptr) return B200K_EINVAL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    switch (dtype) {
        case B200K_F32: return run_bwd<float>(a, b, g, da, db, n, st);
        case B200K_BF16: return run_bwd<__nv_bfloat16>(a, b, g, da, db, n, st);
        case B200K_F16: return run_bwd<__half>(a, b, g, da, db, n, st);
        default: return B200K_EDTYPE;
    }
}
