// Shared device helpers for the b200k kernels (sm_100a only).
//
// Every kernel in this library is HBM-bound: the helpers here exist to (1) move 16 bytes per
// thread per instruction with streaming cache hints, (2) convert between the storage type and
// fp32 with round-to-nearest-even exactly where the reference's trace rounds, and (3) reduce
// across a warp / CTA with shuffles and one shared-memory hop.
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200k.h"

namespace b200k {

constexpr int kWarp = 32;

// ---------------------------------------------------------------------------------------------
// 16-byte global memory access with cache hints.
//   ld_stream : read-once data (activations).  Non-coherent path, do not allocate in L1.
//   ld_keep   : small re-used vectors (weights, cos/sin).  Default caching.
//   st_stream : write-once outputs.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 ld_stream(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
// ld_stream_rw: like ld_stream for buffers the same kernel also writes (in-place updates): the non-coherent
// (.nc) path is only defined for data that stays read-only for the whole kernel.
__device__ __forceinline__ uint4 ld_stream_rw(const void* p) {
    uint4 r;
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p)
                 : "memory");
    return r;
}
__device__ __forceinline__ uint4 ld_keep(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }
__device__ __forceinline__ void st_stream(void* p, const uint4& v) {
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}

// ---------------------------------------------------------------------------------------------
// Storage-type traits: a 16-byte vector holds kElems elements of T.
// unpack(): storage -> fp32 (exact).  pack(): fp32 -> storage (round-to-nearest-even).
// rnd(): round an fp32 value through the storage type and back (a trace-level rounding point).
// ---------------------------------------------------------------------------------------------
template <typename T>
struct Vec;

template <>
struct Vec<float> {
    static constexpr int kElems = 4;
    static constexpr int kDtype = B200K_F32;
    __device__ static __forceinline__ void unpack(const uint4& v, float* f) {
        f[0] = __uint_as_float(v.x);
        f[1] = __uint_as_float(v.y);
        f[2] = __uint_as_float(v.z);
        f[3] = __uint_as_float(v.w);
    }
    __device__ static __forceinline__ uint4 pack(const float* f) {
        return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]), __float_as_uint(f[3]));
    }
    __device__ static __forceinline__ float rnd(float x) { return x; }
    __device__ static __forceinline__ void rnd2(float&, float&) {}
    // float2 views (kElems / 2 pairs) for the packed fp32x2 arithmetic of sm_100 (FMUL2 / FADD2 / FFMA2)
    __device__ static __forceinline__ void unpack2(const uint4& v, float2* f) {
        f[0] = make_float2(__uint_as_float(v.x), __uint_as_float(v.y));
        f[1] = make_float2(__uint_as_float(v.z), __uint_as_float(v.w));
    }
    __device__ static __forceinline__ uint4 pack2(const float2* f) {
        return make_uint4(__float_as_uint(f[0].x), __float_as_uint(f[0].y), __float_as_uint(f[1].x), __float_as_uint(f[1].y));
    }
    __device__ static __forceinline__ uint32_t pack_word(float2 x) { return __float_as_uint(x.x); }  // never executed
    __device__ static __forceinline__ float2 rnd2f(float2 x) { return x; }
    __device__ static __forceinline__ float load1(const void* p, int64_t i) { return reinterpret_cast<const float*>(p)[i]; }
    __device__ static __forceinline__ void store1(void* p, int64_t i, float x) { reinterpret_cast<float*>(p)[i] = x; }
};

template <>
struct Vec<__nv_bfloat16> {
    static constexpr int kElems = 8;
    static constexpr int kDtype = B200K_BF16;
    __device__ static __forceinline__ void unpack(const uint4& v, float* f) {
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            f[2 * i] = __uint_as_float(w[i] << 16);
            f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
        }
    }
    __device__ static __forceinline__ uint4 pack(const float* f) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            __nv_bfloat162 p = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
            w[i] = *reinterpret_cast<uint32_t*>(&p);
        }
        return make_uint4(w[0], w[1], w[2], w[3]);
    }
    __device__ static __forceinline__ float rnd(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }
    // Round a pair through bf16 with one packed conversion (F2FP, full rate) instead of two F2F.
    __device__ static __forceinline__ void rnd2(float& a, float& b) {
        __nv_bfloat162 p = __floats2bfloat162_rn(a, b);
        const uint32_t u = *reinterpret_cast<uint32_t*>(&p);
        a = __uint_as_float(u << 16);
        b = __uint_as_float(u & 0xffff0000u);
    }
    __device__ static __forceinline__ float2 word2(uint32_t w) { return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u)); }
    __device__ static __forceinline__ void unpack2(const uint4& v, float2* f) {
        f[0] = word2(v.x), f[1] = word2(v.y), f[2] = word2(v.z), f[3] = word2(v.w);
    }
    __device__ static __forceinline__ uint32_t pack_word(float2 x) {
        __nv_bfloat162 p = __floats2bfloat162_rn(x.x, x.y);
        return *reinterpret_cast<uint32_t*>(&p);
    }
    __device__ static __forceinline__ uint4 pack2(const float2* f) {
        return make_uint4(pack_word(f[0]), pack_word(f[1]), pack_word(f[2]), pack_word(f[3]));
    }
    __device__ static __forceinline__ float2 rnd2f(float2 x) { return word2(pack_word(x)); }
    __device__ static __forceinline__ float load1(const void* p, int64_t i) {
        return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[i]);
    }
    __device__ static __forceinline__ void store1(void* p, int64_t i, float x) {
        reinterpret_cast<__nv_bfloat16*>(p)[i] = __float2bfloat16_rn(x);
    }
};

template <>
struct Vec<__half> {
    static constexpr int kElems = 8;
    static constexpr int kDtype = B200K_F16;
    __device__ static __forceinline__ void unpack(const uint4& v, float* f) {
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float2 p = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
            f[2 * i] = p.x;
            f[2 * i + 1] = p.y;
        }
    }
    __device__ static __forceinline__ uint4 pack(const float* f) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            __half2 p = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
            w[i] = *reinterpret_cast<uint32_t*>(&p);
        }
        return make_uint4(w[0], w[1], w[2], w[3]);
    }
    __device__ static __forceinline__ float rnd(float x) { return __half2float(__float2half_rn(x)); }
    __device__ static __forceinline__ void rnd2(float& a, float& b) {
        const float2 r = __half22float2(__floats2half2_rn(a, b));
        a = r.x;
        b = r.y;
    }
    __device__ static __forceinline__ void unpack2(const uint4& v, float2* f) {
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) f[i] = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
    }
    __device__ static __forceinline__ uint32_t pack_word(float2 x) {
        __half2 p = __floats2half2_rn(x.x, x.y);
        return *reinterpret_cast<uint32_t*>(&p);
    }
    __device__ static __forceinline__ uint4 pack2(const float2* f) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            __half2 p = __floats2half2_rn(f[i].x, f[i].y);
            w[i] = *reinterpret_cast<uint32_t*>(&p);
        }
        return make_uint4(w[0], w[1], w[2], w[3]);
    }
    __device__ static __forceinline__ float2 rnd2f(float2 x) { return __half22float2(__floats2half2_rn(x.x, x.y)); }
    __device__ static __forceinline__ float load1(const void* p, int64_t i) {
        return __half2float(reinterpret_cast<const __half*>(p)[i]);
    }
    __device__ static __forceinline__ void store1(void* p, int64_t i, float x) {
        reinterpret_cast<__half*>(p)[i] = __float2half_rn(x);
    }
};

// ---------------------------------------------------------------------------------------------
// Packed 2 x 16-bit arithmetic on the 32-bit words of a vector (HMUL2 / HADD2, .BF16 or .F16).
// The reference's trace computes every 16-bit op as: upcast both operands to fp32, operate, round to the
// storage type.  When both operands are storage-type VALUES the fp32 product is exact (8 + 8 or 11 + 11
// significand bits) and the fp32 sum is exact wherever the smaller operand can still affect the rounded
// result, so the single rounding of the packed instruction gives the same bits — with no unpack / repack.
// The _rn intrinsics are required: plain __hmul2 + __hadd2 may be contracted into one HFMA2, which skips the
// product's rounding.
// ---------------------------------------------------------------------------------------------
template <typename T>
struct Pk;
template <>
struct Pk<__nv_bfloat16> {
    using P = __nv_bfloat162;
    __device__ static __forceinline__ P as(uint32_t w) { return *reinterpret_cast<const P*>(&w); }
    __device__ static __forceinline__ uint32_t u(P v) { return *reinterpret_cast<const uint32_t*>(&v); }
    __device__ static __forceinline__ uint32_t mul(uint32_t a, uint32_t b) { return u(__hmul2_rn(as(a), as(b))); }
    __device__ static __forceinline__ uint32_t add(uint32_t a, uint32_t b) { return u(__hadd2_rn(as(a), as(b))); }
    __device__ static __forceinline__ uint32_t sub(uint32_t a, uint32_t b) { return u(__hsub2_rn(as(a), as(b))); }
};
template <>
struct Pk<__half> {
    using P = __half2;
    __device__ static __forceinline__ P as(uint32_t w) { return *reinterpret_cast<const P*>(&w); }
    __device__ static __forceinline__ uint32_t u(P v) { return *reinterpret_cast<const uint32_t*>(&v); }
    __device__ static __forceinline__ uint32_t mul(uint32_t a, uint32_t b) { return u(__hmul2_rn(as(a), as(b))); }
    __device__ static __forceinline__ uint32_t add(uint32_t a, uint32_t b) { return u(__hadd2_rn(as(a), as(b))); }
    __device__ static __forceinline__ uint32_t sub(uint32_t a, uint32_t b) { return u(__hsub2_rn(as(a), as(b))); }
};
template <>
struct Pk<float> {  // never executed: fp32 vectors take the scalar paths
    __device__ static __forceinline__ uint32_t mul(uint32_t a, uint32_t) { return a; }
    __device__ static __forceinline__ uint32_t add(uint32_t a, uint32_t) { return a; }
    __device__ static __forceinline__ uint32_t sub(uint32_t a, uint32_t) { return a; }
};
// r = op(a * b, c * d) word by word
template <typename T, bool kSub>
__device__ __forceinline__ uint4 pk_mul_mul_add(const uint4& a, const uint4& b, const uint4& c, const uint4& d) {
    using K = Pk<T>;
    auto one = [](uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
        const uint32_t x = K::mul(a, b), y = K::mul(c, d);
        return kSub ? K::sub(x, y) : K::add(x, y);
    };
    return make_uint4(one(a.x, b.x, c.x, d.x), one(a.y, b.y, c.y, d.y), one(a.z, b.z, c.z, d.z), one(a.w, b.w, c.w, d.w));
}

// ---------------------------------------------------------------------------------------------
// Reductions.
// ---------------------------------------------------------------------------------------------
struct SumOp {
    __device__ __forceinline__ float operator()(float a, float b) const { return a + b; }
};
struct MaxOp {
    __device__ __forceinline__ float operator()(float a, float b) const { return fmaxf(a, b); }
};

template <int kLanes = 32, typename Op>
__device__ __forceinline__ float warp_reduce(float v, Op op) {
#pragma unroll
    for (int o = kLanes / 2; o > 0; o >>= 1) v = op(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Reduce kVals independent values across a CTA of kThreads threads in one shared-memory hop.
// `smem` must hold kVals * (kThreads / 32) floats and must not be in use by another reduction
// that has not passed its trailing __syncthreads (callers alternate two buffers instead).
// Every thread gets the result.
template <int kThreads, int kVals, typename Op>
__device__ __forceinline__ void block_reduce(float (&v)[kVals], float* smem, Op op, float identity) {
    constexpr int kWarps = kThreads / kWarp;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < kVals; ++i) v[i] = warp_reduce(v[i], op);
    if constexpr (kWarps > 1) {
        if (lane == 0) {
#pragma unroll
            for (int i = 0; i < kVals; ++i) smem[i * kWarps + warp] = v[i];
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < kVals; ++i) {
            float x = lane < kWarps ? smem[i * kWarps + lane] : identity;
            v[i] = warp_reduce<(kWarps > 1 ? kWarps : 2)>(x, op);
            v[i] = __shfl_sync(0xffffffffu, v[i], 0);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// mbarrier + bulk asynchronous copy (TMA, 1-D): global -> shared row staging.
// One elected thread arms the barrier with the byte count and issues the copies; consumers wait
// on the barrier's phase parity.  Addresses and sizes must be multiples of 16 bytes.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// L2 eviction policy for read-once streams.
__device__ __forceinline__ uint64_t policy_evict_first() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
            smem_u32(smem_dst)),
        "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
        : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ uint4 lds128(const void* p) { return *reinterpret_cast<const uint4*>(p); }

// ---------------------------------------------------------------------------------------------
// Host-side helpers.
// ---------------------------------------------------------------------------------------------
inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

inline int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

inline int launch_status() {
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : static_cast<int>(e);
}

}  // namespace b200k
