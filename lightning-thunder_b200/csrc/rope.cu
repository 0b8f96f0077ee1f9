// QKV split + rotary embedding (+ GQA expand) forward / backward.
//
// Input layout (LitGPT fused qkv): qkv[B, T, G, q_per_kv + 2, hs]  (G = n_query_groups; the last two
// slots of every group are k and v).  Outputs q[B, G*q_per_kv, T, hs], k and v[B, Hkv, T, hs] with
// Hkv = G*q_per_kv when `expand` (each k/v head repeated q_per_kv times, as the reference's
// expand+reshape materialises it) else G.  RoPE (rotate-half form) acts on the first rope_n features.
//
// One work item = one (b, t, input head, p) with p a vector position in the first half of the rotary
// span: it loads vectors p and p + half (the rotation partners), so every element is read once and
// written once; items past the rotary span copy one vector.  Consecutive threads take consecutive p,
// so each warp reads and writes 128-byte runs.  kW = elements per vector (8/4 for 16-byte access, 1
// for the scalar fallback when sizes or pointers are not 16-byte friendly).
#include "common.cuh"

namespace b200k {
namespace {

struct RopeShape {
    int64_t B, T;
    int G, qpk, hs, rope_n, expand;
};

template <typename T, int kW>
__device__ __forceinline__ void ldv(const T* p, int64_t elem, float* f, bool stream) {
    using V = Vec<T>;
    if constexpr (kW == 1) {
        f[0] = V::load1(p, elem);
    } else {
        const uint4 v = stream ? ld_stream(p + elem) : ld_keep(p + elem);
        V::unpack(v, f);
    }
}
template <typename T, int kW>
__device__ __forceinline__ void stv(T* p, int64_t elem, const float* f) {
    using V = Vec<T>;
    if constexpr (kW == 1) {
        V::store1(p, elem, f[0]);
    } else {
        st_stream(p + elem, V::pack(f));
    }
}
template <typename T, int kW>
__device__ __forceinline__ void rndv(float* f) {
    using V = Vec<T>;
    if constexpr (kW == 1) {
        f[0] = V::rnd(f[0]);
    } else {
#pragma unroll
        for (int e = 0; e < kW; e += 2) V::rnd2(f[e], f[e + 1]);
    }
}

struct Item {
    int64_t bt;
    int g, slot, p;
};
__device__ __forceinline__ Item decode(int64_t item, int items_per_head, int slots, int G) {
    Item it;
    it.p = static_cast<int>(item % items_per_head);
    int64_t r = item / items_per_head;
    it.slot = static_cast<int>(r % slots);
    r /= slots;
    it.g = static_cast<int>(r % G);
    it.bt = r / G;
    return it;
}

// (__fmul_rn keeps the products from being contracted into FMAs, so fp32 is bit-identical to mul-then-add.)
// roped(lo, hi) for partner vectors lo = x[p], hi = x[p + half]:
//   out_lo = rnd(rnd(lo * cos_lo) + rnd(-hi * sin_lo)),  out_hi = rnd(rnd(hi * cos_hi) + rnd(lo * sin_hi))
template <typename T, int kW>
__device__ __forceinline__ void rope_pair(const float* lo, const float* hi, const float* clo, const float* chi, const float* slo,
                                          const float* shi, float* olo, float* ohi) {
    float a[kW], b[kW];
#pragma unroll
    for (int e = 0; e < kW; ++e) {
        a[e] = __fmul_rn(lo[e], clo[e]);
        b[e] = __fmul_rn(-hi[e], slo[e]);
    }
    rndv<T, kW>(a);
    rndv<T, kW>(b);
#pragma unroll
    for (int e = 0; e < kW; ++e) olo[e] = a[e] + b[e];
#pragma unroll
    for (int e = 0; e < kW; ++e) {
        a[e] = __fmul_rn(hi[e], chi[e]);
        b[e] = __fmul_rn(lo[e], shi[e]);
    }
    rndv<T, kW>(a);
    rndv<T, kW>(b);
#pragma unroll
    for (int e = 0; e < kW; ++e) ohi[e] = a[e] + b[e];
}

// vjp of rope_pair w.r.t. (lo, hi) given cotangents (glo, ghi), rounded where the reference rounds:
//   d_lo = rnd(rnd(cos_lo * glo) + rnd(sin_hi * ghi)),  d_hi = rnd(rnd(cos_hi * ghi) - rnd(sin_lo * glo))
template <typename T, int kW>
__device__ __forceinline__ void rope_pair_bwd(const float* glo, const float* ghi, const float* clo, const float* chi,
                                              const float* slo, const float* shi, float* dlo, float* dhi) {
    float a[kW], b[kW];
#pragma unroll
    for (int e = 0; e < kW; ++e) {
        a[e] = __fmul_rn(clo[e], glo[e]);
        b[e] = __fmul_rn(shi[e], ghi[e]);
    }
    rndv<T, kW>(a);
    rndv<T, kW>(b);
#pragma unroll
    for (int e = 0; e < kW; ++e) dlo[e] = a[e] + b[e];
#pragma unroll
    for (int e = 0; e < kW; ++e) {
        a[e] = __fmul_rn(chi[e], ghi[e]);
        b[e] = __fmul_rn(slo[e], glo[e]);
    }
    rndv<T, kW>(a);
    rndv<T, kW>(b);
#pragma unroll
    for (int e = 0; e < kW; ++e) dhi[e] = a[e] - b[e];
}

template <typename T, int kW>
__global__ void __launch_bounds__(256)
rope_fwd_kernel(const T* __restrict__ qkv, const T* __restrict__ cosp, const T* __restrict__ sinp, T* __restrict__ q,
                T* __restrict__ k, T* __restrict__ v, RopeShape s, int64_t nitems) {
    const int half = s.rope_n / 2 / kW;            // partner distance in vectors
    const int tail = (s.hs - s.rope_n) / kW;       // pass-through vectors per head
    const int iph = half + tail;
    const int slots = s.qpk + 2;
    const int hkv = s.expand ? s.G * s.qpk : s.G;
    for (int64_t item = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; item < nitems;
         item += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const Item it = decode(item, iph, slots, s.G);
        const int64_t b = it.bt / s.T, t = it.bt % s.T;
        const T* src = qkv + ((it.bt * s.G + it.g) * slots + it.slot) * s.hs;
        const bool is_q = it.slot < s.qpk, is_k = it.slot == s.qpk;
        const bool paired = it.p < half;
        const int e0 = paired ? it.p * kW : (s.rope_n + (it.p - half) * kW);
        const int e1 = e0 + half * kW;
        float lo[kW], hi[kW] = {}, olo[kW], ohi[kW];
        ldv<T, kW>(src, e0, lo, true);
        if (paired) ldv<T, kW>(src, e1, hi, true);
        if (paired && (is_q || is_k)) {
            float clo[kW], chi[kW], slo[kW], shi[kW];
            ldv<T, kW>(cosp, t * s.rope_n + e0, clo, false);
            ldv<T, kW>(cosp, t * s.rope_n + e1, chi, false);
            ldv<T, kW>(sinp, t * s.rope_n + e0, slo, false);
            ldv<T, kW>(sinp, t * s.rope_n + e1, shi, false);
            rope_pair<T, kW>(lo, hi, clo, chi, slo, shi, olo, ohi);
        } else {
#pragma unroll
            for (int e = 0; e < kW; ++e) {
                olo[e] = lo[e];
                ohi[e] = hi[e];
            }
        }
        if (is_q) {
            T* dst = q + ((b * (s.G * s.qpk) + it.g * s.qpk + it.slot) * s.T + t) * s.hs;
            stv<T, kW>(dst, e0, olo);
            if (paired) stv<T, kW>(dst, e1, ohi);
        } else {
            T* base = is_k ? k : v;
            const int reps = s.expand ? s.qpk : 1;
            for (int r = 0; r < reps; ++r) {
                const int head = s.expand ? it.g * s.qpk + r : it.g;
                T* dst = base + ((b * hkv + head) * s.T + t) * s.hs;
                stv<T, kW>(dst, e0, olo);
                if (paired) stv<T, kW>(dst, e1, ohi);
            }
        }
    }
}

template <typename T, int kW>
__global__ void __launch_bounds__(256)
rope_bwd_kernel(const T* __restrict__ dq, const T* __restrict__ dk, const T* __restrict__ dv, const T* __restrict__ cosp,
                const T* __restrict__ sinp, T* __restrict__ dqkv, RopeShape s, int64_t nitems) {
    const int half = s.rope_n / 2 / kW;
    const int tail = (s.hs - s.rope_n) / kW;
    const int iph = half + tail;
    const int slots = s.qpk + 2;
    const int hkv = s.expand ? s.G * s.qpk : s.G;
    for (int64_t item = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; item < nitems;
         item += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const Item it = decode(item, iph, slots, s.G);
        const int64_t b = it.bt / s.T, t = it.bt % s.T;
        T* dst = dqkv + ((it.bt * s.G + it.g) * slots + it.slot) * s.hs;
        const bool is_q = it.slot < s.qpk, is_k = it.slot == s.qpk;
        const bool paired = it.p < half;
        const bool roped = paired && (is_q || is_k);
        const int e0 = paired ? it.p * kW : (s.rope_n + (it.p - half) * kW);
        const int e1 = e0 + half * kW;
        float clo[kW], chi[kW], slo[kW], shi[kW];
        if (roped) {
            ldv<T, kW>(cosp, t * s.rope_n + e0, clo, false);
            ldv<T, kW>(cosp, t * s.rope_n + e1, chi, false);
            ldv<T, kW>(sinp, t * s.rope_n + e0, slo, false);
            ldv<T, kW>(sinp, t * s.rope_n + e1, shi, false);
        }
        float olo[kW], ohi[kW];
        if (is_q) {
            const T* src = dq + ((b * (s.G * s.qpk) + it.g * s.qpk + it.slot) * s.T + t) * s.hs;
            float glo[kW], ghi[kW] = {};
            ldv<T, kW>(src, e0, glo, true);
            if (paired) ldv<T, kW>(src, e1, ghi, true);
            if (roped) {
                rope_pair_bwd<T, kW>(glo, ghi, clo, chi, slo, shi, olo, ohi);
            } else {
#pragma unroll
                for (int e = 0; e < kW; ++e) olo[e] = glo[e], ohi[e] = ghi[e];
            }
        } else {
            // k / v: the expand's vjp sums the q_per_kv copies in fp32 and rounds once
            const T* base = is_k ? dk : dv;
            const int reps = s.expand ? s.qpk : 1;
#pragma unroll
            for (int e = 0; e < kW; ++e) olo[e] = 0.f, ohi[e] = 0.f;
            for (int r = 0; r < reps; ++r) {
                const int head = s.expand ? it.g * s.qpk + r : it.g;
                const T* src = base + ((b * hkv + head) * s.T + t) * s.hs;
                float glo[kW], ghi[kW] = {}, dlo[kW], dhi[kW];
                ldv<T, kW>(src, e0, glo, true);
                if (paired) ldv<T, kW>(src, e1, ghi, true);
                if (roped) {
                    rope_pair_bwd<T, kW>(glo, ghi, clo, chi, slo, shi, dlo, dhi);
                    rndv<T, kW>(dlo);
                    rndv<T, kW>(dhi);
                } else {
#pragma unroll
                    for (int e = 0; e < kW; ++e) dlo[e] = glo[e], dhi[e] = ghi[e];
                }
#pragma unroll
                for (int e = 0; e < kW; ++e) olo[e] += dlo[e], ohi[e] += dhi[e];
            }
        }
        stv<T, kW>(dst, e0, olo);
        if (paired) stv<T, kW>(dst, e1, ohi);
    }
}

bool shape_ok(const RopeShape& s) {
    return s.B >= 0 && s.T >= 0 && s.G > 0 && s.qpk > 0 && s.hs > 0 && s.rope_n >= 0 && s.rope_n <= s.hs && s.rope_n % 2 == 0;
}

template <typename T>
int run(bool fwd, const void* a0, const void* a1, const void* a2, const void* cosp, const void* sinp, void* o0, void* o1, void* o2,
        const RopeShape& s, cudaStream_t st) {
    constexpr int kE = Vec<T>::kElems;
    const bool vec = (s.rope_n / 2) % kE == 0 && s.hs % kE == 0 && aligned16(a0) && aligned16(a1) && aligned16(a2) &&
                     aligned16(cosp) && aligned16(sinp) && aligned16(o0) && aligned16(o1) && aligned16(o2);
    const int w = vec ? kE : 1;
    const int64_t iph = s.rope_n / 2 / w + (s.hs - s.rope_n) / w;
    const int64_t nitems = s.B * s.T * s.G * (s.qpk + 2) * iph;
    if (nitems == 0) return 0;
    const int64_t want = (nitems + 255) / 256;
    const int64_t cap = static_cast<int64_t>(sm_count()) * 32;
    const int grid = static_cast<int>(want < cap ? want : cap);
    if (fwd) {
        if (vec)
            rope_fwd_kernel<T, kE><<<grid, 256, 0, st>>>(static_cast<const T*>(a0), static_cast<const T*>(cosp), static_cast<const T*>(sinp),
                                                         static_cast<T*>(o0), static_cast<T*>(o1), static_cast<T*>(o2), s, nitems);
        else
            rope_fwd_kernel<T, 1><<<grid, 256, 0, st>>>(static_cast<const T*>(a0), static_cast<const T*>(cosp), static_cast<const T*>(sinp),
                                                        static_cast<T*>(o0), static_cast<T*>(o1), static_cast<T*>(o2), s, nitems);
    } else {
        if (vec)
            rope_bwd_kernel<T, kE><<<grid, 256, 0, st>>>(static_cast<const T*>(a0), static_cast<const T*>(a1), static_cast<const T*>(a2),
                                                         static_cast<const T*>(cosp), static_cast<const T*>(sinp), static_cast<T*>(o0), s, nitems);
        else
            rope_bwd_kernel<T, 1><<<grid, 256, 0, st>>>(static_cast<const T*>(a0), static_cast<const T*>(a1), static_cast<const T*>(a2),
                                                        static_cast<const T*>(cosp), static_cast<const T*>(sinp), static_cast<T*>(o0), s, nitems);
    }
    return launch_status();
}

int dispatch(bool fwd, const void* a0, const void* a1, const void* a2, const void* cosp, const void* sinp, void* o0, void* o1, void* o2,
             const RopeShape& s, int dtype, cudaStream_t st) {
    switch (dtype) {
        case B200K_F32: return run<float>(fwd, a0, a1, a2, cosp, sinp, o0, o1, o2, s, st);
        case B200K_BF16: return run<__nv_bfloat16>(fwd, a0, a1, a2, cosp, sinp, o0, o1, o2, s, st);
        case B200K_F16: return run<__half>(fwd, a0, a1, a2, cosp, sinp, o0, o1, o2, s, st);
        default: return B200K_EDTYPE;
    }
}

}  // namespace
}  // namespace b200k

using namespace b200k;

extern "C" int b200k_qkv_rope_fwd(const void* qkv, const void* cos, const void* sin, void* q, void* k, void* v, int64_t B, int64_t T,
                                  int n_query_groups, int q_per_kv, int head_size, int rope_n_elem, int expand_kv, int dtype,
                                  void* stream) {
    RopeShape s{B, T, n_query_groups, q_per_kv, head_size, rope_n_elem, expand_kv ? 1 : 0};
    if (!shape_ok(s)) return B200K_EINVAL;
    if (B * T == 0) return 0;
    if (qkv == nullptr || q == nullptr || k == nullptr || v == nullptr || (rope_n_elem > 0 && (cos == nullptr || sin == nullptr)))
        return B200K_EINVAL;
    return dispatch(true, qkv, qkv, qkv, cos, sin, q, k, v, s, dtype, static_cast<cudaStream_t>(stream));
}

extern "C" int b200k_qkv_rope_bwd(const void* dq, const void* dk, const void* dv, const void* cos, const void* sin, void* dqkv,
                                  int64_t B, int64_t T, int n_query_groups, int q_per_kv, int head_size, int rope_n_elem,
                                  int expand_kv, int dtype, void* stream) {
    RopeShape s{B, T, n_query_groups, q_per_kv, head_size, rope_n_elem, expand_kv ? 1 : 0};
    if (!shape_ok(s)) return B200K_EINVAL;
    if (B * T == 0) return 0;
    if (dq == nullptr || dk == nullptr || dv == nullptr || dqkv == nullptr || (rope_n_elem > 0 && (cos == nullptr || sin == nullptr)))
        return B200K_EINVAL;
    return dispatch(false, dq, dk, dv, cos, sin, dqkv, dqkv, dqkv, s, dtype, static_cast<cudaStream_t>(stream));
}
