// QKV split + rotary embedding (+ GQA expand) forward / backward.
//
// Input layout (LitGPT fused qkv): qkv[B, T, G, q_per_kv + 2, hs]  (G = n_query_groups; the last two
// slots of every group are k and v).  Outputs q[B, G*q_per_kv, T, hs], k and v[B, Hkv, T, hs] with
// Hkv = G*q_per_kv when `expand` (each k/v head repeated q_per_kv times, as the reference's
// expand+reshape materialises it) else G.  RoPE (rotate-half form) acts on the first rope_n features.
//
// A work item = (token, group, p) with p a vector position in the first half of the rotary span: for each head
// slot of the group it loads vectors p and p + half (the rotation partners), so every element is read once and
// written once; items past the rotary span copy one vector per slot.  Consecutive threads take consecutive p,
// so each warp reads and writes 128-byte runs.  kW = elements per vector (8/4 for 16-byte access, 1
// for the scalar fallback when sizes or pointers are not 16-byte friendly).
#include <cstdlib>

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

// Loads kept packed (one uint4 per vector) until they are consumed, so that many can be in flight per thread.
template <int kW>
struct Raw {
    uint4 v;
};
template <>
struct Raw<1> {
    float v;
};
template <typename T, int kW>
__device__ __forceinline__ Raw<kW> ldraw(const T* p, int64_t elem) {
    Raw<kW> r;
    if constexpr (kW == 1) r.v = Vec<T>::load1(p, elem);
    else r.v = ld_stream(p + elem);
    return r;
}
template <typename T, int kW>
__device__ __forceinline__ void unraw(const Raw<kW>& r, float* f) {
    if constexpr (kW == 1) f[0] = r.v;
    else Vec<T>::unpack(r.v, f);
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

// Work item = (token, group g, vector position p): the thread walks the group's q_per_kv + 2 head slots with the
// cos/sin vectors of (t, p) held in registers, so the tables are fetched once per q_per_kv + 2 heads and every
// thread has 2 * (q_per_kv + 2) independent 16-byte activations loads to overlap.
constexpr int kTokTile = 16;  // tokens handled by neighbouring threads in the backward
constexpr int kBwdChunk = 4;  // expanded copies gathered together in the backward
constexpr int kMaxQpk = 6;  // slots whose loads are issued together; larger groups loop in chunks

// kPk (16-bit storage, 16-byte vectors): the rotation runs on packed 2 x 16-bit words (Pk<T> in common.cuh) —
// vectors are never unpacked to fp32, which takes the kernel from 18 to 6 ALU instructions per word and frees
// the registers for more loads in flight; bit-identical to the fp32 form (B200K_ROPE_PACKED=0, compared in the tests).
template <typename T, int kW, bool kPk>
__global__ void __launch_bounds__(128)
rope_fwd_kernel(const T* __restrict__ qkv, const T* __restrict__ cosp, const T* __restrict__ sinp, T* __restrict__ q,
                T* __restrict__ k, T* __restrict__ v, RopeShape s, int64_t nitems) {
    const int half = s.rope_n / 2 / kW;            // partner distance in vectors
    const int tail = (s.hs - s.rope_n) / kW;       // pass-through vectors per head
    const int iph = half + tail;
    const int slots = s.qpk + 2;
    const int hkv = s.expand ? s.G * s.qpk : s.G;
    const int nq = s.G * s.qpk;
    for (int64_t item = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; item < nitems;
         item += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int p = static_cast<int>(item % iph);
        const int64_t r = item / iph;
        const int g = static_cast<int>(r % s.G);
        const int64_t bt = r / s.G;
        const int64_t b = bt / s.T, t = bt % s.T;
        const bool paired = p < half;
        const int e0 = paired ? p * kW : (s.rope_n + (p - half) * kW);
        const int e1 = e0 + half * kW;
        const T* grp = qkv + (bt * s.G + g) * (static_cast<int64_t>(slots) * s.hs);
        float clo[kPk ? 1 : kW], chi[kPk ? 1 : kW], slo[kPk ? 1 : kW], shi[kPk ? 1 : kW];
        uint4 pclo, pchi, pslo, pshi;
        if (paired) {
            if constexpr (kPk) {
                pclo = ld_keep(cosp + t * s.rope_n + e0), pchi = ld_keep(cosp + t * s.rope_n + e1);
                pslo = ld_keep(sinp + t * s.rope_n + e0), pshi = ld_keep(sinp + t * s.rope_n + e1);
            } else {
                ldv<T, kW>(cosp, t * s.rope_n + e0, clo, false);
                ldv<T, kW>(cosp, t * s.rope_n + e1, chi, false);
                ldv<T, kW>(sinp, t * s.rope_n + e0, slo, false);
                ldv<T, kW>(sinp, t * s.rope_n + e1, shi, false);
            }
        }
        for (int s0 = 0; s0 < slots; s0 += kMaxQpk) {
            Raw<kW> rlo[kMaxQpk], rhi[kMaxQpk];
#pragma unroll
            for (int j = 0; j < kMaxQpk; ++j) {
                if (s0 + j < slots) {
                    const T* src = grp + static_cast<int64_t>(s0 + j) * s.hs;
                    rlo[j] = ldraw<T, kW>(src, e0);
                    rhi[j] = ldraw<T, kW>(src, paired ? e1 : e0);
                }
            }
#pragma unroll
            for (int j = 0; j < kMaxQpk; ++j) {
                const int slot = s0 + j;
                if (slot < slots) {
                    const bool is_q = slot < s.qpk, is_k = slot == s.qpk;
                    if constexpr (kPk) {
                        uint4 olo = rlo[j].v, ohi = rhi[j].v;
                        if (paired && (is_q || is_k)) {
                            // out_lo = rnd(rnd(lo * cos_lo) - rnd(hi * sin_lo)),  out_hi = rnd(rnd(hi * cos_hi) + rnd(lo * sin_hi))
                            olo = pk_mul_mul_add<T, true>(rlo[j].v, pclo, rhi[j].v, pslo);
                            ohi = pk_mul_mul_add<T, false>(rhi[j].v, pchi, rlo[j].v, pshi);
                        }
                        if (is_q) {
                            T* dst = q + ((b * nq + g * s.qpk + slot) * s.T + t) * s.hs;
                            st_stream(dst + e0, olo);
                            if (paired) st_stream(dst + e1, ohi);
                        } else {
                            T* base = is_k ? k : v;
                            const int reps = s.expand ? s.qpk : 1;
                            for (int rr = 0; rr < reps; ++rr) {
                                const int head = s.expand ? g * s.qpk + rr : g;
                                T* dst = base + ((b * hkv + head) * s.T + t) * s.hs;
                                st_stream(dst + e0, olo);
                                if (paired) st_stream(dst + e1, ohi);
                            }
                        }
                        continue;
                    }
                    float lo[kW], hi[kW], olo[kW], ohi[kW];
                    unraw<T, kW>(rlo[j], lo);
                    unraw<T, kW>(rhi[j], hi);
                    if (paired && (is_q || is_k)) {
                        rope_pair<T, kW>(lo, hi, clo, chi, slo, shi, olo, ohi);
                    } else {
#pragma unroll
                        for (int e = 0; e < kW; ++e) olo[e] = lo[e], ohi[e] = hi[e];
                    }
                    if (is_q) {
                        T* dst = q + ((b * nq + g * s.qpk + slot) * s.T + t) * s.hs;
                        stv<T, kW>(dst, e0, olo);
                        if (paired) stv<T, kW>(dst, e1, ohi);
                    } else {
                        T* base = is_k ? k : v;
                        const int reps = s.expand ? s.qpk : 1;
                        for (int rr = 0; rr < reps; ++rr) {
                            const int head = s.expand ? g * s.qpk + rr : g;
                            T* dst = base + ((b * hkv + head) * s.T + t) * s.hs;
                            stv<T, kW>(dst, e0, olo);
                            if (paired) stv<T, kW>(dst, e1, ohi);
                        }
                    }
                }
            }
        }
    }
}

// Backward work item = (token, group, pass, p): pass 0 turns the group's q cotangents into dq (one output per
// head), pass 1 / 2 fold the q_per_kv expanded copies of dk / dv into the single k / v slot (fp32 sum, rounded
// once, as the expand's vjp does).  Passes are separate threads so their gathers overlap.
template <typename T, int kW, bool kPk>
__global__ void __launch_bounds__(128, 4)
rope_bwd_kernel(const T* __restrict__ dq, const T* __restrict__ dk, const T* __restrict__ dv, const T* __restrict__ cosp,
                const T* __restrict__ sinp, T* __restrict__ dqkv, RopeShape s, int64_t nitems) {
    const int half = s.rope_n / 2 / kW;
    const int tail = (s.hs - s.rope_n) / kW;
    const int iph = half + tail;
    const int slots = s.qpk + 2;
    const int hkv = s.expand ? s.G * s.qpk : s.G;
    const int nq = s.G * s.qpk;
    for (int64_t item = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; item < nitems;
         item += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        // p fastest, then kTokTile consecutive tokens: the cotangents are [B, heads, T, hs], so neighbouring threads
        // read neighbouring tokens of one head — contiguous kTokTile * hs runs instead of isolated head rows
        const int p = static_cast<int>(item % iph);
        int64_t r = item / iph;
        const int tt = static_cast<int>(r % kTokTile);
        r /= kTokTile;
        const int pass = static_cast<int>(r % 3);
        r /= 3;
        const int g = static_cast<int>(r % s.G);
        const int64_t bt = (r / s.G) * kTokTile + tt;
        if (bt >= s.B * s.T) continue;
        const int64_t b = bt / s.T, t = bt % s.T;
        const bool paired = p < half;
        const int e0 = paired ? p * kW : (s.rope_n + (p - half) * kW);
        const int e1 = e0 + half * kW;
        T* grp = dqkv + (bt * s.G + g) * (static_cast<int64_t>(slots) * s.hs);
        const T* base = pass == 0 ? dq : (pass == 1 ? dk : dv);
        const int heads_total = pass == 0 ? nq : hkv;
        const int nsrc = pass == 0 ? s.qpk : (s.expand ? s.qpk : 1);
        const int head0 = pass == 0 ? g * s.qpk : (s.expand ? g * s.qpk : g);
        const bool roped = paired && pass < 2;
        float clo[kPk ? 1 : kW], chi[kPk ? 1 : kW], slo[kPk ? 1 : kW], shi[kPk ? 1 : kW];
        uint4 pclo, pchi, pslo, pshi;
        if (roped) {
            if constexpr (kPk) {
                pclo = ld_keep(cosp + t * s.rope_n + e0), pchi = ld_keep(cosp + t * s.rope_n + e1);
                pslo = ld_keep(sinp + t * s.rope_n + e0), pshi = ld_keep(sinp + t * s.rope_n + e1);
            } else {
                ldv<T, kW>(cosp, t * s.rope_n + e0, clo, false);
                ldv<T, kW>(cosp, t * s.rope_n + e1, chi, false);
                ldv<T, kW>(sinp, t * s.rope_n + e0, slo, false);
                ldv<T, kW>(sinp, t * s.rope_n + e1, shi, false);
            }
        }
        float acc_lo[kW], acc_hi[kW];
#pragma unroll
        for (int e = 0; e < kW; ++e) acc_lo[e] = 0.f, acc_hi[e] = 0.f;
        for (int s0 = 0; s0 < nsrc; s0 += kBwdChunk) {
            Raw<kW> rlo[kBwdChunk], rhi[kBwdChunk];
#pragma unroll
            for (int j = 0; j < kBwdChunk; ++j) {
                if (s0 + j < nsrc) {
                    const T* src = base + ((b * heads_total + head0 + s0 + j) * s.T + t) * s.hs;
                    rlo[j] = ldraw<T, kW>(src, e0);
                    rhi[j] = ldraw<T, kW>(src, paired ? e1 : e0);
                }
            }
#pragma unroll
            for (int j = 0; j < kBwdChunk; ++j) {
                if (s0 + j < nsrc) {
                    if constexpr (kPk) {
                        uint4 dlo = rlo[j].v, dhi = rhi[j].v;
                        if (roped) {
                            // d_lo = rnd(rnd(cos_lo * glo) + rnd(sin_hi * ghi)),  d_hi = rnd(rnd(cos_hi * ghi) - rnd(sin_lo * glo))
                            dlo = pk_mul_mul_add<T, false>(pclo, rlo[j].v, pshi, rhi[j].v);
                            dhi = pk_mul_mul_add<T, true>(pchi, rhi[j].v, pslo, rlo[j].v);
                        }
                        if (pass == 0) {
                            T* dst = grp + static_cast<int64_t>(s0 + j) * s.hs;
                            st_stream(dst + e0, dlo);
                            if (paired) st_stream(dst + e1, dhi);
                        } else {
                            float fl[kW], fh[kW];
                            Vec<T>::unpack(dlo, fl);
                            Vec<T>::unpack(dhi, fh);
#pragma unroll
                            for (int e = 0; e < kW; ++e) acc_lo[e] += fl[e], acc_hi[e] += fh[e];
                        }
                        continue;
                    }
                    float glo[kW], ghi[kW], dlo[kW], dhi[kW];
                    unraw<T, kW>(rlo[j], glo);
                    unraw<T, kW>(rhi[j], ghi);
                    if (roped) {
                        rope_pair_bwd<T, kW>(glo, ghi, clo, chi, slo, shi, dlo, dhi);
                    } else {
#pragma unroll
                        for (int e = 0; e < kW; ++e) dlo[e] = glo[e], dhi[e] = ghi[e];
                    }
                    if (pass == 0) {
                        T* dst = grp + static_cast<int64_t>(s0 + j) * s.hs;
                        stv<T, kW>(dst, e0, dlo);
                        if (paired) stv<T, kW>(dst, e1, dhi);
                    } else {
                        if (roped) {
                            rndv<T, kW>(dlo);
                            rndv<T, kW>(dhi);
                        }
#pragma unroll
                        for (int e = 0; e < kW; ++e) acc_lo[e] += dlo[e], acc_hi[e] += dhi[e];
                    }
                }
            }
        }
        if (pass > 0) {
            T* dst = grp + static_cast<int64_t>(s.qpk + pass - 1) * s.hs;
            stv<T, kW>(dst, e0, acc_lo);
            if (paired) stv<T, kW>(dst, e1, acc_hi);
        }
    }
}

// B200K_ROPE_PACKED=0 selects the fp32 form for 16-bit storage (bit-for-bit comparison in the tests).
bool packed_enabled() {
    static const bool on = [] {
        const char* e = getenv("B200K_ROPE_PACKED");
        return !(e != nullptr && e[0] == '0');
    }();
    return on;
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
    // forward: (token, group, vector position); backward: (token tile, group, pass, token in tile, vector position)
    const int64_t nitems = fwd ? s.B * s.T * s.G * iph : ((s.B * s.T + kTokTile - 1) / kTokTile) * kTokTile * s.G * iph * 3;
    if (nitems == 0) return 0;
    const int64_t want = (nitems + 127) / 128;
    const int64_t cap = static_cast<int64_t>(sm_count()) * 64;
    const int grid = static_cast<int>(want < cap ? want : cap);
    auto launch = [&](auto kernel_fwd, auto kernel_bwd) {
        if (fwd)
            kernel_fwd<<<grid, 128, 0, st>>>(static_cast<const T*>(a0), static_cast<const T*>(cosp), static_cast<const T*>(sinp),
                                             static_cast<T*>(o0), static_cast<T*>(o1), static_cast<T*>(o2), s, nitems);
        else
            kernel_bwd<<<grid, 128, 0, st>>>(static_cast<const T*>(a0), static_cast<const T*>(a1), static_cast<const T*>(a2),
                                             static_cast<const T*>(cosp), static_cast<const T*>(sinp), static_cast<T*>(o0), s, nitems);
    };
    if (vec && kE == 8 && packed_enabled()) launch(rope_fwd_kernel<T, kE, (kE == 8)>, rope_bwd_kernel<T, kE, (kE == 8)>);
    else if (vec) launch(rope_fwd_kernel<T, kE, false>, rope_bwd_kernel<T, kE, false>);
    else launch(rope_fwd_kernel<T, 1, false>, rope_bwd_kernel<T, 1, false>);
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
