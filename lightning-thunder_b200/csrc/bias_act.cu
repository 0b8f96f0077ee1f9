// bias + activation forward / backward.
//
// x, g, out, dx: [rows, hidden] row-major; bias, dbias: [hidden].  Thread t of a CTA owns the 16-byte
// vectors t, t + blockDim.x, ... of a column tile, the CTA strides over rows, so bias stays in
// registers and the backward can keep fp32 partial column sums of dx per thread (dbias), written as
// one partial row per CTA and folded by colsum_finalize (deterministic, no atomics).
// act: 0 = none (bias add only), 1 = relu, 2 = gelu (erf form).
#include <cstdlib>
#include <type_traits>

#include "common.cuh"
#include "reduce.cuh"

namespace b200k {
namespace {

constexpr int kBaThreads = 416;
constexpr int kBaVpl = 2;
constexpr int kBaRowParts = 4;  // CTAs per SM along rows (sizes the scratch)
constexpr float kInvSqrt2 = 0.70710678118654752440f;
constexpr float kTwoOverSqrtPi = 1.1283791670955126f;

// All activation math runs on float2 pairs (sm_100 packed FMUL2 / FADD2 / FFMA2: half the fp32 instructions of
// this ALU-bound chain, results identical to scalar round-to-nearest ops).
__device__ __forceinline__ float2 f2(float x) { return make_float2(x, x); }
__device__ __forceinline__ float hw_rcp(float x) {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float hw_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// erf for bf16 storage: Abramowitz-Stegun 7.1.26 (|error| <= 1.5e-7, far below half a bf16 ulp of a value in
// [-1, 1]) on the hardware exp2 / reciprocal units; fp16 and fp32 storage keep erff.
template <bool kFast>
__device__ __forceinline__ float2 erf2(float2 x) {
    if constexpr (kFast) {
        const float2 ax = make_float2(fabsf(x.x), fabsf(x.y));
        const float2 d = __ffma2_rn(f2(0.3275911f), ax, f2(1.0f));
        const float2 t = make_float2(hw_rcp(d.x), hw_rcp(d.y));
        const float2 q = __fmul2_rn(__fmul2_rn(ax, ax), f2(-1.4426950408889634f));
        const float2 e = make_float2(hw_ex2(q.x), hw_ex2(q.y));
        // the polynomial carries the minus sign of 1 - p(t) * t * e (negation is exact, so the bits are the same)
        float2 p = __ffma2_rn(f2(-1.061405429f), t, f2(1.453152027f));
        p = __ffma2_rn(p, t, f2(-1.421413741f));
        p = __ffma2_rn(p, t, f2(0.284496736f));
        p = __ffma2_rn(p, t, f2(-0.254829592f));
        const float2 r = __ffma2_rn(__fmul2_rn(p, t), e, f2(1.0f));
        return make_float2(copysignf(r.x, x.x), copysignf(r.y, x.y));
    } else {
        return make_float2(erff(x.x), erff(x.y));
    }
}
// 2/sqrt(pi) * exp(-z^2)
template <bool kFast>
__device__ __forceinline__ float2 gauss2(float2 z) {
    if constexpr (kFast) {
        const float2 q = __fmul2_rn(__fmul2_rn(z, z), f2(-1.4426950408889634f));
        return __fmul2_rn(f2(kTwoOverSqrtPi), make_float2(hw_ex2(q.x), hw_ex2(q.y)));
    } else {
        return make_float2(kTwoOverSqrtPi * expf(-(z.x * z.x)), kTwoOverSqrtPi * expf(-(z.y * z.y)));
    }
}

// gelu(a) with one storage rounding per ltorch op: a * rnd(0.5 + rnd(0.5 * rnd(erf(rnd(a / sqrt2)))))
template <typename T>
__device__ __forceinline__ float2 gelu_cdf2(float2 a, float2& z_out) {
    using V = Vec<T>;
    constexpr bool kFast = V::kDtype == B200K_BF16;
    const float2 z = V::rnd2f(__fmul2_rn(a, f2(kInvSqrt2)));
    z_out = z;
    float2 c = V::rnd2f(erf2<kFast>(z));
    c = V::rnd2f(__fmul2_rn(c, f2(0.5f)));
    return V::rnd2f(__fadd2_rn(c, f2(0.5f)));
}

template <typename T, int kAct>
__device__ __forceinline__ void act_fwd(float* a /* rounded pre-activation, in/out */, int n) {
#pragma unroll
    for (int e = 0; e < n; e += 2) {
        const float2 v = make_float2(a[e], a[e + 1]);
        if constexpr (kAct == 1) {
            a[e] = v.x > 0.f ? v.x : 0.f;
            a[e + 1] = v.y > 0.f ? v.y : 0.f;
        } else if constexpr (kAct == 2) {
            float2 z;
            const float2 o = __fmul2_rn(v, gelu_cdf2<T>(v, z));
            a[e] = o.x, a[e + 1] = o.y;
        }
    }
}

template <typename T, int kAct>
__device__ __forceinline__ void act_bwd(const float* a, const float* g, float* dx, int n) {
    using V = Vec<T>;
    constexpr bool kFast = V::kDtype == B200K_BF16;
#pragma unroll
    for (int e = 0; e < n; e += 2) {
        const float2 av = make_float2(a[e], a[e + 1]), gv = make_float2(g[e], g[e + 1]);
        if constexpr (kAct == 0) {
            dx[e] = gv.x, dx[e + 1] = gv.y;
        } else if constexpr (kAct == 1) {
            dx[e] = av.x > 0.f ? gv.x : 0.f;
            dx[e + 1] = av.y > 0.f ? gv.y : 0.f;
        } else {
            float2 z;
            const float2 cdf = gelu_cdf2<T>(av, z);
            const float2 p = V::rnd2f(__fmul2_rn(cdf, gv));                      // cdf * g
            float2 q = V::rnd2f(__fmul2_rn(av, gv));                             // a * g
            q = V::rnd2f(__fmul2_rn(q, f2(0.5f)));
            float2 r = V::rnd2f(__fmul2_rn(gauss2<kFast>(z), q));                // pdf term
            r = V::rnd2f(__fmul2_rn(r, f2(kInvSqrt2)));
            const float2 o = __fadd2_rn(p, r);
            dx[e] = o.x, dx[e + 1] = o.y;
        }
    }
}

// ---- bf16 storage, GELU: the chain on packed bf16x2 words --------------------------------------------------
// Every ltorch op of the reference's trace takes bf16 tensors, computes in fp32 and rounds the result to bf16.
// When both operands are bf16 VALUES (tensor x tensor, or tensor x a constant bf16 holds exactly, like 0.5) the
// fp32 result is exact for a product (8 + 8 significand bits) and within fp32 for a sum wherever the small
// operand can still move the result, so one rounding of the exact result — what HMUL2.BF16 / HADD2.BF16 do —
// equals the reference's two-step rounding bit for bit.  Those ops then need no unpack / repack at all: one
// instruction per pair instead of five.  Ops with an fp32-only constant (1/sqrt2) and the transcendental parts
// stay in fp32.  The _rn intrinsics matter: plain __hmul2 + __hadd2 may be contracted into one HFMA2, which skips
// the product's rounding.  The test-suite compares this path with the fp32 form bit for bit (B200K_BIAS_ACT_PACKED=0).
__device__ __forceinline__ uint32_t bf2_add(uint32_t a, uint32_t b) {
    const __nv_bfloat162 r = __hadd2_rn(*reinterpret_cast<const __nv_bfloat162*>(&a), *reinterpret_cast<const __nv_bfloat162*>(&b));
    return *reinterpret_cast<const uint32_t*>(&r);
}
__device__ __forceinline__ uint32_t bf2_mul(uint32_t a, uint32_t b) {
    const __nv_bfloat162 r = __hmul2_rn(*reinterpret_cast<const __nv_bfloat162*>(&a), *reinterpret_cast<const __nv_bfloat162*>(&b));
    return *reinterpret_cast<const uint32_t*>(&r);
}
constexpr uint32_t kBf2Half = 0x3F003F00u;  // (0.5, 0.5)

// erf of the bf16 pair in word zw by Abramowitz-Stegun 7.1.26 as in erf2<true>, rounded to a bf16 pair; also hands
// back e = exp(-z^2), which the backward's density term needs again.  |z| and the final copysign are one LOP3 each
// on the packed word (rounding is symmetric in the sign, so copysign-then-round = round-then-copysign).
__device__ __forceinline__ uint32_t erf_word_with_exp(uint32_t zw, float2& e_out) {
    using V = Vec<__nv_bfloat16>;
    const float2 ax = V::word2(zw & 0x7fff7fffu);
    const float2 d = __ffma2_rn(f2(0.3275911f), ax, f2(1.0f));
    const float2 t = make_float2(hw_rcp(d.x), hw_rcp(d.y));
    const float2 q = __fmul2_rn(__fmul2_rn(ax, ax), f2(-1.4426950408889634f));
    const float2 e = make_float2(hw_ex2(q.x), hw_ex2(q.y));
    e_out = e;
    float2 p = __ffma2_rn(f2(-1.061405429f), t, f2(1.453152027f));
    p = __ffma2_rn(p, t, f2(-1.421413741f));
    p = __ffma2_rn(p, t, f2(0.284496736f));
    p = __ffma2_rn(p, t, f2(-0.254829592f));
    const float2 r = __ffma2_rn(__fmul2_rn(p, t), e, f2(1.0f));
    return (V::pack_word(r) & 0x7fff7fffu) | (zw & 0x80008000u);
}

// cdf word = 0.5 + 0.5 * erf(a / sqrt2), each step rounded; e_out = exp(-z^2) of the rounded z
__device__ __forceinline__ uint32_t gelu_cdf_word(uint32_t a2, float2& e_out) {
    using V = Vec<__nv_bfloat16>;
    const uint32_t zw = V::pack_word(__fmul2_rn(V::word2(a2), f2(kInvSqrt2)));
    const uint32_t erf_w = erf_word_with_exp(zw, e_out);
    return bf2_add(bf2_mul(erf_w, kBf2Half), kBf2Half);
}
__device__ __forceinline__ uint32_t gelu_fwd_word(uint32_t a2) {
    float2 e;
    return bf2_mul(a2, gelu_cdf_word(a2, e));
}
__device__ __forceinline__ uint32_t gelu_bwd_word(uint32_t a2, uint32_t g2) {
    using V = Vec<__nv_bfloat16>;
    float2 e;
    const uint32_t cdf = gelu_cdf_word(a2, e);
    const uint32_t p = bf2_mul(cdf, g2);                          // cdf * g
    const uint32_t q = bf2_mul(bf2_mul(a2, g2), kBf2Half);        // 0.5 * (a * g)
    const float2 dens = __fmul2_rn(f2(kTwoOverSqrtPi), e);        // 2/sqrt(pi) * exp(-z^2), an fp32 tensor in the trace
    const uint32_t r = V::pack_word(__fmul2_rn(dens, V::word2(q)));
    const uint32_t r2 = V::pack_word(__fmul2_rn(V::word2(r), f2(kInvSqrt2)));
    return bf2_add(p, r2);
}

// kAccum: keep fp32 column sums of the (rounded) backward result for dbias; kStore: write `out`;
// pre_out (forward, nullable): also write the rounded pre-activation a = x + bias, which is what autodiff saves.
template <typename T, int kAct, bool kHasBias, bool kBwd, bool kAccum, bool kStore = true, bool kPacked = false>
__global__ void __launch_bounds__(kBaThreads, 2)
bias_act_kernel(const T* __restrict__ x, const T* __restrict__ bias, const T* __restrict__ g, T* __restrict__ out,
                T* __restrict__ pre_out, float* __restrict__ partial, int64_t rows, int nvec) {
    using V = Vec<T>;
    constexpr int kE = V::kElems;
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int v0 = blockIdx.y * (nthr * kBaVpl) + tid;
    float bf[kBaVpl][kE], acc[kBaVpl][kE];
    uint4 braw[kBaVpl];  // the bias vectors as loaded (the packed path adds them word by word)
#pragma unroll
    for (int i = 0; i < kBaVpl; ++i) braw[i] = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
    for (int i = 0; i < kBaVpl; ++i) {
        const int v = v0 + i * nthr;
#pragma unroll
        for (int e = 0; e < kE; ++e) bf[i][e] = 0.f, acc[i][e] = 0.f;
        if (kHasBias && v < nvec) {
            braw[i] = ld_keep(reinterpret_cast<const uint4*>(bias) + v);
            V::unpack(braw[i], bf[i]);
        }
    }
    // software pipeline: the next row's vectors are requested before the current row's ALU chain starts, so
    // every thread always has loads in flight
    uint4 xv[kBaVpl], gv[kBaVpl], xn[kBaVpl], gn[kBaVpl];
    auto fetch = [&](int64_t row, uint4* xd, uint4* gd) {
#pragma unroll
        for (int i = 0; i < kBaVpl; ++i) {
            const int v = v0 + i * nthr;
            if (row < rows && v < nvec) {
                xd[i] = ld_stream(reinterpret_cast<const uint4*>(x) + row * nvec + v);
                if constexpr (kBwd) gd[i] = ld_stream(reinterpret_cast<const uint4*>(g) + row * nvec + v);
            }
        }
    };
    fetch(blockIdx.x, xv, gv);
    for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
        fetch(row + gridDim.x, xn, gn);
#pragma unroll
        for (int i = 0; i < kBaVpl; ++i) {
            const int v = v0 + i * nthr;
            if constexpr (kPacked) {
                if (v < nvec) {
                    const uint32_t xw[4] = {xv[i].x, xv[i].y, xv[i].z, xv[i].w};
                    const uint32_t gw[4] = {gv[i].x, gv[i].y, gv[i].z, gv[i].w};
                    const uint32_t bw[4] = {braw[i].x, braw[i].y, braw[i].z, braw[i].w};
                    uint32_t aw[4], ow[4];
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        aw[k] = kHasBias ? bf2_add(xw[k], bw[k]) : xw[k];
                        if constexpr (kBwd) {
                            ow[k] = gelu_bwd_word(aw[k], gw[k]);
                            if constexpr (kAccum) {
                                const float2 sum = __fadd2_rn(make_float2(acc[i][2 * k], acc[i][2 * k + 1]), V::word2(ow[k]));
                                acc[i][2 * k] = sum.x;
                                acc[i][2 * k + 1] = sum.y;
                            }
                        } else {
                            ow[k] = gelu_fwd_word(aw[k]);
                        }
                    }
                    if constexpr (!kBwd) {
                        if (pre_out != nullptr)
                            st_stream(reinterpret_cast<uint4*>(pre_out) + row * nvec + v, make_uint4(aw[0], aw[1], aw[2], aw[3]));
                    }
                    if constexpr (kStore) st_stream(reinterpret_cast<uint4*>(out) + row * nvec + v, make_uint4(ow[0], ow[1], ow[2], ow[3]));
                }
            } else if (v < nvec) {
                float a[kE], o[kE];
                V::unpack(xv[i], a);
                if constexpr (kHasBias) {
#pragma unroll
                    for (int e = 0; e < kE; e += 2) {
                        a[e] += bf[i][e];
                        a[e + 1] += bf[i][e + 1];
                        V::rnd2(a[e], a[e + 1]);
                    }
                }
                if constexpr (kBwd) {
                    float gf[kE];
                    V::unpack(gv[i], gf);
                    act_bwd<T, kAct>(a, gf, o, kE);
                    if constexpr (kAccum) {
#pragma unroll
                        for (int e = 0; e < kE; e += 2) {
                            float d0 = o[e], d1 = o[e + 1];
                            V::rnd2(d0, d1);
                            acc[i][e] += d0;
                            acc[i][e + 1] += d1;
                        }
                    }
                } else {
                    if (pre_out != nullptr) st_stream(reinterpret_cast<uint4*>(pre_out) + row * nvec + v, V::pack(a));
                    act_fwd<T, kAct>(a, kE);
#pragma unroll
                    for (int e = 0; e < kE; ++e) o[e] = a[e];
                }
                if constexpr (kStore) st_stream(reinterpret_cast<uint4*>(out) + row * nvec + v, V::pack(o));
            }
        }
#pragma unroll
        for (int i = 0; i < kBaVpl; ++i) {
            xv[i] = xn[i];
            if constexpr (kBwd) gv[i] = gn[i];
        }
    }
    if constexpr (kBwd && kAccum) {
        float* prow = partial + static_cast<int64_t>(blockIdx.x) * nvec * kE;
#pragma unroll
        for (int i = 0; i < kBaVpl; ++i) {
            const int v = v0 + i * nthr;
            if (v < nvec) {
#pragma unroll
                for (int q = 0; q < kE / 4; ++q)
                    reinterpret_cast<float4*>(prow + static_cast<int64_t>(v) * kE)[q] =
                        make_float4(acc[i][4 * q], acc[i][4 * q + 1], acc[i][4 * q + 2], acc[i][4 * q + 3]);
            }
        }
    }
    if constexpr (kBwd) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

// Scalar fallback for hidden not a multiple of the vector width / unaligned pointers: one thread per column.
template <typename T, bool kBwd>
__global__ void __launch_bounds__(256)
bias_act_generic_kernel(const T* __restrict__ x, const T* __restrict__ bias, const T* __restrict__ g, T* __restrict__ out,
                        T* __restrict__ pre_out, T* __restrict__ dbias, int64_t rows, int64_t hidden, int act) {
    using V = Vec<T>;
    const int64_t col = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
    if (col >= hidden) return;
    const float b = bias != nullptr ? V::load1(bias, col) : 0.f;
    float acc = 0.f;
    for (int64_t r = 0; r < rows; ++r) {
        float a[2] = {V::load1(x, r * hidden + col), 0.f};
        if (bias != nullptr) a[0] = V::rnd(a[0] + b);
        float o[2] = {0.f, 0.f};
        if constexpr (kBwd) {
            const float gf[2] = {V::load1(g, r * hidden + col), 0.f};
            if (act == 0) act_bwd<T, 0>(a, gf, o, 2);
            else if (act == 1) act_bwd<T, 1>(a, gf, o, 2);
            else act_bwd<T, 2>(a, gf, o, 2);
            acc += V::rnd(o[0]);
        } else {
            if (pre_out != nullptr) V::store1(pre_out, r * hidden + col, a[0]);
            if (act == 1) act_fwd<T, 1>(a, 2);
            else if (act == 2) act_fwd<T, 2>(a, 2);
            o[0] = a[0];
        }
        V::store1(out, r * hidden + col, o[0]);
    }
    if (kBwd && dbias != nullptr) V::store1(dbias, col, acc);
}

// B200K_BIAS_ACT_PACKED=0 selects the fp32 form of the bf16 GELU chain (bit-for-bit comparison in the tests).
bool packed_bf16_enabled() {
    static const bool on = [] {
        const char* e = getenv("B200K_BIAS_ACT_PACKED");
        return !(e != nullptr && e[0] == '0');
    }();
    return on;
}

int row_parts(int64_t rows, int col_tiles) {
    int64_t parts = static_cast<int64_t>(sm_count()) * kBaRowParts / (col_tiles > 0 ? col_tiles : 1);
    if (parts < 1) parts = 1;
    return static_cast<int>(rows < parts ? rows : parts);
}

template <typename T, bool kBwd>
int run(const void* x, const void* bias, const void* g, void* out, void* pre_out, void* dbias, void* scratch, int64_t scratch_bytes,
        int64_t rows, int64_t hidden, int act, cudaStream_t st) {
    using V = Vec<T>;
    const bool vec_ok = hidden % V::kElems == 0 && aligned16(x) && aligned16(out) && (bias == nullptr || aligned16(bias)) &&
                        (!kBwd || aligned16(g)) && (pre_out == nullptr || aligned16(pre_out));
    if (!vec_ok) {
        bias_act_generic_kernel<T, kBwd><<<static_cast<int>((hidden + 255) / 256), 256, 0, st>>>(
            static_cast<const T*>(x), static_cast<const T*>(bias), static_cast<const T*>(g), static_cast<T*>(out), static_cast<T*>(pre_out),
            static_cast<T*>(dbias), rows, hidden, act);
        return launch_status();
    }
    const int nvec = static_cast<int>(hidden / V::kElems);
    // balanced column tiles: as few tiles as fit kBaThreads * kBaVpl vectors, each with just enough threads
    const int col_tiles = (nvec + kBaThreads * kBaVpl - 1) / (kBaThreads * kBaVpl);
    const int per_tile = (nvec + col_tiles - 1) / col_tiles;
    const int threads = ((per_tile + kBaVpl - 1) / kBaVpl + 31) / 32 * 32;
    const int parts = row_parts(rows, col_tiles);
    const bool accum = kBwd && dbias != nullptr;
    if (accum && (scratch == nullptr || !aligned16(scratch) || scratch_bytes < static_cast<int64_t>(parts) * hidden * 4))
        return B200K_ESCRATCH;
    const dim3 grid(parts, col_tiles);
    auto go = [&](auto kernel) {
        kernel<<<grid, threads, 0, st>>>(static_cast<const T*>(x), static_cast<const T*>(bias), static_cast<const T*>(g),
                                        static_cast<T*>(out), static_cast<T*>(pre_out), static_cast<float*>(scratch), rows, nvec);
        return launch_status();
    };
    auto pick = [&](auto has_bias, auto acc) {
        constexpr bool kB = decltype(has_bias)::value, kA = decltype(acc)::value;
        if constexpr (V::kDtype == B200K_BF16) {
            if (act == 2 && packed_bf16_enabled()) return go(bias_act_kernel<T, 2, kB, kBwd, kA, true, true>);
        }
        return act == 0 ? go(bias_act_kernel<T, 0, kB, kBwd, kA>) : act == 1 ? go(bias_act_kernel<T, 1, kB, kBwd, kA>) : go(bias_act_kernel<T, 2, kB, kBwd, kA>);
    };
    int rc;
    if (bias != nullptr) rc = accum ? pick(std::true_type{}, std::true_type{}) : pick(std::true_type{}, std::false_type{});
    else rc = accum ? pick(std::false_type{}, std::true_type{}) : pick(std::false_type{}, std::false_type{});
    if (rc != 0 || !accum) return rc;
    return launch_colsum_finalize<T>(scratch, dbias, parts, hidden, st);
}

template <typename T>
__global__ void __launch_bounds__(256) colsum_generic_kernel(const T* __restrict__ x, T* __restrict__ out, int64_t rows, int64_t hidden) {
    const int64_t col = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
    if (col >= hidden) return;
    float acc = 0.f;
    for (int64_t r = 0; r < rows; ++r) acc += Vec<T>::load1(x, r * hidden + col);
    Vec<T>::store1(out, col, acc);
}

// Column sum of a [rows, hidden] tensor: out[c] = round_T(sum_r f32(x[r, c])) — the reduction a bias gradient or any
// `sum` over the leading dims of a region output needs.  Same kernel as the bias backward with the activation and
// the dx store compiled out: one streaming read, fp32 partials per CTA, deterministic fold.
template <typename T>
int run_colsum(const void* x, void* out, void* scratch, int64_t scratch_bytes, int64_t rows, int64_t hidden, cudaStream_t st) {
    using V = Vec<T>;
    const bool vec_ok = hidden % V::kElems == 0 && aligned16(x);
    if (!vec_ok) {
        colsum_generic_kernel<T><<<static_cast<int>((hidden + 255) / 256), 256, 0, st>>>(static_cast<const T*>(x), static_cast<T*>(out), rows, hidden);
        return launch_status();
    }
    const int nvec = static_cast<int>(hidden / V::kElems);
    const int col_tiles = (nvec + kBaThreads * kBaVpl - 1) / (kBaThreads * kBaVpl);
    const int per_tile = (nvec + col_tiles - 1) / col_tiles;
    const int threads = ((per_tile + kBaVpl - 1) / kBaVpl + 31) / 32 * 32;
    const int parts = row_parts(rows, col_tiles);
    if (scratch == nullptr || !aligned16(scratch) || scratch_bytes < static_cast<int64_t>(parts) * hidden * 4) return B200K_ESCRATCH;
    bias_act_kernel<T, 0, false, true, true, false><<<dim3(parts, col_tiles), threads, 0, st>>>(
        static_cast<const T*>(x), nullptr, static_cast<const T*>(x), nullptr, nullptr, static_cast<float*>(scratch), rows, nvec);
    if (int rc = launch_status()) return rc;
    return launch_colsum_finalize<T>(scratch, out, parts, hidden, st);
}

}  // namespace
}  // namespace b200k

using namespace b200k;

extern "C" int64_t b200k_bias_act_bwd_scratch_bytes(int64_t rows, int64_t hidden) {
    if (rows <= 0 || hidden <= 0) return 0;
    int64_t parts = static_cast<int64_t>(sm_count()) * kBaRowParts;
    if (rows < parts) parts = rows;
    return parts * hidden * 4;
}

extern "C" int b200k_bias_act_fwd(const void* x, const void* bias, void* out, void* pre_out, int64_t rows, int64_t hidden, int act,
                                  int dtype, void* stream) {
    if (rows < 0 || hidden <= 0 || act < 0 || act > 2) return B200K_EINVAL;
    if (rows == 0) return 0;
    if (x == nullptr || out == nullptr) return B200K_EINVAL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    switch (dtype) {
        case B200K_F32: return run<float, false>(x, bias, nullptr, out, pre_out, nullptr, nullptr, 0, rows, hidden, act, st);
        case B200K_BF16: return run<__nv_bfloat16, false>(x, bias, nullptr, out, pre_out, nullptr, nullptr, 0, rows, hidden, act, st);
        case B200K_F16: return run<__half, false>(x, bias, nullptr, out, pre_out, nullptr, nullptr, 0, rows, hidden, act, st);
        default: return B200K_EDTYPE;
    }
}

extern "C" int b200k_bias_act_bwd(const void* x, const void* bias, const void* g, void* dx, void* dbias, void* scratch,
                                  int64_t scratch_bytes, int64_t rows, int64_t hidden, int act, int dtype, void* stream) {
    if (rows < 0 || hidden <= 0 || act < 0 || act > 2) return B200K_EINVAL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (rows == 0) {
        if (dbias != nullptr) {
            cudaError_t e = cudaMemsetAsync(dbias, 0, static_cast<size_t>(hidden) * (dtype == B200K_F32 ? 4 : 2), st);
            return e == cudaSuccess ? 0 : static_cast<int>(e);
        }
        return 0;
    }
    if (x == nullptr || g == nullptr || dx == nullptr) return B200K_EINVAL;
    switch (dtype) {
        case B200K_F32: return run<float, true>(x, bias, g, dx, nullptr, dbias, scratch, scratch_bytes, rows, hidden, act, st);
        case B200K_BF16: return run<__nv_bfloat16, true>(x, bias, g, dx, nullptr, dbias, scratch, scratch_bytes, rows, hidden, act, st);
        case B200K_F16: return run<__half, true>(x, bias, g, dx, nullptr, dbias, scratch, scratch_bytes, rows, hidden, act, st);
        default: return B200K_EDTYPE;
    }
}

extern "C" int b200k_colsum(const void* x, void* out, void* scratch, int64_t scratch_bytes, int64_t rows, int64_t hidden, int dtype,
                            void* stream) {
    if (rows < 0 || hidden <= 0 || out == nullptr) return B200K_EINVAL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (rows == 0) {
        cudaError_t e = cudaMemsetAsync(out, 0, static_cast<size_t>(hidden) * (dtype == B200K_F32 ? 4 : 2), st);
        return e == cudaSuccess ? 0 : static_cast<int>(e);
    }
    if (x == nullptr) return B200K_EINVAL;
    switch (dtype) {
        case B200K_F32: return run_colsum<float>(x, out, scratch, scratch_bytes, rows, hidden, st);
        case B200K_BF16: return run_colsum<__nv_bfloat16>(x, out, scratch, scratch_bytes, rows, hidden, st);
        case B200K_F16: return run_colsum<__half>(x, out, scratch, scratch_bytes, rows, hidden, st);
        default: return B200K_EDTYPE;
    }
}
