// Generic pointwise region interpreter (see b200k_pointwise in include/b200k.h).
//
// One thread evaluates the region's register program for kW consecutive elements along the inner
// dimension (kW = 4 when every unit-stride operand is suitably aligned, else 1).  The program lives in
// the kernel parameter block (constant bank), so instruction fetch is a uniform constant load and the
// opcode switch never diverges.  The register file is a per-thread local array; short programs (the
// common residual adds, scale/shift chains) stay HBM-bound, long ones are bounded by L1 bandwidth —
// hot long programs have specialised kernels (swiglu.cu, bias_act.cu) that the executor routes to.
#include "common.cuh"

namespace b200k {
namespace {

// Walks the work items a thread owns (first, first + step, ...) over the collapsed [d0, d1, d2 / width] box while
// keeping the 3-D coordinates incrementally: 64-bit divisions cost ~100 instructions each, far more than the
// handful of arithmetic instructions a short program spends per item, so they happen once per thread, not per item.
struct Walker {
    int64_t idx, i0, i1, i2v;        // current item and its coordinates (i2v in vectors)
    int64_t s0, s1, s2v, d1, d2v;    // per-step increments and extents
    __device__ __forceinline__ void init(int64_t first, int64_t step, int64_t d1_, int64_t d2v_) {
        d1 = d1_, d2v = d2v_;
        idx = first;
        i2v = first % d2v;
        const int64_t r = first / d2v;
        i1 = r % d1, i0 = r / d1;
        s2v = step % d2v;
        const int64_t q = step / d2v;
        s1 = q % d1, s0 = q / d1;
        step_ = step;
    }
    __device__ __forceinline__ void next() {
        idx += step_;
        i2v += s2v;
        if (i2v >= d2v) i2v -= d2v, ++i1;
        i1 += s1;
        if (i1 >= d1) i1 -= d1, ++i0;
        i0 += s0;
    }
    int64_t step_;
};

template <int kW>
__device__ __forceinline__ void load_operand(const b200k_pw_operand& t, int64_t off, float* r) {
    const bool bc = t.inner == 0;
    if (t.dtype == B200K_F32) {
        const float* p = static_cast<const float*>(t.ptr) + off;
        if (bc) {
            const float v = *p;
#pragma unroll
            for (int e = 0; e < kW; ++e) r[e] = v;
        } else if constexpr (kW == 4) {
            const float4 v = *reinterpret_cast<const float4*>(p);
            r[0] = v.x, r[1] = v.y, r[2] = v.z, r[3] = v.w;
        } else {
            r[0] = *p;
        }
    } else if (t.dtype == B200K_BF16 || t.dtype == B200K_F16) {
        const uint16_t* p = static_cast<const uint16_t*>(t.ptr) + off;
        uint16_t h[kW];
        if (bc) {
            const uint16_t v = *p;
#pragma unroll
            for (int e = 0; e < kW; ++e) h[e] = v;
        } else if constexpr (kW == 4) {
            const uint2 v = *reinterpret_cast<const uint2*>(p);
            h[0] = v.x & 0xffff, h[1] = v.x >> 16, h[2] = v.y & 0xffff, h[3] = v.y >> 16;
        } else {
            h[0] = *p;
        }
#pragma unroll
        for (int e = 0; e < kW; ++e)
            r[e] = t.dtype == B200K_BF16 ? __uint_as_float(static_cast<uint32_t>(h[e]) << 16) : __half2float(__ushort_as_half(h[e]));
    } else {  // bool
        const uint8_t* p = static_cast<const uint8_t*>(t.ptr) + off;
#pragma unroll
        for (int e = 0; e < kW; ++e) r[e] = p[bc ? 0 : e] != 0 ? 1.f : 0.f;
    }
}

template <int kW>
__device__ __forceinline__ void store_operand(const b200k_pw_operand& t, int64_t off, const float* r) {
    if (t.dtype == B200K_F32) {
        float* p = static_cast<float*>(t.ptr) + off;
        if constexpr (kW == 4) *reinterpret_cast<float4*>(p) = make_float4(r[0], r[1], r[2], r[3]);
        else *p = r[0];
    } else if (t.dtype == B200K_BF16 || t.dtype == B200K_F16) {
        uint16_t h[kW];
#pragma unroll
        for (int e = 0; e < kW; ++e)
            h[e] = t.dtype == B200K_BF16 ? __bfloat16_as_ushort(__float2bfloat16_rn(r[e])) : __half_as_ushort(__float2half_rn(r[e]));
        uint16_t* p = static_cast<uint16_t*>(t.ptr) + off;
        if constexpr (kW == 4) *reinterpret_cast<uint2*>(p) = make_uint2(h[0] | (static_cast<uint32_t>(h[1]) << 16), h[2] | (static_cast<uint32_t>(h[3]) << 16));
        else *p = h[0];
    } else {
        uint8_t* p = static_cast<uint8_t*>(t.ptr) + off;
#pragma unroll
        for (int e = 0; e < kW; ++e) p[e] = r[e] != 0.f ? 1 : 0;
    }
}

__device__ __forceinline__ float pw_unary(int op, float a) {
    switch (op) {
        case B200K_PW_MOV: return a;
        case B200K_PW_NEG: return -a;
        case B200K_PW_ABS: return fabsf(a);
        case B200K_PW_EXP: return expf(a);
        case B200K_PW_LOG: return logf(a);
        case B200K_PW_SQRT: return sqrtf(a);
        case B200K_PW_RSQRT: return rsqrtf(a);
        case B200K_PW_RECIP: return 1.0f / a;
        case B200K_PW_ERF: return erff(a);
        case B200K_PW_TANH: return tanhf(a);
        case B200K_PW_RND_BF16: return __bfloat162float(__float2bfloat16_rn(a));
        case B200K_PW_RND_F16: return __half2float(__float2half_rn(a));
        case B200K_PW_NOT: return a != 0.f ? 0.f : 1.f;
        case B200K_PW_SIGN: return a > 0.f ? 1.f : (a < 0.f ? -1.f : a);
        case B200K_PW_FLOOR: return floorf(a);
        case B200K_PW_CEIL: return ceilf(a);
        case B200K_PW_EXP2: return exp2f(a);
        case B200K_PW_LOG2: return log2f(a);
        case B200K_PW_LOG1P: return log1pf(a);
        case B200K_PW_EXPM1: return expm1f(a);
        case B200K_PW_SIN: return sinf(a);
        case B200K_PW_COS: return cosf(a);
        case B200K_PW_ERFC: return erfcf(a);
        case B200K_PW_SIGMOID: return 1.0f / (1.0f + expf(-a));
        case B200K_PW_TRUNC: return truncf(a);
        case B200K_PW_ROUND: return nearbyintf(a);
        default: return a;
    }
}

__device__ __forceinline__ float pw_binary(int op, float a, float b) {
    switch (op) {
        case B200K_PW_ADD: return __fadd_rn(a, b);
        case B200K_PW_SUB: return __fsub_rn(a, b);
        case B200K_PW_MUL: return __fmul_rn(a, b);
        case B200K_PW_DIV: return a / b;
        case B200K_PW_POW: return powf(a, b);
        case B200K_PW_MAX: return (a != a || b != b) ? NAN : fmaxf(a, b);
        case B200K_PW_MIN: return (a != a || b != b) ? NAN : fminf(a, b);
        case B200K_PW_GT: return a > b ? 1.f : 0.f;
        case B200K_PW_GE: return a >= b ? 1.f : 0.f;
        case B200K_PW_LT: return a < b ? 1.f : 0.f;
        case B200K_PW_LE: return a <= b ? 1.f : 0.f;
        case B200K_PW_EQ: return a == b ? 1.f : 0.f;
        case B200K_PW_NE: return a != b ? 1.f : 0.f;
        case B200K_PW_AND: return (a != 0.f && b != 0.f) ? 1.f : 0.f;
        case B200K_PW_OR: return (a != 0.f || b != 0.f) ? 1.f : 0.f;
        default: return a;
    }
}

__host__ __device__ inline bool is_binary(int op) {
    return op == B200K_PW_ADD || op == B200K_PW_SUB || op == B200K_PW_MUL || op == B200K_PW_DIV || op == B200K_PW_POW ||
           (op >= B200K_PW_MAX && op <= B200K_PW_NE) || op == B200K_PW_AND || op == B200K_PW_OR;
}

template <int kW>
__global__ void __launch_bounds__(256)
pointwise_kernel(const __grid_constant__ b200k_pw_program P) {
    float R[B200K_PW_MAX_REG][kW];
    const int64_t d2v = P.d2 / kW;
    const int64_t total = P.d0 * P.d1 * d2v;
    Walker wk;
    wk.init(blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x, static_cast<int64_t>(gridDim.x) * blockDim.x, P.d1, d2v);
    for (; wk.idx < total; wk.next()) {
        const int64_t i2 = wk.i2v * kW, i1 = wk.i1, i0 = wk.i0;
        for (int k = 0; k < P.n_in; ++k) {
            const b200k_pw_operand& t = P.in[k];
            load_operand<kW>(t, i0 * t.s0 + i1 * t.s1 + i2 * t.inner, R[k]);
        }
#pragma unroll 1
        for (int pc = 0; pc < P.n_code; ++pc) {
            const uint32_t w = P.code[pc];
            const int op = w & 0xff, dst = (w >> 8) & 0xff, ra = (w >> 16) & 0xff, rb = w >> 24;
            if (op == B200K_PW_CONST) {
                const float c = __uint_as_float(P.code[++pc]);
#pragma unroll
                for (int e = 0; e < kW; ++e) R[dst][e] = c;
            } else if (op == B200K_PW_WHERE) {
                const int rc = P.code[++pc] & 0xff;
#pragma unroll
                for (int e = 0; e < kW; ++e) R[dst][e] = R[ra][e] != 0.f ? R[rb][e] : R[rc][e];
            } else if (is_binary(op)) {
#pragma unroll
                for (int e = 0; e < kW; ++e) R[dst][e] = pw_binary(op, R[ra][e], R[rb][e]);
            } else {
#pragma unroll
                for (int e = 0; e < kW; ++e) R[dst][e] = pw_unary(op, R[ra][e]);
            }
        }
        for (int k = 0; k < P.n_out; ++k) {
            const b200k_pw_operand& t = P.out[k];
            store_operand<kW>(t, i0 * t.s0 + i1 * t.s1 + i2, R[P.out_reg[k]]);
        }
    }
}

// -------------------------------------------------------------------------------------------
// Register-resident interpreter (the `small_code` form): 8 elements per thread, a value file of
// B200K_PW_SMALL_REGS x 8 floats that is only ever indexed with compile-time constants (operand selection is a
// uniform switch over the register number), so it lives in hardware registers instead of local memory.
// -------------------------------------------------------------------------------------------
constexpr int kSW = 8;
constexpr int kSR = B200K_PW_SMALL_REGS;

__device__ __forceinline__ void load8(const b200k_pw_operand& t, int64_t off, float* r) {
    if (t.inner == 0) {
        float v;
        if (t.dtype == B200K_F32) v = static_cast<const float*>(t.ptr)[off];
        else if (t.dtype == B200K_BF16) v = __uint_as_float(static_cast<uint32_t>(static_cast<const uint16_t*>(t.ptr)[off]) << 16);
        else if (t.dtype == B200K_F16) v = __half2float(__ushort_as_half(static_cast<const uint16_t*>(t.ptr)[off]));
        else v = static_cast<const uint8_t*>(t.ptr)[off] != 0 ? 1.f : 0.f;
#pragma unroll
        for (int e = 0; e < kSW; ++e) r[e] = v;
        return;
    }
    if (t.dtype == B200K_F32) {
        const float* p = static_cast<const float*>(t.ptr) + off;
        const uint4 a = ld_stream(p), b = ld_stream(p + 4);
        Vec<float>::unpack(a, r);
        Vec<float>::unpack(b, r + 4);
    } else if (t.dtype == B200K_BF16) {
        Vec<__nv_bfloat16>::unpack(ld_stream(static_cast<const uint16_t*>(t.ptr) + off), r);
    } else if (t.dtype == B200K_F16) {
        Vec<__half>::unpack(ld_stream(static_cast<const uint16_t*>(t.ptr) + off), r);
    } else {
        const uint2 v = *reinterpret_cast<const uint2*>(static_cast<const uint8_t*>(t.ptr) + off);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            r[e] = ((v.x >> (8 * e)) & 0xff) != 0 ? 1.f : 0.f;
            r[4 + e] = ((v.y >> (8 * e)) & 0xff) != 0 ? 1.f : 0.f;
        }
    }
}

__device__ __forceinline__ void store8(const b200k_pw_operand& t, int64_t off, const float* r) {
    if (t.dtype == B200K_F32) {
        float* p = static_cast<float*>(t.ptr) + off;
        st_stream(p, Vec<float>::pack(r));
        st_stream(p + 4, Vec<float>::pack(r + 4));
    } else if (t.dtype == B200K_BF16) {
        st_stream(static_cast<uint16_t*>(t.ptr) + off, Vec<__nv_bfloat16>::pack(r));
    } else if (t.dtype == B200K_F16) {
        st_stream(static_cast<uint16_t*>(t.ptr) + off, Vec<__half>::pack(r));
    } else {
        uint32_t lo = 0, hi = 0;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            lo |= (r[e] != 0.f ? 1u : 0u) << (8 * e);
            hi |= (r[4 + e] != 0.f ? 1u : 0u) << (8 * e);
        }
        *reinterpret_cast<uint2*>(static_cast<uint8_t*>(t.ptr) + off) = make_uint2(lo, hi);
    }
}

// v = register `idx` of the value file, or the immediate.  Every R[...] index is a literal.
__device__ __forceinline__ void sm_fetch(const float (&R)[kSR][kSW], int idx, float imm, float* v) {
#define B200K_CASE(i)                                  \
    case i:                                            \
        _Pragma("unroll") for (int e = 0; e < kSW; ++e) v[e] = R[i][e]; \
        break;
    switch (idx) {
        B200K_CASE(0) B200K_CASE(1) B200K_CASE(2) B200K_CASE(3) B200K_CASE(4) B200K_CASE(5) B200K_CASE(6) B200K_CASE(7)
        default:
#pragma unroll
            for (int e = 0; e < kSW; ++e) v[e] = imm;
    }
#undef B200K_CASE
}

__device__ __forceinline__ void sm_put(float (&R)[kSR][kSW], int idx, const float* v) {
#define B200K_CASE(i)                                  \
    case i:                                            \
        _Pragma("unroll") for (int e = 0; e < kSW; ++e) R[i][e] = v[e]; \
        break;
    switch (idx) {
        B200K_CASE(0) B200K_CASE(1) B200K_CASE(2) B200K_CASE(3) B200K_CASE(4) B200K_CASE(5) B200K_CASE(6) B200K_CASE(7)
        default: break;
    }
#undef B200K_CASE
}

// Unit-stride operands are prefetched kAhead work items ahead with cp.async into a per-thread shared-memory
// ring (32 bytes per operand per item), so the bytes in flight per SM are set by the ring — tens of KB — and
// not by how many registers a thread can dedicate to outstanding loads while it interprets the program.
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async8(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int kN>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(kN) : "memory");
}

// fp32 view of one staged operand slot
__device__ __forceinline__ void unstage8(const b200k_pw_operand& t, const unsigned char* slot, float* r) {
    if (t.dtype == B200K_F32) {
        Vec<float>::unpack(lds128(slot), r);
        Vec<float>::unpack(lds128(slot + 16), r + 4);
    } else if (t.dtype == B200K_BF16) {
        Vec<__nv_bfloat16>::unpack(lds128(slot), r);
    } else if (t.dtype == B200K_F16) {
        Vec<__half>::unpack(lds128(slot), r);
    } else {
        const uint2 v = *reinterpret_cast<const uint2*>(slot);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            r[e] = ((v.x >> (8 * e)) & 0xff) != 0 ? 1.f : 0.f;
            r[4 + e] = ((v.y >> (8 * e)) & 0xff) != 0 ? 1.f : 0.f;
        }
    }
}

template <int kAhead>
__global__ void __launch_bounds__(256)
pointwise_reg_kernel(const __grid_constant__ b200k_pw_program P) {
    extern __shared__ __align__(16) unsigned char pw_ring[];
    const int64_t d2v = P.d2 / kSW;
    const int64_t total = P.d0 * P.d1 * d2v;
    const int64_t first = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
    const int64_t step = static_cast<int64_t>(gridDim.x) * blockDim.x;
    auto slot = [&](int stage, int k) { return pw_ring + ((static_cast<size_t>(stage) * P.n_in + k) * blockDim.x + threadIdx.x) * 32; };
    auto issue = [&](const Walker& w, int stage) {
        if (w.idx < total) {
            const int64_t i2 = w.i2v * kSW, i1 = w.i1, i0 = w.i0;
#pragma unroll
            for (int k = 0; k < kSR; ++k) {
                if (k < P.n_in && P.in[k].inner == 1) {
                    const b200k_pw_operand& t = P.in[k];
                    const int64_t off = i0 * t.s0 + i1 * t.s1 + i2;
                    unsigned char* dst = slot(stage, k);
                    if (t.dtype == B200K_F32) {
                        const float* src = static_cast<const float*>(t.ptr) + off;
                        cp_async16(dst, src);
                        cp_async16(dst + 16, src + 4);
                    } else if (t.dtype == B200K_BOOL) {
                        cp_async8(dst, static_cast<const uint8_t*>(t.ptr) + off);
                    } else {
                        cp_async16(dst, static_cast<const uint16_t*>(t.ptr) + off);
                    }
                }
            }
        }
        cp_async_commit();  // one group per item, empty or not, so that wait_group counts items
    };
    Walker cur, pre;  // the item being computed and the one being prefetched (kAhead items further on)
    cur.init(first, step, P.d1, d2v);
    pre = cur;
#pragma unroll
    for (int j = 0; j < kAhead; ++j) {
        issue(pre, j);
        pre.next();
    }

    int stage = 0;
    for (; cur.idx < total; cur.next()) {
        const int64_t i2 = cur.i2v * kSW, i1 = cur.i1, i0 = cur.i0;
        cp_async_wait<kAhead - 1>();
        float R[kSR][kSW];
#pragma unroll
        for (int k = 0; k < kSR; ++k) {
            if (k < P.n_in) {
                const b200k_pw_operand& t = P.in[k];
                if (t.inner == 1) unstage8(t, slot(stage, k), R[k]);
                else load8(t, i0 * t.s0 + i1 * t.s1 + i2 * t.inner, R[k]);
            }
        }
        issue(pre, stage);  // refill the slot just consumed
        pre.next();
        stage = stage + 1 == kAhead ? 0 : stage + 1;
#pragma unroll 1
        for (int pc = 0; pc < P.n_small; ++pc) {
            const uint32_t w0 = P.small_code[3 * pc];
            const int op = w0 & 0xff, dst = (w0 >> 8) & 0xff, ra = (w0 >> 16) & 0xff, rb = w0 >> 24;
            const float imm = __uint_as_float(P.small_code[3 * pc + 2]);
            float a[kSW], b[kSW], o[kSW];
            sm_fetch(R, ra, imm, a);
            if (is_binary(op) || op == B200K_PW_WHERE) sm_fetch(R, rb, imm, b);
            switch (op) {
                case B200K_PW_ADD:
#pragma unroll
                    for (int e = 0; e < kSW; ++e) o[e] = __fadd_rn(a[e], b[e]);
                    break;
                case B200K_PW_SUB:
#pragma unroll
                    for (int e = 0; e < kSW; ++e) o[e] = __fsub_rn(a[e], b[e]);
                    break;
                case B200K_PW_MUL:
#pragma unroll
                    for (int e = 0; e < kSW; ++e) o[e] = __fmul_rn(a[e], b[e]);
                    break;
                case B200K_PW_DIV:
#pragma unroll
                    for (int e = 0; e < kSW; ++e) o[e] = a[e] / b[e];
                    break;
                case B200K_PW_RND_BF16:
#pragma unroll
                    for (int e = 0; e < kSW; e += 2) {
                        o[e] = a[e], o[e + 1] = a[e + 1];
                        Vec<__nv_bfloat16>::rnd2(o[e], o[e + 1]);
                    }
                    break;
                case B200K_PW_RND_F16:
#pragma unroll
                    for (int e = 0; e < kSW; e += 2) {
                        o[e] = a[e], o[e + 1] = a[e + 1];
                        Vec<__half>::rnd2(o[e], o[e + 1]);
                    }
                    break;
                case B200K_PW_WHERE: {
                    float c[kSW];
                    sm_fetch(R, P.small_code[3 * pc + 1] & 0xff, imm, c);
#pragma unroll
                    for (int e = 0; e < kSW; ++e) o[e] = a[e] != 0.f ? b[e] : c[e];
                    break;
                }
                default:
                    if (is_binary(op)) {
#pragma unroll
                        for (int e = 0; e < kSW; ++e) o[e] = pw_binary(op, a[e], b[e]);
                    } else {
#pragma unroll
                        for (int e = 0; e < kSW; ++e) o[e] = pw_unary(op, a[e]);
                    }
            }
            sm_put(R, dst, o);
        }
        for (int k = 0; k < P.n_out; ++k) {
            const b200k_pw_operand& t = P.out[k];
            float o[kSW];
            sm_fetch(R, P.small_out_reg[k], 0.f, o);
            store8(t, i0 * t.s0 + i1 * t.s1 + i2, o);
        }
    }
}

bool operand_vec8_ok(const b200k_pw_operand& t) {
    if (t.inner == 0) return true;
    if (t.inner != 1) return false;
    const uintptr_t a = reinterpret_cast<uintptr_t>(t.ptr);
    return a % 16 == 0 && t.s0 % 8 == 0 && t.s1 % 8 == 0;
}

// Validates the compact form; returns false if it cannot be used for this launch.
bool small_form_ok(const b200k_pw_program* P) {
    if (P->n_small <= 0 || P->n_small > B200K_PW_SMALL_MAX_INSTR || P->n_in > kSR || P->d2 % kSW != 0) return false;
    auto reg_ok = [](int r) { return (r >= 0 && r < kSR) || r == B200K_PW_IMM; };
    for (int pc = 0; pc < P->n_small; ++pc) {
        const uint32_t w0 = P->small_code[3 * pc];
        const int op = w0 & 0xff, dst = (w0 >> 8) & 0xff, ra = (w0 >> 16) & 0xff, rb = w0 >> 24;
        if (op >= B200K_PW_NOP_COUNT || op == B200K_PW_CONST || dst >= kSR || !reg_ok(ra)) return false;
        if ((is_binary(op) || op == B200K_PW_WHERE) && !reg_ok(rb)) return false;
        if (op == B200K_PW_WHERE && !reg_ok(static_cast<int>(P->small_code[3 * pc + 1] & 0xff))) return false;
    }
    for (int k = 0; k < P->n_in; ++k)
        if (!operand_vec8_ok(P->in[k])) return false;
    for (int k = 0; k < P->n_out; ++k)
        if (!operand_vec8_ok(P->out[k]) || P->small_out_reg[k] < 0 || P->small_out_reg[k] >= kSR) return false;
    return true;
}

int elem_size(int dtype) { return dtype == B200K_F32 ? 4 : (dtype == B200K_BOOL ? 1 : 2); }

bool operand_vec4_ok(const b200k_pw_operand& t) {
    if (t.inner == 0) return true;
    if (t.inner != 1) return false;  // strided inner dimension: scalar kernel
    const int es = elem_size(t.dtype);
    if (t.dtype == B200K_BOOL) return true;  // byte loads
    const uintptr_t a = reinterpret_cast<uintptr_t>(t.ptr);
    return a % (4 * es) == 0 && t.s0 % 4 == 0 && t.s1 % 4 == 0;
}

}  // namespace
}  // namespace b200k

using namespace b200k;

extern "C" int b200k_pointwise(const b200k_pw_program* P, void* stream) {
    if (P == nullptr) return B200K_EINVAL;
    if (P->n_in < 0 || P->n_in > B200K_PW_MAX_IN || P->n_out < 1 || P->n_out > B200K_PW_MAX_OUT || P->n_code < 0 ||
        P->n_code > B200K_PW_MAX_CODE || P->n_reg < P->n_in || P->n_reg > B200K_PW_MAX_REG)
        return B200K_ELIMIT;
    if (P->d0 < 0 || P->d1 < 0 || P->d2 < 0) return B200K_EINVAL;
    if (P->d0 * P->d1 * P->d2 == 0) return 0;
    // validate the program once on the host: register indices in range, operands well formed
    for (int pc = 0; pc < P->n_code; ++pc) {
        const uint32_t w = P->code[pc];
        const int op = w & 0xff, dst = (w >> 8) & 0xff, ra = (w >> 16) & 0xff, rb = w >> 24;
        if (op >= B200K_PW_NOP_COUNT || dst >= P->n_reg) return B200K_EINVAL;
        if (op == B200K_PW_CONST) {
            if (++pc >= P->n_code) return B200K_EINVAL;
            continue;
        }
        if (ra >= P->n_reg) return B200K_EINVAL;
        if (op == B200K_PW_WHERE) {
            if (++pc >= P->n_code || rb >= P->n_reg || static_cast<int>(P->code[pc] & 0xff) >= P->n_reg) return B200K_EINVAL;
        } else if (is_binary(op) && rb >= P->n_reg) {
            return B200K_EINVAL;
        }
    }
    bool vec = P->d2 % 4 == 0;
    for (int k = 0; k < P->n_in; ++k) {
        if (P->in[k].ptr == nullptr || P->in[k].dtype < 0 || P->in[k].dtype > B200K_BOOL) return B200K_EINVAL;
        vec = vec && operand_vec4_ok(P->in[k]);
    }
    for (int k = 0; k < P->n_out; ++k) {
        if (P->out[k].ptr == nullptr || P->out[k].inner != 1 || P->out_reg[k] < 0 || P->out_reg[k] >= P->n_reg) return B200K_EINVAL;
        if (P->out[k].dtype < 0 || P->out[k].dtype > B200K_BOOL) return B200K_EINVAL;
        vec = vec && operand_vec4_ok(P->out[k]);
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (small_form_ok(P)) {
        const int64_t total8 = P->d0 * P->d1 * (P->d2 / kSW);
        // ring: 32 bytes per operand per item per thread; as deep as ~96 KB per CTA allows (two CTAs per SM)
        const int n_in = P->n_in > 0 ? P->n_in : 1;
        const int per_stage = n_in * 256 * 32;
        const int ahead = 96 * 1024 / per_stage >= 4 ? 4 : (96 * 1024 / per_stage >= 2 ? 2 : 1);
        const size_t smem = static_cast<size_t>(ahead) * per_stage;
        const int64_t want8 = (total8 + 255) / 256;
        const int64_t cap8 = static_cast<int64_t>(sm_count()) * 2;  // persistent: every CTA walks many items
        const int grid8 = static_cast<int>(want8 < cap8 ? want8 : cap8);
        auto launch = [&](auto kernel) {
            cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
            if (e != cudaSuccess) return static_cast<int>(e);
            kernel<<<grid8, 256, smem, st>>>(*P);
            return launch_status();
        };
        return ahead == 4 ? launch(pointwise_reg_kernel<4>) : ahead == 2 ? launch(pointwise_reg_kernel<2>) : launch(pointwise_reg_kernel<1>);
    }
    const int w = vec ? 4 : 1;
    const int64_t total = P->d0 * P->d1 * (P->d2 / w);
    const int64_t want = (total + 255) / 256;
    const int64_t cap = static_cast<int64_t>(sm_count()) * 16;
    const int grid = static_cast<int>(want < cap ? want : cap);
    if (vec) pointwise_kernel<4><<<grid, 256, 0, st>>>(*P);
    else pointwise_kernel<1><<<grid, 256, 0, st>>>(*P);
    return launch_status();
}
