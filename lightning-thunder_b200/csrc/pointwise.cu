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
    for (int64_t idx = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; idx < total;
         idx += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int64_t i2 = (idx % d2v) * kW;
        const int64_t rest = idx / d2v;
        const int64_t i1 = rest % P.d1, i0 = rest / P.d1;
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
    const int w = vec ? 4 : 1;
    const int64_t total = P->d0 * P->d1 * (P->d2 / w);
    const int64_t want = (total + 255) / 256;
    const int64_t cap = static_cast<int64_t>(sm_count()) * 16;
    const int grid = static_cast<int>(want < cap ? want : cap);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (vec) pointwise_kernel<4><<<grid, 256, 0, st>>>(*P);
    else pointwise_kernel<1><<<grid, 256, 0, st>>>(*P);
    return launch_status();
}
