// Multi-tensor AdamW / Adam step (the optimizer step on the far side of the fused regions).
//
// One launch updates every parameter of a (device, dtype) group.  The host hands over a device table of
// {param, grad, exp_avg, exp_avg_sq, numel} and a chunk list (tensor index, chunk index): a persistent grid walks
// the chunks, each CTA streaming 8192 elements of the four tensors with 16 loads in flight per thread.  Per element
// 8 bytes are read and 6 written (16-bit parameters and states; fp32: 16 / 12), nothing else touches memory.
//
// The arithmetic is the one of ATen's fused optimizer (torch.optim.AdamW(fused=True), which the reference's training
// benchmark uses, thunder/benchmarks/benchmark_litgpt.py:280): fp32 throughout, storage dtype only on load / store.
#include "common.cuh"

namespace b200k {
namespace {

constexpr int kAdamThreads = 256;
constexpr int kAdamVpt = 4;                                   // 16-byte vectors per thread and tensor
constexpr int64_t kAdamChunkVecs = kAdamThreads * kAdamVpt;   // 1024 vectors per chunk

struct AdamArgs {
    float lr, beta1, beta2, eps, weight_decay;
    float bias_correction1, bias_correction2_sqrt;
    int decoupled;
};

// ATen's adam_math with every hyper-parameter narrowed to the fp32 math type first (measured: the moments of
// torch.optim.AdamW(fused=True) are reproduced bit for bit by the fp32 form, not by a double evaluation).  The
// expressions keep ATen's shape so that the compiler's multiply-add contraction falls the same way.
__device__ __forceinline__ void adam_math(float& param, float grad, float& m, float& v, const AdamArgs& a) {
    if (a.weight_decay != 0.f) {
        if (a.decoupled) param -= a.lr * a.weight_decay * param;
        else grad += param * a.weight_decay;
    }
    m = a.beta1 * m + (1 - a.beta1) * grad;
    v = a.beta2 * v + (1 - a.beta2) * grad * grad;
    const float step_size = a.lr / a.bias_correction1;
    const float denom = (sqrtf(v) / a.bias_correction2_sqrt) + a.eps;
    param -= step_size * m / denom;
}

template <typename T>
__global__ void __launch_bounds__(kAdamThreads)
adamw_kernel(const b200k_adam_tensor* __restrict__ table, const int32_t* __restrict__ chunk_tensor,
             const int32_t* __restrict__ chunk_index, int64_t n_chunks, AdamArgs a) {
    using V = Vec<T>;
    constexpr int kE = V::kElems;
    for (int64_t c = blockIdx.x; c < n_chunks; c += gridDim.x) {
        const b200k_adam_tensor t = table[chunk_tensor[c]];
        T* p = static_cast<T*>(t.param);
        const T* g = static_cast<const T*>(t.grad);
        T* m = static_cast<T*>(t.exp_avg);
        T* v = static_cast<T*>(t.exp_avg_sq);
        const int64_t vec0 = static_cast<int64_t>(chunk_index[c]) * kAdamChunkVecs;
        // 16-byte access needs all four tensors on 16-byte addresses; an offset view (a shard of a flat buffer, a DDP
        // bucket view, p[1:]) is updated element by element instead
        const bool aligned = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
                               reinterpret_cast<uintptr_t>(v)) & 15) == 0;
        if (!aligned) {
            const int64_t e0 = vec0 * kE, e1 = e0 + kAdamChunkVecs * kE < t.numel ? e0 + kAdamChunkVecs * kE : t.numel;
            for (int64_t i = e0 + threadIdx.x; i < e1; i += kAdamThreads) {
                float pf = V::load1(p, i), mf = V::load1(m, i), vf = V::load1(v, i);
                adam_math(pf, V::load1(g, i), mf, vf, a);
                V::store1(p, i, pf), V::store1(m, i, mf), V::store1(v, i, vf);
            }
            continue;
        }
        const int64_t nvec = t.numel / kE;  // whole vectors; the tail (< kE elements) is handled below
        uint4 pv[kAdamVpt], gv[kAdamVpt], mv[kAdamVpt], vv[kAdamVpt];
#pragma unroll
        for (int k = 0; k < kAdamVpt; ++k) {
            const int64_t i = vec0 + threadIdx.x + k * kAdamThreads;
            if (i < nvec) {
                // p, m, v are rewritten below: plain (coherent) loads; only the gradient is read-only here
                pv[k] = ld_stream_rw(p + i * kE), gv[k] = ld_stream(g + i * kE);
                mv[k] = ld_stream_rw(m + i * kE), vv[k] = ld_stream_rw(v + i * kE);
            }
        }
#pragma unroll
        for (int k = 0; k < kAdamVpt; ++k) {
            const int64_t i = vec0 + threadIdx.x + k * kAdamThreads;
            if (i < nvec) {
                float pf[kE], gf[kE], mf[kE], vf[kE];
                V::unpack(pv[k], pf), V::unpack(gv[k], gf), V::unpack(mv[k], mf), V::unpack(vv[k], vf);
#pragma unroll
                for (int e = 0; e < kE; ++e) adam_math(pf[e], gf[e], mf[e], vf[e], a);
                st_stream(p + i * kE, V::pack(pf)), st_stream(m + i * kE, V::pack(mf)), st_stream(v + i * kE, V::pack(vf));
            }
        }
        // tail elements of this tensor: done by the CTA that owns its last chunk
        const int64_t tail0 = nvec * kE;
        if (tail0 < t.numel && vec0 <= nvec && nvec < vec0 + kAdamChunkVecs) {
            const int64_t i = tail0 + threadIdx.x;
            if (i < t.numel) {
                float pf = V::load1(p, i), mf = V::load1(m, i), vf = V::load1(v, i);
                adam_math(pf, V::load1(g, i), mf, vf, a);
                V::store1(p, i, pf), V::store1(m, i, mf), V::store1(v, i, vf);
            }
        }
    }
}

}  // namespace
}  // namespace b200k

using namespace b200k;

extern "C" int64_t b200k_adamw_chunk_elems(int dtype) { return kAdamChunkVecs * (dtype == B200K_F32 ? 4 : 8); }

extern "C" int b200k_adamw(const b200k_adam_tensor* table, const int32_t* chunk_tensor, const int32_t* chunk_index, int64_t n_chunks,
                           double lr, double beta1, double beta2, double eps, double weight_decay, float bias_correction1,
                           float bias_correction2_sqrt, int decoupled_weight_decay, int dtype, void* stream) {
    if (n_chunks < 0) return B200K_EINVAL;
    if (n_chunks == 0) return 0;
    if (table == nullptr || chunk_tensor == nullptr || chunk_index == nullptr) return B200K_EINVAL;
    const AdamArgs a{static_cast<float>(lr), static_cast<float>(beta1), static_cast<float>(beta2), static_cast<float>(eps), static_cast<float>(weight_decay),
                     bias_correction1, bias_correction2_sqrt, decoupled_weight_decay ? 1 : 0};
    const int64_t cap = static_cast<int64_t>(sm_count()) * 8;
    const int grid = static_cast<int>(n_chunks < cap ? n_chunks : cap);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    switch (dtype) {
        case B200K_F32: adamw_kernel<float><<<grid, kAdamThreads, 0, st>>>(table, chunk_tensor, chunk_index, n_chunks, a); break;
        case B200K_BF16: adamw_kernel<__nv_bfloat16><<<grid, kAdamThreads, 0, st>>>(table, chunk_tensor, chunk_index, n_chunks, a); break;
        case B200K_F16: adamw_kernel<__half><<<grid, kAdamThreads, 0, st>>>(table, chunk_tensor, chunk_index, n_chunks, a); break;
        default: return B200K_EDTYPE;
    }
    return launch_status();
}
