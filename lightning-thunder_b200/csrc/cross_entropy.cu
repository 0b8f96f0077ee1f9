// Softmax cross-entropy forward / backward.
//
// Forward: one CTA streams one row of logits exactly once (online softmax: running max and
// rescaled sum per thread, merged across the CTA), so the [rows, vocab] log-probability tensor the
// reference materialises is never written: only max, log-sum and the picked -log p per row.
// A one-CTA finalize kernel folds the per-row values in a fixed order (deterministic) into the loss
// and the valid-target count.
// Backward: one pass, reads logits, writes dlogits; log p is recomputed from the saved max and
// log-sum and rounded to the storage dtype exactly where the reference holds it as a tensor.
#include <cooperative_groups.h>

#include <cstdio>
#include <cstdlib>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace b200k {
namespace {

// forward: 256 threads x 8 loads in flight (latency-bound on the single pass over a 256 KB row);
// backward: 512 x 4 (write-heavy, more CTAs in flight does not help)
constexpr int kCeThreads = 256;
constexpr int kCeUnroll = 8;
constexpr int kCeBwdThreads = 512;
constexpr int kCeBwdUnroll = 4;
constexpr float kLog2e = 1.4426950408889634f;

// 2^x, hardware approximation (2 ulp), flushing results below 2^-126 to zero: such terms cannot change an fp32 sum.
__device__ __forceinline__ float fast_exp2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

struct MS {
    float m, s;
};
__device__ __forceinline__ MS ms_merge(MS a, MS b) {
    const float m = fmaxf(a.m, b.m);
    // exp(-inf - -inf) is NaN: an all -inf operand contributes nothing.
    const float sa = a.m == -INFINITY ? 0.f : a.s * exp2f((a.m - m) * kLog2e);
    const float sb = b.m == -INFINITY ? 0.f : b.s * exp2f((b.m - m) * kLog2e);
    return {m, sa + sb};
}

// kCluster CTAs share one row (thread-block cluster, launched with cudaLaunchAttributeClusterDimension): each streams
// a 1/kCluster slice of the vocabulary, the non-leading CTAs deposit their (max, sum) pair in the leading CTA's shared
// memory through distributed shared memory, and one cluster barrier later the leading CTA merges and writes the row's
// results.  Used when there are too few rows to occupy the GPU with one CTA per row (a 128k-vocabulary row is 256 KB:
// with a few hundred rows the single-CTA form leaves SMs idle or a ragged last wave); kCluster = 1 is the plain form.
template <typename T, int kCluster>
__global__ void __launch_bounds__(kCeThreads)
ce_fwd_kernel(const T* __restrict__ logits, const int64_t* __restrict__ target, float* __restrict__ row_max,
              float* __restrict__ row_logsum, float* __restrict__ row_nll, int64_t rows, int64_t vocab, int64_t ignore_index,
              int vec_ok) {
    using V = Vec<T>;
    constexpr int kE = V::kElems;
    __shared__ float sm_m[kCeThreads / 32], sm_s[kCeThreads / 32];
    __shared__ float peer_m[8], peer_s[8];  // slot r: the pair of cluster rank r (written remotely)
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t nvec = vec_ok ? vocab / kE : 0;
    const int crank = kCluster > 1 ? static_cast<int>(cg::this_cluster().block_rank()) : 0;
    const int64_t first_row = kCluster > 1 ? blockIdx.x / kCluster : blockIdx.x;
    const int64_t row_step = kCluster > 1 ? gridDim.x / kCluster : gridDim.x;
    // this CTA's slice of the row's vectors
    const int64_t v0 = kCluster > 1 ? nvec * crank / kCluster : 0;
    const int64_t v1 = kCluster > 1 ? nvec * (crank + 1) / kCluster : nvec;

    for (int64_t row = first_row; row < rows; row += row_step) {
        const T* x = logits + row * vocab;
        MS acc = {-INFINITY, 0.f};
        // vectorised body
        for (int64_t base = v0 + tid; base < v1; base += kCeThreads * kCeUnroll) {
            uint4 v[kCeUnroll];
#pragma unroll
            for (int u = 0; u < kCeUnroll; ++u) {
                const int64_t i = base + static_cast<int64_t>(u) * kCeThreads;
                if (i < v1) v[u] = ld_stream(reinterpret_cast<const uint4*>(x) + i);
            }
            // One running-max update (and one rescale of the running sum) per batch of kCeUnroll vectors, not per
            // vector: the kernel is bound by the special-function unit (ncu: XU pipe 72 %), and the rescale's ex2 is
            // one MUFU in nine when it is paid per vector (391 -> 364 us).  Moving a share of the exponentials onto the
            // FMA pipe (Cody-Waite split + degree-6 polynomial) was measured twice and is slower (401 / 413 us).
            float mx = -INFINITY;
#pragma unroll
            for (int u = 0; u < kCeUnroll; ++u) {
                const int64_t i = base + static_cast<int64_t>(u) * kCeThreads;
                if (i < v1) {
                    float f[kE];
                    V::unpack(v[u], f);
#pragma unroll
                    for (int e = 0; e < kE; ++e) mx = fmaxf(mx, f[e]);
                }
            }
            const float m = fmaxf(acc.m, mx);
            if (m != -INFINITY) {
                float s = acc.s * fast_exp2((acc.m - m) * kLog2e);
                const bool packed = kE == 8 && fabsf(m) <= 64.f;
                // Packed form (two logits per FFMA2 / FADD2): x * log2e - m * log2e.  Scaling x and m separately
                // loses up to |m| * log2e * 2^-24 in the exponent — below 6e-6 for |m| <= 64, i.e. under 4e-6 relative
                // on a term, far inside the fp32 tolerance of the log-sum.  Otherwise (x - m) first: exact for the
                // terms that matter.
                const float2 c2 = make_float2(kLog2e, kLog2e);
                const float nm = -(m * kLog2e);
                const float2 nm2 = make_float2(nm, nm);
                float2 s2 = make_float2(s, 0.f);
#pragma unroll
                for (int u = 0; u < kCeUnroll; ++u) {
                    const int64_t i = base + static_cast<int64_t>(u) * kCeThreads;
                    if (i < v1) {
                        float f[kE];
                        V::unpack(v[u], f);
                        if (packed) {
#pragma unroll
                            for (int e = 0; e + 1 < kE; e += 2) {
                                const float2 a = __ffma2_rn(make_float2(f[e], f[e + 1]), c2, nm2);
                                s2 = __fadd2_rn(s2, make_float2(fast_exp2(a.x), fast_exp2(a.y)));
                            }
                        } else {
#pragma unroll
                            for (int e = 0; e < kE; ++e) s2.x += fast_exp2((f[e] - m) * kLog2e);
                        }
                    }
                }
                acc = {m, s2.x + s2.y};
            }
        }
        // scalar tail (vocab not a multiple of the vector width, or unaligned rows): the leading CTA's
        if (crank == 0) {
            for (int64_t j = nvec * kE + tid; j < vocab; j += kCeThreads) {
                const float f = V::load1(x, j);
                acc = ms_merge(acc, MS{f, 1.f});
            }
        }
        // CTA merge
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            MS other = {__shfl_xor_sync(0xffffffffu, acc.m, o), __shfl_xor_sync(0xffffffffu, acc.s, o)};
            acc = ms_merge(acc, other);
        }
        if (lane == 0) {
            sm_m[warp] = acc.m;
            sm_s[warp] = acc.s;
        }
        __syncthreads();
        MS t = {-INFINITY, 0.f};
        if (warp == 0) {
            t = lane < kCeThreads / 32 ? MS{sm_m[lane], sm_s[lane]} : MS{-INFINITY, 0.f};
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) {
                MS other = {__shfl_xor_sync(0xffffffffu, t.m, o), __shfl_xor_sync(0xffffffffu, t.s, o)};
                t = ms_merge(t, other);
            }
            if (kCluster > 1 && lane == 0 && crank != 0) {
                // hand this CTA's pair to the leading CTA of the cluster (distributed shared memory)
                cg::cluster_group cl = cg::this_cluster();
                *cl.map_shared_rank(&peer_m[crank], 0) = t.m;
                *cl.map_shared_rank(&peer_s[crank], 0) = t.s;
            }
        }
        // every thread of every CTA of the cluster takes part in the cluster barrier; after it the remote stores are
        // visible to the leading CTA
        if (kCluster > 1) cg::this_cluster().sync();
        if (warp == 0 && lane == 0 && crank == 0) {
#pragma unroll
            for (int r = 1; r < kCluster; ++r) t = ms_merge(t, MS{peer_m[r], peer_s[r]});
            const float logsum = logf(t.s);
            row_max[row] = t.m;
            row_logsum[row] = logsum;
            const int64_t tg = target[row];
            float nll = 0.f;
            if (tg != ignore_index && (tg < 0 || tg >= vocab)) {
                // ATen's nll_loss device-asserts on a class index outside [0, vocab) (the reference's oracle path,
                // torchex.py:1731-1783); same contract here: report and trap, never score the row as 0.
                printf("b200k_cross_entropy_fwd: target %lld of row %lld is outside [0, %lld)\n", (long long)tg, (long long)row, (long long)vocab);
                __trap();
            }
            if (tg != ignore_index) {
                // -log p of the target, rounded through the storage dtype like the reference's log_softmax output
                nll = -V::rnd((V::load1(x, tg) - t.m) - logsum);
            }
            row_nll[row] = nll;
        }
        // the peers' slots are free for the next row only after the leading CTA has read them
        if (kCluster > 1) cg::this_cluster().sync();
        __syncthreads();
    }
}

// loss = round_T(round_T(sum_i nll_i) / n_valid)  (reduction = mean), round_T(sum) (sum).
// One CTA, fixed summation order.
template <typename T>
__global__ void __launch_bounds__(1024)
ce_finalize_kernel(const float* __restrict__ row_nll, const int64_t* __restrict__ target, T* __restrict__ loss,
                   int64_t* __restrict__ n_valid_out, int64_t rows, int64_t ignore_index, int reduction) {
    using V = Vec<T>;
    __shared__ float ssum[32];
    __shared__ long long scnt[32];
    float s = 0.f;
    long long c = 0;
    for (int64_t i = threadIdx.x; i < rows; i += blockDim.x) {
        s += row_nll[i];
        c += target[i] != ignore_index ? 1 : 0;
    }
    s = warp_reduce(s, SumOp{});
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
        ssum[warp] = s;
        scnt[warp] = c;
    }
    __syncthreads();
    if (warp == 0) {
        const int nw = blockDim.x >> 5;
        s = lane < nw ? ssum[lane] : 0.f;
        c = lane < nw ? scnt[lane] : 0;
        s = warp_reduce(s, SumOp{});
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
        if (lane == 0) {
            if (n_valid_out != nullptr) *n_valid_out = c;
            const float total = V::rnd(s);
            V::store1(loss, 0, reduction == 1 ? total / static_cast<float>(c) : total);
        }
    }
}

template <typename T>
__global__ void __launch_bounds__(kCeBwdThreads)
ce_bwd_kernel(const T* __restrict__ g, const T* __restrict__ logits, const int64_t* __restrict__ target,
              const float* __restrict__ row_max, const float* __restrict__ row_logsum, const int64_t* __restrict__ n_valid,
              T* __restrict__ dlogits, int64_t rows, int64_t vocab, int64_t ignore_index, int reduction, int vec_ok) {
    using V = Vec<T>;
    constexpr int kE = V::kElems;
    const int tid = threadIdx.x;
    const int64_t nvec = vec_ok ? vocab / kE : 0;
    // d = -(g / total_weight) evaluated in the storage dtype as ATen's nll_loss_backward does
    // (total_weight is the valid count converted to the storage dtype).
    const float gf = V::load1(g, 0);
    float d_all = gf;
    if (reduction == 1) d_all = V::rnd(gf / V::rnd(static_cast<float>(*n_valid)));
    d_all = -d_all;

    for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
        const T* x = logits + row * vocab;
        T* dx = dlogits + row * vocab;
        const int64_t tg = target[row];
        const bool valid = tg != ignore_index;
        if (valid && (tg < 0 || tg >= vocab)) {
            if (tid == 0) printf("b200k_cross_entropy_bwd: target %lld of row %lld is outside [0, %lld)\n", (long long)tg, (long long)row, (long long)vocab);
            __trap();
        }
        const float d = valid ? d_all : 0.f;
        const float m = row_max[row], ls = row_logsum[row];
        if (!valid) {
            // ignored row: gradient is exactly zero
            const uint4 z = make_uint4(0, 0, 0, 0);
            for (int64_t i = tid; i < nvec; i += kCeBwdThreads) st_stream(reinterpret_cast<uint4*>(dx) + i, z);
            for (int64_t j = nvec * kE + tid; j < vocab; j += kCeBwdThreads) V::store1(dx, j, 0.f);
            continue;
        }
        const int64_t tvec = tg / kE;
        for (int64_t base = tid; base < nvec; base += kCeBwdThreads * kCeBwdUnroll) {
            uint4 v[kCeBwdUnroll];
#pragma unroll
            for (int u = 0; u < kCeBwdUnroll; ++u) {
                const int64_t i = base + static_cast<int64_t>(u) * kCeBwdThreads;
                if (i < nvec) v[u] = ld_stream(reinterpret_cast<const uint4*>(x) + i);
            }
#pragma unroll
            for (int u = 0; u < kCeBwdUnroll; ++u) {
                const int64_t i = base + static_cast<int64_t>(u) * kCeBwdThreads;
                if (i < nvec) {
                    float f[kE], o[kE];
                    V::unpack(v[u], f);
#pragma unroll
                    for (int e = 0; e < kE; e += 2) {
                        float lp0 = (f[e] - m) - ls, lp1 = (f[e + 1] - m) - ls;
                        V::rnd2(lp0, lp1);  // log p is a storage-dtype tensor in the reference
                        o[e] = -(fast_exp2(lp0 * kLog2e) * d);
                        o[e + 1] = -(fast_exp2(lp1 * kLog2e) * d);
                    }
                    if (i == tvec) {
                        const int e = static_cast<int>(tg - tvec * kE);
#pragma unroll
                        for (int k = 0; k < kE; ++k)
                            if (k == e) o[k] = d + o[k];
                    }
                    st_stream(reinterpret_cast<uint4*>(dx) + i, V::pack(o));
                }
            }
        }
        for (int64_t j = nvec * kE + tid; j < vocab; j += kCeBwdThreads) {
            const float lp = V::rnd((V::load1(x, j) - m) - ls);
            float o = -(exp2f(lp * kLog2e) * d);
            if (j == tg) o = d + o;
            V::store1(dx, j, o);
        }
    }
}

template <typename T>
int run_fwd(const void* logits, const int64_t* target, float* row_max, float* row_logsum, float* row_nll, void* loss,
            int64_t* n_valid, int64_t rows, int64_t vocab, int64_t ignore_index, int reduction, cudaStream_t st) {
    constexpr int kE = Vec<T>::kElems;
    const int vec_ok = aligned16(logits) && vocab % kE == 0;
    if (rows > 0) {
        const int64_t cap = static_cast<int64_t>(sm_count()) * 4;
        // Few long rows: a 2- or 4-CTA cluster per row (DSMEM merge) instead of idle SMs.  B200K_CE_CLUSTER=1|2|4 pins it.
        int cluster = 1;
        if (vec_ok && vocab >= 32768) cluster = rows * 4 <= cap ? 4 : (rows * 2 <= cap ? 2 : 1);
        if (const char* e = getenv("B200K_CE_CLUSTER")) {
            const int v = atoi(e);
            if (vec_ok && (v == 1 || v == 2 || v == 4)) cluster = v;
        }
        if (cluster == 1) {
            const int grid = static_cast<int>(rows < cap ? rows : cap);
            ce_fwd_kernel<T, 1><<<grid, kCeThreads, 0, st>>>(static_cast<const T*>(logits), target, row_max, row_logsum, row_nll, rows,
                                                             vocab, ignore_index, vec_ok);
        } else {
            const int64_t want = rows * cluster, lim = cap / cluster * cluster;
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(static_cast<unsigned>(want < lim ? want : lim));
            cfg.blockDim = dim3(kCeThreads);
            cfg.stream = st;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.x = static_cast<unsigned>(cluster);
            attr[0].val.clusterDim.y = 1;
            attr[0].val.clusterDim.z = 1;
            cfg.attrs = attr;
            cfg.numAttrs = 1;
            const T* lg = static_cast<const T*>(logits);
            cudaError_t e = cluster == 2 ? cudaLaunchKernelEx(&cfg, ce_fwd_kernel<T, 2>, lg, target, row_max, row_logsum, row_nll, rows, vocab, ignore_index, vec_ok)
                                         : cudaLaunchKernelEx(&cfg, ce_fwd_kernel<T, 4>, lg, target, row_max, row_logsum, row_nll, rows, vocab, ignore_index, vec_ok);
            if (e != cudaSuccess) return static_cast<int>(e);
        }
        if (int rc = launch_status()) return rc;
    }
    if (loss != nullptr) {
        ce_finalize_kernel<T><<<1, 1024, 0, st>>>(row_nll, target, static_cast<T*>(loss), n_valid, rows, ignore_index, reduction);
        return launch_status();
    }
    return 0;
}

template <typename T>
int run_bwd(const void* g, const void* logits, const int64_t* target, const float* row_max, const float* row_logsum,
            const int64_t* n_valid, void* dlogits, int64_t rows, int64_t vocab, int64_t ignore_index, int reduction, cudaStream_t st) {
    constexpr int kE = Vec<T>::kElems;
    const int vec_ok = aligned16(logits) && aligned16(dlogits) && vocab % kE == 0;
    const int64_t cap = static_cast<int64_t>(sm_count()) * 4;
    const int grid = static_cast<int>(rows < cap ? rows : cap);
    ce_bwd_kernel<T><<<grid, kCeBwdThreads, 0, st>>>(static_cast<const T*>(g), static_cast<const T*>(logits), target, row_max, row_logsum,
                                                  n_valid, static_cast<T*>(dlogits), rows, vocab, ignore_index, reduction, vec_ok);
    return launch_status();
}

}  // namespace
}  // namespace b200k

using namespace b200k;

extern "C" int b200k_cross_entropy_fwd(const void* logits, const int64_t* target, float* row_max, float* row_logsum, float* row_nll,
                                       void* loss, int64_t* n_valid, int64_t rows, int64_t vocab, int64_t ignore_index,
                                       int reduction, int dtype, void* stream) {
    if (rows < 0 || vocab <= 0 || (reduction != 1 && reduction != 2)) return B200K_EINVAL;
    if (rows > 0 && (logits == nullptr || target == nullptr || row_max == nullptr || row_logsum == nullptr || row_nll == nullptr))
        return B200K_EINVAL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    switch (dtype) {
        case B200K_F32: return run_fwd<float>(logits, target, row_max, row_logsum, row_nll, loss, n_valid, rows, vocab, ignore_index, reduction, st);
        case B200K_BF16: return run_fwd<__nv_bfloat16>(logits, target, row_max, row_logsum, row_nll, loss, n_valid, rows, vocab, ignore_index, reduction, st);
        case B200K_F16: return run_fwd<__half>(logits, target, row_max, row_logsum, row_nll, loss, n_valid, rows, vocab, ignore_index, reduction, st);
        default: return B200K_EDTYPE;
    }
}

extern "C" int b200k_cross_entropy_bwd(const void* g, const void* logits, const int64_t* target, const float* row_max,
                                       const float* row_logsum, const int64_t* n_valid, void* dlogits, int64_t rows, int64_t vocab,
                                       int64_t ignore_index, int reduction, int dtype, void* stream) {
    if (rows < 0 || vocab <= 0 || (reduction != 1 && reduction != 2)) return B200K_EINVAL;
    if (rows == 0) return 0;
    if (g == nullptr || logits == nullptr || target == nullptr || row_max == nullptr || row_logsum == nullptr || dlogits == nullptr ||
        (reduction == 1 && n_valid == nullptr))
        return B200K_EINVAL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    switch (dtype) {
        case B200K_F32: return run_bwd<float>(g, logits, target, row_max, row_logsum, n_valid, dlogits, rows, vocab, ignore_index, reduction, st);
        case B200K_BF16: return run_bwd<__nv_bfloat16>(g, logits, target, row_max, row_logsum, n_valid, dlogits, rows, vocab, ignore_index, reduction, st);
        case B200K_F16: return run_bwd<__half>(g, logits, target, row_max, row_logsum, n_valid, dlogits, rows, vocab, ignore_index, reduction, st);
        default: return B200K_EDTYPE;
    }
}
