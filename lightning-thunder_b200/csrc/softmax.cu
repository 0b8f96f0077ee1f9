// softmax over the last dimension, forward and backward.  include/b200k.h states the math.
//
// Rows are [rows, cols] row-major.  One CTA owns one row at a time with the row in registers (thread t owns the
// 16-byte vectors t, t + blockDim.x, ...), the next row's loads are issued before the two reductions of the
// current one, every element is read once and written once.  The forward is what torch.softmax computes (fp32
// inside, one rounding); the backward follows the reference's trace op by op — mul, sum, sub, mul, each rounded
// to the storage dtype — and for 16-bit storage runs the three tensor x tensor steps on packed 2 x 16-bit
// arithmetic (Pk<T> in common.cuh), unpacking only for the fp32 row sum.
#include <cfloat>

#include "common.cuh"

namespace b200k {
namespace {

constexpr int sm_max_threads(int vpl) { return vpl >= 4 ? 512 : 256; }

template <typename Op>
__device__ __forceinline__ float sm_block_reduce(float v, float* smem, int nwarps, Op op, float identity) {
    v = warp_reduce(v, op);
    if (nwarps > 1) {
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        if (lane == 0) smem[warp] = v;
        __syncthreads();
        float x = lane < nwarps ? smem[lane] : identity;
        v = warp_reduce(x, op);
    }
    return v;
}

// exp(x) for x <= 0: bf16 storage rounds the result to 8 bits right away, so the 2-ulp hardware exp2 is used
// there (as in the cross-entropy kernels); fp16 / fp32 storage keep expf.
template <bool kFast>
__device__ __forceinline__ float exp_nonpos(float x) {
    if constexpr (kFast) {
        float y;
        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x * 1.4426950408889634f));
        return y;
    } else {
        return expf(x);
    }
}

template <typename T, int kVpl>
__global__ void __launch_bounds__(sm_max_threads(kVpl))
softmax_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t rows, int nvec) {
    using V = Vec<T>;
    constexpr int kE = V::kElems;
    constexpr bool kFast = V::kDtype == B200K_BF16;
    __shared__ float red[4][kWarp];
    const int tid = threadIdx.x, nthr = blockDim.x, nwarps = nthr >> 5;
    auto load_row = [&](int64_t row, uint4 (&dst)[kVpl]) {
        const uint4* xr = reinterpret_cast<const uint4*>(x) + row * nvec;
#pragma unroll
        for (int i = 0; i < kVpl; ++i) {
            const int v = tid + i * nthr;
            dst[i] = v < nvec && row < rows ? ld_stream(xr + v) : make_uint4(0, 0, 0, 0);
        }
    };
    constexpr bool kPrefetch = kVpl <= 4;
    uint4 xv[kVpl], nx[kPrefetch ? kVpl : 1];
    if constexpr (kPrefetch) load_row(blockIdx.x, xv);
    int parity = 0;
    for (int64_t row = blockIdx.x; row < rows; row += gridDim.x, parity ^= 2) {
        if constexpr (kPrefetch) load_row(row + gridDim.x, nx);
        else load_row(row, xv);
        float xf[kVpl][kE];
        float m = -FLT_MAX;
#pragma unroll
        for (int i = 0; i < kVpl; ++i) {
            V::unpack(xv[i], xf[i]);
            if (tid + i * nthr < nvec) {
#pragma unroll
                for (int e = 0; e < kE; ++e) m = fmaxf(m, xf[i][e]);
            }
        }
        m = sm_block_reduce(m, red[parity], nwarps, MaxOp{}, -FLT_MAX);
        // a row of -inf has no finite maximum: torch yields NaN there (exp(-inf - -inf)); keep that
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < kVpl; ++i) {
            const bool live = tid + i * nthr < nvec;
#pragma unroll
            for (int e = 0; e < kE; ++e) {
                xf[i][e] = exp_nonpos<kFast>(xf[i][e] - m);
                s += live ? xf[i][e] : 0.f;
            }
        }
        s = sm_block_reduce(s, red[parity + 1], nwarps, SumOp{}, 0.f);
        const float inv = __frcp_rn(s);
#pragma unroll
        for (int i = 0; i < kVpl; ++i) {
            const int v = tid + i * nthr;
#pragma unroll
            for (int e = 0; e < kE; ++e) xf[i][e] = kFast ? xf[i][e] * inv : xf[i][e] / s;
            if (v < nvec) st_stream(reinterpret_cast<uint4*>(y) + row * nvec + v, V::pack(xf[i]));
        }
        if constexpr (kPrefetch) {
#pragma unroll
            for (int i = 0; i < kVpl; ++i) xv[i] = nx[i];
        }
    }
}

// dx = rnd(y * rnd(g - rnd(sum_j rnd(y * g))))
template <typename T, int kVpl>
__global__ void __launch_bounds__(sm_max_threads(kVpl))
softmax_bwd_kernel(const T* __restrict__ g, const T* __restrict__ y, T* __restrict__ dx, int64_t rows, int nvec) {
    using V = Vec<T>;
    constexpr int kE = V::kElems;
    constexpr bool kPacked = kE == 8;
    __shared__ float red[2][kWarp];
    const int tid = threadIdx.x, nthr = blockDim.x, nwarps = nthr >> 5;
    auto load_row = [&](int64_t row, uint4 (&yd)[kVpl], uint4 (&gd)[kVpl]) {
#pragma unroll
        for (int i = 0; i < kVpl; ++i) {
            const int v = tid + i * nthr;
            const bool ok = v < nvec && row < rows;
            yd[i] = ok ? ld_stream(reinterpret_cast<const uint4*>(y) + row * nvec + v) : make_uint4(0, 0, 0, 0);
            gd[i] = ok ? ld_stream(reinterpret_cast<const uint4*>(g) + row * nvec + v) : make_uint4(0, 0, 0, 0);
        }
    };
    constexpr bool kPrefetch = kVpl <= 2;
    uint4 yv[kVpl], gv[kVpl], ny[kPrefetch ? kVpl : 1], ng[kPrefetch ? kVpl : 1];
    if constexpr (kPrefetch) load_row(blockIdx.x, yv, gv);
    int parity = 0;
    for (int64_t row = blockIdx.x; row < rows; row += gridDim.x, parity ^= 1) {
        if constexpr (kPrefetch) load_row(row + gridDim.x, ny, ng);
        else load_row(row, yv, gv);
        float acc = 0.f;
        uint4 pv[kVpl];  // rnd(y * g): packed words for 16-bit storage, fp32 bits otherwise
#pragma unroll
        for (int i = 0; i < kVpl; ++i) {
            float pf[kE];
            if constexpr (kPacked) {
                pv[i] = make_uint4(Pk<T>::mul(yv[i].x, gv[i].x), Pk<T>::mul(yv[i].y, gv[i].y), Pk<T>::mul(yv[i].z, gv[i].z), Pk<T>::mul(yv[i].w, gv[i].w));
                V::unpack(pv[i], pf);
            } else {
                float yf[kE], gf[kE];
                V::unpack(yv[i], yf);
                V::unpack(gv[i], gf);
#pragma unroll
                for (int e = 0; e < kE; ++e) pf[e] = __fmul_rn(yf[e], gf[e]);
            }
#pragma unroll
            for (int e = 0; e < kE; ++e) acc += pf[e];  // padding vectors are zeros
        }
        const float s = V::rnd(sm_block_reduce(acc, red[parity], nwarps, SumOp{}, 0.f));
#pragma unroll
        for (int i = 0; i < kVpl; ++i) {
            const int v = tid + i * nthr;
            uint4 o;
            if constexpr (kPacked) {
                float s2[2] = {s, s};
                float dummy[kE] = {s2[0], s2[1], s2[0], s2[1], s2[0], s2[1], s2[0], s2[1]};
                const uint32_t sw = V::pack(dummy).x;  // (s, s) in the storage dtype; exact, s is already rounded
                auto one = [&](uint32_t yw, uint32_t gw) { return Pk<T>::mul(yw, Pk<T>::sub(gw, sw)); };
                o = make_uint4(one(yv[i].x, gv[i].x), one(yv[i].y, gv[i].y), one(yv[i].z, gv[i].z), one(yv[i].w, gv[i].w));
            } else {
                float yf[kE], gf[kE], of[kE];
                V::unpack(yv[i], yf);
                V::unpack(gv[i], gf);
#pragma unroll
                for (int e = 0; e < kE; ++e) of[e] = __fmul_rn(yf[e], __fsub_rn(gf[e], s));
                o = V::pack(of);
            }
            if (v < nvec) st_stream(reinterpret_cast<uint4*>(dx) + row * nvec + v, o);
        }
        if constexpr (kPrefetch) {
#pragma unroll
            for (int i = 0; i < kVpl; ++i) yv[i] = ny[i], gv[i] = ng[i];
        }
    }
}

// Any `cols` (no alignment requirement, any width): CTA per row, several passes over the row (L1 / L2 hits).
template <typename T, bool kBwd>
__global__ void __launch_bounds__(256)
softmax_generic_kernel(const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ out, int64_t rows, int64_t cols) {
    using V = Vec<T>;
    constexpr bool kFast = V::kDtype == B200K_BF16;
    __shared__ float red[4][kWarp];
    int parity = 0;
    const int nwarps = blockDim.x >> 5;
    for (int64_t row = blockIdx.x; row < rows; row += gridDim.x, parity ^= 2) {
        const int64_t base = row * cols;
        if constexpr (!kBwd) {
            float m = -FLT_MAX;
            for (int64_t j = threadIdx.x; j < cols; j += blockDim.x) m = fmaxf(m, V::load1(a, base + j));
            m = sm_block_reduce(m, red[parity], nwarps, MaxOp{}, -FLT_MAX);
            float s = 0.f;
            for (int64_t j = threadIdx.x; j < cols; j += blockDim.x) s += exp_nonpos<kFast>(V::load1(a, base + j) - m);
            s = sm_block_reduce(s, red[parity + 1], nwarps, SumOp{}, 0.f);
            const float inv = __frcp_rn(s);
            for (int64_t j = threadIdx.x; j < cols; j += blockDim.x) {
                const float e = exp_nonpos<kFast>(V::load1(a, base + j) - m);
                V::store1(out, base + j, kFast ? e * inv : e / s);
            }
        } else {
            float acc = 0.f;  // a = g, b = y
            for (int64_t j = threadIdx.x; j < cols; j += blockDim.x) acc += V::rnd(__fmul_rn(V::load1(b, base + j), V::load1(a, base + j)));
            const float s = V::rnd(sm_block_reduce(acc, red[parity], nwarps, SumOp{}, 0.f));
            for (int64_t j = threadIdx.x; j < cols; j += blockDim.x)
                V::store1(out, base + j, __fmul_rn(V::load1(b, base + j), V::rnd(__fsub_rn(V::load1(a, base + j), s))));
        }
    }
}

template <typename T, bool kBwd>
int run_softmax(const void* a, const void* b, void* out, int64_t rows, int64_t cols, cudaStream_t st) {
    using V = Vec<T>;
    const bool vec_ok = cols % V::kElems == 0 && aligned16(a) && aligned16(out) && (!kBwd || aligned16(b));
    const int64_t nvec = cols / V::kElems;
    int vpl = 0, threads = 0;
    if (vec_ok) {
        for (int v : {1, 2, 4, 8}) {
            if (v < 4 && nvec > static_cast<int64_t>(kWarp) * v) continue;  // wide rows: >= 4 vectors in flight per thread
            int64_t thr = ((nvec + v - 1) / v + kWarp - 1) / kWarp * kWarp;
            if (thr <= sm_max_threads(v)) {
                vpl = v, threads = static_cast<int>(thr);
                break;
            }
        }
        if (kBwd && vpl == 8) vpl = 0;  // two row tensors in registers: the widest rows take the generic kernel
    }
    auto grid_for = [&](auto kernel, int thr) {
        int per_sm = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, thr, 0) != cudaSuccess || per_sm < 1) per_sm = 1;
        const int64_t cap = static_cast<int64_t>(sm_count()) * per_sm;
        if (rows <= cap) return static_cast<int>(rows);
        const int64_t per_cta = (rows + cap - 1) / cap;
        return static_cast<int>((rows + per_cta - 1) / per_cta);
    };
    auto go = [&](auto kernel) {
        const int grid = grid_for(kernel, threads);
        if constexpr (kBwd)
            kernel<<<grid, threads, 0, st>>>(static_cast<const T*>(a), static_cast<const T*>(b), static_cast<T*>(out), rows, static_cast<int>(nvec));
        else
            kernel<<<grid, threads, 0, st>>>(static_cast<const T*>(a), static_cast<T*>(out), rows, static_cast<int>(nvec));
        return launch_status();
    };
    if constexpr (kBwd) {
        switch (vpl) {
            case 1: return go(softmax_bwd_kernel<T, 1>);
            case 2: return go(softmax_bwd_kernel<T, 2>);
            case 4: return go(softmax_bwd_kernel<T, 4>);
            default: break;
        }
    } else {
        switch (vpl) {
            case 1: return go(softmax_fwd_kernel<T, 1>);
            case 2: return go(softmax_fwd_kernel<T, 2>);
            case 4: return go(softmax_fwd_kernel<T, 4>);
            case 8: return go(softmax_fwd_kernel<T, 8>);
            default: break;
        }
    }
    const int64_t cap = static_cast<int64_t>(sm_count()) * 8;
    softmax_generic_kernel<T, kBwd><<<static_cast<int>(rows < cap ? rows : cap), 256, 0, st>>>(
        static_cast<const T*>(a), static_cast<const T*>(b), static_cast<T*>(out), rows, cols);
    return launch_status();
}

}  // namespace
}  // namespace b200k

using namespace b200k;

extern "C" int b200k_softmax_fwd(const void* x, void* y, int64_t rows, int64_t cols, int dtype, void* stream) {
    if (rows < 0 || cols <= 0) return B200K_EINVAL;
    if (rows == 0) return 0;
    if (x == nullptr || y == nullptr) return B200K_EINVAL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    switch (dtype) {
        case B200K_F32: return run_softmax<float, false>(x, nullptr, y, rows, cols, st);
        case B200K_BF16: return run_softmax<__nv_bfloat16, false>(x, nullptr, y, rows, cols, st);
        case B200K_F16: return run_softmax<__half, false>(x, nullptr, y, rows, cols, st);
        default: return B200K_EDTYPE;
    }
}

extern "C" int b200k_softmax_bwd(const void* dy, const void* y, void* dx, int64_t rows, int64_t cols, int dtype, void* stream) {
    if (rows < 0 || cols <= 0) return B200K_EINVAL;
    if (rows == 0) return 0;
    if (dy == nullptr || y == nullptr || dx == nullptr) return B200K_EINVAL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    switch (dtype) {
        case B200K_F32: return run_softmax<float, true>(dy, y, dx, rows, cols, st);
        case B200K_BF16: return run_softmax<__nv_bfloat16, true>(dy, y, dx, rows, cols, st);
        case B200K_F16: return run_softmax<__half, true>(dy, y, dx, rows, cols, st);
        default: return B200K_EDTYPE;
    }
}
