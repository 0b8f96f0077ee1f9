// RMSNorm (+ residual add) forward and backward for sm_100a.  See include/b200k.h for the math,
// which follows the reference's torch-executor trace op for op.
//
// Layout: activations are [rows, hidden] row-major.  A CTA of `blockDim.x` threads owns one row
// at a time; thread t owns the 16-byte vectors t, t + blockDim.x, ... (kVpl of them), so every
// warp-level access is a run of 512 contiguous bytes.  Rows are streamed exactly once: x, r -> h, y
// forward; dy, h, dh_in -> dx backward.  The backward keeps per-thread fp32 partial sums of dw for
// the columns it owns across all rows the (persistent) CTA processes, writes one partial row per
// CTA to scratch, and a second tiny kernel folds the partials (deterministic, no atomics).
#include <cstdlib>
#include <mutex>

#include "common.cuh"
#include "reduce.cuh"

namespace b200k {
namespace {

constexpr int fwd_max_threads(int vpl) { return vpl >= 4 ? 512 : 256; }
constexpr int kBwdMaxThreads = 256;
constexpr int kBwdCtasPerSm = 4;  // upper bound on resident backward CTAs per SM (sizes the scratch)

__device__ __forceinline__ float block_sum(float v, float* smem, int nwarps) {
    v = warp_reduce(v, SumOp{});
    if (nwarps > 1) {
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        if (lane == 0) smem[warp] = v;
        __syncthreads();
        float x = lane < nwarps ? smem[lane] : 0.f;
        v = warp_reduce(x, SumOp{});
    }
    return v;
}

// -------------------------------------------------------------------------------------------
// Forward.
// -------------------------------------------------------------------------------------------
template <typename T, int kVpl, bool kHasRes, bool kHasW>
__global__ void __launch_bounds__(fwd_max_threads(kVpl))
rmsnorm_fwd_kernel(const T* __restrict__ x, const T* __restrict__ res, const T* __restrict__ w,
                   T* __restrict__ h_out, T* __restrict__ y, float* __restrict__ rstd_out,
                   int64_t rows, int nvec, float inv_hidden, float eps) {
    using V = Vec<T>;
    constexpr int kE = V::kElems;
    __shared__ float red[2][kWarp];
    const int tid = threadIdx.x, nthr = blockDim.x, nwarps = nthr >> 5;

    // Weight vectors for the columns this thread owns stay in registers for the CTA's lifetime.
    uint4 wv[kVpl];
    if constexpr (kHasW) {
#pragma unroll
        for (int i = 0; i < kVpl; ++i) {
            const int v = tid + i * nthr;
            wv[i] = v < nvec ? ld_keep(reinterpret_cast<const uint4*>(w) + v) : make_uint4(0, 0, 0, 0);
        }
    }

    int parity = 0;
    for (int64_t row = blockIdx.x; row < rows; row += gridDim.x, parity ^= 1) {
        const uint4* xr = reinterpret_cast<const uint4*>(x) + row * nvec;
        uint4 xv[kVpl], rv[kVpl];
#pragma unroll
        for (int i = 0; i < kVpl; ++i) {
            const int v = tid + i * nthr;
            xv[i] = v < nvec ? ld_stream(xr + v) : make_uint4(0, 0, 0, 0);
        }
        if constexpr (kHasRes) {
            const uint4* rr = reinterpret_cast<const uint4*>(res) + row * nvec;
#pragma unroll
            for (int i = 0; i < kVpl; ++i) {
                const int v = tid + i * nthr;
                rv[i] = v < nvec ? ld_stream(rr + v) : make_uint4(0, 0, 0, 0);
            }
        }

        float hf[kVpl][kE];
        float ss = 0.f;
#pragma unroll
        for (int i = 0; i < kVpl; ++i) {
            V::unpack(xv[i], hf[i]);
            if constexpr (kHasRes) {
                float rf[kE];
                V::unpack(rv[i], rf);
#pragma unroll
                for (int e = 0; e < kE; ++e) hf[i][e] = V::rnd(hf[i][e] + rf[e]);  // h is a trace-level tensor of T
                const int v = tid + i * nthr;
                if (v < nvec) st_stream(reinterpret_cast<uint4*>(h_out) + row * nvec + v, V::pack(hf[i]));
            }
#pragma unroll
            for (int e = 0; e < kE; ++e) ss += hf[i][e] * hf[i][e];
        }

        ss = block_sum(ss, red[parity], nwarps);
        const float rstd = rsqrtf(ss * inv_hidden + eps);
        if (rstd_out != nullptr && tid == 0) rstd_out[row] = rstd;

#pragma unroll
        for (int i = 0; i < kVpl; ++i) {
            const int v = tid + i * nthr;
            float wf[kE];
            if constexpr (kHasW) V::unpack(wv[i], wf);
#pragma unroll
            for (int e = 0; e < kE; ++e) {
                float n = hf[i][e] * rstd;
                if constexpr (kHasW) n *= wf[e];
                hf[i][e] = n;
            }
            if (v < nvec) st_stream(reinterpret_cast<uint4*>(y) + row * nvec + v, V::pack(hf[i]));
        }
    }
}

// Any `hidden` (no alignment requirement): CTA per row, two passes over the row (second pass hits L2).
template <typename T, bool kHasRes, bool kHasW>
__global__ void __launch_bounds__(256)
rmsnorm_fwd_generic_kernel(const T* __restrict__ x, const T* __restrict__ res, const T* __restrict__ w,
                           T* __restrict__ h_out, T* __restrict__ y, float* __restrict__ rstd_out,
                           int64_t rows, int64_t hidden, float eps) {
    using V = Vec<T>;
    __shared__ float red[2][kWarp];
    int parity = 0;
    for (int64_t row = blockIdx.x; row < rows; row += gridDim.x, parity ^= 1) {
        const int64_t base = row * hidden;
        float ss = 0.f;
        for (int64_t j = threadIdx.x; j < hidden; j += blockDim.x) {
            float h = V::load1(x, base + j);
            if constexpr (kHasRes) {
                h = V::rnd(h + V::load1(res, base + j));
                V::store1(h_out, base + j, h);
            }
            ss += h * h;
        }
        ss = block_sum(ss, red[parity], blockDim.x >> 5);
        const float rstd = rsqrtf(ss / static_cast<float>(hidden) + eps);
        if (rstd_out != nullptr && threadIdx.x == 0) rstd_out[row] = rstd;
        for (int64_t j = threadIdx.x; j < hidden; j += blockDim.x) {
            float h = V::load1(x, base + j);
            if constexpr (kHasRes) h = V::rnd(h + V::load1(res, base + j));
            float n = h * rstd;
            if constexpr (kHasW) n *= V::load1(w, j);
            V::store1(y, base + j, n);
        }
    }
}

// -------------------------------------------------------------------------------------------
// Backward.
// -------------------------------------------------------------------------------------------
template <typename T, int kVpl, int kRows, bool kHasW, bool kHasDhIn>
__global__ void __launch_bounds__(kBwdMaxThreads)
rmsnorm_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ h, const T* __restrict__ w,
                   const float* __restrict__ rstd, const T* __restrict__ dh_in, T* __restrict__ dx,
                   float* __restrict__ dw_partial, int64_t rows, int nvec, float hidden_f) {
    using V = Vec<T>;
    constexpr int kE = V::kElems;
    constexpr bool kIsF32 = V::kDtype == B200K_F32;
    __shared__ float red[2][kRows][kWarp];
    const int tid = threadIdx.x, nthr = blockDim.x, nwarps = nthr >> 5;
    const int lane = tid & 31, warp = tid >> 5;

    uint4 wv[kVpl];
    float dwacc[kVpl][kE];
#pragma unroll
    for (int i = 0; i < kVpl; ++i) {
        const int v = tid + i * nthr;
        if constexpr (kHasW) wv[i] = v < nvec ? ld_keep(reinterpret_cast<const uint4*>(w) + v) : make_uint4(0, 0, 0, 0);
#pragma unroll
        for (int e = 0; e < kE; ++e) dwacc[i][e] = 0.f;
    }

    int parity = 0;
    for (int64_t row0 = static_cast<int64_t>(blockIdx.x) * kRows; row0 < rows;
         row0 += static_cast<int64_t>(gridDim.x) * kRows, parity ^= 1) {
        uint4 dyv[kRows][kVpl], hv[kRows][kVpl], div[kRows][kVpl];
        float rs[kRows];
#pragma unroll
        for (int r = 0; r < kRows; ++r) {
            const int64_t row = row0 + r;
            const bool live = row < rows;
            rs[r] = live ? rstd[row] : 0.f;
#pragma unroll
            for (int i = 0; i < kVpl; ++i) {
                const int v = tid + i * nthr;
                const bool ok = live && v < nvec;
                const int64_t off = row * nvec + v;
                dyv[r][i] = ok ? ld_stream(reinterpret_cast<const uint4*>(dy) + off) : make_uint4(0, 0, 0, 0);
                hv[r][i] = ok ? ld_stream(reinterpret_cast<const uint4*>(h) + off) : make_uint4(0, 0, 0, 0);
                if constexpr (kHasDhIn)
                    div[r][i] = ok ? ld_stream(reinterpret_cast<const uint4*>(dh_in) + off) : make_uint4(0, 0, 0, 0);
            }
        }

        // c[r] = sum_j h * (w * dy)
        float c[kRows];
#pragma unroll
        for (int r = 0; r < kRows; ++r) {
            float acc = 0.f;
#pragma unroll
            for (int i = 0; i < kVpl; ++i) {
                float dyf[kE], hf[kE], wf[kE];
                V::unpack(dyv[r][i], dyf);
                V::unpack(hv[r][i], hf);
                if constexpr (kHasW) V::unpack(wv[i], wf);
#pragma unroll
                for (int e = 0; e < kE; ++e) {
                    const float g = kHasW ? wf[e] * dyf[e] : dyf[e];
                    acc += hf[e] * g;
                }
            }
            c[r] = warp_reduce(acc, SumOp{});
        }
        if (nwarps > 1) {
            if (lane == 0) {
#pragma unroll
                for (int r = 0; r < kRows; ++r) red[parity][r][warp] = c[r];
            }
            __syncthreads();
#pragma unroll
            for (int r = 0; r < kRows; ++r) {
                float xsum = lane < nwarps ? red[parity][r][lane] : 0.f;
                c[r] = warp_reduce(xsum, SumOp{});
            }
        }

#pragma unroll
        for (int r = 0; r < kRows; ++r) {
            const int64_t row = row0 + r;
            if (row >= rows) break;
            const float rstd_r = rs[r];
            // s = ((-0.5 * c) * rstd^3) / hidden   (op order of the reference's backward trace)
            const float s = ((-0.5f * c[r]) * (rstd_r * rstd_r * rstd_r)) / hidden_f;
#pragma unroll
            for (int i = 0; i < kVpl; ++i) {
                const int v = tid + i * nthr;
                float dyf[kE], hf[kE], wf[kE], o[kE];
                V::unpack(dyv[r][i], dyf);
                V::unpack(hv[r][i], hf);
                if constexpr (kHasW) V::unpack(wv[i], wf);
#pragma unroll
                for (int e = 0; e < kE; ++e) {
                    const float g = kHasW ? wf[e] * dyf[e] : dyf[e];
                    const float hs = hf[e] * s;
                    o[e] = rstd_r * g;
                    o[e + 0] = kHasDhIn && kIsF32 ? o[e] : (o[e] + hs) + hs;
                    if constexpr (kHasW) dwacc[i][e] += V::rnd((hf[e] * rstd_r) * dyf[e]);
                }
                if constexpr (kHasDhIn) {
                    float df[kE];
                    V::unpack(div[r][i], df);
#pragma unroll
                    for (int e = 0; e < kE; ++e) {
                        if constexpr (kIsF32) {
                            // fp32 has no rounding boundary: the reference accumulates dh_in first.
                            const float hs = hf[e] * s;
                            o[e] = ((df[e] + o[e]) + hs) + hs;
                        } else {
                            o[e] = df[e] + V::rnd(o[e]);
                        }
                    }
                }
                if (v < nvec) st_stream(reinterpret_cast<uint4*>(dx) + row * nvec + v, V::pack(o));
            }
        }
    }

    if constexpr (kHasW) {
        // One fp32 partial row per CTA; float4 stores (kE floats = kE/4 float4 per vector).
        float* prow = dw_partial + static_cast<int64_t>(blockIdx.x) * nvec * kE;
#pragma unroll
        for (int i = 0; i < kVpl; ++i) {
            const int v = tid + i * nthr;
            if (v < nvec) {
#pragma unroll
                for (int q = 0; q < kE / 4; ++q) {
                    reinterpret_cast<float4*>(prow + static_cast<int64_t>(v) * kE)[q] =
                        make_float4(dwacc[i][4 * q], dwacc[i][4 * q + 1], dwacc[i][4 * q + 2], dwacc[i][4 * q + 3]);
                }
            }
        }
    }
}

// -------------------------------------------------------------------------------------------
// Backward, TMA-staged: the streaming variant used whenever a row group fits shared memory and
// blockDim.x * kVpl == nvec exactly (no column predicates in the hot loop).
//
// A ring of `stages` row groups is kept in flight with 1-D bulk copies (cp.async.bulk -> mbarrier
// complete_tx), so the bytes in flight per SM are set by the ring (up to ~200 KB), not by registers
// or occupancy.  Thread t owns the 16-byte vectors t, t + blockDim.x, ...; it reads them from shared
// memory, hands the slot back (one mbarrier arrive per warp) and thread 0 refills it right after the
// CTA-wide reduction barrier, at which point every warp is known to have released it.  dx goes
// straight from registers to global memory.
// Stage layout: [dy rows | h rows | dh_in rows], each kRows * row_bytes.
// -------------------------------------------------------------------------------------------
template <typename T, int kVpl, int kRows, bool kHasW, bool kHasDhIn>
__global__ void __launch_bounds__(512 / kVpl, kVpl >= 2 ? 2 : 1)
rmsnorm_bwd_tma_kernel(const T* __restrict__ dy, const T* __restrict__ h, const T* __restrict__ w,
                       const float* __restrict__ rstd, const T* __restrict__ dh_in, T* __restrict__ dx,
                       float* __restrict__ dw_partial, int64_t rows, int nvec, float hidden_f, float inv_hidden_exact,
                       int stages) {
    using V = Vec<T>;
    constexpr int kE = V::kElems;
    constexpr bool kIsF32 = V::kDtype == B200K_F32;
    constexpr int kTensors = kHasDhIn ? 3 : 2;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const uint32_t row_bytes = static_cast<uint32_t>(nvec) * 16u;
    const uint32_t group_bytes = row_bytes * kRows;          // one tensor's rows in a stage
    const uint32_t stage_bytes = group_bytes * kTensors;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw);  // [stages]
    uint64_t* empty = full + stages;                         // [stages]
    float* red = reinterpret_cast<float*>(empty + stages);   // [2][kRows][32]
    unsigned char* ring = smem_raw + ((2 * stages * 8 + 2 * kRows * 32 * 4 + 127) / 128) * 128;

    const int nc = blockDim.x;
    const int ncw = nc >> 5;
    const int tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5;
    const int64_t ngroups = (rows + kRows - 1) / kRows;
    uint64_t pol = 0;

    auto issue = [&](int64_t g, int s) {
        const int64_t row0 = g * kRows;
        const int64_t live = rows - row0 < kRows ? rows - row0 : kRows;
        const uint32_t bytes = static_cast<uint32_t>(live) * row_bytes;
        unsigned char* st = ring + static_cast<size_t>(s) * stage_bytes;
        mbar_arrive_expect_tx(&full[s], bytes * kTensors);
        const size_t goff = static_cast<size_t>(row0) * row_bytes;
        bulk_g2s(st, reinterpret_cast<const unsigned char*>(dy) + goff, bytes, &full[s], pol);
        bulk_g2s(st + group_bytes, reinterpret_cast<const unsigned char*>(h) + goff, bytes, &full[s], pol);
        if constexpr (kHasDhIn)
            bulk_g2s(st + 2 * group_bytes, reinterpret_cast<const unsigned char*>(dh_in) + goff, bytes, &full[s], pol);
    };

    if (tid == 0) {
        for (int s = 0; s < stages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], ncw);
        }
        mbar_fence_init();
        pol = policy_evict_first();
        for (int s = 0; s < stages; ++s) {
            const int64_t g = blockIdx.x + static_cast<int64_t>(s) * gridDim.x;
            if (g < ngroups) issue(g, s);
        }
    }
    __syncthreads();

    // All arithmetic below is on float2 pairs: sm_100's packed FMUL2 / FADD2 / FFMA2 halve the fp32 instruction
    // count of this issue-bound kernel, with results identical to the scalar round-to-nearest operations.
    constexpr int kP = kE / 2;
    float2 wf[kVpl][kP];
    float2 dwacc[kVpl][kP];
#pragma unroll
    for (int i = 0; i < kVpl; ++i) {
        if constexpr (kHasW) V::unpack2(ld_keep(reinterpret_cast<const uint4*>(w) + tid + i * nc), wf[i]);
#pragma unroll
        for (int e = 0; e < kP; ++e) dwacc[i][e] = make_float2(0.f, 0.f);
    }

    // Let the dependent dw-reduce grid be scheduled early; it blocks in griddepcontrol.wait until this grid is done.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    int s = 0;
    uint32_t phase = 0;
    int parity = 0;
    for (int64_t g = blockIdx.x; g < ngroups; g += gridDim.x, parity ^= 1) {
        const int64_t row0 = g * kRows;
        float rs[kRows];
#pragma unroll
        for (int r = 0; r < kRows; ++r) rs[r] = row0 + r < rows ? rstd[row0 + r] : 0.f;

        mbar_wait(&full[s], phase);
        const unsigned char* st = ring + static_cast<size_t>(s) * stage_bytes;

        uint4 dyv[kRows][kVpl], hv[kRows][kVpl], div[kRows][kVpl];
#pragma unroll
        for (int r = 0; r < kRows; ++r) {
#pragma unroll
            for (int i = 0; i < kVpl; ++i) {
                const uint32_t off = static_cast<uint32_t>(r) * row_bytes + static_cast<uint32_t>(tid + i * nc) * 16u;
                dyv[r][i] = lds128(st + off);
                hv[r][i] = lds128(st + group_bytes + off);
                if constexpr (kHasDhIn) div[r][i] = lds128(st + 2 * group_bytes + off);
            }
        }
        // Everything this warp needs from the stage is now in registers: hand the slot back.
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);

        float c[kRows];
#pragma unroll
        for (int r = 0; r < kRows; ++r) {
            float2 acc = make_float2(0.f, 0.f);
#pragma unroll
            for (int i = 0; i < kVpl; ++i) {
                float2 dyf[kP], hf[kP];
                V::unpack2(dyv[r][i], dyf);
                V::unpack2(hv[r][i], hf);
#pragma unroll
                for (int e = 0; e < kP; ++e) acc = __ffma2_rn(hf[e], kHasW ? __fmul2_rn(wf[i][e], dyf[e]) : dyf[e], acc);
            }
            c[r] = warp_reduce(acc.x + acc.y, SumOp{});
        }
        if (ncw > 1) {
            float* rb = red + parity * (kRows * 32);
            if (lane == 0) {
#pragma unroll
                for (int r = 0; r < kRows; ++r) rb[r * 32 + warp] = c[r];
            }
            __syncthreads();
#pragma unroll
            for (int r = 0; r < kRows; ++r) {
                float xsum = lane < ncw ? rb[r * 32 + lane] : 0.f;
                c[r] = warp_reduce<16>(xsum, SumOp{});
                c[r] = __shfl_sync(0xffffffffu, c[r], 0);
            }
        }
        // Every warp has released slot s (its arrive precedes the barrier above): refill it.
        if (tid == 0) {
            const int64_t gn = g + static_cast<int64_t>(stages) * gridDim.x;
            if (gn < ngroups) {
                mbar_wait(&empty[s], phase);
                issue(gn, s);
            }
        }
        if (++s == stages) {
            s = 0;
            phase ^= 1u;
        }

#pragma unroll
        for (int r = 0; r < kRows; ++r) {
            const int64_t row = row0 + r;
            if (row >= rows) break;
            const float rstd_r = rs[r];
            const float t3 = (-0.5f * c[r]) * (rstd_r * rstd_r * rstd_r);
            // t3 / hidden; a multiply by the exact reciprocal when hidden is a power of two (bit-identical).
            const float sc = inv_hidden_exact != 0.f ? t3 * inv_hidden_exact : t3 / hidden_f;
            const float2 rstd2 = make_float2(rstd_r, rstd_r), sc2 = make_float2(sc, sc);
#pragma unroll
            for (int i = 0; i < kVpl; ++i) {
                float2 dyf[kP], hf[kP], o[kP];
                V::unpack2(dyv[r][i], dyf);
                V::unpack2(hv[r][i], hf);
#pragma unroll
                for (int e = 0; e < kP; ++e) {
                    const float2 gg = kHasW ? __fmul2_rn(wf[i][e], dyf[e]) : dyf[e];
                    const float2 hs = __fmul2_rn(hf[e], sc2);
                    o[e] = __fmul2_rn(rstd2, gg);
                    if constexpr (!(kHasDhIn && kIsF32)) o[e] = __fadd2_rn(__fadd2_rn(o[e], hs), hs);
                    if constexpr (kHasW) dwacc[i][e] = __fadd2_rn(dwacc[i][e], V::rnd2f(__fmul2_rn(__fmul2_rn(hf[e], rstd2), dyf[e])));
                }
                if constexpr (kHasDhIn && kIsF32) {
                    float2 df[kP];
                    V::unpack2(div[r][i], df);
#pragma unroll
                    for (int e = 0; e < kP; ++e) {
                        // fp32 has no rounding boundary: the reference accumulates dh_in first
                        const float2 hs = __fmul2_rn(hf[e], sc2);
                        o[e] = __fadd2_rn(__fadd2_rn(__fadd2_rn(df[e], o[e]), hs), hs);
                    }
                }
                if constexpr (kHasDhIn && !kIsF32) {
                    // 16-bit: round(dh_in + round(o)) is one packed convert and one packed add on the words as loaded
                    // (the packed add rounds the exact sum once, which equals the fp32 add followed by the rounding)
                    const uint32_t dw_[4] = {div[r][i].x, div[r][i].y, div[r][i].z, div[r][i].w};
                    uint32_t ow[4] = {0u, 0u, 0u, 0u};
#pragma unroll
                    for (int e = 0; e < kP; ++e) ow[e] = Pk<T>::add(dw_[e], V::pack_word(o[e]));
                    st_stream(reinterpret_cast<uint4*>(dx) + row * nvec + tid + i * nc, make_uint4(ow[0], ow[1], ow[2], ow[3]));
                } else {
                    st_stream(reinterpret_cast<uint4*>(dx) + row * nvec + tid + i * nc, V::pack2(o));
                }
            }
        }
    }

    if constexpr (kHasW) {
        float* prow = dw_partial + static_cast<int64_t>(blockIdx.x) * nvec * kE;
#pragma unroll
        for (int i = 0; i < kVpl; ++i) {
            const int v = tid + i * nc;
#pragma unroll
            for (int q = 0; q < kE / 4; ++q) {
                reinterpret_cast<float4*>(prow + static_cast<int64_t>(v) * kE)[q] =
                    make_float4(dwacc[i][2 * q].x, dwacc[i][2 * q].y, dwacc[i][2 * q + 1].x, dwacc[i][2 * q + 1].y);
            }
        }
    }
}

// Generic backward (any hidden): CTA per row for dx, then a column kernel for dw that re-reads h, dy.
template <typename T, bool kHasW, bool kHasDhIn>
__global__ void __launch_bounds__(256)
rmsnorm_bwd_generic_dx_kernel(const T* __restrict__ dy, const T* __restrict__ h, const T* __restrict__ w,
                              const float* __restrict__ rstd, const T* __restrict__ dh_in, T* __restrict__ dx,
                              int64_t rows, int64_t hidden) {
    using V = Vec<T>;
    __shared__ float red[2][kWarp];
    int parity = 0;
    for (int64_t row = blockIdx.x; row < rows; row += gridDim.x, parity ^= 1) {
        const int64_t base = row * hidden;
        float acc = 0.f;
        for (int64_t j = threadIdx.x; j < hidden; j += blockDim.x) {
            const float g = kHasW ? V::load1(w, j) * V::load1(dy, base + j) : V::load1(dy, base + j);
            acc += V::load1(h, base + j) * g;
        }
        const float c = block_sum(acc, red[parity], blockDim.x >> 5);
        const float rstd_r = rstd[row];
        const float s = ((-0.5f * c) * (rstd_r * rstd_r * rstd_r)) / static_cast<float>(hidden);
        for (int64_t j = threadIdx.x; j < hidden; j += blockDim.x) {
            const float g = kHasW ? V::load1(w, j) * V::load1(dy, base + j) : V::load1(dy, base + j);
            const float hs = V::load1(h, base + j) * s;
            float o = rstd_r * g;
            if constexpr (kHasDhIn && V::kDtype == B200K_F32) {
                o = ((V::load1(dh_in, base + j) + o) + hs) + hs;
            } else {
                o = (o + hs) + hs;
                if constexpr (kHasDhIn) o = V::load1(dh_in, base + j) + V::rnd(o);
            }
            V::store1(dx, base + j, o);
        }
    }
}

template <typename T>
__global__ void __launch_bounds__(256)
rmsnorm_bwd_generic_dw_kernel(const T* __restrict__ dy, const T* __restrict__ h, const float* __restrict__ rstd,
                              T* __restrict__ dw, int64_t rows, int64_t hidden) {
    using V = Vec<T>;
    __shared__ float tile[8][kWarp + 1];
    const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
    const int64_t col = static_cast<int64_t>(blockIdx.x) * 32 + cx;
    float acc = 0.f;
    if (col < hidden) {
        for (int64_t r = ry; r < rows; r += 8)
            acc += V::rnd((V::load1(h, r * hidden + col) * rstd[r]) * V::load1(dy, r * hidden + col));
    }
    tile[ry][cx] = acc;
    __syncthreads();
    if (ry == 0 && col < hidden) {
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) s += tile[k][cx];
        V::store1(dw, col, s);
    }
}

// -------------------------------------------------------------------------------------------
// Host-side dispatch.
// -------------------------------------------------------------------------------------------
struct RowPlan {
    int vpl;      // 16-byte vectors per thread
    int threads;  // CTA size (multiple of 32, <= 1024)
};

// Pick (vpl, threads) so that threads * vpl covers nvec with >= 2 vectors in flight per thread
// for wide rows.  Returns vpl = 0 when the row does not fit the register-resident kernels.
RowPlan plan_row(int64_t nvec, int min_vpl, bool fwd) {
    for (int vpl : {1, 2, 4, 8}) {
        if (vpl < min_vpl && nvec > static_cast<int64_t>(kWarp) * vpl) continue;
        int64_t thr = (nvec + vpl - 1) / vpl;
        thr = (thr + kWarp - 1) / kWarp * kWarp;
        const int64_t cap = fwd ? fwd_max_threads(vpl) : kBwdMaxThreads;
        if (thr <= cap) return {vpl, static_cast<int>(thr)};
    }
    return {0, 0};
}

template <typename Kernel>
int persistent_grid(Kernel kernel, int threads, int64_t work_items, int max_ctas_per_sm) {
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, 0) != cudaSuccess || per_sm < 1) per_sm = 1;
    if (per_sm > max_ctas_per_sm) per_sm = max_ctas_per_sm;
    const int64_t cap = static_cast<int64_t>(sm_count()) * per_sm;
    return static_cast<int>(work_items < cap ? work_items : cap);
}

template <typename T, bool kHasRes, bool kHasW>
int launch_fwd(const void* x, const void* res, const void* w, void* h_out, void* y, float* rstd, int64_t rows,
               int64_t hidden, float eps, cudaStream_t st) {
    using V = Vec<T>;
    const bool vec_ok = hidden % V::kElems == 0 && aligned16(x) && aligned16(y) && (!kHasRes || (aligned16(res) && aligned16(h_out))) &&
                        (!kHasW || aligned16(w));
    const int64_t nvec = hidden / V::kElems;
    const RowPlan p = vec_ok ? plan_row(nvec, 4, true) : RowPlan{0, 0};
    auto go = [&](auto kernel) {
        const int grid = persistent_grid(kernel, p.threads, rows, 32);
        kernel<<<grid, p.threads, 0, st>>>(static_cast<const T*>(x), static_cast<const T*>(res), static_cast<const T*>(w),
                                           static_cast<T*>(h_out), static_cast<T*>(y), rstd, rows, static_cast<int>(nvec),
                                           1.0f / static_cast<float>(hidden), eps);
        return launch_status();
    };
    switch (p.vpl) {
        case 1: return go(rmsnorm_fwd_kernel<T, 1, kHasRes, kHasW>);
        case 2: return go(rmsnorm_fwd_kernel<T, 2, kHasRes, kHasW>);
        case 4: return go(rmsnorm_fwd_kernel<T, 4, kHasRes, kHasW>);
        case 8: return go(rmsnorm_fwd_kernel<T, 8, kHasRes, kHasW>);
        default: break;
    }
    const int64_t cap = static_cast<int64_t>(sm_count()) * 8;
    const int grid = static_cast<int>(rows < cap ? rows : cap);
    rmsnorm_fwd_generic_kernel<T, kHasRes, kHasW><<<grid, 256, 0, st>>>(
        static_cast<const T*>(x), static_cast<const T*>(res), static_cast<const T*>(w), static_cast<T*>(h_out),
        static_cast<T*>(y), rstd, rows, hidden, eps);
    return launch_status();
}

template <typename T>
int dispatch_fwd(const void* x, const void* res, const void* w, void* h_out, void* y, float* rstd, int64_t rows,
                 int64_t hidden, float eps, cudaStream_t st) {
    if (res != nullptr) {
        return w != nullptr ? launch_fwd<T, true, true>(x, res, w, h_out, y, rstd, rows, hidden, eps, st)
                            : launch_fwd<T, true, false>(x, res, w, h_out, y, rstd, rows, hidden, eps, st);
    }
    return w != nullptr ? launch_fwd<T, false, true>(x, res, w, h_out, y, rstd, rows, hidden, eps, st)
                        : launch_fwd<T, false, false>(x, res, w, h_out, y, rstd, rows, hidden, eps, st);
}

constexpr int kBwdRows = 2;

// Backward tuning knobs, read once from the environment (A/B measurements only):
//   B200K_RMSNORM_BWD_TMA=0      use the register-resident kernel instead of the TMA-staged one
//   B200K_RMSNORM_BWD_VPL=1|2    16-byte vectors per thread
//   B200K_RMSNORM_BWD_CTAS=1|2   resident CTAs per SM (splits the shared-memory ring)
struct TmaTune {
    bool enabled;
    int vpl, rows, ctas_per_sm;
};
TmaTune tma_tune() {
    static TmaTune t = [] {
        auto geti = [](const char* name, int dflt) {
            const char* e = getenv(name);
            return e != nullptr && e[0] != 0 ? atoi(e) : dflt;
        };
        TmaTune r;
        r.enabled = geti("B200K_RMSNORM_BWD_TMA", 1) != 0;
        r.vpl = geti("B200K_RMSNORM_BWD_VPL", 1);
        r.rows = kBwdRows;  // fixed: the partial-sum scratch (b200k_rmsnorm_bwd_scratch_bytes) is sized for two rows per stage
        r.ctas_per_sm = geti("B200K_RMSNORM_BWD_CTAS", 1);
        if (r.vpl != 1 && r.vpl != 2) r.vpl = 1;
        if (r.rows != 1 && r.rows != 2) r.rows = 2;
        if (r.ctas_per_sm < 1 || r.ctas_per_sm > kBwdCtasPerSm) r.ctas_per_sm = 1;
        return r;
    }();
    return t;
}

template <typename T, bool kHasW, bool kHasDhIn>
int launch_bwd(const void* dy, const void* h, const void* w, const float* rstd, const void* dh_in, void* dx, void* dw,
               void* scratch, int64_t scratch_bytes, int64_t rows, int64_t hidden, cudaStream_t st) {
    using V = Vec<T>;
    const bool vec_ok = hidden % V::kElems == 0 && aligned16(dy) && aligned16(h) && aligned16(dx) &&
                        (!kHasDhIn || aligned16(dh_in)) && (!kHasW || aligned16(w));
    const int64_t nvec = hidden / V::kElems;
    const RowPlan p = vec_ok ? plan_row(nvec, 2, false) : RowPlan{0, 0};
    if (p.vpl == 0) {
        const int64_t cap = static_cast<int64_t>(sm_count()) * 8;
        const int grid = static_cast<int>(rows < cap ? rows : cap);
        rmsnorm_bwd_generic_dx_kernel<T, kHasW, kHasDhIn><<<grid, 256, 0, st>>>(
            static_cast<const T*>(dy), static_cast<const T*>(h), static_cast<const T*>(w), rstd,
            static_cast<const T*>(dh_in), static_cast<T*>(dx), rows, hidden);
        int rc = launch_status();
        if (rc != 0 || !kHasW) return rc;
        rmsnorm_bwd_generic_dw_kernel<T><<<static_cast<int>((hidden + 31) / 32), 256, 0, st>>>(
            static_cast<const T*>(dy), static_cast<const T*>(h), rstd, static_cast<T*>(dw), rows, hidden);
        return launch_status();
    }
    if (kHasW && (scratch == nullptr || !aligned16(scratch) || scratch_bytes < b200k_rmsnorm_bwd_scratch_bytes(rows, hidden)))
        return B200K_ESCRATCH;
    int grid = 0;
    {
        // TMA-staged streaming kernel.  Tunables (env, read once): consumer vectors per thread, rows per
        // stage, CTAs per SM.  Defaults are the measured best for [8192 x 4096] bf16 on B200.
        const TmaTune t = tma_tune();
        int rc = -1000;
        auto try_tma = [&](auto kernel, int vpl, int kr) {
            constexpr int kTensors = kHasDhIn ? 3 : 2;
            const int64_t stage_bytes = nvec * 16 * kr * kTensors;
            const int64_t budget = (t.ctas_per_sm >= 2 ? 100 : 200) * 1024;
            int64_t stages = budget / stage_bytes;
            if (stages > 8) stages = 8;
            const int64_t cthreads = nvec / vpl;
            if (nvec % (32 * vpl) != 0 || stages < 2 || cthreads > 512 / vpl || cthreads < 32) return;
            const size_t smem = ((2 * stages * 8 + 2 * kr * 32 * 4 + 127) / 128) * 128 + stages * stage_bytes;
            cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
            if (e != cudaSuccess) {
                rc = static_cast<int>(e);
                return;
            }
            const int64_t groups = (rows + kr - 1) / kr;
            const int64_t cap = static_cast<int64_t>(sm_count()) * t.ctas_per_sm;
            grid = static_cast<int>(groups < cap ? groups : cap);
            // 1/hidden is exact in fp32 iff hidden is a power of two; then x * (1/hidden) == x / hidden bit for bit.
            const float inv_exact = (hidden & (hidden - 1)) == 0 ? 1.0f / static_cast<float>(hidden) : 0.f;
            kernel<<<grid, static_cast<int>(cthreads), smem, st>>>(
                static_cast<const T*>(dy), static_cast<const T*>(h), static_cast<const T*>(w), rstd, static_cast<const T*>(dh_in),
                static_cast<T*>(dx), static_cast<float*>(scratch), rows, static_cast<int>(nvec), static_cast<float>(hidden),
                inv_exact, static_cast<int>(stages));
            rc = launch_status();
        };
        if (t.enabled) {
            if (t.vpl == 1 && t.rows == 1) try_tma(rmsnorm_bwd_tma_kernel<T, 1, 1, kHasW, kHasDhIn>, 1, 1);
            else if (t.vpl == 1) try_tma(rmsnorm_bwd_tma_kernel<T, 1, 2, kHasW, kHasDhIn>, 1, 2);
            else if (t.vpl == 2 && t.rows == 1) try_tma(rmsnorm_bwd_tma_kernel<T, 2, 1, kHasW, kHasDhIn>, 2, 1);
            else try_tma(rmsnorm_bwd_tma_kernel<T, 2, 2, kHasW, kHasDhIn>, 2, 2);
        }
        if (rc != -1000) {
            if (rc != 0 || !kHasW) return rc;
            return launch_colsum_finalize<T>(scratch, dw, grid, hidden, st);
        }
    }
    auto go = [&](auto kernel) {
        grid = persistent_grid(kernel, p.threads, (rows + kBwdRows - 1) / kBwdRows, kBwdCtasPerSm);
        kernel<<<grid, p.threads, 0, st>>>(static_cast<const T*>(dy), static_cast<const T*>(h), static_cast<const T*>(w), rstd,
                                           static_cast<const T*>(dh_in), static_cast<T*>(dx), static_cast<float*>(scratch), rows,
                                           static_cast<int>(nvec), static_cast<float>(hidden));
        return launch_status();
    };
    int rc;
    switch (p.vpl) {
        case 1: rc = go(rmsnorm_bwd_kernel<T, 1, kBwdRows, kHasW, kHasDhIn>); break;
        case 2: rc = go(rmsnorm_bwd_kernel<T, 2, kBwdRows, kHasW, kHasDhIn>); break;
        case 4: rc = go(rmsnorm_bwd_kernel<T, 4, kBwdRows, kHasW, kHasDhIn>); break;
        default: rc = go(rmsnorm_bwd_kernel<T, 8, 1, kHasW, kHasDhIn>); break;
    }
    if (rc != 0 || !kHasW) return rc;
    return launch_colsum_finalize<T>(scratch, dw, grid, hidden, st);
}

template <typename T>
int dispatch_bwd(const void* dy, const void* h, const void* w, const float* rstd, const void* dh_in, void* dx, void* dw,
                 void* scratch, int64_t scratch_bytes, int64_t rows, int64_t hidden, cudaStream_t st) {
    if (w != nullptr) {
        return dh_in != nullptr
                   ? launch_bwd<T, true, true>(dy, h, w, rstd, dh_in, dx, dw, scratch, scratch_bytes, rows, hidden, st)
                   : launch_bwd<T, true, false>(dy, h, w, rstd, dh_in, dx, dw, scratch, scratch_bytes, rows, hidden, st);
    }
    return dh_in != nullptr
               ? launch_bwd<T, false, true>(dy, h, w, rstd, dh_in, dx, dw, scratch, scratch_bytes, rows, hidden, st)
               : launch_bwd<T, false, false>(dy, h, w, rstd, dh_in, dx, dw, scratch, scratch_bytes, rows, hidden, st);
}

}  // namespace
}  // namespace b200k

using namespace b200k;

extern "C" int b200k_rmsnorm_fwd(const void* x, const void* residual, const void* weight, void* h_out, void* y,
                                 float* rstd_out, int64_t rows, int64_t hidden, float eps, int dtype, void* stream) {
    if (rows < 0 || hidden <= 0) return B200K_EINVAL;
    if (rows == 0) return 0;
    if (x == nullptr || y == nullptr || (residual != nullptr && h_out == nullptr)) return B200K_EINVAL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    switch (dtype) {
        case B200K_F32: return dispatch_fwd<float>(x, residual, weight, h_out, y, rstd_out, rows, hidden, eps, st);
        case B200K_BF16: return dispatch_fwd<__nv_bfloat16>(x, residual, weight, h_out, y, rstd_out, rows, hidden, eps, st);
        case B200K_F16: return dispatch_fwd<__half>(x, residual, weight, h_out, y, rstd_out, rows, hidden, eps, st);
        default: return B200K_EDTYPE;
    }
}

extern "C" int64_t b200k_rmsnorm_bwd_scratch_bytes(int64_t rows, int64_t hidden) {
    if (rows <= 0 || hidden <= 0) return 0;
    int64_t parts = static_cast<int64_t>(sm_count()) * kBwdCtasPerSm;
    const int64_t groups = (rows + kBwdRows - 1) / kBwdRows;
    if (groups < parts) parts = groups;
    return parts * hidden * static_cast<int64_t>(sizeof(float));
}

extern "C" int b200k_rmsnorm_bwd(const void* dy, const void* h, const void* weight, const float* rstd, const void* dh_in,
                                 void* dx, void* dw, void* scratch, int64_t scratch_bytes, int64_t rows, int64_t hidden,
                                 int dtype, void* stream) {
    if (rows < 0 || hidden <= 0) return B200K_EINVAL;
    if (dy == nullptr && rows > 0) return B200K_EINVAL;
    if (rows > 0 && (h == nullptr || rstd == nullptr || dx == nullptr)) return B200K_EINVAL;
    if (weight != nullptr && dw == nullptr) return B200K_EINVAL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (rows == 0) {
        // An empty batch still defines dw = 0.
        if (weight != nullptr) {
            const size_t esz = dtype == B200K_F32 ? 4 : 2;
            cudaError_t e = cudaMemsetAsync(dw, 0, esz * static_cast<size_t>(hidden), st);
            return e == cudaSuccess ? 0 : static_cast<int>(e);
        }
        return 0;
    }
    switch (dtype) {
        case B200K_F32: return dispatch_bwd<float>(dy, h, weight, rstd, dh_in, dx, dw, scratch, scratch_bytes, rows, hidden, st);
        case B200K_BF16:
            return dispatch_bwd<__nv_bfloat16>(dy, h, weight, rstd, dh_in, dx, dw, scratch, scratch_bytes, rows, hidden, st);
        case B200K_F16: return dispatch_bwd<__half>(dy, h, weight, rstd, dh_in, dx, dw, scratch, scratch_bytes, rows, hidden, st);
        default: return B200K_EDTYPE;
    }
}


// -------------------------------------------------------------------------------------------
// b200k_rmsnorm_bwd_noalloc: the same backward, for callers that cannot allocate device memory.
//
// The dw partial rows live in a fixed array inside this library's device image instead of a buffer the
// caller provides.  The array's name is only a host-side handle: its device address is looked up once
// per device with cudaGetSymbolAddress (with that device current) and cached; that pointer, never the
// symbol itself, is what the kernels receive.
// -------------------------------------------------------------------------------------------
namespace {

static __device__ __align__(256) unsigned char g_rmsnorm_bwd_static_scratch[B200K_RMSNORM_BWD_NOALLOC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
void* g_static_scratch_addr[kMaxDevices] = {};
std::mutex g_static_scratch_mutex;

// Device address of the static scratch array on the current device (resolved on first use per device).
int static_scratch_address(void** out) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return static_cast<int>(e);
    if (dev < 0 || dev >= kMaxDevices) return B200K_ELIMIT;
    std::lock_guard<std::mutex> guard(g_static_scratch_mutex);
    if (g_static_scratch_addr[dev] == nullptr) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_rmsnorm_bwd_static_scratch);
        if (e != cudaSuccess) return static_cast<int>(e);
        if (p == nullptr) return B200K_EINVAL;
        g_static_scratch_addr[dev] = p;
    }
    *out = g_static_scratch_addr[dev];
    return 0;
}

}  // namespace

extern "C" int b200k_rmsnorm_bwd_noalloc(const void* dy, const void* h, const void* weight, const float* rstd, const void* dh_in,
                                         void* dx, void* dw, int64_t rows, int64_t hidden, int dtype, void* stream) {
    if (rows < 0 || hidden <= 0) return B200K_EINVAL;
    // Reject shapes whose partial rows would not fit the fixed array before touching anything.
    if (weight != nullptr && b200k_rmsnorm_bwd_scratch_bytes(rows, hidden) > static_cast<int64_t>(B200K_RMSNORM_BWD_NOALLOC_SCRATCH_BYTES))
        return B200K_ELIMIT;
    void* scratch = nullptr;
    if (weight != nullptr && rows > 0) {
        const int rc = static_scratch_address(&scratch);
        if (rc != 0) return rc;
    }
    // Every partial row that the fold kernel reads is fully written by the producing grid first, so the
    // array needs no clearing between calls even though it keeps the previous call's contents.
    return b200k_rmsnorm_bwd(dy, h, weight, rstd, dh_in, dx, dw, scratch, static_cast<int64_t>(B200K_RMSNORM_BWD_NOALLOC_SCRATCH_BYTES),
                             rows, hidden, dtype, stream);
}
