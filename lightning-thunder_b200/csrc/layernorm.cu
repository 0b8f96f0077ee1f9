// LayerNorm over the last dimension, forward and backward, for sm_100a.  include/b200k.h states the math,
// which follows the reference's torch-executor trace of ltorch.layer_norm op for op.
//
// Layout: activations are [rows, hidden] row-major, mean / rstd are fp32 [rows].  Forward: a CTA owns one row
// at a time with the whole row in registers (x is read once, y written once; two CTA-wide reductions per row:
// the sum, then the sum of squared deviations — the two-pass form, no E[x^2] - E[x]^2 cancellation).
// Backward: the same TMA-staged ring as the RMSNorm backward (rmsnorm.cu): 1-D bulk copies keep `stages` row
// groups of dy and x in flight per CTA, thread t owns the 16-byte vectors t, t + blockDim.x, ..., two row sums
// (c1 = sum xc * w * dy, c2 = -sum rstd * w * dy) share one shared-memory hop, dx goes from registers to global
// memory, and dw / db are per-thread fp32 column partials written once per CTA and folded by a second tiny grid
// launched with programmatic stream serialization.
#include <cstdlib>

#include <type_traits>

#include "common.cuh"

namespace b200k {
namespace {

constexpr int kLnBwdCtasPerSm = 8;  // upper bound on resident backward CTAs per SM (sizes the scratch)
constexpr int kLnBwdRows = 2;       // rows per pipeline stage

constexpr int ln_fwd_max_threads(int vpl) { return vpl >= 4 ? 512 : 256; }

__device__ __forceinline__ float ln_block_sum(float v, float* smem, int nwarps) {
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

// n / d, correctly rounded, for a divisor whose reciprocal r = RN(1 / d) is known: one Newton correction of
// q = RN(n * r) with the exact fma residual (Markstein).  Bit-identical to the IEEE division for the row widths
// used here unless the quotient is subnormal.
__device__ __forceinline__ float2 div_by_const2(float2 n, float2 neg_d, float2 r) {
    const float2 q = __fmul2_rn(n, r);
    const float2 e = __ffma2_rn(q, neg_d, n);
    return __ffma2_rn(e, r, q);
}

// -------------------------------------------------------------------------------------------
// Forward.
// -------------------------------------------------------------------------------------------
template <typename T, int kVpl, bool kHasW, bool kHasB>
__global__ void __launch_bounds__(ln_fwd_max_threads(kVpl))
layernorm_fwd_kernel(const T* __restrict__ x, const T* __restrict__ w, const T* __restrict__ b, T* __restrict__ y,
                     float* __restrict__ mean_out, float* __restrict__ rstd_out, int64_t rows, int nvec, float hidden_f,
                     float eps) {
    using V = Vec<T>;
    constexpr int kE = V::kElems;
    __shared__ float red[4][kWarp];
    const int tid = threadIdx.x, nthr = blockDim.x, nwarps = nthr >> 5;

    uint4 wv[kVpl], bv[kVpl];
#pragma unroll
    for (int i = 0; i < kVpl; ++i) {
        const int v = tid + i * nthr;
        if constexpr (kHasW) wv[i] = v < nvec ? ld_keep(reinterpret_cast<const uint4*>(w) + v) : make_uint4(0, 0, 0, 0);
        if constexpr (kHasB) bv[i] = v < nvec ? ld_keep(reinterpret_cast<const uint4*>(b) + v) : make_uint4(0, 0, 0, 0);
    }

    // The next row's vectors are requested before this row's two reductions, so the loads of row k + 1 are in
    // flight while row k waits at its barriers.
    auto load_row = [&](int64_t row, uint4 (&dst)[kVpl]) {
        const uint4* xr = reinterpret_cast<const uint4*>(x) + row * nvec;
#pragma unroll
        for (int i = 0; i < kVpl; ++i) {
            const int v = tid + i * nthr;
            dst[i] = v < nvec && row < rows ? ld_stream(xr + v) : make_uint4(0, 0, 0, 0);
        }
    };
    constexpr bool kPrefetch = kVpl <= 4;  // the widest rows have no registers to spare for a second row
    constexpr int kP = kE / 2;
    uint4 xv[kVpl], nx[kPrefetch ? kVpl : 1];
    if constexpr (kPrefetch) load_row(blockIdx.x, xv);
    int parity = 0;
    for (int64_t row = blockIdx.x; row < rows; row += gridDim.x, parity ^= 2) {
        if constexpr (kPrefetch) load_row(row + gridDim.x, nx);
        else load_row(row, xv);
        // float2 pairs throughout: packed FADD2 / FMUL2 / FFMA2 halve the instruction count and the length of
        // the accumulation chains; the results are those of the scalar round-to-nearest operations.
        float2 xf[kVpl][kP];
        float2 s2 = make_float2(0.f, 0.f);
#pragma unroll
        for (int i = 0; i < kVpl; ++i) {
            V::unpack2(xv[i], xf[i]);  // padding vectors are zeros
#pragma unroll
            for (int e = 0; e < kP; ++e) s2 = __fadd2_rn(s2, xf[i][e]);
        }
        const float mean = ln_block_sum(s2.x + s2.y, red[parity], nwarps) / hidden_f;
        const float2 nmean2 = make_float2(-mean, -mean);
        float2 ss2 = make_float2(0.f, 0.f);
#pragma unroll
        for (int i = 0; i < kVpl; ++i) {
            float2 p = make_float2(0.f, 0.f);
#pragma unroll
            for (int e = 0; e < kP; ++e) {
                xf[i][e] = __fadd2_rn(xf[i][e], nmean2);
                p = __ffma2_rn(xf[i][e], xf[i][e], p);
            }
            if (tid + i * nthr < nvec) ss2 = __fadd2_rn(ss2, p);  // a padding vector would add (0 - mean)^2
        }
        const float var = ln_block_sum(ss2.x + ss2.y, red[parity + 1], nwarps) / hidden_f;
        const float rstd = rsqrtf(var + eps);
        if (tid == 0) {
            if (mean_out != nullptr) mean_out[row] = mean;
            if (rstd_out != nullptr) rstd_out[row] = rstd;
        }
        const float2 rstd2 = make_float2(rstd, rstd);
#pragma unroll
        for (int i = 0; i < kVpl; ++i) {
            const int v = tid + i * nthr;
            float2 wf[kP], bf[kP];
            if constexpr (kHasW) V::unpack2(wv[i], wf);
            if constexpr (kHasB) V::unpack2(bv[i], bf);
#pragma unroll
            for (int e = 0; e < kP; ++e) {
                float2 n = __fmul2_rn(xf[i][e], rstd2);
                if constexpr (kHasW) n = __fmul2_rn(n, wf[e]);
                if constexpr (kHasB) n = __fadd2_rn(n, bf[e]);
                xf[i][e] = n;
            }
            if (v < nvec) st_stream(reinterpret_cast<uint4*>(y) + row * nvec + v, V::pack2(xf[i]));
        }
        if constexpr (kPrefetch) {
#pragma unroll
            for (int i = 0; i < kVpl; ++i) xv[i] = nx[i];
        }
    }
}

// Any `hidden` (no alignment requirement): CTA per row, three passes over the row (the later ones hit L1/L2).
template <typename T, bool kHasW, bool kHasB>
__global__ void __launch_bounds__(256)
layernorm_fwd_generic_kernel(const T* __restrict__ x, const T* __restrict__ w, const T* __restrict__ b, T* __restrict__ y,
                             float* __restrict__ mean_out, float* __restrict__ rstd_out, int64_t rows, int64_t hidden, float eps) {
    using V = Vec<T>;
    __shared__ float red[4][kWarp];
    int parity = 0;
    const float hidden_f = static_cast<float>(hidden);
    for (int64_t row = blockIdx.x; row < rows; row += gridDim.x, parity ^= 2) {
        const int64_t base = row * hidden;
        float s = 0.f;
        for (int64_t j = threadIdx.x; j < hidden; j += blockDim.x) s += V::load1(x, base + j);
        const float mean = ln_block_sum(s, red[parity], blockDim.x >> 5) / hidden_f;
        float ss = 0.f;
        for (int64_t j = threadIdx.x; j < hidden; j += blockDim.x) {
            const float d = __fsub_rn(V::load1(x, base + j), mean);
            ss += d * d;
        }
        const float rstd = rsqrtf(ln_block_sum(ss, red[parity + 1], blockDim.x >> 5) / hidden_f + eps);
        if (threadIdx.x == 0) {
            if (mean_out != nullptr) mean_out[row] = mean;
            if (rstd_out != nullptr) rstd_out[row] = rstd;
        }
        for (int64_t j = threadIdx.x; j < hidden; j += blockDim.x) {
            float n = __fmul_rn(__fsub_rn(V::load1(x, base + j), mean), rstd);
            if constexpr (kHasW) n = __fmul_rn(n, V::load1(w, j));
            if constexpr (kHasB) n = __fadd_rn(n, V::load1(b, j));
            V::store1(y, base + j, n);
        }
    }
}

// -------------------------------------------------------------------------------------------
// Backward, TMA-staged.  Stage layout: [dy rows | x rows], each kRows * row_bytes.
// Requires blockDim.x * kVpl == nvec (no column predicates in the hot loop).
// -------------------------------------------------------------------------------------------
template <typename T, int kVpl, int kRows, bool kHasW, bool kHasB, int kMaxThreads = (kVpl == 1 ? 512 : 256), int kMinBlocks = 1>
__global__ void __launch_bounds__(kMaxThreads, kMinBlocks)
layernorm_bwd_tma_kernel(const T* __restrict__ dy, const T* __restrict__ x, const T* __restrict__ w,
                         const float* __restrict__ mean, const float* __restrict__ rstd, T* __restrict__ dx,
                         float* __restrict__ dw_partial, float* __restrict__ db_partial, int64_t rows, int nvec,
                         float hidden_f, float inv_hidden, int stages) {
    using V = Vec<T>;
    constexpr int kE = V::kElems;
    constexpr int kP = kE / 2;
    constexpr bool kIsF32 = V::kDtype == B200K_F32;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const uint32_t row_bytes = static_cast<uint32_t>(nvec) * 16u;
    const uint32_t group_bytes = row_bytes * kRows;
    const uint32_t stage_bytes = group_bytes * 2;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw);  // [stages]
    uint64_t* empty = full + stages;                         // [stages]
    float* red = reinterpret_cast<float*>(empty + stages);   // [2][2 * kRows][32]
    unsigned char* ring = smem_raw + ((2 * stages * 8 + 2 * 2 * kRows * 32 * 4 + 127) / 128) * 128;

    const int nc = blockDim.x, ncw = nc >> 5;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t ngroups = (rows + kRows - 1) / kRows;
    uint64_t pol = 0;

    auto issue = [&](int64_t g, int s) {
        const int64_t row0 = g * kRows;
        const int64_t live = rows - row0 < kRows ? rows - row0 : kRows;
        const uint32_t bytes = static_cast<uint32_t>(live) * row_bytes;
        unsigned char* st = ring + static_cast<size_t>(s) * stage_bytes;
        mbar_arrive_expect_tx(&full[s], bytes * 2);
        const size_t goff = static_cast<size_t>(row0) * row_bytes;
        bulk_g2s(st, reinterpret_cast<const unsigned char*>(dy) + goff, bytes, &full[s], pol);
        bulk_g2s(st + group_bytes, reinterpret_cast<const unsigned char*>(x) + goff, bytes, &full[s], pol);
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

    float2 wf[kVpl][kP], dwacc[kVpl][kP], dbacc[kVpl][kP];
#pragma unroll
    for (int i = 0; i < kVpl; ++i) {
        if constexpr (kHasW) V::unpack2(ld_keep(reinterpret_cast<const uint4*>(w) + tid + i * nc), wf[i]);
#pragma unroll
        for (int e = 0; e < kP; ++e) dwacc[i][e] = dbacc[i][e] = make_float2(0.f, 0.f);
    }
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    const float2 hid2 = make_float2(-hidden_f, -hidden_f), rcp2 = make_float2(1.0f / hidden_f, 1.0f / hidden_f);
    const bool hidden_pow2 = (__float_as_uint(hidden_f) & 0x007fffffu) == 0u;  // mantissa bits all zero
    float rs_next[kRows], mu_next[kRows];
    auto load_stats = [&](int64_t g) {
#pragma unroll
        for (int r = 0; r < kRows; ++r) {
            const int64_t row = g * kRows + r;
            const bool live = row < rows;
            rs_next[r] = live ? __ldg(rstd + row) : 0.f;
            mu_next[r] = live ? __ldg(mean + row) : 0.f;
        }
    };
    load_stats(blockIdx.x);

    int s = 0;
    uint32_t phase = 0;
    int parity = 0;
    for (int64_t g = blockIdx.x; g < ngroups; g += gridDim.x, parity ^= 1) {
        const int64_t row0 = g * kRows;
        // This group's statistics were requested one iteration ago; request the next group's now, so that their
        // global-memory latency is covered by a whole iteration instead of stalling the first use below.
        float rs[kRows], mu[kRows];
#pragma unroll
        for (int r = 0; r < kRows; ++r) rs[r] = rs_next[r], mu[r] = mu_next[r];
        load_stats(g + gridDim.x);

        mbar_wait(&full[s], phase);
        const unsigned char* st = ring + static_cast<size_t>(s) * stage_bytes;
        uint4 dyv[kRows][kVpl], xv[kRows][kVpl];
#pragma unroll
        for (int r = 0; r < kRows; ++r) {
#pragma unroll
            for (int i = 0; i < kVpl; ++i) {
                const uint32_t off = static_cast<uint32_t>(r) * row_bytes + static_cast<uint32_t>(tid + i * nc) * 16u;
                dyv[r][i] = lds128(st + off);
                xv[r][i] = lds128(st + group_bytes + off);
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);

        // c[2r] = sum_j (x - mean) * (w * dy),  c[2r + 1] = sum_j rstd * (w * dy)   (negated below)
        float c[2 * kRows];
#pragma unroll
        for (int r = 0; r < kRows; ++r) {
            float2 a1 = make_float2(0.f, 0.f), a2 = make_float2(0.f, 0.f);
            const float2 mu2 = make_float2(-mu[r], -mu[r]), rstd2 = make_float2(rs[r], rs[r]);
#pragma unroll
            for (int i = 0; i < kVpl; ++i) {
                float2 dyf[kP], xf[kP];
                V::unpack2(dyv[r][i], dyf);
                V::unpack2(xv[r][i], xf);
#pragma unroll
                for (int e = 0; e < kP; ++e) {
                    const float2 gw = kHasW ? __fmul2_rn(wf[i][e], dyf[e]) : dyf[e];
                    a1 = __ffma2_rn(__fadd2_rn(xf[e], mu2), gw, a1);
                    a2 = __fadd2_rn(a2, gw);
                }
            }
            // rstd is a row constant: it multiplies the sum once instead of every term (the sum is reassociated
            // across lanes anyway, so this is within the same fp32 rounding noise)
            c[2 * r] = warp_reduce(a1.x + a1.y, SumOp{});
            c[2 * r + 1] = warp_reduce((a2.x + a2.y) * rs[r], SumOp{});
        }
        if (ncw > 1) {
            float* rb = red + parity * (2 * kRows * 32);
            if (lane == 0) {
#pragma unroll
                for (int k = 0; k < 2 * kRows; ++k) rb[k * 32 + warp] = c[k];
            }
            __syncthreads();
#pragma unroll
            for (int k = 0; k < 2 * kRows; ++k) {
                float xsum = lane < ncw ? rb[k * 32 + lane] : 0.f;
                c[k] = warp_reduce<16>(xsum, SumOp{});
                c[k] = __shfl_sync(0xffffffffu, c[k], 0);
            }
        }
        // Every warp has released slot s (its arrive precedes the barrier above; a one-warp CTA is its own
        // release): refill it.
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

        // A power-of-two row width divides exactly by one multiplication (its reciprocal is exact and the product
        // is the correctly rounded quotient, subnormals included); other widths take the corrected division.
        auto emit = [&](auto pow2_tag) {
            constexpr bool kPow2 = decltype(pow2_tag)::value;
#pragma unroll
            for (int r = 0; r < kRows; ++r) {
                const int64_t row = row0 + r;
                if (row >= rows) break;
                const float rstd_r = rs[r];
                // var_mean's vjp in the reference's order: gvar = (-0.5 * c1) * rstd^3, mean_grad = (1/H) * c2,
                // var_grad = ((2 * gvar) * (x - mean)) / H
                const float gvar2 = 2.f * ((-0.5f * c[2 * r]) * (rstd_r * rstd_r * rstd_r));
                const float mg = inv_hidden * -c[2 * r + 1];
                const float2 mu2 = make_float2(-mu[r], -mu[r]), rstd2 = make_float2(rstd_r, rstd_r);
                const float2 gv2 = make_float2(gvar2, gvar2), mg2 = make_float2(mg, mg);
#pragma unroll
                for (int i = 0; i < kVpl; ++i) {
                    float2 dyf[kP], xf[kP], o[kP];
                    uint32_t ow[4] = {0u, 0u, 0u, 0u};
                    V::unpack2(dyv[r][i], dyf);
                    V::unpack2(xv[r][i], xf);
#pragma unroll
                    for (int e = 0; e < kP; ++e) {
                        const float2 gw = kHasW ? __fmul2_rn(wf[i][e], dyf[e]) : dyf[e];
                        const float2 xc = __fadd2_rn(xf[e], mu2);
                        const float2 direct = __fmul2_rn(rstd2, gw);
                        const float2 t = __fmul2_rn(gv2, xc);
                        const float2 stats = __fadd2_rn(mg2, kPow2 ? __fmul2_rn(t, rcp2) : div_by_const2(t, hid2, rcp2));
                        // 16-bit: round(round(direct) + round(stats)) as two packed converts and one packed add — the
                        // packed add rounds the exact sum once, which equals the fp32 add followed by the rounding
                        if constexpr (kIsF32) o[e] = __fadd2_rn(direct, stats);
                        else ow[e] = Pk<T>::add(V::pack_word(direct), V::pack_word(stats));
                        if constexpr (kHasW) dwacc[i][e] = __fadd2_rn(dwacc[i][e], V::rnd2f(__fmul2_rn(__fmul2_rn(xc, rstd2), dyf[e])));
                        if constexpr (kHasB) dbacc[i][e] = __fadd2_rn(dbacc[i][e], dyf[e]);
                    }
                    st_stream(reinterpret_cast<uint4*>(dx) + row * nvec + tid + i * nc, kIsF32 ? V::pack2(o) : make_uint4(ow[0], ow[1], ow[2], ow[3]));
                }
            }
        };
        if (hidden_pow2) emit(std::true_type{});
        else emit(std::false_type{});
    }

    auto flush = [&](float* partial, float2 (&acc)[kVpl][kP]) {
        float* prow = partial + static_cast<int64_t>(blockIdx.x) * nvec * kE;
#pragma unroll
        for (int i = 0; i < kVpl; ++i) {
            const int v = tid + i * nc;
#pragma unroll
            for (int q = 0; q < kE / 4; ++q) {
                reinterpret_cast<float4*>(prow + static_cast<int64_t>(v) * kE)[q] =
                    make_float4(acc[i][2 * q].x, acc[i][2 * q].y, acc[i][2 * q + 1].x, acc[i][2 * q + 1].y);
            }
        }
    };
    if constexpr (kHasW) flush(dw_partial, dwacc);
    if constexpr (kHasB) flush(db_partial, dbacc);
}

// out[col] = round_T(sum_g partial[g][col]) for two partial blocks at once (blockIdx.y picks dw or db); the
// same tile shape and dependency handling as colsum_finalize_kernel in reduce.cuh.
template <typename T>
__global__ void __launch_bounds__(256)
layernorm_finalize_kernel(const float* __restrict__ dw_partial, const float* __restrict__ db_partial, T* __restrict__ dw,
                          T* __restrict__ db, int nparts, int64_t hidden) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const float* partial = blockIdx.y == 0 ? dw_partial : db_partial;
    T* out = blockIdx.y == 0 ? dw : db;
    if (out == nullptr) return;
    __shared__ float4 tile[16][17];
    const int cx = threadIdx.x & 15, ry = threadIdx.x >> 4;
    const int64_t col = (static_cast<int64_t>(blockIdx.x) * 16 + cx) * 4;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (col < hidden) {
#pragma unroll 4
        for (int g = ry; g < nparts; g += 16) {
            const float4 v = *reinterpret_cast<const float4*>(partial + static_cast<int64_t>(g) * hidden + col);
            acc.x += v.x, acc.y += v.y, acc.z += v.z, acc.w += v.w;
        }
    }
    tile[ry][cx] = acc;
    __syncthreads();
    if (ry == 0 && col < hidden) {
        float4 t = tile[0][cx];
#pragma unroll
        for (int k = 1; k < 16; ++k) {
            const float4 v = tile[k][cx];
            t.x += v.x, t.y += v.y, t.z += v.z, t.w += v.w;
        }
        Vec<T>::store1(out, col, t.x);
        Vec<T>::store1(out, col + 1, t.y);
        Vec<T>::store1(out, col + 2, t.z);
        Vec<T>::store1(out, col + 3, t.w);
    }
}

// Generic backward (any hidden): CTA per row for dx, then a column kernel for dw / db that re-reads x, dy.
template <typename T, bool kHasW>
__global__ void __launch_bounds__(256)
layernorm_bwd_generic_dx_kernel(const T* __restrict__ dy, const T* __restrict__ x, const T* __restrict__ w,
                                const float* __restrict__ mean, const float* __restrict__ rstd, T* __restrict__ dx,
                                int64_t rows, int64_t hidden, float inv_hidden) {
    using V = Vec<T>;
    __shared__ float red[4][kWarp];
    const float hidden_f = static_cast<float>(hidden);
    int parity = 0;
    for (int64_t row = blockIdx.x; row < rows; row += gridDim.x, parity ^= 2) {
        const int64_t base = row * hidden;
        const float mu = mean[row], rstd_r = rstd[row];
        float a1 = 0.f, a2 = 0.f;
        for (int64_t j = threadIdx.x; j < hidden; j += blockDim.x) {
            const float gw = kHasW ? __fmul_rn(V::load1(w, j), V::load1(dy, base + j)) : V::load1(dy, base + j);
            a1 += __fsub_rn(V::load1(x, base + j), mu) * gw;
            a2 += __fmul_rn(rstd_r, gw);
        }
        const float c1 = ln_block_sum(a1, red[parity], blockDim.x >> 5);
        const float c2 = -ln_block_sum(a2, red[parity + 1], blockDim.x >> 5);
        const float gvar2 = 2.f * ((-0.5f * c1) * (rstd_r * rstd_r * rstd_r));
        const float mg = inv_hidden * c2;
        for (int64_t j = threadIdx.x; j < hidden; j += blockDim.x) {
            const float gw = kHasW ? __fmul_rn(V::load1(w, j), V::load1(dy, base + j)) : V::load1(dy, base + j);
            const float xc = __fsub_rn(V::load1(x, base + j), mu);
            const float direct = __fmul_rn(rstd_r, gw);
            const float stats = __fadd_rn(mg, __fdiv_rn(__fmul_rn(gvar2, xc), hidden_f));
            const float o = V::kDtype == B200K_F32 ? __fadd_rn(direct, stats) : __fadd_rn(V::rnd(direct), V::rnd(stats));
            V::store1(dx, base + j, o);
        }
    }
}

template <typename T>
__global__ void __launch_bounds__(256)
layernorm_bwd_generic_cols_kernel(const T* __restrict__ dy, const T* __restrict__ x, const float* __restrict__ mean,
                                  const float* __restrict__ rstd, T* __restrict__ dw, T* __restrict__ db, int64_t rows,
                                  int64_t hidden) {
    using V = Vec<T>;
    __shared__ float tw[8][kWarp + 1], tb[8][kWarp + 1];
    const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
    const int64_t col = static_cast<int64_t>(blockIdx.x) * 32 + cx;
    float aw = 0.f, ab = 0.f;
    if (col < hidden) {
        for (int64_t r = ry; r < rows; r += 8) {
            const float g = V::load1(dy, r * hidden + col);
            aw += V::rnd(__fmul_rn(__fmul_rn(__fsub_rn(V::load1(x, r * hidden + col), mean[r]), rstd[r]), g));
            ab += g;
        }
    }
    tw[ry][cx] = aw;
    tb[ry][cx] = ab;
    __syncthreads();
    if (ry == 0 && col < hidden) {
        float sw = 0.f, sb = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) sw += tw[k][cx], sb += tb[k][cx];
        if (dw != nullptr) V::store1(dw, col, sw);
        if (db != nullptr) V::store1(db, col, sb);
    }
}

// -------------------------------------------------------------------------------------------
// Host-side dispatch.
// -------------------------------------------------------------------------------------------
struct LnRowPlan {
    int vpl;      // 16-byte vectors per thread
    int threads;  // CTA size (multiple of 32)
};

// (vpl, threads) with threads * vpl covering nvec; wide rows keep >= min_vpl vectors in flight per thread.
// vpl = 0 when the row does not fit the register-resident kernel.
LnRowPlan ln_plan_row(int64_t nvec, int min_vpl) {
    for (int vpl : {1, 2, 4, 8}) {
        if (vpl < min_vpl && nvec > static_cast<int64_t>(kWarp) * vpl) continue;
        int64_t thr = (nvec + vpl - 1) / vpl;
        thr = (thr + kWarp - 1) / kWarp * kWarp;
        if (thr <= ln_fwd_max_threads(vpl)) return {vpl, static_cast<int>(thr)};
    }
    return {0, 0};
}

template <typename Kernel>
int ln_persistent_grid(Kernel kernel, int threads, int64_t work_items, int max_ctas_per_sm) {
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, 0) != cudaSuccess || per_sm < 1) per_sm = 1;
    if (per_sm > max_ctas_per_sm) per_sm = max_ctas_per_sm;
    const int64_t cap = static_cast<int64_t>(sm_count()) * per_sm;
    if (work_items <= cap) return static_cast<int>(work_items);
    // every CTA gets the same number of items (+-1): ceil(items / ceil(items / cap)) CTAs
    const int64_t per_cta = (work_items + cap - 1) / cap;
    return static_cast<int>((work_items + per_cta - 1) / per_cta);
}

template <typename T, bool kHasW, bool kHasB>
int ln_launch_fwd(const void* x, const void* w, const void* b, void* y, float* mean, float* rstd, int64_t rows, int64_t hidden,
                  float eps, cudaStream_t st) {
    using V = Vec<T>;
    const bool vec_ok = hidden % V::kElems == 0 && aligned16(x) && aligned16(y) && (!kHasW || aligned16(w)) && (!kHasB || aligned16(b));
    const int64_t nvec = hidden / V::kElems;
    const LnRowPlan p = vec_ok ? ln_plan_row(nvec, 4) : LnRowPlan{0, 0};
    auto go = [&](auto kernel) {
        const int grid = ln_persistent_grid(kernel, p.threads, rows, 32);
        kernel<<<grid, p.threads, 0, st>>>(static_cast<const T*>(x), static_cast<const T*>(w), static_cast<const T*>(b),
                                           static_cast<T*>(y), mean, rstd, rows, static_cast<int>(nvec), static_cast<float>(hidden), eps);
        return launch_status();
    };
    switch (p.vpl) {
        case 1: return go(layernorm_fwd_kernel<T, 1, kHasW, kHasB>);
        case 2: return go(layernorm_fwd_kernel<T, 2, kHasW, kHasB>);
        case 4: return go(layernorm_fwd_kernel<T, 4, kHasW, kHasB>);
        case 8: return go(layernorm_fwd_kernel<T, 8, kHasW, kHasB>);
        default: break;
    }
    const int64_t cap = static_cast<int64_t>(sm_count()) * 8;
    layernorm_fwd_generic_kernel<T, kHasW, kHasB><<<static_cast<int>(rows < cap ? rows : cap), 256, 0, st>>>(
        static_cast<const T*>(x), static_cast<const T*>(w), static_cast<const T*>(b), static_cast<T*>(y), mean, rstd, rows, hidden, eps);
    return launch_status();
}

template <typename T>
int ln_dispatch_fwd(const void* x, const void* w, const void* b, void* y, float* mean, float* rstd, int64_t rows, int64_t hidden,
                    float eps, cudaStream_t st) {
    if (w != nullptr) {
        return b != nullptr ? ln_launch_fwd<T, true, true>(x, w, b, y, mean, rstd, rows, hidden, eps, st)
                            : ln_launch_fwd<T, true, false>(x, w, b, y, mean, rstd, rows, hidden, eps, st);
    }
    return b != nullptr ? ln_launch_fwd<T, false, true>(x, w, b, y, mean, rstd, rows, hidden, eps, st)
                        : ln_launch_fwd<T, false, false>(x, w, b, y, mean, rstd, rows, hidden, eps, st);
}

int64_t ln_bwd_parts(int64_t rows) {
    // one partial row per backward CTA; a CTA owns at least one row
    const int64_t parts = static_cast<int64_t>(sm_count()) * kLnBwdCtasPerSm;
    return rows < parts ? rows : parts;
}

// Backward tuning knobs, read once from the environment (A/B measurements only; 0 = choose from the row width):
//   B200K_LN_BWD_VPL=1|2|4  16-byte vectors per thread
//   B200K_LN_BWD_ROWS=1|2   rows per pipeline stage
//   B200K_LN_BWD_CTAS=1..4  resident CTAs per SM (splits the shared-memory ring)
struct LnTune {
    int vpl, rows, ctas_per_sm, minb;
};
LnTune ln_tune() {
    static LnTune t = [] {
        auto geti = [](const char* name) {
            const char* e = getenv(name);
            return e != nullptr && e[0] != 0 ? atoi(e) : 0;
        };
        return LnTune{geti("B200K_LN_BWD_VPL"), geti("B200K_LN_BWD_ROWS"), geti("B200K_LN_BWD_CTAS"), geti("B200K_LN_BWD_MINB")};
    }();
    return t;
}

template <typename T, bool kHasW, bool kHasB>
int ln_launch_bwd(const void* dy, const void* x, const void* w, const float* mean, const float* rstd, void* dx, void* dw, void* db,
                  void* scratch, int64_t scratch_bytes, int64_t rows, int64_t hidden, cudaStream_t st) {
    using V = Vec<T>;
    const float inv_hidden = static_cast<float>(1.0 / static_cast<double>(hidden));  // the trace's Python-side 1.0 / H
    const bool vec_ok = hidden % V::kElems == 0 && aligned16(dy) && aligned16(x) && aligned16(dx) && (!kHasW || aligned16(w));
    const int64_t nvec = hidden / V::kElems;
    const LnTune t = ln_tune();
    // Consumer vectors per thread, rows per stage, CTAs per SM.  Measured on B200 ([8192 x 4096] bf16): small CTAs
    // with several vectors per thread and two or more CTAs per SM win — a CTA's two barriers per iteration stall
    // only its own few warps while the other CTAs on the SM keep computing; more threads with one vector each
    // spend their time in the reductions.  Registers (dw / db accumulators, 16 floats per owned vector) cap a
    // thread at 4 vectors.
    int vpl = 0, kr = kLnBwdRows;
    auto fits = [&](int v, int min_threads) {
        return nvec % (32 * v) == 0 && nvec / v >= min_threads && nvec / v <= (v == 1 ? 512 : 256);
    };
    for (int min_threads : {64, 32}) {
        for (int v : {4, 2, 1}) {
            if (vpl == 0 && fits(v, min_threads)) vpl = v;
        }
    }
    if (t.vpl > 0 && fits(t.vpl, 32)) vpl = t.vpl;
    if (vpl == 4) kr = t.minb == 22 && nvec <= 512 ? 2 : 1;
    else if (t.rows == 1 || t.rows == 2) kr = t.rows;
    const int64_t cthreads = vpl > 0 ? nvec / vpl : 1;
    const int64_t stage_bytes = nvec * 16 * kr * 2;
    const int64_t threads_per_sm = vpl == 4 ? 256 : vpl == 2 ? 384 : 512;
    int ctas_per_sm = t.ctas_per_sm > 0 ? t.ctas_per_sm : static_cast<int>(threads_per_sm / cthreads);
    if (ctas_per_sm < 1) ctas_per_sm = 1;
    if (ctas_per_sm > kLnBwdCtasPerSm) ctas_per_sm = kLnBwdCtasPerSm;
    while (ctas_per_sm > 1 && (200 * 1024 / ctas_per_sm) / (stage_bytes > 0 ? stage_bytes : 1) < 3) --ctas_per_sm;
    int64_t stages = (200 * 1024 / ctas_per_sm) / (stage_bytes > 0 ? stage_bytes : 1);
    if (stages > 8) stages = 8;
    const bool tma_ok = vec_ok && vpl > 0 && stages >= 2;
    if (!tma_ok) {
        const int64_t cap = static_cast<int64_t>(sm_count()) * 8;
        layernorm_bwd_generic_dx_kernel<T, kHasW><<<static_cast<int>(rows < cap ? rows : cap), 256, 0, st>>>(
            static_cast<const T*>(dy), static_cast<const T*>(x), static_cast<const T*>(w), mean, rstd, static_cast<T*>(dx), rows,
            hidden, inv_hidden);
        int rc = launch_status();
        if (rc != 0 || (!kHasW && !kHasB)) return rc;
        layernorm_bwd_generic_cols_kernel<T><<<static_cast<int>((hidden + 31) / 32), 256, 0, st>>>(
            static_cast<const T*>(dy), static_cast<const T*>(x), mean, rstd, static_cast<T*>(dw), static_cast<T*>(db), rows, hidden);
        return launch_status();
    }
    if ((kHasW || kHasB) && (scratch == nullptr || !aligned16(scratch) || scratch_bytes < 2 * ln_bwd_parts(rows) * hidden * 4))
        return B200K_ESCRATCH;
    const int64_t groups = (rows + kr - 1) / kr;
    const int64_t cap = static_cast<int64_t>(sm_count()) * ctas_per_sm;
    const int grid = static_cast<int>(groups < cap ? groups : cap);
    float* dw_partial = static_cast<float*>(scratch);
    float* db_partial = dw_partial + static_cast<int64_t>(grid) * hidden;
    const size_t smem = ((2 * stages * 8 + 2 * 2 * kr * 32 * 4 + 127) / 128) * 128 + stages * stage_bytes;
    auto go = [&](auto kernel) {
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        if (e != cudaSuccess) return static_cast<int>(e);
        kernel<<<grid, static_cast<int>(cthreads), smem, st>>>(
            static_cast<const T*>(dy), static_cast<const T*>(x), static_cast<const T*>(w), mean, rstd, static_cast<T*>(dx), dw_partial,
            db_partial, rows, static_cast<int>(nvec), static_cast<float>(hidden), inv_hidden, static_cast<int>(stages));
        return launch_status();
    };
    int rc;
    if (vpl == 1) rc = kr == 1 ? go(layernorm_bwd_tma_kernel<T, 1, 1, kHasW, kHasB>) : go(layernorm_bwd_tma_kernel<T, 1, 2, kHasW, kHasB>);
    else if (vpl == 2 && kr == 1 && t.minb == 2) rc = go(layernorm_bwd_tma_kernel<T, 2, 1, kHasW, kHasB, 256, 2>);
    else if (vpl == 2) rc = kr == 1 ? go(layernorm_bwd_tma_kernel<T, 2, 1, kHasW, kHasB>) : go(layernorm_bwd_tma_kernel<T, 2, 2, kHasW, kHasB>);
    else if (cthreads <= 128 && t.minb == 3) rc = go(layernorm_bwd_tma_kernel<T, 4, 1, kHasW, kHasB, 128, 3>);
    else if (cthreads <= 128 && t.minb == 4) rc = go(layernorm_bwd_tma_kernel<T, 4, 1, kHasW, kHasB, 128, 4>);
    else if (cthreads <= 128 && t.minb == 22) rc = go(layernorm_bwd_tma_kernel<T, 4, 2, kHasW, kHasB, 128, 2>);
    else rc = go(layernorm_bwd_tma_kernel<T, 4, 1, kHasW, kHasB>);
    if (rc != 0 || (!kHasW && !kHasB)) return rc;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(static_cast<unsigned>((hidden + 63) / 64), 2);
    cfg.blockDim = dim3(256);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, layernorm_finalize_kernel<T>, static_cast<const float*>(dw_partial),
                                       static_cast<const float*>(db_partial), static_cast<T*>(kHasW ? dw : nullptr),
                                       static_cast<T*>(kHasB ? db : nullptr), grid, hidden);
    return e == cudaSuccess ? launch_status() : static_cast<int>(e);
}

template <typename T>
int ln_dispatch_bwd(const void* dy, const void* x, const void* w, const float* mean, const float* rstd, void* dx, void* dw, void* db,
                    void* scratch, int64_t scratch_bytes, int64_t rows, int64_t hidden, cudaStream_t st) {
    if (w != nullptr) {
        return db != nullptr ? ln_launch_bwd<T, true, true>(dy, x, w, mean, rstd, dx, dw, db, scratch, scratch_bytes, rows, hidden, st)
                             : ln_launch_bwd<T, true, false>(dy, x, w, mean, rstd, dx, dw, db, scratch, scratch_bytes, rows, hidden, st);
    }
    return db != nullptr ? ln_launch_bwd<T, false, true>(dy, x, w, mean, rstd, dx, dw, db, scratch, scratch_bytes, rows, hidden, st)
                         : ln_launch_bwd<T, false, false>(dy, x, w, mean, rstd, dx, dw, db, scratch, scratch_bytes, rows, hidden, st);
}

}  // namespace
}  // namespace b200k

using namespace b200k;

extern "C" int b200k_layernorm_fwd(const void* x, const void* weight, const void* bias, void* y, float* mean_out, float* rstd_out,
                                   int64_t rows, int64_t hidden, float eps, int dtype, void* stream) {
    if (rows < 0 || hidden <= 0) return B200K_EINVAL;
    if (rows == 0) return 0;
    if (x == nullptr || y == nullptr) return B200K_EINVAL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    switch (dtype) {
        case B200K_F32: return ln_dispatch_fwd<float>(x, weight, bias, y, mean_out, rstd_out, rows, hidden, eps, st);
        case B200K_BF16: return ln_dispatch_fwd<__nv_bfloat16>(x, weight, bias, y, mean_out, rstd_out, rows, hidden, eps, st);
        case B200K_F16: return ln_dispatch_fwd<__half>(x, weight, bias, y, mean_out, rstd_out, rows, hidden, eps, st);
        default: return B200K_EDTYPE;
    }
}

extern "C" int64_t b200k_layernorm_bwd_scratch_bytes(int64_t rows, int64_t hidden) {
    if (rows <= 0 || hidden <= 0) return 0;
    return 2 * ln_bwd_parts(rows) * hidden * static_cast<int64_t>(sizeof(float));
}

extern "C" int b200k_layernorm_bwd(const void* dy, const void* x, const void* weight, const float* mean, const float* rstd, void* dx,
                                   void* dw, void* db, void* scratch, int64_t scratch_bytes, int64_t rows, int64_t hidden, int dtype,
                                   void* stream) {
    if (rows < 0 || hidden <= 0) return B200K_EINVAL;
    if (rows > 0 && (dy == nullptr || x == nullptr || mean == nullptr || rstd == nullptr || dx == nullptr)) return B200K_EINVAL;
    if (weight != nullptr && dw == nullptr) return B200K_EINVAL;
    if (weight == nullptr && dw != nullptr) return B200K_EINVAL;
    if (dtype != B200K_F32 && dtype != B200K_BF16 && dtype != B200K_F16) return B200K_EDTYPE;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (rows == 0) {
        // An empty batch still defines dw = db = 0.
        const size_t bytes = (dtype == B200K_F32 ? 4 : 2) * static_cast<size_t>(hidden);
        for (void* p : {dw, db}) {
            if (p == nullptr) continue;
            cudaError_t e = cudaMemsetAsync(p, 0, bytes, st);
            if (e != cudaSuccess) return static_cast<int>(e);
        }
        return 0;
    }
    switch (dtype) {
        case B200K_F32: return ln_dispatch_bwd<float>(dy, x, weight, mean, rstd, dx, dw, db, scratch, scratch_bytes, rows, hidden, st);
        case B200K_BF16:
            return ln_dispatch_bwd<__nv_bfloat16>(dy, x, weight, mean, rstd, dx, dw, db, scratch, scratch_bytes, rows, hidden, st);
        default: return ln_dispatch_bwd<__half>(dy, x, weight, mean, rstd, dx, dw, db, scratch, scratch_bytes, rows, hidden, st);
    }
}
