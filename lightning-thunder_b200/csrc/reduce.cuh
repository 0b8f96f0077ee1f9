// Second stage of the column reductions (RMSNorm dw, bias gradients): fold per-CTA fp32 partial rows.
#pragma once

#include "common.cuh"

namespace b200k {

// out[col] = round_T(sum_g partial[g][col]).  CTA = 16 float4 column groups (64 columns) x 16 part slices;
// every thread has all of its (nparts / 16) 16-byte loads in flight at once, so the kernel costs about
// one L2 round trip.  Launched with programmatic stream serialization: griddepcontrol.wait holds it
// until the producing grid has completed and flushed, so its launch latency hides under that grid.
template <typename T>
__global__ void __launch_bounds__(256)
colsum_finalize_kernel(const float* __restrict__ partial, T* __restrict__ dw, int nparts, int64_t hidden) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
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
        Vec<T>::store1(dw, col, t.x);
        Vec<T>::store1(dw, col + 1, t.y);
        Vec<T>::store1(dw, col + 2, t.z);
        Vec<T>::store1(dw, col + 3, t.w);
    }
}

template <typename T>
inline int launch_colsum_finalize(const void* scratch, void* dw, int nparts, int64_t hidden, cudaStream_t st) {
    // hidden is a multiple of 4 on every path that reaches here (vectorised kernels only).
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(static_cast<unsigned>((hidden + 63) / 64));
    cfg.blockDim = dim3(256);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, colsum_finalize_kernel<T>, static_cast<const float*>(scratch), static_cast<T*>(dw), nparts, hidden);
    return e == cudaSuccess ? launch_status() : static_cast<int>(e);
}

}  // namespace b200k
