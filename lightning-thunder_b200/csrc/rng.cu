// Counter-based uniform random numbers (prims.uniform_philox, SURVEY §8 row f.1: dropout masks).
//
// Convention of the reference's fusion executor ("the nvFuser way", thunder/executors/torchex.py:478-493): the
// `offset` argument is the generator's Philox offset divided by 4 and is used as the 64-bit counter as it is;
// element i (row-major) takes component i % 4 of the Philox-4x32-10 block whose subsequence is i / 4.  A thread
// owns 8 consecutive elements = two blocks, every generated word is used, and the stores are 16 bytes.
// Integer-bound by design (10 rounds x 2 wide multiplies per block): ~15 integer instructions per element.
#include "common.cuh"

namespace b200k {
namespace {

constexpr uint32_t kM0 = 0xD2511F53u, kM1 = 0xCD9E8D57u, kW0 = 0x9E3779B9u, kW1 = 0xBB67AE85u;

__device__ __forceinline__ uint4 philox4x32_10(uint2 key, uint4 c) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(kM0, c.x), lo0 = kM0 * c.x, hi1 = __umulhi(kM1, c.z), lo1 = kM1 * c.z;
        c = make_uint4(hi1 ^ c.y ^ key.x, lo1, hi0 ^ c.w ^ key.y, lo0);
        key.x += kW0;
        key.y += kW1;
    }
    return c;
}

// (0, 1] as curand_uniform does it, then the closed end is mapped to 0
__device__ __forceinline__ float to_unit(uint32_t x) {
    const float u = __fmaf_rn(static_cast<float>(x), 2.3283064365386963e-10f, 1.1641532182693481e-10f);
    return u == 1.0f ? 0.0f : u;
}

template <typename T, bool kUnit>
__global__ void __launch_bounds__(256)
uniform_philox_kernel(T* __restrict__ out, int64_t numel, uint64_t seed, uint64_t offset, float lo, float hi) {
    using V = Vec<T>;
    const uint2 key = make_uint2(static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
    const uint32_t c_lo = static_cast<uint32_t>(offset), c_hi = static_cast<uint32_t>(offset >> 32);
    const int64_t items = (numel + 7) / 8;
    for (int64_t item = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; item < items;
         item += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const uint64_t s0 = static_cast<uint64_t>(item) * 2;
        const uint4 a = philox4x32_10(key, make_uint4(c_lo, c_hi, static_cast<uint32_t>(s0), static_cast<uint32_t>(s0 >> 32)));
        const uint4 b = philox4x32_10(key, make_uint4(c_lo, c_hi, static_cast<uint32_t>(s0 + 1), static_cast<uint32_t>((s0 + 1) >> 32)));
        const uint32_t x[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            float u = to_unit(x[e]);
            if constexpr (!kUnit) u = __fmaf_rn(hi - lo, u, lo);
            u = V::rnd(u);
            v[e] = u == hi ? lo : u;  // a value that rounds up to the open end
        }
        const int64_t base = item * 8;
        if (base + 8 <= numel && (reinterpret_cast<uintptr_t>(out + base) & 15u) == 0) {
            if constexpr (V::kElems == 8) {
                st_stream(out + base, V::pack(v));
            } else {
                st_stream(out + base, V::pack(v));
                st_stream(out + base + 4, V::pack(v + 4));
            }
        } else {
            for (int e = 0; e < 8 && base + e < numel; ++e) V::store1(out, base + e, v[e]);
        }
    }
}

template <typename T>
int run_uniform(void* out, int64_t numel, uint64_t seed, uint64_t offset, float lo, float hi, cudaStream_t st) {
    const int64_t items = (numel + 7) / 8;
    const int64_t want = (items + 255) / 256;
    const int64_t cap = static_cast<int64_t>(sm_count()) * 8;
    const int grid = static_cast<int>(want < cap ? want : cap);
    if (lo == 0.0f && hi == 1.0f)
        uniform_philox_kernel<T, true><<<grid, 256, 0, st>>>(static_cast<T*>(out), numel, seed, offset, lo, hi);
    else
        uniform_philox_kernel<T, false><<<grid, 256, 0, st>>>(static_cast<T*>(out), numel, seed, offset, lo, hi);
    return launch_status();
}

}  // namespace
}  // namespace b200k

using namespace b200k;

extern "C" int b200k_uniform_philox(void* out, int64_t numel, uint64_t seed, uint64_t offset, float minval, float maxval, int dtype,
                                    void* stream) {
    if (numel < 0 || !(minval < maxval)) return B200K_EINVAL;
    if (numel == 0) return 0;
    if (out == nullptr) return B200K_EINVAL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    switch (dtype) {
        case B200K_F32: return run_uniform<float>(out, numel, seed, offset, minval, maxval, st);
        case B200K_BF16: return run_uniform<__nv_bfloat16>(out, numel, seed, offset, minval, maxval, st);
        case B200K_F16: return run_uniform<__half>(out, numel, seed, offset, minval, maxval, st);
        default: return B200K_EDTYPE;
    }
}
