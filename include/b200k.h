/*
 * b200k — C-ABI kernel library behind the `b200` Thunder fusion executor.
 *
 * Every entry point is what the reference's fusion executor would bind for one fused region
 * family: where the reference hands a region to an external kernel provider
 * (thunder/executors/nvfuserex_impl.py:567 `fd.execute(args)`), b200ex calls one of these.
 * Plain pointers and sizes only: no torch types.  All pointers are DEVICE pointers unless
 * stated otherwise, `stream` is a cudaStream_t passed as void*, and every call is asynchronous
 * on that stream.  Nothing here allocates or synchronises.
 *
 * Return value: 0 on success, a positive cudaError_t value if the launch failed, or a negative
 * B200K_E* code for a rejected argument.  `b200k_error_string` decodes either.
 *
 * dtype codes (B200K_F32 / B200K_BF16 / B200K_F16) name the storage type of the activations;
 * arithmetic is always fp32, rounding to the storage type exactly where the reference's
 * torch-executor trace rounds (thunder/executors/torchex.py:80-87 `convert_element_type`).
 */
#ifndef B200K_H
#define B200K_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define B200K_API __attribute__((visibility("default")))
#else
#define B200K_API
#endif

#define B200K_F32 0
#define B200K_BF16 1
#define B200K_F16 2

#define B200K_EINVAL (-1)     /* bad pointer / size / flag */
#define B200K_EDTYPE (-2)     /* unsupported dtype code */
#define B200K_EALIGN (-3)     /* pointer or row pitch not 16-byte aligned */
#define B200K_ESCRATCH (-4)   /* scratch buffer too small */
#define B200K_ELIMIT (-5)     /* size beyond a documented maximum */

/* Library ABI version (bumped on any signature change). */
B200K_API int b200k_version(void);
/* Static string for a return code of any entry point. Never NULL. */
B200K_API const char* b200k_error_string(int code);
/* Number of SMs of the current device (cached per device; used by callers to size scratch). */
B200K_API int b200k_sm_count(void);

/* ---------------------------------------------------------------------------------------------
 * RMSNorm (+ residual add) forward — SURVEY §8 row a5.
 * Replaces the fused region  `h = x + r; y = rms_norm(h, w, eps)`  that the reference builds from
 * ltorch.add + ltorch.rms_norm (thunder/torch/__init__.py:4450-4479) and runs through
 * nvfuserex_impl.py:544-573.
 *   h    = round_T(f32(x) + f32(residual))                (h = x when residual == NULL)
 *   rstd = rsqrt(sum_j f32(h)^2 / hidden + eps)           fp32 [rows]
 *   y    = round_T((f32(h) * rstd) * f32(weight))         (weight == NULL -> no scale)
 * x, residual, h_out, y: [rows, hidden] row-major contiguous of `dtype`; weight: [hidden] of `dtype`.
 * h_out may be NULL when residual is NULL.  rstd_out may be NULL.
 * --------------------------------------------------------------------------------------------- */
B200K_API int b200k_rmsnorm_fwd(const void* x, const void* residual, const void* weight,
                      void* h_out, void* y, float* rstd_out,
                      int64_t rows, int64_t hidden, float eps, int dtype, void* stream);

/* Size in bytes of the scratch buffer b200k_rmsnorm_bwd needs for (rows, hidden). */
B200K_API int64_t b200k_rmsnorm_bwd_scratch_bytes(int64_t rows, int64_t hidden);

/* ---------------------------------------------------------------------------------------------
 * RMSNorm (+ residual-gradient add) backward — SURVEY §8 row a6.
 * Replaces the backward region autodiff derives from the prim vjps of a5
 * (thunder/core/transforms.py vjp rules; region census SURVEY Appendix C).
 *   g    = f32(weight) * f32(dy)
 *   c    = sum_j f32(h) * g
 *   s    = ((-0.5 * c) * rstd^3) / hidden
 *   dh   = round_T((rstd * g + f32(h) * s) + f32(h) * s)
 *   dx   = round_T(f32(dh_in) + f32(dh))                  (dx = dh when dh_in == NULL)
 *   dw   = round_T(sum_rows f32(round_T((f32(h) * rstd) * f32(dy))))
 * dy, h, dh_in, dx: [rows, hidden] of `dtype`; rstd: fp32 [rows]; weight, dw: [hidden] of `dtype`.
 * weight == NULL means "no scale": g = dy and dw is not written (dw may then be NULL).
 * scratch: device buffer of at least b200k_rmsnorm_bwd_scratch_bytes(rows, hidden) bytes, 16-byte
 * aligned; contents on entry are ignored, contents on exit are unspecified.
 * --------------------------------------------------------------------------------------------- */
B200K_API int b200k_rmsnorm_bwd(const void* dy, const void* h, const void* weight, const float* rstd,
                      const void* dh_in, void* dx, void* dw,
                      void* scratch, int64_t scratch_bytes,
                      int64_t rows, int64_t hidden, int dtype, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200K_H */
