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

/* ---------------------------------------------------------------------------------------------
 * RMSNorm backward for callers that must not allocate.  (No counterpart in the reference: it exists for
 * hosts of this library that cannot call cudaMalloc or any allocating routine at the point where the
 * backward runs — a runtime that has handed all device memory to its own pool and forbids driver
 * allocations, a real-time or signal-handling context, or a launch path that must stay allocation-free
 * between two points of a stream.)
 *
 * Same computation, arguments and results as b200k_rmsnorm_bwd, except that the scratch for the dw partial
 * rows is a fixed array of B200K_RMSNORM_BWD_NOALLOC_SCRATCH_BYTES inside the library's device image; every
 * GPU that loads the library holds that much memory until the process exits.  The caller still passes all
 * inputs and outputs.  The call itself performs no cudaMalloc / cudaFree and calls no allocating routine.
 *
 *  - Limit: returns B200K_ELIMIT ("size exceeds a documented maximum") when
 *    b200k_rmsnorm_bwd_scratch_bytes(rows, hidden) exceeds the array (with weight != NULL).  On a 148-SM
 *    B200 the array covers hidden <= 7084 elements for any number of rows, e.g. [8192 x 4096].
 *  - First call on each device: resolves the array's device address (cudaGetSymbolAddress, with that
 *    device current) and caches it.  Make that first call before starting any stream capture.
 *  - The array is shared by all calls and keeps the previous call's contents.  Nothing in it has to start
 *    at zero (every partial row is written before it is read), but two calls whose kernels can overlap on
 *    the GPU — different streams without an ordering between them — race on it.  Serialise them.
 * --------------------------------------------------------------------------------------------- */
#define B200K_RMSNORM_BWD_NOALLOC_SCRATCH_BYTES (16u << 20)
B200K_API int b200k_rmsnorm_bwd_noalloc(const void* dy, const void* h, const void* weight, const float* rstd,
                                        const void* dh_in, void* dx, void* dw, int64_t rows, int64_t hidden,
                                        int dtype, void* stream);

/* ---------------------------------------------------------------------------------------------
 * SwiGLU forward — SURVEY §8 row a7.  Replaces the region `silu(a) * b` built from ltorch.silu
 * (thunder/torch/__init__.py:2266 -> thunder/clang/__init__.py:1417) + ltorch.mul.
 *   u = round_T(a / (1 + exp(-a)));  out = round_T(f32(u) * f32(b))
 * a, b, out: n contiguous elements of `dtype`.
 * --------------------------------------------------------------------------------------------- */
B200K_API int b200k_swiglu_fwd(const void* a, const void* b, void* out, int64_t n, int dtype, void* stream);

/* ---------------------------------------------------------------------------------------------
 * SwiGLU backward — SURVEY §8 row a8 (autodiff of a7; silu recomputed in the region by
 * thunder/core/rematerialization.py:407).  g is the cotangent of out.
 *   db = round_T(u * g);  gu = round_T(b * g)
 *   s  = round_T(1 / round_T(1 + round_T(exp(-a))))
 *   da = round_T(round_T(gu * s) * round_T(1 + round_T(a * round_T(1 - s))))
 * --------------------------------------------------------------------------------------------- */
B200K_API int b200k_swiglu_bwd(const void* a, const void* b, const void* g, void* da, void* db, int64_t n, int dtype,
                               void* stream);

/* ---------------------------------------------------------------------------------------------
 * Softmax cross-entropy forward — SURVEY §8 row a12.  Replaces ltorch.cross_entropy
 * (thunder/torch/__init__.py:5226-5270; log_softmax :5839-5851 + nll_loss :6018-6031), the op the
 * reference claims in thunder/executors/nvfuserex_impl.py:2967-3018.
 * logits [rows, vocab] of `dtype` row-major; target int64 [rows]; reduction 1 = mean, 2 = sum.
 * Per row (streamed once, online softmax): row_max = max_j x, row_logsum = log sum_j exp(x - max),
 * row_nll = -round_T((x[target] - max) - logsum), or 0 when target == ignore_index.
 * loss (one element of `dtype`, may be NULL to skip) = round_T(round_T(sum row_nll) / n_valid) for mean,
 * round_T(sum row_nll) for sum; n_valid (int64, may be NULL) = #rows with target != ignore_index.
 * row_max, row_logsum, row_nll: fp32 [rows] outputs (the first two are what backward needs).
 * --------------------------------------------------------------------------------------------- */
B200K_API int b200k_cross_entropy_fwd(const void* logits, const int64_t* target, float* row_max, float* row_logsum,
                                      float* row_nll, void* loss, int64_t* n_valid, int64_t rows, int64_t vocab,
                                      int64_t ignore_index, int reduction, int dtype, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Softmax cross-entropy backward — SURVEY §8 row a13.  Replaces nll_loss_backward
 * (thunder/torch/__init__.py:6035-6047, thunder/executors/torchex.py:1731-1783) + log_softmax_backward
 * (:5859-5864); the reference's fused variant is nvfuserex_impl.py:3043-3094.
 *   d = -round_T(g / round_T(n_valid))  (mean)   or  -g  (sum);   logp = round_T((x - max) - logsum)
 *   dlogits[i, j] = round_T([j == target_i] * d - exp(logp) * d)   for valid rows, 0 for ignored rows
 * g: one element of `dtype` (device pointer); n_valid: the int64 the forward wrote (mean only).
 * --------------------------------------------------------------------------------------------- */
B200K_API int b200k_cross_entropy_bwd(const void* g, const void* logits, const int64_t* target, const float* row_max,
                                      const float* row_logsum, const int64_t* n_valid, void* dlogits, int64_t rows,
                                      int64_t vocab, int64_t ignore_index, int reduction, int dtype, void* stream);

/* ---------------------------------------------------------------------------------------------
 * QKV split + RoPE (+ GQA expand) forward — SURVEY §8 row a9.  Replaces the region traced from
 * thunder/benchmarks/__init__.py:2282-2328 (vendored LitGPT attention slice; litgpt apply_rope),
 * which the reference routes to a fusion executor (thunder/executors/torch_compile.py:198-234).
 * qkv [B, T, G, q_per_kv + 2, hs] contiguous; cos, sin [T, rope_n_elem];
 * q [B, G*q_per_kv, T, hs]; k, v [B, Hkv, T, hs] with Hkv = G*q_per_kv if expand_kv else G.
 * For q and k heads, over the first rope_n_elem features (half = rope_n_elem / 2):
 *   out[j] = round_T(round_T(x[j] * cos[j]) + round_T(rot[j] * sin[j])),  rot = cat(-x[half:], x[:half])
 * the remaining features and all of v are copied.
 * --------------------------------------------------------------------------------------------- */
B200K_API int b200k_qkv_rope_fwd(const void* qkv, const void* cos, const void* sin, void* q, void* k, void* v,
                                 int64_t B, int64_t T, int n_query_groups, int q_per_kv, int head_size, int rope_n_elem,
                                 int expand_kv, int dtype, void* stream);

/* ---------------------------------------------------------------------------------------------
 * QKV split + RoPE backward — SURVEY §8 row a10 (slice vjp = pad, cat vjp = slice, expand vjp = sum
 * over q_per_kv in fp32 rounded once, final cat into dqkv).  Shapes as in the forward; writes
 * dqkv [B, T, G, q_per_kv + 2, hs].
 * --------------------------------------------------------------------------------------------- */
B200K_API int b200k_qkv_rope_bwd(const void* dq, const void* dk, const void* dv, const void* cos, const void* sin,
                                 void* dqkv, int64_t B, int64_t T, int n_query_groups, int q_per_kv, int head_size,
                                 int rope_n_elem, int expand_kv, int dtype, void* stream);

/* ---------------------------------------------------------------------------------------------
 * bias + activation forward — SURVEY §8 row a11.  act: 0 = none, 1 = relu (ltorch.relu ->
 * where(a > 0, a, 0)), 2 = gelu, erf form (thunder/torch/__init__.py:2071-2082).
 *   a = round_T(x + bias)  (a = x when bias == NULL);  out = act(a) with one round_T per ltorch op.
 * x, out [rows, hidden]; bias [hidden]; pre_out (may be NULL): also receives a, the pre-activation that
 * autodiff saves for the backward.
 * --------------------------------------------------------------------------------------------- */
B200K_API int b200k_bias_act_fwd(const void* x, const void* bias, void* out, void* pre_out, int64_t rows, int64_t hidden,
                                 int act, int dtype, void* stream);

/* Scratch bytes b200k_bias_act_bwd needs when bias != NULL. */
B200K_API int64_t b200k_bias_act_bwd_scratch_bytes(int64_t rows, int64_t hidden);

/* ---------------------------------------------------------------------------------------------
 * bias + activation backward — SURVEY §8 row a11.  With a = round_T(x + bias) (a = x when bias == NULL, i.e.
 * when the caller passes the saved pre-activation): dx = act'(a) applied to g with the reference's rounding
 * points; dbias = round_T(sum_rows f32(dx)), written iff dbias != NULL (then scratch is required).
 * scratch as for b200k_rmsnorm_bwd (contents ignored on entry, unspecified on exit).
 * --------------------------------------------------------------------------------------------- */
B200K_API int b200k_bias_act_bwd(const void* x, const void* bias, const void* g, void* dx, void* dbias, void* scratch,
                                 int64_t scratch_bytes, int64_t rows, int64_t hidden, int act, int dtype, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Column sum — the `sum` over the leading dims that follows a pointwise region (bias gradients and the like;
 * prims.sum claimed by nvFuser at thunder/executors/nvfuserex_impl.py:2112-2200).
 *   out[c] = round_T(sum_r f32(x[r, c]))     x: [rows, hidden] of `dtype`, out: [hidden] of `dtype`
 * scratch: b200k_bias_act_bwd_scratch_bytes(rows, hidden) bytes, contents ignored / unspecified.
 * --------------------------------------------------------------------------------------------- */
B200K_API int b200k_colsum(const void* x, void* out, void* scratch, int64_t scratch_bytes, int64_t rows, int64_t hidden,
                           int dtype, void* stream);

/* ---------------------------------------------------------------------------------------------
 * LayerNorm over the last dimension, forward — SURVEY §8 row f.1.
 * Replaces the region the reference builds from ltorch.layer_norm -> _native_layer_norm -> _normalize
 * (thunder/torch/__init__.py:4406-4447): var_mean, rsqrt, scale, shift.
 *   mean = sum_j f32(x) / hidden,  var = sum_j (f32(x) - mean)^2 / hidden     fp32 [rows]
 *   rstd = rsqrt(var + eps)                                                    fp32 [rows]
 *   y    = round_T((((f32(x) - mean) * rstd) * f32(weight)) + f32(bias))       (NULL weight / bias: step skipped)
 * x, y: [rows, hidden] row-major contiguous of `dtype`; weight, bias: [hidden] of `dtype`.
 * mean_out / rstd_out may be NULL.
 * --------------------------------------------------------------------------------------------- */
B200K_API int b200k_layernorm_fwd(const void* x, const void* weight, const void* bias, void* y, float* mean_out,
                        float* rstd_out, int64_t rows, int64_t hidden, float eps, int dtype, void* stream);

/* Size in bytes of the scratch buffer b200k_layernorm_bwd needs for (rows, hidden). */
B200K_API int64_t b200k_layernorm_bwd_scratch_bytes(int64_t rows, int64_t hidden);

/* ---------------------------------------------------------------------------------------------
 * LayerNorm backward — SURVEY §8 row f.1.  The program autodiff derives from the forward above: the
 * mul / add / sub / rsqrt vjps plus the var_mean rule (thunder/core/transforms.py:1292-1312).
 *   g      = f32(dy),  gw = f32(weight) * g,  xc = f32(x) - mean
 *   direct = rstd * gw
 *   c1     = sum_j xc * gw,  c2 = sum_j -direct
 *   stats  = (1/hidden) * c2 + ((2 * ((-0.5 * c1) * rstd^3)) * xc) / hidden
 *   dx     = round_T(round_T(direct) + round_T(stats))      (fp32: direct + stats)
 *   dw     = round_T(sum_rows round_T((xc * rstd) * g)),  db = round_T(sum_rows g)
 * weight NULL <-> dw NULL; db is written iff non-NULL.  mean, rstd: fp32 [rows] from the forward.
 * scratch: device buffer of at least b200k_layernorm_bwd_scratch_bytes(rows, hidden) bytes, 16-byte aligned
 * (contents ignored on entry, unspecified on exit); may be NULL when dw and db are both NULL.
 * --------------------------------------------------------------------------------------------- */
B200K_API int b200k_layernorm_bwd(const void* dy, const void* x, const void* weight, const float* mean, const float* rstd,
                        void* dx, void* dw, void* db, void* scratch, int64_t scratch_bytes, int64_t rows,
                        int64_t hidden, int dtype, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Multi-tensor AdamW / Adam step — SURVEY §8 row f.4 (the optimizer step next to the fused regions).
 * Replaces torch.optim.AdamW(fused=True) as the reference's training benchmark uses it
 * (thunder/benchmarks/benchmark_litgpt.py:278-282), i.e. ATen's _fused_adamw_: per element, with the scalar
 * hyper-parameters narrowed to fp32 as there,
 *   p -= lr * weight_decay * p                               (decoupled; else g += p * weight_decay)
 *   m  = beta1 * m + (1 - beta1) * g,   v = beta2 * v + (1 - beta2) * g * g
 *   p -= (lr / bias_correction1) * m / (sqrt(v) / bias_correction2_sqrt + eps)          all in fp32
 * p, m, v are rounded to `dtype` on store.  table: DEVICE array of tensors (all of `dtype`, 16-byte aligned);
 * chunk_tensor / chunk_index: DEVICE arrays of n_chunks entries listing every chunk of b200k_adamw_chunk_elems(dtype)
 * elements of every tensor (tensor index, chunk number within the tensor).  bias_correction1 = 1 - beta1^step,
 * bias_correction2_sqrt = sqrt(1 - beta2^step), computed by the caller in double and narrowed to float.
 * --------------------------------------------------------------------------------------------- */
typedef struct {
    void* param;
    const void* grad;
    void* exp_avg;
    void* exp_avg_sq;
    int64_t numel;
} b200k_adam_tensor;
B200K_API int64_t b200k_adamw_chunk_elems(int dtype);
B200K_API int b200k_adamw(const b200k_adam_tensor* table, const int32_t* chunk_tensor, const int32_t* chunk_index,
                int64_t n_chunks, double lr, double beta1, double beta2, double eps, double weight_decay,
                float bias_correction1, float bias_correction2_sqrt, int decoupled_weight_decay, int dtype, void* stream);

/* ---------------------------------------------------------------------------------------------
 * softmax over the last dimension — SURVEY §8 row f.1.
 * Forward replaces ltorch.softmax (thunder/torch/__init__.py `_softmax`), which the reference's torch executor runs
 * as one torch.softmax call: fp32 inside, one rounding.
 *   y = round_T(exp(f32(x) - max_j f32(x)) / sum_j exp(f32(x) - max_j f32(x)))
 * Backward is the program autodiff derives for it, one storage-dtype rounding per op (mul, sum, sub, mul):
 *   s  = round_T(sum_j round_T(y * dy))          fp32 accumulation inside the sum
 *   dx = round_T(y * round_T(dy - s))
 * x, y, dy, dx: [rows, cols] row-major contiguous of `dtype`.
 * --------------------------------------------------------------------------------------------- */
B200K_API int b200k_softmax_fwd(const void* x, void* y, int64_t rows, int64_t cols, int dtype, void* stream);
B200K_API int b200k_softmax_bwd(const void* dy, const void* y, void* dx, int64_t rows, int64_t cols, int dtype,
                      void* stream);

/* ---------------------------------------------------------------------------------------------
 * Counter-based uniform random numbers — SURVEY §8 row f.1 (dropout masks).
 * Replaces prims.uniform_philox (thunder/core/prims.py:3007-3045) as the reference's fusion executor evaluates it
 * (thunder/executors/nvfuserex_impl.py:1147-1181), with the seed / offset pair of prims.get_and_update_rng_state
 * (thunder/executors/torchex.py:478-493: offset = the generator's Philox offset / 4, advanced by one per call).
 *   block(s) = Philox-4x32-10(key = seed, counter = (offset, subsequence s))              four 32-bit words
 *   u(i)     = word (i % 4) of block(i / 4) * 2^-32 + 2^-33, with 1.0 mapped to 0.0      fp32 in [0, 1)
 *   out[i]   = round_T(minval + (maxval - minval) * u(i)), a value that rounds to maxval becomes minval
 * out: `numel` contiguous elements of `dtype`.  The same (seed, offset) always gives the same values, on any grid.
 * --------------------------------------------------------------------------------------------- */
B200K_API int b200k_uniform_philox(void* out, int64_t numel, uint64_t seed, uint64_t offset, float minval, float maxval,
                         int dtype, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Generic pointwise region — SURVEY §8 rows a3/a4.  Executes a fused region made only of elementwise
 * prims, convert_element_type and broadcast_in_dim (the claimable vocabulary of
 * thunder/executors/nvfuserex_impl.py:971-2514 restricted to pointwise work) as ONE pass over memory:
 * a small register program interpreted per 4-element vector, fp32 arithmetic, explicit rounding
 * instructions where the trace converts to a 16-bit dtype.  The b200 executor compiles the region's
 * bound symbols to this program (b200ex/pointwise.py); regions it recognises as a known family go to
 * the specialised entry points above instead.
 *
 * Iteration space: a 3-D box [d0, d1, d2] (the broadcast output shape with adjacent dims merged).
 * Every operand gives element strides (s0, s1) for the outer dims and `inner`, its element stride along d2
 * (1 = unit stride, 0 = broadcast; any other value selects the scalar kernel).  Outputs need inner = 1.
 * Registers 0 .. n_in-1 are preloaded with the inputs converted to fp32; output k stores register
 * out_reg[k] converted (round-to-nearest-even) to its dtype.  Bool tensors use B200K_BOOL (1 byte, 0/1);
 * comparison results are 0.0f / 1.0f.
 *
 * Instruction word: op | dst << 8 | a << 16 | b << 24.  B200K_PW_CONST and B200K_PW_WHERE take one extra
 * word (the fp32 bit pattern of the constant / the register index c of `where(a, b, c)`).
 * --------------------------------------------------------------------------------------------- */
#define B200K_BOOL 3
#define B200K_PW_MAX_IN 10
#define B200K_PW_MAX_OUT 6
#define B200K_PW_MAX_REG 48
#define B200K_PW_MAX_CODE 192
/* Register-resident form: at most B200K_PW_SMALL_REGS live values (constants are immediates and take none),
 * 8 elements per thread, the value file stays in hardware registers.  Instruction = 3 words:
 *   w0 = op | dst << 8 | a << 16 | b << 24   (a, b: register 0..7, or B200K_PW_IMM = the immediate in w2)
 *   w1 = c  (third operand register of `where(a, b, c)`, or B200K_PW_IMM)
 *   w2 = fp32 bit pattern of the immediate
 * Registers 0 .. n_in-1 hold the inputs on entry.  Short programs (residual adds, scalings, bias + activation
 * chains) run at memory speed in this form; b200k_pointwise uses it whenever n_small > 0 and every unit-stride
 * operand is 16-byte aligned with d2 a multiple of 8, and falls back to `code` otherwise. */
#define B200K_PW_SMALL_REGS 8
#define B200K_PW_SMALL_MAX_INSTR 40
#define B200K_PW_IMM 0xFE

enum b200k_pw_op {
    B200K_PW_CONST = 0, B200K_PW_MOV, B200K_PW_ADD, B200K_PW_SUB, B200K_PW_MUL, B200K_PW_DIV, B200K_PW_NEG, B200K_PW_ABS,
    B200K_PW_EXP, B200K_PW_LOG, B200K_PW_SQRT, B200K_PW_RSQRT, B200K_PW_RECIP, B200K_PW_ERF, B200K_PW_TANH, B200K_PW_POW,
    B200K_PW_MAX, B200K_PW_MIN, B200K_PW_GT, B200K_PW_GE, B200K_PW_LT, B200K_PW_LE, B200K_PW_EQ, B200K_PW_NE,
    B200K_PW_WHERE, B200K_PW_RND_BF16, B200K_PW_RND_F16, B200K_PW_AND, B200K_PW_OR, B200K_PW_NOT, B200K_PW_SIGN,
    B200K_PW_FLOOR, B200K_PW_CEIL, B200K_PW_EXP2, B200K_PW_LOG2, B200K_PW_LOG1P, B200K_PW_EXPM1, B200K_PW_SIN, B200K_PW_COS,
    B200K_PW_ERFC, B200K_PW_SIGMOID, B200K_PW_TRUNC, B200K_PW_ROUND, B200K_PW_NOP_COUNT
};

typedef struct b200k_pw_operand {
    void* ptr;
    int64_t s0, s1; /* element strides of the two outer dims */
    int32_t inner;  /* element stride along d2: 1 unit stride, 0 broadcast, else strided (scalar kernel) */
    int32_t dtype;  /* B200K_F32 / BF16 / F16 / BOOL */
} b200k_pw_operand;

typedef struct b200k_pw_program {
    int64_t d0, d1, d2;
    int32_t n_in, n_out, n_code, n_reg;
    b200k_pw_operand in[B200K_PW_MAX_IN];
    b200k_pw_operand out[B200K_PW_MAX_OUT];
    int32_t out_reg[B200K_PW_MAX_OUT];
    uint32_t code[B200K_PW_MAX_CODE];
    /* Optional second encoding of the same program for the register-resident interpreter (see below);
     * n_small = 0 when the program does not fit it. */
    int32_t n_small;                                 /* number of 3-word instructions */
    int32_t small_out_reg[B200K_PW_MAX_OUT];
    uint32_t small_code[3 * B200K_PW_SMALL_MAX_INSTR];
} b200k_pw_program;

B200K_API int b200k_pointwise(const b200k_pw_program* program /* host pointer */, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200K_H */
