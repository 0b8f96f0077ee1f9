"""Run-time specialisation of generic pointwise regions (NVRTC), the fast path of ``PwLaunch``.

The ahead-of-time interpreter (``b200k_pointwise``) executes any pointwise program but pays for decoding
it per element; this module instead emits the program as straight-line CUDA C inside a hand-written
streaming skeleton (16-byte coalesced loads with several work items in flight per thread, fp32 math with
the explicit rounding points of the program, 16-byte streaming stores), compiles it once per distinct
program with NVRTC for ``sm_100a`` and launches the cubin on the caller's stream — the job nvFuser's
pointwise scheduler does behind ``FusionDefinition.execute`` (thunder/executors/nvfuserex_impl.py:567).
Kernels are cached per (program signature, operand layout class) and shared by every region with the
same dataflow (all layers of a model, all FSDP parameter shards, ...).

It is used when ``cuda.bindings`` (cuda-python) with NVRTC is importable and the launch is vectorisable
(inner extent a multiple of 8, 16-byte aligned unit-stride operands); anything else runs on the
interpreter.  ``B200EX_POINTWISE_JIT=0`` disables it.
"""

from __future__ import annotations

import ctypes
import os
import threading

import torch

MAX_IN, MAX_OUT = 10, 6
MAX_GEN = 4  # random leaves generated inside one kernel (two 64-bit scalars each)
_lock = threading.Lock()
_kernels: dict = {}
_state = {"checked": False, "ok": False, "why": ""}


class JitUnavailable(Exception):
    pass


def enabled() -> bool:
    if os.environ.get("B200EX_POINTWISE_JIT", "1") == "0":
        return False
    if not _state["checked"]:
        with _lock:
            if not _state["checked"]:
                try:
                    from cuda.bindings import driver, nvrtc  # noqa: F401

                    _state["ok"] = True
                except Exception as e:  # pragma: no cover - depends on the host
                    _state["why"] = f"{type(e).__name__}: {e}"
                _state["checked"] = True
    return _state["ok"]


# --------------------------------------------------------------------------------------------------
# Code generation
# --------------------------------------------------------------------------------------------------
_SKELETON_HEAD = r"""
typedef unsigned int u32; typedef unsigned short u16; typedef unsigned char u8; typedef long long i64;
struct Op { void* ptr; i64 s0, s1; int inner; int dtype; };
struct Params { i64 d0, d1, d2; i64 sc[%(MAX_SC)d]; Op in[%(MAX_IN)d]; Op out[%(MAX_OUT)d]; };
struct V4 { u32 x, y, z, w; };
__device__ __forceinline__ V4 ldg16(const void* p) { V4 r; asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%%0,%%1,%%2,%%3}, [%%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p)); return r; }
__device__ __forceinline__ void stg16(void* p, V4 v) { asm volatile("st.global.L1::no_allocate.v4.u32 [%%0], {%%1,%%2,%%3,%%4};" :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory"); }
__device__ __forceinline__ float bf2f(u32 h) { return __uint_as_float(h << 16); }
__device__ __forceinline__ u32 f2bf2(float lo, float hi) { u32 r; asm("cvt.rn.bf16x2.f32 %%0, %%1, %%2;" : "=r"(r) : "f"(hi), "f"(lo)); return r; }
__device__ __forceinline__ float h2f(u32 h) { float f; asm("{.reg .b16 t; cvt.u16.u32 t, %%1; cvt.f32.f16 %%0, t;}" : "=f"(f) : "r"(h)); return f; }
__device__ __forceinline__ u32 f2h2(float lo, float hi) { u32 r; asm("cvt.rn.f16x2.f32 %%0, %%1, %%2;" : "=r"(r) : "f"(hi), "f"(lo)); return r; }
__device__ __forceinline__ float rnd_bf16(float x) { return __uint_as_float(f2bf2(x, 0.f) << 16); }
__device__ __forceinline__ float rnd_f16(float x) { return h2f(f2h2(x, 0.f) & 0xffffu); }
__device__ __forceinline__ float nanmax(float a, float b) { return (a != a || b != b) ? __int_as_float(0x7fc00000) : fmaxf(a, b); }
__device__ __forceinline__ float nanmin(float a, float b) { return (a != a || b != b) ? __int_as_float(0x7fc00000) : fminf(a, b); }
__device__ __forceinline__ float sgn(float a) { return a > 0.f ? 1.f : (a < 0.f ? -1.f : a); }
// operand loads: 8 consecutive elements -> fp32 (unit stride), or one broadcast element
__device__ __forceinline__ void ld_f32(const Op& t, i64 off, float* r) { const float* p = (const float*)t.ptr + off; V4 a = ldg16(p), b = ldg16(p + 4);
  r[0]=__uint_as_float(a.x); r[1]=__uint_as_float(a.y); r[2]=__uint_as_float(a.z); r[3]=__uint_as_float(a.w); r[4]=__uint_as_float(b.x); r[5]=__uint_as_float(b.y); r[6]=__uint_as_float(b.z); r[7]=__uint_as_float(b.w); }
__device__ __forceinline__ void ld_bf16(const Op& t, i64 off, float* r) { V4 a = ldg16((const u16*)t.ptr + off); u32 w[4] = {a.x, a.y, a.z, a.w};
  #pragma unroll
  for (int i = 0; i < 4; ++i) { r[2*i] = __uint_as_float(w[i] << 16); r[2*i+1] = __uint_as_float(w[i] & 0xffff0000u); } }
__device__ __forceinline__ void ld_f16(const Op& t, i64 off, float* r) { V4 a = ldg16((const u16*)t.ptr + off); u32 w[4] = {a.x, a.y, a.z, a.w};
  #pragma unroll
  for (int i = 0; i < 4; ++i) { r[2*i] = h2f(w[i] & 0xffffu); r[2*i+1] = h2f(w[i] >> 16); } }
__device__ __forceinline__ void ld_bool(const Op& t, i64 off, float* r) { const u8* p = (const u8*)t.ptr + off;
  #pragma unroll
  for (int i = 0; i < 8; ++i) r[i] = p[i] != 0 ? 1.f : 0.f; }
__device__ __forceinline__ float ld1(const Op& t, i64 off) {
  if (t.dtype == 0) return ((const float*)t.ptr)[off];
  if (t.dtype == 1) return bf2f(((const u16*)t.ptr)[off]);
  if (t.dtype == 2) return h2f(((const u16*)t.ptr)[off]);
  return ((const u8*)t.ptr)[off] != 0 ? 1.f : 0.f; }
__device__ __forceinline__ void st_f32(const Op& t, i64 off, const float* r) { float* p = (float*)t.ptr + off;
  V4 a = {__float_as_uint(r[0]), __float_as_uint(r[1]), __float_as_uint(r[2]), __float_as_uint(r[3])}, b = {__float_as_uint(r[4]), __float_as_uint(r[5]), __float_as_uint(r[6]), __float_as_uint(r[7])};
  stg16(p, a); stg16(p + 4, b); }
__device__ __forceinline__ void st_bf16(const Op& t, i64 off, const float* r) { V4 a = {f2bf2(r[0], r[1]), f2bf2(r[2], r[3]), f2bf2(r[4], r[5]), f2bf2(r[6], r[7])}; stg16((u16*)t.ptr + off, a); }
__device__ __forceinline__ void st_f16(const Op& t, i64 off, const float* r) { V4 a = {f2h2(r[0], r[1]), f2h2(r[2], r[3]), f2h2(r[4], r[5]), f2h2(r[6], r[7])}; stg16((u16*)t.ptr + off, a); }
__device__ __forceinline__ void st_bool(const Op& t, i64 off, const float* r) { u8* p = (u8*)t.ptr + off;
  #pragma unroll
  for (int i = 0; i < 8; ++i) p[i] = r[i] != 0.f ? 1 : 0; }
// uniform_philox leaves (b200k_uniform_philox convention): elements li0 .. li0+7 of the row-major tensor are the two
// Philox-4x32-10 blocks with subsequence li0/4 and li0/4 + 1, counter = offset, key = seed; u = x * 2^-32 + 2^-33, 1 -> 0
__device__ __forceinline__ void philox_block(u32 k0, u32 k1, u32 c0, u32 c1, u32 c2, u32 c3, u32* w) {
  #pragma unroll
  for (int r = 0; r < 10; ++r) { const u32 h0 = __umulhi(0xD2511F53u, c0), l0 = 0xD2511F53u * c0, h1 = __umulhi(0xCD9E8D57u, c2), l1 = 0xCD9E8D57u * c2;
    c0 = h1 ^ c1 ^ k0; c1 = l1; c2 = h0 ^ c3 ^ k1; c3 = l0; k0 += 0x9E3779B9u; k1 += 0xBB67AE85u; }
  w[0] = c0; w[1] = c1; w[2] = c2; w[3] = c3; }
__device__ __forceinline__ void gen_unit8(i64 seed, i64 offset, i64 li0, float* r) { u32 w[8]; const unsigned long long s = (unsigned long long)li0 >> 2;
  philox_block((u32)seed, (u32)((unsigned long long)seed >> 32), (u32)offset, (u32)((unsigned long long)offset >> 32), (u32)s, (u32)(s >> 32), w);
  philox_block((u32)seed, (u32)((unsigned long long)seed >> 32), (u32)offset, (u32)((unsigned long long)offset >> 32), (u32)(s + 1), (u32)((s + 1) >> 32), w + 4);
  #pragma unroll
  for (int i = 0; i < 8; ++i) { const float u = __fmaf_rn((float)w[i], 2.3283064365386963e-10f, 1.1641532182693481e-10f); r[i] = u == 1.0f ? 0.0f : u; } }
"""

_DT = {torch.float32: "f32", torch.bfloat16: "bf16", torch.float16: "f16", torch.bool: "bool"}

_EXPR = {
    "add": "__fadd_rn({a}, {b})", "sub": "__fsub_rn({a}, {b})", "mul": "__fmul_rn({a}, {b})", "div": "({a} / {b})", "pow": "powf({a}, {b})",
    "maximum": "nanmax({a}, {b})", "minimum": "nanmin({a}, {b})",
    "gt": "({a} > {b} ? 1.f : 0.f)", "ge": "({a} >= {b} ? 1.f : 0.f)", "lt": "({a} < {b} ? 1.f : 0.f)", "le": "({a} <= {b} ? 1.f : 0.f)",
    "eq": "({a} == {b} ? 1.f : 0.f)", "ne": "({a} != {b} ? 1.f : 0.f)", "and": "(({a} != 0.f && {b} != 0.f) ? 1.f : 0.f)",
    "or": "(({a} != 0.f || {b} != 0.f) ? 1.f : 0.f)", "not": "({a} != 0.f ? 0.f : 1.f)",
    "mov": "{a}", "neg": "(-{a})", "abs": "fabsf({a})", "exp": "expf({a})", "log": "logf({a})", "sqrt": "sqrtf({a})", "rsqrt": "rsqrtf({a})",
    "reciprocal": "(1.0f / {a})", "erf": "erff({a})", "tanh": "tanhf({a})", "sign": "sgn({a})", "floor": "floorf({a})", "ceil": "ceilf({a})",
    "exp2": "exp2f({a})", "log2": "log2f({a})", "log1p": "log1pf({a})", "expm1": "expm1f({a})", "sin": "sinf({a})", "cos": "cosf({a})",
    "erfc": "erfcf({a})", "sigmoid": "(1.0f / (1.0f + expf(-{a})))", "trunc": "truncf({a})", "round": "nearbyintf({a})",
    "rnd_bf16": "rnd_bf16({a})", "rnd_f16": "rnd_f16({a})", "where": "({a} != 0.f ? {b} : {c})",
}


def generate_source(spec, inner_flags: tuple, one_d: bool, unroll: int, in_kernel_rng: bool = True) -> str:
    """CUDA C for one program.  ``spec.ssa`` is the SSA list of the compiler; ``inner_flags[k]`` says whether input k
    is unit-stride (True) or broadcast (False) along the inner dimension."""
    n_in, n_out = len(spec.leaves), len(spec.outputs)
    in_dt = [_DT[dt] for dt in spec.leaf_dtypes]
    out_dt = [_DT[dt] for _, dt, _ in spec.outputs]
    gen = (spec.generated or {}) if in_kernel_rng else {}
    slot_of = {idx: slot for slot, idx in enumerate(sorted(gen))}
    L = [_SKELETON_HEAD % {"MAX_IN": MAX_IN, "MAX_OUT": MAX_OUT, "MAX_SC": 2 * MAX_GEN}]
    L.append('extern "C" __global__ void __launch_bounds__(256) pw_kernel(const __grid_constant__ Params P) {')
    L.append("  const i64 d2v = P.d2 / 8, total = P.d0 * P.d1 * d2v;")
    L.append(f"  const i64 first = ((i64)blockIdx.x * blockDim.x) * {unroll} + threadIdx.x, step = (i64)gridDim.x * blockDim.x * {unroll};")
    L.append("  for (i64 base = first; base < total; base += step) {")
    # coordinates and loads of every item first (all requests in flight), then the math, then the stores
    for u in range(unroll):
        L.append(f"    const i64 idx{u} = base + {u} * (i64)blockDim.x; const bool live{u} = idx{u} < total;")
        if one_d:
            L.append(f"    const i64 i2_{u} = idx{u} * 8, i1_{u} = 0, i0_{u} = 0;")
        else:
            L.append(f"    const i64 i2_{u} = (idx{u} % d2v) * 8, r_{u} = idx{u} / d2v, i1_{u} = r_{u} % P.d1, i0_{u} = r_{u} / P.d1;")
    for u in range(unroll):
        for k in range(n_in):
            L.append(f"    float v{u}_{k}[8];")
            off = f"i0_{u} * P.in[{k}].s0 + i1_{u} * P.in[{k}].s1"
            if k in gen:
                lo, hi = gen[k]
                rnd = {"bf16": "rnd_bf16", "f16": "rnd_f16"}.get(in_dt[k], "")
                L.append(f"    if (live{u}) {{ gen_unit8(P.sc[{2 * slot_of[k]}], P.sc[{2 * slot_of[k] + 1}], (i0_{u} * P.d1 + i1_{u}) * P.d2 + i2_{u}, v{u}_{k});")
                L.append("      #pragma unroll")
                scale = "" if (lo, hi) == (0.0, 1.0) else f" x = __fmaf_rn({hi!r}f - {lo!r}f, x, {lo!r}f);"
                L.append(f"      for (int e = 0; e < 8; ++e) {{ float x = v{u}_{k}[e];{scale} x = {rnd}(x); v{u}_{k}[e] = x == {hi!r}f ? {lo!r}f : x; }} }}")
            elif inner_flags[k]:
                L.append(f"    if (live{u}) ld_{in_dt[k]}(P.in[{k}], {off} + i2_{u}, v{u}_{k});")
            else:
                L.append(f"    if (live{u}) {{ const float b = ld1(P.in[{k}], {off}); for (int e = 0; e < 8; ++e) v{u}_{k}[e] = b; }}")
    for u in range(unroll):
        L.append(f"    if (live{u}) {{")
        for k in range(n_out):
            L.append(f"      float o{k}[8];")
        L.append("      #pragma unroll")
        L.append("      for (int e = 0; e < 8; ++e) {")
        for i, (op, a, b, c, imm) in enumerate(spec.ssa):
            if op == "leaf":
                L.append(f"        const float t{i} = v{u}_{imm}[e];")
            elif op == "const":
                L.append(f"        const float t{i} = __uint_as_float(0x{imm:08x}u);")
            else:
                L.append(f"        const float t{i} = " + _EXPR[op].format(a=f"t{a}", b=f"t{b}", c=f"t{c}") + ";")
        for k, oid in enumerate(spec.out_ids):
            L.append(f"        o{k}[e] = t{oid};")
        L.append("      }")
        for k in range(n_out):
            L.append(f"      st_{out_dt[k]}(P.out[{k}], i0_{u} * P.out[{k}].s0 + i1_{u} * P.out[{k}].s1 + i2_{u}, o{k});")
        L.append("    }")
    L.append("  }")
    L.append("}")
    return "\n".join(L)


# --------------------------------------------------------------------------------------------------
# Compilation and launch
# --------------------------------------------------------------------------------------------------
class _Operand(ctypes.Structure):
    _fields_ = [("ptr", ctypes.c_void_p), ("s0", ctypes.c_int64), ("s1", ctypes.c_int64), ("inner", ctypes.c_int32), ("dtype", ctypes.c_int32)]


class Params(ctypes.Structure):
    _fields_ = [("d0", ctypes.c_int64), ("d1", ctypes.c_int64), ("d2", ctypes.c_int64), ("sc", ctypes.c_int64 * (2 * MAX_GEN)),
                ("inp", _Operand * MAX_IN), ("out", _Operand * MAX_OUT)]


def _check(res, what):
    err = res[0] if isinstance(res, tuple) else res
    if int(err) != 0:
        raise JitUnavailable(f"{what} failed: {err}")
    return res[1:] if isinstance(res, tuple) else ()


def compile_cubin(source: str) -> bytes:
    """NVRTC: CUDA C -> sm_100a cubin (works without a GPU, which is how the CPU tests check the generator)."""
    from cuda.bindings import nvrtc

    (prog,) = _check(nvrtc.nvrtcCreateProgram(source.encode(), b"b200_pointwise.cu", 0, [], []), "nvrtcCreateProgram")
    opts = [b"--gpu-architecture=sm_100a", b"-lineinfo", b"--std=c++17"]
    res = nvrtc.nvrtcCompileProgram(prog, len(opts), opts)
    if int(res[0]) != 0:
        (n,) = _check(nvrtc.nvrtcGetProgramLogSize(prog), "nvrtcGetProgramLogSize")
        log = b" " * n
        nvrtc.nvrtcGetProgramLog(prog, log)
        nvrtc.nvrtcDestroyProgram(prog)
        raise JitUnavailable("NVRTC compilation failed:\n" + log.decode(errors="replace"))
    (n,) = _check(nvrtc.nvrtcGetCUBINSize(prog), "nvrtcGetCUBINSize")
    cubin = b" " * n
    _check(nvrtc.nvrtcGetCUBIN(prog, cubin), "nvrtcGetCUBIN")
    nvrtc.nvrtcDestroyProgram(prog)
    return cubin


_culaunch = {"fn": None, "tried": False}


def fast_launch(driver, func, grid: int, block: int, stream: int, params_addr: int, smem: int = 0) -> None:
    """cuLaunchKernel(func, grid, 1, 1, block, 1, 1, smem, stream, {&params}, 0).  Goes straight to libcuda through
    ctypes — a couple of microseconds — and only falls back to the cuda-python wrapper (ten times that per call, which
    is as long as some of the kernels) when libcuda cannot be opened that way."""
    if not _culaunch["tried"]:
        _culaunch["tried"] = True
        try:
            lib = ctypes.CDLL("libcuda.so.1")
            fn = lib.cuLaunchKernel
            fn.restype = ctypes.c_int
            fn.argtypes = [ctypes.c_void_p] + [ctypes.c_uint] * 7 + [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
            _culaunch["fn"] = fn
        except Exception:
            _culaunch["fn"] = None
    argv = (ctypes.c_void_p * 1)(params_addr)
    fn = _culaunch["fn"]
    if fn is not None:
        err = fn(int(func), grid, 1, 1, block, 1, 1, smem, stream, ctypes.addressof(argv), None)
    else:
        (err,) = driver.cuLaunchKernel(func, grid, 1, 1, block, 1, 1, smem, stream, ctypes.addressof(argv), 0)
    if int(err) != 0:
        raise RuntimeError(f"b200ex jit: cuLaunchKernel failed: {err}")


class JitKernel:
    """One specialised kernel, loaded on one device."""

    def __init__(self, cubin: bytes, unroll: int):
        from cuda.bindings import driver

        self._driver = driver
        (self.module,) = _check(driver.cuModuleLoadData(cubin), "cuModuleLoadData")
        (self.func,) = _check(driver.cuModuleGetFunction(self.module, b"pw_kernel"), "cuModuleGetFunction")
        self.unroll = unroll
        self.launches = 0

    def launch(self, params: Params, total_items: int, sm_count: int, stream: int) -> None:
        per_cta = 256 * self.unroll
        want = (total_items + per_cta - 1) // per_cta
        grid = max(1, min(want, sm_count * 8))
        fast_launch(self._driver, self.func, grid, 256, stream, ctypes.addressof(params))
        self.launches += 1


def unroll_for(n_in: int) -> int:
    """Work items per thread per trip: enough 16-byte requests in flight (4-8 per thread) without spilling."""
    return 4 if n_in <= 2 else (2 if n_in <= 4 else 1)


def get_kernel(spec, inner_flags: tuple, one_d: bool, device_index: int, in_kernel_rng: bool = True) -> JitKernel:
    in_kernel_rng = bool(in_kernel_rng and spec.generated)
    key = (spec.signature, inner_flags, one_d, device_index, in_kernel_rng)
    k = _kernels.get(key)
    if k is not None:
        return k
    with _lock:
        k = _kernels.get(key)
        if k is None:
            unroll = unroll_for(len(spec.leaves) - (len(spec.generated) if in_kernel_rng else 0))
            src = generate_source(spec, inner_flags, one_d, unroll, in_kernel_rng)
            with torch.cuda.device(device_index):
                k = JitKernel(compile_cubin(src), unroll)
            _kernels[key] = k
    return k


def cache_info() -> dict:
    return {"kernels": len(_kernels), "launches": sum(k.launches for k in _kernels.values())}
