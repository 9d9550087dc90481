"""B200-native GGUF inference path: ctypes binding of the C ABI in include/b2.h (libb2.so).

This package is the Python face of the drop-in for the one hot path graz-dev/llms-on-kubernetes
deploys -- GGUF quantised-model inference behind the ramalama chart's
``llama-server --host --port --model --alias`` (ramalama-models/helm-chart/templates/
model-deployments.yaml:26-35).  The product is the C++/CUDA code in ``csrc/``; this module only binds
it.  There is no CPU fallback: if libb2.so is missing the import fails, and compute calls on a machine
without a CUDA device raise ``B2Error``.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb2.so")
SERVER_PATH = os.environ.get("B2_SERVER_BIN") or os.path.join(_HERE, "bin", "llama-server")   # B2_SERVER_BIN: an instrumented build (bin/llama-server-tsan)

F32, F16, Q8_0, Q4_K, Q5_K, Q6_K, BF16 = 0, 1, 8, 12, 13, 14, 30


class B2Error(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"b2 error {code}: {msg}")
        self.code = code


class TensorInfo(C.Structure):
    _fields_ = [("name", C.c_char_p), ("n_dims", C.c_uint32), ("dims", C.c_uint64 * 4), ("type", C.c_uint32),
                ("offset", C.c_uint64), ("n_bytes", C.c_uint64)]


class Sampling(C.Structure):
    """b2_sampling: llama-server's sampling parameters (temperature <= 0 or top_k == 1: greedy)."""
    _fields_ = [("temperature", C.c_float), ("top_p", C.c_float), ("min_p", C.c_float), ("top_k", C.c_int32), ("seed", C.c_uint64)]


class ModelInfo(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("arch", "n_layer", "d_model", "d_ff", "n_head", "n_kv", "head_dim", "vocab",
                                          "n_ctx_train", "rope_dim")] + \
               [("eps", C.c_float), ("rope_base", C.c_float), ("bos_id", C.c_int32), ("eos_id", C.c_int32),
                ("eot_id", C.c_int32), ("tied", C.c_int32), ("weight_bytes", C.c_uint64), ("stream_bytes", C.c_uint64)]


# every symbol include/b2.h declares: name -> (restype, argtypes)
_P, _I, _I64, _U64, _SZ = C.c_void_p, C.c_int, C.c_int64, C.c_uint64, C.c_size_t
SYMBOLS = {
    "b2_last_error": (C.c_char_p, []),
    "b2_abi_version": (_I, []),
    "b2_device_count": (_I, []),
    "b2_gguf_open": (_I, [C.c_char_p, C.POINTER(_P)]),
    "b2_gguf_close": (None, [_P]),
    "b2_gguf_version": (C.c_uint32, [_P]),
    "b2_gguf_alignment": (C.c_uint32, [_P]),
    "b2_gguf_data_offset": (_U64, [_P]),
    "b2_gguf_n_kv": (_I64, [_P]),
    "b2_gguf_n_tensors": (_I64, [_P]),
    "b2_gguf_kv_key": (C.c_char_p, [_P, _I64]),
    "b2_gguf_kv_type": (_I, [_P, _I64]),
    "b2_gguf_kv_array_type": (_I, [_P, _I64]),
    "b2_gguf_kv_array_len": (_I64, [_P, _I64]),
    "b2_gguf_find_kv": (_I64, [_P, C.c_char_p]),
    "b2_gguf_kv_u64": (_I, [_P, _I64, C.POINTER(C.c_uint64)]),
    "b2_gguf_kv_i64": (_I, [_P, _I64, C.POINTER(C.c_int64)]),
    "b2_gguf_kv_f64": (_I, [_P, _I64, C.POINTER(C.c_double)]),
    "b2_gguf_kv_str": (_I, [_P, _I64, C.POINTER(_P), C.POINTER(C.c_uint64)]),
    "b2_gguf_kv_arr_str": (_I, [_P, _I64, _I64, C.POINTER(_P), C.POINTER(C.c_uint64)]),
    "b2_gguf_kv_arr_f64": (_I, [_P, _I64, _I64, C.POINTER(C.c_double)]),
    "b2_gguf_kv_arr_i64": (_I, [_P, _I64, _I64, C.POINTER(C.c_int64)]),
    "b2_gguf_tensor_info": (_I, [_P, _I64, C.POINTER(TensorInfo)]),
    "b2_gguf_find_tensor": (_I64, [_P, C.c_char_p]),
    "b2_gguf_tensor_data": (_P, [_P, _I64]),
    "b2_dev_malloc": (_I, [C.POINTER(_P), _SZ]),
    "b2_dev_free": (_I, [_P]),
    "b2_dev_memset": (_I, [_P, _I, _SZ, _P]),
    "b2_memcpy_h2d": (_I, [_P, _P, _SZ, _P]),
    "b2_memcpy_d2h": (_I, [_P, _P, _SZ, _P]),
    "b2_stream_sync": (_I, [_P]),
    "b2_row_bytes": (_I64, [_I, _I64]),
    "b2_dequantize": (_I, [_I, _P, _P, _I64, _P]),
    "b2_dequantize_host": (_I, [_I, _P, _P, _I64]),
    "b2_gemv": (_I, [_I, _P, _P, _P, _I64, _I64, _I, _P]),
    "b2_gemm_scratch_bytes": (_SZ, [_I64, _I]),
    "b2_gemm": (_I, [_I, _P, _P, _P, _I64, _I64, _I, _P, _P]),
    "b2_tc_gemm_scratch_bytes": (_SZ, [_I64, _I64, _I]),
    "b2_tc_gemm": (_I, [_I, _P, _P, _P, _I64, _I64, _I, _I, _P, _P]),
    "b2_tcq_gemm_scratch_bytes": (_SZ, [_I64, _I64, _I]),
    "b2_tcq_plan": (_I, [_P, _I, _I64, _I, _P, _P, _P, _P, _P]),
    "b2_tcq_gemm": (_I, [_I, _P, _P, _P, _I64, _I64, _I, _I, _P, _P]),
    "b2_rmsnorm": (_I, [_P, _P, _P, _I64, _I64, C.c_float, _P]),
    "b2_rope": (_I, [_P, _I, _I, _I, _I, _I, C.c_float, _P, _P]),
    "b2_attn_scratch_bytes": (_SZ, [_I, _I, _I]),
    "b2_attn_decode": (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _I, _P, _P]),
    "b2_attn_decode_static": (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
    "b2_argmax_scratch_bytes": (_SZ, [_I64]),
    "b2_sample_greedy": (_I, [_P, _I64, _P, _P, _P]),
    "b2_sample": (_I, [_P, _I, _I64, _I64, _P, _I64, _P, _P, _P]),
    "b2_ctx_set_sampler": (_I, [_P, _I, _P]),
    "b2_model_load": (_I, [C.c_char_p, _I, _I, _I, C.POINTER(_P)]),
    "b2_allreduce_unique_id": (_I, [_P, _SZ]),
    "b2_allreduce_init": (_I, [_P, _P, _SZ]),
    "b2_model_free": (None, [_P]),
    "b2_model_get_info": (_I, [_P, C.POINTER(ModelInfo)]),
    "b2_model_gguf": (_P, [_P]),
    "b2_model_tensor": (_I, [_P, C.c_char_p, C.POINTER(_P), C.POINTER(TensorInfo)]),
    "b2_ctx_create": (_I, [_P, _I, C.POINTER(_P)]),
    "b2_ctx_free": (None, [_P]),
    "b2_ctx_n_ctx": (_I, [_P]),
    "b2_ctx_reset": (_I, [_P]),
    "b2_ctx_decode": (_I, [_P, _P, _I, _I, _P, C.POINTER(C.c_int32)]),
    "b2_ctx_create_slots": (_I, [_P, _I, _I, C.POINTER(_P)]),
    "b2_ctx_n_slots": (_I, [_P]),
    "b2_ctx_decode_batch": (_I, [_P, _P, _P, _P, _I, _I, _P, _P]),
    "b2_ctx_batch_logits": (_I, [_P, _I, _P]),
    "b2_ctx_prefill": (_I, [_P, _I, _P, _I, _I, _P, _P]),
    "b2_ctx_max_batch": (_I, [_P]),
    "b2_ctx_prefill_multi": (_I, [_P, _I, _P, _P, _P, _P, _P, _P]),
    "b2_ctx_slot_truncate": (_I, [_P, _I, _I]),
    "b2_ctx_kv_pages": (_I, [_P, C.POINTER(_I), C.POINTER(_I), C.POINTER(_I)]),
    "b2_ctx_generate_greedy": (_I, [_P, _I, _P]),
    "b2_ctx_enqueue_steps": (_I, [_P, _I]),
    "b2_ctx_stream": (_P, [_P]),
    "b2_ctx_sync": (_I, [_P]),
    "b2_ctx_get_activation": (_I, [_P, _I, _P, _I64]),
    "b2_ctx_get_kv": (_I, [_P, _I, _I, _P, _P]),
    "b2_ctx_pos": (_I, [_P]),
    "b2_ctx_launches_per_step": (_I, [_P]),
    "b2_ctx_time_steps": (_I, [_P, _I, C.POINTER(C.c_float)]),
    "b2_ctx_time_matvec_only": (_I, [_P, _I, C.POINTER(C.c_float), C.POINTER(C.c_uint64), C.POINTER(_I)]),
}


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `make -C {_HERE}` (or __graft_entry__.build()); "
                          "there is no fallback path")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError here = header and library out of sync
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()


def check(rc: int) -> int:
    if rc < 0:
        raise B2Error(rc, (lib.b2_last_error() or b"").decode(errors="replace"))
    return rc


def sample(logits: np.ndarray, n_draws: int, temperature=0.8, top_k=40, top_p=0.95, min_p=0.05, seed=0, first_draw=0) -> np.ndarray:
    """n_draws tokens drawn on the device from ONE logits row (b2_sample with row_stride 0): draw indices first_draw .. ."""
    lg = DevBuf.from_numpy(np.ascontiguousarray(logits, np.float32))
    par = np.frombuffer(bytes(Sampling(temperature, top_p, min_p, top_k, seed)), np.uint8)
    dpar = DevBuf.from_numpy(par)
    ddraw = DevBuf.from_numpy(np.arange(first_draw, first_draw + n_draws, dtype=np.uint64))
    out = DevBuf(4 * n_draws, zero=True)
    check(lib.b2_sample(lg.ptr, n_draws, logits.size, 0, dpar.ptr, 0, ddraw.ptr, out.ptr, None))
    check(lib.b2_stream_sync(None))
    return out.to_numpy(np.int32)


def device_count() -> int:
    return check(lib.b2_device_count())


class DevBuf:
    """Device allocation through the C ABI (tests use this instead of a CUDA binding of their own)."""

    def __init__(self, nbytes: int, zero: bool = False):
        self.ptr = _P()
        self.nbytes = nbytes
        check(lib.b2_dev_malloc(C.byref(self.ptr), max(nbytes, 1)))
        if zero:
            check(lib.b2_dev_memset(self.ptr, 0, nbytes, None))
            check(lib.b2_stream_sync(None))

    @classmethod
    def from_numpy(cls, a: np.ndarray) -> "DevBuf":
        a = np.ascontiguousarray(a)
        b = cls(a.nbytes)
        if a.nbytes:
            check(lib.b2_memcpy_h2d(b.ptr, a.ctypes.data, a.nbytes, None))
            check(lib.b2_stream_sync(None))
        return b

    def to_numpy(self, dtype, count: int | None = None) -> np.ndarray:
        dt = np.dtype(dtype)
        n = self.nbytes // dt.itemsize if count is None else count
        out = np.empty(n, dt)
        if n:
            check(lib.b2_memcpy_d2h(out.ctypes.data, self.ptr, n * dt.itemsize, None))
            check(lib.b2_stream_sync(None))
        return out

    def offset(self, nbytes: int) -> C.c_void_p:
        return _P(self.ptr.value + nbytes)

    def free(self):
        if self.ptr:
            lib.b2_dev_free(self.ptr)
            self.ptr = _P()

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def dequantize(raw: np.ndarray, qtype: int) -> np.ndarray:
    """raw uint8 blocks (host) -> float32, through b2_dequantize_host (H2D, sm_100a kernel, D2H)."""
    raw = np.ascontiguousarray(raw, np.uint8)
    rb1 = lib.b2_row_bytes(qtype, {F32: 1, F16: 1, BF16: 1, Q8_0: 32}.get(qtype, 256))
    blk = {F32: 1, F16: 1, BF16: 1, Q8_0: 32}.get(qtype, 256)
    if rb1 <= 0 or raw.size % rb1:
        raise B2Error(-1, f"{raw.size} bytes is not a whole number of type-{qtype} blocks")
    n = raw.size // rb1 * blk
    out = np.empty(n, np.float32)
    check(lib.b2_dequantize_host(qtype, raw.ctypes.data, out.ctypes.data, n))
    return out


def gemv(qtype: int, W: np.ndarray, x: np.ndarray, N: int, K: int) -> np.ndarray:
    """y = W.x with host buffers in and out (x may be [B, K])."""
    x = np.ascontiguousarray(x, np.float32)
    B = 1 if x.ndim == 1 else x.shape[0]
    dW, dx = DevBuf.from_numpy(np.ascontiguousarray(W, np.uint8)), DevBuf.from_numpy(x)
    dy = DevBuf(4 * N * B)
    check(lib.b2_gemv(qtype, dW.ptr, dx.ptr, dy.ptr, N, K, B, None))
    check(lib.b2_stream_sync(None))
    y = dy.to_numpy(np.float32)
    return y if x.ndim == 1 else y.reshape(B, N)


def gemm(qtype: int, W: np.ndarray, x: np.ndarray, N: int, K: int) -> np.ndarray:
    """Y[B, N] = X[B, K] . W^T in one pass over the weights (tensor-core kernel, 1 <= B <= 32), host buffers."""
    x = np.ascontiguousarray(x, np.float32)
    assert x.ndim == 2
    B = x.shape[0]
    dW, dx = DevBuf.from_numpy(np.ascontiguousarray(W, np.uint8)), DevBuf.from_numpy(x)
    dy = DevBuf(4 * N * B)
    scratch = DevBuf(max(16, lib.b2_gemm_scratch_bytes(K, B)))
    check(lib.b2_gemm(qtype, dW.ptr, dx.ptr, dy.ptr, N, K, B, scratch.ptr, None))
    check(lib.b2_stream_sync(None))
    return dy.to_numpy(np.float32).reshape(B, N)


def tc_gemm(qtype: int, W: np.ndarray, x: np.ndarray, N: int, K: int, y0: np.ndarray | None = None) -> np.ndarray:
    """Y[T, N] (= y0 +) X[T, K] . W^T on the tcgen05 tensor cores (prompt-phase kernel), host buffers."""
    x = np.ascontiguousarray(x, np.float32)
    assert x.ndim == 2
    T = x.shape[0]
    dW, dx = DevBuf.from_numpy(np.ascontiguousarray(W, np.uint8)), DevBuf.from_numpy(x)
    dy = DevBuf(4 * N * T) if y0 is None else DevBuf.from_numpy(np.ascontiguousarray(y0, np.float32))
    scratch = DevBuf(lib.b2_tc_gemm_scratch_bytes(N, K, T))
    check(lib.b2_tc_gemm(qtype, dW.ptr, dx.ptr, dy.ptr, N, K, T, 0 if y0 is None else 1, scratch.ptr, None))
    check(lib.b2_stream_sync(None))
    return dy.to_numpy(np.float32).reshape(T, N)


def tcq_plan(seg_rows, K: int, n_sms: int = 148) -> dict:
    """The stream-K decomposition b2_tcq_gemm uses for a stack of tensors (host arithmetic; works without a GPU)."""
    rows = np.ascontiguousarray(seg_rows, np.int32)
    tiles, grid, slices = C.c_int32(), C.c_int32(), C.c_int32()
    check(lib.b2_tcq_plan(rows.ctypes.data, rows.size, K, n_sms, C.byref(tiles), C.byref(grid), C.byref(slices), None, None))
    shares = np.zeros(2 * grid.value, np.int64)
    first = np.zeros(tiles.value, np.int32)
    check(lib.b2_tcq_plan(rows.ctypes.data, rows.size, K, n_sms, None, None, None, shares.ctypes.data, first.ctypes.data))
    return {"tiles": tiles.value, "grid": grid.value, "slices": slices.value, "shares": shares.reshape(-1, 2), "first_cta": first}


def tcq_gemm(qtype: int, W: np.ndarray, x: np.ndarray, N: int, K: int, y0: np.ndarray | None = None) -> np.ndarray:
    """Y[T, N] (= y0 +) X[T, K] . W^T for a decode batch: quantised blocks dequantised inside the tcgen05 kernel's producer."""
    x = np.ascontiguousarray(x, np.float32)
    assert x.ndim == 2
    T = x.shape[0]
    dW, dx = DevBuf.from_numpy(np.ascontiguousarray(W, np.uint8)), DevBuf.from_numpy(x)
    dy = DevBuf(4 * N * T) if y0 is None else DevBuf.from_numpy(np.ascontiguousarray(y0, np.float32))
    scratch = DevBuf(lib.b2_tcq_gemm_scratch_bytes(N, K, T))
    check(lib.b2_tcq_gemm(qtype, dW.ptr, dx.ptr, dy.ptr, N, K, T, 0 if y0 is None else 1, scratch.ptr, None))
    check(lib.b2_stream_sync(None))
    return dy.to_numpy(np.float32).reshape(T, N)


class Gguf:
    """b2_gguf_*: the C++ GGUF parser (csrc/gguf.cpp)."""

    def __init__(self, path: str):
        self.h = _P()
        check(lib.b2_gguf_open(os.fsencode(path), C.byref(self.h)))
        self._own = True

    @classmethod
    def _view(cls, handle) -> "Gguf":
        g = cls.__new__(cls)
        g.h, g._own = _P(handle), False
        return g

    def close(self):
        if self.h and self._own:
            lib.b2_gguf_close(self.h)
        self.h = _P()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    version = property(lambda s: lib.b2_gguf_version(s.h))
    alignment = property(lambda s: lib.b2_gguf_alignment(s.h))
    data_offset = property(lambda s: lib.b2_gguf_data_offset(s.h))
    n_kv = property(lambda s: lib.b2_gguf_n_kv(s.h))
    n_tensors = property(lambda s: lib.b2_gguf_n_tensors(s.h))

    def keys(self) -> list[str]:
        return [lib.b2_gguf_kv_key(self.h, i).decode() for i in range(self.n_kv)]

    def _scalar(self, i: int, t: int):
        if t == 8:
            p, n = _P(), C.c_uint64()
            check(lib.b2_gguf_kv_str(self.h, i, C.byref(p), C.byref(n)))
            return C.string_at(p, n.value)
        if t in (6, 12):
            v = C.c_double()
            check(lib.b2_gguf_kv_f64(self.h, i, C.byref(v)))
            return v.value
        if t == 10:
            u = C.c_uint64()
            check(lib.b2_gguf_kv_u64(self.h, i, C.byref(u)))
            return u.value
        v = C.c_int64()
        check(lib.b2_gguf_kv_i64(self.h, i, C.byref(v)))
        return bool(v.value) if t == 7 else v.value

    def get(self, key: str, default=None):
        i = lib.b2_gguf_find_kv(self.h, key.encode())
        if i < 0:
            return default
        t = lib.b2_gguf_kv_type(self.h, i)
        if t != 9:
            return self._scalar(i, t)
        et, n = lib.b2_gguf_kv_array_type(self.h, i), lib.b2_gguf_kv_array_len(self.h, i)
        out = []
        for j in range(n):
            if et == 8:
                p, ln = _P(), C.c_uint64()
                check(lib.b2_gguf_kv_arr_str(self.h, i, j, C.byref(p), C.byref(ln)))
                out.append(C.string_at(p, ln.value))
            elif et in (6, 12):
                v = C.c_double()
                check(lib.b2_gguf_kv_arr_f64(self.h, i, j, C.byref(v)))
                out.append(v.value)
            else:
                v = C.c_int64()
                check(lib.b2_gguf_kv_arr_i64(self.h, i, j, C.byref(v)))
                out.append(v.value)
        return out

    def tensor_info(self, i: int) -> dict:
        ti = TensorInfo()
        check(lib.b2_gguf_tensor_info(self.h, i, C.byref(ti)))
        return {"name": ti.name.decode(), "dims": [ti.dims[d] for d in range(ti.n_dims)], "type": ti.type,
                "offset": ti.offset, "n_bytes": ti.n_bytes}

    def tensor_bytes(self, i: int) -> bytes:
        info = self.tensor_info(i)
        return C.string_at(lib.b2_gguf_tensor_data(self.h, i), info["n_bytes"])


def allreduce_unique_id() -> bytes:
    """NCCL id for a tensor-parallel group: create on rank 0, hand the same bytes to every rank's Context.allreduce_init."""
    buf = C.create_string_buffer(128)
    n = check(lib.b2_allreduce_unique_id(buf, 128))
    return buf.raw[:n]


class Model:
    def __init__(self, path: str, device: int = 0, tp_rank: int = 0, tp_size: int = 1):
        self.h = _P()
        check(lib.b2_model_load(os.fsencode(path), device, tp_rank, tp_size, C.byref(self.h)))
        self.info = ModelInfo()
        check(lib.b2_model_get_info(self.h, C.byref(self.info)))

    def gguf(self) -> Gguf:
        return Gguf._view(lib.b2_model_gguf(self.h))

    def tensor_device_bytes(self, name: str) -> bytes:
        p, ti = _P(), TensorInfo()
        check(lib.b2_model_tensor(self.h, name.encode(), C.byref(p), C.byref(ti)))
        out = np.empty(ti.n_bytes, np.uint8)
        check(lib.b2_memcpy_d2h(out.ctypes.data, p, ti.n_bytes, None))
        check(lib.b2_stream_sync(None))
        return out.tobytes()

    def free(self):
        if self.h:
            lib.b2_model_free(self.h)
            self.h = _P()

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Context:
    def __init__(self, model: Model, n_ctx: int = 0, n_slots: int = 1):
        self.model = model
        self.h = _P()
        check(lib.b2_ctx_create_slots(model.h, n_ctx, n_slots, C.byref(self.h)))
        self.n_ctx = lib.b2_ctx_n_ctx(self.h)
        self.n_slots = lib.b2_ctx_n_slots(self.h)

    def decode_batch(self, tokens, pos, slots=None, last_only: bool = False, want_logits: bool = True):
        """One pass over the weights for up to 32 (token, slot, position) columns -> (greedy ids, logits [n, vocab])."""
        toks = np.ascontiguousarray(tokens, np.int32)
        p = np.ascontiguousarray(pos, np.int32)
        s = None if slots is None else np.ascontiguousarray(slots, np.int32)
        n_out = 1 if last_only else toks.size
        nxt = np.empty(n_out, np.int32)
        logits = np.empty((n_out, self.model.info.vocab), np.float32) if want_logits else None
        check(lib.b2_ctx_decode_batch(self.h, toks.ctypes.data, None if s is None else s.ctypes.data, p.ctypes.data, toks.size,
                                      1 if last_only else 0, nxt.ctypes.data, logits.ctypes.data if want_logits else None))
        return nxt.tolist(), logits

    def allreduce_init(self, unique_id: bytes):
        """Collective over the tensor-parallel ranks: joins the NCCL communicator and finishes the context's set-up."""
        check(lib.b2_allreduce_init(self.h, unique_id, len(unique_id)))

    def prefill_multi(self, prompts, slots, pos0=None, want_logits: bool = True):
        """Several sequences' prompts in one pass -> (greedy ids [n_seq], logits [n_seq, vocab] of each last position)."""
        n = len(prompts)
        sl = np.ascontiguousarray(slots, np.int32)
        p0 = np.zeros(n, np.int32) if pos0 is None else np.ascontiguousarray(pos0, np.int32)
        ln = np.ascontiguousarray([len(p) for p in prompts], np.int32)
        toks = np.ascontiguousarray(np.concatenate([np.asarray(p, np.int32) for p in prompts]), np.int32)
        nxt = np.empty(n, np.int32)
        logits = np.empty((n, self.model.info.vocab), np.float32) if want_logits else None
        check(lib.b2_ctx_prefill_multi(self.h, n, sl.ctypes.data, p0.ctypes.data, ln.ctypes.data, toks.ctypes.data, nxt.ctypes.data,
                                       logits.ctypes.data if want_logits else None))
        return nxt.tolist(), logits

    def slot_truncate(self, slot: int, n_keep: int = 0):
        check(lib.b2_ctx_slot_truncate(self.h, slot, n_keep))

    def kv_pages(self):
        """(total pages, free pages, positions per page, mapped on demand?)"""
        t, f, p = _I(), _I(), _I()
        dyn = check(lib.b2_ctx_kv_pages(self.h, C.byref(t), C.byref(f), C.byref(p)))
        return t.value, f.value, p.value, bool(dyn)

    def max_batch(self) -> int:
        return check(lib.b2_ctx_max_batch(self.h))

    def prefill(self, tokens, pos0: int = 0, slot: int = 0, want_logits: bool = True):
        """A whole prompt of one sequence slot on the tensor-core prompt path -> (greedy id, logits of the last position)."""
        toks = np.ascontiguousarray(tokens, np.int32)
        logits = np.empty(self.model.info.vocab, np.float32) if want_logits else None
        nxt = C.c_int32()
        check(lib.b2_ctx_prefill(self.h, slot, toks.ctypes.data, toks.size, pos0, C.byref(nxt),
                                 logits.ctypes.data if want_logits else None))
        return nxt.value, logits

    def decode(self, tokens, pos0: int, want_logits: bool = True):
        toks = np.ascontiguousarray(tokens, np.int32)
        logits = np.empty(self.model.info.vocab, np.float32) if want_logits else None
        nxt = C.c_int32()
        check(lib.b2_ctx_decode(self.h, toks.ctypes.data, toks.size, pos0,
                                logits.ctypes.data if want_logits else None, C.byref(nxt)))
        return nxt.value, logits

    def generate_greedy(self, n: int) -> list[int]:
        out = np.empty(n, np.int32)
        check(lib.b2_ctx_generate_greedy(self.h, n, out.ctypes.data))
        return out.tolist()

    def set_sampler(self, slot: int = 0, temperature: float = 0.8, top_k: int = 40, top_p: float = 0.95, min_p: float = 0.05, seed: int = 0,
                    greedy: bool = False):
        """Device-side sampling for a sequence slot (llama-server's defaults); greedy=True switches it off."""
        if greedy:
            check(lib.b2_ctx_set_sampler(self.h, slot, None))
        else:
            s = Sampling(temperature, top_p, min_p, top_k, seed)
            check(lib.b2_ctx_set_sampler(self.h, slot, C.byref(s)))

    def greedy(self, prompt, n_new: int):
        """Same contract as oracle.ref_model.RefGGUFModel.greedy."""
        first, logits = self.decode(prompt, 0)
        rest = self.generate_greedy(n_new - 1) if n_new > 1 else []
        return [first] + rest, logits

    ACT = {"hidden": 0, "q": 1, "attn_out": 2, "ffn_h": 3, "logits": 4}

    def activation(self, which: str) -> np.ndarray:
        i = self.model.info
        cap = max(i.vocab, i.d_ff, i.n_head * i.head_dim, i.d_model)
        out = np.empty(cap, np.float32)
        n = check(lib.b2_ctx_get_activation(self.h, self.ACT[which], out.ctypes.data, cap))
        return out[:n].copy()

    def hidden(self) -> np.ndarray:
        return self.activation("hidden")

    def kv(self, layer: int, pos: int) -> tuple[np.ndarray, np.ndarray]:
        i = self.model.info
        k = np.empty(i.n_kv * i.head_dim, np.float16)
        v = np.empty_like(k)
        check(lib.b2_ctx_get_kv(self.h, layer, pos, k.ctypes.data, v.ctypes.data))
        return k, v

    def enqueue_steps(self, n: int):
        check(lib.b2_ctx_enqueue_steps(self.h, n))

    def sync(self):
        check(lib.b2_ctx_sync(self.h))

    def time_steps(self, n: int) -> float:
        ms = C.c_float()
        check(lib.b2_ctx_time_steps(self.h, n, C.byref(ms)))
        return ms.value

    def time_matvec_only(self, reps: int) -> tuple[float, int, int]:
        ms, nbytes, launches = C.c_float(), C.c_uint64(), C.c_int()
        check(lib.b2_ctx_time_matvec_only(self.h, reps, C.byref(ms), C.byref(nbytes), C.byref(launches)))
        return ms.value, nbytes.value, launches.value

    @property
    def stream(self) -> int:
        return lib.b2_ctx_stream(self.h) or 0

    @property
    def launches_per_step(self) -> int:
        return lib.b2_ctx_launches_per_step(self.h)

    def free(self):
        if self.h:
            lib.b2_ctx_free(self.h)
            self.h = _P()

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass
