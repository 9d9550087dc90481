"""oracle/ref_model.py -- TEST INFRASTRUCTURE, NOT PRODUCT.

ctypes driver for oracle/libggml_ref.so (the C restatement of ggml's CPU path).  Parses a
GGUF with llama.cpp's own ``gguf.GGUFReader`` (gguf_reader.py:132-185) -- deliberately NOT
with the product's C++ parser -- and hands raw tensor pointers to the C code.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "libggml_ref.so")
    src = os.path.join(_HERE, "ggml_ref.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return so


class RefTensor(C.Structure):
    _fields_ = [("data", C.c_void_p), ("type", C.c_int32), ("_pad", C.c_int32), ("rows", C.c_int64), ("cols", C.c_int64)]


class RefLayer(C.Structure):
    _fields_ = [(n, RefTensor) for n in ("attn_norm", "wq", "wk", "wv", "wo", "ffn_norm", "w_gate", "w_up", "w_down")] + \
               [("bq", C.c_void_p), ("bk", C.c_void_p), ("bv", C.c_void_p)]


class RefModel(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("arch", "n_layer", "d_model", "d_ff", "n_head", "n_kv", "head_dim", "vocab", "n_ctx", "rope_dim")] + \
               [("eps", C.c_float), ("rope_base", C.c_float),
                ("tok_embd", RefTensor), ("output_norm", RefTensor), ("output", RefTensor),
                ("rope_freqs", C.c_void_p), ("layers", C.POINTER(RefLayer))]


def lib() -> C.CDLL:
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        L = _LIB
        L.ref_dequantize.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64]
        L.ref_dequantize.restype = C.c_int
        L.ref_quantize_q8_0.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        L.ref_quantize_q8_K.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        L.ref_gemv.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64]
        L.ref_gemv.restype = C.c_int
        L.ref_rmsnorm.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_float]
        L.ref_rope_cache.argtypes = [C.c_int32, C.c_int, C.c_float, C.c_void_p, C.c_void_p]
        L.ref_rope.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.ref_ctx_new.argtypes = [C.POINTER(RefModel)]
        L.ref_ctx_new.restype = C.c_void_p
        L.ref_ctx_free.argtypes = [C.c_void_p]
        L.ref_decode.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]
        L.ref_decode.restype = C.c_int32
        L.ref_ctx_hidden.argtypes = [C.c_void_p]
        L.ref_ctx_hidden.restype = C.POINTER(C.c_float)
        L.ref_ctx_buffer.argtypes = [C.c_void_p, C.c_int]
        L.ref_ctx_buffer.restype = C.POINTER(C.c_float)
        for f in (L.ref_ctx_k, L.ref_ctx_v):
            f.argtypes = [C.c_void_p, C.c_int, C.c_int]
            f.restype = C.POINTER(C.c_uint16)
        L.ref_row_bytes.argtypes = [C.c_int, C.c_int64]
        L.ref_row_bytes.restype = C.c_int64
        L.ref_num_threads.restype = C.c_int
        L.ref_set_num_threads.argtypes = [C.c_int]
        L.ref_set_perturb.argtypes = [C.c_float, C.c_uint32]
    return _LIB


def _p(a: np.ndarray) -> int:
    return a.ctypes.data


def dequantize(raw: np.ndarray, qtype: int, n: int) -> np.ndarray:
    raw = np.ascontiguousarray(raw)
    out = np.empty(n, np.float32)
    assert lib().ref_dequantize(qtype, _p(raw), _p(out), n) == 0
    return out


def quantize_q8_K(x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, np.float32)
    out = np.zeros(x.size // 256 * 292, np.uint8)
    lib().ref_quantize_q8_K(_p(x), _p(out), x.size)
    return out


def quantize_q8_0(x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, np.float32)
    out = np.zeros(x.size // 32 * 34, np.uint8)
    lib().ref_quantize_q8_0(_p(x), _p(out), x.size)
    return out


def gemv(qtype: int, W: np.ndarray, x: np.ndarray, N: int, K: int) -> np.ndarray:
    W = np.ascontiguousarray(W)
    x = np.ascontiguousarray(x, np.float32)
    y = np.empty(N, np.float32)
    assert lib().ref_gemv(qtype, _p(W), _p(x), _p(y), N, K) == 0
    return y


def rmsnorm(x: np.ndarray, w: np.ndarray, eps: float) -> np.ndarray:
    x = np.ascontiguousarray(x, np.float32)
    w = np.ascontiguousarray(w, np.float32)
    y = np.empty_like(x)
    lib().ref_rmsnorm(_p(x), _p(w), _p(y), x.size, eps)
    return y


def rope_cache(pos: int, n_dims: int, base: float, freqs: np.ndarray | None) -> np.ndarray:
    out = np.empty(n_dims, np.float32)
    f = np.ascontiguousarray(freqs, np.float32) if freqs is not None else None
    lib().ref_rope_cache(pos, n_dims, base, _p(f) if f is not None else None, _p(out))
    return out


def rope(x: np.ndarray, n_heads: int, head_dim: int, n_dims: int, mode: int, cache: np.ndarray) -> np.ndarray:
    y = np.ascontiguousarray(x, np.float32).copy()
    lib().ref_rope(_p(y), n_heads, head_dim, n_dims, mode, _p(cache))
    return y


class RefGGUFModel:
    """A GGUF parsed by gguf-py and run by the C restatement, one token at a time."""

    def __init__(self, path: str, n_ctx: int = 512):
        import gguf
        self.reader = r = gguf.GGUFReader(path)
        self.t = {t.name: t for t in r.tensors}

        def kv(key, default=None):
            f = r.fields.get(key)
            if f is None:
                return default
            v = f.parts[f.data[0]]
            if f.types[0] == gguf.GGUFValueType.STRING:
                return bytes(v).decode()
            return v[0].item()

        arch = kv("general.architecture")
        self.arch = arch
        m = RefModel()
        m.arch = {"llama": 0, "qwen2": 1}[arch]
        m.n_layer = kv(f"{arch}.block_count")
        m.d_model = kv(f"{arch}.embedding_length")
        m.d_ff = kv(f"{arch}.feed_forward_length")
        m.n_head = kv(f"{arch}.attention.head_count")
        m.n_kv = kv(f"{arch}.attention.head_count_kv", m.n_head)
        m.head_dim = kv(f"{arch}.attention.key_length", m.d_model // m.n_head)
        m.rope_dim = kv(f"{arch}.rope.dimension_count", m.head_dim)
        m.eps = kv(f"{arch}.attention.layer_norm_rms_epsilon", 1e-5)
        m.rope_base = kv(f"{arch}.rope.freq_base", 10000.0)
        m.n_ctx = n_ctx
        self._keep = []

        def T(name):
            t = self.t[name]
            rt = RefTensor()
            arr = np.ascontiguousarray(t.data)  # memmap view of the file bytes
            self._keep.append(arr)
            rt.data = arr.ctypes.data
            rt.type = int(t.tensor_type)
            shape = [int(s) for s in t.shape]  # fastest-first
            rt.cols = shape[0]
            rt.rows = shape[1] if len(shape) > 1 else 1
            return rt

        def B(name):
            if name not in self.t:
                return None
            arr = np.ascontiguousarray(self.t[name].data).view(np.float32)
            self._keep.append(arr)
            return arr.ctypes.data

        m.tok_embd = T("token_embd.weight")
        m.vocab = m.tok_embd.rows
        m.output_norm = T("output_norm.weight")
        m.output = T("output.weight") if "output.weight" in self.t else m.tok_embd
        m.rope_freqs = B("rope_freqs.weight")
        layers = (RefLayer * m.n_layer)()
        for i in range(m.n_layer):
            p = f"blk.{i}."
            L = layers[i]
            L.attn_norm, L.wq, L.wk, L.wv = T(p + "attn_norm.weight"), T(p + "attn_q.weight"), T(p + "attn_k.weight"), T(p + "attn_v.weight")
            L.wo, L.ffn_norm = T(p + "attn_output.weight"), T(p + "ffn_norm.weight")
            L.w_gate, L.w_up, L.w_down = T(p + "ffn_gate.weight"), T(p + "ffn_up.weight"), T(p + "ffn_down.weight")
            L.bq, L.bk, L.bv = B(p + "attn_q.bias"), B(p + "attn_k.bias"), B(p + "attn_v.bias")
        self._layers = layers
        m.layers = C.cast(layers, C.POINTER(RefLayer))
        self.m = m
        self.ctx = lib().ref_ctx_new(C.byref(m))
        self.vocab = m.vocab
        self.d_model = m.d_model

    def decode(self, token: int, pos: int, want_logits: bool = True):
        logits = np.empty(self.vocab, np.float32) if want_logits else None
        tok = lib().ref_decode(self.ctx, int(token), int(pos), _p(logits) if want_logits else None)
        assert tok >= 0, "ref_decode rejected its input"
        return tok, logits

    def hidden(self) -> np.ndarray:
        return np.ctypeslib.as_array(lib().ref_ctx_hidden(self.ctx), shape=(self.d_model,)).copy()

    def activation(self, which: str) -> np.ndarray:
        m = self.m
        idx, n = {"hidden": (0, m.d_model), "q": (1, m.n_head * m.head_dim), "attn_out": (2, m.n_head * m.head_dim),
                  "ffn_h": (3, m.d_ff)}[which]
        return np.ctypeslib.as_array(lib().ref_ctx_buffer(self.ctx, idx), shape=(n,)).copy()

    def kv(self, layer: int, pos: int) -> tuple[np.ndarray, np.ndarray]:
        n = self.m.n_kv * self.m.head_dim
        k = np.ctypeslib.as_array(lib().ref_ctx_k(self.ctx, layer, pos), shape=(n,)).copy().view(np.float16)
        v = np.ctypeslib.as_array(lib().ref_ctx_v(self.ctx, layer, pos), shape=(n,)).copy().view(np.float16)
        return k, v

    def greedy(self, prompt: list[int], n_new: int) -> tuple[list[int], np.ndarray]:
        """Feed the prompt token by token, then generate n_new greedy tokens.
        Returns (generated ids, logits after the last prompt token)."""
        pos, tok, first = 0, None, None
        for t in prompt:
            tok, lg = self.decode(t, pos, want_logits=(pos == len(prompt) - 1))
            pos += 1
        first = lg
        out = []
        for _ in range(n_new):
            out.append(tok)
            tok, _ = self.decode(tok, pos, want_logits=False)
            pos += 1
        return out, first

    def close(self):
        if self.ctx:
            lib().ref_ctx_free(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
