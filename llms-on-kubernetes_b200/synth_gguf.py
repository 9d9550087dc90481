"""Synthetic GGUF files at the real models' shapes.

There are no model files and no network on either machine, so every input to
the parity tests and to bench.py is a GGUF written here: random *valid* quant
blocks (every byte pattern with a finite fp16 scale is a legal block, SURVEY
Appendix B), fp16 super-scales drawn so the dequantised weights have
std ~= 0.02, norm weights 1.0, and a synthetic vocabulary.

The writer below is independent of ``gguf.GGUFWriter`` on purpose: it streams
tensor data to disk as it is generated (the 8B file is 4.9 GB) and gives the
C++ parser a second, independent producer to agree with.  The on-disk layout it
emits is the one ``gguf.GGUFReader`` parses (gguf-py 0.19.0,
gguf_reader.py:132-185,259-287; constants.py:10-13) and tests/test_gguf_format.py
checks it field-for-field against that reader.
"""
from __future__ import annotations

import struct
import zlib
from dataclasses import dataclass, field

import numpy as np

GGUF_MAGIC = 0x46554747
GGUF_VERSION = 3
ALIGN = 32

# ggml type ids (gguf constants.py:4059-4093) and (block elems, block bytes)
F32, F16, Q8_0, Q4_K, Q5_K, Q6_K, BF16 = 0, 1, 8, 12, 13, 14, 30
BLOCK = {F32: (1, 4), F16: (1, 2), BF16: (1, 2), Q8_0: (32, 34),
         Q4_K: (256, 144), Q5_K: (256, 176), Q6_K: (256, 210)}
TYPE_NAME = {F32: "F32", F16: "F16", BF16: "BF16", Q8_0: "Q8_0", Q4_K: "Q4_K",
             Q5_K: "Q5_K", Q6_K: "Q6_K"}

# GGUF metadata value types (constants.py:4157-4170)
T_U8, T_I8, T_U16, T_I16, T_U32, T_I32, T_F32, T_BOOL, T_STR, T_ARR, T_U64, T_I64, T_F64 = range(13)
_SCALAR_FMT = {T_U8: "<B", T_I8: "<b", T_U16: "<H", T_I16: "<h", T_U32: "<I", T_I32: "<i",
               T_F32: "<f", T_BOOL: "<?", T_U64: "<Q", T_I64: "<q", T_F64: "<d"}


@dataclass
class ModelSpec:
    name: str
    arch: str            # "llama" | "qwen2"
    n_layer: int
    d_model: int
    d_ff: int
    n_head: int
    n_kv: int
    head_dim: int
    vocab: int
    rope_base: float
    eps: float
    ftype: str           # "Q4_K_M" | "Q8_0" | "Q5_K_M" | "Q6_K" | "F16"
    ctx: int = 8192
    tied: bool = False
    qkv_bias: bool = False
    rope_freqs: bool = False
    extra_kv: dict = field(default_factory=dict)


SPECS = {
    # real shapes
    "tinyllama-1.1b-q4_k_m": ModelSpec("TinyLlama-1.1B-Chat", "llama", 22, 2048, 5632, 32, 4, 64, 32000, 1e4, 1e-5, "Q4_K_M", ctx=2048),
    "qwen2.5-1.5b-q8_0": ModelSpec("Qwen2.5-1.5B-Instruct", "qwen2", 28, 1536, 8960, 12, 2, 128, 151936, 1e6, 1e-6, "Q8_0", ctx=32768, tied=True, qkv_bias=True),
    "llama-3.1-8b-q4_k_m": ModelSpec("Llama-3.1-8B-Instruct", "llama", 32, 4096, 14336, 32, 8, 128, 128256, 5e5, 1e-5, "Q4_K_M", ctx=131072, rope_freqs=True),
    "llama-3.1-70b-q4_k_m": ModelSpec("Llama-3.1-70B-Instruct", "llama", 80, 8192, 28672, 64, 8, 128, 128256, 5e5, 1e-5, "Q4_K_M", ctx=131072, rope_freqs=True),
    # small shapes for parity tests (every K a multiple of 512 so Q6_K rows stay 16-byte multiples)
    "tiny-llama-q4_k_m": ModelSpec("tiny-llama", "llama", 3, 512, 1024, 8, 2, 64, 1024, 1e4, 1e-5, "Q4_K_M", ctx=512),
    "tiny-llama3-q4_k_m": ModelSpec("tiny-llama3", "llama", 2, 512, 1536, 4, 2, 128, 768, 5e5, 1e-5, "Q4_K_M", ctx=512, rope_freqs=True),
    "tiny-llama-q5_k_m": ModelSpec("tiny-llama-q5", "llama", 2, 512, 1024, 8, 8, 64, 512, 1e4, 1e-5, "Q5_K_M", ctx=512),
    "tiny-qwen2-q8_0": ModelSpec("tiny-qwen2", "qwen2", 2, 512, 1280, 4, 2, 128, 1024, 1e6, 1e-6, "Q8_0", ctx=512, tied=True, qkv_bias=True),
    "tiny-llama-f16": ModelSpec("tiny-llama-f16", "llama", 2, 256, 512, 4, 2, 64, 512, 1e4, 1e-5, "F16", ctx=256),
}


def use_more_bits(i: int, n: int) -> bool:
    """llama.cpp's Q4_K_M / Q5_K_M layer policy ([UPSTREAM-MEMORY])."""
    return i < n // 8 or i >= 7 * n // 8 or (i - n // 8) % 3 == 2


def tensor_type(spec: ModelSpec, name: str, layer: int | None) -> int:
    ft = spec.ftype
    if ft == "F16":
        return F16
    if ft == "Q8_0":
        return Q8_0
    if ft == "Q6_K":
        return Q6_K
    base = Q4_K if ft == "Q4_K_M" else Q5_K
    if name == "output.weight":
        return Q6_K
    if layer is not None and (name.endswith("attn_v.weight") or name.endswith("ffn_down.weight")):
        if use_more_bits(layer, spec.n_layer):
            return Q6_K
        if name.endswith("attn_v.weight") and spec.n_head // spec.n_kv >= 8 and ft == "Q4_K_M" and spec.n_layer >= 80:
            return Q5_K
    return base


def _rng_for(name: str, seed: int) -> np.random.Generator:
    return np.random.default_rng((zlib.crc32(name.encode()) ^ (seed * 0x9E3779B1)) & 0xFFFFFFFF)


def _f16_bytes(vals: np.ndarray) -> np.ndarray:
    return vals.astype(np.float16).view(np.uint8).reshape(vals.shape[0], 2)


def random_blocks(qtype: int, n_blocks: int, rng: np.random.Generator, std: float = 0.02,
                  edge_cases: bool = False) -> np.ndarray:
    """n_blocks random valid blocks of `qtype` as a uint8 array [n_blocks, block_bytes].

    Scale fields are finite fp16; all other bytes are uniform.  The scale ranges give
    dequantised values of roughly zero mean and the requested std.
    """
    _, bsz = BLOCK[qtype]
    raw = rng.integers(0, 256, size=(n_blocks, bsz), dtype=np.uint8)
    u = rng.uniform(0.6, 1.4, size=n_blocks)
    if qtype == Q8_0:
        d = std / 73.0 * u
        raw[:, 0:2] = _f16_bytes(d)
    elif qtype == Q4_K:
        d = std / 258.0 * u
        raw[:, 0:2] = _f16_bytes(d)
        raw[:, 2:4] = _f16_bytes(d * 7.5 * rng.uniform(0.9, 1.1, size=n_blocks))
    elif qtype == Q5_K:
        d = std / 527.0 * u
        raw[:, 0:2] = _f16_bytes(d)
        raw[:, 2:4] = _f16_bytes(d * 15.5 * rng.uniform(0.9, 1.1, size=n_blocks))
    elif qtype == Q6_K:
        d = std / 1369.0 * u
        raw[:, 208:210] = _f16_bytes(d)
    else:
        raise ValueError(qtype)
    if edge_cases and n_blocks >= 16:
        # d = 0, fp16 subnormal d, max 6-bit scales, Q6_K scale -128 / q = -32, negative d
        def setd(i, v):
            b = np.array([v], dtype=np.float16).view(np.uint8)
            off = 208 if qtype == Q6_K else 0
            raw[i, off:off + 2] = b
        setd(0, 0.0)
        setd(1, 5.96e-8)      # smallest fp16 subnormal
        setd(2, 6.0e-5)       # just below the smallest normal
        setd(3, -1.5e-3)      # negative scale is legal
        setd(4, 65504.0)      # fp16 max
        if qtype in (Q4_K, Q5_K):
            raw[5, 4:16] = 0xFF   # every 6-bit scale/min = 63
            raw[6, 4:16] = 0x00
            raw[7, 16:] = 0xFF    # every quant at its maximum
            raw[8, 16:] = 0x00
            raw[9, 2:4] = np.array([0.0], dtype=np.float16).view(np.uint8)  # dmin = 0
        if qtype == Q6_K:
            raw[5, 192:208] = 0x80  # scale = -128
            raw[6, 0:192] = 0x00    # q = -32 everywhere
            raw[7, 0:192] = 0xFF    # q = +31 everywhere
            raw[8, 192:208] = 0x7F
        if qtype == Q8_0:
            raw[5, 2:] = 0x80       # q = -128
            raw[6, 2:] = 0x7F
            raw[7, 2:] = 0x00
    return raw


def random_tensor_bytes(qtype: int, n_rows: int, n_cols: int, rng: np.random.Generator, std: float = 0.02) -> np.ndarray:
    be, _ = BLOCK[qtype]
    if qtype == F32:
        return (rng.standard_normal((n_rows, n_cols), dtype=np.float32) * std).view(np.uint8).reshape(-1)
    if qtype == F16:
        return (rng.standard_normal((n_rows, n_cols), dtype=np.float32) * std).astype(np.float16).view(np.uint8).reshape(-1)
    assert n_cols % be == 0, (n_cols, be)
    return random_blocks(qtype, n_rows * (n_cols // be), rng, std).reshape(-1)


def llama3_rope_factors(head_dim: int, base: float, factor=8.0, low=1.0, high=4.0, old_ctx=8192) -> np.ndarray:
    """rope_freqs.weight as llama.cpp's converter derives it for Llama-3.1 ([UPSTREAM-MEMORY]);
    the same smooth-interpolation rule HF applies for rope_type='llama3'."""
    freqs = 1.0 / (base ** (np.arange(0, head_dim, 2, dtype=np.float32) / head_dim))
    low_wl, high_wl = old_ctx / low, old_ctx / high
    out = []
    for f in freqs:
        wl = 2 * np.pi / f
        if wl < high_wl:
            out.append(1.0)
        elif wl > low_wl:
            out.append(factor)
        else:
            smooth = (old_ctx / wl - low) / (high - low)
            out.append(1.0 / ((1 - smooth) / factor + smooth))
    return np.array(out, dtype=np.float32)


# ----------------------------------------------------------------------------------------
# streaming GGUF writer
# ----------------------------------------------------------------------------------------
def _gstr(s: str | bytes) -> bytes:
    b = s.encode("utf-8") if isinstance(s, str) else s
    return struct.pack("<Q", len(b)) + b


def _kv(key: str, vtype: int, val, sub: int | None = None) -> bytes:
    out = _gstr(key) + struct.pack("<I", vtype)
    if vtype == T_STR:
        out += _gstr(val)
    elif vtype == T_ARR:
        out += struct.pack("<IQ", sub, len(val))
        if sub == T_STR:
            out += b"".join(_gstr(v) for v in val)
        else:
            out += np.asarray(val).astype(np.dtype(_SCALAR_FMT[sub][1:]).newbyteorder("<")).tobytes()
    else:
        out += struct.pack(_SCALAR_FMT[vtype], val)
    return out


class GGUFStreamWriter:
    def __init__(self, path: str, alignment: int = ALIGN):
        self.path, self.align = path, alignment
        self.kvs: list[bytes] = []
        self.tensors: list[tuple[str, tuple[int, ...], int, int]] = []  # name, dims(fastest first), type, nbytes
        if alignment != ALIGN:
            self.add("general.alignment", T_U32, alignment)

    def add(self, key, vtype, val, sub=None):
        self.kvs.append(_kv(key, vtype, val, sub))

    def declare(self, name: str, shape_rows_cols: tuple[int, ...], qtype: int):
        """shape in numpy order (rows, cols); GGUF stores dims fastest-first."""
        dims = tuple(reversed(shape_rows_cols))
        be, bb = BLOCK[qtype]
        n = 1
        for d in dims:
            n *= d
        assert dims[0] % be == 0
        self.tensors.append((name, dims, qtype, n // be * bb))

    def write(self, gen):
        """gen(name, dims, qtype, nbytes) -> uint8 ndarray of exactly nbytes, called in order."""
        a = self.align
        with open(self.path, "wb") as f:
            f.write(struct.pack("<IIQQ", GGUF_MAGIC, GGUF_VERSION, len(self.tensors), len(self.kvs)))
            for kv in self.kvs:
                f.write(kv)
            off = 0
            for name, dims, qt, nb in self.tensors:
                f.write(_gstr(name) + struct.pack("<I", len(dims)) + struct.pack(f"<{len(dims)}Q", *dims)
                        + struct.pack("<IQ", qt, off))
                off += (nb + a - 1) // a * a
            pos = f.tell()
            f.write(b"\0" * ((a - pos % a) % a))
            for name, dims, qt, nb in self.tensors:
                data = gen(name, dims, qt, nb)
                assert data.dtype == np.uint8 and data.size == nb, (name, data.size, nb)
                f.write(memoryview(data))
                f.write(b"\0" * ((a - nb % a) % a))


def synth_vocab(n: int) -> tuple[list[bytes], list[float], list[int]]:
    """SPM-style vocabulary: <unk>, <s>, </s>, 256 byte tokens, a few words, then filler."""
    toks: list[bytes] = [b"<unk>", b"<s>", b"</s>"]
    types = [2, 3, 3]  # UNKNOWN, CONTROL, CONTROL
    for b in range(256):
        toks.append(f"<0x{b:02X}>".encode())
        types.append(6)  # BYTE
    words = ["▁", "▁the", "▁a", "▁hello", "▁world", "e", "t", "a", "o", "th", "he", "▁t",
             "▁h", "ll", "llo", "▁he", "▁hel", "▁w", "or", "ld", "▁wor", ".", ",", "!", "?",
             "<|user|>", "<|assistant|>", "<|system|>", "\n"]
    for w in words:
        if len(toks) < n:
            toks.append(w.encode())
            types.append(4 if w.startswith("<|") else 1)  # chat markers are USER_DEFINED: matched verbatim
    i = 0
    while len(toks) < n:
        toks.append(f"▁tok{i}".encode())
        types.append(1)
        i += 1
    toks, types = toks[:n], types[:n]
    scores = [0.0, 0.0, 0.0] + [-1000.0 - k for k in range(len(toks) - 3)]
    # longer known pieces get better scores so SPM merging prefers them
    for k, t in enumerate(toks):
        if types[k] == 1:
            scores[k] = -10.0 + len(t.decode()) if not t.startswith("▁tok".encode()) else -50.0
    return toks, scores, types


ZEPHYR_TEMPLATE = ("{% for message in messages %}{% if message['role'] == 'user' %}{{ '<|user|>\n' + message['content'] + eos_token }}"
                   "{% elif message['role'] == 'system' %}{{ '<|system|>\n' + message['content'] + eos_token }}"
                   "{% elif message['role'] == 'assistant' %}{{ '<|assistant|>\n'  + message['content'] + eos_token }}{% endif %}"
                   "{% if loop.last and add_generation_prompt %}{{ '<|assistant|>' }}{% endif %}{% endfor %}")


def tensor_plan(spec: ModelSpec) -> list[tuple[str, tuple[int, ...], int]]:
    """[(name, numpy-order shape, ggml type)] in file order."""
    d, ff, hd = spec.d_model, spec.d_ff, spec.head_dim
    q_out, kv_out = spec.n_head * hd, spec.n_kv * hd
    plan = [("token_embd.weight", (spec.vocab, d), tensor_type(spec, "token_embd.weight", None))]
    if spec.rope_freqs:
        plan.append(("rope_freqs.weight", (hd // 2,), F32))
    for i in range(spec.n_layer):
        p = f"blk.{i}."
        plan.append((p + "attn_norm.weight", (d,), F32))
        for nm, rows in (("attn_q", q_out), ("attn_k", kv_out), ("attn_v", kv_out)):
            plan.append((p + nm + ".weight", (rows, d), tensor_type(spec, p + nm + ".weight", i)))
            if spec.qkv_bias:
                plan.append((p + nm + ".bias", (rows,), F32))
        plan.append((p + "attn_output.weight", (d, q_out), tensor_type(spec, p + "attn_output.weight", i)))
        plan.append((p + "ffn_norm.weight", (d,), F32))
        plan.append((p + "ffn_gate.weight", (ff, d), tensor_type(spec, p + "ffn_gate.weight", i)))
        plan.append((p + "ffn_up.weight", (ff, d), tensor_type(spec, p + "ffn_up.weight", i)))
        plan.append((p + "ffn_down.weight", (d, ff), tensor_type(spec, p + "ffn_down.weight", i)))
    plan.append(("output_norm.weight", (d,), F32))
    if not spec.tied:
        plan.append(("output.weight", (spec.vocab, d), tensor_type(spec, "output.weight", None)))
    return plan


def weight_stream_bytes(spec: ModelSpec) -> int:
    """Bytes a decode step must read once (SURVEY §8d W_stream): every layer tensor + output
    norm + lm_head (token_embd when tied); the embedding table itself is not streamed."""
    total = 0
    for name, shape, qt in tensor_plan(spec):
        if name == "token_embd.weight" and not spec.tied:
            continue
        if name == "rope_freqs.weight":
            continue
        be, bb = BLOCK[qt]
        n = int(np.prod(shape))
        total += n // be * bb
    return total


def kv_bytes_per_token(spec: ModelSpec) -> int:
    return 2 * spec.n_layer * spec.n_kv * spec.head_dim * 2  # K and V, fp16


def matmul_params(spec: ModelSpec) -> int:
    """Weights in the layer matrices every prompt token multiplies (Q, K, V, O, gate, up, down; not the lm_head)."""
    qdim, kvdim = spec.n_head * spec.head_dim, spec.n_kv * spec.head_dim
    return spec.n_layer * (spec.d_model * (qdim + 2 * kvdim) + qdim * spec.d_model + 3 * spec.d_model * spec.d_ff)


def write_model(path: str, spec: ModelSpec | str, seed: int = 0, vocab_strings: bool = True,
                norm_jitter: float = 0.0) -> ModelSpec:
    if isinstance(spec, str):
        spec = SPECS[spec]
    w = GGUFStreamWriter(path)
    a = spec.arch
    w.add("general.architecture", T_STR, a)
    w.add("general.name", T_STR, spec.name)
    w.add("general.file_type", T_U32, {"Q4_K_M": 15, "Q8_0": 7, "Q5_K_M": 17, "Q6_K": 18, "F16": 1}[spec.ftype])
    w.add("general.quantization_version", T_U32, 2)
    w.add(f"{a}.block_count", T_U32, spec.n_layer)
    w.add(f"{a}.context_length", T_U32, spec.ctx)
    w.add(f"{a}.embedding_length", T_U32, spec.d_model)
    w.add(f"{a}.feed_forward_length", T_U32, spec.d_ff)
    w.add(f"{a}.attention.head_count", T_U32, spec.n_head)
    w.add(f"{a}.attention.head_count_kv", T_U32, spec.n_kv)
    w.add(f"{a}.attention.layer_norm_rms_epsilon", T_F32, spec.eps)
    w.add(f"{a}.rope.freq_base", T_F32, spec.rope_base)
    if a == "llama":
        w.add(f"{a}.rope.dimension_count", T_U32, spec.head_dim)
        w.add(f"{a}.vocab_size", T_U32, spec.vocab)
    if spec.head_dim * spec.n_head != spec.d_model:
        w.add(f"{a}.attention.key_length", T_U32, spec.head_dim)
        w.add(f"{a}.attention.value_length", T_U32, spec.head_dim)
    for k, (t, v) in spec.extra_kv.items():
        w.add(k, t, v)
    # tokenizer (synthetic SPM-style vocabulary; throughput/parity runs feed token ids)
    w.add("tokenizer.ggml.model", T_STR, "llama")
    if vocab_strings:
        toks, scores, types = synth_vocab(spec.vocab)
        w.add("tokenizer.ggml.tokens", T_ARR, toks, T_STR)
        w.add("tokenizer.ggml.scores", T_ARR, scores, T_F32)
        w.add("tokenizer.ggml.token_type", T_ARR, types, T_I32)
    w.add("tokenizer.ggml.bos_token_id", T_U32, 1)
    w.add("tokenizer.ggml.eos_token_id", T_U32, 2)
    w.add("tokenizer.ggml.unknown_token_id", T_U32, 0)
    w.add("tokenizer.ggml.add_bos_token", T_BOOL, True)
    w.add("tokenizer.ggml.add_eos_token", T_BOOL, False)
    w.add("tokenizer.chat_template", T_STR, ZEPHYR_TEMPLATE)

    plan = tensor_plan(spec)
    for name, shape, qt in plan:
        w.declare(name, shape, qt)
    shapes = {name: (shape, qt) for name, shape, qt in plan}

    def gen(name, dims, qt, nb):
        rng = _rng_for(name, seed)
        shape, _ = shapes[name]
        if name == "rope_freqs.weight":
            return llama3_rope_factors(spec.head_dim, spec.rope_base).view(np.uint8)
        if name.endswith("_norm.weight"):
            v = np.ones(shape, dtype=np.float32)
            if norm_jitter:
                v += rng.uniform(-norm_jitter, norm_jitter, size=shape).astype(np.float32)
            return v.view(np.uint8).reshape(-1)
        if name.endswith(".bias"):
            return (rng.standard_normal(shape, dtype=np.float32) * 0.02).view(np.uint8).reshape(-1)
        rows = int(np.prod(shape[:-1]))
        return random_tensor_bytes(qt, rows, shape[-1], rng)

    w.write(gen)
    return spec


if __name__ == "__main__":
    import argparse
    import time
    ap = argparse.ArgumentParser(description="write a synthetic GGUF at a real model's shape")
    ap.add_argument("spec", choices=sorted(SPECS))
    ap.add_argument("out")
    ap.add_argument("--seed", type=int, default=0)
    args = ap.parse_args()
    t0 = time.time()
    s = write_model(args.out, args.spec, args.seed)
    print(f"{args.out}: {s.name} W_stream={weight_stream_bytes(s)/1e9:.4f} GB kv/token={kv_bytes_per_token(s)} B "
          f"({time.time()-t0:.1f}s)")
