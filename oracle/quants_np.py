"""oracle/quants_np.py -- TEST INFRASTRUCTURE, NOT PRODUCT.

numpy restatement of the ggml block formats on the hot path, written element-by-element
from the layout description rather than by reshaping tricks, so
it is an independent check of both gguf-py and the CUDA kernels.

Pinned: tests/test_oracle_blocks.py compares every function here bit-for-bit with
``gguf.dequantize`` / ``gguf.quantize`` (gguf-py 0.19.0 = llama.cpp's own Python
package; gguf/quants.py:395-401 Q8_0, :475-522 Q4_K, :525-549 Q5_K, :552-572 Q6_K,
:378-393 Q8_0 quantiser) and with the committed fixtures tests/golden/blocks_*.npz.

Not pinned: ``quantize_q8_K`` and the integer dot products restate ggml's CPU
activation path (quantize_row_q8_K_ref, ggml_vec_dot_q4_K_q8_K, ...) from upstream
knowledge; no upstream statement of them exists on this machine ("parity unpinned",
SURVEY.md section 8c tier T3).  They are cross-checked against oracle/ggml_ref.c only.
"""
from __future__ import annotations

import numpy as np

QK_K = 256
F32, F16, Q8_0, Q4_K, Q5_K, Q6_K, BF16 = 0, 1, 8, 12, 13, 14, 30
BLOCK = {F32: (1, 4), F16: (1, 2), BF16: (1, 2), Q8_0: (32, 34), Q4_K: (256, 144), Q5_K: (256, 176), Q6_K: (256, 210)}


def _f16(b: np.ndarray) -> np.ndarray:
    return np.ascontiguousarray(b).view(np.float16).astype(np.float32)


def scale_min_k4(sc12: np.ndarray) -> tuple[np.ndarray, np.ndarray]:
    """[n,12] packed -> (sc[n,8], m[n,8]); ggml get_scale_min_k4 / gguf quants.py:479-501."""
    q = sc12.astype(np.uint8)
    sc = np.empty((q.shape[0], 8), np.uint8)
    m = np.empty((q.shape[0], 8), np.uint8)
    for j in range(8):
        if j < 4:
            sc[:, j] = q[:, j] & 63
            m[:, j] = q[:, j + 4] & 63
        else:
            sc[:, j] = (q[:, j + 4] & 0xF) | ((q[:, j - 4] >> 6) << 4)
            m[:, j] = (q[:, j + 4] >> 4) | ((q[:, j] >> 6) << 4)
    return sc, m


def unpack_q4_K(blocks: np.ndarray) -> np.ndarray:
    """[n,144] -> integer quants [n,256] in element order (gguf quants.py:519-520)."""
    qs = blocks[:, 16:144]
    q = np.empty((blocks.shape[0], 256), np.uint8)
    for c in range(4):
        chunk = qs[:, 32 * c:32 * c + 32]
        q[:, 64 * c:64 * c + 32] = chunk & 0xF
        q[:, 64 * c + 32:64 * c + 64] = chunk >> 4
    return q


def unpack_q5_K(blocks: np.ndarray) -> np.ndarray:
    qh = blocks[:, 16:48]
    qs = blocks[:, 48:176]
    q = np.empty((blocks.shape[0], 256), np.uint8)
    for c in range(4):
        chunk = qs[:, 32 * c:32 * c + 32]
        q[:, 64 * c:64 * c + 32] = (chunk & 0xF) | (((qh >> (2 * c)) & 1) << 4)
        q[:, 64 * c + 32:64 * c + 64] = (chunk >> 4) | (((qh >> (2 * c + 1)) & 1) << 4)
    return q


def unpack_q6_K(blocks: np.ndarray) -> np.ndarray:
    """[n,210] -> signed quants [n,256] (gguf quants.py:565-570)."""
    ql = blocks[:, 0:128]
    qh = blocks[:, 128:192]
    q = np.empty((blocks.shape[0], 256), np.int16)
    for h in range(2):
        l0 = ql[:, 64 * h:64 * h + 32]
        l1 = ql[:, 64 * h + 32:64 * h + 64]
        hh = qh[:, 32 * h:32 * h + 32]
        base = 128 * h
        q[:, base + 0:base + 32] = (l0 & 0xF) | (((hh >> 0) & 3) << 4)
        q[:, base + 32:base + 64] = (l1 & 0xF) | (((hh >> 2) & 3) << 4)
        q[:, base + 64:base + 96] = (l0 >> 4) | (((hh >> 4) & 3) << 4)
        q[:, base + 96:base + 128] = (l1 >> 4) | (((hh >> 6) & 3) << 4)
    return (q - 32).astype(np.int8)


def dequantize(raw: np.ndarray, qtype: int) -> np.ndarray:
    """raw uint8 [..., row_bytes] -> float32 [..., K]; bit-exact with gguf.dequantize."""
    raw = np.ascontiguousarray(raw, dtype=np.uint8)
    be, bb = BLOCK[qtype]
    lead = raw.shape[:-1]
    if qtype == F32:
        return raw.view(np.float32).reshape(*lead, -1)
    if qtype == F16:
        return raw.view(np.float16).astype(np.float32).reshape(*lead, -1)
    if qtype == BF16:
        return (raw.view(np.uint16).astype(np.uint32) << 16).view(np.float32).reshape(*lead, -1)
    b = raw.reshape(-1, bb)
    n = b.shape[0]
    if qtype == Q8_0:
        d = _f16(b[:, 0:2])
        y = b[:, 2:].view(np.int8).astype(np.float32) * d
    elif qtype in (Q4_K, Q5_K):
        d = _f16(b[:, 0:2])
        dmin = _f16(b[:, 2:4])
        sc, m = scale_min_k4(b[:, 4:16])
        q = (unpack_q4_K(b) if qtype == Q4_K else unpack_q5_K(b)).astype(np.float32).reshape(n, 8, 32)
        dd = (d * sc.astype(np.float32)).reshape(n, 8, 1)
        mm = (dmin * m.astype(np.float32)).reshape(n, 8, 1)
        y = (dd * q - mm).reshape(n, 256)
    elif qtype == Q6_K:
        d = _f16(b[:, 208:210])
        sc = b[:, 192:208].view(np.int8).astype(np.float32)
        q = unpack_q6_K(b).astype(np.float32).reshape(n, 16, 16)
        y = ((d * sc).reshape(n, 16, 1) * q).reshape(n, 256)
    else:
        raise ValueError(qtype)
    return y.astype(np.float32).reshape(*lead, -1)


def roundf(x: np.ndarray) -> np.ndarray:
    """C roundf: half away from zero."""
    a = np.abs(x)
    f = np.floor(a)
    return np.sign(x) * (f + np.floor(2 * (a - f)))


def quantize_q8_0(x: np.ndarray) -> np.ndarray:
    """float32 [..., K] -> uint8 [..., K/32*34]; ggml quantize_row_q8_0_ref (gguf quants.py:381-393)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    b = x.reshape(-1, 32)
    d = np.abs(b).max(axis=1, keepdims=True) / np.float32(127)
    with np.errstate(divide="ignore"):
        inv = np.where(d == 0, np.float32(0), np.float32(1) / d).astype(np.float32)
    qs = roundf(b * inv).astype(np.int8)
    out = np.concatenate([d.astype(np.float16).view(np.uint8), qs.view(np.uint8)], axis=1)
    return out.reshape(*x.shape[:-1], -1)


def nearest_int(v: np.ndarray) -> np.ndarray:
    """ggml nearest_int: fp32 magic-number rounding (round-half-to-even) [UPSTREAM-MEMORY]."""
    val = (v.astype(np.float32) + np.float32(12582912.0)).astype(np.float32)
    return (val.view(np.int32) & 0x007FFFFF) - 0x00400000


def quantize_q8_K(x: np.ndarray) -> tuple[np.ndarray, np.ndarray, np.ndarray]:
    """float32 [K] -> (d[K/256] f32, qs[K] i8, bsums[K/16] i16); ggml quantize_row_q8_K_ref [UPSTREAM-MEMORY]."""
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, 256)
    n = x.shape[0]
    idx = np.abs(x).argmax(axis=1)  # first maximum, like the strict '>' scan
    mx = x[np.arange(n), idx]
    d = np.zeros(n, np.float32)
    qs = np.zeros((n, 256), np.int8)
    nz = mx != 0
    with np.errstate(divide="ignore", invalid="ignore"):
        iscale = np.where(nz, np.float32(-127.0) / mx, np.float32(0)).astype(np.float32)
    v = nearest_int(iscale[:, None] * x)
    qs[nz] = np.minimum(127, v[nz]).astype(np.int8)
    with np.errstate(divide="ignore"):
        d[nz] = (np.float32(1) / iscale[nz]).astype(np.float32)
    bsums = qs.reshape(n, 16, 16).astype(np.int32).sum(axis=2).astype(np.int16)
    return d, qs.reshape(-1), bsums.reshape(-1)


def int_dot_k(raw_row: np.ndarray, qtype: int, d8: np.ndarray, q8: np.ndarray) -> np.ndarray:
    """Per-super-block exact pieces of ggml_vec_dot_{q4_K,q5_K,q6_K}_q8_K for one weight row:
    returns float64 [nblocks] of  d*d8*sum_s sc_s*dot_s - dmin*d8*sum_s m_s*bsum_s  (exact in f64)."""
    bb = BLOCK[qtype][1]
    b = raw_row.reshape(-1, bb)
    n = b.shape[0]
    q8 = q8.reshape(n, 256).astype(np.int64)
    if qtype in (Q4_K, Q5_K):
        d = _f16(b[:, 0:2]).reshape(n).astype(np.float64)
        dmin = _f16(b[:, 2:4]).reshape(n).astype(np.float64)
        sc, m = scale_min_k4(b[:, 4:16])
        q = (unpack_q4_K(b) if qtype == Q4_K else unpack_q5_K(b)).astype(np.int64)
        dots = (q * q8).reshape(n, 8, 32).sum(axis=2)
        bs = q8.reshape(n, 8, 32).sum(axis=2)
        isum = (dots * sc.astype(np.int64)).sum(axis=1)
        msum = (bs * m.astype(np.int64)).sum(axis=1)
        return d * d8.astype(np.float64) * isum - dmin * d8.astype(np.float64) * msum
    if qtype == Q6_K:
        d = _f16(b[:, 208:210]).reshape(n).astype(np.float64)
        sc = b[:, 192:208].view(np.int8).astype(np.int64)
        q = unpack_q6_K(b).astype(np.int64)
        dots = (q * q8).reshape(n, 16, 16).sum(axis=2)
        return d * d8.astype(np.float64) * (dots * sc).sum(axis=1)
    raise ValueError(qtype)
