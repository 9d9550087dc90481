"""Generates tests/golden/blocks_golden.npz: seeded random valid blocks per quant type, the sha256 of
`gguf.dequantize` on them (gguf-py 0.19.0 = llama.cpp's own Python statement of the formats) and the
first rows of that output.  Run from the repo root:  python tests/golden/make_block_golden.py"""
import hashlib
import importlib
import os
import sys

import gguf
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sg = importlib.import_module("llms-on-kubernetes_b200.synth_gguf")

GG = {sg.Q8_0: gguf.GGMLQuantizationType.Q8_0, sg.Q4_K: gguf.GGMLQuantizationType.Q4_K,
      sg.Q5_K: gguf.GGMLQuantizationType.Q5_K, sg.Q6_K: gguf.GGMLQuantizationType.Q6_K}
out = {"generator": np.array(f"gguf {getattr(gguf, '__version__', '0.19.0')}, numpy {np.__version__}")}
for qt, gt in GG.items():
    rng = np.random.default_rng(20261106 + qt)
    blocks = sg.random_blocks(qt, 64, rng, edge_cases=True)
    y = gguf.dequantize(blocks, gt).astype(np.float32)
    n = sg.TYPE_NAME[qt]
    out[f"{n}_blocks"] = blocks
    out[f"{n}_in_sha256"] = np.array(hashlib.sha256(blocks.tobytes()).hexdigest())
    out[f"{n}_out_sha256"] = np.array(hashlib.sha256(y.tobytes()).hexdigest())
    out[f"{n}_first_rows"] = y[:4]
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "blocks_golden.npz"), **out)
print("wrote blocks_golden.npz")
