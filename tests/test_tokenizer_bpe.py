"""Byte-level BPE end to end: a BPE vocabulary trained here with the `tokenizers` library (byte-level
alphabet, llama-3 / qwen2 / gpt-2 pre-tokeniser, merges) is written into a GGUF the way llama.cpp's converter stores one
(tokenizer.ggml.model = gpt2, .tokens, .merges, .pre), and `llama-server --tokenize` must produce the ids the library
produces -- pre-split, byte mapping and merge order together -- and decode them back to the text."""
import json
import subprocess

import numpy as np
import pytest

tokenizers = pytest.importorskip("tokenizers")
gguf = pytest.importorskip("gguf")

from test_tokenizer_presplit import PATTERNS, corpus   # noqa: E402

PRE_NAME = {"llama3": "llama-bpe", "qwen2": "qwen2", "gpt2": "gpt-2"}
TRAIN = ["The quick brown fox jumps over the lazy dog. " * 3, "Hello, world! It's 2024 and we'll see 1234567 things, won't we?",
         "def f(x):\n    return x * 2  # comment\n\n", "naïve café, Grüße aus München; 日本語のテキストと漢字 123", "tokens and tokenizers, tokenization, tokenised",
         "  indented\n\tlines\r\nwith  spaces   everywhere  ", "emoji 🙂 and symbols €™ ±≠"] * 4


@pytest.mark.parametrize("kind", list(PATTERNS))
def test_bpe_encode_matches_the_tokenizers_library(b2, tmp_path, kind):
    from tokenizers import Regex, Tokenizer, decoders, models, pre_tokenizers, trainers
    tok = Tokenizer(models.BPE())
    tok.pre_tokenizer = pre_tokenizers.Sequence([pre_tokenizers.Split(Regex(PATTERNS[kind]), behavior="isolated", invert=False),
                                                 pre_tokenizers.ByteLevel(add_prefix_space=False, use_regex=False)])
    tok.decoder = decoders.ByteLevel()
    trainer = trainers.BpeTrainer(vocab_size=600, initial_alphabet=pre_tokenizers.ByteLevel.alphabet(), special_tokens=["<|bos|>", "<|eos|>"],
                                  show_progress=False)
    tok.train_from_iterator(TRAIN, trainer)
    model = json.loads(tok.to_str())["model"]
    vocab = sorted(model["vocab"].items(), key=lambda kv: kv[1])
    assert [i for _, i in vocab] == list(range(len(vocab)))
    merges = [m if isinstance(m, str) else " ".join(m) for m in model["merges"]]
    path = str(tmp_path / f"bpe-{kind}.gguf")
    w = gguf.GGUFWriter(path, "llama")
    w.add_tokenizer_model("gpt2")
    w.add_tokenizer_pre(PRE_NAME[kind])
    w.add_token_list([t for t, _ in vocab])
    w.add_token_types([3 if t in ("<|bos|>", "<|eos|>") else 1 for t, _ in vocab])
    w.add_token_merges(merges)
    w.add_bos_token_id(0)
    w.add_eos_token_id(1)
    w.add_add_bos_token(False)
    w.write_header_to_file(); w.write_kv_data_to_file(); w.write_tensors_to_file(); w.close()

    texts = [t for t in corpus()[:400] if t] + TRAIN[:7]
    bad = []
    for t in texts:
        r = subprocess.run([b2.SERVER_PATH, "-m", path, "--tokenize", t], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        got = json.loads(r.stdout)["tokens"]
        want = tok.encode(t, add_special_tokens=False).ids
        if got != want:
            bad.append((t, want, got))
            continue
        r = subprocess.run([b2.SERVER_PATH, "-m", path, "--detokenize", ",".join(map(str, got))], capture_output=True, text=True)
        assert r.returncode == 0 and json.loads(r.stdout)["content"] == t, (t, r.stdout)
    assert not bad, f"{len(bad)} of {len(texts)} differ, first: {bad[0]!r}"
