"""The BPE pre-tokenisers: `bpe_pre_split` in csrc/server/tokenizer.cpp matches the
llama-3, qwen2 and gpt-2 regexes by hand over Unicode code points.  Checked piece for piece against the `tokenizers`
library (the engine HF's and llama.cpp's vocabularies were built with) running the same regexes, on hand-picked and
random strings: letters of several scripts, digits of several kinds, punctuation, emoji, combining marks, all sorts
of whitespace, contractions in both cases."""
import json
import subprocess

import numpy as np
import pytest

tokenizers = pytest.importorskip("tokenizers")

PATTERNS = {
    "llama3": r"(?i:'s|'t|'re|'ve|'m|'ll|'d)|[^\r\n\p{L}\p{N}]?\p{L}+|\p{N}{1,3}| ?[^\s\p{L}\p{N}]+[\r\n]*|\s*[\r\n]+|\s+(?!\S)|\s+",
    "qwen2": r"(?i:'s|'t|'re|'ve|'m|'ll|'d)|[^\r\n\p{L}\p{N}]?\p{L}+|\p{N}| ?[^\s\p{L}\p{N}]+[\r\n]*|\s*[\r\n]+|\s+(?!\S)|\s+",
    "gpt2": r"'s|'t|'re|'ve|'m|'ll|'d| ?\p{L}+| ?\p{N}+| ?[^\s\p{L}\p{N}]+|\s+(?!\S)|\s+",
}
ALPHABET = (list("abzAZ") + list("0159") + list(".,!?'\"-_()[]{}<>/\\@#$%^&*+=~`|;:") + [" ", " ", " ", "\t", "\n", "\r", "\r\n", " ", "　", " "] +
            list("éßñÇ") + list("жЯ") + list("漢字かなカナ한") + list("٣۷२") + list("²½Ⅷ") + ["́", "‍", "🙂", "👍🏽", "€", "™", "'s", "'T", "'re", "'LL", "'d", "'x"])
FIXED = ["", " ", "  ", "\n", " \n ", "Hello, world's  123456 tests!!\n  done", "I'LL be there, won't I?", "a\tb\t\tc", "x  \n\n  y", "1234567 12 3",
         "  leading and trailing  ", "tab\t", "end with newline\n", "\r\n\r\n", "price: $1,234.50!!!\n\nnext", "naïve café № 5", "日本語のテキスト123と４５６", "don't'll've",
         "a'sb'Tc", "   \t  x", "!!!   ???", "emoji 🙂🙂 ok", "x y　z", "UPPER lower MiXeD", "___a___", "'", "''s", "' s"]


def corpus():
    rng = np.random.default_rng(42)
    out = list(FIXED)
    for _ in range(1500):
        out.append("".join(ALPHABET[i] for i in rng.integers(0, len(ALPHABET), size=int(rng.integers(1, 30)))))
    return out


@pytest.mark.parametrize("kind", list(PATTERNS))
def test_pre_split_equals_the_regex(b2, tmp_path, kind):
    texts = corpus()
    pt = tokenizers.pre_tokenizers.Split(tokenizers.Regex(PATTERNS[kind]), behavior="isolated", invert=False)
    want = [[p for p, _ in pt.pre_tokenize_str(t)] for t in texts]
    path = tmp_path / "texts.json"
    path.write_text(json.dumps(texts), encoding="utf-8")
    r = subprocess.run([b2.SERVER_PATH, "--pre-split", kind, "@" + str(path)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    got = json.loads(r.stdout)["pieces"]
    assert len(got) == len(texts)
    bad = [(t, w, g) for t, w, g in zip(texts, want, got) if w != g]
    assert not bad, f"{len(bad)} of {len(texts)} differ, first: {bad[0]!r}"
    for t, g in zip(texts, got):
        assert "".join(g) == t          # the pieces partition the text
