"""SentencePiece ("llama") tokenisation end to end (config 1's TinyLlama and the llama-2 family): a
BPE-type SentencePiece model with byte fallback is trained here with the `sentencepiece` library, written into a GGUF
as llama.cpp's converter stores it (tokenizer.ggml.model = llama, .tokens, .scores, .token_type), and
`llama-server --tokenize` must give the library's ids and decode them back."""
import io
import json
import subprocess

import pytest

spm = pytest.importorskip("sentencepiece")
gguf = pytest.importorskip("gguf")

from test_tokenizer_bpe import TRAIN   # noqa: E402

TEXTS = ["Hello world", "The quick brown fox jumps over the lazy dog.", "It's 2024 and we'll see 1234567 things, won't we?", "  two  spaces", "tab\tand\nnewline",
         "naïve café Grüße München", "日本語のテキスト", "unseen: ☃ ∑ 🙂", "tokenization tokenizers tokenised", "a", " ", "UPPER lower MiXeD 42", "x=y+z*(3-1)/2",
         "trailing space ", "def f(x):\n    return x * 2"]


def test_spm_encode_matches_sentencepiece(b2, tmp_path):
    model = io.BytesIO()
    spm.SentencePieceTrainer.train(sentence_iterator=iter(TRAIN * 8), model_writer=model, vocab_size=400, model_type="bpe", byte_fallback=True,
                                   character_coverage=0.99, normalization_rule_name="identity", add_dummy_prefix=True, remove_extra_whitespaces=False,
                                   split_digits=True, bos_id=1, eos_id=2, unk_id=0, pad_id=-1, minloglevel=2)
    sp = spm.SentencePieceProcessor(model_proto=model.getvalue())
    n = sp.get_piece_size()
    toks, scores, types = [], [], []
    for i in range(n):
        toks.append(sp.id_to_piece(i))
        scores.append(sp.get_score(i))
        types.append(2 if sp.is_unknown(i) else 3 if sp.is_control(i) else 6 if sp.is_byte(i) else 1)
    path = str(tmp_path / "spm.gguf")
    w = gguf.GGUFWriter(path, "llama")
    w.add_tokenizer_model("llama")
    w.add_token_list(toks)
    w.add_token_scores(scores)
    w.add_token_types(types)
    w.add_bos_token_id(1)
    w.add_eos_token_id(2)
    w.add_unk_token_id(0)
    w.add_add_bos_token(True)
    w.write_header_to_file(); w.write_kv_data_to_file(); w.write_tensors_to_file(); w.close()
    bad = []
    for t in TEXTS + TRAIN[:7]:
        r = subprocess.run([b2.SERVER_PATH, "-m", path, "--tokenize", t], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        got = json.loads(r.stdout)["tokens"]
        want = [1] + sp.encode(t)
        if got != want:
            bad.append((t, want, got))
            continue
        r = subprocess.run([b2.SERVER_PATH, "-m", path, "--detokenize", ",".join(map(str, got))], capture_output=True, text=True)
        assert r.returncode == 0 and json.loads(r.stdout)["content"] == " " + t, (t, r.stdout)
    assert not bad, f"{len(bad)} differ, first: {bad[0]!r}"
