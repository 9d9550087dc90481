"""The C++ GGUF parser (csrc/gguf.cpp, through the C ABI) against llama.cpp's own gguf-py reader,
field for field, on files from two independent writers (gguf.GGUFWriter and synth_gguf's streaming
writer).  Also malformed / truncated files.  CPU only: parsing needs no GPU."""
import os
import struct

import gguf
import numpy as np
import pytest


def _kv_python(reader, key):
    f = reader.fields[key]
    if f.types[0] == gguf.GGUFValueType.STRING:
        return bytes(f.parts[f.data[0]])
    if f.types[0] == gguf.GGUFValueType.ARRAY:
        if f.types[1] == gguf.GGUFValueType.STRING:
            return [bytes(f.parts[i]) for i in f.data]
        return [f.parts[i][0].item() for i in f.data] if len(f.data) and f.parts[f.data[0]].size == 1 else \
            [v.item() for i in f.data for v in f.parts[i]]
    return f.parts[f.data[0]][0].item()


def _compare(b2, path):
    r = gguf.GGUFReader(path)
    g = b2.Gguf(path)
    assert g.version == 3
    assert g.alignment == r.alignment
    assert g.data_offset == r.data_offset
    assert g.n_tensors == len(r.tensors)
    keys = [k for k in r.fields if not k.startswith("GGUF.")]
    assert g.keys() == keys
    for k in keys:
        want, got = _kv_python(r, k), g.get(k)
        if isinstance(want, float):
            assert got == pytest.approx(want, rel=0, abs=0) or np.float32(got) == np.float32(want), k
        elif isinstance(want, list) and want and isinstance(want[0], float):
            assert np.array_equal(np.array(got, np.float64), np.array(want, np.float64)), k
        else:
            assert got == want, k
    for i, t in enumerate(r.tensors):
        info = g.tensor_info(i)
        assert info["name"] == t.name
        assert info["dims"] == [int(s) for s in t.shape]
        assert info["type"] == int(t.tensor_type)
        assert info["offset"] == t.data_offset - r.data_offset
        assert info["n_bytes"] == t.n_bytes
    # raw bytes of a few tensors
    for i in (0, len(r.tensors) // 2, len(r.tensors) - 1):
        assert g.tensor_bytes(i) == bytes(np.ascontiguousarray(r.tensors[i].data).view(np.uint8))
    g.close()


def test_parser_matches_gguf_py_on_synth_writer(b2, tiny_models):
    for p in tiny_models.values():
        _compare(b2, p)


def test_parser_matches_gguf_py_on_gguf_writer(b2, synth, tmp_path):
    """A file produced by llama.cpp's own GGUFWriter, with every scalar KV type, nested-free arrays, a custom
    alignment and raw quantised tensors."""
    p = str(tmp_path / "w.gguf")
    w = gguf.GGUFWriter(p, "llama")
    w.add_custom_alignment(64)
    w.add_uint8("t.u8", 200); w.add_int8("t.i8", -100); w.add_uint16("t.u16", 60000); w.add_int16("t.i16", -30000)
    w.add_uint32("t.u32", 4000000000); w.add_int32("t.i32", -2000000000); w.add_float32("t.f32", 1.5)
    w.add_uint64("t.u64", 2**63 + 5); w.add_int64("t.i64", -2**62); w.add_float64("t.f64", 1e-300)
    w.add_bool("t.bool", True); w.add_string("t.str", "héllo ▁world")
    w.add_array("t.arr_i", [1, 2, 3]); w.add_array("t.arr_f", [0.5, -1.25]); w.add_array("t.arr_s", ["a", "", "ccc"])
    rng = np.random.default_rng(0)
    w.add_tensor("a.weight", synth.random_blocks(synth.Q4_K, 6, rng).reshape(3, -1), raw_dtype=gguf.GGMLQuantizationType.Q4_K)
    w.add_tensor("b.weight", synth.random_blocks(synth.Q6_K, 4, rng).reshape(2, -1), raw_dtype=gguf.GGMLQuantizationType.Q6_K)
    w.add_tensor("c.weight", rng.standard_normal(17).astype(np.float32))
    w.write_header_to_file(); w.write_kv_data_to_file(); w.write_tensors_to_file(); w.close()
    _compare(b2, p)
    g = b2.Gguf(p)
    assert g.alignment == 64 and g.get("t.u64") == 2**63 + 5 and g.get("t.bool") is True
    assert g.get("missing.key", "dflt") == "dflt"


def test_rejects_bad_files(b2, tiny_models, tmp_path):
    with pytest.raises(b2.B2Error) as e:
        b2.Gguf(str(tmp_path / "nope.gguf"))
    assert e.value.code == -2
    src = open(tiny_models["tiny-llama-q4_k_m"], "rb").read()
    cases = {
        "magic": b"GGUX" + src[4:],
        "version": src[:4] + struct.pack("<I", 9) + src[8:],
        "short": src[:20],
        "truncated_kv": src[:200],
        "truncated_data": src[:len(src) // 2],
        "huge_counts": src[:8] + struct.pack("<QQ", 2**40, 2**40) + src[24:],
    }
    for name, data in cases.items():
        p = tmp_path / (name + ".gguf")
        p.write_bytes(data)
        with pytest.raises(b2.B2Error) as e:
            b2.Gguf(str(p))
        assert e.value.code in (-3, -4), (name, e.value)
        assert str(e.value)


def test_synth_stream_bytes_match_survey(synth):
    """W_stream per SURVEY section 8d / BASELINE.md section 3."""
    want = {"tinyllama-1.1b-q4_k_m": 0.6302, "qwen2.5-1.5b-q8_0": 1.6406, "llama-3.1-8b-q4_k_m": 4.6174,
            "llama-3.1-70b-q4_k_m": 41.9215}
    for k, gb in want.items():
        assert round(synth.weight_stream_bytes(synth.SPECS[k]) / 1e9, 4) == gb
    assert synth.kv_bytes_per_token(synth.SPECS["llama-3.1-8b-q4_k_m"]) == 131072
