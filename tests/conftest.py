"""pytest configuration: the `gpu` marker, import paths and shared synthetic-GGUF fixtures."""
import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

PKG = "llms-on-kubernetes_b200"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def synth():
    return importlib.import_module(PKG + ".synth_gguf")


@pytest.fixture(scope="session")
def b2():
    """The ctypes binding of libb2.so.  Importing it needs the built library, not a GPU."""
    return importlib.import_module(PKG)


@pytest.fixture(scope="session")
def gguf_dir(tmp_path_factory):
    return tmp_path_factory.mktemp("gguf")


@pytest.fixture(scope="session")
def tiny_models(synth, gguf_dir):
    """name -> path of the small synthetic models, written once per session."""
    out = {}
    for name in ("tiny-llama-q4_k_m", "tiny-llama3-q4_k_m", "tiny-llama-q5_k_m", "tiny-qwen2-q8_0"):
        p = str(gguf_dir / (name + ".gguf"))
        synth.write_model(p, name, seed=7, norm_jitter=0.1)
        out[name] = p
    return out
