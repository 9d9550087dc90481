"""deploy/: the values overlay keeps the shape the reference chart's templates read (ramalama-models/helm-chart/values.yaml:
image.repository / tag / pullPolicy, models[].modelPath / modelName / replicas / resources), and the image puts
llama-server on PATH.  CPU only."""
import os

import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_values_overlay_has_the_charts_shape():
    v = yaml.safe_load(open(os.path.join(ROOT, "deploy", "values-b200.yaml")))
    assert set(v["image"]) >= {"repository", "tag", "pullPolicy"}
    assert v["models"], "at least one model"
    for m in v["models"]:
        assert m["modelPath"].startswith("/mnt/models/") and m["modelPath"].endswith(".gguf")
        assert m["modelName"] and int(m["replicas"]) >= 1
        assert m["resources"]["limits"]["nvidia.com/gpu"] == 1      # one replica = one GPU
    ref = "/root/reference/ramalama-models/helm-chart/values.yaml"
    if os.path.exists(ref):
        r = yaml.safe_load(open(ref))
        assert set(r["image"]) <= set(v["image"])
        assert set(r["models"][0]) - {"resources"} <= set(v["models"][0])


def test_image_puts_the_server_on_path():
    text = open(os.path.join(ROOT, "deploy", "Dockerfile")).read()
    assert "bin/llama-server" in text and "libb2.so" in text and "ENV PATH=/opt/b2/bin" in text
