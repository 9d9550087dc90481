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


def test_probes_patch_targets_what_the_chart_renders():
    """deploy/probes-patch.yaml is a strategic-merge patch: it must name the container and the port of the chart's
    Deployment template, and probe the path the server answers while loading (503) and when ready (200)."""
    patch = yaml.safe_load(open(os.path.join(ROOT, "deploy", "probes-patch.yaml")))
    cont = patch["spec"]["template"]["spec"]["containers"]
    assert len(cont) == 1 and cont[0]["name"] == "ramalama"
    for probe in ("readinessProbe", "startupProbe", "livenessProbe"):
        assert cont[0][probe]["httpGet"] == {"path": "/health", "port": 8080}
    ann = patch["spec"]["template"]["metadata"]["annotations"]
    assert ann["prometheus.io/path"] == "/metrics" and ann["prometheus.io/port"] == "8080"
    tpl = "/root/reference/ramalama-models/helm-chart/templates/model-deployments.yaml"
    if os.path.exists(tpl):
        text = open(tpl).read()
        assert "- name: ramalama" in text and "containerPort: 8080" in text and "name: ramalama-{{ .modelName }}" in text
    server = open(os.path.join(ROOT, "llms-on-kubernetes_b200", "csrc", "server", "main.cpp")).read()
    assert '"/health"' in server and '"/metrics"' in server
