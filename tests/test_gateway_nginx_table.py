"""the vLLM chart's nginx gateway routes by the JSON `model` field in an OpenResty Lua block
(vllm-models/helm-chart/templates/model-gateway.yaml:51-82).  Neither nginx nor Lua exists on this machine, so the block
is restated as a function (tests/gateway_restated.py: nginx_upstream_for) and this file
  * pins the restatement to the template's TEXT where /root/reference is present (the constructs it relies on are there,
    in that order, and nothing else touches `upstream_name`), and
  * runs a table of requests through it and through the two Python gateways (the reference's gateway.py restated, and
    deploy/gateway.py): the three must pick the same backend for every row, because the model servers behind them are
    the same `llama-server` drop-in.
CPU only."""
import importlib.util
import json
import os
import re

import pytest

import gateway_restated as gr

REF = "/root/reference/vllm-models/helm-chart/templates/model-gateway.yaml"
MODELS = ["llama-3.1-8b", "qwen2.5-1.5b", "tinyllama"]

TABLE = [
    # body, expected upstream
    (json.dumps({"model": "qwen2.5-1.5b", "messages": []}).encode(), "vllm-qwen2.5-1.5b"),
    (json.dumps({"model": "llama-3.1-8b"}).encode(), "vllm-llama-3.1-8b"),
    (json.dumps({"model": "tinyllama", "stream": True}).encode(), "vllm-tinyllama"),
    (json.dumps({"model": "unknown-model"}).encode(), "default_backend"),
    (json.dumps({"messages": [{"role": "user", "content": "hi"}]}).encode(), "default_backend"),
    (json.dumps({"model": ""}).encode(), "default_backend"),
    (json.dumps({"model": None}).encode(), "default_backend"),
    (json.dumps(["model", "tinyllama"]).encode(), "default_backend"),
    (b"{not json", "default_backend"),
    (b"", "default_backend"),
    (None, "default_backend"),
    (json.dumps({"model": "TinyLlama"}).encode(), "default_backend"),        # comparison is exact, case included
    (json.dumps({"model": "tinyllama "}).encode(), "default_backend"),
]


def test_routing_table():
    for body, want in TABLE:
        assert gr.nginx_upstream_for(MODELS, body) == want, body
    assert gr.nginx_backend_host(MODELS, "default_backend") == "vllm-llama-3.1-8b:8080"
    assert gr.nginx_backend_host(MODELS, "vllm-tinyllama") == "vllm-tinyllama:8080"


@pytest.mark.skipif(not os.path.exists(REF), reason="the reference tree is not on this machine")
def test_restatement_follows_the_template_text():
    src = open(REF).read()
    block = src[src.index("access_by_lua_block"):src.index("proxy_pass")]
    # the default, the body read, the guarded decode, one equality test per model, the final assignment -- in this order
    order = ['local upstream_name = "default_backend"', "ngx.req.read_body()", "ngx.req.get_body_data()", "if body then",
             "pcall(cjson.decode, body)", "if success and json_data and json_data.model then", "{{- range .Values.models }}",
             'if model_name == "{{ .modelName }}" then', 'upstream_name = "vllm-{{ .modelName }}"', "ngx.var.upstream = upstream_name"]
    pos = -1
    for needle in order:
        nxt = block.find(needle, pos + 1)
        assert nxt > pos, f"template changed: '{needle}' not found after offset {pos}"
        pos = nxt
    assert len(re.findall(r"upstream_name\s*=", block)) == 2          # the default and the per-model assignment only
    assert 'server vllm-{{ (index .Values.models 0).modelName }}:8080;' in src       # default_backend = models[0]
    assert "server vllm-{{ .modelName }}:8080;" in src
    for h in gr.NGINX_FORWARDED_HEADERS:
        assert f"proxy_set_header {h} " in src
    assert "proxy_pass http://$upstream;" in src


def test_python_gateways_pick_the_same_backend():
    """The ramalama chart's gateway.py (restated) and deploy/gateway.py route the table the same way."""
    here = os.path.dirname(os.path.abspath(__file__))
    spec = importlib.util.spec_from_file_location("b2_deploy_gateway", os.path.join(here, "..", "deploy", "gateway.py"))
    dg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(dg)
    backends = {m: f"http://vllm-{m}:8080" for m in MODELS}
    default = backends[MODELS[0]]
    pick_restated = getattr(gr, "pick_backend", None)
    pick_deploy = getattr(dg, "pick_backend", None)
    for body, want in TABLE:
        host = "http://" + gr.nginx_backend_host(MODELS, want)
        if pick_restated:
            assert pick_restated(backends, default, body or b"") == host, ("gateway.py restated", body)
        if pick_deploy:
            assert pick_deploy(backends, default, body or b"") == host, ("deploy/gateway.py", body)
    assert pick_restated or pick_deploy, "no gateway exposes pick_backend(backends, default, body)"
