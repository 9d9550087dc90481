"""The caller side of the boundary: the reference's own gateway.py, rendered out of its
ConfigMap and run unchanged, against a stub backend -- and the check that tests/gateway_restated.py (used on the
GPU box, where /root/reference does not exist) behaves identically.  CPU only."""
import http.client
import http.server
import json
import os
import re
import socket
import subprocess
import sys
import threading
import time

import pytest

import gateway_restated

REF_YAML = "/root/reference/ramalama-models/helm-chart/templates/api-gateway.yaml"
MODELS = ["tinyllama", "phi-3-mini"]


class StubBackend(http.server.BaseHTTPRequestHandler):
    """Answers like a model server and records what reached it."""
    seen = []

    def _go(self):
        n = int(self.headers.get("Content-Length", 0))
        body = self.rfile.read(n) if n else b""
        StubBackend.seen.append({"name": self.server.name, "method": self.command, "path": self.path,
                                 "headers": {k.lower(): v for k, v in self.headers.items()}, "body": body,
                                 "version": self.request_version})
        if self.path == "/boom":
            self.send_response(500); self.send_header("Content-Type", "application/json"); self.end_headers()
            self.wfile.write(b'{"error":{"message":"kaput","type":"server_error","code":500}}')
            return
        out = json.dumps({"served_by": self.server.name, "path": self.path}).encode()
        self.send_response(200)
        self.send_header("Content-Type", "application/json")
        self.send_header("Content-Length", str(len(out)))
        self.send_header("X-Backend", self.server.name)
        self.end_headers()
        self.wfile.write(out)

    do_GET = do_POST = _go

    def log_message(self, *a):
        pass


def start_stub(name):
    srv = http.server.ThreadingHTTPServer(("127.0.0.1", 0), StubBackend)
    srv.name = name
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    return srv, srv.server_address[1]


def render_reference_gateway(backends: dict) -> str:
    """The ConfigMap holds gateway.py with three template constructs: two `range .Values.models` loops and one
    `index .Values.models 0` (api-gateway.yaml:22-26,47-54).  Render them for the given models, pointing the
    backend URLs at the stub servers instead of the cluster DNS names."""
    text = open(REF_YAML).read()
    start = text.index("  gateway.py: |") + len("  gateway.py: |\n")
    end = text.index("\n---", start)
    lines = [l[4:] if l.startswith("    ") else l for l in text[start:end].split("\n")]
    out, i = [], 0
    while i < len(lines):
        l = lines[i]
        if "{{- range .Values.models }}" in l:
            j = i + 1
            block = []
            while "{{- end }}" not in lines[j]:
                block.append(lines[j]); j += 1
            for name, url in backends.items():
                for b in block:
                    b = b.replace("{{ .modelName }}", name)
                    b = re.sub(r'"http://ramalama-[^"]*"', f'"{url}"', b)
                    out.append(b)
            i = j + 1
            continue
        if "(index .Values.models 0)" in l:
            l = re.sub(r'"http://ramalama-[^"]*"', f'"{list(backends.values())[0]}"', l)
        out.append(l)
        i += 1
    return "\n".join(out)


def free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def wait_port(port, timeout=10):
    t0 = time.time()
    while time.time() - t0 < timeout:
        try:
            socket.create_connection(("127.0.0.1", port), timeout=0.2).close(); return
        except OSError:
            time.sleep(0.05)
    raise RuntimeError("gateway did not start")


def call(port, method, path, body=None, headers=None):
    c = http.client.HTTPConnection("127.0.0.1", port, timeout=10)
    c.request(method, path, body=body, headers=headers or {})
    r = c.getresponse()
    data = r.read()
    c.close()
    return r.status, {k.lower(): v for k, v in r.getheaders()}, data


CASES = [
    ("GET", "/v1/models", None, {}),
    ("GET", "/health", None, {}),
    ("POST", "/v1/chat/completions", json.dumps({"model": "phi-3-mini", "messages": []}).encode(), {"Content-Type": "application/json", "Authorization": "Bearer x"}),
    ("POST", "/v1/chat/completions", json.dumps({"model": "tinyllama", "messages": []}).encode(), {"Content-Type": "application/json"}),
    ("POST", "/v1/chat/completions", json.dumps({"model": "not-deployed"}).encode(), {"Content-Type": "application/json"}),
    ("POST", "/v1/completions", b"this is not json", {"Content-Type": "application/json"}),
    ("POST", "/v1/completions", json.dumps(["a", "list"]).encode(), {}),
    ("GET", "/anything/else?x=1", None, {}),
    ("POST", "/boom", json.dumps({"model": "tinyllama"}).encode(), {}),
]


@pytest.fixture(scope="module")
def stubs():
    servers = {n: start_stub(n) for n in MODELS}
    yield {n: f"http://127.0.0.1:{p}" for n, (_, p) in servers.items()}
    for s, _ in servers.values():
        s.shutdown()


def run_cases(port):
    results = []
    for method, path, body, headers in CASES:
        StubBackend.seen.clear()
        status, hdrs, data = call(port, method, path, body, headers)
        seen = [{k: (v if k != "headers" else {h: x for h, x in v.items() if h in ("content-type", "authorization", "content-length")})
                 for k, v in s.items()} for s in StubBackend.seen]
        try:
            payload = json.loads(data)
        except Exception:
            payload = data
        if isinstance(payload, dict) and "error" in payload and isinstance(payload["error"], str):
            payload = {"error": "<text>"}   # the exact exception text is not part of the contract
        results.append({"status": status, "ctype": hdrs.get("content-type"), "payload": payload, "seen": seen})
    return results


def test_restated_gateway_contract(stubs):
    srv, port = gateway_restated.start(stubs)
    res = run_cases(port)
    srv.shutdown()
    assert res[0]["payload"] == {"object": "list", "data": [{"id": n, "object": "model", "created": 0, "owned_by": "ramalama"} for n in MODELS]}
    assert res[0]["seen"] == []                      # /v1/models never reaches a backend
    assert res[1]["status"] == 200 and res[1]["payload"] == b"OK" and res[1]["seen"] == []
    assert res[2]["payload"]["served_by"] == "phi-3-mini" and res[2]["seen"][0]["headers"]["authorization"] == "Bearer x"
    assert res[3]["payload"]["served_by"] == "tinyllama"
    assert res[4]["payload"]["served_by"] == "tinyllama"     # unknown model -> default backend, body untouched
    assert json.loads(res[4]["seen"][0]["body"])["model"] == "not-deployed"
    assert res[5]["payload"]["served_by"] == "tinyllama" and res[5]["seen"][0]["body"] == b"this is not json"
    assert res[6]["payload"]["served_by"] == "tinyllama"     # JSON that is not an object -> default
    assert res[7]["seen"][0]["method"] == "GET" and res[7]["seen"][0]["path"] == "/anything/else?x=1"
    assert res[8]["status"] == 502 and res[8]["payload"] == {"error": "<text>"}   # backend 500 is hidden behind 502
    for r in res[2:8]:
        assert r["seen"][0]["version"] == "HTTP/1.1" and r["status"] == 200


@pytest.mark.skipif(not os.path.exists(REF_YAML), reason="the reference tree is only present on the authoring machine")
def test_reference_gateway_matches_restatement(stubs, tmp_path):
    script = tmp_path / "gateway.py"
    script.write_text(render_reference_gateway(stubs))
    port = free_port()
    proc = subprocess.Popen([sys.executable, str(script)], env={**os.environ, "PORT": str(port)})
    try:
        wait_port(port)
        ref = run_cases(port)
    finally:
        proc.terminate(); proc.wait(timeout=5)
    srv, rport = gateway_restated.start(stubs)
    mine = run_cases(rport)
    srv.shutdown()
    assert ref == mine
