"""Restatement of the reference's API gateway for tests that run where /root/reference does not exist (the GPU box).

What it restates, in this repo's own words: the Python gateway the ramalama chart ships inside a ConfigMap
(ramalama-models/helm-chart/templates/api-gateway.yaml:9-111):
  * BACKENDS maps every modelName to its backend URL; the first model is the default (:21-26);
  * GET /v1/models is answered by the gateway itself from that list with created 0 / owned_by "ramalama" (:43-61);
  * GET /health answers 200 "OK" (:32-36);
  * everything else is proxied (:63-104): read Content-Length bytes, pick the backend whose name equals the JSON
    body's "model" (anything unparsable or unknown -> default), re-issue the request with urllib (method POST iff
    there is a body), forward all headers except Host and Connection, 300 s timeout, buffer the whole response,
    relay status + headers (minus Connection / Transfer-Encoding) + body; ANY exception, including a backend
    4xx/5xx, becomes 502 {"error": "<text>"};
  * single-threaded socketserver.TCPServer (:110-111).
tests/test_gateway_contract.py checks, where the reference is available, that this restatement and the real script
(rendered from the ConfigMap) answer identically.  It is test infrastructure, not product.
"""
from __future__ import annotations

import http.server
import json
import socketserver
import threading
import urllib.error
import urllib.request


def pick_backend(backends: dict[str, str], default: str, body: bytes) -> str:
    """api-gateway.yaml:67-77: the backend of the JSON body's `model`, else the default; any parse error keeps the default."""
    target = default
    if body:
        try:
            wanted = json.loads(body.decode()).get("model", "")
            if wanted in backends:
                target = backends[wanted]
        except Exception:
            pass
    return target


def make_gateway(backends: dict[str, str], port: int = 0):
    """backends: ordered {modelName: base_url}.  Returns (server, port); call serve_forever in a thread."""
    names = list(backends)
    default = backends[names[0]]

    class Handler(http.server.BaseHTTPRequestHandler):
        def do_GET(self):
            if self.path == "/v1/models":
                payload = {"object": "list",
                           "data": [{"id": n, "object": "model", "created": 0, "owned_by": "ramalama"} for n in names]}
                self._reply(200, "application/json", json.dumps(payload).encode())
            elif self.path == "/health":
                self._reply(200, "text/plain", b"OK")
            else:
                self._forward()

        def do_POST(self):
            self._forward()

        def _reply(self, status, ctype, body):
            self.send_response(status)
            self.send_header("Content-type", ctype)
            self.end_headers()
            self.wfile.write(body)

        def _forward(self):
            n = int(self.headers.get("Content-Length", 0))
            body = self.rfile.read(n) if n > 0 else b""
            target = pick_backend(backends, default, body)
            req = urllib.request.Request(target.rstrip("/") + self.path, data=body if body else None)
            for k, v in self.headers.items():
                if k.lower() not in ("host", "connection"):
                    req.add_header(k, v)
            try:
                with urllib.request.urlopen(req, timeout=300) as resp:
                    self.send_response(resp.status)
                    for k, v in resp.headers.items():
                        if k.lower() not in ("connection", "transfer-encoding"):
                            self.send_header(k, v)
                    self.end_headers()
                    self.wfile.write(resp.read())
            except Exception as e:  # backend errors included: the client only ever sees 502
                self._reply(502, "application/json", json.dumps({"error": str(e)}).encode())

        def log_message(self, *args):
            pass

    socketserver.TCPServer.allow_reuse_address = True
    srv = socketserver.TCPServer(("127.0.0.1", port), Handler)
    return srv, srv.server_address[1]


def start(backends: dict[str, str]):
    srv, port = make_gateway(backends)
    t = threading.Thread(target=srv.serve_forever, daemon=True)
    t.start()
    return srv, port


# ---- the vLLM chart's nginx gateway (vllm-models/helm-chart/templates/model-gateway.yaml:51-82), restated ----
# nginx / OpenResty / Lua are not on this machine, so the routing block is restated as a table-driven function and
# checked against the template's text in tests/test_gateway_nginx_table.py.
def nginx_upstream_for(models: list[str], body: bytes | None) -> str:
    """access_by_lua_block, model-gateway.yaml:53-75: the upstream a request goes to.  `default_backend` is the first
    model's server (:20-22); a body that is absent, not JSON, not an object, without `model`, or naming an unknown
    model stays on it (cjson.decode under pcall: errors leave the default)."""
    import json
    upstream = "default_backend"
    if body:
        try:
            data = json.loads(body)
        except Exception:
            return upstream
        # Lua: `json_data and json_data.model` -- indexing a non-table raises inside the block only for non-indexable
        # values; cjson arrays are tables without a `model` key.  Only an object with a truthy `model` matches.
        if isinstance(data, dict) and data.get("model"):
            name = data["model"]
            for m in models:                      # one `if model_name == "<m>"` per model, in values order (:66-70)
                if name == m:
                    upstream = "vllm-" + m
    return upstream


def nginx_backend_host(models: list[str], upstream: str) -> str:
    """upstream blocks, model-gateway.yaml:14-22: every upstream is one server `vllm-<model>:8080`."""
    if upstream == "default_backend":
        return f"vllm-{models[0]}:8080"
    assert upstream.startswith("vllm-") and upstream[5:] in models
    return f"{upstream}:8080"


NGINX_FORWARDED_HEADERS = ("Host", "X-Real-IP", "X-Forwarded-For", "X-Forwarded-Proto")   # proxy_set_header, :78-81
