"""Optional replacement for the API gateway script the ramalama chart ships in its ConfigMap.

The reference's gateway (ramalama-models/helm-chart/templates/api-gateway.yaml:9-111) routes by the JSON `model` field,
answers /v1/models and /health itself -- and serves ONE request at a time on a single thread, buffers every response
whole (no SSE reaches the client before the generation ends) and turns every backend error into 502.  With replicas that
batch hundreds of sequences that one thread is the bottleneck of the deployment.  This script keeps the contract and
removes the three limits:

  * same routing: BACKENDS maps modelName -> base URL in chart order, the first entry is the default; an unparsable
    body or an unknown / missing `model` goes to the default backend; GET /v1/models lists the names with created 0 /
    owned_by "ramalama"; GET /health -> 200 "OK"; every other path and method is proxied;
  * one thread per connection (ThreadingHTTPServer), so requests to the replicas overlap;
  * the backend's status code and headers pass through (a 404 stays a 404; only a backend that cannot be reached is 502);
  * the response body is relayed as it arrives: a `text/event-stream` answer is forwarded chunk by chunk, everything
    else with its Content-Length.

Configuration, in this order: the BACKENDS literal below (the chart's template can fill it exactly like the
reference's, `{{- range .Values.models }}`), else the environment: B2_GATEWAY_BACKENDS='{"name": "http://host:8080", ...}'
(JSON object, order kept), B2_GATEWAY_PORT (default 8080).  Python standard library only.
"""
import http.client
import http.server
import json
import os
import sys
import urllib.parse

BACKENDS = {}   # filled by the chart template, or from B2_GATEWAY_BACKENDS
HOP_BY_HOP = {"connection", "keep-alive", "proxy-authenticate", "proxy-authorization", "te", "trailers", "transfer-encoding", "upgrade", "host"}


def load_backends():
    if BACKENDS:
        return dict(BACKENDS)
    raw = os.environ.get("B2_GATEWAY_BACKENDS", "")
    table = json.loads(raw) if raw else {}
    if not isinstance(table, dict) or not table:
        raise SystemExit("gateway: no backends configured (BACKENDS literal or B2_GATEWAY_BACKENDS)")
    return table


def pick_backend(backends, default, body):
    """The reference's routing rule (api-gateway.yaml:67-77; the nginx gateway's Lua block, model-gateway.yaml:53-75, is
    the same rule): the backend named by the JSON body's `model`, else the default; a body that is not a JSON object, or
    whose `model` is not a known name, stays on the default."""
    if body:
        try:
            doc = json.loads(body.decode())
        except (ValueError, UnicodeDecodeError):
            return default
        if isinstance(doc, dict):
            name = doc.get("model")
            if isinstance(name, str) and name in backends:
                return backends[name]
    return default


def make_handler(backends):
    names = list(backends)
    default = backends[names[0]]

    class Handler(http.server.BaseHTTPRequestHandler):
        protocol_version = "HTTP/1.1"

        def log_message(self, *a):
            pass

        def do_GET(self):
            if self.path == "/v1/models":
                body = json.dumps({"object": "list", "data": [{"id": n, "object": "model", "created": 0, "owned_by": "ramalama"} for n in names]}).encode()
                return self.reply(200, "application/json", body)
            if self.path == "/health":
                return self.reply(200, "text/plain", b"OK")
            self.forward()

        def do_POST(self):
            self.forward()

        do_PUT = do_DELETE = do_PATCH = do_OPTIONS = do_POST

        def reply(self, status, ctype, body):
            self.send_response(status)
            self.send_header("Content-Type", ctype)
            self.send_header("Content-Length", str(len(body)))
            self.end_headers()
            self.wfile.write(body)

        def forward(self):
            n = int(self.headers.get("Content-Length") or 0)
            body = self.rfile.read(n) if n > 0 else b""
            target = pick_backend(backends, default, body)
            u = urllib.parse.urlsplit(target)
            headers = {k: v for k, v in self.headers.items() if k.lower() not in HOP_BY_HOP}
            if body:
                headers["Content-Length"] = str(len(body))
                if not any(k.lower() == "content-type" for k in headers):
                    # what the reference's urllib client sends for a body without a type: backends see the same request
                    headers["Content-Type"] = "application/x-www-form-urlencoded"
            try:
                conn = http.client.HTTPConnection(u.hostname, u.port or 80, timeout=300)
                conn.request(self.command, self.path, body=body if body else None, headers=headers)
                resp = conn.getresponse()
            except OSError as e:
                return self.reply(502, "application/json", json.dumps({"error": f"backend unreachable: {e}"}).encode())
            try:
                self.send_response(resp.status)
                length = resp.getheader("Content-Length")
                for k, v in resp.getheaders():
                    if k.lower() not in HOP_BY_HOP and k.lower() != "content-length":
                        self.send_header(k, v)
                if length is not None:
                    # a sized answer: relay it in pieces as they arrive
                    self.send_header("Content-Length", length)
                    self.end_headers()
                    while True:
                        piece = resp.read(65536)
                        if not piece:
                            break
                        self.wfile.write(piece)
                else:
                    # a stream (SSE, or a server that closes to delimit): chunked to 1.1 clients, close-delimited otherwise
                    chunked = self.request_version == "HTTP/1.1"
                    if chunked:
                        self.send_header("Transfer-Encoding", "chunked")
                    else:
                        self.send_header("Connection", "close")
                        self.close_connection = True
                    self.end_headers()
                    while True:
                        piece = resp.read1(65536)
                        if not piece:
                            break
                        self.wfile.write(b"%x\r\n%s\r\n" % (len(piece), piece) if chunked else piece)
                        self.wfile.flush()
                    if chunked:
                        self.wfile.write(b"0\r\n\r\n")
            except (BrokenPipeError, ConnectionResetError):
                self.close_connection = True
            finally:
                conn.close()

    return Handler


class GatewayServer(http.server.ThreadingHTTPServer):
    daemon_threads = True
    request_queue_size = 1024   # hundreds of clients connect at once; the default backlog of 5 refuses most of them


def make_gateway(backends, port=0, host="0.0.0.0"):
    srv = GatewayServer((host, port), make_handler(backends))
    return srv, srv.server_address[1]


if __name__ == "__main__":
    server, p = make_gateway(load_backends(), int(os.environ.get("B2_GATEWAY_PORT", "8080")))
    print(f"gateway: listening on {p}, backends {list(load_backends())}", file=sys.stderr, flush=True)
    server.serve_forever()
