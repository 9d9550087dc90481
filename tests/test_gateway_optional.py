"""deploy/gateway.py -- the optional, concurrent replacement for the chart's gateway script: the same
routing contract as the reference's (checked case by case against the restatement of the reference script), plus what it
adds: overlapping requests, status pass-through, streamed responses.  CPU only."""
import http.client
import http.server
import importlib.util
import json
import os
import threading
import time

import gateway_restated
import test_gateway_contract as ref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("b2_gateway", os.path.join(ROOT, "deploy", "gateway.py"))
gw = importlib.util.module_from_spec(spec)
spec.loader.exec_module(gw)


def serve(server):
    threading.Thread(target=server.serve_forever, daemon=True).start()


def test_same_routing_contract_as_the_reference_gateway():
    servers = {n: ref.start_stub(n) for n in ref.MODELS}
    backends = {n: f"http://127.0.0.1:{p}" for n, (_, p) in servers.items()}
    theirs, tport = gateway_restated.make_gateway(backends)
    ours, oport = gw.make_gateway(backends, host="127.0.0.1")
    serve(theirs); serve(ours)
    try:
        a, b = ref.run_cases(tport), ref.run_cases(oport)
        for case, x, y in zip(ref.CASES, a, b):
            if case[1] == "/boom":
                # the one designed difference: the backend's own status and body reach the client
                assert x["status"] == 502 and y["status"] == 500 and y["payload"]["error"]["message"] == "kaput"
            else:
                assert x["status"] == y["status"] and x["payload"] == y["payload"], case
                assert (x["ctype"] or "").lower() == (y["ctype"] or "").lower(), case
            assert x["seen"] == y["seen"], case      # what reached the backend: method, path, body, forwarded headers
    finally:
        theirs.shutdown(); ours.shutdown()
        for s, _ in servers.values():
            s.shutdown()


class SlowBackend(http.server.BaseHTTPRequestHandler):
    protocol_version = "HTTP/1.1"

    def do_POST(self):
        n = int(self.headers.get("Content-Length", 0))
        body = json.loads(self.rfile.read(n) or b"{}")
        if body.get("stream"):
            self.send_response(200)
            self.send_header("Content-Type", "text/event-stream")
            self.send_header("Connection", "close")
            self.end_headers()
            for i in range(4):
                self.wfile.write(f"data: {{\"i\": {i}}}\n\n".encode()); self.wfile.flush()
                time.sleep(0.15)
            self.wfile.write(b"data: [DONE]\n\n")
            self.close_connection = True
            return
        time.sleep(0.4)
        out = b'{"ok": true}'
        self.send_response(200)
        self.send_header("Content-Type", "application/json")
        self.send_header("Content-Length", str(len(out)))
        self.end_headers()
        self.wfile.write(out)

    def log_message(self, *a):
        pass


def test_requests_overlap_and_streams_are_relayed_as_they_arrive():
    backend = http.server.ThreadingHTTPServer(("127.0.0.1", 0), SlowBackend)
    serve(backend)
    ours, port = gw.make_gateway({"m": f"http://127.0.0.1:{backend.server_address[1]}"}, host="127.0.0.1")
    serve(ours)
    try:
        # four 0.4 s requests at once: a single-threaded gateway needs 1.6 s
        t0 = time.time()
        res = [None] * 4
        ts = [threading.Thread(target=lambda i=i: res.__setitem__(i, ref.call(port, "POST", "/v1/completions", b'{"model": "m"}',
                                                                                   {"Content-Type": "application/json"})[0])) for i in range(4)]
        for t in ts: t.start()
        for t in ts: t.join()
        assert res == [200] * 4 and time.time() - t0 < 1.2
        # SSE: the first event reaches the client long before the last one is produced
        c = http.client.HTTPConnection("127.0.0.1", port, timeout=10)
        t0 = time.time()
        c.request("POST", "/v1/chat/completions", body=b'{"model": "m", "stream": true}', headers={"Content-Type": "application/json"})
        r = c.getresponse()
        assert r.status == 200 and r.getheader("Content-Type") == "text/event-stream"
        first = r.read1(4096)
        t_first = time.time() - t0
        rest = r.read()
        t_all = time.time() - t0
        c.close()
        assert first.startswith(b"data: {\"i\": 0}") and rest.rstrip().endswith(b"data: [DONE]")
        assert t_first < 0.3 and t_all > 0.5
    finally:
        ours.shutdown(); backend.shutdown()


def test_unreachable_backend_is_a_502():
    ours, port = gw.make_gateway({"m": f"http://127.0.0.1:{ref.free_port()}"}, host="127.0.0.1")
    serve(ours)
    try:
        status, _, data = ref.call(port, "POST", "/v1/completions", b'{"model": "m"}', {"Content-Type": "application/json"})
        assert status == 502 and "error" in json.loads(data)
    finally:
        ours.shutdown()
