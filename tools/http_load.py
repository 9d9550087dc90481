"""Concurrent-request throughput of one llama-server replica (one GPU of it, direct to the replica).
Usage: python tools/http_load.py [concurrency ...]   -- starts bin/llama-server on the synthetic Llama-3.1-8B Q4_K_M GGUF,
then for every concurrency level sends that many /v1/completions requests at once (128 token ids in, 128 new tokens,
greedy, EOS ignored) and prints the aggregate generated tok/s and the request latency.
B2_LOAD_GATEWAY=ours puts deploy/gateway.py in front of the replica, =reference the restatement of the chart's own
single-threaded gateway script (tests/gateway_restated.py): SURVEY 8d config 4 asks for both paths."""
import http.client, importlib, json, os, socket, subprocess, sys, threading, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
b2 = importlib.import_module("llms-on-kubernetes_b200")
synth = importlib.import_module("llms-on-kubernetes_b200.synth_gguf")
levels = [int(a) for a in sys.argv[1:]] or [1, 8, 32, 64]
name = os.environ.get("B2_LOAD_MODEL", "llama-3.1-8b-q4_k_m")
path = os.path.join(os.environ.get("B2_BENCH_DIR", "/tmp"), f"b2_synth_{name}_s0.gguf")
if not os.path.exists(path):
    synth.write_model(path + ".tmp", name, seed=0); os.replace(path + ".tmp", path)
s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
env = dict(os.environ, B2_CTX_SIZE="512", B2_N_SLOTS=os.environ.get("B2_N_SLOTS", "32"))
proc = subprocess.Popen([b2.SERVER_PATH, "--host", "127.0.0.1", "--port", str(port), "--model", path, "--alias", "m"], env=env,
                        stderr=subprocess.DEVNULL)
try:
    t0 = time.time()
    while True:
        try:
            c = http.client.HTTPConnection("127.0.0.1", port, timeout=2); c.request("GET", "/health"); r = c.getresponse(); r.read(); c.close()
            if r.status == 200: break
        except OSError:
            pass
        assert proc.poll() is None and time.time() - t0 < 600
        time.sleep(0.2)
    via = os.environ.get("B2_LOAD_GATEWAY", "")
    target = port
    gw_proc = None
    if via == "ours":
        s2 = socket.socket(); s2.bind(("127.0.0.1", 0)); target = s2.getsockname()[1]; s2.close()
        gw_proc = subprocess.Popen([sys.executable, os.path.join(ROOT, "deploy", "gateway.py")],
                                   env=dict(os.environ, B2_GATEWAY_BACKENDS=json.dumps({"m": f"http://127.0.0.1:{port}"}), B2_GATEWAY_PORT=str(target)),
                                   stderr=subprocess.DEVNULL)
        time.sleep(1.0)
    elif via == "reference":
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import gateway_restated
        gsrv, target = gateway_restated.make_gateway({"m": f"http://127.0.0.1:{port}"})
        threading.Thread(target=gsrv.serve_forever, daemon=True).start()
    vocab = synth.SPECS[name].vocab
    rng = np.random.default_rng(0)

    def one(out, i, req):
        t = time.time()
        c = http.client.HTTPConnection("127.0.0.1", target, timeout=600)
        c.request("POST", "/v1/completions", body=json.dumps(req), headers={"Content-Type": "application/json"})
        r = c.getresponse(); d = json.loads(r.read()); c.close()
        out[i] = (r.status, d["usage"]["completion_tokens"] if r.status == 200 else 0, time.time() - t)

    for n in levels:
        reqs = [{"model": "m", "prompt": rng.integers(3, vocab, size=128).tolist(), "max_tokens": 128, "temperature": 0, "ignore_eos": True}
                for _ in range(n)]
        out = [None] * n
        ts = [threading.Thread(target=one, args=(out, i, reqs[i])) for i in range(n)]
        t = time.time()
        for th in ts: th.start()
        for th in ts: th.join()
        dt = time.time() - t
        ok = sum(1 for o in out if o[0] == 200)
        toks = sum(o[1] for o in out)
        lat = sorted(o[2] for o in out)
        print(json.dumps({"via": via or "direct", "concurrency": n, "ok": ok, "generated_tokens": toks, "seconds": round(dt, 3), "tok_per_s": round(toks / dt, 1),
                          "latency_p50_s": round(lat[len(lat) // 2], 3), "latency_max_s": round(lat[-1], 3)}), flush=True)
finally:
    if 'gw_proc' in dir() and gw_proc is not None:
        gw_proc.terminate()
    proc.terminate()
    proc.wait(timeout=20)
