"""SURVEY 8d config 4 on a bare box: N llama-server replicas, one per GPU, behind a gateway that routes by the JSON
`model` field (replica i is model "llama-i", as 8e prescribes), C concurrent /v1/completions requests spread over them
(128 token ids in, 128 new tokens, greedy, EOS ignored).  Prints aggregate generated tok/s through the gateway.

  python tools/replica_load.py N_REPLICAS [concurrency ...]     B2_LOAD_GATEWAY=ours (deploy/gateway.py, default) | reference
"""
import http.client, importlib, json, os, socket, subprocess, sys, threading, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
b2 = importlib.import_module("llms-on-kubernetes_b200")
synth = importlib.import_module("llms-on-kubernetes_b200.synth_gguf")
n_rep = int(sys.argv[1])
levels = [int(a) for a in sys.argv[2:]] or [32 * n_rep]
name = os.environ.get("B2_LOAD_MODEL", "llama-3.1-8b-q4_k_m")
path = os.path.join(os.environ.get("B2_BENCH_DIR", "/tmp"), f"b2_synth_{name}_s0.gguf")
if not os.path.exists(path):
    synth.write_model(path + ".tmp", name, seed=0); os.replace(path + ".tmp", path)


def free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


procs, backends = [], {}
for i in range(n_rep):
    port = free_port()
    env = dict(os.environ, CUDA_VISIBLE_DEVICES=str(i), B2_CTX_SIZE="512", B2_N_SLOTS=os.environ.get("B2_N_SLOTS", "256"))
    procs.append(subprocess.Popen([b2.SERVER_PATH, "--host", "127.0.0.1", "--port", str(port), "--model", path, "--alias", f"llama-{i}"], env=env,
                                  stderr=subprocess.DEVNULL))
    backends[f"llama-{i}"] = f"http://127.0.0.1:{port}"
gw_proc = None
try:
    t0 = time.time()
    for url in backends.values():
        port = int(url.rsplit(":", 1)[1])
        while True:
            try:
                c = http.client.HTTPConnection("127.0.0.1", port, timeout=2); c.request("GET", "/health"); r = c.getresponse(); r.read(); c.close()
                if r.status == 200: break
            except OSError:
                pass
            assert time.time() - t0 < 900
            time.sleep(0.2)
    via = os.environ.get("B2_LOAD_GATEWAY", "ours")
    if via == "ours":
        gport = free_port()
        gw_proc = subprocess.Popen([sys.executable, os.path.join(ROOT, "deploy", "gateway.py")],
                                   env=dict(os.environ, B2_GATEWAY_BACKENDS=json.dumps(backends), B2_GATEWAY_PORT=str(gport)), stderr=subprocess.DEVNULL)
        time.sleep(1.0)
    else:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import gateway_restated
        gsrv, gport = gateway_restated.make_gateway(backends)
        threading.Thread(target=gsrv.serve_forever, daemon=True).start()
    vocab = synth.SPECS[name].vocab
    rng = np.random.default_rng(0)

    def one(out, i, req):
        t = time.time()
        c = http.client.HTTPConnection("127.0.0.1", gport, timeout=900)
        c.request("POST", "/v1/completions", body=json.dumps(req), headers={"Content-Type": "application/json"})
        r = c.getresponse(); d = json.loads(r.read()); c.close()
        out[i] = (r.status, d["usage"]["completion_tokens"] if r.status == 200 else 0, time.time() - t)

    for n in [min(8, levels[0])] + levels:      # the first, small round warms every replica up (untimed in the report)
        reqs = [{"model": f"llama-{i % n_rep}", "prompt": rng.integers(3, vocab, size=128).tolist(), "max_tokens": 128, "temperature": 0,
                 "ignore_eos": True} for i in range(n)]
        out = [(0, 0, 0.0)] * n
        ts = [threading.Thread(target=one, args=(out, i, reqs[i])) for i in range(n)]
        t = time.time()
        for th in ts: th.start()
        for th in ts: th.join()
        dt = time.time() - t
        lat = sorted(o[2] for o in out)
        print(json.dumps({"replicas": n_rep, "via": via, "concurrency": n, "ok": sum(1 for o in out if o[0] == 200), "generated_tokens": sum(o[1] for o in out),
                          "seconds": round(dt, 3), "tok_per_s": round(sum(o[1] for o in out) / dt, 1), "latency_p50_s": round(lat[len(lat) // 2], 3)}),
              flush=True)
finally:
    if gw_proc is not None:
        gw_proc.terminate()
    for p in procs:
        p.terminate()
    for p in procs:
        p.wait(timeout=30)
