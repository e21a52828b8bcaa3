"""The HTTP front end (server/b2_server.cpp): the process the operator's pod template runs (`serve`, OLLAMA_HOST, port
11434, probes on GET /api/tags -- reference pkg/model/pod.go:14-66).  The probe endpoints are exercised without a GPU;
generation (GPU) is checked token for token against the oracle through the ids the API itself returns."""
import hashlib
import json
import os
import socket
import subprocess
import time
import urllib.request

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "server", "b2-server")


def _store(tmp, gguf_path, name="phi"):
    root = os.path.join(tmp, "models")
    data = open(gguf_path, "rb").read()
    os.makedirs(os.path.join(root, "blobs"), exist_ok=True)
    layers = []
    for mt, blob in (("application/vnd.ollama.image.model", data), ("application/vnd.ollama.image.template", b"{{ if .System }}{{ .System }} {{ end }}{{ .Prompt }}"),
                     ("application/vnd.ollama.image.params", json.dumps({"stop": ["<never>"]}).encode())):
        d = "sha256:" + hashlib.sha256(blob).hexdigest()
        with open(os.path.join(root, "blobs", d.replace(":", "-")), "wb") as f:
            f.write(blob)
        layers.append({"mediaType": mt, "digest": d, "size": len(blob)})
    mdir = os.path.join(root, "manifests", "registry.ollama.ai", "library", name)
    os.makedirs(mdir, exist_ok=True)
    with open(os.path.join(mdir, "latest"), "w") as f:
        json.dump({"schemaVersion": 2, "layers": layers}, f)
    return tmp


class Server:
    def __init__(self, home):
        s = socket.socket(); s.bind(("127.0.0.1", 0)); self.port = s.getsockname()[1]; s.close()
        env = dict(os.environ, OLLAMA_HOST=f"127.0.0.1:{self.port}", OLLAMA_MODELS=home, B2_MAX_BATCH="4", B2_MAX_CTX="256")
        self.p = subprocess.Popen([BIN, "serve"], env=env, stderr=subprocess.PIPE)
        for _ in range(100):
            try:
                socket.create_connection(("127.0.0.1", self.port), timeout=0.2).close(); break
            except OSError:
                time.sleep(0.05)

    def get(self, path):
        with urllib.request.urlopen(f"http://127.0.0.1:{self.port}{path}", timeout=5) as r:
            return r.status, r.read()

    def post(self, path, body, timeout=120):
        req = urllib.request.Request(f"http://127.0.0.1:{self.port}{path}", data=json.dumps(body).encode(), headers={"Content-Type": "application/json"})
        try:
            with urllib.request.urlopen(req, timeout=timeout) as r:
                return r.status, r.read()
        except urllib.error.HTTPError as e:
            return e.code, e.read()

    def close(self):
        self.p.terminate()
        try:
            self.p.wait(timeout=10)
        except subprocess.TimeoutExpired:
            self.p.kill()


@pytest.fixture()
def server(pkg, model_dir, tmp_path):
    if not os.path.exists(BIN):
        pytest.skip("server/b2-server is not built (./build.sh)")
    home = _store(str(tmp_path), pkg.synth.cached_model("tiny-llama", model_dir))
    s = Server(home)
    yield s
    s.close()


def test_probe_endpoints_answer_without_a_gpu(server):
    code, body = server.get("/")
    assert code == 200 and body == b"Ollama is running"
    code, body = server.get("/api/tags")                      # the readiness and liveness probe of the pod template
    tags = json.loads(body)
    assert code == 200 and [m["name"] for m in tags["models"]] == ["phi:latest"] and tags["models"][0]["size"] > 0
    code, body = server.get("/api/version")
    assert code == 200 and "version" in json.loads(body)
    code, body = server.post("/api/show", {"name": "phi"})
    show = json.loads(body)
    assert code == 200 and "{{ .Prompt }}" in show["template"] and show["details"]["format"] == "gguf"
    code, body = server.post("/api/show", {"name": "nosuchmodel"})
    assert code == 404 and "error" in json.loads(body)
    code, body = server.post("/api/pull", {"name": "phi"})
    assert code == 405 and "read-only" in json.loads(body)["error"]
    code, body = server.post("/api/generate", {"prompt": "no model named"})
    assert code == 400


@pytest.mark.gpu
def test_generate_streams_the_oracles_tokens(server, pkg, oracle, model_dir):
    path = pkg.synth.cached_model("tiny-llama", model_dir)
    code, body = server.post("/api/generate", {"model": "phi", "prompt": "Hello B200", "raw": True, "stream": True,
                                               "options": {"temperature": 0, "num_predict": 24}})
    assert code == 200
    frames = [json.loads(l) for l in body.decode().splitlines() if l.strip()]
    assert frames[-1]["done"] is True and all(not f["done"] for f in frames[:-1])
    last = frames[-1]
    assert last["eval_count"] > 0 and last["prompt_eval_count"] > 0 and last["eval_duration"] > 0 and last["prompt_eval_duration"] > 0
    ctx = last["context"]
    n_prompt = last["prompt_eval_count"]
    assert ctx[0] == 1                                                   # BOS from the file's tokenizer defaults
    assert bytes(t - 3 for t in ctx[1:n_prompt]) == "\u2581Hello\u2581B200".encode()      # byte-fallback pieces (the vocabulary has no word pieces for it)
    om = oracle.OracleModel(path, faithful=2)
    want, _ = om.generate(np.array(ctx[:n_prompt], np.int32), len(ctx) - n_prompt)
    assert ctx[n_prompt:] == [int(t) for t in want][: len(ctx) - n_prompt]
    # the same request again, not streamed, and through the chat and OpenAI surfaces
    code, body = server.post("/api/generate", {"model": "phi", "prompt": "Hello B200", "raw": True, "stream": False, "options": {"temperature": 0, "num_predict": 24}})
    whole = json.loads(body)
    assert code == 200 and whole["done"] and whole["context"] == ctx and whole["response"] == "".join(f["response"] for f in frames)
    code, body = server.post("/api/chat", {"model": "phi", "messages": [{"role": "user", "content": "Hi"}], "stream": False, "options": {"temperature": 0, "num_predict": 8}})
    chat = json.loads(body)
    assert code == 200 and chat["done"] and chat["message"]["role"] == "assistant" and chat["eval_count"] > 0
    code, body = server.post("/v1/chat/completions", {"model": "phi", "messages": [{"role": "user", "content": "Hi"}], "max_tokens": 8, "temperature": 0})
    oa = json.loads(body)
    assert code == 200 and oa["choices"][0]["message"]["role"] == "assistant" and oa["usage"]["completion_tokens"] > 0


@pytest.mark.gpu
def test_concurrent_requests_share_decode_steps(server):
    """Four clients at once: the engine thread admits them into sequence slots and advances them together; every one
    must get what it gets alone (greedy, same prompt -> same tokens)."""
    import threading
    body = {"model": "phi", "prompt": "batch me", "raw": True, "stream": False, "options": {"temperature": 0, "num_predict": 16}}
    alone = json.loads(server.post("/api/generate", body)[1])["context"]
    out = [None] * 4

    def run(i):
        out[i] = json.loads(server.post("/api/generate", body)[1])["context"]
    th = [threading.Thread(target=run, args=(i,)) for i in range(4)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert all(o == alone for o in out)
