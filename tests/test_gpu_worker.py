"""lmstudio.chat_model end to end over the wire (BASELINE.json configs[2]: concurrent chat requests through a NATS queue
group, continuous batching): stub broker (tools/nats_stub.py) <- C++ worker (csrc/worker.cpp) <- engine on the GPU.
64 concurrent requests; every reply equals the envelope the same request gets through b200_handle when it is alone."""
import json
import os
import threading
import time

import pytest

import nats_stub
import nats_llm_studio_b200 as b200
from nats_llm_studio_b200.engine import lib

pytestmark = pytest.mark.gpu


def _strip(rep):
    """the parts of a chat reply that do not depend on wall-clock time or on a counter"""
    r = json.loads(rep); resp = r["data"]["response"]
    return r["ok"], r["data"]["http_status"], resp["choices"], resp["usage"], resp["model"]


def test_64_concurrent_chats_over_nats(golden_dir):
    lib().b200_shutdown()
    eng = b200.Engine(max_ctx=1024, max_seqs=16)
    m = eng.load_file("nats/tiny", os.path.join(golden_dir, "tiny-llama.gguf"))
    reqs = [json.dumps({"model": "nats/tiny", "messages": [{"role": "system", "content": "s" * (i % 5)}, {"role": "user", "content": ("w%d " % i) * (2 + i % 9)}],
                        "temperature": 0 if i % 3 else 0.7, "seed": 40 + i, "max_tokens": 6 + i % 5, "ignore_eos": True}).encode() for i in range(64)]
    solo = [_strip(eng.handle("lmstudio.chat_model", r)) for r in reqs]
    broker = nats_stub.Broker().start()
    rc = {}
    t = threading.Thread(target=lambda: rc.setdefault("rc", b200.Engine.worker_run(broker.url, None, 0)), daemon=True); t.start()
    for _ in range(300):
        with broker.lock:
            if broker.subs.get("lmstudio.chat_model"):
                break
        time.sleep(0.01)
    client = nats_stub.Client(broker.url)
    models = json.loads(client.request("lmstudio.list_models", b""))
    assert [e["id"] for e in models["data"]["models"]["data"]] == ["nats/tiny"]
    out = [None] * 64
    def work(i):
        out[i] = client.request("lmstudio.chat_model", reqs[i], timeout=120)
    th = [threading.Thread(target=work, args=(i,)) for i in range(64)]
    l0 = eng.launch_count()
    [x.start() for x in th]; [x.join(timeout=180) for x in th]
    launches = eng.launch_count() - l0
    for i in range(64):
        assert out[i] is not None, i
        assert _strip(out[i]) == solo[i], i
    rep = json.loads(client.request("lmstudio.delete_model", json.dumps({"model_id": "nats/tiny"}).encode()))
    assert rep["ok"] and rep["data"] == {"deleted_dir": "hbm://nats/tiny", "model_id": "nats/tiny"}
    assert json.loads(client.request("lmstudio.list_models", b""))["data"]["models"]["data"] == []
    client.close()
    assert b200.Engine.worker_stop() == 0
    t.join(timeout=20); assert rc.get("rc") == 0
    broker.stop()
    m.h = None
    lib().b200_shutdown()
    assert launches > 0


def test_sync_model_from_bucket_into_hbm(golden_dir):
    """Config 5's path at test size: a .gguf stored as Object Store chunks -> worker -> pinned host buffers -> HBM -> on-GPU
    repack; the synced model answers chat requests exactly like the same file loaded from disk; a corrupted digest is refused."""
    lib().b200_shutdown()
    eng = b200.Engine(max_ctx=512)
    path = os.path.join(golden_dir, "tiny-llama.gguf")
    data = open(path, "rb").read()
    broker = nats_stub.Broker(chunk=128 * 1024).start()     # nats.go's default object chunk size
    meta = broker.put_object("llm-models", "acme/tiny-llama/model.gguf", data)
    broker.put_object("llm-models", "acme/bad/model.gguf", data, corrupt_digest=True)
    t = threading.Thread(target=lambda: b200.Engine.worker_run(broker.url, None, 4), daemon=True); t.start()
    for _ in range(300):
        with broker.lock:
            if broker.subs.get("lmstudio.sync_model_from_bucket"):
                break
        time.sleep(0.01)
    client = nats_stub.Client(broker.url)
    r = json.loads(client.request("lmstudio.sync_model_from_bucket", json.dumps({"bucket": "llm-models", "object_name": "acme/tiny-llama/model.gguf"}).encode()))
    assert r["ok"], r
    assert r["data"]["local_path"] == "hbm://acme/tiny-llama" and r["data"]["bytes"] == len(data) and r["data"]["chunks"] == meta["chunks"]
    assert r["data"]["digest"].rstrip("=") == meta["digest"].rstrip("=")
    ref = eng.load_file("disk/tiny", path)
    chat = lambda mid: json.loads(client.request("lmstudio.chat_model", json.dumps({"model": mid, "messages": [{"role": "user", "content": "compare me"}], "temperature": 0, "max_tokens": 12, "ignore_eos": True}).encode()))
    a, b = chat("acme/tiny-llama"), chat("disk/tiny")
    assert a["ok"] and a["data"]["response"]["choices"] == b["data"]["response"]["choices"]
    bad = json.loads(client.request("lmstudio.sync_model_from_bucket", json.dumps({"bucket": "llm-models", "object_name": "acme/bad/model.gguf"}).encode()))
    assert not bad["ok"] and "digest mismatch" in bad["error"]
    ids = [e["id"] for e in json.loads(client.request("lmstudio.list_models", b""))["data"]["models"]["data"]]
    assert "acme/tiny-llama" in ids and "acme/bad" not in ids
    client.close(); b200.Engine.worker_stop(); t.join(timeout=20); broker.stop()
    ref.close()
    lib().b200_shutdown()
