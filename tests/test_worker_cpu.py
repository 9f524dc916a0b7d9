"""The NATS worker wiring (csrc/worker.cpp; /root/reference/README.md:475-502) on CPU: a stub broker speaking the NATS
core protocol (tools/nats_stub.py), the C++ worker queue-subscribed to the lmstudio.* subjects, and a request/reply
client — the shape of the reference's own client example (README.md:506-562).  No GPU: list_models and the validation
paths of the other handlers answer without touching the device; chat generation through NATS is in test_gpu_worker.py."""
import json
import threading
import time

import pytest

import nats_stub
import nats_llm_studio_b200 as b200


@pytest.fixture()
def served():
    broker = nats_stub.Broker().start()
    rc = {}
    t = threading.Thread(target=lambda: rc.setdefault("rc", b200.Engine.worker_run(broker.url, "lmstudio-workers", 4)), daemon=True)
    t.start()
    client = nats_stub.Client(broker.url)
    for _ in range(200):                                   # wait until the worker's subscriptions are in place
        with broker.lock:
            if len(broker.subs.get("lmstudio.chat_model", [])) == 1:
                break
        time.sleep(0.01)
    yield broker, client
    client.close()
    assert b200.Engine.worker_stop() == 0
    t.join(timeout=10)
    assert rc.get("rc") == 0
    broker.stop()


def test_subjects_answer_with_the_reference_envelopes(served):
    broker, client = served
    with broker.lock:
        for s in ("lmstudio.list_models", "lmstudio.pull_model", "lmstudio.delete_model", "lmstudio.chat_model"):
            (conn, sid, queue), = broker.subs[s]
            assert queue == "lmstudio-workers"            # QueueSubscribe with NATS_QUEUE_GROUP's default (README.md:494)
    assert client.request("lmstudio.list_models", b"") == b'{"ok":true,"data":{"http_status":200,"models":{"data":[],"object":"list"}}}'
    assert client.request("lmstudio.chat_model", b"") == b'{"ok":false,"error":"payload vazio em ChatModel","data":null}'
    assert client.request("lmstudio.chat_model", b'{"messages":[]}') == b'{"ok":false,"error":"\'model\' is required in ChatModel","data":null}'
    assert client.request("lmstudio.pull_model", b'{}') == b'{"ok":false,"error":"\'identifier\' is required","data":null}'
    r = json.loads(client.request("lmstudio.delete_model", b'{"model_id":"nobody/none"}'))
    assert r == {"ok": False, "error": "model directory not found: hbm://nobody/none", "data": {"dir": "hbm://nobody/none", "model_id": "nobody/none"}}
    r = json.loads(client.request("lmstudio.chat_model", b'{"model":"nobody/none","messages":[]}'))
    assert r["ok"] and r["data"]["http_status"] == 404      # the backend's answer is forwarded, status unchecked (nats_llm_studio.go:173-178)


def test_concurrent_requests_are_all_answered(served):
    broker, client = served
    out = [None] * 48
    def work(i):
        out[i] = client.request("lmstudio.list_models" if i % 2 else "lmstudio.chat_model", b"" if i % 2 else b'{"model":""}')
    th = [threading.Thread(target=work, args=(i,)) for i in range(48)]
    [t.start() for t in th]; [t.join(timeout=30) for t in th]
    for i, o in enumerate(out):
        assert o is not None
        assert json.loads(o)["ok"] == bool(i % 2)


def test_worker_reports_an_unreachable_server():
    assert b200.Engine.worker_run("nats://127.0.0.1:1", None, 1) < 0
    assert "NATS" in b200.lib().b200_last_error().decode()


def test_queue_group_delivers_to_one_member():
    """broker semantics the deployment relies on (README.md:484): one member of a queue group gets each request"""
    broker = nats_stub.Broker().start()
    got = {"a": 0, "b": 0}
    import socket
    def member(name):
        s = socket.create_connection(("127.0.0.1", broker.port)); f = s.makefile("rwb")
        f.readline(); f.write(b"CONNECT {}\r\nSUB work grp 1\r\n"); f.flush()
        while True:
            line = f.readline()
            if not line:
                return
            if line.startswith(b"MSG"):
                p = line.split(); n = int(p[-1]); f.read(n + 2); got[name] += 1
                f.write(b"PUB " + p[3] + b" 2\r\nok\r\n"); f.flush()
    for n in "ab":
        threading.Thread(target=member, args=(n,), daemon=True).start()
    time.sleep(0.2)
    c = nats_stub.Client(broker.url)
    for _ in range(20):
        assert c.request("work", b"x", timeout=5) == b"ok"
    assert got["a"] + got["b"] == 20 and got["a"] > 0 and got["b"] > 0
    c.close(); broker.stop()


def test_sync_from_bucket_reads_the_object_store(served, golden_dir):
    """lmstudio.sync_model_from_bucket (README.md:286-318) against the stub's Object Store: validation, a missing object,
    and — without a GPU — the transfer itself up to the point where HBM is needed (metadata record, chunk consumer and the
    GGUF header all arrive; the engine then refuses for lack of a device, loudly, as ok:false)."""
    import os
    broker, client = served
    r = json.loads(client.request("lmstudio.sync_model_from_bucket", b"{}"))
    assert r == {"ok": False, "error": "'bucket' and 'object_name' are required", "data": None}
    r = json.loads(client.request("lmstudio.sync_model_from_bucket", b"nonsense"))
    assert not r["ok"] and r["error"].startswith("invalid JSON in SyncModelFromBucket")
    req = {"bucket": "llm-models", "object_name": "acme/tiny-llama/model.gguf", "publisher": "acme", "model_dir": "tiny-llama"}
    r = json.loads(client.request("lmstudio.sync_model_from_bucket", json.dumps(req).encode()))
    assert not r["ok"] and "object not found" in r["error"] and r["data"]["model"] == "acme/tiny-llama"
    broker.chunk = 100_000
    broker.put_object("llm-models", req["object_name"], open(os.path.join(golden_dir, "tiny-llama.gguf"), "rb").read())
    r = json.loads(client.request("lmstudio.sync_model_from_bucket", json.dumps(req).encode()))
    import torch
    if not torch.cuda.is_available():
        assert not r["ok"] and "CUDA" in r["error"], r
    else:
        assert r["ok"], r
