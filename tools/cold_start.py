#!/usr/bin/env python
"""Config 5: cold start.  A GGUF arrives as JetStream-Object-Store-sized chunks (default 128 KiB), is
copied into pinned host buffers and streamed into HBM through b200_model_open / b200_model_upload / b200_model_commit
(on-GPU repack), then the first token of a chat request is produced.  Prints one JSON line: seconds from the first chunk
to the first generated token, split into stream / commit / first token.  The chunk source is a file in /dev/shm (there is
no NATS server on the box); what is measured is everything after the chunks reach host memory."""
import argparse, ctypes as C, json, mmap, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import make_gguf  # noqa: E402
import nats_llm_studio_b200 as b200  # noqa: E402
from nats_llm_studio_b200.engine import lib, _check  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--preset", default="llama3-70b"); ap.add_argument("--mix", default="q4_k_m"); ap.add_argument("--n-layer", type=int, default=0)
ap.add_argument("--chunk", type=int, default=128 << 10); ap.add_argument("--batch", type=int, default=32 << 20, help="bytes handed to one b200_model_upload call")
ap.add_argument("--prompt", type=int, default=16)
a = ap.parse_args()
d = os.environ.get("B200_BENCH_DIR") or "/dev/shm"
path = os.path.join(d, f"b200tp-{a.preset}-{a.mix}" + (f"-L{a.n_layer}" if a.n_layer else "") + ".gguf")
if not os.path.exists(path):
    over = dict(mix=a.mix, seed=2, reuse_layers=True)
    if a.n_layer:
        over["n_layer"] = a.n_layer
    make_gguf.write_model(path + ".tmp", make_gguf.spec_from_preset(a.preset, **over)); os.replace(path + ".tmp", path)
size = os.path.getsize(path)
eng = b200.Engine(max_ctx=2048)
L = lib()
pin = [L.b200_pinned_alloc(a.batch) for _ in range(2)]
assert all(pin), "pinned allocation failed"
views = [np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(a.batch,)) for p in pin]
f = open(path, "rb"); mm = mmap.mmap(f.fileno(), 0, access=mmap.ACCESS_READ); src = np.frombuffer(mm, dtype=np.uint8)
_ = int(src[:: 1 << 20].sum())                     # page the file in: the clock starts when chunks are in host memory

t0 = time.perf_counter()
h = C.c_void_p(); n = 1 << 20
while True:
    rc = L.b200_model_open(b"cold/start", src[:n].tobytes(), n, C.byref(h))
    if rc == -6 and n < size:
        n = min(size, n * 4); continue
    _check(rc); break
t_open = time.perf_counter() - t0
off = 0; i = 0; n_chunks = 0
while off < size:
    nb = min(a.batch, size - off); v = views[i & 1]
    for c0 in range(0, nb, a.chunk):               # object-store chunks land in the pinned buffer one by one
        c1 = min(nb, c0 + a.chunk); v[c0:c1] = src[off + c0:off + c1]; n_chunks += 1
    _check(L.b200_model_upload(h, off, pin[i & 1], nb))
    off += nb; i += 1
t_stream = time.perf_counter() - t0
_check(L.b200_model_commit(h))
t_commit = time.perf_counter() - t0
m = b200.Model(h, eng)
prompt = np.random.default_rng(0).integers(0, m.n_vocab, a.prompt).astype(np.int32)
toks, ttft_ms, _ = m.generate(prompt, 1)
t_first = time.perf_counter() - t0
dec = None
try:
    ms, _ = m.bench_decode(512, 32, 3)
    dec = round(1e3 / ms, 1)
except Exception as ex:                            # reported, not fatal
    dec = str(ex)
print(json.dumps({"workload": f"cold start {a.preset} {a.mix}" + (f" L{a.n_layer}" if a.n_layer else ""), "file_bytes": size, "chunk_bytes": a.chunk, "chunks": n_chunks,
                  "upload_call_bytes": a.batch, "open_s": round(t_open, 3), "stream_s": round(t_stream, 3), "stream_gbs": round(size / t_stream / 1e9, 2),
                  "commit_s": round(t_commit - t_stream, 3), "first_token_s": round(t_first - t_commit, 3), "ttft_from_first_chunk_s": round(t_first, 3),
                  "bs1_decode_tok_s_ctx512": dec}))
m.close()
