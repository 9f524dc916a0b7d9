"""ctypes binding of include/b200_engine.h (mirrors the cgo binding shown in INTEGRATION.md)."""
import ctypes as C
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get("B200_LIB") or os.path.join(HERE, "libb200engine.so")      # B200_LIB: an experimental build variant (tools/)
GGML_TYPES = {"F32": 0, "F16": 1, "Q8_0": 8, "Q4_K": 12, "Q5_K": 13, "Q6_K": 14}
SUBJECTS = ("lmstudio.list_models", "lmstudio.pull_model", "lmstudio.delete_model", "lmstudio.chat_model")

_lib = None


EV_TOKEN, EV_FINISH, EV_ERROR = 1, 2, 3                     # b200_event.kind
FINISH_STOP, FINISH_LENGTH, FINISH_CANCELLED, FINISH_TIMEOUT = 1, 2, 3, 4      # b200_event.finish_reason


class EngineError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"b200 engine error {code}: {msg}")
        self.code = code
        self.msg = msg


class Config(C.Structure):
    _fields_ = [("device", C.c_int), ("max_ctx", C.c_int), ("max_seqs", C.c_int), ("kv_pages", C.c_uint32)]


class Sampling(C.Structure):
    _fields_ = [("temperature", C.c_float), ("top_k", C.c_int), ("top_p", C.c_float), ("seed", C.c_uint64),
                ("max_tokens", C.c_int), ("stop_token", C.c_int), ("ignore_eos", C.c_int), ("timeout_ms", C.c_int)]


class Event(C.Structure):
    _fields_ = [("req_id", C.c_uint64), ("kind", C.c_int), ("token", C.c_int), ("finish_reason", C.c_int),
                ("n_prompt", C.c_int), ("n_generated", C.c_int), ("ttft_ms", C.c_float), ("decode_ms", C.c_float)]


# every symbol include/b200_engine.h declares: name -> (restype, argtypes)
vp, ci, sz = C.c_void_p, C.c_int, C.c_size_t
SYMBOLS = {
    "b200_init": (ci, [C.POINTER(Config)]), "b200_shutdown": (ci, []), "b200_last_error": (C.c_char_p, []),
    "b200_pinned_alloc": (vp, [sz]), "b200_pinned_free": (None, [vp]),
    "b200_model_open": (ci, [C.c_char_p, vp, sz, C.POINTER(vp)]), "b200_model_upload": (ci, [vp, C.c_uint64, vp, sz]),
    "b200_model_commit": (ci, [vp]), "b200_model_close": (ci, [vp]),
    "b200_model_load_file": (ci, [C.c_char_p, C.c_char_p, C.POINTER(vp)]), "b200_model_find": (ci, [C.c_char_p, C.POINTER(vp)]),
    "b200_model_list": (ci, [vp, sz, C.POINTER(sz)]), "b200_model_info": (ci, [vp, vp, sz, C.POINTER(sz)]),
    "b200_submit": (ci, [vp, vp, ci, C.POINTER(Sampling), C.c_uint64]), "b200_poll": (ci, [C.POINTER(Event), ci, ci]),
    "b200_cancel": (ci, [C.c_uint64]),
    "b200_model_generate": (ci, [vp, vp, ci, ci, C.POINTER(Sampling), vp, C.POINTER(ci), C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "b200_model_eval": (ci, [vp, vp, ci, vp, vp]), "b200_model_n_vocab": (ci, [vp]),
    "b200_tokenize": (ci, [vp, C.c_char_p, sz, vp, ci, C.POINTER(ci)]), "b200_detokenize": (ci, [vp, vp, ci, vp, sz, C.POINTER(sz)]),
    "b200_gguf_tokenize": (ci, [vp, sz, C.c_char_p, sz, ci, vp, ci, C.POINTER(ci)]),
    "b200_gguf_detokenize": (ci, [vp, sz, vp, ci, vp, sz, C.POINTER(sz)]),
    "b200_gguf_render_chat": (ci, [vp, sz, C.c_char_p, sz, vp, ci, C.POINTER(ci)]),
    "b200_worker_run": (ci, [C.c_char_p, C.c_char_p, ci]), "b200_worker_stop": (ci, []),
    "b200_handle": (ci, [C.c_char_p, vp, sz, vp, sz, C.POINTER(sz)]), "b200_set_models_dir": (ci, [C.c_char_p]),
    "b200_sample": (ci, [vp, ci, C.POINTER(Sampling), C.c_uint64, C.POINTER(ci)]),
    "b200_dequantize": (ci, [ci, vp, C.c_int64, ci, vp]), "b200_dequantize_noalloc": (ci, [ci, vp, C.c_int64, ci, vp]), "b200_gemv": (ci, [ci, vp, C.c_int64, ci, vp, vp, ci, C.POINTER(C.c_float)]),
    "b200_bench_decode": (ci, [vp, ci, ci, ci, C.POINTER(C.c_float), C.POINTER(C.c_double)]),
    "b200_bench_batch_step": (ci, [vp, ci, ci, ci, ci, C.POINTER(C.c_float), C.POINTER(C.c_double)]),
    "b200_gemm_bf16": (ci, [vp, vp, vp, ci, ci, ci]), "b200_qgemm": (ci, [ci, vp, C.c_int64, ci, vp, ci, vp, ci, C.POINTER(C.c_float)]), "b200_model_prefill_slot": (ci, [vp, ci, vp, ci, ci]), "b200_model_release_slot": (ci, [vp, ci]), "b200_model_kv_free_pages": (ci, [vp]),
    "b200_model_batch_step": (ci, [vp, vp, vp, vp, ci, vp, vp]),
    "b200_model_batch_step_sampled": (ci, [vp, vp, vp, vp, ci, vp, vp, vp, vp]),
    "b200_tp_shard_window": (ci, [C.c_char_p, C.c_int64, ci, ci, ci, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(ci), C.POINTER(ci)]),
    "b200_tp_unique_id": (ci, [vp, sz]), "b200_tp_init": (ci, [ci, ci, vp, sz]), "b200_model_set_prefill_min": (ci, [vp, ci]),
    "b200_launch_count": (C.c_uint64, []), "b200_layout_selftest": (ci, []),
}


def lib():
    """Loads libb200engine.so (building it if the sources are newer).  No fallback of any kind."""
    global _lib
    if _lib is None:
        from . import build as _build
        if not os.environ.get("B200_LIB") and _build.needs_build():
            _build.build()
        L = C.CDLL(SO_PATH)
        for name, (res, args) in SYMBOLS.items():
            f = getattr(L, name)          # AttributeError if the library does not export a declared symbol
            f.restype = res; f.argtypes = args
        _lib = L
    return _lib


def _check(rc):
    if rc < 0:
        raise EngineError(rc, lib().b200_last_error().decode("utf-8", "replace"))
    return rc


def _string_call(fn, *args, cap=1 << 20):
    """Calls fn(*args, buf, cap, &need) once with a generous buffer (handlers EXECUTE the request on every
    call, so a size-probing first call would run it twice); grows and retries only if it did not fit."""
    need = sz(0)
    while True:
        buf = C.create_string_buffer(cap)
        rc = fn(*args, buf, cap, C.byref(need))
        if rc == -6 and need.value >= cap:
            cap = need.value + 1
            continue
        _check(rc)
        return buf.raw[:need.value]


class Model:
    def __init__(self, handle, engine):
        self.h = handle
        self.engine = engine
        self.n_vocab = lib().b200_model_n_vocab(handle)

    def info(self):
        return json.loads(_string_call(lib().b200_model_info, self.h))

    def close(self):
        if self.h:
            _check(lib().b200_model_close(self.h)); self.h = None

    def eval(self, toks, all_logits=True):
        toks = np.ascontiguousarray(toks, dtype=np.int32)
        if all_logits:
            out = np.empty((len(toks), self.n_vocab), dtype=np.float32)
            _check(lib().b200_model_eval(self.h, toks.ctypes.data, len(toks), out.ctypes.data, None))
        else:
            out = np.empty(self.n_vocab, dtype=np.float32)
            _check(lib().b200_model_eval(self.h, toks.ctypes.data, len(toks), None, out.ctypes.data))
        return out

    def generate(self, prompt, n_gen, temperature=0.0, top_k=0, top_p=1.0, seed=0, ignore_eos=True, stop_token=-1):
        prompt = np.ascontiguousarray(prompt, dtype=np.int32)
        out = np.empty(n_gen, dtype=np.int32)
        s = Sampling(temperature, top_k, top_p, seed, n_gen, stop_token, int(ignore_eos), 0)
        n_out, ttft, dec = ci(0), C.c_float(0), C.c_float(0)
        _check(lib().b200_model_generate(self.h, prompt.ctypes.data, len(prompt), n_gen, C.byref(s), out.ctypes.data,
                                         C.byref(n_out), C.byref(ttft), C.byref(dec)))
        return out[:n_out.value], ttft.value, dec.value

    def submit(self, prompt, req_id, max_tokens=256, temperature=0.0, top_k=0, top_p=1.0, seed=0, ignore_eos=True, stop_token=-1, timeout_ms=0):
        prompt = np.ascontiguousarray(prompt, dtype=np.int32)
        s = Sampling(temperature, top_k, top_p, seed, max_tokens, stop_token, int(ignore_eos), timeout_ms)
        _check(lib().b200_submit(self.h, prompt.ctypes.data, len(prompt), C.byref(s), req_id))

    def prefill_slot(self, slot, toks, pos0=0):
        toks = np.ascontiguousarray(toks, dtype=np.int32)
        _check(lib().b200_model_prefill_slot(self.h, slot, toks.ctypes.data, len(toks), pos0))

    def release_slot(self, slot):
        """return the KV pages of sequence slot `slot` to the model's pool"""
        _check(lib().b200_model_release_slot(self.h, slot))

    def kv_free_pages(self):
        n = lib().b200_model_kv_free_pages(self.h)
        if n < 0:
            _check(n)
        return n

    def batch_step(self, toks, slots, pos, want_logits=False):
        toks = np.ascontiguousarray(toks, dtype=np.int32); slots = np.ascontiguousarray(slots, dtype=np.int32); pos = np.ascontiguousarray(pos, dtype=np.int32)
        nxt = np.empty(len(toks), dtype=np.int32)
        lg = np.empty((len(toks), self.n_vocab), dtype=np.float32) if want_logits else None
        _check(lib().b200_model_batch_step(self.h, toks.ctypes.data, slots.ctypes.data, pos.ctypes.data, len(toks), nxt.ctypes.data,
                                           lg.ctypes.data if want_logits else None))
        return (nxt, lg) if want_logits else nxt

    def batch_step_sampled(self, toks, slots, pos, samp, draw, want_logits=False):
        """samp: list of dicts (temperature, top_k, top_p, seed) per row; draw: tokens each row's request has generated so far"""
        toks = np.ascontiguousarray(toks, dtype=np.int32); slots = np.ascontiguousarray(slots, dtype=np.int32); pos = np.ascontiguousarray(pos, dtype=np.int32)
        draw = np.ascontiguousarray(draw, dtype=np.int32)
        arr = (Sampling * len(toks))(*[Sampling(d.get("temperature", 0.0), d.get("top_k", 0), d.get("top_p", 1.0), d.get("seed", 0), 0, -1, 1, 0) for d in samp])
        nxt = np.empty(len(toks), dtype=np.int32)
        lg = np.empty((len(toks), self.n_vocab), dtype=np.float32) if want_logits else None
        _check(lib().b200_model_batch_step_sampled(self.h, toks.ctypes.data, slots.ctypes.data, pos.ctypes.data, len(toks), arr, draw.ctypes.data,
                                                   nxt.ctypes.data, lg.ctypes.data if want_logits else None))
        return (nxt, lg) if want_logits else nxt

    def set_prefill_min(self, n):
        _check(lib().b200_model_set_prefill_min(self.h, n))

    def bench_decode(self, ctx_len, n_steps, warmup):
        ms, by = C.c_float(0), C.c_double(0)
        _check(lib().b200_bench_decode(self.h, ctx_len, n_steps, warmup, C.byref(ms), C.byref(by)))
        return ms.value, by.value

    def bench_batch_step(self, n_seqs, ctx, iters=20, warmup=3):
        """device-timed batched decode step: (ms per step, algorithmic HBM bytes per step on this GPU)"""
        ms, by = C.c_float(0), C.c_double(0)
        _check(lib().b200_bench_batch_step(self.h, n_seqs, ctx, iters, warmup, C.byref(ms), C.byref(by)))
        return ms.value, by.value

    def tokenize(self, text):
        b = text.encode("utf-8") if isinstance(text, str) else text
        n = ci(0)
        _check(lib().b200_tokenize(self.h, b, len(b), None, 0, C.byref(n)))
        out = np.empty(max(n.value, 1), dtype=np.int32)
        _check(lib().b200_tokenize(self.h, b, len(b), out.ctypes.data, n.value, C.byref(n)))
        return out[:n.value]

    def detokenize(self, toks):
        toks = np.ascontiguousarray(toks, dtype=np.int32)
        return _string_call(lambda *a: lib().b200_detokenize(self.h, toks.ctypes.data, len(toks), *a))


class GgufTokenizer:
    """Host-only view of a GGUF's tokenizer and chat template (no GPU): pass the file's bytes (the header suffices)."""

    def __init__(self, gguf_bytes):
        self.buf = bytes(gguf_bytes)

    def _tokens(self, fn, *args):
        n = ci(0)
        _check(fn(self.buf, len(self.buf), *args, None, 0, C.byref(n)))
        out = np.empty(max(n.value, 1), dtype=np.int32)
        _check(fn(self.buf, len(self.buf), *args, out.ctypes.data, n.value, C.byref(n)))
        return out[:n.value]

    def tokenize(self, text, parse_special=False):
        b = text.encode("utf-8") if isinstance(text, str) else text
        return self._tokens(lib().b200_gguf_tokenize, b, len(b), int(parse_special))

    def detokenize(self, toks):
        toks = np.ascontiguousarray(toks, dtype=np.int32)
        return _string_call(lambda *a: lib().b200_gguf_detokenize(self.buf, len(self.buf), toks.ctypes.data, len(toks), *a))

    def render_chat(self, messages):
        b = json.dumps(messages).encode()
        return self._tokens(lib().b200_gguf_render_chat, b, len(b))


class Engine:
    """Process-wide engine (one per GPU process)."""

    def __init__(self, device=0, max_ctx=4096, max_seqs=1, kv_pages=0):
        cfg = Config(device, max_ctx, max_seqs, kv_pages)
        _check(lib().b200_init(C.byref(cfg)))

    def load_file(self, model_id, path):
        h = vp()
        _check(lib().b200_model_load_file(model_id.encode(), path.encode(), C.byref(h)))
        return Model(h, self)

    def load_bytes(self, model_id, data, chunk=1 << 20):
        """open + chunked upload + commit: the path a JetStream Object Store reader drives."""
        h = vp()
        n = min(len(data), 1 << 16)
        while True:
            rc = lib().b200_model_open(model_id.encode(), data[:n], n, C.byref(h))
            if rc == -6 and n < len(data):
                n = min(len(data), n * 4); continue
            _check(rc); break
        for off in range(0, len(data), chunk):
            part = data[off:off + chunk]
            _check(lib().b200_model_upload(h, off, part, len(part)))
        _check(lib().b200_model_commit(h))
        return Model(h, self)

    def handle(self, subject, payload=b""):
        if isinstance(payload, str):
            payload = payload.encode()
        return _string_call(lambda *a: lib().b200_handle(subject.encode(), payload, len(payload), *a))

    @staticmethod
    def qgemm(type_name, raw, N, K, x, iters=1):
        """x[B x K] . W[N x K]^T with W given as raw ggml rows: the dequant-in-GEMM kernel of batched decode"""
        raw = np.ascontiguousarray(raw, dtype=np.uint8); x = np.ascontiguousarray(x, dtype=np.float32)
        y = np.empty((x.shape[0], N), dtype=np.float32)
        ms = C.c_float(0)
        _check(lib().b200_qgemm(GGML_TYPES[type_name], raw.ctypes.data, N, K, x.ctypes.data, x.shape[0], y.ctypes.data, iters, C.byref(ms)))
        return (y, ms.value) if iters > 1 else y

    @staticmethod
    def sample(logits, temperature, top_k=0, top_p=1.0, seed=0, step=0):
        logits = np.ascontiguousarray(logits, dtype=np.float32)
        s = Sampling(temperature, top_k, top_p, seed, 0, -1, 1, 0); t = ci(0)
        _check(lib().b200_sample(logits.ctypes.data, len(logits), C.byref(s), step, C.byref(t)))
        return t.value

    @staticmethod
    def cancel(req_id):
        """ask the scheduler to stop a request; it answers with a FINISH event (reason CANCELLED) at its next iteration"""
        return lib().b200_cancel(req_id)

    @staticmethod
    def poll(max_events=64, timeout_ms=1000):
        """events the scheduler thread produced since the last poll (blocks up to timeout_ms for the first one)"""
        buf = (Event * max_events)()
        n = lib().b200_poll(buf, max_events, timeout_ms)
        if n < 0:
            _check(n)
        return [Event.from_buffer_copy(buf[i]) for i in range(n)]

    @staticmethod
    def tp_shard_window(kind, N, K, rank, world):
        """(row0, nrows, col0, ncols) of the tensor [N x K] that tensor-parallel rank `rank` of `world` keeps (host only)"""
        r0, nr, c0, nc = C.c_int64(0), C.c_int64(0), ci(0), ci(0)
        _check(lib().b200_tp_shard_window(kind.encode(), N, K, rank, world, C.byref(r0), C.byref(nr), C.byref(c0), C.byref(nc)))
        return r0.value, nr.value, c0.value, nc.value

    @staticmethod
    def tp_unique_id():
        """rank 0: the 128-byte NCCL id the other ranks need for tp_init"""
        buf = C.create_string_buffer(128)
        _check(lib().b200_tp_unique_id(buf, 128))
        return buf.raw

    @staticmethod
    def tp_init(rank, world, nccl_id):
        _check(lib().b200_tp_init(rank, world, nccl_id, len(nccl_id) if nccl_id else 0))

    @staticmethod
    def worker_run(url=None, queue_group=None, n_threads=0):
        """blocking: serve the lmstudio.* subjects on a NATS server until worker_stop() (run it on a thread)"""
        return lib().b200_worker_run(url.encode() if url else None, queue_group.encode() if queue_group else None, n_threads)

    @staticmethod
    def worker_stop():
        return lib().b200_worker_stop()

    def set_models_dir(self, d):
        _check(lib().b200_set_models_dir(d.encode()))

    def list_models(self):
        return json.loads(_string_call(lib().b200_model_list))

    @staticmethod
    def dequantize(type_name, raw, n_rows, K):
        raw = np.ascontiguousarray(raw, dtype=np.uint8)
        out = np.empty((n_rows, K), dtype=np.float32)
        _check(lib().b200_dequantize(GGML_TYPES[type_name], raw.ctypes.data, n_rows, K, out.ctypes.data))
        return out

    @staticmethod
    def gemv(type_name, raw, N, K, x, iters=1):
        raw = np.ascontiguousarray(raw, dtype=np.uint8)
        x = np.ascontiguousarray(x, dtype=np.float32)
        y = np.empty(N, dtype=np.float32)
        ms = C.c_float(0)
        _check(lib().b200_gemv(GGML_TYPES[type_name], raw.ctypes.data, N, K, x.ctypes.data, y.ctypes.data, iters, C.byref(ms)))
        return y, ms.value

    @staticmethod
    def gemm_bf16(A, B):
        A = np.ascontiguousarray(A, dtype=np.float32); B = np.ascontiguousarray(B, dtype=np.float32)
        M, K = A.shape; N = B.shape[0]
        Cm = np.empty((M, N), dtype=np.float32)
        _check(lib().b200_gemm_bf16(A.ctypes.data, B.ctypes.data, Cm.ctypes.data, M, N, K))
        return Cm

    @staticmethod
    def launch_count():
        return int(lib().b200_launch_count())

    def shutdown(self):
        lib().b200_shutdown()
