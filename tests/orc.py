"""ctypes wrapper around oracle/liboracle.so — the CPU parity checker (test infrastructure only)."""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_SO = os.path.join(ROOT, "oracle", "liboracle.so")
_lib = None

GGML = {"F32": 0, "F16": 1, "Q8_0": 8, "Q4_K": 12, "Q5_K": 13, "Q6_K": 14, "BF16": 30}


def lib():
    global _lib
    if _lib is None:
        src = os.path.join(ROOT, "oracle", "oracle.c")
        if not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")], stdout=subprocess.DEVNULL)
        L = ctypes.CDLL(_SO)
        vp, ci, cf = ctypes.c_void_p, ctypes.c_int, ctypes.c_float
        L.orc_load.restype = vp; L.orc_load.argtypes = [ctypes.c_char_p]
        L.orc_free.argtypes = [vp]
        L.orc_dequantize.argtypes = [ci, vp, ctypes.c_int64, vp]
        for f in ("orc_n_vocab", "orc_n_embd", "orc_n_layer", "orc_n_past"):
            getattr(L, f).argtypes = [vp]
        L.orc_set_act_quant.argtypes = [vp, ci]; L.orc_set_kv_f16.argtypes = [vp, ci]; L.orc_set_xchg_f16.argtypes = [vp, ci]
        L.orc_set_threads.argtypes = [ci]
        L.orc_reset.argtypes = [vp]
        L.orc_eval_token.argtypes = [vp, ctypes.c_int32, vp]
        L.orc_forward.argtypes = [vp, vp, ci, vp, ci]
        L.orc_generate.argtypes = [vp, vp, ci, ci, vp, vp]
        L.orc_tensor_f32.restype = ctypes.c_int64; L.orc_tensor_f32.argtypes = [vp, ctypes.c_char_p, vp, ctypes.c_int64]
        _lib = L
    return _lib


def dequantize(type_name, raw, n_elems):
    raw = np.ascontiguousarray(raw, dtype=np.uint8)
    out = np.empty(n_elems, dtype=np.float32)
    rc = lib().orc_dequantize(GGML[type_name], raw.ctypes.data, n_elems, out.ctypes.data)
    if rc:
        raise ValueError("orc_dequantize failed")
    return out


class Model:
    def __init__(self, path, act_quant=0, kv_f16=0, xchg_f16=None):
        self.h = lib().orc_load(path.encode())
        if not self.h:
            raise RuntimeError(f"oracle could not load {path}")
        self.n_vocab = lib().orc_n_vocab(self.h)
        self.n_embd = lib().orc_n_embd(self.h)
        lib().orc_set_act_quant(self.h, act_quant); lib().orc_set_kv_f16(self.h, kv_f16)
        # engine numerics = int8 activations + fp16 KV + fp16 exchange vectors; xchg follows act_quant unless given
        lib().orc_set_xchg_f16(self.h, int(act_quant if xchg_f16 is None else xchg_f16))

    def close(self):
        if self.h:
            lib().orc_free(self.h); self.h = None

    def __del__(self):
        self.close()

    def forward(self, toks, all_logits=True):
        toks = np.ascontiguousarray(toks, dtype=np.int32)
        out = np.empty((len(toks), self.n_vocab) if all_logits else (self.n_vocab,), dtype=np.float32)
        if lib().orc_forward(self.h, toks.ctypes.data, len(toks), out.ctypes.data, int(all_logits)):
            raise RuntimeError("orc_forward failed")
        return out

    def reset(self):
        lib().orc_reset(self.h)

    def eval_token(self, tok, want_logits=True):
        out = np.empty(self.n_vocab, dtype=np.float32) if want_logits else None
        if lib().orc_eval_token(self.h, int(tok), out.ctypes.data if want_logits else None):
            raise RuntimeError("orc_eval_token failed")
        return out

    def generate(self, prompt, n_gen, want_logits=False):
        prompt = np.ascontiguousarray(prompt, dtype=np.int32)
        out = np.empty(n_gen, dtype=np.int32)
        lg = np.empty((n_gen, self.n_vocab), dtype=np.float32) if want_logits else None
        if lib().orc_generate(self.h, prompt.ctypes.data, len(prompt), n_gen, out.ctypes.data,
                              lg.ctypes.data if want_logits else None):
            raise RuntimeError("orc_generate failed")
        return (out, lg) if want_logits else out

    def tensor(self, name):
        n = lib().orc_tensor_f32(self.h, name.encode(), None, 0)
        if n < 0:
            raise KeyError(name)
        out = np.empty(n, dtype=np.float32)
        lib().orc_tensor_f32(self.h, name.encode(), out.ctypes.data, n)
        return out


def matvec(type_name, raw, N, K, x, act_quant=0):
    L = lib()
    L.orc_matvec.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
    raw = np.ascontiguousarray(raw, dtype=np.uint8); x = np.ascontiguousarray(x, dtype=np.float32)
    y = np.empty(N, dtype=np.float32)
    if L.orc_matvec(GGML[type_name], raw.ctypes.data, N, K, x.ctypes.data, y.ctypes.data, act_quant):
        raise ValueError("orc_matvec failed")
    return y
