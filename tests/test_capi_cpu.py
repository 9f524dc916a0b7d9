"""CPU-side checks of the C-ABI library: it loads, exports every declared symbol, the layout maps are
bijections, the handlers' validation + envelope bytes follow the reference, and compute entry points
fail loudly without a GPU (no CPU fallback)."""
import json
import os
import re

import pytest

import nats_llm_studio_b200 as b200
from nats_llm_studio_b200 import engine as eng

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "b200_engine.h")).read()
    declared = set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", hdr))
    L = b200.lib()
    for name in declared:
        assert hasattr(L, name), f"{name} declared in include/b200_engine.h but not exported"
    assert declared == set(eng.SYMBOLS), "python binding table and header disagree"


def test_layout_selftest():
    assert b200.lib().b200_layout_selftest() == 0


def _handle(subject, payload=b""):
    import ctypes as C
    L = b200.lib()
    need = C.c_size_t(0)
    L.b200_handle(subject.encode(), payload, len(payload), None, 0, C.byref(need))
    buf = C.create_string_buffer(need.value + 1)
    assert L.b200_handle(subject.encode(), payload, len(payload), buf, need.value + 1, C.byref(need)) == 0
    return buf.raw[:need.value]


# golden envelope bytes derived from nats_llm_studio.go (respondError with a nil map emits "data":null)
ENVELOPES = [
    ("lmstudio.chat_model", b"", b'{"ok":false,"error":"payload vazio em ChatModel","data":null}'),                       # :331-334
    ("lmstudio.chat_model", b"{", b'{"ok":false,"error":"invalid JSON in ChatModel: unexpected end of JSON input","data":null}'),  # :339-342
    ("lmstudio.chat_model", b"x", b'{"ok":false,"error":"invalid JSON in ChatModel: invalid character \'x\' looking for beginning of value","data":null}'),
    ("lmstudio.chat_model", b'{"model":""}', b'{"ok":false,"error":"\'model\' is required in ChatModel","data":null}'),  # :343-346
    ("lmstudio.chat_model", b'{"messages":[]}', b'{"ok":false,"error":"\'model\' is required in ChatModel","data":null}'),
    ("lmstudio.chat_model", b"null", b'{"ok":false,"error":"\'model\' is required in ChatModel","data":null}'),
    ("lmstudio.chat_model", b'{"model":7}', b'{"ok":false,"error":"invalid JSON in ChatModel: json: cannot unmarshal number into Go struct field .model of type string","data":null}'),
    ("lmstudio.pull_model", b"", b'{"ok":false,"error":"invalid JSON in PullModel: unexpected end of JSON input","data":null}'),   # :255-258
    ("lmstudio.pull_model", b"{}", b'{"ok":false,"error":"\'identifier\' is required","data":null}'),                    # :259-262
    ("lmstudio.pull_model", b'{"identifier":[1]}', b'{"ok":false,"error":"invalid JSON in PullModel: json: cannot unmarshal array into Go struct field PullModelRequest.identifier of type string","data":null}'),
    ("lmstudio.delete_model", b"{}", b'{"ok":false,"error":"\'model_id\' is required","data":null}'),                    # :297-300
    ("lmstudio.delete_model", b"]", b'{"ok":false,"error":"invalid JSON in DeleteModel: invalid character \']\' looking for beginning of value","data":null}'),
    ("lmstudio.delete_model", b'{"model_id":"ghost"}', b'{"ok":false,"error":"model directory not found: hbm://ghost","data":{"dir":"hbm://ghost","model_id":"ghost"}}'),  # :304-311
    ("lmstudio.list_models", b"", b'{"ok":true,"data":{"http_status":200,"models":{"data":[],"object":"list"}}}'),        # :240-246
]


@pytest.mark.parametrize("subject,payload,want", ENVELOPES)
def test_envelope_bytes(subject, payload, want):
    assert _handle(subject, payload) == want


def test_html_escaping_and_key_order():
    # Go's json.Marshal escapes <, >, & and sorts map keys (http_status < response)
    out = _handle("lmstudio.chat_model", '{"model":"a<b>&c "}'.encode())
    assert out.startswith(b'{"ok":true,"data":{"http_status":404,"response":{"error":{"message":"model \'a\\u003cb\\u003e\\u0026c\\u2028\' not found"')
    assert json.loads(out)["data"]["response"]["error"]["code"] == "model_not_found"


def test_pull_missing_file_keeps_model_and_output_keys(tmp_path):
    b200.lib().b200_set_models_dir(str(tmp_path).encode())
    out = json.loads(_handle("lmstudio.pull_model", b'{"identifier":"pub/none"}'))
    assert out["ok"] is False and set(out["data"]) == {"model", "output"} and out["data"]["model"] == "pub/none"   # :266-276


@pytest.mark.skipif(_has_gpu(), reason="checks the no-GPU failure mode")
def test_compute_fails_loudly_without_gpu(golden_dir):
    with pytest.raises(b200.EngineError) as e:
        b200.Engine()
    assert "no CPU fallback" in str(e.value)
    import numpy as np
    with pytest.raises(b200.EngineError):
        b200.Engine.dequantize("Q8_0", np.zeros(34 * 8, np.uint8), 1, 256)
    out = json.loads(_handle("lmstudio.pull_model", json.dumps({"identifier": os.path.join(golden_dir, "tiny-llama.gguf")}).encode()))
    assert out["ok"] is False and "no CPU fallback" in out["error"]
