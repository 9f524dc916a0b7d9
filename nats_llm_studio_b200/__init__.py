"""B200-native GGUF inference engine behind the nats-llm-studio NATS subjects.

The product is `libb200engine.so` (CUDA kernels for sm_100a + C ABI, sources in `csrc/`, header in
`include/b200_engine.h`).  This package is the thin ctypes binding used by tests, smoke and bench;
there is no Python or CPU compute path: every call fails loudly if the library or a GPU is missing.
"""
from .engine import (Engine, EngineError, Model, GgufTokenizer, lib, GGML_TYPES, EV_TOKEN, EV_FINISH, EV_ERROR,  # noqa: F401
                     FINISH_STOP, FINISH_LENGTH, FINISH_CANCELLED, FINISH_TIMEOUT)
