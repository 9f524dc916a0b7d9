"""The GGUF tokenizer and chat templates (csrc/tokenizer.hpp) against Hugging Face `tokenizers` golden vectors
(tests/golden/make_tokenizer_golden.py): byte-level BPE with the Llama-3 and GPT-2 pre-tokeniser patterns, special tokens,
detokenisation, and the Llama-3 chat layout.  Host only (b200_gguf_* entry points)."""
import json
import os

import numpy as np
import pytest

import nats_llm_studio_b200 as b200


@pytest.fixture(scope="module")
def gold(golden_dir):
    return json.load(open(os.path.join(golden_dir, "bpe_golden.json")))


@pytest.mark.parametrize("family", ["llama3", "gpt2"])
def test_bpe_matches_hf_tokenizers(gold, golden_dir, family):
    g = gold[family]
    tk = b200.GgufTokenizer(open(os.path.join(golden_dir, g["gguf"]), "rb").read())
    for text, ids in g["encode"].items():
        got = tk.tokenize(text).tolist()
        assert got == ids, (family, text, got, ids)
        assert tk.detokenize(ids).decode("utf-8") == g["decode"][text] == text


def test_special_tokens_and_chat_template(gold, golden_dir):
    g = gold["llama3"]
    tk = b200.GgufTokenizer(open(os.path.join(golden_dir, g["gguf"]), "rb").read())
    sp = g["specials"]
    ids = tk.tokenize("<|start_header_id|>user<|end_header_id|>", parse_special=True).tolist()
    assert ids[0] == sp["<|start_header_id|>"] and ids[-1] == sp["<|end_header_id|>"] and len(ids) > 2
    assert sp["<|eot_id|>"] not in tk.tokenize("<|eot_id|>", parse_special=False).tolist()      # user text never becomes a control token
    assert tk.detokenize([sp["<|eot_id|>"]]) == b""                                               # control tokens print nothing
    for chat in g["chats"]:
        assert tk.render_chat(chat["messages"]).tolist() == chat["ids"], chat["messages"]
    with pytest.raises(b200.EngineError):
        tk.render_chat([{"role": "user"}])


def test_synthetic_byte_vocabulary_still_works(golden_dir):
    """the random-init GGUFs of tools/make_gguf.py carry a byte-level vocabulary and the minimal <|role|> template"""
    tk = b200.GgufTokenizer(open(os.path.join(golden_dir, "tiny-llama.gguf"), "rb").read()[:200000])
    ids = tk.tokenize("hi <|user|>", parse_special=True)
    assert tk.detokenize(ids[:3]) == b"hi " and len(ids) == 4
    chat = tk.render_chat([{"role": "user", "content": "yo"}]).tolist()
    assert len(chat) == 1 + 1 + 2 + 1 + 1                   # bos, <|user|>, y, o, <|end|>, <|assistant|>


def test_sentencepiece_vocabulary_matches_sentencepiece(golden_dir):
    """tokenizer.ggml.model = "llama" (Llama-2 / Mistral style GGUFs): score-ordered bigram merges, U+2581 spaces with the dummy
    prefix, byte fallback — against `sentencepiece` itself (tests/golden/make_spm_golden.py), and the [INST] chat layout."""
    g = json.load(open(os.path.join(golden_dir, "spm_golden.json")))
    tk = b200.GgufTokenizer(open(os.path.join(golden_dir, g["gguf"]), "rb").read())
    for text, ids in g["encode"].items():
        got = tk.tokenize(text).tolist()
        assert got == ids, (text, got, ids)
        assert tk.detokenize(ids).decode("utf-8") == g["decode"][text], text
    assert tk.tokenize("<s>", parse_special=True).tolist() == [g["bos"]]
    assert g["bos"] not in tk.tokenize("<s>", parse_special=False).tolist()                        # user text never becomes a control token
    for chat in g["chats"]:
        assert tk.render_chat(chat["messages"]).tolist() == chat["ids"], chat["messages"]
