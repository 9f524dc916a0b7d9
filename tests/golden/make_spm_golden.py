#!/usr/bin/env python3
"""Golden vectors for the GGUF SentencePiece ("llama") tokenizer path of csrc/tokenizer.hpp, made with an IMPORTED
implementation: `sentencepiece` 0.2 (run here, in the build container; outputs committed).  No real Llama-2 / Mistral
vocabulary exists offline, so a BPE-type SentencePiece model is TRAINED on a small synthetic corpus with Llama-2's trainer
settings (identity normalisation, byte fallback, digits split, dummy prefix), written into a GGUF header
(tokenizer.ggml.model = "llama", tokens, scores, token types, a Llama-2-style [INST] chat template) with llama.cpp's gguf-py
writer, and a set of texts is encoded with `sentencepiece`; the C++ tokenizer must reproduce the ids from the GGUF alone.

Outputs: tests/golden/spm-llama2.gguf (header only: no tensors), spm_golden.json"""
import io
import json
import os

import sentencepiece as spm
from gguf import GGUFWriter

HERE = os.path.dirname(os.path.abspath(__file__))
CORPUS = ["The quick brown fox jumps over the lazy dog.", "Hello, world! This is a test of the tokenizer; it isn't hard.",
          "def f(x): return x * 2 + 1  # comment", "Numbers: 1 22 333 4444 55555 3.14159 2024-11-07", "We're here, they've gone, I'll stay, she'd know, I'm fine.",
          "naïve café résumé über straße — ¿qué? «oui»", "   leading spaces and  double  spaces  ", "UPPER lower MiXeD CamelCaseWords snake_case_words kebab-case",
          "email@example.com http://example.org/path?query=1&b=2 #hashtag @mention $100 50% (parens) [brackets] {braces}",
          "[INST] Say hello from a NATS client in Go. [/INST] Hello from NATS!", "<<SYS>> You are terse. <</SYS>>"] * 6
TEXTS = ["Hello world", "Hello, world! It's a test.", "  two spaces   three", "line1\nline2\n\n\nline5", "I'M SHOUTING, DON'T YOU'LL WE'VE",
         "12345678 9 10 007", "x=1;y=2;//c", "tabs\tand\r\nCRLF", "naïve café über", "日本語 text 中文!", "trailing space ", " leading", "a", "", "!!!???...",
         "snake_case CamelCase kebab-case 3.14", "The year 2024 was… fine — really", "[INST] hi [/INST]"]
CHATS = [[{"role": "user", "content": "Hi there"}], [{"role": "system", "content": "You are terse."}, {"role": "user", "content": "Say hello from NATS client in Go."}],
         [{"role": "user", "content": "a"}, {"role": "assistant", "content": "b"}, {"role": "user", "content": "c d"}]]
TEMPLATE = ("{% for message in messages %}{% if message['role'] == 'user' %}{{ bos_token + '[INST] ' + message['content'] + ' [/INST]' }}"
            "{% elif message['role'] == 'assistant' %}{{ ' ' + message['content'] + ' ' + eos_token }}{% endif %}{% endfor %}")


def main():
    model = io.BytesIO()
    spm.SentencePieceTrainer.train(sentence_iterator=iter(CORPUS), model_writer=model, vocab_size=420, model_type="bpe", byte_fallback=True,
                                   normalization_rule_name="identity", remove_extra_whitespaces=False, add_dummy_prefix=True, split_digits=True,
                                   allow_whitespace_only_pieces=True, character_coverage=0.99, unk_id=0, bos_id=1, eos_id=2, pad_id=-1, minloglevel=2)
    sp = spm.SentencePieceProcessor(model_proto=model.getvalue())
    n = sp.get_piece_size()
    tokens, scores, types = [], [], []
    for i in range(n):
        tokens.append(sp.id_to_piece(i)); scores.append(float(sp.get_score(i)))
        types.append(2 if sp.is_unknown(i) else 3 if sp.is_control(i) else 6 if sp.is_byte(i) else 1)      # gguf TokenType: NORMAL 1, UNKNOWN 2, CONTROL 3, BYTE 6
    path = os.path.join(HERE, "spm-llama2.gguf")
    w = GGUFWriter(path, "llama")
    w.add_tokenizer_model("llama"); w.add_token_list(tokens); w.add_token_scores(scores); w.add_token_types(types)
    w.add_bos_token_id(sp.bos_id()); w.add_eos_token_id(sp.eos_id()); w.add_unk_token_id(sp.unk_id()); w.add_add_bos_token(True); w.add_add_space_prefix(True)
    w.add_chat_template(TEMPLATE)
    w.write_header_to_file(); w.write_kv_data_to_file(); w.write_tensors_to_file(); w.close()
    enc = {t: sp.encode(t) for t in TEXTS}
    dec = {t: sp.decode(sp.encode(t)) for t in TEXTS}
    chats = []
    for msgs in CHATS:                                              # the Llama-2 layout, applied by hand with `sentencepiece` for the text parts
        ids, sys_block = [], ""
        for m in msgs:
            if m["role"] == "system":
                sys_block = "<<SYS>>\n" + m["content"] + "\n<</SYS>>\n\n"
            elif m["role"] == "user":
                ids += [sp.bos_id()] + sp.encode("[INST] " + sys_block + m["content"] + " [/INST]"); sys_block = ""
            else:
                ids += sp.encode(m["content"]) + [sp.eos_id()]
        chats.append({"messages": msgs, "ids": ids})
    out = {"made_with": f"sentencepiece {spm.__version__}", "gguf": "spm-llama2.gguf", "n_vocab": n, "bos": sp.bos_id(), "eos": sp.eos_id(), "encode": enc, "decode": dec, "chats": chats}
    json.dump(out, open(os.path.join(HERE, "spm_golden.json"), "w"), ensure_ascii=False, indent=0)
    print(n, "pieces;", sum(1 for t in types if t == 6), "byte tokens")


if __name__ == "__main__":
    main()
