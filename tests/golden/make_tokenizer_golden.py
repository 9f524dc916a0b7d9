#!/usr/bin/env python3
"""Golden vectors for the GGUF BPE tokenizer (csrc/tokenizer.hpp), made with an IMPORTED implementation: Hugging Face
`tokenizers` 0.22 (run here, in the build container; outputs committed).  No real Llama-3 / granite vocabulary exists
offline, so a byte-level BPE vocabulary is TRAINED on a small synthetic corpus with the Llama-3 pre-tokeniser pattern,
written into a GGUF header (tokenizer.ggml.model = "gpt2", tokens, merges, token types, a Llama-3-style chat template)
with llama.cpp's gguf-py writer, and a set of texts is encoded with `tokenizers`; the C++ tokenizer must reproduce the
ids from the GGUF alone.  A second header uses GPT-2's pattern (tokenizer.ggml.pre = "gpt-2").

Outputs: tests/golden/bpe-llama3.gguf, bpe-gpt2.gguf (headers only: no tensors), bpe_golden.json"""
import json
import os

from gguf import GGUFWriter
from tokenizers import Regex, Tokenizer, decoders, models, pre_tokenizers, trainers

HERE = os.path.dirname(os.path.abspath(__file__))
LLAMA3 = r"(?i:'s|'t|'re|'ve|'m|'ll|'d)|[^\r\n\p{L}\p{N}]?\p{L}+|\p{N}{1,3}| ?[^\s\p{L}\p{N}]+[\r\n]*|\s*[\r\n]+|\s+(?!\S)|\s+"
GPT2 = r"'s|'t|'re|'ve|'m|'ll|'d| ?\p{L}+| ?\p{N}+| ?[^\s\p{L}\p{N}]+|\s+(?!\S)|\s+"
SPECIALS = ["<|begin_of_text|>", "<|end_of_text|>", "<|start_header_id|>", "<|end_header_id|>", "<|eot_id|>"]
CORPUS = ["The quick brown fox jumps over the lazy dog. ", "Hello, world! This is a test of the tokenizer; it isn't hard.\n",
          "def f(x):\n    return x * 2 + 1  # comment\n\n", "Numbers: 1 22 333 4444 55555 3.14159 2024-11-07\n", "We're here, they've gone, I'll stay, she'd know, I'm fine.\n",
          "naïve café résumé über straße — ¿qué? «oui» 日本語のテキスト 中文 한국어\n", "   leading spaces\tand\ttabs  \n\n\nmultiple newlines  \n  ", "UPPER lower MiXeD CamelCaseWords snake_case_words kebab-case\n",
          "email@example.com http://example.org/path?query=1&b=2 #hashtag @mention $100 50% (parens) [brackets] {braces}\n"] * 4
TEXTS = ["Hello world", "Hello, world! It's a test.", "  two spaces   three", "line1\nline2\n\n\nline5   \n  indented", "I'M SHOUTING, DON'T YOU'LL WE'VE",
         "12345678 9 10 007", "x=1;y=2;//c", "tabs\tand\r\nCRLF", "naïve café über", "日本語 text 中文!", "trailing space ", " leading", "a", "", "!!!???...", "end.\n",
         "snake_case CamelCase kebab-case 3.14", "The year 2024 was… fine — really",
         # scripts, numerals and symbols beyond ASCII: \\p{L} / \\p{N} classes of the pre-tokeniser
         "Привет, мир! Ελληνικά עברית العربية हिन्दी ไทย", "x² + y³ = ½ ٣٤٥ ३४५ Ⅻ ① 12345", "price: 100€ ±5% → ok ✓ 😀👍 done", "a\u0301 e\u0308 combining marks", "ｆｕｌｌwidth １２３ 𝒜𝓑 𝟙𝟚𝟛"]
CHATS = [[{"role": "user", "content": "Hi there"}], [{"role": "system", "content": "You are terse."}, {"role": "user", "content": "Say hello from NATS client in Go."}],
         [{"role": "user", "content": "a"}, {"role": "assistant", "content": "b\n"}, {"role": "user", "content": "<|eot_id|> is text here"}]]
TEMPLATE = ("{% set loop_messages = messages %}{% for message in loop_messages %}{% set content = '<|start_header_id|>' + message['role'] + '<|end_header_id|>\n\n'+ message['content'] + '<|eot_id|>' %}"
            "{% if loop.index0 == 0 %}{% set content = bos_token + content %}{% endif %}{{ content }}{% endfor %}{{ '<|start_header_id|>assistant<|end_header_id|>\n\n' }}")


def make(name, pattern, pre):
    tok = Tokenizer(models.BPE())
    tok.pre_tokenizer = pre_tokenizers.Sequence([pre_tokenizers.Split(Regex(pattern), behavior="isolated"), pre_tokenizers.ByteLevel(add_prefix_space=False, use_regex=False)])
    tok.decoder = decoders.ByteLevel()
    tok.train_from_iterator(CORPUS, trainers.BpeTrainer(vocab_size=700, initial_alphabet=pre_tokenizers.ByteLevel.alphabet(), special_tokens=[], show_progress=False))
    model = json.loads(tok.to_str())["model"]
    vocab = sorted(model["vocab"].items(), key=lambda kv: kv[1])
    tokens = [t for t, _ in vocab] + SPECIALS
    merges = [m if isinstance(m, str) else " ".join(m) for m in model["merges"]]
    types = [1] * len(vocab) + [3] * len(SPECIALS)
    path = os.path.join(HERE, f"{name}.gguf")
    w = GGUFWriter(path, "llama")
    w.add_tokenizer_model("gpt2"); w.add_tokenizer_pre(pre); w.add_token_list(tokens); w.add_token_types(types); w.add_token_merges(merges)
    w.add_bos_token_id(len(vocab)); w.add_eos_token_id(len(vocab) + 1); w.add_chat_template(TEMPLATE)
    w.write_header_to_file(); w.write_kv_data_to_file(); w.write_tensors_to_file(); w.close()
    sp = {s: len(vocab) + i for i, s in enumerate(SPECIALS)}
    enc = {t: tok.encode(t).ids for t in TEXTS}
    dec = {t: tok.decode(tok.encode(t).ids) for t in TEXTS}
    chats = []
    for msgs in CHATS:                                              # the template, applied by hand with `tokenizers` for the text parts
        ids = [sp["<|begin_of_text|>"]]
        for m in msgs:
            ids += [sp["<|start_header_id|>"]] + tok.encode(m["role"]).ids + [sp["<|end_header_id|>"]] + tok.encode("\n\n").ids + tok.encode(m["content"]).ids + [sp["<|eot_id|>"]]
        ids += [sp["<|start_header_id|>"]] + tok.encode("assistant").ids + [sp["<|end_header_id|>"]] + tok.encode("\n\n").ids
        chats.append({"messages": msgs, "ids": ids})
    return {"gguf": f"{name}.gguf", "n_vocab": len(tokens), "specials": sp, "encode": enc, "decode": dec, "chats": chats}


if __name__ == "__main__":
    import tokenizers
    out = {"made_with": f"tokenizers {tokenizers.__version__}", "llama3": make("bpe-llama3", LLAMA3, "llama-bpe"), "gpt2": make("bpe-gpt2", GPT2, "gpt-2")}
    json.dump(out, open(os.path.join(HERE, "bpe_golden.json"), "w"), ensure_ascii=False, indent=0)
    print({k: v["n_vocab"] for k, v in out.items() if isinstance(v, dict)})
