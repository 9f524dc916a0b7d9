// tokenizer.hpp — the text side of chat_model: the reference forwards `messages` verbatim to a runtime that applies the
// model's chat template and tokenizer (/root/reference/nats_llm_studio.go:158-167); here that runtime is in-process, so
// the GGUF's own tokenizer is implemented:
//   * tokenizer.ggml.model == "gpt2": byte-level BPE (tokens + merges from the GGUF; Llama-3 and granite-3 use it) with
//     the pre-tokeniser split named by tokenizer.ggml.pre ("llama-bpe"/"llama3": the Llama-3 pattern; anything else:
//     GPT-2's).  The split is hand-coded (no regex engine, no ICU); \p{L} and \p{N} come from generated range tables
//     (unicode_tables.hpp, tools/make_unicode_tables.py), \s is the Unicode White_Space set.
//     Checked against Hugging Face `tokenizers` on a trained vocabulary (tests/golden/make_tokenizer_golden.py).
//   * tokenizer.ggml.model == "llama" with scores: SentencePiece BPE as llama.cpp applies it to Llama-2 / Mistral style
//     vocabularies (spaces -> U+2581 with the dummy prefix, score-ordered bigram merges, <0xXX> byte fallback).  Checked
//     against `sentencepiece` on a trained model (tests/golden/make_spm_golden.py).
//   * anything else (the synthetic byte-level vocabulary of tools/make_gguf.py): one token per byte.
// Chat templates: tokenizer.chat_template is Jinja; it is not executed — the template's family is recognised from its
// marker tokens (Llama-3 headers, granite roles, ChatML) or its [INST] markers (Llama-2) and rendered natively; unknown templates fall back to the
// minimal <|role|> form of the synthetic vocabulary.
#pragma once
#include <algorithm>
#include <cstdint>
#include <queue>
#include <string>
#include <unordered_map>
#include <vector>

#include "unicode_tables.hpp"

namespace b200 {

struct Tokenizer {
    enum Kind { BYTES, BPE, SPM } kind = BYTES;
    enum Template { TPL_MINIMAL, TPL_LLAMA3, TPL_GRANITE, TPL_CHATML, TPL_LLAMA2 } tpl = TPL_MINIMAL;
    bool llama3_split = false;
    std::vector<std::string> piece;                        // token id -> text it stands for (detokenisation)
    std::unordered_map<std::string, int> special;          // control / user-defined tokens, matched verbatim when parse_special
    std::vector<std::string> special_sorted;               // longest first
    int byte_tok[256];
    std::unordered_map<std::string, int> vocab;            // BPE: byte-unicode token string -> id
    std::unordered_map<std::string, int> rank;             // BPE: "left right" -> merge priority
    std::string b2u[256];                                  // GPT-2 byte -> unicode character (UTF-8)
    std::unordered_map<uint32_t, int> u2b;
    std::vector<float> score;                              // SPM: piece score (merge priority)
    int n_vocab = 0, bos = -1, eos = -1; bool add_bos = true, add_space_prefix = true;

    // ---------------------------------------------------------------- construction
    static void utf8_put(uint32_t c, std::string& s) {
        if (c < 0x80) s.push_back((char)c);
        else if (c < 0x800) { s.push_back((char)(0xC0 | (c >> 6))); s.push_back((char)(0x80 | (c & 63))); }
        else if (c < 0x10000) { s.push_back((char)(0xE0 | (c >> 12))); s.push_back((char)(0x80 | ((c >> 6) & 63))); s.push_back((char)(0x80 | (c & 63))); }
        else { s.push_back((char)(0xF0 | (c >> 18))); s.push_back((char)(0x80 | ((c >> 12) & 63))); s.push_back((char)(0x80 | ((c >> 6) & 63))); s.push_back((char)(0x80 | (c & 63))); }
    }
    static uint32_t utf8_get(const std::string& s, size_t& i) {          // decodes one code point, advances i (invalid bytes come out as themselves)
        const unsigned char c = (unsigned char)s[i];
        auto cont = [&](size_t k) { return i + k < s.size() && ((unsigned char)s[i + k] & 0xC0) == 0x80; };
        if (c < 0x80) { i += 1; return c; }
        if ((c & 0xE0) == 0xC0 && cont(1)) { uint32_t r = ((c & 31u) << 6) | ((unsigned char)s[i + 1] & 63u); i += 2; return r; }
        if ((c & 0xF0) == 0xE0 && cont(1) && cont(2)) { uint32_t r = ((c & 15u) << 12) | (((unsigned char)s[i + 1] & 63u) << 6) | ((unsigned char)s[i + 2] & 63u); i += 3; return r; }
        if ((c & 0xF8) == 0xF0 && cont(1) && cont(2) && cont(3)) { uint32_t r = ((c & 7u) << 18) | (((unsigned char)s[i + 1] & 63u) << 12) | (((unsigned char)s[i + 2] & 63u) << 6) | ((unsigned char)s[i + 3] & 63u); i += 4; return r; }
        i += 1; return c;
    }
    void build(const std::vector<std::string>& tokens, const std::vector<int>& types, const std::vector<std::string>& merges, const std::string& model,
               const std::string& pre, const std::string& chat_template, int bos_, int eos_, bool add_bos_, int n_vocab_,
               const std::vector<float>& scores = {}, bool add_space_prefix_ = true) {
        n_vocab = n_vocab_; bos = bos_; eos = eos_; add_bos = add_bos_; add_space_prefix = add_space_prefix_; score = scores;
        for (int i = 0; i < 256; i++) byte_tok[i] = -1;
        {   // GPT-2's reversible byte <-> printable-character map
            int n = 0;
            for (int b = 0; b < 256; b++) {
                const bool keep = (b >= 33 && b <= 126) || (b >= 161 && b <= 172) || (b >= 174 && b <= 255);
                const uint32_t cp = keep ? (uint32_t)b : 256u + (uint32_t)n++;
                b2u[b].clear(); utf8_put(cp, b2u[b]); u2b[cp] = b;
            }
        }
        kind = (model == "gpt2" && !merges.empty()) ? BPE : (model == "llama" && scores.size() == tokens.size() && !tokens.empty()) ? SPM : BYTES;
        llama3_split = pre == "llama-bpe" || pre == "llama3" || pre == "llama-v3";
        piece.assign(tokens.size(), ""); special.clear(); vocab.clear(); rank.clear();
        for (size_t i = 0; i < tokens.size(); i++) {
            const std::string& s = tokens[i]; const int ty = i < types.size() ? types[i] : 1;
            if (ty == 3 || ty == 4) { special[s] = (int)i; piece[i] = ty == 4 ? s : ""; if (kind == BPE) vocab[s] = (int)i; continue; }   // CONTROL prints nothing
            if (kind == SPM && ty == 2) { piece[i] = ""; continue; }                                       // UNKNOWN: never produced (byte fallback), prints nothing
            if (kind == BPE) {
                vocab[s] = (int)i;
                std::string raw; size_t k = 0;
                while (k < s.size()) { const uint32_t cp = utf8_get(s, k); auto it = u2b.find(cp); if (it != u2b.end()) raw.push_back((char)it->second); else utf8_put(cp, raw); }
                piece[i] = raw;
            } else if (ty == 6 && s.size() == 6 && s.compare(0, 3, "<0x") == 0 && s[5] == '>') {      // BYTE
                const int b = (int)strtol(s.substr(3, 2).c_str(), nullptr, 16); byte_tok[b & 255] = (int)i; piece[i] = std::string(1, (char)b);
            } else { if (kind == SPM) vocab[s] = (int)i; std::string p = s; size_t k; while ((k = p.find("\xE2\x96\x81")) != std::string::npos) p.replace(k, 3, " "); piece[i] = p; }
        }
        for (size_t i = 0; i < merges.size(); i++) rank[merges[i]] = (int)i;
        special_sorted.clear(); for (auto& kv : special) if (!kv.first.empty()) special_sorted.push_back(kv.first);
        std::sort(special_sorted.begin(), special_sorted.end(), [](const std::string& a, const std::string& b) { return a.size() != b.size() ? a.size() > b.size() : a < b; });
        auto has = [&](const char* t) { return chat_template.find(t) != std::string::npos && special.count(t); };
        tpl = has("<|start_header_id|>") ? TPL_LLAMA3 : has("<|start_of_role|>") ? TPL_GRANITE : has("<|im_start|>") ? TPL_CHATML
            : chat_template.find("[INST]") != std::string::npos ? TPL_LLAMA2 : TPL_MINIMAL;
    }

    // ---------------------------------------------------------------- character classes (see the header comment)
    static bool is_space(uint32_t c) { return c == ' ' || (c >= 9 && c <= 13) || c == 0x85 || c == 0xA0 || c == 0x1680 || (c >= 0x2000 && c <= 0x200A) || c == 0x2028 || c == 0x2029 || c == 0x202F || c == 0x205F || c == 0x3000; }
    static bool is_digit(uint32_t c) { return c < 0x80 ? (c >= '0' && c <= '9') : uni_in(UNI_NUMBER, (int)(sizeof(UNI_NUMBER) / sizeof(UNI_NUMBER[0])), c); }      // \p{N}
    static bool is_letter(uint32_t c) {                                                                                                                 // \p{L}
        if (c < 0x80) return (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z');
        return uni_in(UNI_LETTER, (int)(sizeof(UNI_LETTER) / sizeof(UNI_LETTER[0])), c);
    }

    // Splits text into pre-tokens (byte ranges) the way the model's pattern does.  Llama-3:
    //   (?i:'s|'t|'re|'ve|'m|'ll|'d) | [^\r\n\p{L}\p{N}]?\p{L}+ | \p{N}{1,3} | ?[^\s\p{L}\p{N}]+[\r\n]* | \s*[\r\n]+ | \s+(?!\S) | \s+
    // GPT-2: 's|'t|'re|'ve|'m|'ll|'d | ?\p{L}+ | ?\p{N}+ | ?[^\s\p{L}\p{N}]+ | \s+(?!\S) | \s+
    void pre_split(const std::string& text, std::vector<std::pair<size_t, size_t>>& out) const {
        std::vector<uint32_t> cp; std::vector<size_t> off;
        for (size_t i = 0; i < text.size();) { off.push_back(i); cp.push_back(utf8_get(text, i)); }
        off.push_back(text.size());
        const size_t n = cp.size();
        auto lower = [](uint32_t c) { return c >= 'A' && c <= 'Z' ? c + 32 : c; };
        size_t i = 0;
        while (i < n) {
            size_t j = i;
            auto emit = [&](size_t e) { out.push_back({off[i], off[e]}); i = e; };
            if (cp[i] == '\'' && i + 1 < n) {                                   // contractions
                const uint32_t a = llama3_split ? lower(cp[i + 1]) : cp[i + 1], b = i + 2 < n ? (llama3_split ? lower(cp[i + 2]) : cp[i + 2]) : 0;
                if (a == 's' || a == 't' || a == 'm' || a == 'd') { emit(i + 2); continue; }
                if ((a == 'r' && b == 'e') || (a == 'v' && b == 'e') || (a == 'l' && b == 'l')) { emit(i + 3); continue; }
            }
            if (llama3_split) {
                j = i; if (cp[j] != '\r' && cp[j] != '\n' && !is_letter(cp[j]) && !is_digit(cp[j]) && j + 1 < n && is_letter(cp[j + 1])) j++;
                if (is_letter(cp[j])) { while (j < n && is_letter(cp[j])) j++; emit(j); continue; }
                if (is_digit(cp[i])) { j = i; while (j < n && j < i + 3 && is_digit(cp[j])) j++; emit(j); continue; }
                j = i; if (cp[j] == ' ' && j + 1 < n) j++;
                if (!is_space(cp[j]) && !is_letter(cp[j]) && !is_digit(cp[j])) {
                    while (j < n && !is_space(cp[j]) && !is_letter(cp[j]) && !is_digit(cp[j])) j++;
                    while (j < n && (cp[j] == '\r' || cp[j] == '\n')) j++;
                    emit(j); continue;
                }
            } else {
                j = i; if (cp[j] == ' ' && j + 1 < n) j++;
                if (is_letter(cp[j])) { while (j < n && is_letter(cp[j])) j++; emit(j); continue; }
                if (is_digit(cp[j])) { while (j < n && is_digit(cp[j])) j++; emit(j); continue; }
                if (!is_space(cp[j])) { while (j < n && !is_space(cp[j]) && !is_letter(cp[j]) && !is_digit(cp[j])) j++; emit(j); continue; }
            }
            // whitespace run starting at i
            j = i; while (j < n && is_space(cp[j])) j++;
            if (llama3_split) {                                                 // \s*[\r\n]+ : up to and including the run's last newline
                size_t last_nl = n; for (size_t k = i; k < j; k++) if (cp[k] == '\r' || cp[k] == '\n') last_nl = k;
                if (last_nl != n) { emit(last_nl + 1); continue; }
            }
            if (j < n && j - i > 1) { emit(j - 1); continue; }                 // \s+(?!\S): leave the last space to the next word
            emit(j);                                                            // end of text, or a single space before a word (\s+)
        }
    }

    void bpe(const std::string& word, std::vector<int32_t>& out) const {
        std::vector<std::string> sym;
        for (unsigned char c : word) sym.push_back(b2u[c]);
        while (sym.size() > 1) {
            int best = INT32_MAX; size_t at = 0;
            for (size_t k = 0; k + 1 < sym.size(); k++) { auto it = rank.find(sym[k] + " " + sym[k + 1]); if (it != rank.end() && it->second < best) { best = it->second; at = k; } }
            if (best == INT32_MAX) break;
            sym[at] += sym[at + 1]; sym.erase(sym.begin() + (long)at + 1);
        }
        for (auto& s : sym) {
            auto it = vocab.find(s);
            if (it != vocab.end()) { out.push_back(it->second); continue; }
            size_t k = 0;                                                       // not in the vocabulary as a unit: its characters (byte tokens always exist)
            while (k < s.size()) { size_t k0 = k; utf8_get(s, k); auto b = vocab.find(s.substr(k0, k - k0)); if (b != vocab.end()) out.push_back(b->second); }
        }
    }

    // SentencePiece BPE as llama.cpp's "llama" vocabulary applies it (llm_tokenizer_spm): spaces become U+2581 (one is prepended
    // to a fragment that starts the text or follows a special token when add_space_prefix), the text is cut into UTF-8
    // characters, and the adjacent pair whose concatenation is a piece with the highest score is merged until none is left
    // (ties: the leftmost); what is still not a piece falls back to <0xXX> byte tokens.
    void spm(const std::string& frag, bool prefix, std::vector<int32_t>& out) const {
        std::string t;
        if (prefix) t += "\xE2\x96\x81";
        for (char c : frag) { if (c == ' ') t += "\xE2\x96\x81"; else t.push_back(c); }
        struct Sym { size_t off, len; int prev, next; };
        std::vector<Sym> sym;
        for (size_t i = 0; i < t.size();) { size_t i0 = i; utf8_get(t, i); sym.push_back({i0, i - i0, (int)sym.size() - 1, -1}); if (sym.size() > 1) sym[sym.size() - 2].next = (int)sym.size() - 1; }
        struct Bi { float sc; int left, right; size_t len; };
        auto worse = [](const Bi& a, const Bi& b) { return a.sc < b.sc || (a.sc == b.sc && a.left > b.left); };
        std::priority_queue<Bi, std::vector<Bi>, decltype(worse)> pq(worse);
        auto try_add = [&](int l, int r) {
            if (l < 0 || r < 0) return;
            auto it = vocab.find(t.substr(sym[l].off, sym[l].len + sym[r].len));
            if (it != vocab.end() && (size_t)it->second < score.size()) pq.push({score[it->second], l, r, sym[l].len + sym[r].len});
        };
        for (int i = 1; i < (int)sym.size(); i++) try_add(i - 1, i);
        while (!pq.empty()) {
            const Bi b = pq.top(); pq.pop();
            Sym& L = sym[b.left]; Sym& R = sym[b.right];
            if (L.len == 0 || R.len == 0 || L.len + R.len != b.len || L.next != b.right) continue;       // one side was merged away since
            L.len += R.len; R.len = 0; L.next = R.next; if (R.next >= 0) sym[R.next].prev = b.left;
            try_add(L.prev, b.left); try_add(b.left, L.next);
        }
        for (int i = sym.empty() ? -1 : 0; i >= 0; i = sym[i].next) {
            if (sym[i].len == 0) continue;
            const std::string p = t.substr(sym[i].off, sym[i].len);
            auto it = vocab.find(p);
            if (it != vocab.end()) { out.push_back(it->second); continue; }
            for (unsigned char c : p) { const int bt = byte_tok[c]; if (bt >= 0) out.push_back(bt); }
        }
    }

    std::vector<int32_t> tokenize(const std::string& text, bool parse_special) const {
        std::vector<int32_t> out;
        bool after_special = true;                                              // SPM: the text start counts as "after a special token" (BOS)
        auto plain = [&](const std::string& seg) {
            if (seg.empty()) return;
            if (kind == SPM) { spm(seg, add_space_prefix && after_special, out); after_special = false; return; }
            if (kind == BPE) { std::vector<std::pair<size_t, size_t>> parts; pre_split(seg, parts); for (auto& p : parts) bpe(seg.substr(p.first, p.second - p.first), out); }
            else for (unsigned char c : seg) { const int t = byte_tok[c]; out.push_back(t >= 0 ? t : (int)c % std::max(1, n_vocab)); }
        };
        if (!parse_special || special_sorted.empty()) { plain(text); return out; }
        size_t i = 0, start = 0;
        while (i < text.size()) {
            bool hit = false;
            if (text[i] == '<' || kind == BPE) for (const std::string& sp : special_sorted)
                if (text.compare(i, sp.size(), sp) == 0) { plain(text.substr(start, i - start)); out.push_back(special.at(sp)); after_special = true; i += sp.size(); start = i; hit = true; break; }
            if (!hit) i++;
        }
        plain(text.substr(start));
        return out;
    }
    std::string detokenize(const int32_t* toks, int n) const {
        std::string s;
        for (int i = 0; i < n; i++) if (toks[i] >= 0 && (size_t)toks[i] < piece.size()) s += piece[toks[i]];
        if (kind == SPM && add_space_prefix && !s.empty() && s[0] == ' ') s.erase(0, 1);      // the dummy prefix of the first fragment
        return s;
    }
    int special_id(const std::string& s) const { auto it = special.find(s); return it == special.end() ? -1 : it->second; }
    // token that ends an assistant turn in this template (-1: none beyond EOS)
    int stop_token() const {
        switch (tpl) { case TPL_LLAMA2: return eos; case TPL_LLAMA3: return special_id("<|eot_id|>"); case TPL_GRANITE: return special_id("<|end_of_text|>"); case TPL_CHATML: return special_id("<|im_end|>"); default: return special_id("<|end|>"); }
    }
};

}  // namespace b200
