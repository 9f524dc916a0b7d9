// handlers.cpp — C++ mirror of the reference's subject handlers and envelope
// (/root/reference/nats_llm_studio.go:181-364): same payloads, same validation order, same error
// strings, same {ok,error,data} bytes as Go's json.Marshal would produce.  The Go worker a
// maintainer builds calls the same engine through cgo; this mirror exists
// because no Go toolchain is present on the build/GPU image, and it is what tests/ and
// bench.py's e2e leg drive.  Also: the byte-level tokenizer and the chat rendering.
#include <algorithm>
#include <atomic>
#include <ctime>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include "engine.hpp"
#include "json.hpp"

namespace b200 {

Registry& registry() { static Registry r; return r; }

// ------------------------------------------------------------------------------------ tokenizer
std::string tokenizer_from_gguf(const GgufMeta& meta, int n_vocab, Tokenizer& out) {
    std::vector<std::string> tokens, merges; std::vector<int> types;
    if (auto v = meta.find("tokenizer.ggml.tokens")) tokens = v->arr_str;
    if (auto v = meta.find("tokenizer.ggml.merges")) merges = v->arr_str;
    std::vector<float> scores;
    if (auto v = meta.find("tokenizer.ggml.scores")) { scores.assign(v->arr_num.size(), 0.f); for (size_t i = 0; i < v->arr_num.size(); i++) scores[i] = (float)v->arr_num[i]; }
    if (auto v = meta.find("tokenizer.ggml.token_type")) { types.assign(v->arr_num.size(), 0); for (size_t i = 0; i < v->arr_num.size(); i++) types[i] = (int)v->arr_num[i]; }
    out.build(tokens, types, merges, meta.str("tokenizer.ggml.model"), meta.str("tokenizer.ggml.pre"), meta.str("tokenizer.chat_template"),
              (int)meta.num("tokenizer.ggml.bos_token_id", -1), (int)meta.num("tokenizer.ggml.eos_token_id", -1), meta.num("tokenizer.ggml.add_bos_token", 1) != 0,
              n_vocab > 0 ? n_vocab : (int)tokens.size(), scores, meta.num("tokenizer.ggml.add_space_prefix", 1) != 0);
    return "";
}
void Model::build_tokenizer() { tokenizer_from_gguf(meta, n_vocab, tok); }
std::vector<int32_t> Model::tokenize(const std::string& text, bool parse_special) const { return tok.tokenize(text, parse_special); }
std::string Model::detokenize(const int32_t* toks, int n) const { return tok.detokenize(toks, n); }

static std::string quant_name(int ftype) {
    switch (ftype) { case 7: return "Q8_0"; case 14: return "Q4_K_S"; case 15: return "Q4_K_M"; case 16: return "Q5_K_S"; case 17: return "Q5_K_M"; case 18: return "Q6_K"; case 1: return "F16"; case 0: return "F32"; default: return "unknown"; }
}
static std::string publisher_of(const std::string& id) { size_t k = id.find('/'); return k == std::string::npos ? "local" : id.substr(0, k); }

static Json model_entry(const Model& m) {
    Json e = Json::object();
    e.set("id", Json::string(m.id)).set("object", Json::string("model")).set("type", Json::string("llm"))
        .set("publisher", Json::string(publisher_of(m.id))).set("arch", Json::string(m.arch)).set("compatibility_type", Json::string("gguf"))
        .set("quantization", Json::string(quant_name((int)m.meta.num("general.file_type", -1)))).set("state", Json::string("loaded"))
        .set("max_context_length", Json::integer(m.n_ctx_train));
    return e;
}
std::string Model::info_json() const {
    Json e = model_entry(*this);
    e.set("n_layer", Json::integer(n_layer)).set("n_embd", Json::integer(n_embd)).set("n_head", Json::integer(n_head)).set("n_head_kv", Json::integer(n_head_kv))
        .set("n_ff", Json::integer(n_ff)).set("n_vocab", Json::integer(n_vocab)).set("weight_bytes", Json::integer((long long)weight_bytes))
        .set("hbm_bytes", Json::integer((long long)hbm_bytes)).set("grid", Json::integer(grid)).set("ring_bytes", Json::integer(ring_bytes))
        .set("n_phases", Json::integer(n_phases)).set("n_splits", Json::integer(n_splits))
        .set("tp_rank", Json::integer(engine().tp_rank)).set("tp_world", Json::integer(engine().tp_world))
        .set("embd_bytes", Json::integer((long long)(meta.tensor("token_embd.weight") ? meta.tensor("token_embd.weight")->nbytes : 0)))
        .set("kv_bytes_per_pos", Json::integer((long long)n_layer * 2 * n_head_kv * head_dim * 2));     // this rank's KV shard, fp16
    return json_dump(e);
}

// ------------------------------------------------------------------------------------ envelope (nats_llm_studio.go:186-226)
// NATSResponse{OK bool `json:"ok"`; Error string `json:"error,omitempty"`; Data interface{} `json:"data,omitempty"`}
static std::string envelope(bool ok, const std::string& error, const Json* data /* nullptr = nil interface (omitted) */) {
    std::string out = ok ? "{\"ok\":true" : "{\"ok\":false";
    if (!error.empty()) { out += ",\"error\":"; json_escape(error, out); }
    if (data) { out += ",\"data\":"; json_write(*data, out); }
    out.push_back('}');
    return out;
}
// respondError(msg, err, nil): a typed-nil map inside the interface is NOT omitted -> "data":null
static std::string respond_error_nil(const std::string& err) { Json n = Json::null(); return envelope(false, err, &n); }

// Go-style decode of {"<field>": string} payloads (json.Unmarshal into a struct with one string field)
static bool decode_string_field(const std::string& payload, const char* field, const char* go_struct, const char* go_type_for_value,
                                std::string& value, std::string& err) {
    Json doc; JsonParser ps(payload);
    if (!ps.parse(doc)) { err = ps.err; return false; }
    value.clear();
    if (doc.kind == Json::Null) return true;
    if (doc.kind != Json::Obj) { err = std::string("json: cannot unmarshal ") + doc.go_kind() + " into Go value of type " + go_type_for_value; return false; }
    const Json* f = doc.get_fold(field);
    if (!f || f->kind == Json::Null) return true;
    if (f->kind != Json::Str) { err = std::string("json: cannot unmarshal ") + f->go_kind() + " into Go struct field " + go_struct + "." + field + " of type string"; return false; }
    value = f->s; return true;
}

// ------------------------------------------------------------------------------------ list / pull / delete
static Json list_models() {
    Json list = Json::array();
    { Registry& R = registry(); std::lock_guard<std::mutex> lk(R.mu); for (auto& kv : R.models) if (kv.second->committed) list.push(model_entry(*kv.second)); }
    Json models = Json::object(); models.set("data", list).set("object", Json::string("list"));
    return models;
}
std::string list_models_json() { return json_dump(list_models()); }

static std::string on_list_models() {                                   // onListModels :228-248
    Json models = list_models();
    Json data = Json::object(true); data.set("http_status", Json::integer(200)).set("models", models);
    return envelope(true, "", &data);
}

static bool file_exists(const std::string& p) { struct stat st; return stat(p.c_str(), &st) == 0 && S_ISREG(st.st_mode); }

// A .gguf file -> HBM: reader threads fill a ring of pinned buffers (pread), the calling thread hands the buffers to the
// engine in file order (upload_async: H2D copy queued, tensors repacked on the GPU as they complete) and a buffer is
// refilled once its copy has finished.  Reading, copying and repacking overlap; the host side was the bound of a
// single-threaded read -> copy -> wait loop (6 GB/s).
int load_file_into(const std::string& id, const std::string& path, std::shared_ptr<Model>& out, std::string& log) {
    auto t0 = std::chrono::steady_clock::now();
    if (int rc0 = engine().ensure()) return rc0;                        // no GPU -> fail loudly, nothing else happens
    const int fd = ::open(path.c_str(), O_RDONLY);
    if (fd < 0) return fail(B200_ERR_NOT_FOUND, "cannot open " + path);
    struct stat st; if (fstat(fd, &st) != 0) { ::close(fd); return fail(B200_ERR_NOT_FOUND, "cannot stat " + path); }
    const uint64_t fsize = (uint64_t)st.st_size;
    auto m = std::make_shared<Model>();
    int rc = B200_OK;
    {   // header (may exceed one read: a 128k-token vocabulary is ~5 MB)
        std::vector<uint8_t> head; size_t want = 1u << 20;
        for (;;) {
            head.resize((size_t)std::min<uint64_t>(want, fsize));
            const ssize_t g = pread(fd, head.data(), head.size(), 0);
            if (g < 0) { ::close(fd); return fail(B200_ERR_INVALID, "read error on " + path); }
            head.resize((size_t)g);
            rc = m->open(id, head.data(), head.size());
            if (rc != B200_ERR_MORE || head.size() >= fsize) break;
            want *= 4;
        }
        if (rc != B200_OK) { ::close(fd); return rc; }
    }
    const size_t CH = 32u << 20; int nt_env = 0; if (const char* e = getenv("B200_LOAD_THREADS")) nt_env = atoi(e);
    const int NT = nt_env > 0 ? std::min(nt_env, 32) : std::max(2, std::min(12, (int)std::thread::hardware_concurrency() * 3 / 4)), NB = NT + 2;      // 70B on a 16-core box: 8 readers 13.7 GB/s, 12: 21.3, 16: 16.9      // a reader does ~3 GB/s out of the page cache
    const uint64_t n_chunks = (fsize + CH - 1) / CH;
    struct Buf { uint8_t* p = nullptr; size_t n = 0; int state = 0; uint64_t chunk = 0; cudaEvent_t ticket = nullptr; };   // 0 free, 1 reading, 2 ready, 3 copying
    std::vector<Buf> bufs((size_t)NB);
    for (auto& b : bufs) { b.p = (uint8_t*)b200_pinned_alloc(CH); if (!b.p) rc = fail(B200_ERR_NOMEM, "pinned staging allocation failed"); }
    std::mutex mu; std::condition_variable cv; uint64_t next_read = 0; bool abort_all = false; const int dev = engine().device;
    auto reader = [&]() {
        cudaSetDevice(dev);
        for (;;) {
            uint64_t c; Buf* b;
            {
                std::unique_lock<std::mutex> lk(mu);
                if (abort_all || next_read >= n_chunks) return;
                c = next_read; b = &bufs[(size_t)(c % NB)];
                cv.wait(lk, [&] { return abort_all || (next_read == c && (b->state == 0 || b->state == 3)); });   // chunk c's turn, its buffer handed back
                if (abort_all || next_read != c) continue;
                next_read++;
                cudaEvent_t tk = b->state == 3 ? b->ticket : nullptr; b->state = 1; b->ticket = nullptr;
                lk.unlock(); cv.notify_all();
                if (tk) { cudaEventSynchronize(tk); cudaEventDestroy(tk); }                              // the copy out of this buffer has finished
            }
            const uint64_t off = c * CH; const size_t n = (size_t)std::min<uint64_t>(CH, fsize - off);
            size_t got = 0; while (got < n) { const ssize_t g = pread(fd, b->p + got, n - got, (off_t)(off + got)); if (g <= 0) break; got += (size_t)g; }
            { std::lock_guard<std::mutex> lk(mu); b->n = got; b->chunk = c; b->state = got == n ? 2 : -1; }
            cv.notify_all();
        }
    };
    std::vector<std::thread> th;
    if (rc == B200_OK) for (int i = 0; i < NT; i++) th.emplace_back(reader);
    for (uint64_t c = 0; c < n_chunks && rc == B200_OK; c++) {
        Buf& b = bufs[(size_t)(c % NB)];
        { std::unique_lock<std::mutex> lk(mu); cv.wait(lk, [&] { return (b.state == 2 || b.state == -1) && b.chunk == c; }); if (b.state == -1) rc = fail(B200_ERR_INVALID, "read error on " + path); }
        if (rc != B200_OK) break;
        cudaEvent_t tk = nullptr;
        rc = m->upload_async(c * CH, b.p, b.n, &tk);
        { std::lock_guard<std::mutex> lk(mu); b.ticket = tk; b.state = 3; }
        cv.notify_all();
    }
    { std::lock_guard<std::mutex> lk(mu); abort_all = true; }
    cv.notify_all();
    for (auto& t : th) t.join();
    ::close(fd);
    if (rc == B200_OK) rc = m->commit();                                 // waits for the copies and the repacks
    cudaStreamSynchronize(engine().copy_stream);
    for (auto& b : bufs) { if (b.ticket) cudaEventDestroy(b.ticket); b200_pinned_free(b.p); }
    if (rc != B200_OK) return rc;
    double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    char buf[256]; snprintf(buf, sizeof buf, "loaded %s: %zu tensors, %llu weight bytes resident in HBM, %.0f ms", id.c_str(), m->meta.tensors.size(), (unsigned long long)m->weight_bytes, ms);
    log = buf;
    { Registry& R = registry(); std::lock_guard<std::mutex> lk(R.mu); R.models[id] = m; }
    out = m;
    return B200_OK;
}

static std::string on_pull_model(const std::string& payload) {          // OnPullModel :250-286
    std::string id, err;
    if (!decode_string_field(payload, "identifier", "PullModelRequest", "nats_llm_studio.PullModelRequest", id, err)) return respond_error_nil("invalid JSON in PullModel: " + err);
    if (id.empty()) return respond_error_nil("'identifier' is required");
    std::string dir; { Registry& R = registry(); std::lock_guard<std::mutex> lk(R.mu); dir = R.models_dir; }
    std::string path;
    for (const std::string& c : {dir + "/" + id + ".gguf", dir + "/" + id, dir + "/" + id + "/model.gguf", id}) if (file_exists(c)) { path = c; break; }
    Json data = Json::object(true); data.set("model", Json::string(id));
    if (path.empty()) { data.set("output", Json::string("")); return envelope(false, "failed to pull model '" + id + "': no GGUF file under " + dir, &data); }
    std::shared_ptr<Model> m; std::string log;
    if (load_file_into(id, path, m, log) != B200_OK) { data.set("output", Json::string("")); return envelope(false, "failed to pull model '" + id + "': " + g_last_error, &data); }
    data.set("output", Json::string(log));
    return envelope(true, "", &data);
}

static std::string on_delete_model(const std::string& payload) {        // OnDeleteModel :288-324, DeleteModel :99-133
    std::string id, err;
    if (!decode_string_field(payload, "model_id", "DeleteModelRequest", "nats_llm_studio.DeleteModelRequest", id, err)) return respond_error_nil("invalid JSON in DeleteModel: " + err);
    if (id.empty()) return respond_error_nil("'model_id' is required");
    std::shared_ptr<Model> m;
    { Registry& R = registry(); std::lock_guard<std::mutex> lk(R.mu); auto it = R.models.find(id); if (it != R.models.end()) { m = it->second; R.models.erase(it); } }
    const std::string where = "hbm://" + id;
    Json data = Json::object(true); data.set("model_id", Json::string(id));
    if (!m) { data.set("dir", Json::string(where)); return envelope(false, "model directory not found: " + where, &data); }
    m->release_locked();
    data.set("deleted_dir", Json::string(where));
    return envelope(true, "", &data);
}

// ------------------------------------------------------------------------------------ chat
// The model's chat template applied to `messages` (OpenAI-style [{role, content}]), ending with the header of the
// assistant's turn.  Families are recognised from tokenizer.chat_template's marker tokens (tokenizer.hpp); the text of a
// message is tokenised as plain text (special-token strings inside user content stay text).
static std::string render_messages(const Tokenizer& tk, const Json& messages, std::vector<int32_t>& toks) {
    if (messages.kind != Json::Arr) return "'messages' must be an array";
    auto sp = [&](const std::string& s) { const int id = tk.special_id(s); if (id >= 0) toks.push_back(id); else { auto t = tk.tokenize(s, false); toks.insert(toks.end(), t.begin(), t.end()); } };
    auto text = [&](const std::string& s) { auto t = tk.tokenize(s, false); toks.insert(toks.end(), t.begin(), t.end()); };
    if (tk.tpl == Tokenizer::TPL_LLAMA2) {
        // Llama-2 / Mistral layout: <s>[INST] <<SYS>>\n{system}\n<</SYS>>\n\n{user} [/INST] {assistant} </s><s>[INST] ...   (BOS per user turn)
        std::string sys;
        for (const Json& msg : messages.arr) {
            if (msg.kind != Json::Obj) return "each message must be an object";
            const Json* role = msg.get("role"); const Json* content = msg.get("content");
            if (!role || role->kind != Json::Str || !content || content->kind != Json::Str) return "each message needs string 'role' and 'content'";
            if (role->s == "system") sys = "<<SYS>>\n" + content->s + "\n<</SYS>>\n\n";
            else if (role->s == "assistant") { text(content->s); if (tk.eos >= 0) toks.push_back(tk.eos); }
            else { if (tk.bos >= 0) toks.push_back(tk.bos); text("[INST] " + sys + content->s + " [/INST]"); sys.clear(); }
        }
        return "";
    }
    if (tk.tpl == Tokenizer::TPL_LLAMA3) sp("<|begin_of_text|>");
    else if (tk.bos >= 0 && tk.add_bos && tk.tpl == Tokenizer::TPL_MINIMAL) toks.push_back(tk.bos);
    for (const Json& msg : messages.arr) {
        if (msg.kind != Json::Obj) return "each message must be an object";
        const Json* role = msg.get("role"); const Json* content = msg.get("content");
        if (!role || role->kind != Json::Str || !content || content->kind != Json::Str) return "each message needs string 'role' and 'content'";
        switch (tk.tpl) {
            case Tokenizer::TPL_LLAMA3: sp("<|start_header_id|>"); text(role->s); sp("<|end_header_id|>"); text("\n\n"); text(content->s); sp("<|eot_id|>"); break;     // (the template trims content; so do we not: verbatim)
            case Tokenizer::TPL_GRANITE: sp("<|start_of_role|>"); text(role->s); sp("<|end_of_role|>"); text(content->s); sp("<|end_of_text|>"); text("\n"); break;
            case Tokenizer::TPL_CHATML: sp("<|im_start|>"); text(role->s + "\n" + content->s); sp("<|im_end|>"); text("\n"); break;
            default: sp("<|" + role->s + "|>"); text(content->s); sp("<|end|>"); break;
        }
    }
    switch (tk.tpl) {
        case Tokenizer::TPL_LLAMA3: sp("<|start_header_id|>"); text("assistant"); sp("<|end_header_id|>"); text("\n\n"); break;
        case Tokenizer::TPL_GRANITE: sp("<|start_of_role|>"); text("assistant"); sp("<|end_of_role|>"); break;
        case Tokenizer::TPL_CHATML: sp("<|im_start|>"); text("assistant\n"); break;
        default: sp("<|assistant|>"); break;
    }
    return "";
}
static std::string render_prompt(const Model& m, const Json& messages, std::vector<int32_t>& toks) { return render_messages(m.tok, messages, toks); }
std::string render_chat(const Tokenizer& tok, const std::string& messages_json, std::vector<int32_t>& toks) {
    Json doc; JsonParser ps(messages_json);
    if (!ps.parse(doc)) return "invalid JSON: " + ps.err;
    return render_messages(tok, doc, toks);
}

static Json lm_error(const std::string& message, const char* type, const char* param, const char* code) {
    Json e = Json::object(); e.set("message", Json::string(message)).set("type", Json::string(type)).set("param", param ? Json::string(param) : Json::null()).set("code", code ? Json::string(code) : Json::null());
    Json r = Json::object(); r.set("error", e); return r;
}

static std::string on_chat_model(const std::string& payload) {          // OnChatModel :327-364
    if (payload.empty()) return respond_error_nil("payload vazio em ChatModel");
    Json doc; JsonParser ps(payload);
    if (!ps.parse(doc)) return respond_error_nil("invalid JSON in ChatModel: " + ps.err);
    std::string model_id;
    if (doc.kind == Json::Obj) {
        const Json* f = doc.get_fold("model");
        if (f && f->kind != Json::Null) { if (f->kind != Json::Str) return respond_error_nil(std::string("invalid JSON in ChatModel: json: cannot unmarshal ") + f->go_kind() + " into Go struct field .model of type string"); model_id = f->s; }
    } else if (doc.kind != Json::Null) return respond_error_nil(std::string("invalid JSON in ChatModel: json: cannot unmarshal ") + doc.go_kind() + " into Go value of type struct { Model string \"json:\\\"model\\\"\" }");
    if (model_id.empty()) return respond_error_nil("'model' is required in ChatModel");

    auto reply = [&](int status, const Json& response) { Json data = Json::object(true); data.set("http_status", Json::integer(status)).set("response", response); return envelope(true, "", &data); };
    std::shared_ptr<Model> m;
    { Registry& R = registry(); std::lock_guard<std::mutex> lk(R.mu); auto it = R.models.find(model_id); if (it != R.models.end() && it->second->committed) m = it->second; }
    // the reference forwards whatever the backend answers and never inspects the status (:173-178, :356-363)
    if (!m) return reply(404, lm_error("model '" + model_id + "' not found", "invalid_request_error", "model", "model_not_found"));
    const Json* messages = doc.get("messages");
    if (!messages) return reply(400, lm_error("'messages' field is required", "invalid_request_error", "messages", nullptr));
    std::vector<int32_t> toks; std::string terr = render_prompt(*m, *messages, toks);
    if (!terr.empty()) return reply(400, lm_error(terr, "invalid_request_error", "messages", nullptr));
    auto numf = [&](const char* k, double d) { const Json* v = doc.get(k); return v && v->kind == Json::Num ? v->num : d; };
    b200_sampling s{}; s.temperature = (float)numf("temperature", 0.8); s.top_k = (int)numf("top_k", 40); s.top_p = (float)numf("top_p", 0.95);
    const Json* seed = doc.get("seed"); s.seed = seed && seed->kind == Json::Num ? (uint64_t)seed->num : (uint64_t)std::chrono::steady_clock::now().time_since_epoch().count();
    int max_tokens = (int)numf("max_tokens", -1);
    const int room = engine().max_ctx - (int)toks.size() - 1;
    if (room < 1) return reply(400, lm_error("prompt of " + std::to_string(toks.size()) + " tokens exceeds the context window", "invalid_request_error", "messages", "context_length_exceeded"));
    if (max_tokens <= 0 || max_tokens > room) max_tokens = std::min(room, max_tokens <= 0 ? 512 : max_tokens);
    s.max_tokens = max_tokens; s.stop_token = m->tok.stop_token();
    const Json* ign = doc.get("ignore_eos"); s.ignore_eos = ign && ign->kind == Json::Bool && ign->b;
    // the reference bounds the handler with a 2-minute context (:328); a generation that outlives it is answered the way
    // a cancelled context would be: ok:false with Go's deadline error text.  B200_CHAT_TIMEOUT_MS overrides (tests).
    s.timeout_ms = 120000; if (const char* e = getenv("B200_CHAT_TIMEOUT_MS")) s.timeout_ms = std::max(1, atoi(e));
    std::vector<int32_t> out; float ttft = 0, dec = 0; int reason = 0;
    auto t0 = std::chrono::steady_clock::now();
    // max_seqs > 1: the request joins the continuous-batching scheduler, so concurrent chat requests share batched steps
    if (sched_generate(m, toks, s, out, &ttft, &dec, &reason) != B200_OK)
        return reply(500, lm_error(g_last_error, "internal_error", nullptr, nullptr));
    if (reason == B200_FINISH_TIMEOUT || reason == B200_FINISH_CANCELLED) {
        Json data = Json::object(true); data.set("http_status", Json::integer(0));
        return envelope(false, std::string("error calling LM Studio: ") + (reason == B200_FINISH_TIMEOUT ? "context deadline exceeded" : "context canceled"), &data);
    }
    const int n_out = (int)out.size();
    double wall_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    bool stopped = reason == B200_FINISH_STOP;
    int n_text = stopped ? n_out - 1 : n_out;
    // OpenAI `stop`: a string or an array of strings; the text ends before the first occurrence of any of them (the stop string
    // itself is not returned) and finish_reason is "stop".  Matched on the detokenised text, so a stop string may span tokens.
    std::vector<std::string> stops;
    if (const Json* st = doc.get("stop")) {
        if (st->kind == Json::Str && !st->s.empty()) stops.push_back(st->s);
        else if (st->kind == Json::Arr) for (const Json& e : st->arr) if (e.kind == Json::Str && !e.s.empty()) stops.push_back(e.s);
    }
    std::string content; bool stop_string = false; int n_used = n_out;
    if (!stops.empty()) {
        for (int i = 0; i < n_text && !stop_string; i++) {
            const std::string acc = m->detokenize(out.data(), i + 1);
            size_t cut = std::string::npos;
            for (const std::string& sx : stops) cut = std::min(cut, acc.find(sx));
            if (cut != std::string::npos) { content = acc.substr(0, cut); stop_string = true; n_used = i + 1; }
        }
    }
    if (!stop_string) content = m->detokenize(out.data(), n_text);
    else stopped = true;
    static std::atomic<uint64_t> counter{0};
    char idbuf[64]; snprintf(idbuf, sizeof idbuf, "chatcmpl-%012llx", (unsigned long long)(++counter));
    Json message = Json::object(); message.set("role", Json::string("assistant")).set("content", Json::string(content));
    Json choice = Json::object(); choice.set("index", Json::integer(0)).set("logprobs", Json::null()).set("finish_reason", Json::string(stopped ? "stop" : "length")).set("message", message);
    Json choices = Json::array(); choices.push(choice);
    Json usage = Json::object(); usage.set("prompt_tokens", Json::integer((long long)toks.size())).set("completion_tokens", Json::integer(n_used)).set("total_tokens", Json::integer((long long)toks.size() + n_used));
    Json stats = Json::object();
    stats.set("tokens_per_second", Json::number(n_out > 1 && dec > 0 ? (n_out - 1) / (dec * 1e-3) : 0.0)).set("time_to_first_token", Json::number(ttft * 1e-3))
        .set("generation_time", Json::number(wall_s)).set("stop_reason", Json::string(stop_string ? "stopStringFound" : stopped ? "eosFound" : "maxPredictedTokensReached"));
    Json info = Json::object(); info.set("arch", Json::string(m->arch)).set("quant", Json::string(quant_name((int)m->meta.num("general.file_type", -1)))).set("format", Json::string("gguf")).set("context_length", Json::integer(engine().max_ctx));
    Json fmts = Json::array(); fmts.push(Json::string("gguf"));
    Json runtime = Json::object(); runtime.set("name", Json::string("b200-gguf-engine")).set("version", Json::string("0.1.0")).set("supported_formats", fmts);
    Json resp = Json::object();
    resp.set("id", Json::string(idbuf)).set("object", Json::string("chat.completion")).set("created", Json::integer((long long)time(nullptr))).set("model", Json::string(model_id))
        .set("choices", choices).set("usage", usage).set("stats", stats).set("model_info", info).set("runtime", runtime);
    return reply(200, resp);
}

std::string handle_subject(const std::string& subject, const std::string& payload) {
    if (subject == "lmstudio.chat_model") return on_chat_model(payload);
    if (subject == "lmstudio.list_models") return on_list_models();
    if (subject == "lmstudio.pull_model") return on_pull_model(payload);
    if (subject == "lmstudio.delete_model") return on_delete_model(payload);
    return respond_error_nil("unknown subject " + subject);
}

}  // namespace b200
