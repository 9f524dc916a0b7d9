// capi.cpp — extern "C" surface declared in include/b200_engine.h.
#include <thread>
#include <atomic>
#include <chrono>
#include <cstring>

#include "engine.hpp"

namespace b200 {
int load_file_into(const std::string& id, const std::string& path, std::shared_ptr<Model>& out, std::string& log);
std::string list_models_json();

// ---- request queue behind b200_submit / b200_poll (one scheduler thread owns the engine).
// max_seqs == 1: requests run one after another through the persistent batch-1 decode kernel.
// max_seqs  > 1: continuous batching — every iteration admits waiting requests into free sequence slots (their prompts
// are ingested by the tensor-core prefill into the slot's KV pages) and then advances ALL active sequences by one token
// with one batched step; finished sequences free their slot immediately.
// A caller blocked inside the library for one request (lmstudio.chat_model through b200_handle): its events are
// delivered here instead of the b200_poll queue.
struct Waiter { std::mutex mu; std::condition_variable cv; std::vector<int32_t> toks; bool done = false; b200_event fin{}; };
struct Request {
    std::shared_ptr<Model> m; std::vector<int32_t> toks; b200_sampling s; uint64_t id; std::atomic<bool> cancelled{false};
    std::chrono::steady_clock::time_point t_submit;
    std::shared_ptr<Waiter> waiter;
    bool expired() const { return s.timeout_ms > 0 && std::chrono::steady_clock::now() - t_submit > std::chrono::milliseconds(s.timeout_ms); }
};
struct Seq { std::shared_ptr<Request> r; int slot = 0, n_fed = 0, pos = 0, cur_tok = 0, n_gen = 0; bool decoding = false; float ttft_ms = 0; std::chrono::steady_clock::time_point t_first; };
// std::mutex is not fair: a thread that wants run_mu to delete the model could wait behind the scheduler's back-to-back
// steps until the model's requests had all finished.  The scheduler steps aside while such a thread is waiting.
static void yield_to_release(Model& m) { while (m.lock_waiters.load(std::memory_order_acquire) > 0) std::this_thread::sleep_for(std::chrono::microseconds(20)); }

struct Scheduler {
    std::mutex mu; std::condition_variable cv_req, cv_ev;
    std::deque<std::shared_ptr<Request>> q; std::deque<b200_event> ev;
    std::map<uint64_t, std::shared_ptr<Request>> live;
    std::thread th; bool stop = false, started = false;
    void start() { if (started) return; started = true; th = std::thread([this] { loop(); }); }
    void push_events(std::vector<b200_event>& out, const std::vector<uint64_t>& done) {
        {
            std::lock_guard<std::mutex> lk(mu);
            for (auto& e : out) {
                auto it = live.find(e.req_id);
                if (it != live.end() && it->second->waiter) {           // a blocked in-library caller owns this request's events
                    Waiter& w = *it->second->waiter; std::lock_guard<std::mutex> wl(w.mu);
                    if (e.kind == B200_EV_TOKEN) w.toks.push_back(e.token); else { w.fin = e; w.done = true; w.cv.notify_all(); }
                } else ev.push_back(e);
            }
            for (auto id : done) live.erase(id);
        }
        out.clear(); cv_ev.notify_all();
    }
    static double ms_since(std::chrono::steady_clock::time_point t) { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t).count(); }

    void run_single(const std::shared_ptr<Request>& r) {
        b200_event fin{}; fin.req_id = r->id; fin.kind = B200_EV_FINISH; fin.n_prompt = (int)r->toks.size();
        std::vector<b200_event> out;
        if (r->cancelled) fin.finish_reason = B200_FINISH_CANCELLED;
        else {
            int n_gen = r->s.max_tokens > 0 ? r->s.max_tokens : 256;
            std::vector<int32_t> toks((size_t)n_gen); int n_out = 0; float ttft = 0, dec = 0;
            int reason = 0;
            int rc = r->expired() ? B200_OK : r->m->generate(r->toks.data(), (int)r->toks.size(), n_gen, &r->s, toks.data(), &n_out, &ttft, &dec, &reason, &r->t_submit);
            if (rc != B200_OK) { fin.kind = B200_EV_ERROR; fin.finish_reason = rc; }
            else {
                for (int i = 0; i < n_out; i++) { b200_event e{}; e.req_id = r->id; e.kind = B200_EV_TOKEN; e.token = toks[i]; e.n_generated = i + 1; out.push_back(e); }
                fin.finish_reason = reason ? reason : B200_FINISH_TIMEOUT; fin.n_generated = n_out; fin.ttft_ms = ttft; fin.decode_ms = dec;
            }
        }
        out.push_back(fin); push_events(out, {r->id});
    }

    void loop() {
        std::deque<std::shared_ptr<Request>> waiting; std::vector<Seq> active; std::vector<int> free_slots; std::shared_ptr<Model> cur;
        for (;;) {
            {
                std::unique_lock<std::mutex> lk(mu);
                cv_req.wait(lk, [&] { return stop || !q.empty() || !active.empty() || !waiting.empty(); });
                if (stop) return;
                while (!q.empty()) { waiting.push_back(q.front()); q.pop_front(); }
            }
            const int max_seqs = engine().max_seqs, max_ctx = engine().max_ctx;
            if (max_seqs <= 1) { while (!waiting.empty()) { auto r = waiting.front(); waiting.pop_front(); run_single(r); } continue; }
            std::vector<b200_event> out; std::vector<uint64_t> done;
            if (active.empty()) { free_slots.clear(); for (int i = max_seqs - 1; i >= 0; i--) free_slots.push_back(i); cur.reset(); }
            auto finish = [&](Seq& s, int kind, int reason) {
                b200_event f{}; f.req_id = s.r->id; f.kind = kind; f.finish_reason = reason; f.n_prompt = (int)s.r->toks.size(); f.n_generated = s.n_gen; f.ttft_ms = s.ttft_ms;
                f.decode_ms = s.n_gen ? (float)ms_since(s.t_first) : 0.f;
                out.push_back(f); done.push_back(s.r->id); free_slots.push_back(s.slot);
                s.r->m->kv_release(s.slot);                              // its KV pages go back to the pool
            };
            // ---- admit: a waiting request gets a sequence slot; its prompt is ingested in chunks below
            while (!waiting.empty() && !free_slots.empty() && (!cur || waiting.front()->m == cur)) {
                auto r = waiting.front(); waiting.pop_front();
                b200_event e{}; e.req_id = r->id; e.n_prompt = (int)r->toks.size();
                if (r->cancelled || r->expired()) { e.kind = B200_EV_FINISH; e.finish_reason = r->cancelled ? B200_FINISH_CANCELLED : B200_FINISH_TIMEOUT; out.push_back(e); done.push_back(r->id); continue; }
                const int n = (int)r->toks.size();
                bool bad = n < 1 || n + 1 > max_ctx;
                for (int t : r->toks) bad = bad || t < 0 || t >= r->m->n_vocab;
                if (bad) { e.kind = B200_EV_ERROR; e.finish_reason = B200_ERR_INVALID; out.push_back(e); done.push_back(r->id); continue; }
                // KV pages of the whole request (prompt + max_tokens, capped by max_ctx) are reserved at admission, so that no
                // sequence can run out of pages half way; with the default pool every slot is fully backed and this always passes
                const int want_pos = std::min(max_ctx, n + (r->s.max_tokens > 0 ? r->s.max_tokens : 256)), want_pages = (want_pos + 15) / 16;
                if (r->m->kv_free_count() < want_pages) {
                    if (active.empty()) { e.kind = B200_EV_ERROR; e.finish_reason = B200_ERR_NOMEM; out.push_back(e); done.push_back(r->id); continue; }   // can never fit
                    waiting.push_front(r); break;                          // wait for running requests to return pages
                }
                cur = r->m;
                Seq s; s.r = r; s.slot = free_slots.back(); free_slots.pop_back();
                { std::lock_guard<std::mutex> lk(cur->run_mu); if (cur->committed && cur->kv_ensure(s.slot, want_pos - 1) != B200_OK) { free_slots.push_back(s.slot); e.kind = B200_EV_ERROR; e.finish_reason = B200_ERR_NOMEM; out.push_back(e); done.push_back(r->id); continue; } }
                if (n == 1) { s.decoding = true; s.cur_tok = r->toks[0]; s.pos = 0; }
                active.push_back(s);
            }
            if (active.empty()) { push_events(out, done); continue; }
            int rc = B200_OK;
            // ---- one prompt chunk: up to PREFILL_CHUNK rows taken, in arrival order, from the sequences still ingesting
            {
                std::vector<int32_t> toks, slots, pos;
                for (Seq& s : active) {
                    if (s.decoding || s.r->cancelled) continue;
                    const int n = (int)s.r->toks.size(), k = std::min(n - 1 - s.n_fed, Model::PREFILL_CHUNK - (int)toks.size());
                    for (int i = 0; i < k; i++) { toks.push_back(s.r->toks[s.n_fed + i]); slots.push_back(s.slot); pos.push_back(s.n_fed + i); }
                    s.n_fed += k;
                    if (s.n_fed == n - 1) { s.decoding = true; s.cur_tok = s.r->toks[n - 1]; s.pos = n - 1; }
                    if ((int)toks.size() == Model::PREFILL_CHUNK) break;
                }
                if (!toks.empty()) { yield_to_release(*cur); std::lock_guard<std::mutex> lk(cur->run_mu); rc = cur->prefill_rows(toks.data(), slots.data(), pos.data(), (int)toks.size()); }
            }
            // ---- one batched decode step for every sequence whose prompt is in the cache
            std::vector<int> idx;
            for (int i = 0; i < (int)active.size(); i++) if (active[i].decoding && !active[i].r->cancelled) idx.push_back(i);
            const int B = (int)idx.size();
            std::vector<int32_t> toks(B), slots(B), pos(B), next(B), draw(B); std::vector<b200_sampling> samp(B);
            for (int j = 0; j < B; j++) { const Seq& s = active[idx[j]]; toks[j] = s.cur_tok; slots[j] = s.slot; pos[j] = s.pos; samp[j] = s.r->s; draw[j] = s.n_gen; }
            if (rc == B200_OK && B > 0) { yield_to_release(*cur); std::lock_guard<std::mutex> lk(cur->run_mu); rc = cur->batch_step(toks.data(), slots.data(), pos.data(), B, next.data(), nullptr, samp.data(), draw.data()); }
            if (rc != B200_OK) {                                   // engine failure: every active request fails
                for (Seq& s : active) finish(s, B200_EV_ERROR, rc);
                active.clear(); push_events(out, done); continue;
            }
            for (int j = 0; j < B; j++) {
                Seq& s = active[idx[j]]; auto& r = s.r;
                s.n_gen++; s.pos++; s.cur_tok = next[j];
                if (s.n_gen == 1) { s.ttft_ms = (float)ms_since(r->t_submit); s.t_first = std::chrono::steady_clock::now(); }
                b200_event e{}; e.req_id = r->id; e.n_prompt = (int)r->toks.size(); e.kind = B200_EV_TOKEN; e.token = next[j]; e.n_generated = s.n_gen; e.ttft_ms = s.ttft_ms; out.push_back(e);
            }
            std::vector<Seq> keep;
            for (Seq& s : active) {
                auto& r = s.r;
                const int max_tokens = r->s.max_tokens > 0 ? r->s.max_tokens : 256;
                const bool stopped = s.n_gen > 0 && !r->s.ignore_eos && (s.cur_tok == r->m->eos || s.cur_tok == r->s.stop_token);
                if (r->cancelled) finish(s, B200_EV_FINISH, B200_FINISH_CANCELLED);
                else if (r->expired()) finish(s, B200_EV_FINISH, B200_FINISH_TIMEOUT);
                else if (stopped) finish(s, B200_EV_FINISH, B200_FINISH_STOP);
                else if (s.n_gen >= max_tokens || (s.decoding && s.pos + 1 >= max_ctx)) finish(s, B200_EV_FINISH, B200_FINISH_LENGTH);
                else keep.push_back(s);
            }
            active.swap(keep);
            push_events(out, done);
        }
    }
    ~Scheduler() { shutdown(); }
    void shutdown() { { std::lock_guard<std::mutex> lk(mu); stop = true; } cv_req.notify_all(); if (th.joinable()) th.join(); started = false; stop = false; }
};
static Scheduler& sched() { static Scheduler s; return s; }

static void enqueue(const std::shared_ptr<Request>& r) {
    Scheduler& S = sched();
    { std::lock_guard<std::mutex> lk(S.mu); S.start(); S.q.push_back(r); S.live[r->id] = r; }
    S.cv_req.notify_one();
}

// One blocking generation.  max_seqs == 1: the persistent batch-1 kernel, directly.  max_seqs > 1: the request joins the
// continuous-batching scheduler like any b200_submit request (so concurrent lmstudio.chat_model calls share batched
// steps) and this thread sleeps until its FINISH event.
int sched_generate(const std::shared_ptr<Model>& m, const std::vector<int32_t>& prompt, const b200_sampling& s, std::vector<int32_t>& out, float* ttft_ms, float* decode_ms, int* finish_reason) {
    const int n_gen = s.max_tokens > 0 ? s.max_tokens : 256;
    if (engine().max_seqs <= 1) {
        out.assign((size_t)n_gen, 0); int n_out = 0;
        const auto t0 = std::chrono::steady_clock::now();
        int rc = m->generate(prompt.data(), (int)prompt.size(), n_gen, &s, out.data(), &n_out, ttft_ms, decode_ms, finish_reason, &t0);
        out.resize(rc == B200_OK ? (size_t)n_out : 0);
        return rc;
    }
    static std::atomic<uint64_t> next_id{0};
    auto r = std::make_shared<Request>(); r->m = m; r->toks = prompt; r->s = s; r->s.max_tokens = n_gen;
    r->id = (1ull << 63) | ++next_id;                                   // internal id space, apart from b200_submit callers' ids
    r->t_submit = std::chrono::steady_clock::now(); r->waiter = std::make_shared<Waiter>();
    enqueue(r);
    Waiter& w = *r->waiter; std::unique_lock<std::mutex> lk(w.mu);
    w.cv.wait(lk, [&] { return w.done; });
    out = w.toks;
    if (ttft_ms) *ttft_ms = w.fin.ttft_ms; if (decode_ms) *decode_ms = w.fin.decode_ms;
    if (w.fin.kind == B200_EV_ERROR) return fail(w.fin.finish_reason, "generation failed in the batched engine (code " + std::to_string(w.fin.finish_reason) + ")");
    if (finish_reason) *finish_reason = w.fin.finish_reason;
    return B200_OK;
}
}  // namespace b200

using namespace b200;

static int copy_out(const std::string& s, char* out, size_t cap, size_t* need) {
    if (need) *need = s.size();
    if (!out || cap < s.size()) return out ? fail(B200_ERR_MORE, "output buffer too small") : B200_OK;
    memcpy(out, s.data(), s.size());
    if (cap > s.size()) out[s.size()] = 0;
    return B200_OK;
}
static std::shared_ptr<Model> find_shared(const b200_model* m) {
    Registry& R = registry(); std::lock_guard<std::mutex> lk(R.mu);
    for (auto& kv : R.models) if ((const b200_model*)kv.second.get() == m) return kv.second;
    return nullptr;
}

extern "C" {

int b200_init(const b200_config* cfg) { return engine().init(cfg); }
int b200_shutdown(void) {
    sched().shutdown();
    { Registry& R = registry(); std::lock_guard<std::mutex> lk(R.mu); R.models.clear(); }
    engine().shutdown();
    return B200_OK;
}
const char* b200_last_error(void) { return g_last_error.c_str(); }
void* b200_pinned_alloc(size_t n) { void* p = nullptr; if (engine().ensure() != B200_OK) return nullptr; return cudaMallocHost(&p, n) == cudaSuccess ? p : nullptr; }
void b200_pinned_free(void* p) { if (p) cudaFreeHost(p); }

int b200_model_open(const char* id, const void* head, size_t n, b200_model** out) {
    if (!id || !head || !out) return fail(B200_ERR_INVALID, "null argument");
    auto m = std::make_shared<Model>();
    int rc = m->open(id, (const uint8_t*)head, n);
    if (rc != B200_OK) return rc;
    { Registry& R = registry(); std::lock_guard<std::mutex> lk(R.mu); R.models[id] = m; }
    *out = (b200_model*)m.get();
    return B200_OK;
}
int b200_model_upload(b200_model* m, uint64_t off, const void* bytes, size_t n) { if (!m || (!bytes && n)) return fail(B200_ERR_INVALID, "null argument"); return ((Model*)m)->upload(off, bytes, n); }
int b200_model_release_slot(b200_model* m, int slot) {
    if (!m) return fail(B200_ERR_INVALID, "null model");
    Model* M = (Model*)m; if (slot < 0 || slot >= engine().max_seqs) return fail(B200_ERR_INVALID, "slot out of range");
    M->kv_release(slot); return B200_OK;
}
int b200_model_kv_free_pages(b200_model* m) { if (!m) return fail(B200_ERR_INVALID, "null model"); return ((Model*)m)->kv_free_count(); }
int b200_model_commit(b200_model* m) { if (!m) return fail(B200_ERR_INVALID, "null model"); return ((Model*)m)->commit(); }
int b200_model_close(b200_model* m) {
    if (!m) return fail(B200_ERR_INVALID, "null model");
    std::shared_ptr<Model> keep;
    { Registry& R = registry(); std::lock_guard<std::mutex> lk(R.mu); for (auto it = R.models.begin(); it != R.models.end(); ++it) if ((b200_model*)it->second.get() == m) { keep = it->second; R.models.erase(it); break; } }
    if (!keep) return fail(B200_ERR_NOT_FOUND, "model is not registered");
    keep->release_locked();
    return B200_OK;
}
int b200_model_load_file(const char* id, const char* path, b200_model** out) {
    if (!id || !path) return fail(B200_ERR_INVALID, "null argument");
    std::shared_ptr<Model> m; std::string log;
    int rc = load_file_into(id, path, m, log);
    if (rc == B200_OK && out) *out = (b200_model*)m.get();
    return rc;
}
int b200_model_find(const char* id, b200_model** out) {
    if (!id || !out) return fail(B200_ERR_INVALID, "null argument");
    Registry& R = registry(); std::lock_guard<std::mutex> lk(R.mu);
    auto it = R.models.find(id); if (it == R.models.end()) return fail(B200_ERR_NOT_FOUND, std::string("model '") + id + "' is not loaded");
    *out = (b200_model*)it->second.get(); return B200_OK;
}
int b200_model_list(char* out, size_t cap, size_t* need) {
    return copy_out(list_models_json(), out, cap, need);
}
int b200_model_info(const b200_model* m, char* out, size_t cap, size_t* need) { if (!m) return fail(B200_ERR_INVALID, "null model"); return copy_out(((const Model*)m)->info_json(), out, cap, need); }
int b200_model_n_vocab(const b200_model* m) { return m ? ((const Model*)m)->n_vocab : 0; }

int b200_model_generate(b200_model* m, const int32_t* prompt, int n_prompt, int n_gen, const b200_sampling* s, int32_t* out, int* n_out, float* ttft, float* dec) {
    if (!m || !prompt || !out) return fail(B200_ERR_INVALID, "null argument");
    return ((Model*)m)->generate(prompt, n_prompt, n_gen, s, out, n_out, ttft, dec);
}
int b200_model_eval(b200_model* m, const int32_t* toks, int n, float* all, float* last) { if (!m || !toks) return fail(B200_ERR_INVALID, "null argument"); return ((Model*)m)->eval(toks, n, all, last); }
int b200_bench_decode(b200_model* m, int ctx, int steps, int warmup, float* ms, double* bytes) { if (!m) return fail(B200_ERR_INVALID, "null model"); return ((Model*)m)->bench_decode(ctx, steps, warmup, ms, bytes); }

int b200_tokenize(const b200_model* m, const char* text, size_t n, int32_t* out, int cap, int* n_out) {
    if (!m || (!text && n)) return fail(B200_ERR_INVALID, "null argument");
    auto t = ((const Model*)m)->tokenize(std::string(text ? text : "", n), true);
    if (n_out) *n_out = (int)t.size();
    if (!out || cap < (int)t.size()) return out ? fail(B200_ERR_MORE, "token buffer too small") : B200_OK;
    memcpy(out, t.data(), t.size() * 4); return B200_OK;
}
int b200_detokenize(const b200_model* m, const int32_t* toks, int n, char* out, size_t cap, size_t* need) {
    if (!m || (!toks && n)) return fail(B200_ERR_INVALID, "null argument");
    return copy_out(((const Model*)m)->detokenize(toks, n), out, cap, need);
}

// host-only forms (no GPU, no loaded model): the tokenizer and chat template of a GGUF header
static int with_gguf_tokenizer(const void* gguf, size_t n, Tokenizer& tk) {
    if (!gguf) return fail(B200_ERR_INVALID, "null argument");
    GgufMeta meta; std::string err = gguf_parse((const uint8_t*)gguf, n, meta);
    if (err == "need more") return fail(B200_ERR_MORE, "GGUF header longer than the supplied prefix");
    if (!err.empty()) return fail(B200_ERR_INVALID, "GGUF: " + err);
    tokenizer_from_gguf(meta, 0, tk);
    return B200_OK;
}
static int copy_tokens(const std::vector<int32_t>& t, int32_t* out, int cap, int* n_out) {
    if (n_out) *n_out = (int)t.size();
    if (!out || cap < (int)t.size()) return out ? fail(B200_ERR_MORE, "token buffer too small") : B200_OK;
    memcpy(out, t.data(), t.size() * 4); return B200_OK;
}
int b200_gguf_tokenize(const void* gguf, size_t n, const char* text, size_t tn, int parse_special, int32_t* out, int cap, int* n_out) {
    Tokenizer tk; if (int rc = with_gguf_tokenizer(gguf, n, tk)) return rc;
    return copy_tokens(tk.tokenize(std::string(text ? text : "", text ? tn : 0), parse_special != 0), out, cap, n_out);
}
int b200_gguf_detokenize(const void* gguf, size_t n, const int32_t* toks, int nt, char* out, size_t cap, size_t* need) {
    Tokenizer tk; if (int rc = with_gguf_tokenizer(gguf, n, tk)) return rc;
    return copy_out(tk.detokenize(toks, nt), out, cap, need);
}
int b200_gguf_render_chat(const void* gguf, size_t n, const char* messages_json, size_t mn, int32_t* out, int cap, int* n_out) {
    Tokenizer tk; if (int rc = with_gguf_tokenizer(gguf, n, tk)) return rc;
    std::vector<int32_t> t; std::string err = render_chat(tk, std::string(messages_json ? messages_json : "", messages_json ? mn : 0), t);
    if (!err.empty()) return fail(B200_ERR_INVALID, err);
    return copy_tokens(t, out, cap, n_out);
}

int b200_submit(b200_model* m, const int32_t* toks, int n, const b200_sampling* s, uint64_t req_id) {
    if (!m || !toks || n < 1) return fail(B200_ERR_INVALID, "bad submit arguments");
    auto sp = find_shared(m); if (!sp || !sp->committed) return fail(B200_ERR_STATE, "model is not loaded");
    auto r = std::make_shared<Request>(); r->m = sp; r->toks.assign(toks, toks + n); r->id = req_id; r->t_submit = std::chrono::steady_clock::now();
    if (s) r->s = *s; else { r->s = b200_sampling{}; r->s.max_tokens = 256; r->s.stop_token = -1; }
    if (req_id >> 63) return fail(B200_ERR_INVALID, "request ids with the top bit set are reserved for the library's own requests");
    enqueue(r);
    return B200_OK;
}
int b200_poll(b200_event* out, int max, int timeout_ms) {
    if (!out || max < 1) return fail(B200_ERR_INVALID, "bad poll arguments");
    Scheduler& S = sched(); std::unique_lock<std::mutex> lk(S.mu);
    if (S.ev.empty() && timeout_ms != 0) S.cv_ev.wait_for(lk, std::chrono::milliseconds(timeout_ms < 0 ? 3600000 : timeout_ms), [&] { return !S.ev.empty(); });
    int n = 0; while (n < max && !S.ev.empty()) { out[n++] = S.ev.front(); S.ev.pop_front(); }
    return n;
}
int b200_cancel(uint64_t req_id) {
    Scheduler& S = sched(); std::lock_guard<std::mutex> lk(S.mu);
    auto it = S.live.find(req_id); if (it == S.live.end()) return fail(B200_ERR_NOT_FOUND, "no such request");
    it->second->cancelled = true; return B200_OK;
}

int b200_handle(const char* subject, const void* payload, size_t n, char* reply, size_t cap, size_t* need) {
    if (!subject) return fail(B200_ERR_INVALID, "null subject");
    return copy_out(handle_subject(subject, std::string((const char*)payload, payload ? n : 0)), reply, cap, need);
}
int b200_worker_run(const char* nats_url, const char* queue_group, int n_threads) { return worker_run(nats_url, queue_group, n_threads); }
int b200_worker_stop(void) { return worker_stop(); }
int b200_set_models_dir(const char* dir) { Registry& R = registry(); std::lock_guard<std::mutex> lk(R.mu); R.models_dir = dir ? dir : ""; while (R.models_dir.size() > 1 && R.models_dir.back() == '/') R.models_dir.pop_back(); return B200_OK; }

int b200_sample(const float* logits, int V, const b200_sampling* s, uint64_t step, int* token) { if (!logits || !s || !token) return fail(B200_ERR_INVALID, "null argument"); return kernel_sample(logits, V, *s, step, token); }
int b200_dequantize(int type, const void* raw, int64_t n_rows, int K, float* out) { if ((!raw || !out) && n_rows) return fail(B200_ERR_INVALID, "null argument"); return kernel_dequantize(type, raw, n_rows, K, out); }
int b200_gemv(int type, const void* raw, int64_t N, int K, const float* x, float* y, int iters, float* ms) { if (!raw || !x || !y) return fail(B200_ERR_INVALID, "null argument"); return kernel_gemv(type, raw, N, K, x, y, iters, ms); }
int b200_dequantize_noalloc(int type, const void* d_raw, int64_t n_rows, int K, float* d_out) { return kernel_dequantize_noalloc(type, d_raw, n_rows, K, d_out); }
int b200_qgemm(int type, const void* raw, int64_t N, int K, const float* x, int B, float* y, int iters, float* ms) { if (!raw || !x || !y) return fail(B200_ERR_INVALID, "null argument"); return kernel_qgemm(type, raw, N, K, x, B, y, iters, ms); }
int b200_gemm_bf16(const float* A, const float* B, float* C, int M, int N, int K) { if (!A || !B || !C) return fail(B200_ERR_INVALID, "null argument"); return kernel_gemm_bf16(A, B, C, M, N, K); }
int b200_model_set_prefill_min(b200_model* m, int n) { if (!m) return fail(B200_ERR_INVALID, "null model"); ((Model*)m)->prefill_min = n; return B200_OK; }
int b200_model_batch_step(b200_model* m, const int32_t* toks, const int32_t* slots, const int32_t* pos, int n, int32_t* next, float* logits) {
    if (!m || !toks || !slots || !pos || !next) return fail(B200_ERR_INVALID, "null argument");
    Model* M = (Model*)m; std::lock_guard<std::mutex> lk(M->run_mu); return M->batch_step(toks, slots, pos, n, next, logits);
}
int b200_model_batch_step_sampled(b200_model* m, const int32_t* toks, const int32_t* slots, const int32_t* pos, int n, const b200_sampling* samp, const int32_t* draw, int32_t* next, float* logits) {
    if (!m || !toks || !slots || !pos || !next) return fail(B200_ERR_INVALID, "null argument");
    Model* M = (Model*)m; std::lock_guard<std::mutex> lk(M->run_mu); return M->batch_step(toks, slots, pos, n, next, logits, samp, draw);
}
int b200_model_prefill_slot(b200_model* m, int slot, const int32_t* toks, int n, int pos0) {
    if (!m || (!toks && n)) return fail(B200_ERR_INVALID, "null argument");
    Model* M = (Model*)m; if (slot < 0 || slot >= engine().max_seqs) return fail(B200_ERR_INVALID, "bad sequence slot");
    std::lock_guard<std::mutex> lk(M->run_mu); return M->prefill_slot(slot, toks, n, pos0);
}
int b200_bench_batch_step(b200_model* m, int n_seqs, int ctx, int iters, int warmup, float* ms, double* bytes) { if (!m) return fail(B200_ERR_INVALID, "null model"); return ((Model*)m)->bench_batch(n_seqs, ctx, iters, warmup, ms, bytes); }
int b200_tp_unique_id(void* out, size_t cap) { if (!out) return fail(B200_ERR_INVALID, "null argument"); if (int rc = engine().ensure()) return rc; return engine().tp_unique_id(out, cap); }
int b200_tp_init(int rank, int world, const void* id, size_t n) { return engine().tp_init(rank, world, id, n); }
int b200_tp_shard_window(const char* kind, int64_t N, int K, int rank, int world, int64_t* row0, int64_t* nrows, int* col0, int* ncols) {
    if (!row0 || !nrows || !col0 || !ncols) return fail(B200_ERR_INVALID, "null argument");
    return tp_shard_window(kind, N, K, rank, world, row0, nrows, col0, ncols);
}
int b200_layout_selftest(void) { return layout_selftest(); }
uint64_t b200_launch_count(void) { return engine().launches; }

}  // extern "C"
