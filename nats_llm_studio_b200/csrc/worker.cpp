// worker.cpp — the subscription wiring the reference describes but does not contain (/root/reference/README.md:475-502:
// connect to NATS_URL, queue-subscribe to the four lmstudio.* subjects with NATS_QUEUE_GROUP, answer every request with
// msg.Respond) plus the bucket-sync subject sketched at README.md:250-318.  The reference would do this with nats.go
// (go.mod:7, un-vendored); no Go toolchain exists on the build/GPU image, so this is a small client of the NATS core
// protocol (text lines over TCP: INFO / CONNECT / SUB / PUB / MSG / PING / PONG) in C++ in front of the same handlers
// (handle_subject) that b200_handle exposes.
//
// Threading: nats.go delivers a subscription's messages one at a time; that serialises chat requests inside one worker.
// Here the reader thread only parses and enqueues; a pool of handler threads runs the requests, so concurrent
// lmstudio.chat_model requests reach the continuous-batching scheduler together (BASELINE.json configs[2]).
#include <arpa/inet.h>
#include <netdb.h>
#include <netinet/tcp.h>
#include <sys/socket.h>
#include <unistd.h>

#include <atomic>
#include <cstring>

#include "engine.hpp"
#include "json.hpp"

namespace b200 {

namespace {

struct NatsConn {
    int fd = -1; std::mutex wmu; std::string rbuf; size_t rpos = 0;
    ~NatsConn() { close_fd(); }
    void close_fd() { if (fd >= 0) { ::shutdown(fd, SHUT_RDWR); ::close(fd); fd = -1; } }

    // nats://host:port (scheme and port optional)
    std::string connect_url(const std::string& url) {
        std::string u = url; size_t k = u.find("://"); if (k != std::string::npos) u = u.substr(k + 3);
        k = u.find('@'); if (k != std::string::npos) u = u.substr(k + 1);            // credentials are not supported: plain connections only
        std::string host = u, port = "4222"; k = u.rfind(':'); if (k != std::string::npos) { host = u.substr(0, k); port = u.substr(k + 1); }
        addrinfo hints{}; hints.ai_family = AF_UNSPEC; hints.ai_socktype = SOCK_STREAM; addrinfo* res = nullptr;
        if (getaddrinfo(host.c_str(), port.c_str(), &hints, &res) != 0 || !res) return "cannot resolve " + host;
        std::string err = "cannot connect to " + host + ":" + port;
        for (addrinfo* a = res; a; a = a->ai_next) {
            int s = ::socket(a->ai_family, a->ai_socktype, a->ai_protocol); if (s < 0) continue;
            if (::connect(s, a->ai_addr, a->ai_addrlen) == 0) { fd = s; err.clear(); break; }
            ::close(s);
        }
        freeaddrinfo(res);
        if (fd >= 0) { int one = 1; setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one); }
        return err;
    }
    bool send_all(const std::string& s) {
        std::lock_guard<std::mutex> lk(wmu);
        size_t off = 0;
        while (off < s.size()) { ssize_t n = ::send(fd, s.data() + off, s.size() - off, MSG_NOSIGNAL); if (n <= 0) return false; off += (size_t)n; }
        return true;
    }
    bool fill() { char tmp[65536]; ssize_t n = ::recv(fd, tmp, sizeof tmp, 0); if (n <= 0) return false; rbuf.append(tmp, (size_t)n); return true; }
    bool read_line(std::string& line) {
        for (;;) {
            size_t k = rbuf.find("\r\n", rpos);
            if (k != std::string::npos) { line = rbuf.substr(rpos, k - rpos); rpos = k + 2; compact(); return true; }
            if (!fill()) return false;
        }
    }
    bool read_bytes(size_t n, std::string& out) {                        // payload + trailing CRLF
        while (rbuf.size() - rpos < n + 2) if (!fill()) return false;
        out = rbuf.substr(rpos, n); rpos += n + 2; compact(); return true;
    }
    void compact() { if (rpos > (1u << 20)) { rbuf.erase(0, rpos); rpos = 0; } }
    bool publish(const std::string& subject, const std::string& payload) {
        return send_all("PUB " + subject + " " + std::to_string(payload.size()) + "\r\n" + payload + "\r\n");
    }
};

struct Job { std::string subject, reply, payload; };

// ---- SHA-256 (FIPS 180-4) and base64: the Object Store records an object's digest as "SHA-256=<base64url>"
struct Sha256 {
    uint32_t h[8] = {0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a, 0x510e527f, 0x9b05688c, 0x1f83d9ab, 0x5be0cd19};
    uint8_t buf[64]; size_t fill = 0; uint64_t total = 0;
    static uint32_t rotr(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }
    void block(const uint8_t* p) {
        static const uint32_t K[64] = {0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5, 0xd807aa98, 0x12835b01, 0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174,
                                       0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da, 0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967,
                                       0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13, 0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85, 0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070,
                                       0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f, 0x682e6ff3, 0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208, 0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2};
        uint32_t w[64];
        for (int i = 0; i < 16; i++) w[i] = (uint32_t)p[4 * i] << 24 | (uint32_t)p[4 * i + 1] << 16 | (uint32_t)p[4 * i + 2] << 8 | p[4 * i + 3];
        for (int i = 16; i < 64; i++) { const uint32_t s0 = rotr(w[i - 15], 7) ^ rotr(w[i - 15], 18) ^ (w[i - 15] >> 3), s1 = rotr(w[i - 2], 17) ^ rotr(w[i - 2], 19) ^ (w[i - 2] >> 10); w[i] = w[i - 16] + s0 + w[i - 7] + s1; }
        uint32_t a = h[0], b = h[1], c = h[2], d = h[3], e = h[4], f = h[5], g = h[6], hh = h[7];
        for (int i = 0; i < 64; i++) {
            const uint32_t S1 = rotr(e, 6) ^ rotr(e, 11) ^ rotr(e, 25), ch = (e & f) ^ (~e & g), t1 = hh + S1 + ch + K[i] + w[i];
            const uint32_t S0 = rotr(a, 2) ^ rotr(a, 13) ^ rotr(a, 22), mj = (a & b) ^ (a & c) ^ (b & c), t2 = S0 + mj;
            hh = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
        }
        h[0] += a; h[1] += b; h[2] += c; h[3] += d; h[4] += e; h[5] += f; h[6] += g; h[7] += hh;
    }
    void update(const void* data, size_t n) {
        const uint8_t* p = (const uint8_t*)data; total += n;
        if (fill) { const size_t k = std::min(n, 64 - fill); memcpy(buf + fill, p, k); fill += k; p += k; n -= k; if (fill == 64) { block(buf); fill = 0; } }
        while (n >= 64) { block(p); p += 64; n -= 64; }
        if (n) { memcpy(buf, p, n); fill = n; }
    }
    std::string digest() {
        const uint64_t bits = total * 8; uint8_t pad[72] = {0x80}; const size_t padn = (fill < 56 ? 56 - fill : 120 - fill);
        uint8_t len[8]; for (int i = 0; i < 8; i++) len[i] = (uint8_t)(bits >> (56 - 8 * i));
        update(pad, padn); update(len, 8);
        std::string out(32, '\0'); for (int i = 0; i < 8; i++) for (int k = 0; k < 4; k++) out[4 * i + k] = (char)(h[i] >> (24 - 8 * k));
        return out;
    }
};
std::string b64_encode(const std::string& in, bool url) {
    const char* T = url ? "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789-_" : "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789+/";
    std::string out; size_t i = 0;
    for (; i + 2 < in.size(); i += 3) { const uint32_t v = (uint8_t)in[i] << 16 | (uint8_t)in[i + 1] << 8 | (uint8_t)in[i + 2]; out += T[v >> 18]; out += T[(v >> 12) & 63]; out += T[(v >> 6) & 63]; out += T[v & 63]; }
    if (i + 1 == in.size()) { const uint32_t v = (uint8_t)in[i] << 16; out += T[v >> 18]; out += T[(v >> 12) & 63]; out += "=="; }
    else if (i + 2 == in.size()) { const uint32_t v = (uint8_t)in[i] << 16 | (uint8_t)in[i + 1] << 8; out += T[v >> 18]; out += T[(v >> 12) & 63]; out += T[(v >> 6) & 63]; out += "="; }
    return out;
}
std::string b64_decode(const std::string& in) {                           // standard or url-safe alphabet, padding optional
    std::string out; uint32_t v = 0; int bits = 0;
    for (char ch : in) {
        int d = ch >= 'A' && ch <= 'Z' ? ch - 'A' : ch >= 'a' && ch <= 'z' ? ch - 'a' + 26 : ch >= '0' && ch <= '9' ? ch - '0' + 52 : (ch == '+' || ch == '-') ? 62 : (ch == '/' || ch == '_') ? 63 : -1;
        if (d < 0) continue;
        v = v << 6 | (uint32_t)d; bits += 6;
        if (bits >= 8) { bits -= 8; out.push_back((char)((v >> bits) & 0xff)); }
    }
    return out;
}
std::string rstrip_eq(std::string s) { while (!s.empty() && s.back() == '=') s.pop_back(); return s; }

// messages of the worker's own inbox subscriptions (JetStream API replies, object chunks), keyed by subscription id
struct Inbox { std::mutex mu; std::condition_variable cv; std::deque<std::string> msgs; };

struct Worker {
    NatsConn nc; std::atomic<bool> stop{false};
    std::mutex mu; std::condition_variable cv; std::deque<Job> jobs; std::vector<std::thread> pool;
    std::atomic<uint64_t> served{0};
    std::mutex imu; std::map<std::string, std::shared_ptr<Inbox>> inboxes; std::atomic<uint64_t> next_sid{100};

    std::pair<std::string, std::shared_ptr<Inbox>> open_inbox(std::string& subject) {
        const std::string sid = std::to_string(next_sid++);
        subject = "_INBOX.b200." + std::to_string((uint64_t)getpid()) + "." + sid;
        auto ib = std::make_shared<Inbox>();
        { std::lock_guard<std::mutex> lk(imu); inboxes[sid] = ib; }
        nc.send_all("SUB " + subject + " " + sid + "\r\n");
        return {sid, ib};
    }
    void close_inbox(const std::string& sid) { nc.send_all("UNSUB " + sid + "\r\n"); std::lock_guard<std::mutex> lk(imu); inboxes.erase(sid); }
    bool next(Inbox& ib, std::string& out, int timeout_ms) {
        std::unique_lock<std::mutex> lk(ib.mu);
        if (!ib.cv.wait_for(lk, std::chrono::milliseconds(timeout_ms), [&] { return !ib.msgs.empty() || stop.load(); }) || ib.msgs.empty()) return false;
        out = std::move(ib.msgs.front()); ib.msgs.pop_front(); return true;
    }
    bool request(const std::string& subject, const std::string& payload, std::string& reply, int timeout_ms) {      // nc.Request
        std::string inbox; auto [sid, ib] = open_inbox(inbox);
        nc.send_all("PUB " + subject + " " + inbox + " " + std::to_string(payload.size()) + "\r\n" + payload + "\r\n");
        const bool ok = next(*ib, reply, timeout_ms);
        close_inbox(sid); return ok;
    }
    std::string sync_from_bucket(const std::string& payload);

    void handler_loop() {
        for (;;) {
            Job j;
            { std::unique_lock<std::mutex> lk(mu); cv.wait(lk, [&] { return stop || !jobs.empty(); }); if (jobs.empty()) return; j = std::move(jobs.front()); jobs.pop_front(); }
            const std::string rep = j.subject == "lmstudio.sync_model_from_bucket" ? sync_from_bucket(j.payload) : handle_subject(j.subject, j.payload);
            if (!j.reply.empty()) nc.publish(j.reply, rep);                 // msg.Respond (nats_llm_studio.go:214)
            served++;
        }
    }
};
Worker* g_worker = nullptr; std::mutex g_worker_mu;

std::string env_error(const std::string& err, const Json* data) {          // respondError (nats_llm_studio.go:219-226)
    std::string out = "{\"ok\":false,\"error\":"; json_escape(err, out); out += ",\"data\":"; if (data) json_write(*data, out); else out += "null"; out += "}"; return out;
}

// lmstudio.sync_model_from_bucket (README.md:286-318): {"bucket","object_name","publisher","model_dir"} -> the object is
// read from the JetStream Object Store (metadata record by direct get, chunks through an ephemeral ordered push consumer on
// $O.<bucket>.C.<nuid>), streamed chunk by chunk through pinned host buffers into HBM (b200_model_open / upload), checked
// against the recorded SHA-256 digest, and committed (on-GPU repack) under the id "<publisher>/<model_dir>".  The reference
// would write the file under LMSTUDIO_MODELS_DIR; here the model lives in HBM, so "local_path" names it there.
std::string Worker::sync_from_bucket(const std::string& payload) {
    Json doc; JsonParser ps(payload);
    if (payload.empty() || !ps.parse(doc) || doc.kind != Json::Obj) return env_error("invalid JSON in SyncModelFromBucket: " + (payload.empty() ? std::string("unexpected end of JSON input") : ps.err), nullptr);
    auto field = [&](const char* k) { const Json* f = doc.get(k); return f && f->kind == Json::Str ? f->s : std::string(); };
    const std::string bucket = field("bucket"), object = field("object_name"); std::string publisher = field("publisher"), dir = field("model_dir");
    if (bucket.empty() || object.empty()) return env_error("'bucket' and 'object_name' are required", nullptr);
    if (publisher.empty()) { size_t k = object.find('/'); publisher = k == std::string::npos ? "local" : object.substr(0, k); }
    if (dir.empty()) { size_t a = object.find('/'), b = object.rfind('/'); dir = (a != std::string::npos && b > a) ? object.substr(a + 1, b - a - 1) : object; }
    const std::string id = publisher + "/" + dir;
    Json data = Json::object(true); data.set("bucket", Json::string(bucket)).set("object_name", Json::string(object)).set("model", Json::string(id));
    auto failed = [&](const std::string& why) { return env_error("failed to sync model '" + id + "' from bucket '" + bucket + "': " + why, &data); };
    const auto t0 = std::chrono::steady_clock::now();
    // ---- 1. metadata record
    std::string rep;
    if (!request("$JS.API.STREAM.MSG.GET.OBJ_" + bucket, "{\"last_by_subj\":\"$O." + bucket + ".M." + rstrip_eq(b64_encode(object, true)) + "\"}", rep, 10000)) return failed("no answer from JetStream (is the Object Store enabled?)");
    Json jr; { JsonParser p2(rep); if (!p2.parse(jr) || jr.kind != Json::Obj) return failed("malformed JetStream answer"); }
    if (const Json* e = jr.get("error")) { const Json* d = e->kind == Json::Obj ? e->get("description") : nullptr; return failed("object not found (" + (d && d->kind == Json::Str ? d->s : std::string("JetStream error")) + ")"); }
    const Json* msg = jr.get("message"); const Json* md = msg && msg->kind == Json::Obj ? msg->get("data") : nullptr;
    if (!md || md->kind != Json::Str) return failed("JetStream answer without a message");
    Json meta; { const std::string raw = b64_decode(md->s); JsonParser p3(raw); if (!p3.parse(meta) || meta.kind != Json::Obj) return failed("malformed object metadata"); }
    const Json *jn = meta.get("nuid"), *js = meta.get("size"), *jc = meta.get("chunks"), *jd = meta.get("digest"), *jdel = meta.get("deleted");
    if (jdel && jdel->kind == Json::Bool && jdel->b) return failed("object was deleted");
    if (!jn || jn->kind != Json::Str || !js || js->kind != Json::Num || !jc || jc->kind != Json::Num) return failed("object metadata lacks nuid/size/chunks");
    const uint64_t size = (uint64_t)js->num; const long n_chunks = (long)jc->num;
    // ---- 2. ordered push consumer over the chunk subject
    std::string inbox; auto [sid, ib] = open_inbox(inbox);
    struct Guard { Worker* w; std::string sid; ~Guard() { w->close_inbox(sid); } } guard{this, sid};
    const std::string cfg = "{\"stream_name\":\"OBJ_" + bucket + "\",\"config\":{\"deliver_subject\":\"" + inbox + "\",\"filter_subject\":\"$O." + bucket + ".C." + jn->s +
                            "\",\"deliver_policy\":\"all\",\"ack_policy\":\"none\",\"replay_policy\":\"instant\",\"flow_control\":false,\"mem_storage\":true,\"num_replicas\":1,\"inactive_threshold\":30000000000}}";
    if (!request("$JS.API.CONSUMER.CREATE.OBJ_" + bucket, cfg, rep, 10000)) return failed("no answer to the consumer request");
    { Json cr; JsonParser p4(rep); if (p4.parse(cr) && cr.kind == Json::Obj && cr.get("error")) return failed("JetStream refused the chunk consumer"); }
    // ---- 3. chunks -> pinned buffers -> HBM
    const size_t PIN = 32u << 20;
    uint8_t* pin = (uint8_t*)b200_pinned_alloc(PIN);
    if (!pin) return failed(g_last_error.empty() ? "pinned staging allocation failed" : g_last_error);
    struct PinGuard { uint8_t* p; ~PinGuard() { b200_pinned_free(p); } } pg{pin};
    b200_model* model = nullptr; std::string head; Sha256 sha; uint64_t got = 0, flushed = 0; size_t fill = 0; std::string chunk;
    auto flush = [&]() -> int { if (!fill) return B200_OK; int rc = b200_model_upload(model, flushed, pin, fill); flushed += fill; fill = 0; return rc; };
    auto feed = [&](const uint8_t* p, size_t n) -> int {                     // after the model is open
        while (n) { const size_t k = std::min(n, PIN - fill); memcpy(pin + fill, p, k); fill += k; p += k; n -= k; if (fill == PIN) if (int rc = flush()) return rc; }
        return B200_OK;
    };
    int rc = B200_OK;
    for (long c = 0; c < n_chunks && got < size; c++) {
        if (!next(*ib, chunk, 30000)) { rc = fail(B200_ERR_STATE, "timed out waiting for chunk " + std::to_string(c) + " of " + std::to_string(n_chunks)); break; }
        sha.update(chunk.data(), chunk.size()); got += chunk.size();
        if (!model) {                                                        // the GGUF header may span several chunks
            head += chunk;
            rc = b200_model_open(id.c_str(), head.data(), head.size(), &model);
            if (rc == B200_ERR_MORE) { rc = B200_OK; model = nullptr; continue; }
            if (rc != B200_OK) break;
            if ((rc = feed((const uint8_t*)head.data(), head.size()))) break;
            head.clear(); head.shrink_to_fit();
        } else if ((rc = feed((const uint8_t*)chunk.data(), chunk.size()))) break;
    }
    if (rc == B200_OK && !model) rc = fail(B200_ERR_INVALID, "object ended inside the GGUF header");
    if (rc == B200_OK) rc = flush();
    if (rc == B200_OK && got != size) rc = fail(B200_ERR_INVALID, "object has " + std::to_string(got) + " bytes, its metadata says " + std::to_string(size));
    const std::string digest = "SHA-256=" + b64_encode(sha.digest(), true);
    if (rc == B200_OK && jd && jd->kind == Json::Str && rstrip_eq(jd->s) != rstrip_eq(digest)) rc = fail(B200_ERR_INVALID, "digest mismatch: the bucket records " + jd->s + ", the received bytes hash to " + digest);
    if (rc == B200_OK) rc = b200_model_commit(model);
    if (rc != B200_OK) { const std::string why = g_last_error; if (model) b200_model_close(model); return failed(why); }
    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    Json ok = Json::object(true);
    ok.set("local_path", Json::string("hbm://" + id)).set("model", Json::string(id)).set("bytes", Json::integer((long long)got)).set("chunks", Json::integer(n_chunks))
        .set("digest", Json::string(digest)).set("seconds", Json::number(ms * 1e-3));
    std::string out = "{\"ok\":true,\"data\":"; json_write(ok, out); out += "}";
    return out;
}

}  // namespace

// Serves the subjects until b200_worker_stop() or the connection drops.  Returns B200_OK on a clean stop.
int worker_run(const char* url_arg, const char* queue_arg, int n_threads) {
    const char* eu = getenv("NATS_URL"); const char* eq = getenv("NATS_QUEUE_GROUP");
    const std::string url = url_arg && *url_arg ? url_arg : (eu && *eu ? eu : "nats://127.0.0.1:4222");                 // README.md:488-494 defaults
    const std::string queue = queue_arg && *queue_arg ? queue_arg : (eq && *eq ? eq : "lmstudio-workers");
    if (const char* md = getenv("LMSTUDIO_MODELS_DIR")) { if (*md) b200_set_models_dir(md); }
    if (n_threads <= 0) n_threads = std::max(4, 2 * engine().max_seqs);       // enough in-flight chats to keep every sequence slot busy
    Worker w;
    if (std::string err = w.nc.connect_url(url); !err.empty()) return fail(B200_ERR_NOT_FOUND, "NATS: " + err);
    std::string line;
    if (!w.nc.read_line(line) || line.compare(0, 4, "INFO") != 0) return fail(B200_ERR_INVALID, "NATS: no INFO from the server");
    if (!w.nc.send_all("CONNECT {\"verbose\":false,\"pedantic\":false,\"name\":\"b200-lmstudio-worker\",\"lang\":\"c++\",\"version\":\"0.2\",\"protocol\":0}\r\n"))
        return fail(B200_ERR_INVALID, "NATS: CONNECT failed");
    static const char* subjects[] = {"lmstudio.list_models", "lmstudio.pull_model", "lmstudio.delete_model", "lmstudio.chat_model", "lmstudio.sync_model_from_bucket"};
    std::string subs; int sid = 1;
    for (const char* s : subjects) subs += std::string("SUB ") + s + " " + queue + " " + std::to_string(sid++) + "\r\n";      // QueueSubscribe
    subs += "PING\r\n";
    if (!w.nc.send_all(subs)) return fail(B200_ERR_INVALID, "NATS: SUB failed");
    for (int i = 0; i < n_threads; i++) w.pool.emplace_back([&w] { w.handler_loop(); });
    { std::lock_guard<std::mutex> lk(g_worker_mu); g_worker = &w; }
    int rc = B200_OK;
    while (!w.stop) {
        if (!w.nc.read_line(line)) { if (!w.stop) rc = fail(B200_ERR_STATE, "NATS: connection closed"); break; }
        if (line.compare(0, 4, "MSG ") == 0) {                               // MSG <subject> <sid> [reply-to] <#bytes>
            std::vector<std::string> f; size_t p = 4;
            while (p < line.size()) { size_t q = line.find(' ', p); if (q == std::string::npos) q = line.size(); if (q > p) f.push_back(line.substr(p, q - p)); p = q + 1; }
            if (f.size() < 3) { rc = fail(B200_ERR_INVALID, "NATS: malformed MSG"); break; }
            Job j; j.subject = f[0]; j.reply = f.size() >= 4 ? f[2] : ""; const size_t n = (size_t)strtoull(f.back().c_str(), nullptr, 10);
            if (!w.nc.read_bytes(n, j.payload)) { if (!w.stop) rc = fail(B200_ERR_STATE, "NATS: connection closed"); break; }
            std::shared_ptr<Inbox> ib;
            { std::lock_guard<std::mutex> lk(w.imu); auto it = w.inboxes.find(f[1]); if (it != w.inboxes.end()) ib = it->second; }
            if (ib) { { std::lock_guard<std::mutex> lk(ib->mu); ib->msgs.push_back(std::move(j.payload)); } ib->cv.notify_all(); continue; }      // a reply / a chunk for a handler thread
            if (atoi(f[1].c_str()) >= 100) continue;                            // a late message of an inbox that is already closed
            { std::lock_guard<std::mutex> lk(w.mu); w.jobs.push_back(std::move(j)); }
            w.cv.notify_one();
        } else if (line == "PING") w.nc.send_all("PONG\r\n");
        else if (line.compare(0, 4, "-ERR") == 0) { rc = fail(B200_ERR_INVALID, "NATS: " + line); break; }
        // PONG, +OK, INFO updates: nothing to do
    }
    { std::lock_guard<std::mutex> lk(g_worker_mu); g_worker = nullptr; }
    { std::lock_guard<std::mutex> lk(w.mu); w.stop = true; }                   // drain: queued requests are still answered
    w.cv.notify_all();
    for (auto& t : w.pool) t.join();
    return rc;
}

int worker_stop() {
    std::lock_guard<std::mutex> lk(g_worker_mu);
    if (!g_worker) return fail(B200_ERR_STATE, "no worker is running");
    g_worker->stop = true; g_worker->nc.close_fd();                           // unblocks the reader
    return B200_OK;
}

}  // namespace b200
