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

struct Worker {
    NatsConn nc; std::atomic<bool> stop{false};
    std::mutex mu; std::condition_variable cv; std::deque<Job> jobs; std::vector<std::thread> pool;
    std::atomic<uint64_t> served{0};

    void handler_loop() {
        for (;;) {
            Job j;
            { std::unique_lock<std::mutex> lk(mu); cv.wait(lk, [&] { return stop || !jobs.empty(); }); if (jobs.empty()) return; j = std::move(jobs.front()); jobs.pop_front(); }
            const std::string rep = handle_subject(j.subject, j.payload);
            if (!j.reply.empty()) nc.publish(j.reply, rep);                 // msg.Respond (nats_llm_studio.go:214)
            served++;
        }
    }
};
Worker* g_worker = nullptr; std::mutex g_worker_mu;

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
