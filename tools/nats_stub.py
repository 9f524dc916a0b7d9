#!/usr/bin/env python3
"""A stand-in for nats-server (test and bench infrastructure; neither nats-server nor the `nats` CLI exists on the build
or GPU image): the subset of the NATS core protocol that the reference's deployment uses (README.md:475-562) —
INFO / CONNECT / PING / PONG / SUB with queue groups / UNSUB [max] / PUB with reply-to / MSG — plus the two JetStream API
requests an Object Store reader needs (README.md:250-318), served from objects put with Broker.put_object:
    $JS.API.STREAM.MSG.GET.OBJ_<bucket>   {"last_by_subj": "$O.<bucket>.M.<b64(name)>"}  -> the object's metadata record
    $JS.API.CONSUMER.CREATE.OBJ_<bucket>  {"config": {"deliver_subject", "filter_subject": "$O.<bucket>.C.<nuid>"}}
                                          -> the object's chunks, in order, on deliver_subject
Also: Client (blocking request/reply over one socket, thread-safe) and `python tools/nats_stub.py --port N` to run a broker.
"""
import argparse
import base64
import hashlib
import json
import random
import socket
import socketserver
import threading
import time


class Broker:
    def __init__(self, host="127.0.0.1", port=0, chunk=128 * 1024):
        self.lock = threading.Lock()
        self.subs = {}                      # subject -> list of (conn, sid, queue)
        self.objects = {}                   # (bucket, name) -> {"meta": dict, "chunks": [bytes]}
        self.chunk = chunk
        self.rr = 0
        broker = self

        class Handler(socketserver.BaseRequestHandler):
            def handle(self):
                broker._serve(self.request)

        class Server(socketserver.ThreadingTCPServer):
            allow_reuse_address = True
            daemon_threads = True

        self.server = Server((host, port), Handler)
        self.port = self.server.server_address[1]
        self.url = f"nats://{host}:{self.port}"
        self.thread = threading.Thread(target=self.server.serve_forever, daemon=True)

    def start(self):
        self.thread.start(); return self

    def stop(self):
        self.server.shutdown(); self.server.server_close()

    # ---- object store content (what `nats object put <bucket> <file>` would have stored)
    def put_object(self, bucket, name, data, corrupt_digest=False):
        nuid = "".join(random.choice("ABCDEFGHIJKLMNOPQRSTUVWXYZ0123456789") for _ in range(22))
        chunks = [data[i:i + self.chunk] for i in range(0, len(data), self.chunk)] or [b""]
        dg = hashlib.sha256(data if not corrupt_digest else data + b"x").digest()
        meta = {"name": name, "bucket": bucket, "nuid": nuid, "size": len(data), "chunks": len(chunks),
                "digest": "SHA-256=" + base64.urlsafe_b64encode(dg).decode(), "mtime": time.strftime("%Y-%m-%dT%H:%M:%SZ", time.gmtime())}
        with self.lock:
            self.objects[(bucket, name)] = {"meta": meta, "chunks": chunks}
        return meta

    # ---- protocol
    class _Conn:
        def __init__(self, sock):
            self.sock = sock; self.wlock = threading.Lock(); self.limits = {}

        def send(self, b):
            with self.wlock:
                try:
                    self.sock.sendall(b); return True
                except OSError:
                    return False

    def _deliver(self, subject, reply, payload):
        with self.lock:
            targets = list(self.subs.get(subject, []))
        plain = [t for t in targets if not t[2]]
        groups = {}
        for t in targets:
            if t[2]:
                groups.setdefault(t[2], []).append(t)
        chosen = plain + [g[self._next() % len(g)] for g in groups.values()]        # one member per queue group
        for conn, sid, _ in chosen:
            head = f"MSG {subject} {sid} {reply + ' ' if reply else ''}{len(payload)}\r\n".encode()
            conn.send(head + payload + b"\r\n")
            lim = conn.limits.get(sid)
            if lim is not None:
                conn.limits[sid] = lim - 1
                if lim - 1 <= 0:
                    self._unsub(conn, sid)
        return len(chosen)

    def _next(self):
        with self.lock:
            self.rr += 1; return self.rr

    def _unsub(self, conn, sid):
        with self.lock:
            for subj in list(self.subs):
                self.subs[subj] = [t for t in self.subs[subj] if not (t[0] is conn and t[1] == sid)]

    def _js_api(self, subject, reply, payload):
        try:
            req = json.loads(payload or b"{}")
        except ValueError:
            req = {}
        if subject.startswith("$JS.API.STREAM.MSG.GET.OBJ_"):
            bucket = subject[len("$JS.API.STREAM.MSG.GET.OBJ_"):]
            want = req.get("last_by_subj", "")
            with self.lock:
                hit = [o for (b, n), o in self.objects.items() if b == bucket and want == f"$O.{bucket}.M.{base64.urlsafe_b64encode(n.encode()).decode().rstrip('=')}"]
            if not hit:
                body = {"type": "io.nats.jetstream.api.v1.stream_msg_get_response", "error": {"code": 404, "err_code": 10037, "description": "no message found"}}
            else:
                body = {"type": "io.nats.jetstream.api.v1.stream_msg_get_response",
                        "message": {"subject": want, "seq": 1, "data": base64.b64encode(json.dumps(hit[0]["meta"]).encode()).decode(), "time": hit[0]["meta"]["mtime"]}}
            self._deliver(reply, "", json.dumps(body).encode())
        elif subject.startswith("$JS.API.CONSUMER.CREATE.OBJ_"):
            bucket = subject[len("$JS.API.CONSUMER.CREATE.OBJ_"):].split(".")[0]
            cfg = req.get("config", {})
            filt, dst = cfg.get("filter_subject", ""), cfg.get("deliver_subject", "")
            with self.lock:
                hit = [o for (b, n), o in self.objects.items() if b == bucket and filt == f"$O.{bucket}.C.{o['meta']['nuid']}"]
            if not hit or not dst:
                self._deliver(reply, "", json.dumps({"type": "io.nats.jetstream.api.v1.consumer_create_response", "error": {"code": 404, "description": "stream or subject not found"}}).encode())
                return
            self._deliver(reply, "", json.dumps({"type": "io.nats.jetstream.api.v1.consumer_create_response", "stream_name": f"OBJ_{bucket}", "name": "stub", "config": cfg,
                                                   "num_pending": len(hit[0]["chunks"])}).encode())
            for ch in hit[0]["chunks"]:                                      # push delivery, in order
                self._deliver(dst, "", ch)

    def _serve(self, sock):
        conn = self._Conn(sock)
        sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
        conn.send(b'INFO {"server_id":"STUB","server_name":"b200-nats-stub","version":"2.10.12","proto":1,"headers":false,"max_payload":67108864}\r\n')
        buf = b""
        try:
            while True:
                while b"\r\n" not in buf:
                    d = sock.recv(1 << 16)
                    if not d:
                        return
                    buf += d
                line, buf = buf.split(b"\r\n", 1)
                op = line.split(b" ", 1)[0].upper()
                if op == b"PUB":
                    parts = line.decode().split()
                    subject, n = parts[1], int(parts[-1]); reply = parts[2] if len(parts) == 4 else ""
                    while len(buf) < n + 2:
                        d = sock.recv(1 << 20)
                        if not d:
                            return
                        buf += d
                    payload, buf = buf[:n], buf[n + 2:]
                    if subject.startswith("$JS.API."):
                        self._js_api(subject, reply, payload)
                    else:
                        self._deliver(subject, reply, payload)
                elif op == b"SUB":
                    parts = line.decode().split()
                    subject, sid = parts[1], parts[-1]; queue = parts[2] if len(parts) == 4 else ""
                    with self.lock:
                        self.subs.setdefault(subject, []).append((conn, sid, queue))
                elif op == b"UNSUB":
                    parts = line.decode().split()
                    if len(parts) == 3:
                        conn.limits[parts[1]] = int(parts[2])
                    else:
                        self._unsub(conn, parts[1])
                elif op == b"PING":
                    conn.send(b"PONG\r\n")
                # CONNECT, PONG: nothing to do
        except OSError:
            return
        finally:
            with self.lock:
                for subj in list(self.subs):
                    self.subs[subj] = [t for t in self.subs[subj] if t[0] is not conn]


class Client:
    """Blocking NATS client: request() publishes with a fresh _INBOX reply subject and waits for one message
    (what nc.Request does in the reference's example client, README.md:506-562).  Thread-safe."""

    def __init__(self, url):
        host, port = url.split("://")[-1].split(":")
        self.sock = socket.create_connection((host, int(port)))
        self.sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
        self.wlock = threading.Lock(); self.lock = threading.Lock()
        self.waiting = {}                   # sid -> [event, payload]
        self.sid = 0
        self.buf = b""
        self._line()                        # INFO
        self._send(b'CONNECT {"verbose":false,"name":"b200-test-client"}\r\n')
        self.reader = threading.Thread(target=self._read_loop, daemon=True); self.reader.start()

    def _send(self, b):
        with self.wlock:
            self.sock.sendall(b)

    def _line(self):
        while b"\r\n" not in self.buf:
            d = self.sock.recv(1 << 16)
            if not d:
                raise ConnectionError("closed")
            self.buf += d
        line, self.buf = self.buf.split(b"\r\n", 1)
        return line

    def _read_loop(self):
        try:
            while True:
                line = self._line()
                if line.startswith(b"MSG "):
                    parts = line.decode().split(); n = int(parts[-1]); sid = parts[2]
                    while len(self.buf) < n + 2:
                        d = self.sock.recv(1 << 20)
                        if not d:
                            return
                        self.buf += d
                    payload, self.buf = self.buf[:n], self.buf[n + 2:]
                    with self.lock:
                        w = self.waiting.pop(sid, None)
                    if w:
                        w[1] = payload; w[0].set()
                elif line == b"PING":
                    self._send(b"PONG\r\n")
        except (OSError, ConnectionError):
            return

    def request(self, subject, payload=b"", timeout=120.0):
        if isinstance(payload, str):
            payload = payload.encode()
        with self.lock:
            self.sid += 1; sid = str(self.sid)
            w = [threading.Event(), None]; self.waiting[sid] = w
        inbox = f"_INBOX.{random.getrandbits(64):016x}.{sid}"
        self._send(f"SUB {inbox} {sid}\r\nUNSUB {sid} 1\r\nPUB {subject} {inbox} {len(payload)}\r\n".encode() + payload + b"\r\n")
        if not w[0].wait(timeout):
            raise TimeoutError(f"no reply on {subject}")
        return w[1]

    def close(self):
        try:
            self.sock.shutdown(socket.SHUT_RDWR)
        except OSError:
            pass
        self.sock.close()


if __name__ == "__main__":
    ap = argparse.ArgumentParser(); ap.add_argument("--port", type=int, default=4222); a = ap.parse_args()
    b = Broker(port=a.port).start()
    print(f"stub broker on {b.url}", flush=True)
    try:
        while True:
            time.sleep(3600)
    except KeyboardInterrupt:
        b.stop()
