"""Test double of the reference dispatcher (server.go `c` command: JobAlloc 161-191, NewConnect 267-308): hands out
pieces highest index first, one per connection, parses "success;id" / "fail;id[;why]" by prefix from 100-byte reads,
never re-queues a failed piece, stops when every piece has been reported.

Opt-in robustness (the Go-side patch is sketched in INTEGRATION.md, defaults = the reference's behaviour):
`requeue=N` hands a piece out again, up to N more times, when its worker reports `fail` or drops the connection without a
report (server.go:276-283 drops it); `buf=N` reads reports of up to N bytes (server.go:271 reads 100)."""
import socket
import threading
import time


class FakeServer:
    def __init__(self, directory, pieces, args, requeue=0, buf=100):
        if args == "264":
            args = "-threads 4 -vcodec libx264"                 # server.go:69-70
        elif args == "265":
            args = "-threads 4 -vcodec libx265 -crf 26"         # server.go:67-68
        self.dir, self.args = directory, args
        self.order = list(range(pieces))                         # initMap
        self.remain = {i: str(i) for i in range(pieces)}         # remainMap
        self.sent, self.success, self.failed = [], [], []
        self.requeue, self.buf = requeue, buf
        self._retry, self._attempts, self._open = [], {}, 0      # pieces waiting for another hand-out, hand-outs per piece, unreported jobs
        self.t_sent, self.t_success = [], []                     # wall-clock times of the hand-outs and of the `success` reports
        self.sock = socket.socket()
        self.sock.bind(("127.0.0.1", 0))
        self.sock.listen(16)
        self.addr = self.sock.getsockname()
        self._next = pieces - 1                                  # JobAlloc: num-- and descending
        self._lock = threading.Lock()
        self._thread = threading.Thread(target=self._accept, daemon=True)
        self._thread.start()

    def _accept(self):
        self.sock.settimeout(0.2)
        while True:
            with self._lock:
                if self._next < 0 and not self._retry and self._open == 0:
                    break
            try:
                c, _ = self.sock.accept()
            except socket.timeout:
                continue
            except OSError:
                break
            with self._lock:
                if self._next >= 0:
                    piece = self.remain[self.order[self._next]]
                    self._next -= 1
                elif self._retry:
                    piece = self._retry.pop(0)
                else:
                    c.close()
                    continue
                self._attempts[piece] = self._attempts.get(piece, 0) + 1
                self._open += 1
            job = self.dir + ";" + piece + ";" + self.args       # server.go:179
            self.sent.append(job)
            self.t_sent.append(time.perf_counter())
            c.sendall(job.encode())
            threading.Thread(target=self._reader, args=(c, piece), daemon=True).start()
        self.sock.close()

    def _reader(self, c, piece):
        try:
            data = c.recv(self.buf).decode(errors="replace")     # server.go:271-272
        except OSError:
            data = ""
        with self._lock:
            self._open -= 1
            if data[0:7] != "success" and self._attempts[piece] <= self.requeue:
                self._retry.append(piece)                        # opt-in: another worker (or this one, later) gets it again
            elif data[0:4] == "fail":                            # server.go:276-283: logged, dropped, not re-queued
                self.failed.append(data.split(";"))
            elif data[0:7] == "success":                         # server.go:284-292
                n = int(data.split(";")[1])
                self.success.append(n)
                self.t_success.append(time.perf_counter())
                self.remain.pop(n, None)
            else:
                self.failed.append(["lost", ""])
        c.close()

    def join(self, timeout=60):
        self._thread.join(timeout)
        return len(self.remain) == 0
