"""CPU: the worker-loop mirror against the dispatcher double, with convert() stubbed (no GPU): the bytes on the wire, the
file conventions and the failure conventions of client.go / server.go."""
import os
import threading

import pytest

from govideocompressor_b200 import worker
from protocol_twin import FakeServer


def stub_convert(calls, fail_ids=()):
    def convert(path, args):
        calls.append((path, args))
        with open("c" + path + ".log", "w") as f:
            f.write("stub log for " + path)
        if path.split(".")[0] not in fail_ids:
            with open(path, "rb") as src, open("c" + path, "wb") as dst:
                dst.write(b"ENC:" + src.read())
        return 0
    return convert


def make_pieces(root, d, n):
    os.makedirs(os.path.join(root, d))
    for i in range(n):
        with open(os.path.join(root, d, f"{i}.mp4"), "wb") as f:
            f.write(b"piece%d" % i)


def test_jobs_descend_and_results_land_under_original_names(tmp_path):
    src, dst, work = (os.path.join(tmp_path, x) for x in ("www", "ftp", "work"))
    os.makedirs(work)
    make_pieces(src, "12clip", 5)
    server = FakeServer("12clip", 5, "264")
    calls = []
    done = worker.run_worker(server.addr, stub_convert(calls), worker.local_fetch(src), worker.local_store(dst), work)
    assert server.join()
    assert server.sent == [f"12clip;{i};-threads 4 -vcodec libx264" for i in (4, 3, 2, 1, 0)]        # highest index first
    assert done == [f"success;{i}" for i in (4, 3, 2, 1, 0)]
    assert calls[0] == ("4.mp4", "-threads 4 -vcodec libx264")                                       # bare file name, args verbatim
    for i in range(5):
        assert open(os.path.join(dst, "12clip", f"{i}.mp4"), "rb").read() == b"ENC:piece%d" % i       # c<id>.mp4 uploaded as <id>.mp4
        assert os.path.exists(os.path.join(dst, "12clip", f"c{i}.mp4.log"))
    assert os.listdir(work) == []                                                                    # worker cleans up


def test_failed_piece_is_reported_and_not_requeued(tmp_path):
    src, dst, work = (os.path.join(tmp_path, x) for x in ("www", "ftp", "work"))
    os.makedirs(work)
    make_pieces(src, "d", 3)
    server = FakeServer("d", 3, "-vcodec libx264 -s 1280x720")
    done = worker.run_worker(server.addr, stub_convert([], fail_ids=("1",)), worker.local_fetch(src), worker.local_store(dst), work)
    server.join(5)
    assert done == ["success;2", "fail;1", "success;0"]           # missing c1.mp4 => "fail;1" (client.go:163-167)
    assert sorted(server.success) == [0, 2] and server.failed == [["fail", "1"]]
    assert list(server.remain) == [1]                            # the dispatcher never hands piece 1 out again
    assert os.path.exists(os.path.join(dst, "d", "c1.mp4.log"))  # the log is uploaded before the missing file is noticed


def test_two_workers_share_the_queue_and_empty_args_fail_fast(tmp_path):
    src, dst = os.path.join(tmp_path, "www"), os.path.join(tmp_path, "ftp")
    make_pieces(src, "d", 8)
    server = FakeServer("d", 8, "264")
    results = [None, None]

    def run(k):
        w = os.path.join(tmp_path, f"work{k}")
        os.makedirs(w)
        results[k] = worker.run_worker(server.addr, stub_convert([]), worker.local_fetch(src), worker.local_store(dst), w)
    ts = [threading.Thread(target=run, args=(k,)) for k in range(2)]
    # workers chdir into their own directory: run them one after the other in-process (cwd is process-wide)
    for t in ts:
        t.start()
        t.join()
    assert server.join()
    assert sorted(int(s.split(";")[1]) for r in results for s in r) == list(range(8))
    st = worker.download_and_convert("d/3.mp4", "", stub_convert([]), worker.local_fetch(src), worker.local_store(dst), str(tmp_path))
    assert st.startswith("fail;3;")                              # client.go:87-91


def test_job_string_longer_than_100_bytes_is_truncated_like_the_reference():
    long_args = "-threads 4 -vcodec libx264 " + "-an " * 30
    server = FakeServer("d", 1, long_args)
    import socket
    c = socket.create_connection(server.addr)
    data = c.recv(100)
    assert len(data) == 100 and not data.decode().endswith(long_args[-10:])   # client.go:43 reads 100 bytes, silently
    c.sendall(b"fail;0;truncated")
    c.close()
    server.join(5)
