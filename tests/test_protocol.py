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


def flaky_convert(calls, fail_first=()):
    """Fails a piece (no output file) the first time it is seen, succeeds afterwards."""
    seen = set()

    def convert(path, args):
        calls.append(path)
        piece = path.split(".")[0]
        with open("c" + path + ".log", "w") as f:
            f.write("log")
        if piece in fail_first and piece not in seen:
            seen.add(piece)
            return 0
        with open(path, "rb") as src, open("c" + path, "wb") as dst:
            dst.write(b"ENC:" + src.read())
        return 0
    return convert


def test_opt_in_requeue_hands_a_failed_piece_out_again(tmp_path):
    """SURVEY 8f.4, behind a flag: the reference drops a failed piece (previous test); with requeue=1 it comes back once."""
    src, dst, work = (os.path.join(tmp_path, x) for x in ("www", "ftp", "work"))
    os.makedirs(work)
    make_pieces(src, "d", 3)
    server = FakeServer("d", 3, "264", requeue=1)
    calls = []
    done = worker.run_worker(server.addr, flaky_convert(calls, fail_first=("1",)), worker.local_fetch(src), worker.local_store(dst), work)
    assert server.join(5)                                         # nothing remains
    assert done == ["success;2", "fail;1", "success;0", "success;1"]
    assert calls == ["2.mp4", "1.mp4", "0.mp4", "1.mp4"] and server.failed == []
    assert open(os.path.join(dst, "d", "1.mp4"), "rb").read() == b"ENC:piece1"


def test_opt_in_requeue_gives_up_after_its_limit_and_covers_lost_reports(tmp_path):
    src, dst, work = (os.path.join(tmp_path, x) for x in ("www", "ftp", "work"))
    os.makedirs(work)
    make_pieces(src, "d", 2)
    server = FakeServer("d", 2, "264", requeue=2)
    done = worker.run_worker(server.addr, stub_convert([], fail_ids=("0",)), worker.local_fetch(src), worker.local_store(dst), work)
    server.join(5)
    assert done == ["success;1", "fail;0", "fail;0", "fail;0"]      # first hand-out + two more, then it is reported as failed
    assert server.failed == [["fail", "0"]] and list(server.remain) == [0]
    # a worker that takes a job and dies without reporting: the piece is handed out again instead of being lost
    import socket
    make_pieces(src, "e", 1)
    server = FakeServer("e", 1, "264", requeue=1)
    c = socket.create_connection(server.addr)
    assert c.recv(100).decode() == "e;0;-threads 4 -vcodec libx264"
    c.close()
    done = worker.run_worker(server.addr, stub_convert([]), worker.local_fetch(src), worker.local_store(dst), work)
    assert server.join(5) and done == ["success;0"]


def test_opt_in_job_buffer_carries_a_long_argument_string(tmp_path, monkeypatch):
    src, dst, work = (os.path.join(tmp_path, x) for x in ("www", "ftp", "work"))
    os.makedirs(work)
    make_pieces(src, "d", 1)
    long_args = "-threads 4 -vcodec libx264 " + "-an " * 30
    monkeypatch.setenv("GVC_JOB_BUF", "4096")
    server = FakeServer("d", 1, long_args, buf=4096)
    calls = []
    done = worker.run_worker(server.addr, stub_convert(calls), worker.local_fetch(src), worker.local_store(dst), work)
    assert server.join(5) and done == ["success;0"]
    assert calls == [("0.mp4", long_args)]                        # nothing was cut off (client.go:43 would have kept 100 bytes)


def test_the_documented_argument_grammar_is_the_parsers_grammar():
    """INTEGRATION.md's option table is the contract a maintainer of the Go side reads (the job string of server.go:179 is free
    form).  Every example in it must parse (together with the mandatory codec option), every listed option name too, and
    options the table does not list must be refused."""
    import os
    import re
    import emu_lib
    doc = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "INTEGRATION.md")).read()
    rows = re.findall(r"^\| (`-[^|]*) \| `([^`]*)` \|", doc, flags=re.M)
    assert len(rows) >= 15, "option table not found in INTEGRATION.md"
    listed = set()
    for names, example in rows:
        opts = re.findall(r"`(-[^`]+)`", names)
        listed.update(opts)
        args = example if example.split()[0] in ("-vcodec", "-c:v", "-codec:v") else "-vcodec libx264 " + example
        rc, opt, err = emu_lib.parse_args(args)
        assert rc == 0, f"documented example {example!r} is refused: {err}"
        for o in opts:                                     # every spelling of the option, with the example's value
            alt = " ".join([o] + example.split()[1:])
            alt = alt if o in ("-vcodec", "-c:v", "-codec:v") else "-vcodec libx264 " + alt
            assert emu_lib.parse_args(alt)[0] == 0, f"documented option {o!r} is refused"
    assert emu_lib.parse_args("-vcodec libx264 -b:v 2M")[1]["bitrate"] == 2000000
    assert emu_lib.parse_args("-vcodec libx264 -b:v 800k")[1]["bitrate"] == 800000
    assert emu_lib.parse_args("-vcodec libx264 -an")[1]["drop_audio"] == 1
    assert emu_lib.parse_args("-threads 4 -vcodec libx264")[0] == 0                    # server.go:69-70, the "264" shorthand
    assert emu_lib.parse_args("-threads 4 -vcodec libx265 -crf 26")[0] == -5           # server.go:67-68, the "265" shorthand: unsupported
    for bad in ("-bogus 1", "-bf 2", "-vcodec libx264 -acodec mp3", "-vcodec libx264 -b:v fast", "-vcodec libx264 -vf hflip", "-crf 23"):
        assert emu_lib.parse_args(bad)[0] != 0, f"{bad!r} should be refused"
    # the parser knows no option the table does not list
    src = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "govideocompressor_b200", "csrc", "ff_args.cpp")).read()
    known = set(re.findall(r'k == "(-[^"]+)"', src))
    assert known == listed, f"parser and document disagree: {sorted(known ^ listed)}"
