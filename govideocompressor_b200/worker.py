"""Host mirror of the reference worker loop (client.go:21-99, 132-175) around convert().

The Go worker stays unchanged in a real deployment (only convert()'s body becomes a cgo call); this
twin exists because neither `go` nor the worker's transports (wget from nginx, FTP upload) exist on the build / GPU
boxes.  It speaks the same bytes on the control connection:

    server -> worker   "<dir>;<id>;<args>"                 one write, read into a 100-byte buffer (client.go:43-45)
    worker -> server   "success;<id>" | "fail;<id>[;why]"  one write, then both sides close (client.go:60-69)

and keeps the same file conventions: input `<id>.mp4` in the working directory, output `c<id>.mp4` + `c<id>.mp4.log`,
uploaded as `<dir>/c<id>.mp4.log` and `<dir>/<id>.mp4`, local files removed afterwards (client.go:133-139).
`fetch` / `store` replace wget / FTP; the defaults are the local-filesystem data plane.
"""
import os
import shutil
import socket


def local_fetch(src_root):
    def fetch(path, dst):                     # client.go:92-94: wget http://SERVER_IP/<dir>/<id>.mp4
        shutil.copyfile(os.path.join(src_root, path), dst)
    return fetch


def local_store(dst_root):
    def store(remote_path, local_file):       # client.go:159,169: ftp.Stor(path+"/"+name, file)
        os.makedirs(os.path.dirname(os.path.join(dst_root, remote_path)), exist_ok=True)
        shutil.copyfile(local_file, os.path.join(dst_root, remote_path))
    return store


def download_and_convert(path, args, convert, fetch, store, workdir):
    """client.go:84-99 + 132-175.  Returns the status string for the server."""
    file_dir, file_name = path.split("/")[0], path.split("/")[1]
    piece = file_name.split(".")[0]
    if len(args) <= 0:
        return "fail;" + piece + ";empty convert args"            # client.go:87-91
    cwd = os.getcwd()
    os.chdir(workdir)
    try:
        try:
            fetch(path, file_name)
        except Exception:
            pass                                                   # client.go:93-94: wget's result is ignored
        convert(file_name, args)                                   # client.go:96
        name = "c" + file_name
        try:
            if not os.path.exists(name + ".log"):
                return "fail;" + piece + ";log file missing"      # client.go:153-157
            try:
                store(file_dir + "/" + name + ".log", name + ".log")
            except Exception:
                return "fail;" + piece + ";log upload failed"     # client.go:159-162
            if not os.path.exists(name):
                return "fail;" + piece                             # client.go:163-167
            try:
                store(file_dir + "/" + name[1:], name)            # uploaded under the ORIGINAL name (client.go:169)
            except Exception:
                return "fail;" + piece
            return "success;" + piece                              # client.go:173
        finally:
            for f in (name, name + ".log", name[1:]):              # client.go:133-139
                if os.path.exists(f):
                    os.remove(f)
    finally:
        os.chdir(cwd)


def job_buffer_bytes():
    """client.go:43 reads the job into 100 bytes, which truncates long argument strings; GVC_JOB_BUF lifts that (opt-in,
    SURVEY.md 8f.4 — the Go-side change is sketched in INTEGRATION.md)."""
    try:
        return max(100, int(os.environ.get("GVC_JOB_BUF", "100")))
    except ValueError:
        return 100


def serve_one_job(server_addr, convert, fetch, store, workdir, timeout=10.0):
    """One iteration of client.go:37-80.  Returns the status sent, or None when no job was received."""
    try:
        conn = socket.create_connection(server_addr, timeout=timeout)
    except OSError:
        return None
    try:
        try:
            data = conn.recv(job_buffer_bytes())                   # client.go:43-45
        except OSError:
            return None                                            # client.go:46-50: a failed read ends this round of the loop
        if not data:
            return None
        parts = data.decode(errors="replace").split(";")
        if len(parts) <= 1:
            return None
        path = "/".join(parts[:2]) + ".mp4"                        # client.go:54
        status = download_and_convert(path, parts[2] if len(parts) > 2 else "", convert, fetch, store, workdir)
        conn.sendall(status.encode())
        return status
    finally:
        conn.close()


def run_worker(server_addr, convert, fetch, store, workdir, max_idle=2):
    """Loops like client.go:37 until the server stops answering `max_idle` times in a row."""
    done, idle = [], 0
    while idle < max_idle:
        st = serve_one_job(server_addr, convert, fetch, store, workdir, timeout=1.0)
        if st is None:
            idle += 1
        else:
            idle = 0
            done.append(st)
    return done
