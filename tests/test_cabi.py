"""CPU: libgvc.so loads and exports every symbol include/gvc.h declares; argument checking that needs no GPU."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    names = []
    for f in os.listdir(os.path.join(ROOT, "include")):
        if f.endswith(".h"):
            text = open(os.path.join(ROOT, "include", f)).read()
            text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
            names += re.findall(r"\b(gvc_[a-z0-9_]+)\s*\(", text)
    return sorted(set(names))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as g
    g.build()
    import govideocompressor_b200 as gvc
    return gvc.load_library()


def test_header_declares_the_boundary():
    names = declared_functions()
    for must in ("gvc_init", "gvc_shutdown", "gvc_strerror", "gvc_free", "gvc_resize_plan_create",
                 "gvc_resize_frames_device", "gvc_resize_frames_host"):
        assert must in names


def test_every_declared_symbol_is_exported(lib):
    for name in declared_functions():
        assert hasattr(lib, name), f"{name} declared in include/*.h but not exported by libgvc.so"


def test_no_cpu_fallback_without_device(lib):
    """In the CPU container there is no device: init must fail loudly and compute calls must refuse."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("device present")
    assert lib.gvc_init(0) == -2  # GVC_ERR_NO_DEVICE
    assert b"no CUDA device" in lib.gvc_last_error()
    h = ctypes.c_void_p()
    assert lib.gvc_resize_plan_create(ctypes.byref(h), 64, 64, 32, 32, 0, 4) == -7  # GVC_ERR_NOT_INIT
    assert not h.value


def test_strerror(lib):
    assert lib.gvc_strerror(0) == b"ok"
    assert b"fallback" in lib.gvc_strerror(-5)


def test_library_does_not_link_the_oracle():
    import subprocess
    out = subprocess.run(["ldd", os.path.join(ROOT, "govideocompressor_b200", "libgvc.so")], capture_output=True, text=True).stdout
    assert "oracle" not in out and "swscale" not in out
