"""The C-ABI from C and C++: the header compiles as plain C, the library exports every declared
symbol, and (GPU) a C++ host program calling it like the shim headers do matches the oracle."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HDR = os.path.join(ROOT, "include", "sr_capi.h")
LIBDIR = os.path.join(ROOT, "stereo_reconstruction_b200")
GXX = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
GCC = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else "gcc"


def test_header_is_plain_c():
    subprocess.check_call([GCC, "-std=c99", "-Wall", "-Werror", "-fsyntax-only", "-x", "c", HDR])


def test_library_exports_every_declared_symbol():
    import stereo_reconstruction_b200 as s
    declared = set(re.findall(r"\b(sr_[a-z0-9_]+)\s*\(", open(HDR).read())) - {"sr_image", "sr_ctx"}
    lib = s.capi.lib()
    missing = [n for n in sorted(declared) if not hasattr(lib, n)]
    assert not missing, missing
    assert set(s.capi.SYMBOLS) == declared, set(s.capi.SYMBOLS) ^ declared
    assert b"sm_100a" in lib.sr_version()


def _build_client(tmp):
    from oracle import orc
    orc.build()
    exe = os.path.join(tmp, "capi_client")
    subprocess.check_call([GXX, "-O1", "-std=c++17", os.path.join(ROOT, "tests", "cpp", "capi_client.cpp"), "-o", exe,
                           "-L" + LIBDIR, "-lsr_b200", "-L" + os.path.join(ROOT, "oracle"), "-lsr_oracle",
                           "-Wl,-rpath," + LIBDIR, "-Wl,-rpath," + os.path.join(ROOT, "oracle")])
    return exe


def test_cpp_client_links(tmp_path):
    exe = _build_client(str(tmp_path))
    r = subprocess.run([exe], capture_output=True, text=True)
    # no GPU here: the library must refuse loudly (no CPU fallback), not compute
    import torch
    if not torch.cuda.is_available():
        assert r.returncode == 2 and "no CUDA device" in r.stderr


@pytest.mark.gpu
def test_cpp_client_parity(tmp_path):
    exe = _build_client(str(tmp_path))
    r = subprocess.run([exe], capture_output=True, text=True)
    print(r.stdout, r.stderr)
    assert r.returncode == 0 and r.stdout.strip().endswith("OK")
