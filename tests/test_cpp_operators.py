"""The C++ host side above the C ABI (include/srb/gpu_operators.h + include/srb/compat/*.h): operators with the reference's
Operator interface, driven by a PipelineDriver-style loop in tests/cpp/test_operators.cpp. On CPU the program must compile,
link against libsrb.so and report the missing GPU through the status convention; on the GPU box it runs its checks."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "build", "test_operators")
SRC = os.path.join(ROOT, "tests", "cpp", "test_operators.cpp")
HEADERS = [os.path.join(ROOT, "include", "srb", h) for h in
           ("gpu_operators.h", "srb_c_api.h", "compat/status.h", "compat/column.h", "compat/chunk.h", "compat/operator.h")]


def _build():
    from starrocks_b200 import build as B
    so = B.build() if hasattr(B, "build") else None
    lib_dir = os.path.join(ROOT, "starrocks_b200")
    assert os.path.exists(os.path.join(lib_dir, "libsrb.so")), so
    newest = max(os.path.getmtime(p) for p in [SRC] + HEADERS)
    if os.path.exists(EXE) and os.path.getmtime(EXE) >= newest:
        return EXE
    os.makedirs(os.path.dirname(EXE), exist_ok=True)
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), SRC, "-L", lib_dir, "-lsrb",
                           "-Wl,-rpath,$ORIGIN/../starrocks_b200", "-o", EXE])
    return EXE


def test_cpp_binding_compiles_and_links():
    exe = _build()
    # without a GPU srb_ctx_create returns a status (no exception, no crash) and the program says so
    import torch
    if not torch.cuda.is_available():
        r = subprocess.run([exe, "--allow-no-gpu"], capture_output=True, text=True, timeout=120)
        assert r.returncode == 0, r.stdout + r.stderr
        assert "no usable GPU" in r.stdout


@pytest.mark.gpu
def test_cpp_operators_on_gpu():
    exe = _build()
    r = subprocess.run([exe], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.count("PASS") == 5 and " 0 failed" in r.stdout, r.stdout
