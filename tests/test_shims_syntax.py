"""shims/*.cc — the three gRPC programs of the reference over the role cores — cannot be linked in this image
(no gRPC, no protoc).  They are syntax-checked against API-shaped stand-in headers (shims/standin/): every call
they make into the generated classes and into include/sl_roles.h / sl_gpu.h must at least type-check.
SURVEY.md section 8(f) n1; the handlers they replace: src/master.cc:70-118,204-215, src/worker.cc:41-104,181-192,
src/file_server.cc:95-148."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("name", ["worker_shim.cc", "master_shim.cc", "file_server_shim.cc"])
def test_shim_type_checks_against_standin_headers(name):
    src = os.path.join(ROOT, "shims", name)
    cmd = ["g++", "-std=c++17", "-fsyntax-only", "-Wall", "-Werror", "-I", os.path.join(ROOT, "shims", "standin"),
           "-I", os.path.join(ROOT, "include"), src]
    p = subprocess.run(cmd, capture_output=True, text=True)
    assert p.returncode == 0, p.stderr[-4000:]
    # and without gRPC on the include path the file compiles to nothing instead of failing
    p = subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-I", os.path.join(ROOT, "include"), src],
                       capture_output=True, text=True)
    assert p.returncode == 0, p.stderr[-2000:]
