"""Builds and runs the C++ host-mirror test (tests/cpp/test_ligero_e2e.cpp).  The build is a CPU check (header compiles,
links against both libraries); the run needs a GPU."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "cpp", "test_ligero_e2e")


def _build():
    from lumenos_b200 import build as lb
    from oracle import oracle as O
    lb.build(); O.build()
    src = os.path.join(ROOT, "tests", "cpp", "test_ligero_e2e.cpp")
    if os.path.exists(EXE) and os.path.getmtime(EXE) > max(os.path.getmtime(src), os.path.getmtime(os.path.join(ROOT, "include", "lumenos_b200.hpp"))):
        return
    cmd = ["g++", "-O2", "-std=c++17", "-Wall", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "oracle"), src, "-o", EXE,
           "-L", os.path.join(ROOT, "lumenos_b200"), "-llumenos_b200", "-L", os.path.join(ROOT, "oracle"), "-llumenos_oracle",
           "-Wl,-rpath," + os.path.join(ROOT, "lumenos_b200"), "-Wl,-rpath," + os.path.join(ROOT, "oracle"), "-ldl", "-lpthread", "-lrt"]
    subprocess.check_call(cmd)


def test_cpp_mirror_builds():
    _build()
    assert os.path.exists(EXE)


@pytest.mark.gpu
@pytest.mark.parametrize("args", [["8", "256", "16", "6"], ["6", "64", "16", "6"]])
def test_cpp_mirror_commit_prove(args):
    _build()
    r = subprocess.run([EXE] + args, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "ok:" in r.stdout
