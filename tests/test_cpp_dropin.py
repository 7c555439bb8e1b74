"""The C++ drop-in host layer (include/blur_algorithms.hpp): the reference's own function names
over the C ABI, exercised by tests/cpp/dropin_test.cpp (a C++ program modelled on the reference's
Test(flag) dispatcher, Source.cpp:574-608) against the CPU oracle."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "cpp", "dropin_test")
SRC = EXE + ".cpp"


def build_dropin_test() -> str:
    deps = [SRC, os.path.join(ROOT, "include", "blur_algorithms.hpp"), os.path.join(ROOT, "include", "blur_b200.h")]
    if os.path.exists(EXE) and all(os.path.getmtime(EXE) >= os.path.getmtime(d) for d in deps):
        return EXE
    env = {k: v for k, v in os.environ.items() if k not in ("CC", "CXX")}
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle")], check=True, env=env)
    subprocess.run(["g++", "-std=c++17", "-O1", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "oracle"), SRC,
                    "-L", os.path.join(ROOT, "blur_algorithms_b200"), "-lblur_b200", "-L", os.path.join(ROOT, "oracle"), "-loracle",
                    "-Wl,-rpath,$ORIGIN/../../blur_algorithms_b200", "-Wl,-rpath,$ORIGIN/../../oracle", "-o", EXE],
                   check=True, env=env)
    return EXE


def test_cpp_dropin_links_and_host_helpers_match_oracle():
    r = subprocess.run([build_dropin_test(), "link"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "FAIL" not in r.stdout


@pytest.mark.gpu
def test_cpp_dropin_parity_on_gpu():
    r = subprocess.run([build_dropin_test()], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "all ok" in r.stdout and "FAIL " not in r.stdout
