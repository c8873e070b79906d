"""The reference's matcher unit tests replayed against the C++ facade (include/libspm_b200/spm.hpp).
CPU: the facade + tests compile against the header with -Wall -Wextra -Werror (like the reference's
own test flags, test/jstmap_test.cmake:44).  GPU: the binary runs and every test passes."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "tests", "cpp", "matcher_tests.bin")


def _build():
    libdir = os.path.join(ROOT, "libspm_b200", "lib")
    gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    cmd = [gxx, "-std=c++20", "-O1", "-pedantic", "-Wall", "-Wextra", "-Werror", f"-I{ROOT}/include",
           f"-I{ROOT}/tests/cpp", os.path.join(ROOT, "tests", "cpp", "matcher_tests.cpp"), "-o", BIN,
           f"-L{libdir}", "-lspmb", f"-Wl,-rpath,{libdir}"]
    subprocess.check_call(cmd)


def test_facade_compiles_with_reference_warning_flags():
    _build()
    assert os.path.exists(BIN)


@pytest.mark.gpu
def test_reference_matcher_tests_pass_on_gpu():
    _build()
    r = subprocess.run([BIN], capture_output=True, text=True, timeout=120)
    print(r.stdout[-3000:])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-1000:]
    assert "0 failures" in r.stdout


def test_host_hit_ordering_equals_std_sort():
    """libspm_b200/csrc/hit_order.hpp (the host's share of a search pass) against std::stable_sort + std::unique: pure C++."""
    gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    exe = os.path.join(ROOT, "tests", "cpp", "hit_order_test.bin")
    subprocess.check_call([gxx, "-std=c++17", "-O2", "-Wall", "-Wextra", "-Werror", os.path.join(ROOT, "tests", "cpp", "hit_order_test.cpp"),
                           "-o", exe])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-1000:]
    assert " 0 failures" in r.stdout
