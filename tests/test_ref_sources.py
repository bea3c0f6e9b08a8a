"""The reference's OWN test sources (finite_fields/tests/{test_fanpaartower,tests,test_kernels}.cu, ntt/tests/test_ntt.cu),
compiled unmodified from where they lie against include/ulvt + libulvt_gpu.a by tests/cpp/ref_tests/Makefile (a Catch2
stand-in supplies TEST_CASE / REQUIRE / BENCHMARK), run as binaries: the drop-in check of SURVEY 8b.  The host-only
executables run anywhere; the GPU ones on the B200 box."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "tests", "cpp", "ref_tests", "_build")


def _run(name, *args, timeout=1500):
    exe = os.path.join(BUILD, name)
    if not os.path.exists(exe):
        pytest.skip(f"{exe} not built (the reference sources were absent at build time)")
    r = subprocess.run([exe, *args], capture_output=True, text=True, timeout=timeout)
    tail = "\n".join((r.stdout + r.stderr).splitlines()[-15:])
    assert r.returncode == 0, tail
    last = [l for l in r.stdout.splitlines() if l.startswith("test cases:")][-1]
    cases = int(last.split("|")[0].split(":")[1])
    assert cases > 0, f"no test case matched {args}: {tail}"
    return last


def test_reference_test_fanpaartower_host():
    """finite_fields/tests/test_fanpaartower.cu: compact, packed and bit-sliced known answers (131 assertions)"""
    assert "failed: 0" in _run("test_fanpaartower")


def test_reference_finite_field_tests_host():
    """finite_fields/tests/tests.cu: interleave, packed / bit-sliced products, CPU benchmarks (35 assertions)"""
    assert "failed: 0" in _run("finite_field_tests")


@pytest.mark.gpu
def test_reference_test_kernels_gpu(cuda):
    """finite_fields/tests/test_kernels.cu: the multiply / inverse kernels over 2^24 elements"""
    assert "failed: 0" in _run("test_kernels")


@pytest.mark.gpu
@pytest.mark.parametrize("flt", ["[cap]", "[additive][0]~[slow]", "[additive][2]", "[ntt]~[additive]", "[roundtrip]"])
def test_reference_ntt_tests_gpu(cuda, flt):
    """ntt/tests/test_ntt.cu: capability check; additive NTT digests log_h 1..28 (rate 1) and 1..27 (rate 4); BabyBear NTT
    digests log 1..27 and its round trip -- the reference's own assertions against its own golden tables"""
    assert "failed: 0" in _run("ntt_tests", flt)
