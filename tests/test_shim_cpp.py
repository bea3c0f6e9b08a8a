"""The C++ drop-in surface include/ulvt/...: tests/cpp/shim_test.cu uses the reference's own
class names (AdditiveNTTConf / AdditiveNTT / NTTData / FanPaarTowerField / BitsliceUtils /
multiply_unrolled) over libb2ntt.so and mirrors run_and_check_additive_ntt (ntt/tests/test_ntt.cu:191-217)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "tests", "cpp", "shim_test")


def test_shim_headers_exist_and_cite_the_reference():
    for rel, cite in [("ulvt/ntt/additive_ntt.cuh", "additive_ntt.cuh:176-270"), ("ulvt/ntt/nttconf.cuh", "nttconf.cuh:9-114"), ("ulvt/ntt/gpuntt.cuh", "gpuntt.cuh:126-209"),
                      ("ulvt/finite_fields/baby_bear.cuh", "risc0_baby_bear.h:40-187"),
                      ("ulvt/finite_fields/binary_tower.cuh", "binary_tower.cuh:19-262"), ("ulvt/utils/bitslicing.cuh", "bitslicing.cuh:8-104"),
                      ("ulvt/utils/common.cuh", "common.cuh:3-32"), ("ulvt/sumcheck/sumcheck.cuh", "sumcheck/sumcheck.cuh:10-300"),
                      ("ulvt/finite_fields/circuit_generator/unrolled/binary_tower_unrolled.cuh", "binary_tower_unrolled.cuh:4-5"),
                      ("ulvt/finite_fields/binary_tower_simd.cuh", "binary_tower_simd.cuh:11-188"),
                      ("ulvt/finite_fields/circuit_generator/utils/utils.hpp", "utils.{hpp,cpp}:8-32")]:
        assert cite in open(os.path.join(ROOT, "include", rel)).read(), rel


def test_shim_test_program_builds():
    subprocess.run(["make", "-C", os.path.join(ROOT, "tests", "cpp")], check=True, stdout=subprocess.DEVNULL)
    assert os.path.exists(BIN)


@pytest.mark.gpu
def test_shim_test_program_passes_on_gpu(cuda):
    if not os.path.exists(BIN):
        pytest.fail("tests/cpp/shim_test was not built (run __graft_entry__.build())")
    r = subprocess.run([BIN, "24", "22"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert " 0 failed" in r.stdout
