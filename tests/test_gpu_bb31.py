"""GPU parity tests of the BabyBear radix-2 NTT (b2ntt_bb31_*) through the C ABI: the reference's own digests
(bb31_ntt_hashes, ntt/tests/test_ntt.cu:21-50 -> tests/golden/bb31_ntt_md5.json), the CPU oracle on seeded inputs, the
reference's round-trip property (test_ntt.cu:152-187) and the error contract of NTTConfRad2 (ntt/nttconf.cuh:32-39)."""
import hashlib
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = json.load(open(os.path.join(ROOT, "tests", "golden", "bb31_ntt_md5.json")))["md5"]


@pytest.mark.parametrize("log_lens", [range(1, 13), range(13, 21), range(21, 25), [25, 26, 27]])
def test_bb31_golden_md5(b2, oracle, cuda, log_lens):
    """TEST_CASE("NTTBB31 all input lengths"): input BB31(mt19937(0xdeadbeef + log_len)()), MD5 over asUInt32() of the output"""
    for log_len in log_lens:
        x = oracle.mt19937(0xDEADBEEF + log_len, 1 << log_len)
        ntt = b2.BabyBearNTT(log_len, 137, 27)
        out = ntt.apply(x)
        assert hashlib.md5(out.tobytes()).hexdigest() == GOLD[str(log_len)], log_len
        ntt.close()


@pytest.mark.parametrize("log_len,gen,lgo", [(1, 137, 27), (5, 137, 27), (11, 137, 27), (12, 137, 27), (16, 31, 27), (17, 137, 20), (18, 137, 27)])
def test_bb31_vs_oracle(b2, oracle, cuda, log_len, gen, lgo):
    x = np.random.default_rng(log_len).integers(0, 2**32, size=1 << log_len, dtype=np.uint32)
    ntt = b2.BabyBearNTT(log_len, gen, lgo)
    assert np.array_equal(ntt.apply(x), oracle.bb31_ntt(x, log_len, gen, lgo))
    rev = np.random.default_rng(log_len + 1).integers(0, 2**32, size=1 << log_len, dtype=np.uint32)
    assert np.array_equal(ntt.apply(rev, bit_reversed=True), oracle.bb31_ntt(rev, log_len, gen, lgo, input_bit_reversed=True))
    ntt.close()


def test_bb31_round_trip_device(b2, cuda):
    """TEST_CASE("NTTBB31 round trip") at log 24, device-resident: forward with g, inverse with 1/g, scale by 1/n"""
    import torch

    L, P = 24, b2.api.BB31_P
    x = np.random.RandomState(0xAABBCCDD).randint(0, 2**32, 1 << L, dtype=np.uint32) % P
    fwd = b2.BabyBearNTT(L, 137, 27)
    inv = b2.BabyBearNTT(L, pow(137, P - 2, P), 27)
    d = torch.from_numpy(b2.bb31_encode(x).view(np.int32)).to(cuda)
    y = fwd.exec(d)
    assert not torch.equal(y, d)
    z = b2.bb31_decode(inv.exec(y).cpu().numpy().view(np.uint32)).astype(np.uint64)
    z = z * pow(1 << L, P - 2, P) % P
    assert np.array_equal(z.astype(np.uint32), x)
    fwd.close(), inv.close()


def test_bb31_errors(b2, cuda):
    for args in ((0, 137, 27), (28, 137, 27), (12, 137, 11)):
        with pytest.raises(b2.B2NTTError):
            b2.BabyBearNTT(*args)
    ntt = b2.BabyBearNTT(8)
    with pytest.raises(b2.B2NTTError):
        ntt.apply(np.zeros(100, dtype=np.uint32))
    ntt.close()
