"""CPU tests: the oracle against the reference's golden vectors (pins parity), the oracle against the
reference's own host field code when oracle/_ref is built, and the CPU emulator of the CUDA pass
phases against the oracle."""
import ctypes as C
import json
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = json.load(open(os.path.join(ROOT, "tests", "golden", "additive_ntt_md5.json")))["md5"]
KATS = json.load(open(os.path.join(ROOT, "tests", "golden", "field_kats.json")))


def _rand(n, seed):
    return np.random.default_rng(seed).integers(0, 2**32, size=n, dtype=np.uint32)


# ---- NTT: golden MD5 table of ntt/tests/test_ntt.cu:52-124 -----------------------------------------
@pytest.mark.parametrize("log_rate,max_h", [(0, 22), (2, 20)])
def test_oracle_reproduces_golden_md5(oracle, log_rate, max_h):
    for log_h in range(1, max_h + 1):
        x = oracle.reference_input(log_h, log_rate)
        y = oracle.antt_forward(x, log_h, log_rate)
        assert oracle.md5_hex(y) == GOLD[str(log_rate)][str(log_h)], (log_rate, log_h)
        assert oracle.golden_md5(log_rate, log_h) == GOLD[str(log_rate)][str(log_h)]


def test_oracle_plain_and_fast_products_agree(oracle):
    x = _rand(1 << 10, 3)
    y = oracle.antt_forward(x, 10, 1)
    oracle.lib.orc_antt_set_plain(1)
    try:
        assert np.array_equal(oracle.antt_forward(x, 10, 1), y)
    finally:
        oracle.lib.orc_antt_set_plain(0)


def test_gf128_ntt_is_limbwise_and_matches_full_product(oracle):
    """with subfield twiddles the GF(2^128) transform is four interleaved GF(2^32)
    transforms; the definitional form multiplies with the full 128-bit tower product."""
    log_h = 8
    x = _rand((1 << log_h) * 4, 11)
    y = oracle.antt_forward(x, log_h, 1, 4)
    assert np.array_equal(y, oracle.antt_forward_full128(x, log_h, 1))
    for l in range(4):
        yl = oracle.antt_forward(np.ascontiguousarray(x.reshape(-1, 4)[:, l]), log_h, 1)
        assert np.array_equal(y.reshape(-1, 4)[:, l], yl)


def test_inverse_inverts(oracle):
    for limbs in (1, 4):
        for log_h, log_rate in [(1, 0), (5, 0), (9, 2), (12, 1)]:
            x = _rand((1 << log_h) * limbs, log_h)
            y = oracle.antt_forward(x, log_h, log_rate, limbs)
            n = (1 << log_h) * limbs
            for c in range(1 << log_rate):
                assert np.array_equal(oracle.antt_inverse(y[c * n:(c + 1) * n], log_h, log_rate, c, limbs), x)


# ---- field: known-answer vectors of finite_fields/tests -----------------------------------------------
def test_field_kats(oracle):
    L = oracle.lib
    for a, b, c in KATS["mul32"]:
        assert L.orc_mul32(a, b) == c and L.orc_mul32_fast(a, b) == c
    for a, c in KATS["square32"]:
        assert L.orc_square32(a) == c
    for a, c in KATS["inverse32"]:
        assert L.orc_inverse32(a) == c
    for h, a, b, c in KATS["packed_mul"]:
        assert L.orc_mul_packed32(a, b, h) == c, (h, hex(a), hex(b))


def test_bitsliced_kats(oracle):
    for case in KATS["bitsliced"]:
        h = case["h"]
        w = 1 << h
        a = np.zeros(w, dtype=np.uint32)
        b = np.zeros(w, dtype=np.uint32)
        a[:len(case["a"])] = case["a"]
        b[:len(case["b"])] = case["b"]
        r = oracle.unbitslice(oracle.mul_bitsliced(oracle.bitslice(a, w), oracle.bitslice(b, w), h), w)
        for i, v in case["result"].items():
            assert r[int(i)] == v, (h, i)
        # the compact product agrees with the bit-sliced one (ties compact GF(2^128) to the KATs)
        if h == 7:
            assert np.array_equal(oracle.mul128(a, b).reshape(-1), r)
        else:
            assert np.array_equal(oracle.mul32(a, b), r)


def test_oracle_matches_reference_host_code(oracle):
    R = oracle.ref_lib()
    if R is None:
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    rng = np.random.default_rng(5)
    for _ in range(300):
        a, b = (int(v) for v in rng.integers(0, 2**32, 2))
        assert R.ref_mul(a, b, 5) == oracle.lib.orc_mul32(a, b)
        assert R.ref_square(a, 5) == oracle.lib.orc_square32(a)
        assert R.ref_inverse(a, 5) == oracle.lib.orc_inverse32(a)
    for h in (5, 7):
        w = 1 << h
        a, b = _rand(w, 1), _rand(w, 2)
        out = np.empty(w, dtype=np.uint32)
        R.ref_mul_unrolled(a, b, out, h)
        assert np.array_equal(out, oracle.mul_bitsliced(a, b, h))
        t = a.copy()
        R.ref_bitslice(t, w)
        assert np.array_equal(t, oracle.bitslice(a, w))
    a, b = _rand(4, 3), _rand(4, 4)
    out = np.empty(4, dtype=np.uint32)
    R.ref_mul128_words(a, b, out)
    assert np.array_equal(out, oracle.mul128(a, b).reshape(-1))


def test_bb31_oracle_matches_reference_digests(oracle):
    """the BabyBear NTT restatement (oracle/bb31.c) against bb31_ntt_hashes (ntt/tests/test_ntt.cu:21-50)"""
    import hashlib
    import json

    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "bb31_ntt_md5.json")))["md5"]
    for log_len in range(1, 21):
        out = oracle.bb31_ntt(oracle.mt19937(0xDEADBEEF + log_len, 1 << log_len), log_len)
        assert hashlib.md5(out.tobytes()).hexdigest() == gold[str(log_len)], log_len
    # the inverse generator undoes the transform up to the factor n (test_ntt.cu:152-187)
    P, L = oracle.BB31_P, 10
    x = _rand(1 << L, 3) % np.uint32(P)
    y = oracle.bb31_ntt(oracle.bb31_ntt(x, L), L, pow(137, P - 2, P))
    assert np.array_equal((y.astype(np.uint64) * pow(1 << L, P - 2, P) % P).astype(np.uint32), x)


# ---- CPU emulation of the CUDA pass phases (same source as the kernels) -------------------------------
@pytest.fixture(scope="module")
def emul():
    p = os.path.join(ROOT, "tests", "emul", "libb2emul.so")
    import shutil
    import subprocess

    # the emulator compiles the kernels' own phase functions: rebuild whenever a header under csrc/ is newer (make decides), so
    # these tests never run a stale copy; on a box without make / g++ the prebuilt library that travelled with the repo is used
    if shutil.which("make") and shutil.which("g++"):
        subprocess.run(["make", "-s", "-C", os.path.dirname(p)], check=True)
    E = C.CDLL(p)
    u32p = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")
    E.emul_antt.argtypes = [C.c_int] * 5 + [u32p, u32p, C.c_int, C.c_int]
    E.emul_antt_sched.argtypes = [C.c_int] * 5 + [u32p, u32p] + [C.c_int] * 5
    E.emul_constants.argtypes = [C.c_int, C.c_int, u32p]
    E.emul_gf32_mul.argtypes = [C.c_uint32, C.c_uint32]
    E.emul_gf32_mul.restype = C.c_uint32
    E.emul_sliced_mul.argtypes = [u32p, u32p, u32p, C.c_int]
    return E


@pytest.mark.parametrize("field_bits,layout,log_h,log_rate", [
    (32, 0, 6, 0), (32, 0, 12, 0), (32, 0, 15, 2), (128, 0, 4, 0), (128, 0, 10, 0), (128, 0, 13, 1),
    (32, 1, 9, 0), (128, 1, 8, 1), (32, 0, 16, 0),
    (128, 0, 15, 1), (32, 1, 17, 0), (128, 1, 16, 0), (128, 0, 17, 0), (128, 0, 12, 0), (32, 0, 14, 1),
    (32, 0, 20, 0), (128, 0, 18, 1)])  # 6-bit upper passes on 8-line super-lines  # the last six reach the tile pass and the upper passes
def test_emulated_kernels_match_oracle(oracle, emul, field_bits, layout, log_h, log_rate):
    limbs = field_bits // 32
    x = _rand((1 << log_h) * limbs, 17 * log_h + log_rate)
    src = oracle.bitslice(x, field_bits) if layout else x
    out = np.zeros((1 << (log_h + log_rate)) * limbs, dtype=np.uint32)
    assert emul.emul_antt(field_bits, layout, log_h, log_rate, 1, src, out, 0, 0) == 0
    got = oracle.unbitslice(out, field_bits) if layout else out
    ref = oracle.antt_forward(x, log_h, log_rate, limbs)
    assert np.array_equal(got, ref)
    # inverse of coset 0
    n = (1 << log_h) * limbs
    back = np.zeros(n, dtype=np.uint32)
    assert emul.emul_antt(field_bits, layout, log_h, log_rate, 1, np.ascontiguousarray(out[:n]), back, 1, 0) == 0
    assert np.array_equal(back, src)


@pytest.mark.parametrize("field_bits,log_h,log_rate", [(128, 13, 0), (128, 14, 1), (128, 16, 0), (32, 15, 0), (32, 16, 2), (32, 18, 0), (128, 17, 0)])
def test_emulated_block_sliced_pass(oracle, emul, field_bits, log_h, log_rate):
    """schedule 5: the block-sliced lowest pass (b2_bsl.cuh) under the upper passes, forward and inverse"""
    limbs = field_bits // 32
    x = _rand((1 << log_h) * limbs, 41 * log_h + log_rate)
    out = np.zeros((1 << (log_h + log_rate)) * limbs, dtype=np.uint32)
    assert emul.emul_antt_sched(field_bits, 0, log_h, log_rate, 1, x, out, 0, 0, 0, 0, 5) == 0
    assert np.array_equal(out, oracle.antt_forward(x, log_h, log_rate, limbs))
    n = (1 << log_h) * limbs
    back = np.zeros(n, dtype=np.uint32)
    assert emul.emul_antt_sched(field_bits, 0, log_h, log_rate, 1, np.ascontiguousarray(out[:n]), back, 1, 0, 0, 0, 5) == 0
    assert np.array_equal(back, x)


def test_plan_constants_and_products_match_oracle(oracle, emul):
    for log_h, log_rate in [(20, 0), (20, 2), (28, 0), (28, 4)]:
        c = np.zeros(32 * 32, dtype=np.uint32)
        emul.emul_constants(log_h, log_rate, c)
        ref = oracle.antt_constants(log_h, log_rate)
        assert np.array_equal(c.reshape(32, 32)[:log_h, :ref.shape[1]], ref)
    rng = np.random.default_rng(9)
    for _ in range(200):
        a, b = (int(v) for v in rng.integers(0, 2**32, 2))
        assert emul.emul_gf32_mul(a, b) == oracle.lib.orc_mul32(a, b)
    for h in (2, 3, 4, 5, 6, 7):
        a, b = _rand(1 << h, h), _rand(1 << h, h + 50)
        out = np.empty(1 << h, dtype=np.uint32)
        emul.emul_sliced_mul(a, b, out, h)
        assert np.array_equal(out, oracle.mul_bitsliced(a, b, h))


# ---- sumcheck restatement: the reference's own test protocol (sumcheck/test/test.cu:16-96) ------------------------
def run_sumcheck_protocol(prover, evals_compact, num_vars, comp, oracle, seed, brute_force=True):
    """Drives `prover` (this_round_messages / move_to_next_round) the way test_sumcheck does and applies its REQUIREs."""
    rng = np.random.default_rng(seed)
    challenges, expected = [], None
    for rnd in range(num_vars):
        s, pts = prover.this_round_messages()
        assert rnd == 0 or np.array_equal(s, expected), ("claim", rnd)
        assert np.array_equal(s, pts[0] ^ pts[1]), ("sum = p(0) + p(1)", rnd)
        ch = rng.integers(0, 2**32, 4, dtype=np.uint32)
        challenges.append(ch)
        expected = oracle.sc_interpolate(ch, pts)
        prover.move_to_next_round(ch)
    s, _ = prover.this_round_messages()
    assert np.array_equal(s, expected), "final claim"
    if brute_force:
        assert np.array_equal(oracle.sc_eval_composition(evals_compact, np.stack(challenges), num_vars, comp), expected)
    return expected


@pytest.mark.parametrize("comp", [2, 3, 4])
def test_sumcheck_restatement_passes_the_reference_verifier(oracle, comp):
    for num_vars in (1, 5, 6, 9):
        evals = _rand(4 * comp << num_vars, 100 * comp + num_vars)
        run_sumcheck_protocol(oracle.Sumcheck(evals, num_vars, comp), evals, num_vars, comp, oracle, seed=num_vars)


class RefHostSumcheck:
    """Replays the data flow of the reference's Sumcheck class (sumcheck/sumcheck.cuh:136-299) on the CPU, calling the
    REFERENCE's own batch helpers (sumcheck/core/core.cu, via oracle/_ref) where the class launches its kernels
    (compute_compositions, kernels.cuh:5-102; fold_large_list_halves, kernels.cu:9-30)."""

    def __init__(self, R, oracle, evals_compact, num_vars, comp):
        self.R, self.nv, self.C, self.n0, self.round = R, num_vars, comp, 1 << num_vars, 0
        self.ev = oracle.bitslice(evals_compact, 128).copy()
        self.cpu = np.zeros(128 * comp, dtype=np.uint32)
        self.coeff = np.zeros(128 * (comp + 1), dtype=np.uint32)
        for t in range(comp + 1):
            R.ref_sc_repeat_value(self.coeff[128 * t:128 * (t + 1)], np.array([t, 0, 0, 0], dtype=np.uint32))
        if self.n0 == 32:
            self.cpu[:] = self.ev

    def this_round_messages(self):
        R, C, IP = self.R, self.C, self.C + 1
        n = self.n0 >> self.round
        padded = max(32, n)
        s, pts = np.zeros(4, dtype=np.uint32), np.zeros((IP, 4), dtype=np.uint32)
        fsum = np.zeros(IP * 128, dtype=np.uint32)
        prod = np.zeros(128, dtype=np.uint32)
        if padded == 32:
            R.ref_sc_compose(self.cpu, prod, C, 32)
            for t in range(IP):
                folded = np.zeros(128 * C, dtype=np.uint32)
                for c in range(C):
                    R.ref_sc_fold_small(self.cpu[128 * c:128 * (c + 1)], folded[128 * c:128 * (c + 1)], self.coeff[128 * t:128 * (t + 1)], n)
                R.ref_sc_compose(folded, fsum[128 * t:128 * (t + 1)], C, 32)
            R.ref_sc_compute_sum(s, prod, n)
            for t in range(IP):
                R.ref_sc_compute_sum(pts[t], fsum[128 * t:128 * (t + 1)], n // 2)
            return s, pts
        rows = padded // 32
        total = np.zeros(128, dtype=np.uint32)
        for row in range(rows):
            R.ref_sc_compose(self.ev[128 * row:], prod, C, self.n0)
            total ^= prod
            if row < rows // 2:
                folded = np.zeros(IP * C * 128, dtype=np.uint32)
                for c in range(C):
                    lo = 128 * ((self.n0 // 32) * c + row)
                    hi = lo + 128 * (rows // 2)
                    for t in range(IP):
                        d = 128 * (c * IP + t)
                        R.ref_sc_fold_batch(self.ev[lo:lo + 128], self.ev[hi:hi + 128], folded[d:d + 128], self.coeff[128 * t:128 * (t + 1)], 1)
                for t in range(IP):
                    R.ref_sc_compose(folded[128 * t:], prod, C, IP * 32)
                    fsum[128 * t:128 * (t + 1)] ^= prod
        R.ref_sc_compute_sum(s, total, 32)
        for t in range(IP):
            R.ref_sc_compute_sum(pts[t], fsum[128 * t:128 * (t + 1)], 32)
        return s, pts

    def move_to_next_round(self, challenge):
        R, C = self.R, self.C
        n = self.n0 >> self.round
        coeff = np.zeros(128, dtype=np.uint32)
        R.ref_sc_repeat_value(coeff, np.ascontiguousarray(challenge, dtype=np.uint32))
        if n <= 32:
            for c in range(C):
                col = self.cpu[128 * c:128 * (c + 1)]
                R.ref_sc_fold_small(col, col, coeff, n)
        else:
            half_rows = n >> 6
            for row in range(half_rows):
                for c in range(C):
                    lo = 128 * row + 4 * self.n0 * c
                    hi = lo + 128 * half_rows
                    R.ref_sc_fold_batch(self.ev[lo:lo + 128], self.ev[hi:hi + 128], self.ev[lo:lo + 128], coeff, 0)
            if n // 2 == 32:
                for c in range(C):
                    self.cpu[128 * c:128 * (c + 1)] = self.ev[4 * self.n0 * c:4 * self.n0 * c + 128]
        self.round += 1


@pytest.mark.parametrize("comp", [2, 3, 4])
def test_sumcheck_restatement_matches_reference_batch_code(oracle, comp):
    """every round message of the compact restatement == the reference's own bit-sliced batch code, same inputs"""
    R = oracle.ref_lib()
    if R is None or not hasattr(R, "ref_sc_compose"):
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    for num_vars in (5, 7, 8):
        evals = _rand(4 * comp << num_vars, 31 * comp + num_vars)
        ref, mine = RefHostSumcheck(R, oracle, evals, num_vars, comp), oracle.Sumcheck(evals, num_vars, comp)
        rng = np.random.default_rng(num_vars + comp)
        for rnd in range(num_vars + 1):
            rs, rp = ref.this_round_messages()
            s, p = mine.this_round_messages()
            assert np.array_equal(s, rs) and np.array_equal(p, rp), (num_vars, comp, rnd)
            if rnd < num_vars:
                ch = rng.integers(0, 2**32, 4, dtype=np.uint32)
                ref.move_to_next_round(ch), mine.move_to_next_round(ch)
