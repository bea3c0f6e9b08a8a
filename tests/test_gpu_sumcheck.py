"""GPU parity tests of the bit-sliced GF(2^128) sumcheck prover (b2ntt_sumcheck_*) through the C ABI:
round-by-round equality with the CPU restatement on the same seeded inputs and challenges, and the reference's own
verifier protocol (sumcheck/test/test.cu:16-96) at the sizes the reference instantiates."""
import numpy as np
import pytest

from test_oracle import run_sumcheck_protocol

pytestmark = pytest.mark.gpu


def _rand(n, seed):
    return np.random.default_rng(seed).integers(0, 2**32, size=n, dtype=np.uint32)


@pytest.mark.parametrize("comp", [2, 3, 4])
@pytest.mark.parametrize("transposed", [False, True])
def test_messages_match_the_restatement_every_round(b2, oracle, cuda, comp, transposed):
    for num_vars in (5, 6, 7, 10, 11, 13):
        evals = _rand(4 * comp << num_vars, 7 * num_vars + comp)
        fed = oracle.bitslice(evals, 128) if transposed else evals
        sc = b2.Sumcheck(fed, num_vars, comp, data_is_transposed=transposed)
        ref = oracle.Sumcheck(evals, num_vars, comp)
        rng = np.random.default_rng(num_vars)
        for rnd in range(num_vars + 1):
            s, p = sc.this_round_messages()
            rs, rp = ref.this_round_messages()
            assert np.array_equal(s, rs) and np.array_equal(p, rp), (num_vars, comp, rnd)
            if rnd < num_vars:
                ch = rng.integers(0, 2**32, 4, dtype=np.uint32)
                sc.move_to_next_round(ch), ref.move_to_next_round(ch)
        sc.close()


@pytest.mark.parametrize("comp", [2, 3, 4])
def test_reference_verifier_protocol_with_brute_force(b2, oracle, cuda, comp):
    """test_sumcheck<NUM_VARS, comp, false> at a size the brute-force evaluation finishes in seconds"""
    num_vars = 14
    evals = _rand(4 * comp << num_vars, comp)
    run_sumcheck_protocol(b2.Sumcheck(evals, num_vars, comp), evals, num_vars, comp, oracle, seed=comp)


@pytest.mark.parametrize("num_vars,comp", [(20, 2), (20, 3), (20, 4), (24, 2), (24, 4)])
def test_reference_sizes_verifier_consistency(b2, oracle, cuda, num_vars, comp):
    """the reference's instantiations (sumcheck.cuh:12): every round's claim checks of test.cu:44-50, device-resident
    input adopted in place; the final claim is compared with the restatement at 20 variables"""
    import torch

    g = torch.Generator(device=cuda).manual_seed(num_vars * comp)
    d = torch.randint(-2**31, 2**31 - 1, (4 * comp << num_vars,), dtype=torch.int32, device=cuda, generator=g)
    host = d.cpu().numpy().view(np.uint32).copy() if num_vars == 20 else None
    sc = b2.Sumcheck(d, num_vars, comp, data_is_transposed=False)
    final = run_sumcheck_protocol(sc, None, num_vars, comp, oracle, seed=num_vars, brute_force=False)
    if host is not None:
        ref = oracle.Sumcheck(host, num_vars, comp)
        rng = np.random.default_rng(num_vars)
        for _ in range(num_vars):
            ref.move_to_next_round(rng.integers(0, 2**32, 4, dtype=np.uint32))
        assert np.array_equal(ref.this_round_messages()[0], final)
    sc.close()


@pytest.mark.parametrize("num_vars,comp", [(24, 2), (24, 3), (28, 2)])
def test_device_p1_equals_the_derived_value_every_round(b2, oracle, cuda, num_vars, comp):
    """with p(1) derived from the previous round the per-round checks are tautological.  set_recompute makes every
    device round compute p(1) from the data (as compute_compositions does) and fail with B2NTT_ECHECK on any difference; the
    messages must also equal those of the default (derived) run word for word.  28 variables = the reference's largest
    instantiation (sumcheck.cuh:12)."""
    import torch

    g = torch.Generator(device=cuda).manual_seed(7 * num_vars + comp)
    d = torch.randint(-2**31, 2**31 - 1, (4 * comp << num_vars,), dtype=torch.int32, device=cuda, generator=g)
    runs = []
    for recompute in (True, False):
        sc = b2.Sumcheck(d.clone(), num_vars, comp, data_is_transposed=False)
        if recompute:
            sc.set_recompute(True)
        rng = np.random.default_rng(num_vars)
        msgs = []
        for _ in range(num_vars):
            s_, p_ = sc.this_round_messages()  # raises B2NTTError(B2NTT_ECHECK) on a mismatch
            msgs.append((s_.copy(), p_.copy()))
            sc.move_to_next_round(rng.integers(0, 2**32, 4, dtype=np.uint32))
        msgs.append(tuple(x.copy() for x in sc.this_round_messages()))
        runs.append(msgs)
        sc.close()
    for (s0, p0), (s1, p1) in zip(*runs):
        assert np.array_equal(s0, s1) and np.array_equal(p0, p1)


def test_sumcheck_argument_errors(b2, cuda):
    with pytest.raises(b2.B2NTTError):
        b2.Sumcheck(np.zeros(4 * 2 << 4, dtype=np.uint32), 4, 2)
    with pytest.raises(b2.B2NTTError):
        b2.Sumcheck(np.zeros(4 * 5 << 6, dtype=np.uint32), 6, 5)
    sc = b2.Sumcheck(_rand(4 * 2 << 5, 1), 5, 2)
    for _ in range(5):
        sc.move_to_next_round(np.ones(4, dtype=np.uint32))
    with pytest.raises(b2.B2NTTError):
        sc.move_to_next_round(np.ones(4, dtype=np.uint32))


@pytest.mark.parametrize("comp", [2, 3, 4])
def test_messages_match_the_reference_cuda_prover(b2, oracle, cuda, comp):
    """oracle/_ref/ref_sumcheck = the reference's own Sumcheck<20, comp, false> (sumcheck/sumcheck.cuh) built for sm_100a
    from /root/reference by oracle/Makefile, run on this GPU with seeded inputs: every round's sum and points must be
    equal word for word."""
    import os
    import subprocess

    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "ref_sumcheck")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/ref_sumcheck not built (needs /root/reference at build time)")
    nv = 20
    out = subprocess.run([exe, str(nv), str(comp)], check=True, capture_output=True, text=True, timeout=600).stdout.splitlines()
    rounds = [ln.split() for ln in out if ln.startswith("round ")]
    assert len(rounds) == nv + 1
    evals = oracle.mt19937(0xDEADBEEF + nv + comp, 4 * comp << nv)
    chal = oracle.mt19937(0xC0FFEE + nv + comp, 4 * nv).reshape(nv, 4)
    sc = b2.Sumcheck(evals, nv, comp)
    for rnd, f in enumerate(rounds):
        assert int(f[1]) == rnd and f[2] == "sum" and f[7] == "points"
        ref_sum = np.array([int(x, 16) for x in f[3:7]], dtype=np.uint32)
        ref_pts = np.array([int(x, 16) for x in f[8:]], dtype=np.uint32).reshape(comp + 1, 4)
        s, p = sc.this_round_messages()
        assert np.array_equal(s, ref_sum) and np.array_equal(p, ref_pts), (comp, rnd)
        if rnd < nv:
            sc.move_to_next_round(chal[rnd])
    sc.close()
