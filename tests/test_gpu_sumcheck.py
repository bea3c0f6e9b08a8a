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
