"""General-basis GF(2^128) additive NTT: b2ntt_general_* through the C ABI against the CPU oracle
(oracle/antt_general.c).  The reference has no full-width transform, so the oracle is pinned to (1) the MD5-pinned
reference-basis restatement and (2) the definition (direct evaluation of the novel-basis polynomial)."""
import ctypes as C

import numpy as np
import pytest


def _rand(n, seed):
    return np.random.default_rng(seed).integers(0, 2**32, size=n, dtype=np.uint32)


def _basis(log_h, seed):
    """random GF(2)-independent basis (rank checked by elimination over the 128-bit integers)"""
    rng = np.random.default_rng(seed)
    rows, piv = [], []
    while len(rows) < log_h:
        w = rng.integers(0, 2**32, size=4, dtype=np.uint32)
        v = int.from_bytes(w.tobytes(), "little")
        r = v
        for p in piv:
            r = min(r, r ^ p)
        if r:
            piv.append(r)
            piv.sort(reverse=True)
            rows.append(w)
    return np.stack(rows)


# ---- CPU: pin the oracle ---------------------------------------------------------------------------
@pytest.mark.parametrize("log_h", [1, 2, 5, 9, 12])
def test_oracle_reference_basis_matches_pinned_restatement(oracle, log_h):
    x = _rand(4 << log_h, 10 + log_h)
    assert np.array_equal(oracle.antt_general(x, log_h), oracle.antt_forward_full128(x, log_h, 0))
    # limb 0 of the input = the reference's test stream => MD5(limb 0 of the output) is the reference's digest (F8)
    x = x.reshape(-1, 4).copy()
    x[:, 0] = oracle.reference_input(log_h, 0)
    y = oracle.antt_general(x, log_h).reshape(-1, 4)
    assert oracle.md5_hex(np.ascontiguousarray(y[:, 0])) == oracle.golden_md5(0, log_h)


@pytest.mark.parametrize("log_h", [1, 2, 3, 5, 7])
def test_oracle_butterflies_match_the_definition(oracle, log_h):
    for seed in (1, 2):
        basis = _basis(log_h, 100 * seed + log_h)
        x = _rand(4 << log_h, seed)
        assert np.array_equal(oracle.antt_general(x, log_h, basis), oracle.antt_general(x, log_h, basis, direct=True))
    x = _rand(4 << log_h, 3)
    assert np.array_equal(oracle.antt_general(x, log_h), oracle.antt_general(x, log_h, direct=True))


def test_oracle_inverse_and_dependent_basis(oracle):
    basis = _basis(10, 5)
    x = _rand(4 << 10, 6)
    y = oracle.antt_general(x, 10, basis)
    assert np.array_equal(oracle.antt_general(y, 10, basis, inverse=True), x)
    bad = basis.copy()
    bad[7] = bad[1] ^ bad[4] ^ bad[6]
    with pytest.raises(ValueError):
        oracle.antt_general(x, 10, bad)


def test_cabi_rejects_bad_general_arguments(b2):
    L = b2.lib()
    h = C.c_void_p()
    assert L.b2ntt_general_create(C.byref(h), 0, None, 0) == -1
    assert L.b2ntt_general_create(C.byref(h), 35, None, 0) == -1
    assert L.b2ntt_general_create(C.byref(h), 33, None, 0) == -1  # the reference basis has 32 elements
    bad = _basis(6, 1)
    bad[5] = bad[0] ^ bad[3]
    assert L.b2ntt_general_create(C.byref(h), 6, bad.ctypes.data, 0) == -1  # dependent basis: EINVAL, with or without a device
    zero = np.zeros((3, 4), dtype=np.uint32)
    assert L.b2ntt_general_create(C.byref(h), 3, zero.ctypes.data, 0) == -1
    assert L.b2ntt_general_exec(None, None, None, 0, None) == -1


# ---- GPU: parity ------------------------------------------------------------------------------------
def _dev(x, cuda):
    import torch

    return torch.from_numpy(x.view(np.int32)).to(cuda)


def _host(t):
    return t.cpu().numpy().view(np.uint32)


@pytest.mark.gpu
@pytest.mark.parametrize("log_h", [1, 2, 3, 6, 9, 10, 11, 12, 13, 16])  # 11 = first size on the CTA-level product
def test_gpu_general_vs_oracle(b2, oracle, cuda, log_h):
    basis = _basis(log_h, 40 + log_h)
    g = b2.GeneralNTT(log_h, basis)
    assert np.array_equal(g.constants(), oracle.antt_general_constants(log_h, basis))
    x = _rand(4 << log_h, log_h)
    d = _dev(x, cuda)
    import torch

    out = torch.empty_like(d)
    g.exec(d, out)
    want = oracle.antt_general(x, log_h, basis)
    assert np.array_equal(_host(out), want)
    assert np.array_equal(_host(d), x)  # out of place leaves the input alone
    g.exec(out, inverse=True)  # in place
    assert np.array_equal(_host(out), x)
    if log_h <= 6:
        assert np.array_equal(want, oracle.antt_general(x, log_h, basis, direct=True))
    g.close()


@pytest.mark.gpu
@pytest.mark.parametrize("log_h", [4, 11, 20])
def test_gpu_general_reference_basis_equals_fast_path(b2, oracle, cuda, log_h):
    """basis = NULL reproduces b2ntt_exec (GF(2^128), log_rate 0) bit for bit, and so the reference's digests."""
    x = _rand(4 << log_h, log_h).reshape(-1, 4)
    x[:, 0] = oracle.reference_input(log_h, 0)
    x = x.reshape(-1)
    g, plan = b2.GeneralNTT(log_h), b2.Plan(128, log_h, 0)
    a = _host(g.exec(_dev(x, cuda)))
    b = _host(plan.exec(_dev(x, cuda)))
    assert np.array_equal(a, b)
    assert oracle.md5_hex(np.ascontiguousarray(a.reshape(-1, 4)[:, 0])) == oracle.golden_md5(0, log_h)
    g.close(), plan.close()


@pytest.mark.gpu
def test_gpu_general_properties_large(b2, cuda):
    """log_h = 22 (64 MiB): linearity over GF(2) and inverse(forward(x)) = x on a random basis."""
    import torch

    log_h = 22
    g = b2.GeneralNTT(log_h, _basis(log_h, 7))
    gen = torch.Generator(device=cuda).manual_seed(5)
    x = torch.randint(-(2**31), 2**31 - 1, (4 << log_h,), dtype=torch.int32, device=cuda, generator=gen)
    y = torch.randint(-(2**31), 2**31 - 1, (4 << log_h,), dtype=torch.int32, device=cuda, generator=gen)
    fx, fy, fxy = torch.empty_like(x), torch.empty_like(x), torch.empty_like(x)
    g.exec(x, fx), g.exec(y, fy), g.exec(x ^ y, fxy)
    assert torch.equal(fx ^ fy, fxy)
    g.exec(fx, inverse=True)
    assert torch.equal(fx, x)
    g.close()
