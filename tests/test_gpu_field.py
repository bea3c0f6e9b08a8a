"""GPU parity tests of the element-wise tower multiply (b2ntt_mul) and the layout conversion (b2ntt_bitslice) through the
C ABI: against the CPU oracle on seeded inputs, against the reference's known-answer vectors
(finite_fields/tests/test_binary_tower.cu, test_binary_tower_unrolled.cu -> tests/golden/field_kats.json), on ragged
sizes, and through field identities at the size of config 5."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KATS = json.load(open(os.path.join(ROOT, "tests", "golden", "field_kats.json")))


def _dev(x, cuda):
    import torch

    return torch.from_numpy(np.ascontiguousarray(x).view(np.int32)).to(cuda)


def _host(t):
    return t.cpu().numpy().view(np.uint32)


def _rand(n, seed):
    return np.random.default_rng(seed).integers(0, 2**32, size=n, dtype=np.uint32)


def test_mul32_known_answers(b2, cuda):
    a = np.array([k[0] for k in KATS["mul32"]], dtype=np.uint32)
    b = np.array([k[1] for k in KATS["mul32"]], dtype=np.uint32)
    c = np.array([k[2] for k in KATS["mul32"]], dtype=np.uint32)
    assert np.array_equal(_host(b2.mul(_dev(a, cuda), _dev(b, cuda), 32)), c)
    sq = np.array([k[0] for k in KATS["square32"]], dtype=np.uint32)
    assert np.array_equal(_host(b2.mul(_dev(sq, cuda), _dev(sq, cuda), 32)), np.array([k[1] for k in KATS["square32"]], dtype=np.uint32))
    x = np.array([k[0] for k in KATS["inverse32"] if k[0]], dtype=np.uint32)
    xi = np.array([k[1] for k in KATS["inverse32"] if k[0]], dtype=np.uint32)
    assert np.all(_host(b2.mul(_dev(x, cuda), _dev(xi, cuda), 32)) == 1)


def test_bitsliced_known_answers(b2, oracle, cuda):
    """the bit-sliced vectors of test_binary_tower_unrolled.cu, run in both layouts"""
    for case in KATS["bitsliced"]:
        h = case["h"]
        if h not in (5, 7):
            continue
        w = 1 << h
        a = np.zeros(w, dtype=np.uint32)
        b = np.zeros(w, dtype=np.uint32)
        a[:len(case["a"])] = case["a"]
        b[:len(case["b"])] = case["b"]
        da, db = _dev(a, cuda), _dev(b, cuda)
        compact = _host(b2.mul(da, db, w))
        b2.bitslice_(da, w), b2.bitslice_(db, w)
        sliced = b2.mul(da, db, w, b2.api.LAYOUT_BITSLICED)
        b2.bitslice_(sliced, w, inverse=True)
        for i, v in case["result"].items():
            assert compact[int(i)] == v and _host(sliced)[int(i)] == v, (h, i)
        assert np.array_equal(compact, _host(sliced))


@pytest.mark.parametrize("n", [1, 2, 31, 32, 33, 257, 1000, 32 * 33 + 5, 4096])
def test_mul128_compact_ragged_sizes(b2, oracle, cuda, n):
    """compact GF(2^128): any element count, the words past the last element stay untouched"""
    import torch

    a, b = _rand(4 * n, n).reshape(-1, 4), _rand(4 * n, n + 1).reshape(-1, 4)
    out = torch.full((4 * n + 64,), 0x5A5A5A5A, dtype=torch.int32, device=cuda)
    b2.mul(_dev(a, cuda), _dev(b, cuda), 128, out=out[:4 * n])
    assert np.array_equal(_host(out[:4 * n]).reshape(-1, 4), oracle.mul128(a, b))
    assert bool((out[4 * n:] == 0x5A5A5A5A).all())


@pytest.mark.parametrize("n", [1, 31, 33, 1023, 1025, 4096 * 3 + 7])
def test_mul32_compact_ragged_sizes(b2, oracle, cuda, n):
    import torch

    a, b = _rand(n, n), _rand(n, n + 1)
    out = torch.full((n + 64,), 0x5A5A5A5A, dtype=torch.int32, device=cuda)
    b2.mul(_dev(a, cuda), _dev(b, cuda), 32, out=out[:n])
    assert np.array_equal(_host(out[:n]), oracle.mul32(a, b))
    assert bool((out[n:] == 0x5A5A5A5A).all())


@pytest.mark.parametrize("field_bits", [8, 16, 32])
def test_mul_compact_small_fields(b2, oracle, cuda, field_bits):
    """FanPaarTowerField<3..5> in 32-bit containers (one element per word)"""
    n = 3000
    mask = np.uint32((1 << field_bits) - 1)
    a, b = _rand(n, 1) & mask, _rand(n, 2) & mask
    got = _host(b2.mul(_dev(a, cuda), _dev(b, cuda), field_bits))
    h = {8: 3, 16: 4, 32: 5}[field_bits]
    ref = np.array([oracle.lib.orc_mul_packed32(int(x), int(y), h) if h < 5 else oracle.lib.orc_mul32_fast(int(x), int(y)) for x, y in zip(a, b)], dtype=np.uint32)
    assert np.array_equal(got, ref)


@pytest.mark.parametrize("field_bits,n_blocks", [(32, 1), (32, 129), (128, 1), (128, 33), (128, 31), (128, 70)])
def test_mul_bitsliced_vs_oracle(b2, oracle, cuda, field_bits, n_blocks):
    a, b = _rand(n_blocks * field_bits, 11), _rand(n_blocks * field_bits, 12)
    got = _host(b2.mul(_dev(a, cuda), _dev(b, cuda), field_bits, b2.api.LAYOUT_BITSLICED))
    assert np.array_equal(got, oracle.mul_bitsliced(a, b, 5 if field_bits == 32 else 7))


@pytest.mark.parametrize("width", [32, 64, 128])
def test_bitslice_roundtrip_vs_oracle(b2, oracle, cuda, width):
    for n_blocks in (1, 3, 8, 33, 100):
        x = _rand(n_blocks * width, width + n_blocks)
        d = _dev(x, cuda)
        b2.bitslice_(d, width)
        assert np.array_equal(_host(d), oracle.bitslice(x, width))
        b2.bitslice_(d, width, inverse=True)
        assert np.array_equal(_host(d), x)


def test_mul128_field_identities_at_config5_size(b2, cuda):
    """config 5 size (2^26 products here, 1 GiB per operand): commutativity, distributivity, the unit, and the compact
    product equals the bit-sliced product of the bit-sliced operands"""
    import torch

    n = 1 << 26
    g = torch.Generator(device=cuda).manual_seed(55)
    a, b, c = (torch.randint(-2**31, 2**31 - 1, (n * 4,), dtype=torch.int32, device=cuda, generator=g) for _ in range(3))
    ab = b2.mul(a, b, 128)
    assert torch.equal(ab, b2.mul(b, a, 128))
    lhs = b2.mul(a, b ^ c, 128)
    lhs ^= ab
    assert torch.equal(lhs, b2.mul(a, c, 128))
    del lhs, c
    one = torch.zeros_like(a)
    one.view(-1, 4)[:, 0] = 1
    assert torch.equal(b2.mul(a, one, 128), a)
    del one
    b2.bitslice_(a, 128), b2.bitslice_(b, 128)
    s = b2.mul(a, b, 128, b2.api.LAYOUT_BITSLICED)
    b2.bitslice_(s, 128, inverse=True)
    assert torch.equal(s, ab)


def test_config5_full_size_vs_oracle(b2, oracle, cuda):
    """config 5 at its stated size: 2^28 bit-sliced GF(2^128) pairs (2^23 blocks of 32, 4 GiB per operand).  The four
    reference products of finite_fields/tests/test_fanpaartower.cu:204-273 are planted in the first and the last block;
    a strided sample of 2^16 blocks is compared with the CPU oracle's bit-sliced multiply."""
    import torch

    n_blocks, W = 1 << 23, 128
    g = torch.Generator(device=cuda).manual_seed(528)
    a = torch.randint(-2**31, 2**31 - 1, (n_blocks * W,), dtype=torch.int32, device=cuda, generator=g)
    b = torch.randint(-2**31, 2**31 - 1, (n_blocks * W,), dtype=torch.int32, device=cuda, generator=g)
    case = next(c for c in KATS["bitsliced"] if c["h"] == 7 and len(c["a"]) == 16)
    ka, kb = np.zeros(W, dtype=np.uint32), np.zeros(W, dtype=np.uint32)
    ka[:16], kb[:16] = case["a"], case["b"]  # four 128-bit operands, then zero lanes
    ka_s, kb_s = oracle.bitslice(ka, W), oracle.bitslice(kb, W)
    for blk in (0, n_blocks - 1):
        a[blk * W:(blk + 1) * W] = _dev(ka_s, cuda)
        b[blk * W:(blk + 1) * W] = _dev(kb_s, cuda)
    out = b2.mul(a, b, 128, b2.api.LAYOUT_BITSLICED)
    for blk in (0, n_blocks - 1):
        got = oracle.unbitslice(_host(out[blk * W:(blk + 1) * W]), W)
        for i, v in case["result"].items():
            assert got[int(i)] == v, (blk, i)
        assert not got[16:].any()
    stride = n_blocks >> 16
    idx = (torch.arange(1 << 16, device=cuda) * stride + 3).clamp_(max=n_blocks - 1)
    sa = _host(a.view(n_blocks, W)[idx].reshape(-1))
    sb = _host(b.view(n_blocks, W)[idx].reshape(-1))
    so = _host(out.view(n_blocks, W)[idx].reshape(-1))
    assert np.array_equal(so, oracle.mul_bitsliced(sa, sb, 7))


@pytest.mark.parametrize("field_bits,layout", [(32, 0), (32, 1), (128, 0), (128, 1), (8, 0), (16, 0)])
def test_mul_in_place_aliasing(b2, oracle, cuda, field_bits, layout):
    """include/b2ntt.h: d_out may alias d_a or d_b (multiply_unrolled's destination may alias an input,
    sumcheck/core/core.cu:21)"""
    limbs = max(1, field_bits // 32)
    n = 32 * 257
    mask = np.uint32(0xFFFFFFFF if field_bits >= 32 else (1 << field_bits) - 1)
    a, b = _rand(n * limbs, 71) & mask, _rand(n * limbs, 72) & mask
    ref = _host(b2.mul(_dev(a, cuda), _dev(b, cuda), field_bits, layout))
    da, db = _dev(a, cuda), _dev(b, cuda)
    b2.mul(da, db, field_bits, layout, out=da)
    assert np.array_equal(_host(da), ref)
    da = _dev(a, cuda)
    b2.mul(da, db, field_bits, layout, out=db)
    assert np.array_equal(_host(db), ref)
    sq = _host(b2.mul(da, da.clone(), field_bits, layout))
    b2.mul(da, da, field_bits, layout, out=da)
    assert np.array_equal(_host(da), sq)


@pytest.mark.parametrize("field_bits,n_threads", [(32, 1), (32, 200), (128, 1), (128, 130)])
def test_mul_repeat_vs_oracle(b2, oracle, cuda, field_bits, n_threads):
    """the dependent-chain probe (bitsliced_repeat's shape, profiling/kernels/bitsliced_repeat.cu:6-33): x <- x * y, `iters`
    times, in both widths; y = x's words rotated (by 1 / by 33) XOR 0x9e3779b9 * (index + 1), as include/b2ntt.h states"""
    W, iters = field_bits, 3
    x = _rand(n_threads * W, 9 + field_bits).reshape(n_threads, W)
    rot = 1 if W == 32 else 33
    idx = (np.arange(W) + rot) % W
    y = x[:, idx] ^ (np.uint32(0x9E3779B9) * (np.arange(W, dtype=np.uint32) + np.uint32(1)))[None, :]
    ref = x.copy()
    for _ in range(iters):
        ref = oracle.mul_bitsliced(ref.reshape(-1), np.ascontiguousarray(y).reshape(-1), 5 if W == 32 else 7).reshape(n_threads, W)
    d = _dev(x.reshape(-1), cuda)
    b2.mul_repeat(d, field_bits, iters)
    assert np.array_equal(_host(d).reshape(n_threads, W), ref)
