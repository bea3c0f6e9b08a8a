"""GPU parity tests of the additive NTT: CUDA path (through the C ABI) vs the CPU oracle on the same
seeded inputs, vs the reference's golden MD5 table (ntt/tests/test_ntt.cu:52-124), and through
size-independent properties at full size."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _dev(x, cuda):
    import torch

    return torch.from_numpy(x.view(np.int32)).to(cuda)


def _host(t):
    return t.cpu().numpy().view(np.uint32)


def _rand(n, seed):
    return np.random.default_rng(seed).integers(0, 2**32, size=n, dtype=np.uint32)


# ---- golden digests: the reference's own test, log_rate 0 and 2 (test_ntt.cu:219-229) ---------------
@pytest.mark.parametrize("log_rate,log_hs", [(0, range(1, 25)), (2, range(1, 23))])
def test_golden_md5_gf32(b2, oracle, cuda, log_rate, log_hs, schedule):
    for log_h in log_hs:
        x = oracle.reference_input(log_h, log_rate)
        plan = b2.Plan(32, log_h, log_rate)
        out = plan.exec_host(x)
        assert oracle.md5_hex(out) == oracle.golden_md5(log_rate, log_h), (log_rate, log_h)
        plan.close()


@pytest.mark.parametrize("log_rate,log_h", [(0, 25), (0, 26), (0, 27), (0, 28), (2, 24), (2, 25), (2, 26)])
def test_golden_md5_gf32_large(b2, oracle, cuda, log_rate, log_h):
    x = oracle.reference_input(log_h, log_rate)
    plan = b2.Plan(32, log_h, log_rate)
    out = plan.exec_host(x)
    assert oracle.md5_hex(out) == oracle.golden_md5(log_rate, log_h)
    plan.close()


@pytest.fixture(params=["auto", "tile", "bundle", "bsl"])
def schedule(request, monkeypatch):
    """the plan picks its pass schedule by size (b2ntt_plan_create); B2NTT_UPPER pins it, so that the upper/tile kernels
    and the bundle-pass kernel each meet every small edge case on the GPU"""
    if request.param != "auto":
        monkeypatch.setenv("B2NTT_UPPER", {"tile": "2", "bundle": "0", "bsl": "5"}[request.param])
    else:
        monkeypatch.delenv("B2NTT_UPPER", raising=False)
    return request.param


# ---- oracle parity on random inputs, every log_rate, both fields, device-resident path ------------
@pytest.mark.parametrize("field_bits", [32, 128])
@pytest.mark.parametrize("log_rate", [0, 1, 2, 3, 4])
def test_oracle_parity(b2, oracle, cuda, field_bits, log_rate, schedule):
    limbs = field_bits // 32
    for log_h in [1, 2, 3, 4, 5, 6, 7, 9, 12, 15, 16, 18]:
        x = _rand((1 << log_h) * limbs, 1000 * log_h + log_rate)
        plan = b2.Plan(field_bits, log_h, log_rate)
        out = _host(plan.exec(_dev(x, cuda)))
        ref = oracle.antt_forward(x, log_h, log_rate, limbs)
        assert np.array_equal(out, ref), (field_bits, log_h, log_rate)
        plan.close()


def test_gf128_limb0_is_golden(b2, oracle, cuda, schedule):
    """limb 0 of a GF(2^128) transform fed with the reference stream reproduces the
    GF(2^32) golden digest."""
    for log_h, log_rate in [(20, 0), (18, 2), (22, 0)]:
        x = np.zeros((1 << log_h, 4), dtype=np.uint32)
        x[:, 0] = oracle.reference_input(log_h, log_rate)
        x[:, 1:] = _rand((1 << log_h) * 3, 5).reshape(-1, 3)
        plan = b2.Plan(128, log_h, log_rate)
        out = plan.exec_host(x.reshape(-1)).reshape(-1, 4)
        assert oracle.md5_hex(np.ascontiguousarray(out[:, 0])) == oracle.golden_md5(log_rate, log_h)
        plan.close()


@pytest.mark.parametrize("field_bits", [32, 128])
def test_batch_columns(b2, oracle, cuda, field_bits, schedule):
    limbs = field_bits // 32
    log_h, log_rate, batch = 11, 2, 5
    x = _rand(batch * (1 << log_h) * limbs, 77)
    plan = b2.Plan(field_bits, log_h, log_rate, batch=batch)
    out = _host(plan.exec(_dev(x, cuda))).reshape(batch, -1)
    for c in range(batch):
        ref = oracle.antt_forward(x.reshape(batch, -1)[c], log_h, log_rate, limbs)
        assert np.array_equal(out[c], ref)


@pytest.mark.parametrize("field_bits", [32, 128])
def test_inverse_roundtrip_and_oracle(b2, oracle, cuda, field_bits, schedule):
    limbs = field_bits // 32
    for log_h, log_rate in [(3, 0), (10, 0), (14, 2), (20, 1)]:
        x = _rand((1 << log_h) * limbs, log_h)
        plan = b2.Plan(field_bits, log_h, log_rate)
        y = plan.exec(_dev(x, cuda))
        n = (1 << log_h) * limbs
        for coset in range(1 << log_rate):
            back = _host(plan.exec_inverse(y[coset * n:(coset + 1) * n].clone(), coset=coset))
            assert np.array_equal(back, x), (log_h, log_rate, coset)
        if log_h <= 14:
            yi = _host(y)[:n]
            assert np.array_equal(oracle.antt_inverse(yi, log_h, log_rate, 0, limbs), x)


@pytest.mark.parametrize("field_bits", [32, 128])
def test_bitsliced_layout(b2, oracle, cuda, field_bits, schedule):
    limbs = field_bits // 32
    for log_h, log_rate in [(5, 0), (6, 1), (8, 0), (13, 2), (17, 0)]:
        x = _rand((1 << log_h) * limbs, 31 * log_h)
        xs = oracle.bitslice(x, field_bits)
        plan = b2.Plan(field_bits, log_h, log_rate, layout=b2.api.LAYOUT_BITSLICED)
        out = _host(plan.exec(_dev(xs, cuda)))
        ref = oracle.antt_forward(x, log_h, log_rate, limbs)
        assert np.array_equal(oracle.unbitslice(out, field_bits), ref), (field_bits, log_h, log_rate)


def test_full_size_properties_gf128(b2, oracle, cuda):
    """log_n = 26 GF(2^128): linearity + limb-0 golden digest + inverse round trip."""
    import torch

    log_h = 26
    n = 1 << log_h
    g = torch.Generator(device=cuda).manual_seed(5)
    a = torch.randint(-2**31, 2**31 - 1, (n * 4,), dtype=torch.int32, device=cuda, generator=g)
    b = torch.randint(-2**31, 2**31 - 1, (n * 4,), dtype=torch.int32, device=cuda, generator=g)
    a.view(-1, 4)[:, 0] = _dev(oracle.reference_input(log_h, 0), cuda)
    plan = b2.Plan(128, log_h, 0)
    fa = plan.exec(a.clone())
    assert oracle.md5_hex(_host(fa.view(-1, 4)[:, 0].contiguous())) == oracle.golden_md5(0, log_h)
    fb = plan.exec(b.clone())
    fab = plan.exec(a ^ b)
    assert torch.equal(fab, fa ^ fb)
    assert torch.equal(plan.exec_inverse(fa), a)


def test_headline_size_properties_gf128(b2, oracle, cuda):
    """BASELINE metric size, log_n = 28 GF(2^128) (4 GiB): limb-0 golden digest (ntt/tests/test_ntt.cu:52-124, F8),
    linearity, inverse round trip."""
    import torch

    log_h = 28
    n = 1 << log_h
    g = torch.Generator(device=cuda).manual_seed(28)
    a = torch.randint(-2**31, 2**31 - 1, (n * 4,), dtype=torch.int32, device=cuda, generator=g)
    b = torch.randint(-2**31, 2**31 - 1, (n * 4,), dtype=torch.int32, device=cuda, generator=g)
    a.view(-1, 4)[:, 0] = _dev(oracle.reference_input(log_h, 0), cuda)
    plan = b2.Plan(128, log_h, 0)
    fa = plan.exec(a.clone())
    assert oracle.md5_hex(_host(fa.view(-1, 4)[:, 0].contiguous())) == oracle.golden_md5(0, log_h)
    fab = plan.exec(a ^ b)
    fab ^= fa
    assert torch.equal(fab, plan.exec(b.clone()))
    del fab, b
    assert torch.equal(plan.exec_inverse(fa), a)
    plan.close()


def test_config2_compact_and_bitsliced_agree(b2, oracle, cuda, schedule):
    """config 2: log_n = 24 GF(2^128), the bit-sliced image transforms to the bit-sliced image of the compact result"""
    import torch

    log_h = 24
    x = torch.randint(-2**31, 2**31 - 1, ((1 << log_h) * 4,), dtype=torch.int32, device=cuda, generator=torch.Generator(device=cuda).manual_seed(24))
    yc = b2.Plan(128, log_h, 0).exec(x.clone())
    xs = x.clone()
    b2.bitslice_(xs, 128)
    ys = b2.Plan(128, log_h, 0, layout=b2.api.LAYOUT_BITSLICED).exec(xs)
    b2.bitslice_(ys, 128, inverse=True)
    assert torch.equal(ys, yc)


def test_config3_rs_columns_golden(b2, oracle, cuda):
    """config 3 shape (log_h = 20, log_rate = 2, GF(2^32)) on a batch of columns: every column carrying the reference
    stream reproduces the reference digest 53e8a086... (ntt/tests/test_ntt.cu:112)"""
    cols = 8
    x = oracle.reference_input(20, 2)
    plan = b2.Plan(32, 20, 2, batch=cols)
    out = _host(plan.exec(_dev(np.tile(x, cols), cuda)))
    for c in range(cols):
        assert oracle.md5_hex(out[c << 22:(c + 1) << 22]) == oracle.golden_md5(2, 20) == "53e8a086be97f61f43930357442272cb"
    plan.close()


def test_error_paths(b2, cuda):
    with pytest.raises(b2.B2NTTError):
        b2.Plan(32, 0, 0)
    with pytest.raises(b2.B2NTTError):
        b2.Plan(32, 30, 4)
    with pytest.raises(b2.B2NTTError):
        b2.Plan(64, 10, 0)
    ntt = b2.AdditiveNTT(b2.AdditiveNTTConf(10, 0))
    bad = b2.NTTData(size=1 << 9)
    out = b2.NTTData()
    assert ntt.apply(bad, out) is False
    good = b2.NTTData(size=1 << 10)
    good.order = b2.DataOrder.BIT_REVERSED
    assert ntt.apply(good, out) is False
    assert b2.check_gpu_capabilities(0)


@pytest.mark.gpu
def test_exec_host_batch_matches_exec_host(b2, oracle, cuda):
    """the pipelined host path gives the same bytes as AdditiveNTT::apply, in order, with buffer sets reused"""
    for field_bits, log_h, r in [(32, 14, 2), (128, 13, 0)]:
        limbs = field_bits // 32
        p = b2.Plan(field_bits, log_h, r)
        xs = [np.random.default_rng(40 + i).integers(0, 2**32, size=(1 << log_h) * limbs, dtype=np.uint32) for i in range(7)]
        outs = [np.zeros(p.out_words, dtype=np.uint32) for _ in xs]
        p.exec_host_batch(xs, outs)
        for x, o in zip(xs, outs):
            assert np.array_equal(o, oracle.antt_forward(x, log_h, r, limbs))
        p.close()


def test_config2_both_layouts_vs_oracle(b2, oracle, cuda):
    """config 2 at its stated size (log_n = 24, GF(2^128)): the compact AND the bit-sliced transform against the CPU
    oracle word for word, plus the reference digest additive_ntt_hashes[0][24] (ntt/tests/test_ntt.cu:79) on limb 0 of
    both layouts."""
    import torch

    log_h = 24
    x = np.empty((1 << log_h, 4), dtype=np.uint32)
    x[:, 0] = oracle.reference_input(log_h, 0)
    x[:, 1:] = _rand((1 << log_h) * 3, 24).reshape(-1, 3)
    ref = oracle.antt_forward(x.reshape(-1), log_h, 0, 4)
    assert oracle.md5_hex(np.ascontiguousarray(ref.reshape(-1, 4)[:, 0])) == oracle.golden_md5(0, log_h)
    dx = _dev(x.reshape(-1), cuda)
    yc = b2.Plan(128, log_h, 0).exec(dx.clone())
    assert np.array_equal(_host(yc), ref)
    xs = dx.clone()
    b2.bitslice_(xs, 128)
    ys = b2.Plan(128, log_h, 0, layout=b2.api.LAYOUT_BITSLICED).exec(xs)
    b2.bitslice_(ys, 128, inverse=True)
    assert torch.equal(ys, yc)
    assert oracle.md5_hex(_host(ys.view(-1, 4)[:, 0].contiguous())) == oracle.golden_md5(0, log_h)


@pytest.mark.parametrize("log_h", [25, 26])
def test_tile_schedule_bitsliced_limb0_digest(b2, oracle, cuda, log_h):
    """the upper + tile schedule on bit-sliced images above log_h = 24: limb 0 carries the reference stream"""
    import torch

    n = 1 << log_h
    x = torch.randint(-2**31, 2**31 - 1, (n * 4,), dtype=torch.int32, device=cuda, generator=torch.Generator(device=cuda).manual_seed(log_h))
    x.view(-1, 4)[:, 0] = _dev(oracle.reference_input(log_h, 0), cuda)
    b2.bitslice_(x, 128)
    y = b2.Plan(128, log_h, 0, layout=b2.api.LAYOUT_BITSLICED).exec(x)
    b2.bitslice_(y, 128, inverse=True)
    assert oracle.md5_hex(_host(y.view(-1, 4)[:, 0].contiguous())) == oracle.golden_md5(0, log_h)


@pytest.mark.parametrize("field_bits,log_h", [(128, 23), (32, 26), (128, 25)])
def test_pipelined_upper_passes_equal_the_default_kernels(b2, cuda, monkeypatch, field_bits, log_h):
    """B2NTT_UPPER_PIPE=1 (b2_upper_pipe.cu: producer warp + mbarrier ring, table-free butterflies of every class) must give the
    same words as upper_kernel, forward and inverse, on the upper + tile schedule"""
    import torch

    monkeypatch.setenv("B2NTT_UPPER", "2")
    limbs = field_bits // 32
    x = torch.randint(-2**31, 2**31 - 1, ((1 << log_h) * limbs,), dtype=torch.int32, device=cuda,
                      generator=torch.Generator(device=cuda).manual_seed(field_bits + log_h))
    plan = b2.Plan(field_bits, log_h, 0)
    monkeypatch.delenv("B2NTT_UPPER_PIPE", raising=False)
    y0 = plan.exec(x.clone())
    monkeypatch.setenv("B2NTT_UPPER_PIPE", "1")
    y1 = plan.exec(x.clone())
    assert torch.equal(y0, y1)
    assert torch.equal(plan.exec_inverse(y1), x)
    plan.close()


def test_wrong_device_pointers_are_rejected(b2, cuda):
    """a plan is bound to one device: host pointers and tensors of another GPU are refused, and no ABI call changes the
    caller's current device"""
    import ctypes as C

    import torch

    plan = b2.Plan(32, 12, 0)
    host = np.zeros(1 << 12, dtype=np.uint32)
    rc = b2.lib().b2ntt_exec(plan._h, host.ctypes.data, host.ctypes.data, C.c_void_p(0))
    assert rc == -1  # B2NTT_EINVAL: not device memory
    if torch.cuda.device_count() >= 2:
        other = torch.zeros(1 << 12, dtype=torch.int32, device="cuda:1")
        with pytest.raises(ValueError):
            plan.exec(other)
        assert b2.lib().b2ntt_exec(plan._h, other.data_ptr(), other.data_ptr(), C.c_void_p(0)) == -1
        # a plan on device 1 used while device 0 is current: right answer, current device untouched
        p1 = b2.Plan(32, 12, 0, device=1)
        x = torch.randint(-2**31, 2**31 - 1, (1 << 12,), dtype=torch.int32, device="cuda:1")
        torch.cuda.set_device(0)
        with torch.cuda.device(1):
            s1 = torch.cuda.current_stream().cuda_stream
        y1 = p1.exec(x.clone(), stream=s1)
        torch.cuda.synchronize(1)
        assert torch.cuda.current_device() == 0
        assert torch.equal(y1.cpu(), plan.exec(x.to("cuda:0")).cpu())
        p1.close()
    plan.close()
