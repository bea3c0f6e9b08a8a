"""GPU test of b2ntt_sumcheck_round_static (include/b2ntt.h): the first-round sumcheck messages computed with no allocation,
the round kernel's accumulator in a static __device__ array.  It has no counterpart in the reference; the check is bit-exact
equality with the existing entry point (b2ntt_sumcheck_round_messages after b2ntt_sumcheck_load_device) on the same device
buffer, on two calls in a row -- the second call meets the first call's sums in the shared array and must clear them.
Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ELIMIT, EINVAL = -6, -1


def _existing(L, d_evals, num_vars, comp):
    """b2ntt_sumcheck_create + load_device(transposed = 1) + round_messages: host sum[4], points[(comp + 1) * 4]"""
    h = C.c_void_p()
    assert L.b2ntt_sumcheck_create(C.byref(h), num_vars, comp, d_evals.device.index or 0) == 0
    try:
        assert L.b2ntt_sumcheck_load_device(h, d_evals.data_ptr(), 1) == 0
        s, p = np.zeros(4, np.uint32), np.zeros(4 * (comp + 1), np.uint32)
        assert L.b2ntt_sumcheck_round_messages(h, s.ctypes.data, p.ctypes.data) == 0
        return s, p
    finally:
        L.b2ntt_sumcheck_destroy(h)


@pytest.mark.parametrize("num_vars,comp", [(6, 2), (11, 3), (16, 4), (20, 2)])
def test_static_round_equals_the_existing_entry_point_twice_in_a_row(b2, cuda, num_vars, comp):
    import torch

    L = b2.lib()
    stream = torch.cuda.Stream(device=cuda)
    outs = []
    for call in range(2):  # different data per call: stale sums of call 0 would corrupt call 1
        g = torch.Generator(device=cuda).manual_seed(1000 * num_vars + 10 * comp + call)
        d = torch.randint(-2**31, 2**31 - 1, (4 * comp << num_vars,), dtype=torch.int32, device=cuda, generator=g)
        d_points = torch.full((4 * (comp + 1),), -1, dtype=torch.int32, device=cuda)
        d_sum = torch.full((4,), -1, dtype=torch.int32, device=cuda)
        torch.cuda.synchronize(cuda)
        rc = L.b2ntt_sumcheck_round_static(d.data_ptr(), num_vars, comp, d_points.data_ptr(), d_sum.data_ptr(), stream.cuda_stream)
        assert rc == 0, (rc, L.b2ntt_last_cuda_error())
        stream.synchronize()
        outs.append((d, d_sum.cpu().numpy().view(np.uint32), d_points.cpu().numpy().view(np.uint32)))
    for d, s, p in outs:  # the existing entry point afterwards, so the two static calls really are back to back
        rs, rp = _existing(L, d, num_vars, comp)
        assert np.array_equal(s, rs) and np.array_equal(p, rp), (num_vars, comp)
        assert p.any()


def test_static_round_rejects_sizes_beyond_its_scratch(b2, cuda):
    import torch

    L = b2.lib()
    d = torch.zeros(4 * 2 << 6, dtype=torch.int32, device=cuda)
    o = torch.zeros(32, dtype=torch.int32, device=cuda)
    assert L.b2ntt_sumcheck_round_static(d.data_ptr(), 6, 5, o.data_ptr(), o.data_ptr(), None) == ELIMIT
    assert b"composition_size > 4" in L.b2ntt_last_cuda_error()
    assert L.b2ntt_sumcheck_round_static(d.data_ptr(), 31, 2, o.data_ptr(), o.data_ptr(), None) == ELIMIT
    assert b"scratch" in L.b2ntt_strerror(ELIMIT)
    assert L.b2ntt_sumcheck_round_static(d.data_ptr(), 5, 2, o.data_ptr(), o.data_ptr(), None) == EINVAL
    assert L.b2ntt_sumcheck_round_static(None, 6, 2, o.data_ptr(), o.data_ptr(), None) == EINVAL
    host = np.zeros(32, np.uint32)
    assert L.b2ntt_sumcheck_round_static(d.data_ptr(), 6, 2, host.ctypes.data, o.data_ptr(), None) == EINVAL
