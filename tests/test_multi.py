"""Multi-GPU partitioning.  CPU part: the host-side logic of binius-ntt_b200/multi.py under
`gloo` with world_size 2 and 4 -- column sharding and the four-step exchange -- with the two local
phases of FourStepNTT replaced by the oracle (test infrastructure; the product phases are libb2ntt.so
calls and are covered by the gpu tests below, which walk all ranks of a split on ONE GPU)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_shard_columns_partitions_exactly(b2):
    for n in (0, 1, 7, 256, 257):
        for world in (1, 2, 3, 4, 8):
            got = [b2.shard_columns(n, world, r) for r in range(world)]
            assert sum(c for _, c in got) == n
            pos = 0
            for first, count in got:
                assert first == pos
                pos += count
            assert max(c for _, c in got) - min(c for _, c in got) <= 1
    with pytest.raises(ValueError):
        b2.shard_columns(4, 2, 2)


def _oracle_top(O, buf, log_n, g, limbs):
    """numpy restatement of phase 1: top g stages across the 2^g local segments (additive_ntt.cuh:127-159)."""
    G = 1 << g
    C = O.antt_constants(log_n, 0)
    pitch = C.shape[1]
    flat = np.ascontiguousarray(C).reshape(-1)
    seg = buf.reshape(G, -1).copy()
    for t in range(g):
        stage = log_n - 1 - t
        bit = g - 1 - t
        for h in range(G):
            if h >> bit & 1:
                continue
            w = O.lib.orc_antt_twiddle(flat, pitch, log_n, 0, 0, stage, h >> (bit + 1))
            u, v = seg[h], seg[h | 1 << bit]
            if w:
                u ^= O.mul32(np.full(v.size, w, dtype=np.uint32), v)
            v ^= u
    return seg.reshape(-1)


def _four_step_worker(rank, world, port, log_n, limbs, ret):
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    sys.path.insert(0, os.path.join(root, "oracle"))
    import binius_ntt_b200 as b2
    import oracle as O

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)

    class OracleFourStep(b2.FourStepNTT):  # local phases on the CPU oracle: exercises split + exchange logic only
        def _make_plans(self):
            self.full_plan = self.local_plan = None

        def top(self, buf, stream=None):
            buf.copy_(torch.from_numpy(_oracle_top(O, buf.numpy().view(np.uint32), self.log_n, self.g, self.limbs).view(np.int32)))
            return buf

        def local(self, buf, stream=None):
            full = O.antt_forward(buf.numpy().view(np.uint32), self.log_local, self.g, self.limbs)
            n = self.local_words
            buf.copy_(torch.from_numpy(full[self.rank * n:(self.rank + 1) * n].view(np.int32).copy()))
            return buf

    x = np.random.default_rng(99).integers(0, 2**32, size=(1 << log_n) * limbs, dtype=np.uint32)
    fs = OracleFourStep(log_n, field_bits=32 * limbs)
    mine = torch.from_numpy(np.ascontiguousarray(b2.FourStepNTT.m_shard(x, world, rank, limbs)).view(np.int32).copy())
    out = fs.forward(mine)
    want = O.antt_forward(x, log_n, 0, limbs)
    n = fs.local_words
    ret[rank] = bool(np.array_equal(out.numpy().view(np.uint32), want[rank * n:(rank + 1) * n]))
    # column sharding: every rank owns a disjoint slab, no collective involved
    first, count = b2.shard_columns(10, world, rank)
    allc = [None] * world
    dist.all_gather_object(allc, (first, count))
    ret[world + rank] = sorted(allc) == [b2.shard_columns(10, world, r) for r in range(world)]
    dist.destroy_process_group()


@pytest.mark.parametrize("world,log_n,limbs", [(2, 10, 4), (4, 11, 1), (4, 10, 4)])
def test_four_step_split_gloo(world, log_n, limbs):
    port = _free_port()
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_four_step_worker, args=(world, port, log_n, limbs, ret), nprocs=world, join=True)
        assert all(ret.get(r) for r in range(2 * world)), dict(ret)


# ---- GPU: the product phases through the C ABI, all ranks of a split walked on one device --------
@pytest.mark.gpu
@pytest.mark.parametrize("field_bits,log_n,g", [(128, 16, 1), (128, 16, 2), (128, 18, 3), (32, 18, 3), (32, 20, 2)])
def test_four_step_phases_on_one_gpu(b2, oracle, cuda, field_bits, log_n, g):
    limbs = field_bits // 32
    G = 1 << g
    x = np.random.default_rng(5 + log_n + g).integers(0, 2**32, size=(1 << log_n) * limbs, dtype=np.uint32)
    want = oracle.antt_forward(x, log_n, 0, limbs)
    full = b2.Plan(field_bits, log_n, 0)
    local = b2.Plan(field_bits, log_n - g, g)
    shards = []
    for m in range(G):
        buf = torch.from_numpy(np.ascontiguousarray(b2.FourStepNTT.m_shard(x, G, m, limbs)).view(np.int32).copy()).to(cuda)
        full.exec_top(buf, g)
        shards.append(buf.view(G, -1))
    for h in range(G):
        block = torch.stack([shards[m][h] for m in range(G)]).reshape(-1).contiguous()  # what the all-to-all delivers to rank h
        got = local.exec_coset(block, block, h).cpu().numpy().view(np.uint32)
        n = block.numel()
        assert np.array_equal(got, want[h * n:(h + 1) * n]), (h,)
    full.close(), local.close()


@pytest.mark.gpu
def test_exec_coset_matches_rows_of_exec(b2, oracle, cuda):
    for field_bits, log_h, r in [(32, 12, 2), (128, 10, 3), (32, 4, 2), (128, 3, 1)]:
        limbs = field_bits // 32
        x = np.random.default_rng(log_h).integers(0, 2**32, size=(1 << log_h) * limbs, dtype=np.uint32)
        want = oracle.antt_forward(x, log_h, r, limbs)
        p = b2.Plan(field_bits, log_h, r)
        for c in range(1 << r):
            d = torch.from_numpy(x.view(np.int32).copy()).to(cuda)
            got = p.exec_coset(d, None, c).cpu().numpy().view(np.uint32)
            assert np.array_equal(got, want[c * x.size:(c + 1) * x.size]), (field_bits, log_h, r, c)
        p.close()


@pytest.mark.gpu
def test_column_sharded_rs_single_rank(b2, oracle, cuda):
    rs = b2.ColumnShardedRS(6, 12, 2, field_bits=32)
    assert (rs.first, rs.count) == (0, 6)
    x = np.random.default_rng(3).integers(0, 2**32, size=6 << 12, dtype=np.uint32)
    out = rs.encode(torch.from_numpy(x.view(np.int32).copy()).to(cuda)).cpu().numpy().view(np.uint32)
    for c in range(6):
        assert np.array_equal(out[c << 14:(c + 1) << 14], oracle.antt_forward(x[c << 12:(c + 1) << 12], 12, 2, 1))
    rs.close()


@pytest.mark.gpu
def test_rs_encode_from_evaluations_is_systematic(b2, oracle, cuda):
    rs = b2.ColumnShardedRS(3, 13, 2, field_bits=32)
    y = np.random.default_rng(8).integers(0, 2**32, size=3 << 13, dtype=np.uint32)
    out = rs.encode_from_evaluations(torch.from_numpy(y.view(np.int32).copy()).to(cuda)).cpu().numpy().view(np.uint32)
    for c in range(3):
        col = y[c << 13:(c + 1) << 13]
        want = oracle.antt_forward(oracle.antt_inverse(col, 13, 2, 0, 1), 13, 2, 1)
        assert np.array_equal(out[c << 15:(c + 1) << 15], want)
        assert np.array_equal(out[c << 15:(c << 15) + (1 << 13)], col)  # coset 0 reproduces the message
    rs.close()
