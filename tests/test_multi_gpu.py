"""Multi-GPU parity on real devices (skipped below two visible GPUs): the single-process drivers b2ntt_mg_*  and the
one-process-per-GPU path under torchrun with NCCL (tools/multi_gpu_check.py --parity-only: fused peer-store exchange and the
all_to_all baseline), each block compared with a single-GPU transform of the whole input."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gpus():
    import torch

    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def test_mg_four_step_single_process(b2, cuda):
    import torch

    n_gpus = 1 << (_gpus().bit_length() - 1) if _gpus() else 0
    if n_gpus < 2:
        pytest.skip("needs at least two GPUs")
    for log_n, field in ((20, 128), (18, 32)):
        limbs = field // 32
        x = torch.randint(-2**31, 2**31 - 1, ((1 << log_n) * limbs,), dtype=torch.int32, device="cuda:0", generator=torch.Generator(device="cuda:0").manual_seed(log_n))
        ref = b2.Plan(field, log_n, 0).exec(x.clone())
        mg = b2.MultiGPU(b2.api.MG_FOUR_STEP, list(range(n_gpus)), field, log_n)
        ins = [b2.FourStepNTT.m_shard(x, n_gpus, r, limbs).contiguous().to(f"cuda:{r}") for r in range(n_gpus)]
        outs = [torch.empty_like(t) for t in ins]
        for _ in range(2):  # twice: the blocks are reused
            mg.exec(ins, outs)
        w = ref.numel() // n_gpus
        for r in range(n_gpus):
            assert torch.equal(outs[r].to("cuda:0"), ref[r * w:(r + 1) * w]), (log_n, field, r)
        mg.close()


def test_mg_column_shard_single_process(b2, oracle, cuda):
    import torch

    n_gpus = min(_gpus(), 4)
    if n_gpus < 2:
        pytest.skip("needs at least two GPUs")
    cols, log_h, log_rate = 7, 12, 2  # uneven split
    x = np.random.default_rng(3).integers(0, 2**32, size=cols << log_h, dtype=np.uint32)
    mg = b2.MultiGPU(b2.api.MG_COLUMN_SHARD, list(range(n_gpus)), 32, log_h, log_rate, cols)
    ins, outs, spans = [], [], []
    for i in range(n_gpus):
        first, count = mg.columns(i)
        spans.append((first, count))
        ins.append(torch.from_numpy(x[first << log_h:(first + count) << log_h].view(np.int32)).to(f"cuda:{i}") if count else None)
        outs.append(torch.empty(count << (log_h + log_rate), dtype=torch.int32, device=f"cuda:{i}") if count else None)
    assert sum(c for _, c in spans) == cols and spans[0][0] == 0
    mg.exec(ins, outs)
    for (first, count), o in zip(spans, outs):
        for c in range(count):
            ref = oracle.antt_forward(x[(first + c) << log_h:(first + c + 1) << log_h], log_h, log_rate, 1)
            assert np.array_equal(o[c << (log_h + log_rate):(c + 1) << (log_h + log_rate)].cpu().numpy().view(np.uint32), ref)
    mg.close()


def test_four_step_torchrun_nccl(cuda):
    n = 1 << (_gpus().bit_length() - 1) if _gpus() else 0
    if n < 2:
        pytest.skip("needs at least two GPUs")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1", "--master-port", "29533",
                        os.path.join(ROOT, "tools", "multi_gpu_check.py"), "--parity-only"], capture_output=True, text=True, timeout=900)
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert r.returncode == 0 and len(lines) == 2, r.stdout[-2000:] + r.stderr[-2000:]
    assert all('"all_blocks_equal": true' in l for l in lines), lines
