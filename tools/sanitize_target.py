"""Small workload for compute-sanitizer (tools/sanitize.sh): every kernel family of libb2ntt.so once, at sizes a sanitizer
can replay -- the transforms under all three pass schedules (B2NTT_UPPER = 0, 1, 2), forward and inverse, both fields, a coset
fan-out, the bit-sliced layout, the element-wise multiplies, the layout conversion, the sumcheck prover and the general-basis
transform.  Results are compared with each other (round trips), not with the oracle: parity is the GPU tests' job."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import binius_ntt_b200 as b2

dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(99)


def rnd(words):
    return torch.randint(-2**31, 2**31 - 1, (words,), dtype=torch.int32, device=dev, generator=g)


for sched in ("0", "1", "2"):
    os.environ["B2NTT_UPPER"] = sched
    for field, log_h, log_rate, layout in ((32, 16, 0, 0), (128, 14, 0, 0), (128, 14, 1, 0), (32, 17, 0, 1), (128, 15, 0, 1), (32, 9, 2, 0), (128, 3, 0, 0)):
        plan = b2.Plan(field, log_h, log_rate, layout=layout)
        x = rnd(plan.in_words)
        y = plan.exec(x.clone(), None if log_rate == 0 else torch.empty(plan.out_words, dtype=torch.int32, device=dev))
        back = plan.exec_inverse(y[:plan.in_words].clone())
        torch.cuda.synchronize()
        assert torch.equal(back, x), (sched, field, log_h, log_rate, layout)
        plan.close()
    print("schedule", sched, "ok", flush=True)
del os.environ["B2NTT_UPPER"]
for field, layout, n in ((32, 0, 4096), (32, 1, 4096), (128, 0, 2048), (128, 1, 2048), (8, 0, 1000), (16, 0, 1000)):
    w = max(1, field // 32)
    a, b = rnd(n * w), rnd(n * w)
    ab, ba = b2.mul(a, b, field, layout), b2.mul(b, a, field, layout)
    torch.cuda.synchronize()
    assert torch.equal(ab, ba)
for width in (32, 64, 128):
    x = rnd(width * 33)
    y = x.clone()
    b2.bitslice_(y, width)
    b2.bitslice_(y, width, inverse=True)
    torch.cuda.synchronize()
    assert torch.equal(x, y)
io = rnd(32 * 256)
b2.mul_repeat(io, 32, 3)
sc = b2.Sumcheck(np.random.default_rng(5).integers(0, 2**32, size=2 * (1 << 10) * 4, dtype=np.uint32), 10, 2)
for _ in range(10):
    s, pts = sc.this_round_messages()
    sc.move_to_next_round(np.arange(4, dtype=np.uint32) + 7)
gn = b2.GeneralNTT(12)
x = rnd(gn.words)
y = gn.exec(x.clone())
z = gn.exec(y.clone(), inverse=True)
torch.cuda.synchronize()
assert torch.equal(z, x)
print("sanitize target ok", flush=True)
