"""Two rounds of the sumcheck prover at (num_vars, C) = argv or (26, 2): the target of the ncu captures of
sc_round_kernel / sc_fold_kernel (profiles/README.md)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import binius_ntt_b200 as b2

nv = int(sys.argv[1]) if len(sys.argv) > 1 else 26
comp = int(sys.argv[2]) if len(sys.argv) > 2 else 2
d = torch.randint(-2**31, 2**31 - 1, (4 * comp << nv,), dtype=torch.int32, device="cuda")
sc = b2.Sumcheck(d, nv, comp, data_is_transposed=True)
for r in range(2):
    s, p = sc.this_round_messages()
    sc.move_to_next_round(np.array([1 + r, 2, 3, 4], dtype=np.uint32))
torch.cuda.synchronize()
print("ok", sc.launches)
