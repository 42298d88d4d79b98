import sys, numpy as np, torch
sys.path.insert(0, '.')
import evox_b200 as E
from oracle import oracle_c as C
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1025
x = np.random.default_rng(n * 31 + 3).random((n, 3), dtype=np.float32)
r = E.operators.non_dominated_sort(torch.from_numpy(x).cuda()).cpu().numpy()
torch.cuda.synchronize()
ref = C.non_dominated_sort(x)
print("mismatch", int((r != ref).sum()), "fronts", ref.max() + 1, np.bincount(ref)[:6], np.bincount(r)[:6])
