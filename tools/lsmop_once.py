import sys; sys.path.insert(0, '.')
import torch, evox_b200 as E
X = torch.rand((100000, 300), device="cuda")
p = E.problems.numerical.LSMOP1(d=300, m=3)
for _ in range(5): p.evaluate(None, X)
torch.cuda.synchronize()
