"""Size-independent exactness check of front ranks, in plain torch (no library kernel involved).

`rank` is the non-dominated-sort rank vector of `x` iff for every row j
    rank[j] == 1 + max(rank[i] for i dominating j)      (0 when nothing dominates j),
which follows from the peel of non_dominate.py:74-113: a row leaves in the front after its last
dominator left.  Checking the identity for a sample of rows j against ALL rows i costs
len(sample) * n compares, so it runs at N = 1M where no dense N x N check can.  Used by the `-m gpu`
tests and by bench.py's `verified` field (each rank of a multi-GPU run checks its share of the sample).
"""
import torch


def rank_mismatches_sampled(x, rank, sample_idx, block_elems=1 << 24):
    """Number of sampled rows whose rank violates the identity above (0 = the sample is exact)."""
    sample_idx = sample_idx.to(device=x.device, dtype=torch.long)
    if sample_idx.numel() == 0:
        return 0
    xs = x[sample_idx]                                            # [s, m]
    best = torch.full((sample_idx.numel(),), -1, dtype=torch.int32, device=x.device)
    step = max(1, block_elems // sample_idx.numel())
    for s in range(0, x.shape[0], step):
        blk = x[s:s + step]                                       # [b, m]
        le = (blk[:, None, :] <= xs[None, :, :]).all(-1)
        lt = (blk[:, None, :] < xs[None, :, :]).any(-1)
        dom = le & lt                                             # blk row dominates sample row
        rk = rank[s:s + step, None].expand(-1, sample_idx.numel())
        rk = torch.where(dom, rk, torch.full_like(rk, -1))
        best = torch.maximum(best, rk.max(dim=0).values)
    return int((rank[sample_idx] != best + 1).sum())


def fronts_contiguous(rank):
    """True iff every front index 0 .. max(rank) is used (no empty front below the last)."""
    if rank.numel() == 0:
        return True
    nf = int(rank.max()) + 1
    return bool((torch.bincount(rank.long(), minlength=nf) > 0).all()) and int(rank.min()) == 0
