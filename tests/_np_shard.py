"""NumPy stand-in for the device side of one rank of the sharded non-dominated sort (the contract of
the evx_nds_shard_* entry points in include/evoxmo.h), so that the host-side driver in
evox_b200/distributed.py - panel ownership, mask exchange layout, batching, termination, rank
assembly - can be exercised with world_size 2 on CPU over gloo.  Test infrastructure only."""
import numpy as np
import torch

from oracle import oracle_np as O

TW, PW = 1024, 32


class NumpyShard:
    def __init__(self, n, m, world, rank):
        self.n, self.m, self.world, self.rank = n, m, world, rank
        self.npad = -(-max(n, 1) // TW) * TW
        self.np_ = self.npad // TW
        npairs = (self.np_ + 1) // 2                  # panels are owned in pairs: pair P -> rank P % world
        self.npl_max = 2 * -(-npairs // world)
        self.mask_words = self.npl_max * PW
        self.device = torch.device("cpu")

    def _owned(self, u):
        return ((u // TW) // 2) % self.world == self.rank

    def _local(self, u):                       # label -> (local word, bit)
        J = u // TW
        l = 2 * ((J // 2) // self.world) + (J % 2)
        c = u - J * TW
        return l * PW + c // 32, c % 32

    def build(self, x):
        x = x.numpy()
        nan = np.isnan(x).any(1)
        uniq, inv = np.unique(x[~nan], axis=0, return_inverse=True)     # lexicographic labels
        self.row2u = np.full(self.n, -1, np.int64)
        self.row2u[~nan] = inv.reshape(-1)
        self.U = uniq.shape[0]
        self.dom = O.dominate_relation(uniq, uniq)
        labels = np.arange(self.U)
        self.mine = labels[self._owned(labels)]
        self.count = self.dom[:, self.mine].sum(0).astype(np.int64)
        self.rank_u = np.zeros(self.U, np.int32)

    def detect(self, k, local_mask):
        m = np.zeros(self.mask_words, np.uint32)
        z = self.count == 0
        for u in self.mine[z]:
            w, b = self._local(int(u))
            m[w] |= np.uint32(1) << np.uint32(b)
        self.rank_u[self.mine[z]] = k
        self.count[z] = -1
        local_mask.copy_(torch.from_numpy(m.view(np.int32)))

    def apply(self, k, all_masks, any_slot):
        masks = all_masks.numpy().view(np.uint32).reshape(self.world, self.mask_words)
        front = []
        for r in range(self.world):
            for wl in np.nonzero(masks[r])[0]:
                l, w = divmod(int(wl), PW)
                J = 2 * ((l // 2) * self.world + r) + (l % 2)
                for b in range(32):
                    if (int(masks[r, wl]) >> b) & 1:
                        front.append(J * TW + w * 32 + b)
        any_slot[0] = 1 if front else 0
        if front:
            self.count -= self.dom[np.array(front)][:, self.mine].sum(0)

    def ranks(self, out):
        r = np.zeros(self.n, np.int32)
        ok = self.row2u >= 0
        u = self.row2u[ok]
        r[ok] = np.where(self._owned(u), self.rank_u[u], 0)
        out.copy_(torch.from_numpy(r))
