"""Sharded non-dominated sort: one process per GPU, `torch.distributed` for the plumbing.

The reference replicates selection on every device and only shards evaluation
(src/evox/workflows/std_workflow.py:160,187-200); its N x N byte matrix cannot exist at N = 1M.
Here the packed bit matrix is sharded by column panels (panel J -> rank J mod world).  A panel's
dominator counts are private to its owner, so the only exchange per front is the all-gather of front
membership (N/8 bytes in total); ranks are assembled with one final all-reduce.  Population
evaluation is sharded by rows with an all-gather of F, exactly the reference's own collective
(src/evox/core/distributed.py:49-56).
"""
import torch
import torch.distributed as dist

from ._dev import as_f32_cuda, stream_ptr
from ._lib import EvoxMoError, check, lib


def recommended_world(n, available=8):
    """How many GPUs to give one non-dominated sort of n rows (measured on 8 x B200, profiles/r2_scaling.md):
    the bit matrix shrinks as 1/G but every front costs one exchange across all GPUs, so small populations
    with many light fronts gain next to nothing (N = 200k: 3.24 ms on one, 3.09 on two, 3.07 on four, 3.28 on
    eight), while N = 1M gains up to eight (28.7 -> 9.1 ms)."""
    if n < 300_000:
        want = 1
    elif n < 500_000:
        want = 2
    else:
        want = available
    w = 1
    while w * 2 <= min(want, available):
        w *= 2
    return w


class _CudaShard:
    """Device side of one rank: thin wrapper over the evx_nds_shard_* entry points."""

    def __init__(self, n, m, world, rank):
        L = lib()
        nbytes = L.evx_nds_shard_workspace_bytes(n, m, world, rank)
        if nbytes == 0:
            raise EvoxMoError(f"sharded nds: unsupported shape n={n} m={m} world={world}")
        self.ws = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
        self.plan = L.evx_nds_shard_create(n, m, world, rank, self.ws.data_ptr(), nbytes)
        if not self.plan:
            raise EvoxMoError(L.evx_last_error().decode())
        self.mask_words = int(L.evx_nds_shard_mask_words(self.plan))
        self.device = torch.device("cuda", torch.cuda.current_device())
        self.n = n

    def __del__(self):
        try:
            if getattr(self, "plan", None):
                lib().evx_nds_shard_destroy(self.plan)
                self.plan = None
        except Exception:
            pass

    def build(self, x):
        check(lib().evx_nds_shard_build(self.plan, x.data_ptr(), stream_ptr()))

    def detect(self, k, local_mask):
        check(lib().evx_nds_shard_detect(self.plan, k, local_mask.data_ptr(), stream_ptr()))

    def apply(self, k, all_masks, any_slot):
        check(lib().evx_nds_shard_apply(self.plan, k, all_masks.data_ptr(), any_slot.data_ptr(), stream_ptr()))

    def ranks(self, out):
        check(lib().evx_nds_shard_ranks(self.plan, out.data_ptr(), stream_ptr()))


class ShardedNonDominatedSort:
    """rank = ShardedNonDominatedSort(n, m)(x) with x [n, m] replicated on every rank."""

    MAX_BATCH = 64

    def __init__(self, n, m, group=None, shard_factory=None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.n, self.m = n, m
        self.shard = (shard_factory or _CudaShard)(n, m, self.world, self.rank)
        dev = self.shard.device
        mw = self.shard.mask_words
        self.local_mask = torch.zeros(mw, dtype=torch.int32, device=dev)
        self.all_masks = torch.zeros(mw * self.world, dtype=torch.int32, device=dev)
        self.any = torch.zeros(self.MAX_BATCH, dtype=torch.int32, device=dev)
        self.out = torch.zeros(n, dtype=torch.int32, device=dev)
        self.num_fronts = 0
        self.collectives = 0

    def _exchange(self):
        if self.world == 1:
            self.all_masks.copy_(self.local_mask)
        else:
            dist.all_gather_into_tensor(self.all_masks, self.local_mask, group=self.group)
            self.collectives += 1

    def __call__(self, x):
        if self.n == 0:
            return self.out
        self.shard.build(x)
        k, batch = 0, 8
        while True:
            for b in range(batch):
                self.shard.detect(k + b, self.local_mask)
                self._exchange()
                self.shard.apply(k + b, self.all_masks, self.any[b:b + 1])
            flags = self.any[:batch].cpu()            # one readback per batch: the trip count is data dependent
            empty = (flags == 0).nonzero()
            if empty.numel():
                self.num_fronts = k + int(empty[0])
                break
            k += batch
            batch = min(batch * 2, self.MAX_BATCH)
        self.shard.ranks(self.out)
        if self.world > 1:
            dist.all_reduce(self.out, op=dist.ReduceOp.SUM, group=self.group)
            self.collectives += 1
        return self.out


class FusedShardedNonDominatedSort:
    """Same result as ShardedNonDominatedSort, but the whole peel is ONE persistent kernel per rank:
    front members travel between GPUs as peer stores over NVLink inside the kernel, and the per-front
    barrier spans all ranks' CTAs - no NCCL call and no kernel launch per front.  Needs peer-mapped
    memory, obtained from torch symmetric memory (plumbing only).  The peer state is cleared once, here;
    calls need no host-level hand-shake between them and never block the host."""

    def __init__(self, n, m, group=None):
        import ctypes

        import torch.distributed._symmetric_memory as symm
        if not dist.is_initialized():
            raise EvoxMoError("FusedShardedNonDominatedSort needs an initialised process group")
        self.group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        self.n, self.m = n, m
        self.shard = _CudaShard(n, m, self.world, self.rank)
        L = lib()
        nbytes = int(L.evx_nds_shard_peer_bytes(self.shard.plan))
        self.peer = symm.empty(nbytes, dtype=torch.uint8, device=self.shard.device)
        self.handle = symm.rendezvous(self.peer, self.group.group_name)
        ptrs = list(self.handle.buffer_ptrs)
        self._ptrs = (ctypes.c_void_p * self.world)(*ptrs)
        self.out = torch.zeros(n, dtype=torch.int32, device=self.shard.device)
        self._reset()

    def _reset(self):
        """Clear this rank's peer state and wait until every rank has done so (once per plan, and again
        only after a time-out)."""
        check(lib().evx_nds_shard_peer_init(self.shard.plan, self.peer.data_ptr(), stream_ptr()))
        torch.cuda.current_stream().synchronize()
        dist.barrier(group=self.group)

    @property
    def num_fronts(self):
        """Fronts found by the last call (waits for that call's kernels)."""
        return int(lib().evx_nds_last_num_fronts())

    def __call__(self, x):
        if self.n == 0:
            return self.out
        L = lib()
        self.shard.build(x)
        check(L.evx_nds_shard_peel_fused(self.shard.plan, self._ptrs, None, stream_ptr()))
        self.shard.ranks(self.out)
        dist.all_reduce(self.out, op=dist.ReduceOp.SUM, group=self.group)
        return self.out


def sharded_evaluate(problem, state, X, group=None):
    """Evaluate this rank's slice of the (replicated) candidate matrix and all-gather the fitness, as
    std_workflow.py:187-200 (slice by process index) + core/distributed.py:49-56 (all_gather) do.
    Rows are dealt in equal slices; a population that does not divide pads the last slice."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return problem.evaluate(state, X)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    n = X.shape[0]
    per = -(-n // world)
    lo = min(rank * per, n)
    Xl = X[lo:min(lo + per, n)]
    if Xl.shape[0] < per:                                   # ragged tail: evaluate copies of a valid row, cut off below
        Xl = torch.cat([Xl, X[:1].expand(per - Xl.shape[0], -1)], dim=0)
    F_local, state = problem.evaluate(state, Xl.contiguous())
    F = torch.empty((per * world, F_local.shape[1]), dtype=F_local.dtype, device=F_local.device)
    dist.all_gather_into_tensor(F, F_local.contiguous(), group=group)
    return F[:n], state


def non_dominated_sort_sharded(x, group=None):
    x = as_f32_cuda(x, "x")
    return ShardedNonDominatedSort(x.shape[0], x.shape[1], group)(x)


class ShardedNSGA2:
    """NSGA-II generation over several GPUs, one process each (BASELINE.json configs[2]).  The state is
    replicated, as in the reference's multi-device workflow (std_workflow.py:160,187-200): every rank
    runs the same variation kernel on the same (seed, generation), evaluates ITS slice of the
    candidates, all-gathers the fitness; then the merged population is ranked by the sharded
    non-dominated sort (bit matrix split by column panels, fused front exchange) and every rank applies
    the same survivor selection to the ranks.  Results are identical to the single-GPU NSGA2."""

    def __init__(self, algorithm, group=None, exchange="fused"):
        self.alg = algorithm
        self.group = group
        n, m = 2 * algorithm.pop_size, algorithm.n_objs
        self.sorter = (FusedShardedNonDominatedSort if exchange == "fused" else ShardedNonDominatedSort)(n, m, group)

    def setup(self, seed=0):
        return self.alg.setup(seed)

    def step(self, problem, state, pstate=None):
        from .operators.selection.non_dominate import select_by_rank
        alg = self.alg
        if state.generation == 0:                            # init_ask / init_tell once (std_workflow.py:204-214)
            fit, pstate = sharded_evaluate(problem, pstate, state.population, self.group)
            state = alg.init_tell(state, fit)
        cand, state = alg.ask(state)
        fit, pstate = sharded_evaluate(problem, pstate, cand, self.group)
        merged_pop = torch.cat([state.population, state.next_generation], dim=0)
        merged_fit = torch.cat([state.fitness, fit], dim=0)
        rank = self.sorter(merged_fit)
        pop, f = select_by_rank(merged_pop, merged_fit, rank, alg.pop_size)
        return state.replace(population=pop, fitness=f), pstate

    def run(self, problem, state, generations, pstate=None):
        for _ in range(generations):
            state, pstate = self.step(problem, state, pstate)
        return state, pstate
