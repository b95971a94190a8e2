"""Multi-shard replay: one ReplayShard per GPU / rank.

The reference shards replay by running one server per port with capacity // n and lets the
learner poll the shards round-robin, one whole batch from one shard at a time, with no
cross-shard normalisation (test/apex-dqn/trainer.py:33-36,52-61; actors are pinned to a shard,
worker.py:94).  Two modes are provided:

  Mode R (reference parity): step t draws its batch from shard t mod n; no collective.
  Mode G (north_star "NCCL-reduced shard totals"): every rank contributes its 16-byte shard root
         {total, min}; after one all-gather each rank knows the exclusive prefix of totals and
         serves the stratified targets of the GLOBAL cdf that fall inside its range, with the
         global minimum in the importance weights.  The rule, identical on host and device
         (reth_b200/csrc/tree.cuh per_sample_kernel):
             total   = ((t_0 + t_1) + ...) in rank order,  seg = total / B
             target_i = (i + u_i) * seg
             rank r serves i in [first_r, end_r):  first_r = min{i : target_i >= prefix_r},
             end_r = B for the last rank, else min{i : target_i >= prefix_r + t_r}
             local target = target_i - prefix_r;  p_min = min over non-empty shards of min_r.

torch.distributed carries the exchange (NCCL on GPUs, gloo in the CPU tests); the data path has
no other collective: frames never leave their shard unless the learner is remote.
"""
import numpy as np


def owner_round_robin(step, n_shards):
    """Mode R: the shard that serves the whole batch of learner step `step` (trainer.py:33-36)."""
    return int(step) % int(n_shards)


def combine_roots(all_roots):
    """all_roots: float64[2*n] = {total_r, min_r} in rank order -> (total, prefix[n], p_min)."""
    r = np.asarray(all_roots, dtype=np.float64).reshape(-1, 2)
    prefix = np.empty(len(r), np.float64)
    total = 0.0
    mn = 0.0
    for s, (t, m) in enumerate(r):
        prefix[s] = total
        total = total + t
        if t != 0.0 and m != 0.0:
            mn = m if (mn == 0.0 or m < mn) else mn
    return total, prefix, (mn if mn != 0.0 else 1.0)


def global_quota(all_roots, shard, batch, uniforms):
    """Host statement of the Mode-G range rule for explicit uniforms u[batch]:
    -> (first, count, local_targets float64[count], p_min)."""
    r = np.asarray(all_roots, dtype=np.float64).reshape(-1, 2)
    n = len(r)
    total, prefix, pmin = combine_roots(all_roots)
    u = np.asarray(uniforms, dtype=np.float64)
    assert len(u) == batch
    seg = total / float(batch)
    targets = (np.arange(batch, dtype=np.float64) + u) * seg

    def lower_bound(bound):
        if not bound > 0.0:
            return 0
        return int(np.searchsorted(targets, bound, side="left"))   # targets are monotone in i
    first = lower_bound(prefix[shard])
    end = batch if shard == n - 1 else lower_bound(prefix[shard] + r[shard, 0])
    end = max(end, first)
    return first, end - first, targets[first:end] - prefix[shard], pmin


def global_quota_multi(all_roots, shard, batch, uniforms, groups):
    """Host statement of the packed Mode-G prefetch group (reth_replay_sample_global_multi): `groups`
    independent global draws of `batch` from the same roots, uniforms u[groups*batch].
    -> (quota int64[groups, 3] = {first, count, packed offset}, local_targets float64[sum(count)], p_min)."""
    u = np.asarray(uniforms, dtype=np.float64).reshape(groups, batch)
    quota = np.zeros((groups, 3), np.int64)
    targets, off, pmin = [], 0, 1.0
    for g in range(groups):
        first, count, t, pmin = global_quota(all_roots, shard, batch, u[g])
        quota[g] = (first, count, off)
        targets.append(t)
        off += count
    return quota, (np.concatenate(targets) if targets else np.empty(0)), pmin


class ShardGroup:
    """This rank's shard inside a torch.distributed process group."""

    def __init__(self, shard, group=None, device=None):
        import torch
        import torch.distributed as dist
        self.shard = shard
        self.dist = dist
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.device = device if device is not None else torch.device("cuda", shard.device)
        self.all_roots = torch.zeros(2 * self.world, dtype=torch.float64, device=self.device)
        self._mine = torch.as_tensor(shard.root_dev(), device=self.device)   # zero-copy view of {total, min}
        self.board = False

    def exchange_roots(self, async_op=False, stream=None):
        """all-gather of the 16-byte shard roots -> device float64[2*world] (self.all_roots).
        Default: torch.distributed (NCCL over NVLink).  After connect_board(): ONE kernel of this library on
        `stream` (peer stores into every rank's roots board + a wait on the own board), no host round trip."""
        if self.board:
            self.shard.roots_exchange(3, stream=stream)
            return None
        return self.dist.all_gather_into_tensor(self.all_roots, self._mine, group=self.group, async_op=async_op)

    def connect_board(self):
        """Map every rank's roots board into this process (CUDA IPC; peers are reached over NVLink) and
        switch exchange_roots() to the device-side exchange.  Collective: every rank must call it."""
        import torch
        blobs = [None] * self.world
        self.dist.all_gather_object(blobs, self.shard.board_export(), group=self.group)
        self.shard.board_connect(blobs, self.rank)
        self.all_roots = torch.as_tensor(self.shard.all_roots_dev(self.world), device=self.device)
        self.dist.barrier(group=self.group)      # every board is mapped before anyone publishes
        self.board = True

    def connect_learner(self, learner_rank):
        """Map the learner rank's batch ring into this process (CUDA IPC; a ring on another GPU is
        reached over NVLink).  Collective: every rank of the group must call it."""
        from .shard import IpcRing
        blobs = [None] * self.world
        self.dist.all_gather_object(blobs, self.shard.ipc_export(), group=self.group)
        self.learner_rank = int(learner_rank)
        self.learner_ring = None if self.rank == learner_rank else IpcRing(blobs[learner_rank])
        return self.learner_ring

    def assemble(self, slot, dst_slot, stream=None):
        """Cross-shard batch assembly: after sample_global(slot) on every rank, each rank writes its
        slice of the global batch into the learner rank's batch slot `dst_slot` (!= slot on the learner
        rank) -- the learner's own slice by a local gather, the others by peer stores from the kernel.  The caller orders
        the learner's reads after a collective on the same stream (e.g. exchange_roots())."""
        self.shard.gather_to(slot, self.learner_ring, dst_slot, stream=stream)

    def sample_global(self, global_batch, beta=0.4, uniforms=None, seed=0, slot=0, stream=None, groups=1):
        """Draw this rank's part of one global stratified batch into batch slot `slot`;
        -> (first, count) of the global batch positions it serves.
        groups > 1 (prefetch group; uniforms on the device or None): `groups` global batches from the same tree
        state, this rank's shares packed back to back from `slot` on -> int64[groups, 3] = {first, count, offset}
        per batch; gather_multi / update_slots over the same slots handle the packed batch."""
        if groups > 1:
            self.shard.sample_global_multi(global_batch, self.all_roots, self.world, self.rank, slot, groups,
                                           beta=beta, uniforms_dev=uniforms, seed=seed, stream=stream)
            return self.shard.slot_quota(slot, groups, stream=stream)
        self.shard.sample_global(global_batch, self.all_roots, self.world, self.rank, beta=beta,
                                 uniforms=uniforms, seed=seed, slot=slot, stream=stream)
        return self.shard.slot_meta(slot, stream=stream)
