"""
Multi-GPU plumbing: one process per GPU (torchrun-style), ``torch.distributed`` over NCCL/NVLink
on the B200 box (``gloo`` on CPU in the host-logic tests).

The reference is single-process, single-device; the partitioning below is
the engine's own:

  * CCSD iteration - rank g owns a contiguous slice [lo, hi) of the VIRTUAL index: it stores
    ``<ab|ef>`` for a in the slice and ``<mb|ef>`` for b in the slice, and computes every term of
    r1 / r2 whose sliced virtual index falls in its range (the unsymmetrised r2 may carry a term
    in either pair orientation, because the final r2 + r2^T (pycc/ccwfn.py:621) restores it).
    One exchange step per iteration: a sum-allreduce of the partial (r1, r2).
  * (T) - the occupied pairs (i >= j), each standing for its j+1 triplets (i, j, k <= j), are
    dealt to the ranks by cost; one allreduce of a single double.
"""
import torch


def virtual_slice(nv, rank, world):
    """[lo, hi) of the virtual index owned by ``rank``.  Boundaries are kept even when nv is even
    so that slices of stride-1 virtual dimensions stay 16-byte aligned (vectorised copies)."""
    if world <= 1:
        return 0, nv
    unit = 2 if (nv % 2 == 0 and nv // 2 >= world) else 1
    n = nv // unit
    lo = (n * rank) // world * unit
    hi = (n * (rank + 1)) // world * unit
    return lo, hi


def flat_chunk(n, rank, world):
    """(lo, hi, chunk): the slice [lo, hi) of a flat n-vector owned by ``rank`` when it is cut into
    ``world`` equal chunks (even length, so slices stay 16-byte aligned; the last one may be short).
    Used for the DIIS history and for host <-> device shards of the amplitude vector."""
    if world <= 1:
        return 0, n, n
    chunk = ((n + world - 1) // world + 1) // 2 * 2
    lo = min(n, rank * chunk)
    return lo, min(n, lo + chunk), chunk


def deal_pairs(no, rank, world):
    """Occupied pairs (i >= j) for ``rank``: longest-processing-time-first by the cost j+1
    (number of triplets k <= j), ties broken by order, so every rank gets the same list."""
    pairs = [(i, j) for i in range(no) for j in range(i + 1)]
    if world <= 1:
        return pairs
    order = sorted(range(len(pairs)), key=lambda q: (-(pairs[q][1] + 1), q))
    load = [0] * world
    mine = []
    for q in order:
        r = min(range(world), key=lambda x: (load[x], x))
        load[r] += pairs[q][1] + 1
        if r == rank:
            mine.append(q)
    return [pairs[q] for q in sorted(mine)]


class Comm(object):
    """Sum-allreduce over a ``torch.distributed`` process group (None = single process)."""

    def __init__(self, group=None):
        self.group = None
        self.rank, self.world = 0, 1
        if group is not None:
            import torch.distributed as dist
            if not dist.is_initialized():
                raise RuntimeError("torch.distributed is not initialised (launch with torchrun / init_process_group)")
            self.group = dist.group.WORLD if group in ("world", True) else group
            self.rank = dist.get_rank(self.group)
            self.world = dist.get_world_size(self.group)
        self.bytes_reduced = 0

    def allreduce_(self, *tensors):
        """In-place sum over the ranks; identical bits on every rank (NCCL / gloo guarantee)."""
        if self.world <= 1:
            return
        import torch.distributed as dist
        for t in tensors:
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
            self.bytes_reduced += t.numel() * t.element_size()

    def reduce_scatter_(self, full, lo, n):
        """Sum ``full`` (world * n elements) over the ranks, leaving full[lo : lo + n] (lo = rank * n) complete
        on this rank; the rest of ``full`` is unspecified afterwards.  In place: NCCL's reduce-scatter with
        recvbuff = sendbuff + rank * n.  gloo has no reduce-scatter: an allreduce does (host-logic tests)."""
        if self.world <= 1:
            return
        import torch.distributed as dist
        if dist.get_backend(self.group) == "nccl":
            dist.reduce_scatter_tensor(full[lo:lo + n], full, op=dist.ReduceOp.SUM, group=self.group)
            self.bytes_reduced += n * full.element_size()
        else:
            dist.all_reduce(full, op=dist.ReduceOp.SUM, group=self.group)
            self.bytes_reduced += full.numel() * full.element_size()

    def allreduce_scalar(self, x):
        if self.world <= 1:
            return float(x)
        import torch.distributed as dist
        dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(self.group) == "nccl" \
            else torch.device("cpu")
        t = torch.tensor([float(x)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return float(t.item())

    def _device(self):
        import torch.distributed as dist
        return torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(self.group) == "nccl" \
            else torch.device("cpu")

    def allreduce_scalars(self, xs):
        """Sum a short list of host doubles over the ranks with ONE collective (the new Gram-matrix
        entries of a DIIS step); every rank receives identical values."""
        if self.world <= 1:
            return [float(x) for x in xs]
        import torch.distributed as dist
        t = torch.tensor([float(x) for x in xs], dtype=torch.float64, device=self._device())
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t.tolist()

    def allgather_into(self, full, chunk, in_place=False):
        """full[r*len(chunk) : (r+1)*len(chunk)] = rank r's ``chunk`` on every rank (equal lengths).
        ``in_place``: ``chunk`` IS this rank's slice of ``full`` (NCCL gathers in place; gloo gets a copy)."""
        if self.world <= 1:
            if not in_place:
                full[:chunk.numel()].copy_(chunk)
            return
        import torch.distributed as dist
        if dist.get_backend(self.group) == "nccl":
            dist.all_gather_into_tensor(full, chunk, group=self.group)
        else:
            dist.all_gather(list(full.split(chunk.numel())), chunk.clone() if in_place else chunk, group=self.group)
        self.bytes_gathered = getattr(self, "bytes_gathered", 0) + full.numel() * full.element_size()

    def barrier(self):
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier(group=self.group)


class nvtx_range(object):
    """NVTX range around one builder / phase (ranges per builder for ncu / nsys timelines).
    Enabled with PCCB_NVTX=1 on a CUDA device; otherwise a no-op with no per-call cost beyond one attribute test."""
    import os as _os
    enabled = _os.environ.get("PCCB_NVTX", "0") == "1"

    def __init__(self, name):
        self.name = name

    def __enter__(self):
        if nvtx_range.enabled and torch.cuda.is_available():
            torch.cuda.nvtx.range_push(self.name)
            self.pushed = True
        else:
            self.pushed = False
        return self

    def __exit__(self, *exc):
        if self.pushed:
            torch.cuda.nvtx.range_pop()
        return False
