"""One ensemble over the GPUs of one box, one process per GPU (torchrun).

torch.distributed is the plumbing only: it carries the 128-byte NCCL id and the 64-byte CUDA IPC
handles between the ranks; the exchanges themselves (NCCL all-gather of the updated colour, or NVLink
peer loads inside the update kernel) are issued by the C-ABI runtime on the sampler's own stream.
Host-side shard arithmetic lives here so it can be tested on CPU with gloo.
"""
import ctypes as C
import os

from . import _capi


def bind_host_to_device(pci_bus_id):
    """One process per GPU: run this process on the CPU cores NVML reports as local to its GPU (the GPU's NUMA node), so that
    host buffers pinned AFTERWARDS are allocated there and the rank's host<->device copies do not cross the socket link.
    With 4-8 ranks on a two-socket box whose buffers all sit on one node, the copies of an end-to-end call share one node's memory
    and the inter-socket link.  Returns the cores bound to, or None when NVML, the device or an allowed core is not there (nothing
    is changed then)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        handle = pynvml.nvmlDeviceGetHandleByPciBusId(pci_bus_id.encode() if isinstance(pci_bus_id, str) else pci_bus_id)
        words = (max(os.cpu_count() or 1, 1) + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(handle, words)
        cores = {64 * i + b for i, word in enumerate(mask) for b in range(64) if (int(word) >> b) & 1}
        cores &= set(os.sched_getaffinity(0))
        if not cores:
            return None
        os.sched_setaffinity(0, cores)
        return sorted(cores)
    except Exception:
        return None


def shard_slots(num_walkers, rank, world):
    """Owned slots [begin, end) of each colour for `rank`: contiguous walker-index shards of
    the K = W/2 walkers per colour (MCMCpp/EnsembleSampler.h:225-226 numbering)."""
    K = num_walkers // 2
    if K % world:
        raise ValueError("walkers per colour (%d) must divide evenly over %d GPUs" % (K, world))
    n = K // world
    return rank * n, (rank + 1) * n


def owned_walkers(num_walkers, rank, world):
    """Global walker indices this rank updates (red slice then black slice)."""
    K = num_walkers // 2
    b, e = shard_slots(num_walkers, rank, world)
    return list(range(b, e)) + list(range(K + b, K + e))


def shard_data_rows(num_rows, rank, world):
    """DATA sharding: contiguous row range [begin, end) of the Calculator's data matrix held by `rank`
    (remainder rows go to the first ranks)."""
    base, rem = divmod(num_rows, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def dim_slice(num_params, rank, world):
    """DIM sharding: columns [begin, end) of a walker row held by `rank` -- pairs of columns dealt out so that every
    slice but possibly the last is even (mirrors mcmcpp_dim_slice in the C-ABI)."""
    pairs = num_params // 2
    base, rem = divmod(pairs, world)
    begin = 2 * (rank * base + min(rank, rem))
    count = 2 * (base + (1 if rank < rem else 0))
    if rank == world - 1:
        count = num_params - begin
    return begin, begin + count


def sequential_draw_index(step, half, slot, num_walkers):
    """First (0-based) draw of the sampler's single pcg64 stream used by (step, colour, slot):
    3 draws per update in red-then-black walker order -- independent of the GPU count."""
    return 3 * (step * num_walkers + half * (num_walkers // 2) + slot)


def env_rank_world():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def connect(sampler, group=None):
    """Collective over all ranks: give `sampler` (created with rank/world/shardMode) what its shard mode exchanges
    through -- ALLGATHER / DATA: an NCCL communicator; PEER / DIM: the other ranks' peer-mapped arrays (CUDA IPC
    handles; no NCCL, so the ranks may be processes sharing one GPU)."""
    import torch.distributed as dist
    if sampler.world == 1:
        return
    if sampler.shardMode in ("allgather", "data"):
        ident = [None]
        if sampler.rank == 0:
            buf = C.create_string_buffer(128)
            _capi.check(_capi.lib().mcmcpp_comm_unique_id(buf))
            ident[0] = buf.raw
        dist.broadcast_object_list(ident, src=0, group=group)
        sampler.commInit(ident[0])
    if sampler.shardMode in ("peer", "dim"):
        handles = [None] * sampler.world
        dist.all_gather_object(handles, sampler.peerExport(), group=group)
        sampler.peerImport(handles)
    dist.barrier(group=group)


def reduce_counts(sampler, group=None):
    """Ensemble-wide (accepted, total): the C-ABI returns this rank's owned walkers only."""
    import torch
    import torch.distributed as dist
    a, t = sampler._counts()
    if sampler.world == 1 or sampler.shardMode in ("data", "dim"):      # DATA / DIM sharding: every rank counts every walker
        return a, t
    dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    v = torch.tensor([a, t], dtype=torch.int64, device=dev)
    dist.all_reduce(v, group=group)
    return int(v[0]), int(v[1])


class RankGroup:
    """All ranks of ONE DIM-sharded ensemble inside one process on one device (a one-GPU box): samplers[r] was created with
    rank=r, world=len(samplers), shardMode="dim".  runMCMC steps them in lockstep, every rank's update kernel in a single
    cooperative launch (ranks that wait on one another must not be separate launches on one GPU).  Same results as one process
    per GPU."""

    def __init__(self, samplers):
        self.samplers = list(samplers)
        self._arr = (C.c_void_p * len(self.samplers))(*[s._h for s in self.samplers])
        _capi.check(_capi.lib().mcmcpp_sampler_peer_import_local(self._arr, len(self.samplers)))

    def runMCMC(self, numSteps):
        rc = _capi.check(_capi.lib().mcmcpp_sampler_run_group(self._arr, len(self.samplers), int(numSteps)))
        return rc != _capi.CHAIN_FULL
