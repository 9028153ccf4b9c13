"""One process per GPU, no data-path collective (the tracker path is "replicas only").

Rank r tracks its own clips; the only communication is the max-over-ranks of the timed interval that bench.py
reports, and the rendezvous itself.  Backend "nccl" on GPUs, "gloo" in the CPU tests.
"""
import os


def init(backend=None):
    """-> (rank, world, local_rank); initialises torch.distributed when WORLD_SIZE > 1"""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if not dist.is_initialized():
            if backend is None:
                backend = "nccl" if torch.cuda.is_available() else "gloo"
            kw = {"device_id": torch.device("cuda", local)} if backend == "nccl" else {}
            dist.init_process_group(backend, **kw)
    return rank, world, local


def clip_shard(n_clips, rank, world):
    """contiguous shard of clip indices for this rank (config 5: 64 clips over 8 GPUs -> 8 each)"""
    per, rem = divmod(n_clips, world)
    lo = rank * per + min(rank, rem)
    return list(range(lo, lo + per + (1 if rank < rem else 0)))


def max_over_ranks(value, device="cpu"):
    """max of a python float over all ranks (identity when not distributed)"""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()


def finalize():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()
