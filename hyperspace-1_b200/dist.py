"""Multi-GPU plumbing: one process per GPU, buckets sharded by owner(b) = b * N / numBuckets (contiguous
ranges, so every GPU writes a contiguous file-id range). Rows reach their owner through the NCCL all-to-all
inside hs_build_index_dist; torch.distributed is used only to hand the NCCL unique id to every rank.
Stands in for Spark's ShuffleExchangeExec under repartition (HS/actions/CreateActionBase.scala:129-130)."""


def owner_of(bucket, nranks, num_buckets):
    return bucket * nranks // num_buckets


def bucket_range(rank, nranks, num_buckets):
    """[lo, hi) of the buckets `rank` owns."""
    owned = [b for b in range(num_buckets) if owner_of(b, nranks, num_buckets) == rank]
    return (owned[0], owned[-1] + 1) if owned else (0, 0)


def init_comm(ctx, device=None):
    """Create the NCCL communicator of a context from an initialised torch.distributed process group."""
    import torch
    import torch.distributed as dist
    rank, world = dist.get_rank(), dist.get_world_size()
    if world == 1:
        ctx.comm_init(0, 1, None)
        return
    backend = dist.get_backend()
    dev = device if (device is not None and backend == "nccl") else torch.device("cpu")
    uid = torch.zeros(128, dtype=torch.uint8, device=dev)
    if rank == 0:
        uid = torch.frombuffer(bytearray(ctx.comm_unique_id()), dtype=torch.uint8).to(dev)
    dist.broadcast(uid, 0)
    ctx.comm_init(rank, world, uid.cpu().numpy().tobytes())


def build_index_dist(ctx, cols, n_indexed, num_buckets):
    """This rank's share of the rows in, the buckets this rank owns out: (arrays, bucket_offsets)."""
    return ctx.build_index(cols, n_indexed, num_buckets, dist=True)
