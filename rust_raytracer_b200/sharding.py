"""Sample-strata sharding over ranks: global sample index
s = tid*k*k + sy*k + sx in [0, T*k*k); rank g of G renders the contiguous block
[g*S/G, (g+1)*S/G) of every pixel. Weak scaling: a job of G ranks has G x the samples."""


def shard_range(total_samples, rank, world):
    b = total_samples * rank // world
    e = total_samples * (rank + 1) // world
    return b, e


def shard_params(params, rank, world):
    """Copy of RtgpuRenderParams restricted to this rank's block of the T*k*k samples."""
    p = type(params).from_buffer_copy(params)
    total = p.replicas * p.sqrt_spt * p.sqrt_spt
    p.sample_begin, p.sample_end = shard_range(total, rank, world)
    return p
