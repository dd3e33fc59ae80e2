"""One process per GPU (torchrun): sample sharding and the film reduce.  No collective runs during rendering -- the only
exchange of the path is one sum of the fp32 films:

    GPU g renders the sample indices s = g (mod G) of every pixel (sample_offset = g, sample_stride = G) into a RAW-SUM
    film; reduce(sum) to rank 0; rank 0 divides by the job's spp.

Streams are keyed by (pixel, global sample index), never by rank, so the G-GPU film equals the 1-GPU film up to fp32
summation order.  `render_fn(sample_offset, sample_stride, film_mode_sum=True)` is whatever produces this rank's film
(api.Scene.render_bdpt on a GPU; the gloo test on CPU plugs in the host build of the device headers).
"""
import torch
import torch.distributed as dist


def shard(rank, world, spp):
    """(sample_offset, sample_stride, n_local): this rank renders s = offset, offset + stride, ... < spp."""
    if not (0 <= rank < world) or spp < 0:
        raise ValueError("bad rank / world / spp")
    n_local = len(range(rank, spp, world))
    return rank, world, n_local


def reduce_film(film_sum, spp_total, dst=0, group=None):
    """Sum the raw-sum films onto `dst` (NCCL over NVLink for CUDA tensors, gloo for CPU ones) and resolve the mean
    there.  `film_sum` is modified in place; ranks other than dst keep an unspecified partial result."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(film_sum, dst=dst, op=dist.ReduceOp.SUM, group=group)
        if dist.get_rank(group) != dst:
            return film_sum
    film_sum.mul_(1.0 / float(spp_total))
    return film_sum


def render_sharded(render_fn, spp_total, film_like, group=None):
    """Render this rank's shard with render_fn(sample_offset, sample_stride) -> raw-sum film (array-like, H x W x 3) and
    reduce.  Returns (film tensor on this rank, n_local samples per pixel)."""
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    off, stride, n_local = shard(rank, world, spp_total)
    out = render_fn(off, stride)
    t = out if isinstance(out, torch.Tensor) else torch.as_tensor(out)
    t = t.to(device=film_like.device, dtype=torch.float32).contiguous()
    return reduce_film(t, spp_total, 0, group), n_local
