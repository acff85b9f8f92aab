"""Host-side logic of the multi-GPU split (one process per GPU, torchrun): which sample indices a rank renders and how
the per-rank films meet.

The reference parallelises the same loop over CPU workers by image blocks (SamplingIntegrator::render ->
BlockedRenderProcess, src/librender/integrator.cpp:103-139) or, in the responsive scheduler, by giving each thread its
own sampler stream over the whole image (src/mitsuba/im_render.cpp:180-247) and summing the targets.  Here rank r
renders the sample indices {s : s mod world == r} of every pixel (the counter-based sampler makes the streams disjoint
by construction), so ranks are perfectly balanced and exchange nothing until the single film reduction
(binding.Comm.film_reduce -> mtsb_film_reduce -> ncclReduce over NVLink).

torch.distributed is used as plumbing only: to ship the 128-byte NCCL id from rank 0 to the others.
"""


def sample_share(spp, world, rank):
    """Number of sample indices in [0, spp) that rank `rank` of `world` renders."""
    if world < 1 or not (0 <= rank < world) or spp < 0:
        raise ValueError("bad split: spp=%r world=%r rank=%r" % (spp, world, rank))
    return (spp - rank + world - 1) // world if spp > rank else 0


def sample_indices(spp, world, rank):
    return list(range(rank, spp, world))


def passes(spp, world, rank, spp_per_pass=0):
    """(spp_begin, spp_count) arguments of mtsb_render_pass(stride=world, phase=rank) that cover this rank's share;
    the k-th local sample has index rank + k * world, so a pass starting at local sample k0 has spp_begin = k0 * world."""
    mine = sample_share(spp, world, rank)
    step = mine if spp_per_pass <= 0 else spp_per_pass
    out, k = [], 0
    while k < mine:
        c = min(step, mine - k)
        out.append((k * world, c))
        k += c
    return out


def pass_indices(spp_begin, spp_count, world, rank):
    """The sample indices one pass touches (what generate_kernel computes: sppBegin + phase + k * stride)."""
    return [spp_begin + rank + k * world for k in range(spp_count)]


def exchange_unique_id(dist, make_id, rank, src=0):
    """Rank `src` creates the NCCL id (make_id() -> 128 bytes); everyone returns the same bytes."""
    box = [make_id() if rank == src else None]
    dist.broadcast_object_list(box, src=src)
    if not isinstance(box[0], (bytes, bytearray)) or len(box[0]) != 128:
        raise RuntimeError("NCCL id exchange failed on rank %d" % rank)
    return bytes(box[0])


def render_distributed(render, comm, spp, world, rank, spp_per_pass=0, root=0):
    """Queues this rank's passes and the film reduction on the context stream (all asynchronous); the caller syncs."""
    for begin, count in passes(spp, world, rank, spp_per_pass):
        render.render_pass(begin, count, spp, stride=world, phase=rank)
    if comm is not None and world > 1:
        comm.film_reduce(render, root)
