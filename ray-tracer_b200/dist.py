"""Multi-GPU rendering: one process per GPU, image rows interleaved over ranks, one framebuffer gather.

The reference's only parallelism is its thread-per-core row partition — thread i renders the rows with
y % cpuCount == i and results funnel through one mpsc channel (examples/book-one.rs:54-88).  Here rank r of G
renders rows y = r + k*G on its own GPU (scene replicated, Philox keys are functions of the GLOBAL pixel so the
image does not depend on G), and the channel becomes ONE collective: an all-gather of the [rows, W, 3] f32
shards over NCCL (NVLink 5 / NVSwitch), followed by the rt_deinterleave_rows kernel that puts row y back at
index y.  `torch.distributed` is plumbing only; on CPU tensors (gloo, used by the tests of this host logic) the
de-interleave is pure index arithmetic.
"""
import torch
import torch.distributed as dist


def shard_rows(height, rank, world):
    """Number of rows rank `rank` of `world` owns (rows y = rank + k*world)."""
    return len(range(rank, height, world))


def padded_rows(height, world):
    return -(-height // world)


def gather_rows(local, height, world, device_ctx=None, group=None):
    """All-gather row shards and de-interleave them.

    local: [padded_rows(height, world), W, 3] f32 tensor holding this rank's rows (rows past shard_rows() are padding).
    Returns the full [height, W, 3] image on every rank.  CUDA tensors need `device_ctx` (ray-tracer_b200.Device):
    the de-interleave runs as the rt_deinterleave_rows kernel on the current stream.
    """
    rows_padded, width = local.shape[0], local.shape[1]
    assert rows_padded == padded_rows(height, world) and local.shape[2] == 3 and local.is_contiguous()
    if world == 1:
        return local[:height]
    gathered = torch.empty((world * rows_padded, width, 3), dtype=local.dtype, device=local.device)  # shards back to back
    dist.all_gather_into_tensor(gathered, local, group=group)
    if local.is_cuda:
        if device_ctx is None:
            raise ValueError("gather_rows on CUDA tensors needs the Device that owns the kernels")
        full = torch.empty((height, width, 3), dtype=local.dtype, device=local.device)
        device_ctx.deinterleave_rows(gathered.data_ptr(), full.data_ptr(), width, height, world, torch.cuda.current_stream().cuda_stream)
        return full
    return gathered.view(world, rows_padded, width, 3).permute(1, 0, 2, 3).reshape(rows_padded * world, width, 3)[:height].contiguous()


def render_sharded(scene, cam, width, height, spp, max_depth, seed, rank, world, queue_capacity=0, flags=0, group=None):
    """This rank's rows through rt_render_device (outputs stay in HBM), then the gather.  Returns (full image sums
    [H, W, 3] on the GPU, RtStats of this rank's render)."""
    rows_padded = padded_rows(height, world)
    local = torch.zeros((rows_padded, width, 3), dtype=torch.float32, device=f"cuda:{scene.device.device_id}")
    prm = scene.params(width, height, spp, max_depth, seed, rank, world, queue_capacity, flags)
    stats = scene.render_device(cam, prm, local.data_ptr(), None, torch.cuda.current_stream().cuda_stream)
    return gather_rows(local, height, world, scene.device, group), stats
