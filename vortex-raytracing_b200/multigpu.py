"""Screen-space sharding of one frame over several PROCESSES, one per GPU — the round-1 arrangement, kept for
hosts that run one Vortex process per GPU.  The backend itself now spans the GPUs of a box behind ONE vx_device_h
(VORTEX_B200_GPUS / vx_b200_select_gpus, csrc/vx_group.inc), which is what bench.py measures; this module keeps the band
arithmetic the tests use and the torch.distributed helpers.

The image is cut into bands of `band_rows` rows; band j belongs to rank j % world (the same rule the CUDA tile scheduler
applies after vx_b200_set_partition).  The scene is replicated: every rank uploads the same buffers through vx_copy_to_dev.
After the render each rank holds its own bands inside a full-size framebuffer; `gather_framebuffer` moves the missing
bands into rank 0's framebuffer with ONE collective (torch.distributed gather: NCCL over NVLink on the GPU box, gloo in the
CPU tests) so that the single-buffer vx_copy_from_dev contract of the reference host (raycast/tracer.cpp:239) still holds
on rank 0.  There is no other data-path communication: rays are independent and the scene is read-only.
`share_framebuffer` removes even that collective: non-owner ranks map the owner's framebuffer (CUDA IPC over NVLink/NVSwitch)
and their kernels write into it directly, so the transfer overlaps the tracing tile by tile.
"""
import numpy as np
import torch
import torch.distributed as dist


def num_bands(height, band_rows=8):
    return (height + band_rows - 1) // band_rows


def band_owner(band, world):
    return band % world


def owned_bands(rank, world, height, band_rows=8):
    return list(range(rank, num_bands(height, band_rows), world))


def owned_rows(rank, world, height, band_rows=8):
    """row indices rendered by `rank` (ascending)"""
    y = np.arange(height)
    return y[(y // band_rows) % world == rank]


def rays_per_rank(width, height, world, band_rows=8):
    return [int(len(owned_rows(r, world, height, band_rows))) * width for r in range(world)]


class _runtime_stream:
    """Run the enclosed collectives on the vx runtime's own CUDA stream (vx_b200_stream) and wait for them before returning: the
    runtime issues all of its work on that non-blocking stream and does not synchronise with torch's current stream, so a
    collective left on torch's stream could race the next vx_start or a download.  No-op for CPU tensors (gloo tests)."""

    def __init__(self, tensor, stream):
        self.ctx = None
        if stream and tensor.is_cuda:
            self.ext = torch.cuda.ExternalStream(int(stream), device=tensor.device)
            self.ctx = torch.cuda.stream(self.ext)

    def __enter__(self):
        if self.ctx:
            self.ctx.__enter__()

    def __exit__(self, *a):
        if self.ctx:
            self.ctx.__exit__(*a)
            self.ext.synchronize()


def _padded_view(fb, band_rows, world):
    """fb [H, W] -> (padded [NBpad, band_rows, W] or a view when no padding is needed, NBpad)"""
    H, W = fb.shape
    nb = num_bands(H, band_rows)
    nb_pad = (nb + world - 1) // world * world
    if nb_pad * band_rows == H:
        return fb.view(nb_pad, band_rows, W), nb_pad, False
    pad = torch.zeros((nb_pad * band_rows, W), dtype=fb.dtype, device=fb.device)
    pad[:H] = fb
    return pad.view(nb_pad, band_rows, W), nb_pad, True


def gather_framebuffer(fb, rank, world, band_rows=8, group=None, dst=0, stream=None):
    """fb: [H, W] int32/uint32-as-int32 tensor (this rank's framebuffer; only its own bands are valid).
    On return rank `dst`'s fb holds every band.  One gather of H*W*4/world bytes per non-root rank.
    stream: the runtime's CUDA stream (Device.stream()) when fb aliases a vx buffer — the gather then runs on it and has
    completed when this returns."""
    if world == 1:
        return fb
    with _runtime_stream(fb, stream):
        return _gather_framebuffer(fb, rank, world, band_rows, group, dst)


def _gather_framebuffer(fb, rank, world, band_rows, group, dst):
    H, W = fb.shape
    bands, nb_pad, padded = _padded_view(fb, band_rows, world)
    mine = bands[rank::world].contiguous()                       # pack my bands: [NBpad/world, band_rows, W]
    if rank == dst:
        recv = [torch.empty_like(mine) for _ in range(world)]
        dist.gather(mine, gather_list=recv, dst=dst, group=group)
        if padded:
            full = torch.empty((nb_pad, band_rows, W), dtype=fb.dtype, device=fb.device)
            for r in range(world):
                full[r::world] = recv[r]
            fb.copy_(full.view(nb_pad * band_rows, W)[:H])
        else:
            for r in range(world):
                if r != rank:
                    bands[r::world] = recv[r]
    else:
        dist.gather(mine, gather_list=None, dst=dst, group=group)
    return fb


def share_framebuffer(tracer, rank, world, owner=0, group=None):
    """Peer-memory alternative to gather_framebuffer: rank `owner` exports its device framebuffer (CUDA IPC handle of its arena +
    offset, vx_b200_export_buffer) and every other rank maps it (vx_b200_set_remote_framebuffer), after which their render
    kernels store their bands' pixels straight into the owner's framebuffer over NVLink while they trace.  No collective
    remains on the data path; the image is complete on `owner` once all ranks' vx_ready_wait returned (barrier before reading)."""
    if world == 1:
        return
    payload = [tracer.out.ipc_export() if rank == owner else None]
    dist.broadcast_object_list(payload, src=owner, group=group)
    if rank != owner:
        handle, offset = payload[0]
        tracer.dev.set_remote_framebuffer(handle, offset, tracer.width * tracer.height * 4)


def unshare_framebuffer(tracer, rank, world, owner=0):
    if world > 1 and rank != owner:
        tracer.dev.set_remote_framebuffer(None)


def scene_tensors(tracer, cuda_device):
    """the tracer's device scene buffers as flat uint8 torch tensors (aliases, no copy), in upload order"""
    out = {}
    for name, b in tracer.bufs.items():
        if b.size:
            out[name] = torch.as_tensor(DeviceArray(b.device_ptr, (b.size,), "|u1"), device=torch.device("cuda", cuda_device))
    return out


def broadcast_scene(tracer, tensors, rank, world, src=0, group=None):
    """Scene replication without N host uploads: rank `src` has uploaded the scene through vx_copy_to_dev; every other rank
    receives the same bytes into its own vx buffers over NVLink (one NCCL broadcast per buffer) and tells its runtime that the
    buffers changed (vx_b200_mark_written), so that its GPU-side records are rebuilt exactly as after an upload."""
    if world == 1:
        return
    for name, t in tensors.items():
        with _runtime_stream(t, tracer.dev.stream()):      # on the runtime's stream, complete before mark_written
            dist.broadcast(t, src=src, group=group)
    if rank != src:
        for name in tensors:
            tracer.bufs[name].mark_written()


class DeviceArray:
    """Wraps a raw device pointer of the vx arena (vx_b200_device_ptr) as a __cuda_array_interface__ provider so that
    torch.as_tensor can alias it without a copy (plumbing for NCCL; torch never computes on it)."""

    def __init__(self, ptr, shape, typestr="<i4"):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 3,
                                         "strides": None}


def framebuffer_tensor(tracer, cuda_device):
    """the tracer's device framebuffer as a [H, W] int32 torch tensor (alias, no copy)"""
    arr = DeviceArray(tracer.out.device_ptr, (tracer.height, tracer.width))
    return torch.as_tensor(arr, device=torch.device("cuda", cuda_device))
