"""Multi-GPU: sample partition + one film all-reduce.

The reference parallelises over whole-frame samples -- each thread renders 1-spp frames
and adds them into a shared accumulator (``include/amber/rendering/parallel.h:50-68``,
``include/amber/prelude/accumulator.h:136-166``).  Here a rank plays the thread: rank r
of G renders global sample indices [r*S/G, (r+1)*S/G) for every pixel into its own film;
the only exchange is ONE sum all-reduce of the film (NCCL over NVLink on GPUs, gloo in
the CPU tests), after which every rank holds the full sum and divides by S.  The sample
count travels IN the film (its trailing element, see ``amber_cuda_film_device``), so a step
costs exactly one collective and no device-to-host read.  Because
the random stream is keyed by (pixel, global sample index), the image does not depend
on G (up to float summation order).
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def partition_samples(total_samples: int, rank: int, world_size: int) -> tuple[int, int]:
    """Contiguous split of [0, total_samples): returns (begin, count) for ``rank``."""
    base, extra = divmod(total_samples, world_size)
    begin = rank * base + min(rank, extra)
    count = base + (1 if rank < extra else 0)
    return begin, count


def count_digits(samples: int) -> list[float]:
    """The film's trailing element: the sample count as three 16-bit digits, each exact in
    float32 under summation over up to 256 ranks (``include/amber_cuda.h``)."""
    return [float(samples & 0xFFFF), float((samples >> 16) & 0xFFFF), float(samples >> 32), 0.0]


def digits_count(tail) -> int:
    return int(round(float(tail[0]))) + (int(round(float(tail[1]))) << 16) + (int(round(float(tail[2]))) << 32)


def all_reduce_film(film: torch.Tensor, samples: int, group=None) -> int:
    """Sum ``film`` (any shape) across ranks together with the sample count: ONE collective
    over the film with the count appended.  In place; returns the total sample count."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return samples
    flat = torch.cat([film.reshape(-1), torch.tensor(count_digits(samples), dtype=film.dtype, device=film.device)])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    film.copy_(flat[:-4].reshape(film.shape))
    return digits_count(flat[-4:].tolist())


class _DevicePointer:
    """Exposes a raw device allocation to torch through __cuda_array_interface__."""

    def __init__(self, ptr: int, n_floats: int):
        self.__cuda_array_interface__ = {
            "shape": (n_floats,), "typestr": "<f4", "data": (ptr, False), "version": 2, "strides": None}


def film_tensor(device) -> torch.Tensor:
    """The context's device film (W*H float4 sums) as a torch view -- no copy."""
    ptr, n, _ = device.film_device()
    return torch.as_tensor(_DevicePointer(ptr, n), device=torch.device("cuda", device.device))


def reduce_device_film(device, group=None) -> None:
    """All-reduce the film of an ``amber_b200.render.Device`` in place: one collective, the sample
    count rides in the film's trailing element and ``resolve`` reads it there on the device."""
    from . import capi
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    film = film_tensor(device)                      # publishes the count into the trailing element
    dist.all_reduce(film, op=dist.ReduceOp.SUM, group=group)
    if film.is_cuda:
        torch.cuda.current_stream(film.device).synchronize()  # the context resolves on its own stream
    device.film_set_samples(capi.SAMPLES_IN_FILM)
