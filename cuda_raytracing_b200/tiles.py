"""Multi-GPU image partition: BAND_ROWS-row bands interleaved over ranks, scene replicated,
one exchange per frame -- a gather of every rank's bands to rank 0 (NCCL on GPUs, gloo in the CPU tests).

The partition itself is applied inside the kernels (rtb_render_params.tile_rank / tile_count); this module is
the host-side bookkeeping: which rows a rank owns, packing them for the collective, re-assembling on rank 0.
Works on any torch device/backend so the logic is covered by world_size-2 gloo tests without a GPU.
"""
import torch
import torch.distributed as dist

BAND_ROWS = 8


def rank_rows(height, rank, world, device="cpu"):
    """Row indices (ascending) owned by `rank`: bands b with b % world == rank."""
    bands = (height + BAND_ROWS - 1) // BAND_ROWS
    ids = torch.arange(rank, max(bands, rank), world, device=device)
    rows = (ids[:, None] * BAND_ROWS + torch.arange(BAND_ROWS, device=device)[None, :]).reshape(-1)
    return rows[rows < height]


def max_rows(height, world):
    bands = (height + BAND_ROWS - 1) // BAND_ROWS
    return ((bands + world - 1) // world) * BAND_ROWS


class FrameGather:
    """Pre-allocated buffers for the per-frame gather; frame is a (H, pitch) uint8 tensor on every rank."""

    def __init__(self, height, pitch, rank, world, device, group=None):
        self.h, self.pitch, self.rank, self.world, self.group = height, pitch, rank, world, group
        self.rows = [rank_rows(height, r, world, device) for r in range(world)]
        n = max_rows(height, world)
        self.send = torch.zeros((n, pitch), dtype=torch.uint8, device=device)
        self.recv = [torch.zeros((n, pitch), dtype=torch.uint8, device=device) for _ in range(world)] if rank == 0 else None

    def gather(self, frame):
        """After every rank rendered its own bands into `frame`, make rank 0's frame complete."""
        if self.world == 1:
            return frame
        mine = self.rows[self.rank]
        self.send[: len(mine)] = frame[mine]
        dist.gather(self.send, self.recv, dst=0, group=self.group)
        if self.rank == 0:
            for r in range(1, self.world):
                rr = self.rows[r]
                frame[rr] = self.recv[r][: len(rr)]
        return frame
