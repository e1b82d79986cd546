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
    """Pre-allocated buffers for the per-frame gather; frame is a (H, pitch) uint8 tensor on every rank.

    One exchange per frame: every rank r > 0 packs its rows (one index_select) and sends them to rank 0; rank 0 posts
    all receives as one batched group (a single NCCL group call on GPUs) and scatters everything it received into
    its frame with one index_select + one index_copy_, whatever the number of ranks."""

    def __init__(self, height, pitch, rank, world, device, group=None):
        self.h, self.pitch, self.rank, self.world, self.group = height, pitch, rank, world, group
        self.rows = [rank_rows(height, r, world, device) for r in range(world)]
        n = max_rows(height, world)
        self.n = n
        self.send = torch.zeros((n, pitch), dtype=torch.uint8, device=device) if rank != 0 else None
        self.recv = None
        if rank == 0 and world > 1:
            self.recv = torch.zeros((world - 1, n, pitch), dtype=torch.uint8, device=device)
            src, dst = [], []
            for r in range(1, world):
                k = len(self.rows[r])
                src.append(torch.arange(k, device=device) + (r - 1) * n)
                dst.append(self.rows[r])
            self.src_idx = torch.cat(src) if src else None
            self.dst_rows = torch.cat(dst) if dst else None

    def gather(self, frame):
        """After every rank rendered its own bands into `frame`, make rank 0's frame complete."""
        if self.world == 1:
            return frame
        if self.rank != 0:
            mine = self.rows[self.rank]
            torch.index_select(frame, 0, mine, out=self.send[: len(mine)])
            dist.send(self.send, dst=0, group=self.group)
            return frame
        ops = [dist.P2POp(dist.irecv, self.recv[r - 1], r, self.group) for r in range(1, self.world)]
        for req in dist.batch_isend_irecv(ops):
            req.wait()
        frame.index_copy_(0, self.dst_rows, self.recv.view(-1, self.pitch).index_select(0, self.src_idx))
        return frame


class _DevicePointer:
    """Minimal __cuda_array_interface__ carrier: lets torch view memory this package allocated (rtb_frame_alloc_shared)."""

    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "|u1", "data": (int(ptr), False), "version": 3, "strides": None}


class PeerFrame:
    """The per-frame exchange over NVLink peer memory instead of a collective (one process per GPU, one NVSwitch node).

    Rank 0 owns the complete frame in memory allocated by the library (`Scene.shared_alloc`) and broadcasts its CUDA IPC
    handle once; every other rank maps it.  Per frame a rank r > 0 either
      * mode "direct": renders its bands straight into the mapped frame -- the kernels' pixel stores are the transfer,
        nothing is packed, copied or unpacked afterwards; or
      * mode "push":   renders into a local frame and enqueues ONE strided peer copy of the bands it owns
        (`rtb_push_bands`, a cudaMemcpy2DAsync over NVLink) behind the render kernel.
    Completion, stream ordered with no host sync, is either
      * signal "peer": every rank r > 0 stores the frame number into its own 32-bit flag word behind the shared frame
        (`rtb_frame_signal`: system-scope fence + store, behind its render / push) and rank 0 enqueues a kernel that waits
        until all world-1 flags reached it (`rtb_frame_wait`) -- no collective on the frame's path; or
      * signal "nccl": a 4-byte all-reduce on the same stream (what the two-processes-on-one-GPU tests must use: a
        spinning wait kernel and the kernel it waits for cannot share a GPU).
    `target` is the device pointer this rank passes to `Camera.render_scene`; `frame` (rank 0) is the complete image."""

    def __init__(self, scene, height, pitch, rank, world, device, mode="direct", group=None, signal="nccl"):
        assert mode in ("direct", "push") and signal in ("peer", "nccl")
        self.scene, self.h, self.pitch, self.rank, self.world, self.mode, self.group = scene, height, pitch, rank, world, mode, group
        self.signal, self.frame_no = signal, 0
        self.shared = self.peer = None
        self.flag = torch.zeros(1, dtype=torch.int32, device=device)
        self.flags_off = ((height * pitch + 255) // 256) * 256          # one 32-bit flag word per rank, behind the frame
        box = [None]
        if rank == 0:
            self.shared, handle = scene.shared_alloc(self.flags_off + 4 * world + 256)
            torch.as_tensor(_DevicePointer(self.shared + self.flags_off, (4 * world,)), device=device).zero_()
            torch.cuda.synchronize(device)
            box[0] = handle
        if world > 1:
            dist.broadcast_object_list(box, src=0, group=group)
        if rank == 0:
            self.frame = torch.as_tensor(_DevicePointer(self.shared, (height, pitch)), device=device)
            self.target = self.shared
        else:
            self.peer = scene.shared_open(box[0])
            self.frame = torch.zeros((height, pitch), dtype=torch.uint8, device=device) if mode == "push" else None
            self.target = self.frame.data_ptr() if mode == "push" else self.peer

    def gather(self, stream=None):
        """Call after this rank's render was enqueued on `stream` (a cudaStream_t value or None for the current torch stream)."""
        if self.world == 1:
            return self.frame
        if self.rank != 0 and self.mode == "push":
            s = stream if stream is not None else torch.cuda.current_stream().cuda_stream
            self.scene.push_bands(self.frame.data_ptr(), self.peer, self.pitch, self.h, self.rank, self.world, s)
        if self.signal == "nccl":
            dist.all_reduce(self.flag, group=self.group)
            return self.frame
        s = stream if stream is not None else torch.cuda.current_stream().cuda_stream
        self.frame_no += 1
        if self.rank == 0:
            self.scene.frame_wait(self.shared + self.flags_off, 1, self.world - 1, self.frame_no, s)
        else:
            self.scene.frame_signal(self.peer + self.flags_off, self.rank, self.frame_no, s)
        return self.frame

    def close(self):
        if self.peer:
            self.scene.shared_close(self.peer)
            self.peer = None
        if self.shared:
            self.frame = None
            self.scene.shared_free(self.shared)
            self.shared = None
