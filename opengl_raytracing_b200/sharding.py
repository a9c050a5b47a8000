"""Multi-GPU sharding of one frame: one process per GPU, torch.distributed for
the plumbing. The path shards with NO data-path collective for row tiles (pixels are
independent) and ONE real exchange for sample ranges (a sum of per-rank partial colours):

  * row tiles      rank g renders rows [g*H/G, (g+1)*H/G) (``row_range``); the tiles are
                   gathered to rank 0 (``gather_rows``), or written straight into rank 0's
                   frame buffer when the output pointers are peer-mapped (rt_render takes
                   any device address).
  * sample ranges  rank g renders frame_count in [g*S/G, (g+1)*S/G) into an FP32 partial sum
                   (RT_FLAG_ACCUMULATE), ``reduce_samples`` sums the partials on rank 0 and
                   scales by 1/S. Position/normal come from the rank that owns sample 0.

The host logic is backend-agnostic (gloo on CPU in tests/test_sharding.py, nccl on GPUs).
"""
from __future__ import annotations

import os


def row_range(height: int, rank: int, world: int, align: int = 4) -> tuple:
    """Rows of ``rank``: contiguous, 4-aligned boundaries (the kernel's tile height) so that
    no warp tile straddles two ranks; the last rank takes the remainder."""
    per = (height + world - 1) // world
    per = (per + align - 1) // align * align
    a = min(rank * per, height)
    b = min(a + per, height)
    return a, b


def sample_range(num_samples: int, rank: int, world: int) -> tuple:
    """[first, first+count) of the frame_count indices rendered by ``rank``."""
    base, rem = divmod(num_samples, world)
    first = rank * base + min(rank, rem)
    return first, base + (1 if rank < rem else 0)


def gather_rows(buf, height: int, world: int, dst: int = 0, group=None):
    """Gather the row tiles of a full-frame (H, W, C) tensor onto ``dst``: every rank sends its
    own rows; ``dst`` receives them in place. Equal-size tiles use one all_gather-free
    gather; ragged last tiles go point to point."""
    import torch.distributed as dist
    rank = dist.get_rank(group)
    if world == 1:
        return buf
    ops = []
    if rank == dst:
        for r in range(world):
            if r == dst:
                continue
            a, b = row_range(height, r, world)
            if b > a:
                ops.append(dist.P2POp(dist.irecv, buf[a:b], r, group))
    else:
        a, b = row_range(height, rank, world)
        if b > a:
            ops.append(dist.P2POp(dist.isend, buf[a:b], dst, group))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    return buf


def reduce_samples(partial_sum, total_samples: int, dst: int = 0, group=None):
    """Sum the per-rank partial colour sums on ``dst`` and turn them into the mean (alpha
    restored to 1). One collective: the path's only real exchange step."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(partial_sum, dst=dst, op=dist.ReduceOp.SUM, group=group)
        if dist.get_rank(group) != dst:
            return partial_sum
    partial_sum[..., :3] *= 1.0 / float(total_samples)
    partial_sum[..., 3] = 1.0
    return partial_sum


# --------------------------------------------------------------------------- fused gather over NVLink
BAND_ROWS = 32      # interleaved row bands: sky rows cost one ray per pixel, object rows dozens


def band_plan(rank: int, world: int, band_rows: int = BAND_ROWS) -> dict:
    """rt_frame_params fields for rank ``rank`` of ``world`` under interleaved row bands: band k of
    the frame (``band_rows`` rows each) belongs to rank k % world."""
    return {"row_band": band_rows, "row_period": world, "row_phase": rank}


class _DevicePtr:
    """Minimal __cuda_array_interface__ view of a raw device allocation (for torch.as_tensor)."""

    def __init__(self, ptr: int, shape: tuple, typestr: str):
        self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (ptr, False), "version": 3}


class PeerFrame:
    """The three output images of one frame, allocated on the root rank's GPU and mapped into every
    other rank's address space through CUDA IPC (rt_device_alloc / rt_ipc_export / rt_ipc_open).
    Each rank passes ``color`` / ``position`` / ``normal`` (plain device addresses) to rt_render:
    the kernel's float4 stores then go straight into the root's frame over NVLink — the gather
    disappears into the kernel epilogue (SURVEY.md 8e "fused form"). ``tensors()`` on the root
    returns torch views of the frame."""

    def __init__(self, ctx, height: int, width: int, root: int = 0, group=None):
        import ctypes as C
        import torch
        import torch.distributed as dist
        self.ctx, self.h, self.w, self.root = ctx, height, width, root
        self.rank = dist.get_rank(group)
        self.is_root = self.rank == root
        sizes = {"color": height * width * 16, "position": height * width * 16, "normal": height * width * 8}
        # the 64-byte IPC handles travel through the process group (CPU tensor under gloo)
        dev = torch.device("cpu") if dist.get_backend(group) == "gloo" else torch.device("cuda", ctx.device)
        handles = torch.zeros(3, 64, dtype=torch.uint8, device=dev)
        self.ptr = {}
        if self.is_root:
            hs = []
            for name, nbytes in sizes.items():
                p = C.c_void_p(0)
                ctx._check(ctx.lib.rt_device_alloc(ctx.handle, C.byref(p), nbytes))
                buf = C.create_string_buffer(64)
                ctx._check(ctx.lib.rt_ipc_export(ctx.handle, p, buf))
                self.ptr[name] = p.value
                hs.append(list(buf.raw))
            handles.copy_(torch.tensor(hs, dtype=torch.uint8))
        dist.broadcast(handles, src=root, group=group)
        if not self.is_root:
            hb = handles.cpu().numpy()
            for i, name in enumerate(sizes):
                p = C.c_void_p(0)
                ctx._check(ctx.lib.rt_ipc_open(ctx.handle, hb[i].tobytes(), C.byref(p)))
                self.ptr[name] = p.value
        self.color, self.position, self.normal = (C.c_void_p(self.ptr[k]) for k in ("color", "position", "normal"))

    def tensors(self):
        import torch
        assert self.is_root
        c = torch.as_tensor(_DevicePtr(self.ptr["color"], (self.h, self.w, 4), "<f4"), device=f"cuda:{self.ctx.device}")
        p = torch.as_tensor(_DevicePtr(self.ptr["position"], (self.h, self.w, 4), "<f4"), device=f"cuda:{self.ctx.device}")
        n = torch.as_tensor(_DevicePtr(self.ptr["normal"], (self.h, self.w, 4), "<f2"), device=f"cuda:{self.ctx.device}")
        return c, p, n

    def close(self):
        for name, p in self.ptr.items():
            if self.is_root:
                self.ctx.lib.rt_device_free(self.ctx.handle, p)
            else:
                self.ctx.lib.rt_ipc_close(self.ctx.handle, p)
        self.ptr = {}


class PeerUnavailable(RuntimeError):
    """The ranks cannot map each other's memory (no CUDA IPC here); raised on EVERY rank together."""


class PeerReducer:
    """The sample-range exchange step without NCCL (one process per GPU): every rank's partial-sum frame, rank 0's
    result frame and a block of flag words per rank are CUDA-IPC mapped into every rank; ``reduce`` enqueues ONE
    kernel (rt_peer_reduce, include/rt_multi.h) that signals "my partial is complete", waits for every peer, sums
    this rank's slice of all partials in rank order (deterministic), scales, and stores it into rank 0's frame.
    THREE partial buffers rotate between consecutive reduces: render into ``partial`` (the buffer of the NEXT
    reduce), then call ``reduce``. ``result`` on rank 0 = the finished mean frame.

    Why three: a reduce kernel passes its entry wait only once every peer has STARTED the reduce of the same epoch,
    i.e. (stream order on the peer) finished the one before. So when this rank's reduce of epoch j has completed,
    every peer is done reading this rank's partial of epoch j-1 — the buffer the render of epoch j+2 writes. The
    caller may therefore overlap the reduce of frame j with the render of frame j+1 on another stream (bench.py's
    pipelined mode), as long as the render of frame j+2 is ordered behind this rank's reduce of frame j (an event).
    With the render and the reduce on ONE stream two buffers would do; the third is what makes the overlap safe."""

    NBUF = 3

    FLAG_BYTES = 2 << 20          # its own 2 MB allocation: CUDA IPC handles map whole allocations

    def __init__(self, ctx, height: int, width: int, group=None):
        import ctypes as C
        import torch
        import torch.distributed as dist
        from . import abi
        self.ctx, self.h, self.w, self.group = ctx, height, width, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        assert self.world <= abi.RT_MAX_PEERS
        self.epoch = 0
        nbytes = height * width * 16
        names = ["partial%d" % i for i in range(self.NBUF)] + ["flags"] + (["result"] if self.rank == 0 else [])
        self.own, handles = {}, torch.zeros(self.NBUF + 2, 64, dtype=torch.uint8)
        self.maps, self._opened = [], []
        dev = torch.device("cpu") if dist.get_backend(group) == "gloo" else torch.device("cuda", ctx.device)
        torch.cuda.set_device(ctx.device)

        def agree(ok: bool, what: str, err):
            """Every rank learns whether EVERY rank got through the step (no rank may wait in a collective for one that
            raised): all raise PeerUnavailable together, after freeing what they hold."""
            flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
            if int(flag.item()) == 0:
                self.close()
                raise PeerUnavailable(f"{what} failed on {'this rank: ' + str(err) if err else 'another rank'}")

        err = None
        try:
            if os.environ.get("RT_PEER_FORCE_FAIL") == "export":      # dev knob: exercises the fallback of the callers
                raise RuntimeError("RT_PEER_FORCE_FAIL=export")
            for i, name in enumerate(names):
                p = C.c_void_p(0)
                ctx._check(ctx.lib.rt_device_alloc(ctx.handle, C.byref(p), self.FLAG_BYTES if name == "flags" else nbytes))
                self.own[name] = p.value
                buf = C.create_string_buffer(64)
                ctx._check(ctx.lib.rt_ipc_export(ctx.handle, p, buf))
                handles[i] = torch.tensor(list(buf.raw), dtype=torch.uint8)
            zero = torch.as_tensor(_DevicePtr(self.own["flags"], (self.FLAG_BYTES // 4,), "<u4"), device=f"cuda:{ctx.device}")
            zero.view(torch.int32).zero_()
            torch.cuda.synchronize(ctx.device)
        except Exception as e:                                        # noqa: BLE001
            err = e
        agree(err is None, "allocating / exporting the peer buffers (CUDA IPC)", err)
        mine = handles.to(dev)
        everyone = [torch.zeros_like(mine) for _ in range(self.world)]
        dist.all_gather(everyone, mine, group=group)
        try:
            if os.environ.get("RT_PEER_FORCE_FAIL") == "open" and self.rank == self.world - 1:
                raise RuntimeError("RT_PEER_FORCE_FAIL=open")
            for r in range(self.world):
                if r == self.rank:
                    self.maps.append(dict(self.own))
                    continue
                hb = everyone[r].cpu().numpy()
                m = {}
                for i, name in enumerate(["partial%d" % k for k in range(self.NBUF)] + ["flags"] + (["result"] if r == 0 else [])):
                    p = C.c_void_p(0)
                    ctx._check(ctx.lib.rt_ipc_open(ctx.handle, hb[i].tobytes(), C.byref(p)))
                    m[name] = p.value
                    self._opened.append(p.value)
                self.maps.append(m)
        except Exception as e:                                        # noqa: BLE001
            err = e
        agree(err is None, "mapping the peers' buffers (CUDA IPC)", err)   # also: every flag block is zeroed and mapped before the first epoch

    def _tensor(self, ptr):
        import torch
        return torch.as_tensor(_DevicePtr(ptr, (self.h, self.w, 4), "<f4"), device=f"cuda:{self.ctx.device}")

    @property
    def partial(self):
        """The buffer this rank renders its partial sum of the NEXT reduce into (a torch view)."""
        return self._tensor(self.own["partial%d" % ((self.epoch + 1) % self.NBUF)])

    @property
    def result(self):
        assert self.rank == 0
        return self._tensor(self.own["result"])

    def reduce(self, stream: int, total_samples: int) -> None:
        import ctypes as C
        from . import abi
        self.epoch += 1
        d = abi.PeerReduceDesc()
        d.rank, d.world = self.rank, self.world
        for r in range(self.world):
            d.partial[r] = self.maps[r]["partial%d" % (self.epoch % self.NBUF)]
            d.flags[r] = self.maps[r]["flags"]
        d.result = self.maps[0]["result"]
        d.pixels, d.scale, d.epoch = self.h * self.w, 1.0 / float(total_samples), self.epoch
        self.ctx._check(self.ctx.lib.rt_peer_reduce(self.ctx.handle, C.byref(d), C.c_void_p(stream)))

    def check(self) -> None:
        """After a synchronise: raises if a wait inside an earlier reduce kernel timed out."""
        import ctypes as C
        self.ctx._check(self.ctx.lib.rt_peer_reduce_status(self.ctx.handle, C.c_void_p(self.own["flags"])))

    def close(self):
        for p in self._opened:
            self.ctx.lib.rt_ipc_close(self.ctx.handle, p)
        for p in self.own.values():
            self.ctx.lib.rt_device_free(self.ctx.handle, p)
        self._opened, self.own = [], {}
