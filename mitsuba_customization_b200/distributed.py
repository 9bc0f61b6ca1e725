"""Multi-GPU host logic for the measured-BRDF path: one process per GPU.

Forward eval / pdf / sample shard by query batch with the table replicated and NO collective.
The backward keeps a private fp32 gradient table per GPU and sums them with one allreduce over
NVLink (libmerl_b200's merl_allreduce_sum_f32, i.e. NCCL bound at run time); a renderer sums its
film the same way.  torch.distributed is used only as the out-of-band channel that ships the
128-byte NCCL id from rank 0 to the others (any backend: gloo on CPU, nccl on the GPU box).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import torch

from . import _cabi
from ._cabi import MerlError, check

ID_BYTES = 128


def shard_bounds(n: int, world: int, rank: int, align: int = 4) -> Tuple[int, int]:
    """Contiguous shard [lo, hi) of n queries for `rank`; interior cuts are multiples of `align`
    so every shard but the last keeps the 16-byte vector path.  Shards cover [0, n) exactly."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"rank {rank} not in [0, {world})")
    per = -(-n // world)              # ceil
    per = -(-per // align) * align    # round up to the alignment
    lo = min(rank * per, n)
    hi = min(lo + per, n)
    return lo, hi


def table_owner(table_id: int, world: int) -> int:
    """By-table sharding for multi-material scenes (BASELINE config 5): round-robin owner.  The same
    rule as merl_render's by-table variant (merl_render_params_t.table_shard_*: table m is evaluated
    by rank m % world), so a host can tell which tables a rank must hold."""
    return table_id % world


def exchange_id(make_id, group=None) -> bytes:
    """Rank 0 calls make_id() -> 128 bytes; everyone returns the same bytes.  Runs over any
    initialised torch.distributed backend (tensors live on the backend's device)."""
    import torch.distributed as dist

    if not dist.is_initialized():
        raise MerlError("torch.distributed is not initialised (needed to ship the communicator id)")
    backend = dist.get_backend(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    buf = torch.zeros(ID_BYTES, dtype=torch.uint8)
    if dist.get_rank(group) == 0:
        raw = make_id()
        if len(raw) != ID_BYTES:
            raise MerlError(f"communicator id must be {ID_BYTES} bytes")
        buf = torch.frombuffer(bytearray(raw), dtype=torch.uint8).clone()
    buf = buf.to(dev)
    dist.broadcast(buf, src=0, group=group)
    return bytes(buf.cpu().numpy().tobytes())


def _nccl_unique_id() -> bytes:
    raw = (C.c_char * ID_BYTES)()
    check(_cabi.lib().merl_comm_unique_id(C.cast(raw, C.c_void_p)))
    return bytes(raw.raw)


class Comm:
    """libmerl_b200 communicator of this rank (wraps merl_comm_create / merl_allreduce_sum_f32)."""

    def __init__(self, rank: int, world: int, device: int, id_bytes: Optional[bytes] = None):
        self.rank, self.world, self.device = rank, world, device
        if id_bytes is None:
            id_bytes = exchange_id(_nccl_unique_id)
        self._h = C.c_void_p()
        buf = (C.c_char * ID_BYTES).from_buffer_copy(id_bytes)
        check(_cabi.lib().merl_comm_create(C.cast(buf, C.c_void_p), world, rank, device, C.byref(self._h)))

    def allreduce_sum_(self, t: torch.Tensor) -> torch.Tensor:
        if t.dtype != torch.float32 or not t.is_cuda or not t.is_contiguous():
            raise MerlError("allreduce_sum_ needs a contiguous float32 CUDA tensor")
        stream = C.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)
        check(_cabi.lib().merl_allreduce_sum_f32(self._h, C.c_void_p(t.data_ptr()), t.numel(), stream))
        return t

    def close(self):
        h, self._h = self._h, None
        if h:
            check(_cabi.lib().merl_comm_destroy(h))

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PeerBuffer:
    """An exchange buffer of `count` floats on this rank's GPU (merl_peerbuf_*): the ranks of `comm` sum
    it in place with ONE kernel over NVLink peer memory (transport "peer", default), with the library's
    multimem kernel through the NVSwitch ("nvls": in-switch reduction, NCCL >= 2.28 symmetric window), or
    with ncclAllReduce on a buffer registered with NCCL ("nccl").  "auto" takes "nvls" on more than two ranks when every
    rank can, else "peer" (the library gives all ranks the same answer, so the choice is collective).
    Creation and destruction are collective.  `tensor(shape)` is a zero-copy torch view of the
    library-owned memory."""

    def __init__(self, comm: Comm, count: int, transport: str = "peer"):
        self.comm = comm
        self._h = C.c_void_p()
        kinds = {"peer": _cabi.XCHG_PEER, "nccl": _cabi.XCHG_NCCL, "nvls": _cabi.XCHG_NVLS}
        if transport == "auto" and comm.world <= 2:
            transport = "peer"   # two copies: nothing for the switch to save (54 vs 78 us per 23 MB)
        if transport == "auto":
            rc = _cabi.lib().merl_peerbuf_create(comm._h, int(count), _cabi.XCHG_NVLS, C.byref(self._h))
            if rc == _cabi.ERR_UNSUPPORTED:
                transport = "peer"
                check(_cabi.lib().merl_peerbuf_create(comm._h, int(count), _cabi.XCHG_PEER, C.byref(self._h)))
            else:
                check(rc)
                transport = "nvls"
        else:
            check(_cabi.lib().merl_peerbuf_create(comm._h, int(count), kinds[transport], C.byref(self._h)))
        self.info = _cabi.PeerbufInfo()
        check(_cabi.lib().merl_peerbuf_info(self._h, C.byref(self.info)))
        self.count = int(self.info.count)
        self.transport = transport
        self.registration = self.info.registration.decode(errors="replace")
        self.device = torch.device("cuda", comm.device)

    def tensor(self, shape=None) -> torch.Tensor:
        from .bsdf import _DevArray
        shape = (self.count,) if shape is None else tuple(shape)
        n = 1
        for d in shape:
            n *= int(d)
        if n > self.count:
            raise MerlError(f"view of {n} floats exceeds the buffer's {self.count}")
        return torch.as_tensor(_DevArray(self.info.data, shape), device=self.device)

    def allreduce_(self, count: Optional[int] = None, mask_plug=None) -> None:
        """Sum the first `count` floats (default: all) over the ranks, in place, on torch's current stream.
        mask_plug: the plugin whose table gradient this is (count = 4 * ncell): the validity mask of a
        deferred-mask backward is applied inside the exchange."""
        count = self.count if count is None else int(count)
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        check(_cabi.lib().merl_peerbuf_allreduce(self._h, count, mask_plug._h if mask_plug is not None else None, stream))

    def status(self) -> int:
        check(_cabi.lib().merl_peerbuf_info(self._h, C.byref(self.info)))
        return int(self.info.status)

    def close(self):
        h, self._h = self._h, None
        if h:
            check(_cabi.lib().merl_peerbuf_destroy(h))

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def sharded_backward(plug, comm: Optional[Comm], si_wi: torch.Tensor, wo: torch.Tensor, grad_rgb: torch.Tensor,
                     grad_cells: Optional[torch.Tensor] = None, warp_aggregate: bool = False,
                     reduce: bool = True) -> torch.Tensor:
    """This rank's shard of the queries -> private gradient table -> one allreduce.

    si_wi / wo / grad_rgb are THIS RANK's shard, shape (3, n_local).  Returns the (ncell, 4)
    gradient table, summed over ranks when `reduce` (the only collective on the path)."""
    from .bsdf import SurfaceInteraction3f

    g = plug.backward(SurfaceInteraction3f(wi=si_wi), wo, grad_rgb, grad_cells=grad_cells,
                      warp_aggregate=warp_aggregate)
    if reduce and comm is not None and comm.world > 1:
        comm.allreduce_sum_(g)
    return g
