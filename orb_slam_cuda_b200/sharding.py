"""Multi-GPU partitioning of the two offline workloads (the reference itself is single-GPU).

  * batched extraction: frames are independent units (ORBextractor::operator() keeps no cross-frame state,
    reference src/ORBextractor.cc:1813-1814 are timing counters only) -> contiguous frame blocks per rank,
    NO collective on the data path;
  * all-vs-all keyframe matching: the keyframe descriptors are produced frame-sharded, every rank all-gathers the
    database (NCCL over NVLink on the GPU box, gloo in the CPU tests) and scores its own block of query keyframes
    against all keyframes.  The gather is chunked so that matching against chunk c overlaps the transfer of chunk
    c + 1 (the collective runs on torch.distributed's own stream, the matcher on the caller's).

One process per GPU; torch.distributed is plumbing only — the scoring itself is the orbx_match_allpairs_device
kernel, injected here as `score_fn` so the host logic can be tested on CPU with gloo (tests/test_sharding_cpu.py).
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple

import torch
import torch.distributed as dist


def block_range(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block [begin, end) of n units owned by `rank`: sizes differ by at most one, earlier ranks larger."""
    if n < 0 or world < 1 or not (0 <= rank < world):
        raise ValueError(f"block_range({n}, {rank}, {world})")
    base, rem = divmod(n, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def owner_of(unit: int, n: int, world: int) -> int:
    """Inverse of block_range: the rank that owns `unit`."""
    if not (0 <= unit < n):
        raise ValueError(f"unit {unit} outside [0, {n})")
    base, rem = divmod(n, world)
    cut = rem * (base + 1)
    return unit // (base + 1) if unit < cut else rem + (unit - cut) // max(base, 1)


def chunk_plan(k_local_max: int, chunks: int):
    """Splits the per-rank keyframe block [0, k_local_max) into `chunks` pieces (the last ones may be empty)."""
    chunks = max(1, min(chunks, max(k_local_max, 1)))
    step = -(-k_local_max // chunks)
    return [(c * step, min((c + 1) * step, k_local_max)) for c in range(chunks) if c * step < k_local_max]


def allpairs_sharded(local_desc: torch.Tensor, k_total: int,
                     score_fn: Callable[[torch.Tensor, torch.Tensor, torch.Tensor], None],
                     group: Optional[dist.ProcessGroup] = None, chunks: int = 4):
    """All-vs-all keyframe scoring, sharded by query block.

    local_desc : uint8 [k_local, n, 32] — descriptors of this rank's keyframes, block_range(k_total, rank, world)
    score_fn(q, db, out) : fills out[kq, kd] (int32) with the score of every query keyframe in q [kq, n, 32]
                           against every database keyframe in db [kd, n, 32]; must be stream-ordered with torch's
                           current stream (the product passes a wrapper around ORBmatcher.match_allpairs_device)
    returns (scores int32 [k_local, k_total] for this rank's query keyframes, gathered-bytes count)

    Database order: global keyframe index.  Ranks may own unequal blocks (k_total not divisible by world): every
    chunk is padded to the largest block so the all-gather is regular; padded keyframes are never scored."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    begin, end = block_range(k_total, rank, world)
    k_local = end - begin
    if local_desc.shape[0] != k_local:
        raise ValueError(f"rank {rank} owns keyframes [{begin}, {end}) but got {local_desc.shape[0]} descriptor sets")
    n = local_desc.shape[1]
    scores = torch.zeros((k_local, k_total), dtype=torch.int32, device=local_desc.device)
    if world == 1:
        score_fn(local_desc, local_desc, scores)
        return scores, 0
    k_max = block_range(k_total, 0, world)[1]  # rank 0 owns the largest block
    plan = chunk_plan(k_max, chunks)
    gathered = []
    works = []
    moved = 0
    for (c0, c1) in plan:  # issue every chunk's all-gather up front; they complete in order on the comm stream
        send = torch.zeros((c1 - c0, n, 32), dtype=torch.uint8, device=local_desc.device)
        hi = min(c1, k_local)
        if hi > c0:
            send[: hi - c0] = local_desc[c0:hi]
        recv = torch.empty((world, c1 - c0, n, 32), dtype=torch.uint8, device=local_desc.device)
        works.append(dist.all_gather_into_tensor(recv.view(-1), send.view(-1), group=group, async_op=True))
        gathered.append(recv)
        moved += recv.numel()
    for (c0, c1), recv, work in zip(plan, gathered, works):
        work.wait()  # stream-orders the consumer after this chunk only; later chunks are still in flight
        if k_local == 0:
            continue
        # ONE scoring launch per chunk over the keyframes of every source rank (world x chunk database keyframes): a launch per
        # (chunk, source) would be a grid of k_local x chunk CTAs — a fraction of the GPU at 8 ranks.  Padded keyframes are
        # scored and dropped by the column selection below.
        width = c1 - c0
        sel, cols = [], []
        for src in range(world):
            sb, se = block_range(k_total, src, world)
            valid = min(c1, se - sb) - c0
            if valid > 0:
                sel.extend(range(src * width, src * width + valid))
                cols.extend(range(sb + c0, sb + c0 + valid))
        if not cols:
            continue
        tmp = torch.empty((k_local, world * width), dtype=torch.int32, device=local_desc.device)
        score_fn(local_desc, recv.view(world * width, n, 32), tmp)
        if len(sel) == world * width and cols == list(range(cols[0], cols[0] + len(cols))):
            scores[:, cols[0]: cols[0] + len(cols)] = tmp
        else:
            scores[:, torch.tensor(cols, device=local_desc.device)] = tmp[:, torch.tensor(sel, device=local_desc.device)]
    return scores, moved


class NcclComm:
    """An ncclComm_t of this process group's ranks for the native entry point orbx_allpairs_sharded (the C ABI takes the
    communicator as a plain pointer; torch.distributed keeps its own private).  Created through ctypes on the NCCL library the
    process has loaded already — the same one liborbx.so resolves at run time — with the unique id broadcast over
    torch.distributed.  One process per GPU; plumbing only."""

    def __init__(self, device: torch.device, group: Optional[dist.ProcessGroup] = None):
        import ctypes as C
        self._C = C
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.comm = C.c_void_p()
        self._lib = None
        if self.world == 1:
            return

        class UniqueId(C.Structure):
            _fields_ = [("internal", C.c_byte * 128)]

        self._lib = C.CDLL("libnccl.so.2")  # resolves to the already loaded copy (torch's)
        uid = UniqueId()
        if self.rank == 0:
            self._check(self._lib.ncclGetUniqueId(C.byref(uid)))
        t = torch.frombuffer(bytearray(bytes(uid)), dtype=torch.uint8).to(device)
        dist.broadcast(t, src=0, group=group)
        C.memmove(C.byref(uid), bytes(t.cpu().numpy().tobytes()), 128)
        self._lib.ncclCommInitRank.argtypes = [C.POINTER(C.c_void_p), C.c_int, UniqueId, C.c_int]
        torch.cuda.set_device(device)
        self._check(self._lib.ncclCommInitRank(C.byref(self.comm), self.world, uid, self.rank))

    def _check(self, rc):
        if rc != 0:
            self._lib.ncclGetErrorString.restype = self._C.c_char_p
            raise RuntimeError("NCCL: " + self._lib.ncclGetErrorString(rc).decode())

    def destroy(self):
        if self._lib is not None and self.comm:
            self._lib.ncclCommDestroy.argtypes = [self._C.c_void_p]
            self._lib.ncclCommDestroy(self.comm)
            self.comm = self._C.c_void_p()


def allpairs_sharded_native(comm: NcclComm, local_desc: torch.Tensor, k_total: int, th: int = 50, ratio: float = 0.6, chunks: int = 4,
                            stream: int = 0, scores: Optional[torch.Tensor] = None):
    """orbx_allpairs_sharded (include/orbx.h): the whole sharded workload — own block first, chunked ncclAllGather on an internal
    stream, one scoring launch per landed chunk — inside liborbx.so.  Returns (scores int32 [k_local, k_total], gathered bytes);
    asynchronous on `stream` (0 = the library's own stream: synchronise the device before reading)."""
    import ctypes as C
    from . import lib, _check, _p
    begin, end = block_range(k_total, comm.rank, comm.world)
    if local_desc.shape[0] != end - begin:
        raise ValueError(f"rank {comm.rank} owns keyframes [{begin}, {end}) but got {local_desc.shape[0]} descriptor sets")
    n = local_desc.shape[1]
    if scores is None:
        scores = torch.empty((end - begin, k_total), dtype=torch.int32, device=local_desc.device)
    moved = C.c_ulonglong(0)
    f = lib().orbx_allpairs_sharded
    f.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_int, C.c_void_p,
                  C.POINTER(C.c_ulonglong)]
    _check(f(comm.comm, comm.rank, comm.world, _p(local_desc.data_ptr()), k_total, n, th, ratio, _p(scores.data_ptr()), chunks,
             _p(stream) if stream else None, C.byref(moved)))
    return scores, int(moved.value)


def frames_for_rank(n_frames: int, rank: int, world: int) -> range:
    """Frame indices of an offline sequence extracted by `rank` (contiguous block, no collective)."""
    b, e = block_range(n_frames, rank, world)
    return range(b, e)
