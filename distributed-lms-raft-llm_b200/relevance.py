"""Corpus relevance search (BASELINE configs[4]): the reference compares a query with ONE assignment text
(REF/lms_server.py:1254-1266); this widens the same cosine to a corpus of assignment-chunk embeddings, sharded
by rows across GPUs.  Every rank scans its shard (fused cosine + top-k kernels), the per-rank candidates are
exchanged with one all-gather (NCCL over NVLink on GPUs; gloo in the CPU tests) and merged on every rank.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist

from . import _lib


def shard_bounds(n_rows: int, world_size: int, rank: int):
    """Contiguous row shards, sizes differing by at most one (first shards take the remainder)."""
    base, rem = divmod(n_rows, world_size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def exchange_candidates(scores, idx, group=None):
    """The path's one exchange step: all-gather every rank's [Q,k] (score, global row) candidates and lay
    them out as [Q, world*k] in rank order, i.e. ascending row order among equal scores.  Works on CUDA
    tensors (NCCL) and on host tensors (gloo, used by the CPU tests of this protocol)."""
    ws = dist.get_world_size(group)
    Q, k = scores.shape
    if scores.is_cuda:
        all_s = torch.empty((ws, Q, k), dtype=scores.dtype, device=scores.device)
        all_i = torch.empty((ws, Q, k), dtype=idx.dtype, device=idx.device)
        dist.all_gather_into_tensor(all_s, scores.contiguous(), group=group)
        dist.all_gather_into_tensor(all_i, idx.contiguous(), group=group)
    else:
        ls = [torch.empty_like(scores) for _ in range(ws)]
        li = [torch.empty_like(idx) for _ in range(ws)]
        dist.all_gather(ls, scores.contiguous(), group=group)
        dist.all_gather(li, idx.contiguous(), group=group)
        all_s, all_i = torch.stack(ls), torch.stack(li)
    return (all_s.permute(1, 0, 2).reshape(Q, ws * k).contiguous(),
            all_i.permute(1, 0, 2).reshape(Q, ws * k).contiguous())


class _Exchange:
    """This process's NCCL communicator for the candidate exchange, owned by libtutor_b200.so (tutor_nccl_*)."""
    comm, world, rank = None, 1, 0


def _nccl_library_path():
    """The NCCL PyTorch ships (site-packages/nvidia/nccl/lib/libnccl.so.2); None lets the loader search."""
    try:
        import nvidia.nccl
        for d in list(nvidia.nccl.__path__):
            p = os.path.join(d, "lib", "libnccl.so.2")
            if os.path.exists(p):
                return p
    except ImportError:
        pass
    return None


def init_exchange(group=None):
    """Create the library's own NCCL communicator over the ranks of `group` (default: the world).  The unique id is made
    on rank 0 and broadcast through torch.distributed; after this, CorpusShard.search runs its exchange as ONE
    ncclAllGather issued by the C library on the caller's stream (tutor_topk_merge_allgather) instead of two
    torch.distributed all-gathers and two layout copies issued from Python."""
    import ctypes as C
    lib = _lib.load()
    ws, rank = dist.get_world_size(group), dist.get_rank(group)
    path = _nccl_library_path()
    _lib.check(lib.tutor_nccl_load(path.encode() if path else None))
    uid = C.create_string_buffer(128)
    if rank == 0:
        _lib.check(lib.tutor_nccl_unique_id(uid))
    box = [uid.raw]
    dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    uid = C.create_string_buffer(box[0], 128)
    comm = C.c_void_p()
    _lib.check(lib.tutor_nccl_comm_create(uid, ws, rank, C.byref(comm)))
    _Exchange.comm, _Exchange.world, _Exchange.rank = comm, ws, rank
    return comm


def close_exchange():
    """Destroy the communicator.  Every rank must call this together, and any CUDA graph that captured a search through it
    must have been deleted first: NCCL keeps a communicator alive while a captured graph still refers to it, and
    ncclCommDestroy then waits for ever."""
    if _Exchange.comm is not None:
        _lib.load().tutor_nccl_comm_destroy(_Exchange.comm)
        _Exchange.comm, _Exchange.world, _Exchange.rank = None, 1, 0


class CorpusShard:
    """One rank's rows [row_lo, row_hi) of the corpus, resident in HBM with their inverse norms."""

    def __init__(self, rows: torch.Tensor, row_lo: int = 0):
        if not rows.is_cuda or rows.dtype != torch.float32 or rows.dim() != 2:
            raise ValueError("corpus shard must be a 2-D fp32 CUDA tensor")
        self.lib = _lib.load()
        self.rows = rows.contiguous()
        self.row_lo = int(row_lo)
        self.inv_norm = torch.empty((rows.shape[0],), dtype=torch.float32, device=rows.device)
        self._scratch = None
        self._xscratch = None
        _lib.check(self.lib.tutor_row_inv_norm(self.rows.data_ptr(), self.inv_norm.data_ptr(), rows.shape[0],
                                               rows.shape[1], torch.cuda.current_stream(rows.device).cuda_stream))

    def topk(self, queries: torch.Tensor, k: int):
        """queries [Q,d] fp32 CUDA -> (scores [Q,k], global row indices [Q,k] int32) for this shard."""
        q = queries.to(self.rows.device, torch.float32).contiguous()
        Q, d = q.shape
        N = self.rows.shape[0]
        need = self.lib.tutor_cosine_topk_scratch_bytes(Q, N, k)
        if self._scratch is None or self._scratch.numel() < need:
            self._scratch = torch.empty((need,), dtype=torch.uint8, device=q.device)
        scores = torch.empty((Q, k), dtype=torch.float32, device=q.device)
        idx = torch.empty((Q, k), dtype=torch.int32, device=q.device)
        _lib.check(self.lib.tutor_cosine_topk(q.data_ptr(), self.rows.data_ptr(), self.inv_norm.data_ptr(), Q, N, d, k,
                                              scores.data_ptr(), idx.data_ptr(), self.row_lo,
                                              self._scratch.data_ptr(), torch.cuda.current_stream(q.device).cuda_stream))
        return scores, idx

    def search(self, queries: torch.Tensor, k: int, group=None):
        """Global top-k over every rank's shard: local top-k, all-gather of [Q,k] candidates, device merge."""
        scores, idx = self.topk(queries, k)
        if _Exchange.comm is not None and _Exchange.world > 1 and scores.is_cuda:
            # the C library's own exchange: pack -> one ncclAllGather -> merge, all on the current stream
            ws, Q = _Exchange.world, scores.shape[0]
            need = self.lib.tutor_topk_merge_allgather_scratch_bytes(Q, k, ws)
            if self._xscratch is None or self._xscratch.numel() < need:
                self._xscratch = torch.empty((need,), dtype=torch.uint8, device=scores.device)
            out_s, out_i = torch.empty_like(scores), torch.empty_like(idx)
            _lib.check(self.lib.tutor_topk_merge_allgather(_Exchange.comm, scores.data_ptr(), idx.data_ptr(), Q, k, ws,
                                                           self._xscratch.data_ptr(), out_s.data_ptr(), out_i.data_ptr(),
                                                           torch.cuda.current_stream(scores.device).cuda_stream))
            return out_s, out_i
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return scores, idx
        ws = dist.get_world_size(group)
        Q = scores.shape[0]
        cand_s, cand_i = exchange_candidates(scores, idx, group)
        out_s, out_i = torch.empty_like(scores), torch.empty_like(idx)
        _lib.check(self.lib.tutor_topk_merge(cand_s.data_ptr(), cand_i.data_ptr(), Q, ws * k, k, out_s.data_ptr(),
                                             out_i.data_ptr(), torch.cuda.current_stream(scores.device).cuda_stream))
        return out_s, out_i
