"""NVLink peer-memory all-reduce for the row-parallel linears (Python side).

Counterpart of the reference's vllm/model_executor/parallel_utils/custom_all_reduce.py:26-227
(init / capture / CustomAllreduce) on top of the kernels of csrc/w4a16_allreduce.cu: flag-in-data
packets for decode-sized messages, the one-shot pull up to a few hundred KiB, reduce-scatter +
all-gather (two-shot) up to the registered size.  Differences that matter:

* no buffer registration step at graph capture: inputs are staged into the rank's own
  registered region by the kernel itself, so any tensor can be reduced, eagerly or inside a
  CUDA graph, with the same call;
* the cross-GPU epoch lives in device memory, so captured launches replay correctly;
* one process per GPU under ``torchrun`` (the reference drives its workers through Ray).

``init_custom_all_reduce`` is collective over the tensor-parallel group.  Messages that are
not fp16, not a multiple of 16 bytes or larger than the registered size fall through to the
group's ``torch.distributed`` backend (``communication_op.tensor_model_parallel_all_reduce``).
"""
from __future__ import annotations

import ctypes
from typing import List, Optional

import torch
import torch.distributed as dist

from .. import _lib
from .parallel_state import (get_tensor_model_parallel_group, get_tensor_model_parallel_rank,
                             get_tensor_model_parallel_world_size)

_MAX_WORLD = 8
_FUSED_NORM_MAX_ROWS = 32
_handle: Optional["CustomAllreduce"] = None


class CustomAllreduce:
    """One registered region per rank, mapped by every peer of the tensor-parallel group."""

    def __init__(self, max_bytes: int, fused_out_features: int = 0, ll_bytes: int = 1 << 20) -> None:
        L = _lib.lib()
        self.rank = get_tensor_model_parallel_rank()
        self.world = get_tensor_model_parallel_world_size()
        if not 2 <= self.world <= _MAX_WORLD:
            raise ValueError(f"custom all-reduce supports 2..{_MAX_WORLD} ranks, got {self.world}")
        self.max_bytes = (int(max_bytes) + 255) // 256 * 256
        # Above this the one-shot pull (every rank reads every peer's whole message) loses to
        # NCCL's ring / NVLS on B200 (profiles/r1e_allreduce_norm_bench_tp{2,8}.jsonl: TP = 2
        # wins at 16 KiB and loses at 256 KiB; TP = 8 wins at 256 KiB and loses at 1 MiB) --
        # the same split the reference makes between its one- and two-stage kernels
        # (csrc/custom_all_reduce.cuh:520-531).  A tuning value: tests and benches may raise it.
        self.oneshot_max_bytes = (128 << 10) if self.world <= 2 else (512 << 10)
        # (the r2 sweeps: the packet kernel below beats the one-shot kernel at every size both
        # fit -- 7.5 vs 21 us at 16 KiB, 18 vs 25 us at 256 KiB on 8 ranks -- so the one-shot
        # kernel only serves messages the packet area cannot hold)
        # Reduce-scatter + all-gather (b200_w4a16_allreduce_twoshot) for the largest messages.
        # Measured per size against NCCL 2.28 (profiles/r2p_allreduce_bench_tp2.jsonl,
        # r2r_allreduce_bench_tp{4,8}.jsonl): at 8 ranks it wins from 4 MiB up (49 vs 58 us, 66 vs
        # 76 us at 8 MiB) and loses at 1-2 MiB (40-42 vs 33-36 us: two NVLink barriers); at 4 ranks
        # NCCL wins everywhere above the packet sizes; at two ranks it moves the same bytes as the
        # one-shot kernel behind two barriers (the reference makes the same exception,
        # custom_all_reduce.cuh:520-524).  Tuning values: tests and benches may change them.
        self.use_twoshot = self.world >= 6
        self.twoshot_min_bytes = (2 << 20) + 1
        self.use_ll = True      # flag-in-data packets when the message fits the packet area
        # packets double the NVLink bytes and every rank receives world - 1 copies: measured to
        # win up to 1 MiB at TP = 2 (r1e_allreduce_norm_bench_tp2.jsonl); not yet measured per
        # size at TP = 8, so stay at the size the one-shot kernel was already trusted with there
        self.ll_max_bytes = (1 << 20) if self.world <= 2 else (512 << 10)
        # Row-parallel decode GEMMs: the fused GEMM + all-reduce kernel at every TP.  (Round 1, when
        # its tile exchange still paid a system fence and a flag round trip per tile, GEMM then
        # packet all-reduce won at TP = 2: 7.37 vs 8.26 ms per Llama-2-70B decode step.  With the
        # tiles travelling as packets the fused kernel is ahead there too: 6.15 vs 6.96 ms,
        # profiles/r4j_bench_tp2{,_fused}.json; TP = 4 / 8: 4.46 / 4.26 vs 6.01 / 6.32 ms, r2r / r3a.)
        self.prefer_fused_gemm = True
        self.device = torch.device("cuda", torch.cuda.current_device())
        self.fused_out_features = int(fused_out_features)
        region_bytes = L.b200_w4a16_ar_region_bytes(self.max_bytes, self.world, self.fused_out_features)
        # flag-in-data packet area of the all-reduce + norm kernel: [2][world][2 x ll_bytes]
        self.ll_bytes = (int(ll_bytes) + 255) // 256 * 256
        self.ll_off = (region_bytes + 255) // 256 * 256
        region_bytes = self.ll_off + 4 * self.world * self.ll_bytes
        own = ctypes.c_void_p()
        _lib.check(L.b200_w4a16_ar_alloc(region_bytes, ctypes.byref(own)), "all-reduce region")
        self._own = own.value
        handle = ctypes.create_string_buffer(64)
        _lib.check(L.b200_w4a16_ar_export(self._own, handle), "all-reduce export")
        handles: List[Optional[bytes]] = [None] * self.world
        dist.all_gather_object(handles, handle.raw, group=get_tensor_model_parallel_group())
        self._imported: List[int] = []
        self._regions = (ctypes.c_void_p * self.world)()
        for q, raw in enumerate(handles):
            if q == self.rank:
                self._regions[q] = self._own
                continue
            peer = ctypes.c_void_p()
            _lib.check(L.b200_w4a16_ar_import(ctypes.create_string_buffer(raw, 64), ctypes.byref(peer)),
                       f"all-reduce import of rank {q}")
            self._regions[q] = peer.value
            self._imported.append(peer.value)
        dist.barrier(group=get_tensor_model_parallel_group())   # every region is mapped everywhere

    def should(self, t: torch.Tensor) -> bool:
        nbytes = t.numel() * t.element_size()
        return (t.is_cuda and t.dtype == torch.float16 and t.is_contiguous()
                and nbytes % 16 == 0 and t.data_ptr() % 16 == 0
                and (0 < nbytes <= min(self.max_bytes, self.oneshot_max_bytes) or self._fits_ll(nbytes)
                     or self._fits_twoshot(nbytes)))

    def _fits_ll(self, nbytes: int) -> bool:
        return self.use_ll and 0 < nbytes <= min(self.ll_bytes, self.ll_max_bytes)

    def _fits_twoshot(self, nbytes: int) -> bool:
        return self.use_twoshot and max(self.oneshot_max_bytes + 1, self.twoshot_min_bytes) <= nbytes <= self.max_bytes

    def all_reduce(self, t: torch.Tensor) -> torch.Tensor:
        out = torch.empty_like(t)
        if not self._fits_ll(t.numel() * 2) and self._fits_twoshot(t.numel() * 2):
            st = _lib.lib().b200_w4a16_allreduce_twoshot(
                self._regions, self.rank, self.world, t.data_ptr(), out.data_ptr(), t.numel(),
                self.max_bytes, 1, torch.cuda.current_stream(t.device).cuda_stream)
            _lib.check(st, "custom all-reduce (two-shot)")
            return out
        if self._fits_ll(t.numel() * 2):
            st = _lib.lib().b200_w4a16_allreduce_ll(
                self._regions, self.rank, self.world, t.data_ptr(), out.data_ptr(), t.numel(),
                self.ll_off, self.ll_bytes, 1, torch.cuda.current_stream(t.device).cuda_stream)
            _lib.check(st, "custom all-reduce (packets)")
            return out
        st = _lib.lib().b200_w4a16_allreduce(
            self._regions, self.rank, self.world, t.data_ptr(), out.data_ptr(), t.numel(),
            self.max_bytes, 1, torch.cuda.current_stream(t.device).cuda_stream)
        _lib.check(st, "custom all-reduce")
        return out

    def all_reduce_add_rms_norm(self, partial: torch.Tensor, residual: torch.Tensor,
                                weight: torch.Tensor, eps: float) -> torch.Tensor:
        """``residual += all_reduce(partial)`` (in place) and the RMSNorm of it (returned), one
        kernel: all-reduce (linear.py:541-542) + fused_add_rms_norm (llama.py:205-222)."""
        from .. import ops
        _, rows, hidden = ops._norm_args(partial, weight, residual)    # fp16, contiguous, same shape
        if hidden % 8 or not 0 < hidden <= 16384:
            raise ValueError("fused all-reduce + RMSNorm needs hidden % 8 == 0 and hidden <= 16384")
        if partial.numel() * 2 > self.max_bytes:
            raise ValueError("message larger than the registered peer buffer")
        out = torch.empty_like(partial)
        if rows == 0:
            return out
        if self.use_ll and 0 < partial.numel() * 2 <= self.ll_bytes:
            st = _lib.lib().b200_w4a16_allreduce_add_rms_norm_ll(
                self._regions, self.rank, self.world, partial.data_ptr(), residual.data_ptr(),
                weight.data_ptr(), out.data_ptr(), residual.data_ptr(), float(eps),
                partial.numel() // hidden, hidden, self.ll_off, self.ll_bytes, 1,
                torch.cuda.current_stream(partial.device).cuda_stream)
            _lib.check(st, "all-reduce + add + rms_norm (packets)")
            return out
        st = _lib.lib().b200_w4a16_allreduce_add_rms_norm(
            self._regions, self.rank, self.world, partial.data_ptr(), residual.data_ptr(),
            weight.data_ptr(), out.data_ptr(), residual.data_ptr(), float(eps),
            partial.numel() // hidden, hidden, self.max_bytes, 1,
            torch.cuda.current_stream(partial.device).cuda_stream)
        _lib.check(st, "all-reduce + add + rms_norm")
        return out

    def can_fuse(self, M: int, N: int, G: int) -> bool:
        """Fused row-parallel GEMM + all-reduce: decode batches whose output fits the slots
        (and, unless forced, only when the packet all-reduce cannot take the message)."""
        if not self.prefer_fused_gemm and self._fits_ll(M * N * 2):
            return False
        return 0 < M <= 16 and G % 128 == 0 and N <= self.fused_out_features and N <= 128 * 128

    def gemm_all_reduce(self, x: torch.Tensor, qweight: torch.Tensor, scales: torch.Tensor,
                        qzeros: torch.Tensor, workspace: torch.Tensor, G: int) -> torch.Tensor:
        M, K = x.shape
        N = scales.shape[1]
        out = torch.empty((M, N), dtype=torch.float16, device=x.device)
        st = _lib.lib().b200_w4a16_gemm_allreduce(
            x.data_ptr(), qweight.data_ptr(), scales.data_ptr(), qzeros.data_ptr(), out.data_ptr(),
            M, K, N, G, 0, workspace.data_ptr(), workspace.numel(), self._regions, self.rank,
            self.world, self.max_bytes, self.fused_out_features,
            torch.cuda.current_stream(x.device).cuda_stream)
        _lib.check(st, "fused gemm + all-reduce")
        return out

    def close(self) -> None:
        L = _lib.lib()
        torch.cuda.synchronize(self.device)
        dist.barrier(group=get_tensor_model_parallel_group())   # nobody still reads our region
        for ptr in self._imported:
            L.b200_w4a16_ar_unimport(ptr)
        self._imported = []
        if self._own:
            L.b200_w4a16_ar_free(self._own)
            self._own = None


def init_custom_all_reduce(max_bytes: int = 8 << 20, fused_out_features: int = 0,
                           ll_bytes: int = 1 << 20) -> None:
    """Collective over the TP group: allocate, exchange and map the regions (default 8 MiB
    messages, the reference's limit, custom_all_reduce.py:143).  ``fused_out_features`` > 0
    also reserves the slots of the fused row-parallel GEMM + all-reduce (the model's hidden
    size: the output width of o_proj / down_proj)."""
    global _handle
    if _handle is not None or get_tensor_model_parallel_world_size() == 1:
        return
    _handle = CustomAllreduce(max_bytes, fused_out_features, ll_bytes)


def fused_handle(M: int, N: int, G: int) -> Optional["CustomAllreduce"]:
    """The handle if a row-parallel GEMM of this shape can run fused with its all-reduce."""
    if _handle is not None and _handle.can_fuse(M, N, G):
        return _handle
    return None


def norm_handle(partial: torch.Tensor) -> Optional["CustomAllreduce"]:
    """The handle if the all-reduce of this tensor can run fused with the add + RMSNorm after it."""
    hidden = partial.shape[-1]
    # with packets the fused kernel wins up to ~128 rows; the barrier variant (message larger
    # than the packet area) up to ~32 rows at TP = 2, ~64 at TP = 8 (scripts/norm_bench.py)
    if (_handle is not None and _handle.should(partial) and hidden % 8 == 0 and hidden <= 16384
            and partial.numel() // hidden <= (4 * _FUSED_NORM_MAX_ROWS if _handle._fits_ll(partial.numel() * 2)
                                              else _FUSED_NORM_MAX_ROWS if _handle.world <= 2
                                              else 2 * _FUSED_NORM_MAX_ROWS)):
        return _handle
    return None


def is_initialized() -> bool:
    return _handle is not None


def destroy_custom_all_reduce() -> None:
    global _handle
    if _handle is not None:
        _handle.close()
        _handle = None


def maybe_all_reduce(input_: torch.Tensor) -> Optional[torch.Tensor]:
    """The reduced tensor if the peer-memory kernel took the message, else ``None``."""
    if _handle is None or not _handle.should(input_):
        return None
    return _handle.all_reduce(input_)
