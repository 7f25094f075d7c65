"""Resident-weight sharded layers for the throughput path (BASELINE.json configs): one process per
GPU, weights pre-sharded and resident in HBM (placement is outside the timed region, as in the
reference's loop, case6_attention.py:234-238), per-shard compute through the C ABI, collectives
on a side stream overlapped with the GEMMs of the next token chunk.

  TensorParallelFFN   W1 P(None,'model') column-parallel, W2 P('model',None) row-parallel; the
                      partial y of token chunk c is reduce-scattered (NCCL over NVLink) while the
                      GEMMs of chunk c+1 run; y comes back sequence-sharded P('model') like x_sharding.
  FsdpFFN             W1/W2 P('data',None): tokens sharded over 'data', each weight all-gathered
                      in column chunks that overlap the previous chunk's GEMM.
"""
from __future__ import annotations

import json
import os
import sys
from typing import Optional

import torch

from . import capi


def _act_code(act: str, bias: bool) -> int:
    if act == "gelu_tanh":
        return capi.EPI_BIAS_GELU_TANH if bias else capi.EPI_GELU_TANH
    if act == "relu":
        return capi.EPI_BIAS_RELU if bias else capi.EPI_RELU
    raise ValueError(f"unknown activation {act!r}")


class TensorParallelFFN:
    """y = act(x.W1 + b1).W2 + b2 with d_ff sharded over the `world` ranks of `group`.

    x: [T, d_model] bf16, replicated over the group.  Returns this rank's sequence block
    y[rank*T/world : (rank+1)*T/world] (reduce-scatter), bf16.  world == 1 is the plain two-GEMM
    layer.  The partial sums cross NVLink in bf16 (NCCL accumulates each hop in f32): half the
    bytes of f32 partials, inside the bf16 parity bar."""

    def __init__(self, w1, b1, w2, b2, act: str, T: int, group=None, chunks: int = 4):
        self.w1, self.b1, self.w2, self.b2 = w1, b1, w2, b2
        self.act = _act_code(act, b1 is not None)
        self.d_model, self.F = w1.shape
        self.T = T
        self.group = group
        if group is not None:
            import torch.distributed as dist
            self.dist = dist
            self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        else:
            self.world, self.rank = 1, 0
        dev = w1.device
        self.h = torch.empty((T, self.F), dtype=torch.bfloat16, device=dev)
        if self.world == 1:
            self.y = torch.empty((T, self.d_model), dtype=torch.bfloat16, device=dev)
            self.chunks = 1
        else:
            while chunks > 1 and (T % (self.world * chunks) or (T // (self.world * chunks)) % 128):
                chunks //= 2
            self.chunks = max(chunks, 1)
            if T % self.world or (T // self.world // self.chunks) % 128:
                raise ValueError(f"T={T} must split into {self.world} x {self.chunks} blocks of a multiple of 128 rows")
            self.ypart = torch.empty((T, self.d_model), dtype=torch.bfloat16, device=dev)
            self.y = torch.empty((T // self.world, self.d_model), dtype=torch.bfloat16, device=dev)
            self.comm_stream = torch.cuda.Stream(device=dev)
            self.chunk_done = [torch.cuda.Event() for _ in range(self.chunks)]
        self.kernel_events = None

    def forward(self, x: torch.Tensor, record=None) -> torch.Tensor:
        if self.world == 1:
            if record is not None:
                record("start")
            capi.gemm_bf16(x, self.w1, bias=self.b1, epilogue=self.act, out=self.h)
            if record is not None:
                record("gemm1")
            capi.gemm_bf16(self.h, self.w2, bias=self.b2, epilogue=capi.EPI_BIAS if self.b2 is not None else capi.EPI_NONE,
                           out=self.y)
            if record is not None:
                record("gemm2")
            return self.y
        W, C, T = self.world, self.chunks, self.T
        rows_blk = T // W // C                  # rows of one (rank block, chunk) cell
        rows_chunk = rows_blk * W               # logical rows of one chunk GEMM
        main = torch.cuda.current_stream(x.device)
        self.comm_stream.wait_stream(main)      # previous step's consumers of y / ypart are done
        for c in range(C):
            hc = self.h[c * rows_chunk:(c + 1) * rows_chunk]
            yc = self.ypart[c * rows_chunk:(c + 1) * rows_chunk]
            # chunk c = rows [c*rows_blk, (c+1)*rows_blk) of EVERY rank's sequence block of x
            xa = x[c * rows_blk:]
            if record is not None:
                record("start")
            capi.gemm_bf16(xa, self.w1, bias=self.b1, epilogue=self.act, out=hc,
                           a_row_blocks=(rows_chunk, rows_blk, T // W))
            if record is not None:
                record("gemm1")
            # b2 enters the sum exactly once: on rank 0's partial
            b2 = self.b2 if (self.rank == 0 and self.b2 is not None) else None
            capi.gemm_bf16(hc, self.w2, bias=b2, epilogue=capi.EPI_BIAS if b2 is not None else capi.EPI_NONE, out=yc)
            if record is not None:
                record("gemm2")
            self.chunk_done[c].record(main)
            with torch.cuda.stream(self.comm_stream):
                self.comm_stream.wait_event(self.chunk_done[c])
                self.dist.reduce_scatter_tensor(self.y[c * rows_blk:(c + 1) * rows_blk], yc, group=self.group)
        main.wait_stream(self.comm_stream)
        return self.y


class PeerReduceScatter:
    """Receive side of the fused GEMM -> reduce-scatter for a row-parallel layer with T rows and N output
    features over the W ranks of `group`: CUDA-IPC-shared receive buffers (two, alternating per step, each
    W slots of [T/W, N] bf16), a table of every owner's slot pointers for the GEMM epilogue to push into,
    a flag barrier and the local f32 reduction of the W received partials.

    `chunks` = 2 (opt-in; measured slower than one launch at 2 and at 8 GPUs -- 0.489 vs 0.466 ms per step at 8: the
    second GEMM at K = d_ff / 8 loses more to two short launches than the hidden half of the tail saves,
    profiles/r02_bench_n8_chunks*.json) pipelines the tail: the product is
    pushed as two GEMM launches, each covering HALF of every owner's row block, and the barrier + reduction
    of the first half run on a side stream underneath the second GEMM, so only half of the reduction (and one
    barrier) is left after the last GEMM tile.  The side stream is joined before `finish` returns: from outside
    the call stays ordered on the caller's stream.

    `batch` > 1: the T rows are a [batch, seq] activation and the result is scattered ALONG SEQ -- owner j receives
    rows (b, j*seq/W + s') for every batch item b, laid out [batch, seq/W, N]: the sequence-sharded out_shardings
    P('data','model') of the reference (case6_attention.py:161,230).  The GEMM reads its A operand through the
    two-level row-blocked view (b200_gemm_args.a_blocks_inner); chunks then split the batch items.  `batch` = 1
    scatters contiguous blocks of the flattened rows."""

    _FLAG_BYTES = 16384        # [0, 1024): barrier flags; [1024, 16384): progress flags [sender][raster group]
    _PROGRESS_OFF = 1024

    def __init__(self, T: int, N: int, group, device, chunks: int = 1, batch: int = 1):
        import torch.distributed as dist
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        W = self.world
        if T % W or (T // W) % 256:
            raise ValueError(f"T={T} must split into {W} blocks of a multiple of 256 rows")
        self.T, self.N, self.R = T, N, T // W
        self.batch = batch
        chunks = int(os.environ.get("B200_TP_CHUNKS", chunks))
        if batch > 1:
            self.seq = T // batch
            self.rpb = self.seq // W                             # rows of one (batch item, owner) cell
            if T % batch or self.seq % W or self.rpb % 128:
                raise ValueError(f"seq={T // batch} must split into {W} blocks of a multiple of 128 rows")
            while chunks > 1 and (batch % chunks or ((batch // chunks) * self.rpb) % 256):
                chunks -= 1
        else:
            while chunks > 1 and (self.R % chunks or (self.R // chunks) % 256):
                chunks -= 1
        self.chunks = max(chunks, 1)
        # tile order of the push: block after block (default), or cycling through the owners tile by tile
        # (B200_TP_INTERLEAVE=1; measured 1 % slower at 8 GPUs, profiles/r02_nvlink_push.md)
        self.interleave = os.environ.get("B200_TP_INTERLEAVE", "0") != "0"
        self.Rc = self.R // self.chunks                         # rows of one chunk inside one owner's block
        self.slot = self.R * N                                  # elements of one receive slot
        total = self._FLAG_BYTES + 2 * W * self.slot * 2        # flags + two receive buffers of W slots
        self.base, handle = capi.peer_alloc(total)
        mine = torch.tensor(list(handle), dtype=torch.uint8, device=device)
        handles = [torch.empty_like(mine) for _ in range(W)]
        dist.all_gather(handles, mine, group=group)
        self.peer_base = [self.base if p == self.rank else capi.peer_open(bytes(handles[p].cpu().tolist()))
                          for p in range(W)]
        self.flag_ptrs = torch.tensor(self.peer_base, dtype=torch.int64, device=device)
        # c_ptrs[buf][chunk][p]: where this rank's partial rows of (owner p, chunk) go in p's memory
        self.c_ptrs = []
        for buf in range(2):
            per_chunk = []
            for c in range(self.chunks):
                ptrs = [self.peer_base[p] + self._FLAG_BYTES + ((buf * W + self.rank) * self.slot + c * self.Rc * N) * 2
                        for p in range(W)]
                per_chunk.append(torch.tensor(ptrs, dtype=torch.int64, device=device))
            self.c_ptrs.append(per_chunk)
        self.buf = 0
        self.epoch = 0
        # progressive form (B200_TP_PROGRESSIVE=1): the pushing GEMM signals every finished raster group (16 row-blocks of
        # 256 rows, tiles cycling through the owners) to all owners, and the reduction runs group by group on a side stream
        # WHILE the GEMM produces the later groups: only the last group's share of the sum is left after the last tile
        self.progressive = (os.environ.get("B200_TP_PROGRESSIVE", "0") != "0" and self.chunks == 1 and 16 % W == 0
                            and T % 4096 == 0 and N >= 256 and W * (T // 4096) * 4 <= self._FLAG_BYTES - self._PROGRESS_OFF)
        if self.progressive:
            self.groups = T // 4096                              # raster groups of the pushing GEMM
            self.rows_per_group = (16 // W) * 256                # rows of every owner's block in one group
            self.interleave = True
            self.counters = torch.zeros(self.groups, dtype=torch.int32, device=device)
            self.progress_ptrs = torch.tensor([b + self._PROGRESS_OFF for b in self.peer_base], dtype=torch.int64, device=device)
            self.pepoch = 0
            self.buf_free = [torch.cuda.Event() for _ in range(2)]
            self.started = torch.cuda.Event()
            self.sum_ctas = int(os.environ.get("B200_TP_SUM_CTAS", "0")) or \
                2 * torch.cuda.get_device_properties(device).multi_processor_count
        self.side = torch.cuda.Stream(device=device) if (self.chunks > 1 or self.progressive) else None
        self.pushed = [torch.cuda.Event() for _ in range(self.chunks)]
        self.reduced = torch.cuda.Event()
        dist.barrier(group=group)

    def global_rows(self, local_rows: torch.Tensor, owner: int) -> torch.Tensor:
        """Rows of the flattened [T, .] input that the given rows of `owner`'s output block correspond to."""
        if self.batch > 1:
            return (local_rows // self.rpb) * self.seq + owner * self.rpb + local_rows % self.rpb
        return owner * self.R + local_rows

    def _push(self, a: torch.Tensor, w: torch.Tensor, c: int):
        if self.batch > 1:
            inner = self.batch // self.chunks                   # batch items of this launch
            capi.gemm_bf16_scatter(a[c * inner * self.seq:], w, self.c_ptrs[self.buf][c], self.Rc,
                                   (self.rank + 1) % self.world, self.N, M=self.world * self.Rc,
                                   a_view=(self.rpb, inner, self.seq, self.rpb), interleave=self.interleave)
        elif self.chunks == 1:
            capi.gemm_bf16_scatter(a, w, self.c_ptrs[self.buf][0], self.R, (self.rank + 1) % self.world, self.N,
                                   interleave=self.interleave)
        else:
            # logical block j of this launch = rows [c*Rc, (c+1)*Rc) of owner j's row block of `a`
            capi.gemm_bf16_scatter(a[c * self.Rc:], w, self.c_ptrs[self.buf][c], self.Rc, (self.rank + 1) % self.world,
                                   self.N, a_block_stride=self.R, M=self.world * self.Rc, interleave=self.interleave)

    def _reduce(self, c: int, bias, out: torch.Tensor):
        """Barrier (every rank's pushes of chunk c have landed), then out rows of chunk c = bf16(sum of the W partials)."""
        self.epoch += 1
        capi.peer_barrier(self.flag_ptrs, self.rank, self.world, self.epoch)
        src = self.base + self._FLAG_BYTES + (self.buf * self.world * self.slot + c * self.Rc * self.N) * 2
        capi.sum_bf16_partials(src, self.slot, self.world, bias, self.N, out[c * self.Rc:(c + 1) * self.Rc])

    def push_and_reduce(self, a: torch.Tensor, w: torch.Tensor, bias, out: torch.Tensor, record=None) -> torch.Tensor:
        """out[T/W, N] = this rank's row block of sum over ranks of (a @ w) (+ bias): the partial product is stored into
        the owners' receive slots from the GEMM epilogue (NVLink), reduced locally after a flag barrier."""
        main = torch.cuda.current_stream(a.device)
        if self.progressive:
            return self._push_and_reduce_progressive(a, w, bias, out, main, record)
        for c in range(self.chunks):
            self._push(a, w, c)
            if record is not None:
                record("gemm2")
                if c + 1 < self.chunks:
                    record("start")
            if c + 1 < self.chunks:
                self.pushed[c].record(main)
                with torch.cuda.stream(self.side):
                    self.side.wait_event(self.pushed[c])
                    self._reduce(c, bias, out)
                    if c + 2 == self.chunks:
                        self.reduced.record(self.side)
            else:
                self._reduce(c, bias, out)
                if record is not None:
                    record("reduce")
        if self.chunks > 1:
            main.wait_event(self.reduced)
        self.buf ^= 1
        return out

    def _push_and_reduce_progressive(self, a, w, bias, out, main, record):
        self.pepoch += 1
        main.wait_event(self.buf_free[self.buf])     # every owner has finished reading this receive buffer (two steps ago)
        progress = (self.counters, self.progress_ptrs, self.rank, self.groups, self.pepoch)
        tune = {"cta_group": 2, "block_n": 256, "group_m": 16}    # the reducer's groups are sized from this tile shape
        rot = (self.rank + 1) % self.world
        if self.batch > 1:
            capi.gemm_bf16_scatter(a, w, self.c_ptrs[self.buf][0], self.R, rot, self.N, M=self.T,
                                   a_view=(self.rpb, self.batch, self.seq, self.rpb), interleave=True, progress=progress, tune=tune)
        else:
            capi.gemm_bf16_scatter(a, w, self.c_ptrs[self.buf][0], self.R, rot, self.N, interleave=True, progress=progress,
                                   tune=tune)
        # the reduction follows the GEMM on the SAME stream as its programmatic dependent: it starts once the whole persistent
        # grid is resident, fills in beside it (no shared memory) and polls the progress flags group by group
        src = self.base + self._FLAG_BYTES + (self.buf * self.world * self.slot) * 2
        capi.peer_progressive_sum(self.base + self._PROGRESS_OFF, self.pepoch, self.world, self.groups, self.rows_per_group,
                                  src, self.slot, bias, self.N, out, self.sum_ctas, main)
        if record is not None:
            record("gemm2")                          # (the GEMM and the reduction beside it: they end together)
        self.reduced.record(main)
        side = self.side
        side.wait_event(self.reduced)
        self.epoch += 1
        capi.peer_barrier(self.flag_ptrs, self.rank, self.world, self.epoch, stream=side)   # off the step's critical path
        self.buf_free[self.buf].record(side)
        self.buf ^= 1
        return out

    # the two halves separately (single chunk): kept for callers that interleave other work
    def push_gemm(self, a: torch.Tensor, w: torch.Tensor):
        """Partial product a @ w with every finished tile stored into its owner's receive slot (NVLink)."""
        if self.chunks != 1:
            raise RuntimeError("push_gemm / finish are the single-chunk form; use push_and_reduce")
        self._push(a, w, 0)

    def finish(self, bias, out: torch.Tensor) -> torch.Tensor:
        """Barrier (all pushes landed), then out[T/W, N] = bf16(sum of the W partials) (+ bias)."""
        self._reduce(0, bias, out)
        self.buf ^= 1
        return out


class FusedTensorParallelFFN:
    """Tensor-parallel FFN whose row-parallel reduction is fused into the second GEMM: every finished tile
    of the partial y is stored from the GEMM epilogue straight into its owner's receive slot over NVLink
    (CUDA-IPC-mapped peer memory), ranks staggered so each destination has one sender at a time; then a
    flag barrier and one local pass summing the W received bf16 partials in f32 (+ b2).  No NCCL on the
    data path; both GEMMs run at full size (`chunks` = 2 splits the second so that the barrier + reduction of the first
    half hide under the second half: measured slower, PeerReduceScatter).  Same result layout as
    TensorParallelFFN: this rank's sequence block y[rank*T/W : (rank+1)*T/W]; with `batch` > 1 the block of every batch
    item along seq, [batch, seq/W, d_model] (PeerReduceScatter)."""

    def __init__(self, w1, b1, w2, b2, act: str, T: int, group, chunks: int = 1, batch: int = 1):
        self.w1, self.b1, self.w2, self.b2 = w1, b1, w2, b2
        self.act = _act_code(act, b1 is not None)
        self.d_model, self.F = w1.shape
        self.T = T
        dev = w1.device
        self.rs = PeerReduceScatter(T, self.d_model, group, dev, chunks=chunks, batch=batch)
        self.world, self.rank = self.rs.world, self.rs.rank
        self.h = torch.empty((T, self.F), dtype=torch.bfloat16, device=dev)
        self.y = torch.empty((self.rs.R, self.d_model), dtype=torch.bfloat16, device=dev)

    def forward(self, x: torch.Tensor, record=None) -> torch.Tensor:
        if record is not None:
            record("start")
        capi.gemm_bf16(x, self.w1, bias=self.b1, epilogue=self.act, out=self.h)
        if record is not None:
            record("gemm1")
        return self.rs.push_and_reduce(self.h, self.w2, self.b2, self.y, record=record)


class ShardedMHA:
    """Multi-head attention layer of the reference module (case6_attention.py:93-143) with heads sharded
    over the `group` ranks (the BASELINE.json case6 layout): to_q/to_k/to_v column-parallel
    and fused into ONE [M, 3*Hl*D] GEMM, attention on the local heads straight out of that buffer (strided
    q/k/v views, no transpose), to_out_0 row-parallel with the fused push reduce-scatter (or a plain GEMM +
    bias on one GPU).  x: [B, S, M] bf16 replicated over the group; returns this rank's sequence block of y,
    [B, S/W, M] flattened to [B*S/W, M]."""

    def __init__(self, wq, wk, wv, wo, bo, heads_local: int, dim_head: int, B: int, S: int, causal: bool, group=None,
                 scatter: str = "seq"):
        self.Hl, self.D, self.B, self.S, self.causal = heads_local, dim_head, B, S, causal
        self.wqkv = torch.cat([wq, wk, wv], dim=1).contiguous()        # [M, 3*Hl*D]
        self.wo, self.bo = wo, bo
        self.M = wq.shape[0]
        dev = wq.device
        T = B * S
        inner = heads_local * dim_head
        self.qkv = torch.empty((T, 3 * inner), dtype=torch.bfloat16, device=dev)
        self.o = torch.empty((T, inner), dtype=torch.bfloat16, device=dev)
        self.rs = None
        self.world = 1
        if group is not None:
            import torch.distributed as dist
            self.world = dist.get_world_size(group)
        if self.world > 1:
            # scatter = "seq": y is this rank's block along S of every batch item, [B, S/W, M] -- the out_shardings
            # P('data','model') of case6_attention.py:230; "rows": a contiguous block of the flattened B*S rows
            self.rs = PeerReduceScatter(T, self.M, group, dev, batch=B if scatter == "seq" else 1)
            self.y = torch.empty((self.rs.R, self.M), dtype=torch.bfloat16, device=dev)
        else:
            self.y = torch.empty((T, self.M), dtype=torch.bfloat16, device=dev)

    def forward(self, x: torch.Tensor, record=None) -> torch.Tensor:
        B, S, Hl, D = self.B, self.S, self.Hl, self.D
        inner = Hl * D
        if record is not None:
            record("start")
        capi.gemm_bf16(x.reshape(B * S, self.M), self.wqkv, out=self.qkv)                  # :96-99
        if record is not None:
            record("qkv")
        q, k, v = (self.qkv[:, i * inner:(i + 1) * inner].view(B, S, Hl, D) for i in range(3))   # :109-112
        capi.attention_fwd(q, k, v, scale=D ** -0.5, causal=self.causal, out=self.o.view(B, S, Hl, D))   # :121-135
        if record is not None:
            record("attn")
        if self.rs is None:
            capi.gemm_bf16(self.o, self.wo, bias=self.bo, epilogue=capi.EPI_BIAS, out=self.y)           # :139
        else:
            self.rs.push_and_reduce(self.o, self.wo, self.bo, self.y)
        if record is not None:
            record("out")
        return self.y


class FsdpFFN:
    """FSDP layout (BASELINE.json case3 config): tokens sharded over 'data' (this rank holds
    x[T/world rows]); W1 [d_model, d_ff] and W2 [d_ff, d_model] are dim0-sharded over 'data'.
    Each weight is all-gathered in `chunks` column chunks on a side stream; the GEMM on column
    chunk j runs while chunk j+1 (and then W2's chunks) are in flight.

    `prefetch=True` (default): the gathered-weight buffers are double-buffered by call parity and a forward call also
    starts the all-gather of W1 for the NEXT call once W2's chunks are queued (FSDP "forward prefetch"), so in a
    steady stream of calls W1's gather runs under GEMM2 of the call before.  Every call still gathers both weights in
    full over NVLink -- nothing gathered is kept beyond the call that uses it.  Dependencies are per-buffer events
    (the side stream never joins the whole main stream).  `gemm_max_ctas` > 0 bounds the persistent GEMM's grid so the
    collective's CTAs find free SMs beside it (0 = env B200_FSDP_GEMM_CTAS or no bound)."""

    def __init__(self, w1_shard, b1, w2_shard, b2, act: str, T_local: int, group, chunks: int = 4, prefetch: bool = True,
                 gemm_max_ctas: int = 0):
        import torch.distributed as dist
        self.dist, self.group = dist, group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.w1s, self.b1, self.w2s, self.b2 = w1_shard, b1, w2_shard, b2
        self.act = _act_code(act, b1 is not None)
        self.d_model = w1_shard.shape[0] * self.world
        self.F = w1_shard.shape[1]
        self.T = T_local
        self.chunks = chunks
        self.prefetch = prefetch
        self.gemm_max_ctas = gemm_max_ctas or int(os.environ.get("B200_FSDP_GEMM_CTAS", "0"))
        dev = w1_shard.device
        nbuf = 2 if prefetch else 1
        # gathered weights, chunk-major: [buffer][chunk][K rows][cols/chunks]
        self.w1g = torch.empty((nbuf, chunks, self.d_model, self.F // chunks), dtype=torch.bfloat16, device=dev)
        self.w2g = torch.empty((nbuf, chunks, self.F, self.d_model // chunks), dtype=torch.bfloat16, device=dev)
        # column-chunked views of the local shards must be contiguous for NCCL: keep a chunk-major copy
        self.w1s_c = w1_shard.reshape(w1_shard.shape[0], chunks, -1).permute(1, 0, 2).contiguous()
        self.w2s_c = w2_shard.reshape(w2_shard.shape[0], chunks, -1).permute(1, 0, 2).contiguous()
        self.h = torch.empty((T_local, self.F), dtype=torch.bfloat16, device=dev)
        self.y = torch.empty((T_local, self.d_model), dtype=torch.bfloat16, device=dev)
        self.comm_stream = torch.cuda.Stream(device=dev)
        self.ready1 = [[torch.cuda.Event() for _ in range(chunks)] for _ in range(nbuf)]   # W1 chunk gathered
        self.ready2 = [[torch.cuda.Event() for _ in range(chunks)] for _ in range(nbuf)]   # W2 chunk gathered
        self.read1 = [torch.cuda.Event() for _ in range(nbuf)]    # the GEMM1s that read w1g[buf] are done
        self.read2 = [torch.cuda.Event() for _ in range(nbuf)]    # the GEMM2s that read w2g[buf] are done
        self._call = 0
        self._w1_inflight = -1        # call index whose W1 gather has already been queued (prefetch)

    def _gather_w1(self, buf):
        self.comm_stream.wait_event(self.read1[buf])
        for j in range(self.chunks):
            self.dist.all_gather_into_tensor(self.w1g[buf, j], self.w1s_c[j], group=self.group)
            self.ready1[buf][j].record(self.comm_stream)

    def _gather_w2(self, buf):
        self.comm_stream.wait_event(self.read2[buf])
        for j in range(self.chunks):
            self.dist.all_gather_into_tensor(self.w2g[buf, j], self.w2s_c[j], group=self.group)
            self.ready2[buf][j].record(self.comm_stream)

    def forward(self, x: torch.Tensor, record=None) -> torch.Tensor:
        C = self.chunks
        main = torch.cuda.current_stream(x.device)
        call = self._call
        self._call += 1
        buf = call & 1 if self.prefetch else 0
        with torch.cuda.stream(self.comm_stream):
            if not self.prefetch:
                self.comm_stream.wait_stream(main)
            if self._w1_inflight != call:
                self._gather_w1(buf)
            self._gather_w2(buf)
            if self.prefetch:
                self._gather_w1(buf ^ 1)          # for the next call: runs under this call's second GEMM
                self._w1_inflight = call + 1
        tune = {"max_ctas": self.gemm_max_ctas} if self.gemm_max_ctas else None   # travels with each call: no library state
        fc, dc = self.F // C, self.d_model // C
        for j in range(C):
            main.wait_event(self.ready1[buf][j])
            if record is not None:
                record("start")
            b = self.b1[j * fc:(j + 1) * fc] if self.b1 is not None else None
            capi.gemm_bf16(x, self.w1g[buf, j], bias=b, epilogue=self.act, out=self.h[:, j * fc:(j + 1) * fc], tune=tune)
            if record is not None:
                record("gemm1")
        self.read1[buf].record(main)
        for j in range(C):
            main.wait_event(self.ready2[buf][j])
            if record is not None:
                record("start")
            b = self.b2[j * dc:(j + 1) * dc] if self.b2 is not None else None
            capi.gemm_bf16(self.h, self.w2g[buf, j], bias=b, epilogue=capi.EPI_BIAS if b is not None else capi.EPI_NONE,
                           out=self.y[:, j * dc:(j + 1) * dc], tune=tune)
            if record is not None:
                record("gemm2")
        self.read2[buf].record(main)
        return self.y


class PeerGatherFFN:
    """FSDP layout without NCCL on the data path: every rank keeps its dim0-shard of W1 and W2 in
    CUDA-IPC-shared memory, stored COLUMN-CHUNK-MAJOR ([chunk][rows][cols/chunks]); a forward pass PULLS the
    other ranks' shards straight out of their HBM, one column chunk at a time, while the GEMM of the previous
    chunk runs.  Because both the source chunk and its place in the chunk-major gathered matrix are contiguous,
    every pull is a plain 1-D peer copy on the copy engines (mode "ce": no SM time) -- or one SM kernel per chunk
    (mode "kernel", b200_peer_gather).  No cross-rank synchronisation is needed: weights are read-only.  Pull
    order is staggered by rank so every source serves one reader at a time."""

    def __init__(self, w1_shard, b1, w2_shard, b2, act: str, T_local: int, group, chunks: int = 4, copy_streams: int = 4,
                 mode: str = "ce", ctas: int = 48, prefetch: bool = True):
        import torch.distributed as dist
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        W = self.world
        self.b1, self.b2 = b1, b2
        self.act = _act_code(act, b1 is not None)
        self.k1, self.F = w1_shard.shape            # rows of W1 held here, d_ff
        self.k2, self.d_model = w2_shard.shape      # rows of W2 held here, d_model
        self.T = T_local
        self.chunks = C = chunks
        self.mode, self.ctas = mode, ctas
        # prefetch: as in FsdpFFN -- gathered buffers double-buffered by call parity, W1 of the next call pulled under
        # this call's second GEMM, per-buffer events instead of stream joins
        self.prefetch = prefetch
        nbuf = 2 if prefetch else 1
        dev = w1_shard.device
        n1, n2 = w1_shard.numel() * 2, w2_shard.numel() * 2
        self.n1 = n1
        self.base, handle = capi.peer_alloc(n1 + n2)
        # chunk-major copies of the local shards inside the shared allocation
        w1c = w1_shard.reshape(self.k1, C, self.F // C).permute(1, 0, 2).contiguous()
        w2c = w2_shard.reshape(self.k2, C, self.d_model // C).permute(1, 0, 2).contiguous()
        cur = torch.cuda.current_stream(dev)
        capi.peer_copy2d(self.base, n1, w1c.data_ptr(), n1, n1, 1, cur)
        capi.peer_copy2d(self.base + n1, n2, w2c.data_ptr(), n2, n2, 1, cur)
        torch.cuda.synchronize(dev)
        del w1c, w2c
        mine = torch.tensor(list(handle), dtype=torch.uint8, device=dev)
        handles = [torch.empty_like(mine) for _ in range(W)]
        dist.all_gather(handles, mine, group=group)
        self.peer_base = [self.base if p == self.rank else capi.peer_open(bytes(handles[p].cpu().tolist()))
                          for p in range(W)]
        self.src_bases = torch.tensor(self.peer_base, dtype=torch.int64, device=dev)
        # gathered weights, chunk-major: [chunk][W * rows][cols / chunks]
        self.w1g = torch.empty((nbuf, C, self.k1 * W, self.F // C), dtype=torch.bfloat16, device=dev)
        self.w2g = torch.empty((nbuf, C, self.k2 * W, self.d_model // C), dtype=torch.bfloat16, device=dev)
        self.h = torch.empty((T_local, self.F), dtype=torch.bfloat16, device=dev)
        self.y = torch.empty((T_local, self.d_model), dtype=torch.bfloat16, device=dev)
        self.streams = [torch.cuda.Stream(device=dev) for _ in range(copy_streams)]
        self.ready = [[[torch.cuda.Event() for _ in range(copy_streams)] for _ in range(2 * chunks)] for _ in range(nbuf)]
        self.read = [[torch.cuda.Event() for _ in range(2)] for _ in range(nbuf)]   # [buf][which]: the GEMMs reading it are done
        self._call = 0
        self._w1_inflight = -1
        dist.barrier(group=group)

    def _pull(self, which: int, j: int, buf: int):
        """Enqueue the pulls of column chunk j of weight `which` (0 = W1, 1 = W2) from every rank into buffer `buf`."""
        W, C = self.world, self.chunks
        rows, cols = (self.k1, self.F) if which == 0 else (self.k2, self.d_model)
        dst = self.w1g[buf] if which == 0 else self.w2g[buf]
        blk = rows * (cols // C) * 2                          # bytes of one rank's chunk (contiguous on both sides)
        src_off = (0 if which == 0 else self.n1) + j * blk
        dst0 = dst.data_ptr() + j * W * blk
        if self.mode == "kernel":
            capi.peer_gather(self.src_bases, W, self.rank, src_off, blk, dst0, blk, 1, blk, self.ctas, self.streams[0])
        else:
            for i in range(W):
                p = (self.rank + 1 + i) % W                   # own shard last
                st = self.streams[i % len(self.streams)]
                capi.peer_copy2d(dst0 + p * blk, blk, self.peer_base[p] + src_off, blk, blk, 1, st)
        for k, st in enumerate(self.streams):
            self.ready[buf][which * C + j][k].record(st)

    def _pull_weight(self, which: int, buf: int):
        for st in self.streams:
            st.wait_event(self.read[buf][which])              # the GEMMs that read this buffer last are done
        for j in range(self.chunks):
            self._pull(which, j, buf)

    def forward(self, x: torch.Tensor, record=None) -> torch.Tensor:
        C = self.chunks
        main = torch.cuda.current_stream(x.device)
        call = self._call
        self._call += 1
        buf = call & 1 if self.prefetch else 0
        if not self.prefetch:
            for st in self.streams:
                st.wait_stream(main)                          # the previous step's GEMMs have read w1g / w2g
        if self._w1_inflight != call:
            self._pull_weight(0, buf)
        self._pull_weight(1, buf)
        if self.prefetch:
            self._pull_weight(0, buf ^ 1)                     # for the next call: runs under this call's second GEMM
            self._w1_inflight = call + 1
        fc, dc = self.F // C, self.d_model // C
        for j in range(C):
            for ev in self.ready[buf][j]:
                main.wait_event(ev)
            if record is not None:
                record("start")
            b = self.b1[j * fc:(j + 1) * fc] if self.b1 is not None else None
            capi.gemm_bf16(x, self.w1g[buf, j], bias=b, epilogue=self.act, out=self.h[:, j * fc:(j + 1) * fc])
            if record is not None:
                record("gemm1")
        self.read[buf][0].record(main)
        for j in range(C):
            for ev in self.ready[buf][C + j]:
                main.wait_event(ev)
            if record is not None:
                record("start")
            b = self.b2[j * dc:(j + 1) * dc] if self.b2 is not None else None
            capi.gemm_bf16(self.h, self.w2g[buf, j], bias=b, epilogue=capi.EPI_BIAS if b is not None else capi.EPI_NONE,
                           out=self.y[:, j * dc:(j + 1) * dc])
            if record is not None:
                record("gemm2")
        self.read[buf][1].record(main)
        return self.y


class _RawCudaBuffer:
    """A device allocation made outside torch (b200_peer_alloc: CUDA-IPC-exported), exposed through
    __cuda_array_interface__ so that torch can view it without copying."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes // 2,), "typestr": "<u2", "data": (ptr, False), "version": 2}


def _bf16_view(ptr: int, rows: int, cols: int, device) -> torch.Tensor:
    return torch.as_tensor(_RawCudaBuffer(ptr, rows * cols * 2), device=device).view(torch.bfloat16).view(rows, cols)


class PushGatherFFN:
    """FSDP layout with the weight all-gather done by PUSHING: every rank stores its dim0-shard of W1 and W2 straight into
    every rank's gathered matrices (CUDA-IPC-shared, double-buffered by call parity) with one small kernel per weight
    (`b200_peer_push`: posted 16-byte stores over NVLink, a few CTAs that co-reside with the persistent GEMM), framed by two
    flag barriers on a side stream -- "the destination buffer is free on every rank" before, "every shard has landed" after.
    A dim0-shard is a contiguous row block of the gathered matrix, so the push is one contiguous copy per peer and the GEMMs
    read plain row-major weights.  As in the other FSDP layers the gather of W1 for the NEXT call runs under this call's
    second GEMM, and nothing gathered outlives the call that uses it."""

    _FLAG_BYTES = 1024

    def __init__(self, w1_shard, b1, w2_shard, b2, act: str, T_local: int, group, ctas: int = 0, prefetch: bool = True):
        import torch.distributed as dist
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        W = self.world
        self.b1, self.b2 = b1, b2
        self.act = _act_code(act, b1 is not None)
        self.k1, self.F = w1_shard.shape
        self.k2, self.d_model = w2_shard.shape
        self.T = T_local
        self.prefetch = prefetch
        self.ctas = ctas or int(os.environ.get("B200_FSDP_PUSH_CTAS", "16"))
        dev = w1_shard.device
        self.w1s, self.w2s = w1_shard.contiguous(), w2_shard.contiguous()
        self.n1, self.n2 = W * self.k1 * self.F * 2, W * self.k2 * self.d_model * 2       # bytes of the gathered matrices
        total = self._FLAG_BYTES + 2 * (self.n1 + self.n2)
        self.base, handle = capi.peer_alloc(total)
        mine = torch.tensor(list(handle), dtype=torch.uint8, device=dev)
        handles = [torch.empty_like(mine) for _ in range(W)]
        dist.all_gather(handles, mine, group=group)
        self.peer_base = [self.base if p == self.rank else capi.peer_open(bytes(handles[p].cpu().tolist()))
                          for p in range(W)]
        self.bases = torch.tensor(self.peer_base, dtype=torch.int64, device=dev)          # flags sit at offset 0 of each
        off = lambda buf, which: self._FLAG_BYTES + buf * (self.n1 + self.n2) + (0 if which == 0 else self.n1)
        self._off = off
        self.w1g = [_bf16_view(self.base + off(b, 0), W * self.k1, self.F, dev) for b in range(2)]
        self.w2g = [_bf16_view(self.base + off(b, 1), W * self.k2, self.d_model, dev) for b in range(2)]
        self.h = torch.empty((T_local, self.F), dtype=torch.bfloat16, device=dev)
        self.y = torch.empty((T_local, self.d_model), dtype=torch.bfloat16, device=dev)
        self.cs = torch.cuda.Stream(device=dev)
        self.ready = [[torch.cuda.Event() for _ in range(2)] for _ in range(2)]      # [buf][which]: gathered and visible
        self.read = [[torch.cuda.Event() for _ in range(2)] for _ in range(2)]       # [buf][which]: my GEMMs have read it
        self.epoch = 0
        self._call = 0
        self._w1_inflight = -1
        dist.barrier(group=group)

    def _gather(self, which: int, buf: int):
        """Queue the all-gather of weight `which` (0 = W1, 1 = W2) into buffer `buf` on the side stream."""
        cs = self.cs
        cs.wait_event(self.read[buf][which])                  # this rank's GEMMs that read the buffer last are done ...
        self.epoch += 1
        capi.peer_barrier(self.bases, self.rank, self.world, self.epoch, stream=cs)     # ... and so are everybody's
        shard = self.w1s if which == 0 else self.w2s
        nbytes = shard.numel() * 2
        capi.peer_push(self.bases, self.world, self.rank, self._off(buf, which) + self.rank * nbytes, shard, self.ctas, cs)
        self.epoch += 1
        capi.peer_barrier(self.bases, self.rank, self.world, self.epoch, stream=cs)     # every rank's shard has landed here
        self.ready[buf][which].record(cs)

    def forward(self, x: torch.Tensor, record=None) -> torch.Tensor:
        main = torch.cuda.current_stream(x.device)
        call = self._call
        self._call += 1
        buf = call & 1 if self.prefetch else 0
        if self._w1_inflight != call:
            self._gather(0, buf)
        self._gather(1, buf)
        if self.prefetch:
            self._gather(0, buf ^ 1)                          # for the next call: runs under this call's second GEMM
            self._w1_inflight = call + 1
        main.wait_event(self.ready[buf][0])
        if record is not None:
            record("start")
        capi.gemm_bf16(x, self.w1g[buf], bias=self.b1, epilogue=self.act, out=self.h)
        if record is not None:
            record("gemm1")
        self.read[buf][0].record(main)
        main.wait_event(self.ready[buf][1])
        if record is not None:
            record("start")
        capi.gemm_bf16(self.h, self.w2g[buf], bias=self.b2, epilogue=capi.EPI_BIAS if self.b2 is not None else capi.EPI_NONE,
                       out=self.y)
        if record is not None:
            record("gemm2")
        self.read[buf][1].record(main)
        return self.y


# ------------------------------------------------------------------------ bench harness glue
def _synthetic(shape, seed, kind, fan_in=None, device="cuda"):
    """Seeded synthetic tensors of SURVEY.md §8d, generated on the host like the oracle's."""
    g = torch.Generator().manual_seed(seed)
    t = torch.randn(shape, generator=g, dtype=torch.float32)
    if kind == "weight":
        t = t / (fan_in if fan_in is not None else shape[0]) ** 0.5
    elif kind == "bias":
        t = t * 0.02
    return t


class FFNBench:
    """Builds the case4 (TP) / case3 (FSDP) FFN workload for this rank and exposes the two step
    functions bench.py times: `step_resident` (inputs in HBM) and `step_e2e` (x from pinned host
    memory, y back to pinned host memory, in token chunks so copies overlap the GEMMs)."""

    def __init__(self, cfg, act, rank, world, dev, chunks=4, host_tensors=None):
        self.cfg, self.rank, self.world, self.dev = cfg, rank, world, dev
        T, d, f = cfg["batch"] * cfg["seq"], cfg["d_model"], cfg["d_ff"]
        self.T, self.d, self.f = T, d, f
        self.mode = cfg.get("parallelism", "tp") if world > 1 else "single"
        group = None
        self._dist = None
        if world > 1:
            import torch.distributed as dist
            group = dist.group.WORLD
            self._dist = dist
        if host_tensors is not None:          # the same seeded tensors, already generated for another sharding
            x, w1, b1, w2, b2 = host_tensors
        else:
            x = _synthetic((T, d), 0, "activation").to(torch.bfloat16)
            w1 = _synthetic((d, f), 1, "weight").to(torch.bfloat16)
            b1 = _synthetic((f,), 2, "bias").to(torch.bfloat16)
            w2 = _synthetic((f, d), 3, "weight").to(torch.bfloat16)
            b2 = _synthetic((d,), 4, "bias").to(torch.bfloat16)
        self._x_full = x
        self.host_weights = (w1, b1, w2, b2)          # full bf16 copies on the host: what bench.py's parity check reads
        if self.mode == "fsdp":
            tl, kl1, kl2 = T // world, d // world, f // world
            self.x_host = x[rank * tl:(rank + 1) * tl].contiguous().pin_memory()
            shard = (w1[rank * kl1:(rank + 1) * kl1].contiguous().to(dev), b1.to(dev),
                     w2[rank * kl2:(rank + 1) * kl2].contiguous().to(dev), b2.to(dev))
            self.layer = None
            # measured with the forward prefetch (profiles/r01_fsdp_gather_modes.md): copy-engine pulls win at 2 GPUs (2706 vs
            # NCCL 2501 TFLOP/s) and at 4 (5152 vs 4496), NCCL all-gather wins at 8 (7094 vs 6660); the SM pull is behind both
            default_mode = "peer" if world <= 4 else "nccl"
            fsdp_mode = os.environ.get("B200_FSDP_COLLECTIVE", default_mode)
            prefetch = os.environ.get("B200_FSDP_PREFETCH", "1") != "0"
            # with the next call's W1 prefetched, each weight has a whole GEMM to arrive under: one chunk keeps the GEMMs at
            # full waves (a d_model/4-column chunk of GEMM2 is 1.7 waves of tiles)
            chunks = int(os.environ.get("B200_FSDP_CHUNKS", "1" if prefetch else str(chunks)))
            if fsdp_mode == "push":
                try:
                    self.layer = PushGatherFFN(*shard, act, tl, group, prefetch=prefetch)
                    self.collective = (f"weight shards PUSHED into every rank's gathered matrices over NVLink peer memory "
                                       f"({self.layer.ctas} CTAs beside the GEMM, flag barriers on a side stream)" +
                                       (", W1 of the next call prefetched under GEMM2" if prefetch else ""))
                except Exception as e:   # noqa: BLE001
                    if rank == 0:
                        print(f"[bench] push gather unavailable ({e!r}); using NCCL", file=sys.stderr, flush=True)
            if fsdp_mode in ("peer", "peer_kernel"):
                try:
                    self.layer = PeerGatherFFN(*shard, act, tl, group, chunks=chunks, prefetch=prefetch,
                                               mode="kernel" if fsdp_mode == "peer_kernel" else "ce",
                                               copy_streams=1 if fsdp_mode == "peer_kernel" else 4)
                    self.collective = ("one-kernel SM pulls" if fsdp_mode == "peer_kernel" else "copy-engine pulls") + \
                        f" of weight column chunks ({chunks}) from peer HBM over NVLink, overlapped with GEMMs" + \
                        (", W1 of the next call prefetched under GEMM2" if prefetch else "")
                except Exception as e:   # noqa: BLE001
                    if rank == 0:
                        print(f"[bench] peer gather unavailable ({e!r}); using NCCL", file=sys.stderr, flush=True)
            if self.layer is None:
                self.layer = FsdpFFN(*shard, act, tl, group, chunks=chunks, prefetch=prefetch)
                self.collective = f"NCCL all-gather of weight column chunks ({chunks}) on a side stream" + \
                    (", W1 of the next call prefetched under GEMM2" if self.layer.prefetch else "") + \
                    (f", GEMM grid {self.layer.gemm_max_ctas} CTAs" if self.layer.gemm_max_ctas else "")
            self.T_in, self.T_out = tl, tl
        else:
            fl = f // world
            self.x_host = x.pin_memory()
            shard = (w1[:, rank * fl:(rank + 1) * fl].contiguous().to(dev), b1[rank * fl:(rank + 1) * fl].contiguous().to(dev),
                     w2[rank * fl:(rank + 1) * fl].contiguous().to(dev), b2.to(dev))
            self.collective = "none"
            if world > 1 and os.environ.get("B200_TP_COLLECTIVE", "fused") == "fused":
                try:
                    self.layer = FusedTensorParallelFFN(*shard, act, T, group)
                    self.collective = "fused GEMM epilogue push over NVLink peer memory + local reduce"
                except Exception as e:   # noqa: BLE001 - e.g. CUDA IPC not permitted in this container
                    if rank == 0:
                        print(f"[bench] fused reduce-scatter unavailable ({e!r}); using NCCL", file=sys.stderr, flush=True)
                    self.layer = None
            else:
                self.layer = None
            if self.layer is None:
                self.layer = TensorParallelFFN(*shard, act, T, group, chunks=chunks)
                if world > 1:
                    self.collective = "NCCL reduce-scatter per token chunk on a side stream"
            self.T_in, self.T_out = T, T // world
        self.x = self.x_host.to(dev)
        self.y_host = torch.empty((self.T_out, d), dtype=torch.bfloat16).pin_memory()
        self.x_stage = torch.empty_like(self.x)
        self.copy_in = torch.cuda.Stream(device=dev)
        self.copy_out = torch.cuda.Stream(device=dev)
        self._events = []
        # streaming host-buffer path: two sets of staging buffers (by step parity) so that the uploads of step i+1 and the
        # downloads of step i-1 run under the GEMMs of step i; on one GPU 2 token chunks per step keep both GEMMs at full
        # waves, on several GPUs a step is one chunk (the layer has its own chunking for the collective)
        self.e2e_chunks = 2 if world == 1 else 1
        self._e2e_step = 0
        # staging depth 2 (B200_E2E_DEPTH to A/B it: 3 and 4 measured no faster at 8 GPUs, where the step is bound by
        # the host link -- profiles/r02_nvlink_push.md)
        self.e2e_depth = D = max(2, int(os.environ.get("B200_E2E_DEPTH", "2")))
        self._x_stages = [self.x_stage] + [torch.empty_like(self.x) for _ in range(D - 1)]
        self._y_devs = [torch.empty((self.T_out, d), dtype=torch.bfloat16, device=dev) for _ in range(D)]
        self._y_hosts = [self.y_host] + [torch.empty((self.T_out, d), dtype=torch.bfloat16).pin_memory() for _ in range(D - 1)]
        mk = lambda: [[torch.cuda.Event() for _ in range(self.e2e_chunks)] for _ in range(D)]
        self._in_ev, self._xfree_ev, self._out_ev, self._ycopied_ev = mk(), mk(), mk(), mk()

    def host_tensors(self):
        return (self._x_full,) + tuple(self.host_weights)

    def parallelism_name(self):
        if self.mode == "single":
            return "1 GPU"
        mesh = "1x%d" % self.world if self.mode == "tp" else "%dx1" % self.world
        coll = getattr(self, "collective", "NCCL all-gather of weight column chunks on a side stream")
        return f"{'tp' if self.mode == 'tp' else 'fsdp'}{self.world} (mesh {mesh}; {coll})"

    # --- kernel-resident step
    def reset_kernel_events(self):
        self._events = []

    def _record(self, tag):
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        self._events.append((tag, ev))

    def step_resident(self, record_kernels=False):
        return self.layer.forward(self.x, record=self._record if record_kernels else None)

    def mean_kernel_ms(self):
        """(mean device time of one GEMM launch, number of GEMM launches timed), from events recorded around
        every GEMM launch of the timed region."""
        total, n = 0.0, 0
        prev = None
        for tag, ev in self._events:
            if tag in ("gemm1", "gemm2") and prev is not None:
                total += prev.elapsed_time(ev)
                n += 1
            prev = ev
        return total / max(n, 1), n

    def kernel_parts_ms(self, steps: int):
        """Mean device time per step of each recorded phase (gemm1, gemm2, reduce = the part of barrier + reduction that is
        left after the last GEMM), from the same events as mean_kernel_ms."""
        parts = {}
        prev = None
        for tag, ev in self._events:
            if tag != "start" and prev is not None:
                parts[tag] = parts.get(tag, 0.0) + prev.elapsed_time(ev)
            prev = ev
        return {k: v / max(steps, 1) for k, v in parts.items()}

    # --- end-to-end step: host x -> device -> FFN -> host y
    def step_e2e(self):
        """The public host-buffer call: x arrives in pinned host memory, y is returned in pinned host
        memory.  On one GPU the tokens go through in chunks so that the H2D copy of chunk c+1, the two
        GEMMs of chunk c and the D2H copy of chunk c-1 overlap (PCIe is full duplex)."""
        main = torch.cuda.current_stream(self.dev)
        if self.world > 1:
            # Several GPUs: the same streaming scheme with one chunk per step.  In TP mode x is replicated over the model
            # axis: every rank uploads only ITS 1/W of the rows over PCIe and the ranks all-gather the rest over NVLink
            # (W x 134 MB of host reads would be PCIe-bound).  The layer writes its own output buffer, so the result is
            # moved to this step's staging buffer (a device copy) before the download starts.
            par = self._e2e_step % self.e2e_depth
            self._e2e_step += 1
            x_stage, y_dev, y_host = self._x_stages[par], self._y_devs[par], self._y_hosts[par]
            if self.mode == "tp":
                rows = self.T // self.world
                sl = slice(self.rank * rows, (self.rank + 1) * rows)
            else:
                sl = slice(0, self.T_in)
            with torch.cuda.stream(self.copy_in):
                self.copy_in.wait_event(self._xfree_ev[par][0])      # the forward of two calls ago has read x_stage
                x_stage[sl].copy_(self.x_host[sl] if self.mode == "tp" else self.x_host, non_blocking=True)
                self._in_ev[par][0].record(self.copy_in)
            main.wait_event(self._in_ev[par][0])
            if self.mode == "tp":
                self._dist.all_gather_into_tensor(x_stage, x_stage[sl])
            y = self.layer.forward(x_stage)
            self._xfree_ev[par][0].record(main)
            main.wait_event(self._ycopied_ev[par][0])                # the download of two calls ago has read y_dev
            y_dev.copy_(y, non_blocking=True)
            self._out_ev[par][0].record(main)
            with torch.cuda.stream(self.copy_out):
                self.copy_out.wait_event(self._out_ev[par][0])
                y_host.copy_(y_dev, non_blocking=True)
                self._ycopied_ev[par][0].record(self.copy_out)
            return y_host
        # One GPU: a streaming call.  Step i uses staging set i % 2; within a step the tokens go through in chunks.  Every
        # dependency is an event on the buffer concerned (no stream-wide joins), so consecutive calls overlap: PCIe is
        # full duplex and the copy engines run beside the GEMMs.  The returned pinned buffer is complete once
        # `e2e_drain()` (or a synchronize) has run; it is overwritten two calls later.
        layer, C = self.layer, self.e2e_chunks
        rows = self.T // C
        par = self._e2e_step % self.e2e_depth
        self._e2e_step += 1
        x_stage, y_dev, y_host = self._x_stages[par], self._y_devs[par], self._y_hosts[par]
        for c in range(C):
            sl = slice(c * rows, (c + 1) * rows)
            with torch.cuda.stream(self.copy_in):
                self.copy_in.wait_event(self._xfree_ev[par][c])      # GEMM1 of two calls ago has read this chunk
                x_stage[sl].copy_(self.x_host[sl], non_blocking=True)
                self._in_ev[par][c].record(self.copy_in)
            main.wait_event(self._in_ev[par][c])
            capi.gemm_bf16(x_stage[sl], layer.w1, bias=layer.b1, epilogue=layer.act, out=layer.h[sl])
            self._xfree_ev[par][c].record(main)
            main.wait_event(self._ycopied_ev[par][c])                # the download of two calls ago has read this chunk
            capi.gemm_bf16(layer.h[sl], layer.w2, bias=layer.b2,
                           epilogue=capi.EPI_BIAS if layer.b2 is not None else capi.EPI_NONE, out=y_dev[sl])
            self._out_ev[par][c].record(main)
            with torch.cuda.stream(self.copy_out):
                self.copy_out.wait_event(self._out_ev[par][c])
                y_host[sl].copy_(y_dev[sl], non_blocking=True)
                self._ycopied_ev[par][c].record(self.copy_out)
        return y_host

    def e2e_drain(self):
        """Joins the copy streams into the current stream: after it, everything step_e2e() has returned is in host memory
        as far as the current stream is concerned (bench.py calls it before closing the timed region)."""
        main = torch.cuda.current_stream(self.dev)
        main.wait_stream(self.copy_out)
        main.wait_stream(self.copy_in)

    def e2e_bytes(self):
        h2d = self.x_host.numel() * 2
        if self.world > 1 and self.mode == "tp":
            h2d //= self.world          # each rank uploads its share of the replicated x
        return h2d, self.y_host.numel() * 2


def ncu_traffic_bytes():
    """dram read+write bytes per GEMM launch from the committed one-GPU ncu --set full capture of the case4 shape."""
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "ncu_traffic.json")
    if os.path.exists(path):
        try:
            with open(path) as f:
                return json.load(f).get("gemm_bf16_kernel_bytes_per_launch")
        except (OSError, ValueError):
            return None
    return None


def ffn_algorithmic_bytes(cfg, world: int):
    """Bytes one GEMM launch of the tensor-parallel FFN must move on ONE rank (operands read once + result written once):
    GEMM1 x[T,d] + W1[d,f/W] + h[T,f/W]; GEMM2 h + W2[f/W,d] + y partial [T,d]; bf16."""
    T, d, f = cfg["batch"] * cfg["seq"], cfg["d_model"], cfg["d_ff"] // world
    return {"gemm1": 2 * (T * d + d * f + T * f), "gemm2": 2 * (T * f + f * d + T * d)}


def attention_mesh(world: int):
    """(data, model) mesh for the attention configs at `world` GPUs: (1,1), (1,2), (2,2), (2,4)."""
    return {1: (1, 1), 2: (1, 2), 4: (2, 2), 8: (2, 4)}.get(world, (1, world))


def run_attention_workload(cfg, rank, world, dev, steps, warmup, keep=False):
    """The MHA layer of BASELINE.json's case5 / case6 configs (projections + attention core + out-projection)
    on a (data x model) mesh: batch over 'data', heads over 'model' (fused push reduce-scatter for the
    row-parallel out-projection).  Returns timing / FLOP figures (max over ranks)."""
    B, S, H, D, M = cfg["batch"], cfg["seq"], cfg["heads"], cfg["dim_head"], cfg["d_model"]
    causal = cfg["causal"]
    d_ax, m_ax = attention_mesh(world)
    if B % d_ax or H % m_ax:
        raise SystemExit(f"batch {B} / heads {H} do not split over mesh {d_ax}x{m_ax}")
    group = None
    dist = None
    if world > 1:
        import torch.distributed as dist
        mine = None
        for g in range(d_ax):
            ranks = list(range(g * m_ax, (g + 1) * m_ax))
            pg = dist.new_group(ranks=ranks) if m_ax > 1 else None
            if rank in ranks:
                mine = pg
        group = mine
    Bl, Hl = B // d_ax, H // m_ax
    m_rank = rank % m_ax
    gen = torch.Generator(device=dev).manual_seed(1234 + rank // m_ax)        # same x inside a model group
    x = torch.randn((Bl, S, M), device=dev, generator=gen, dtype=torch.float32).to(torch.bfloat16)
    wgen = torch.Generator(device=dev).manual_seed(99)                         # same weights on every rank

    def weight(rows, cols, fan_in):
        return (torch.randn((rows, cols), device=dev, generator=wgen, dtype=torch.float32) / fan_in ** 0.5).to(torch.bfloat16)

    inner = H * D
    wq, wk, wv = weight(M, inner, M), weight(M, inner, M), weight(M, inner, M)
    wo = weight(inner, M, inner)
    bo = (torch.randn((M,), device=dev, generator=wgen, dtype=torch.float32) * 0.02).to(torch.bfloat16)
    cs = slice(m_rank * Hl * D, (m_rank + 1) * Hl * D)
    layer = ShardedMHA(wq[:, cs].contiguous(), wk[:, cs].contiguous(), wv[:, cs].contiguous(), wo[cs].contiguous(), bo,
                       Hl, D, Bl, S, causal, group if m_ax > 1 else None)
    del wq, wk, wv, wo
    for _ in range(warmup):
        layer.forward(x)
    torch.cuda.synchronize(dev)
    if dist is not None:
        dist.barrier()
        torch.cuda.synchronize(dev)
    events = []

    def record(tag):
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        events.append((tag, ev))

    l0 = capi.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        layer.forward(x, record=record)
    e1.record()
    torch.cuda.synchronize(dev)
    if dist is not None:
        dist.barrier()
    ms = e0.elapsed_time(e1) / steps
    parts = {"qkv": 0.0, "attn": 0.0, "out": 0.0}
    prev = None
    for tag, ev in events:
        if tag != "start":
            parts[tag] += prev.elapsed_time(ev)
        prev = ev
    parts = {k: v / steps for k, v in parts.items()}
    if dist is not None:
        t = torch.tensor([ms, parts["qkv"], parts["attn"], parts["out"]], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, parts = float(t[0]), {"qkv": float(t[1]), "attn": float(t[2]), "out": float(t[3])}
    core = 2.0 * B * H * S * (S + 1) * D if causal else 4.0 * B * H * S * S * D
    proj = 8.0 * B * S * M * inner
    res = {"ms_per_step": ms, "parts_ms": parts, "core_flops": core, "proj_flops": proj, "mesh": (d_ax, m_ax),
           "launches": int(capi.launch_count() - l0), "steps": steps}
    if keep:      # for bench.py's stage-wise parity check of what was just timed
        res.update(layer=layer, x=x, group=group if m_ax > 1 else None)
    return res


def attention_core_sample(dev, B, S, H, D, causal, peaks, iters=5):
    """CUDA-event timing of the attention core alone at one shape (the head dim 64 figure of the default bench line)."""
    q, k, v = (torch.randn(B, S, H, D, device=dev).to(torch.bfloat16) for _ in range(3))
    o = torch.empty_like(q)
    for _ in range(2):
        capi.attention_fwd(q, k, v, causal=causal, out=o)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        capi.attention_fwd(q, k, v, causal=causal, out=o)
    e1.record()
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / iters
    flops = 2.0 * B * H * S * (S + 1) * D if causal else 4.0 * B * H * S * S * D
    tf = flops / (ms * 1e-3) / 1e12
    return {"shape": [B, S, H, D], "causal": causal, "ms": ms, "core_tflops": tf, "core_frac_of_peak": tf / peaks["bf16_burst"]}


def _attention_e2e(r, dev, world, steps=3):
    """The same layer with x in pinned HOST memory and y returned to pinned host memory, every step (no pipelining across
    steps: upload, forward, download).  Returns (ms per step max over ranks, h2d bytes, d2h bytes) summed over ranks."""
    layer, x = r["layer"], r["x"]
    x_host = x.cpu().pin_memory()
    y0 = layer.forward(x)
    y_host = torch.empty(y0.shape, dtype=y0.dtype).pin_memory()
    x_dev = torch.empty_like(x)

    def step():
        x_dev.copy_(x_host, non_blocking=True)
        y = layer.forward(x_dev)
        y_host.copy_(y, non_blocking=True)

    step()
    torch.cuda.synchronize(dev)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / steps
    h2d, d2h = x_host.numel() * 2, y_host.numel() * 2
    if dist is not None:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        b = torch.tensor([h2d, d2h], device=dev, dtype=torch.float64)
        dist.all_reduce(b)
        h2d, d2h = int(b[0]), int(b[1])
    return ms, h2d, d2h


def bench_attention(args, cfg, rank, world, dev, metric, unit, peaks, sampler, attn_flops, cpu_sample=None):
    """`bench.py --workload case5_attention_dense|case6_attention`: the full MHA layer at the config shape."""
    sampler.start()
    r = run_attention_workload(cfg, rank, world, dev, args.steps, args.warmup, keep=True)
    clocks = sampler.stop()
    e2e_ms, h2d, d2h = _attention_e2e(r, dev, world)
    if rank != 0:
        return
    total0 = r["core_flops"] + r["proj_flops"]
    e2e = {"value": total0 / (e2e_ms * 1e-3) / 1e12, "unit": unit, "ms_per_step": e2e_ms, "h2d_bytes_per_step": h2d,
           "d2h_bytes_per_step": d2h, "what": "x uploaded from pinned host memory and y downloaded every step, not pipelined"}
    cpu = cpu_sample(cfg) if (cpu_sample is not None and world == 1 and not args.no_cpu_baseline) else None
    total = r["core_flops"] + r["proj_flops"]
    val = total / (r["ms_per_step"] * 1e-3) / 1e12
    core_tf = r["core_flops"] / (r["parts_ms"]["attn"] * 1e-3) / 1e12 / world
    d_ax, m_ax = r["mesh"]
    print(json.dumps({
        "metric": metric, "value": val, "unit": unit, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": {"workload": args.workload, "batch": cfg["batch"], "seq_len": cfg["seq"], "heads": cfg["heads"],
                   "dim_head": cfg["dim_head"], "d_model": cfg["d_model"], "causal": cfg["causal"],
                   "parallelism": f"mesh data x model = {d_ax}x{m_ax} (batch over data, heads over model)",
                   "flops_per_step": total, "core_flops": r["core_flops"], "proj_flops": r["proj_flops"],
                   "parts_ms": r["parts_ms"]},
        "clocks": clocks, "gpu_launches": r["launches"],
        "roofline": {"bound": "tensor", "kernel": "attention_fwd9_kernel (CTA pairs; per GPU, attention core only)",
                     "achieved": core_tf, "peak": peaks["bf16_burst"], "unit": unit,
                     "frac": core_tf / peaks["bf16_burst"], "peak_source": peaks["source"], "traffic": None},
        "cpu_baseline": cpu, "e2e": e2e}), flush=True)
