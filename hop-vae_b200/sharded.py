"""Multi-GPU modes of the Hopfield retrieval (new design; the reference is single-device).

One process per GPU (`torchrun`), `torch.distributed` for the plumbing.

* **batch-parallel** (`batch_parallel_rows`): query rows are independent, so N is split across ranks and
  every rank holds all patterns; the forward needs no collective at all.
* **pattern-sharded** (`PatternShardedHopfield`): the stored patterns are split into G contiguous shards;
  every rank sees all N queries and evaluates the fused retrieval on its shard, producing per row the
  shard-local normalised read-out O_g, `lse_g`, the row max and its (global) index.  Softmax over P is a
  (max, sum) monoid, so one exchange per pass finishes the job:

      lse = log sum_g exp(lse_g),  w_g = exp(lse_g - lse),  O = sum_g w_g O_g

  `all_gather` of the 3 per-row statistics (12 bytes/row/shard), the library's merge kernel scales the local
  O_g by w_g in place, one `all_reduce(SUM)` over NVLink gives O (or `reduce_scatter` when the consumer is
  row-sharded: `scatter_rows=True`).  With T > 0 retrieve updates the merge happens after every pass
  (the state q <- beta * A K^ is itself a read-out over the sharded K^), so it is one launch per pass.

On one NVLink / NVSwitch node the exchange is fused into ONE kernel per rank over peer memory
(`PeerExchange`, `hop_sm100_peer_merge`): every rank's fused forward leaves (O_g, lse_g, max_g, arg_g) in a
buffer all peers have mapped; rank r waits on a flag barrier in peer memory, pulls the G partial rows of its
N/G rows with P2P loads, merges them and pushes the finished rows into every rank's result buffer with P2P
stores -- no NCCL call, no (G, N, Dv) staging, no separate scale pass.  The NCCL form above stays as the
portable fallback (`exchange="collective"`, other topologies, and the CPU tests).

The arithmetic is injected as a small backend object: `CudaBackend` calls the C-ABI ops (the product);
the CPU `gloo` tests inject an oracle-backed stand-in to exercise the sharding / exchange logic.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import torch
import torch.distributed as dist
from torch import Tensor

from . import cabi


def shard_range(n_patterns: int, n_shards: int, shard: int, align: int = 128) -> Tuple[int, int]:
    """Contiguous [start, start+count) of shard `shard`; boundaries on multiples of `align` patterns (one
    score tile) where P allows, so no shard pays for a padded tile but the last."""
    if not (0 <= shard < n_shards):
        raise ValueError(f"shard {shard} out of range for {n_shards} shards")
    if n_patterns < n_shards:
        raise ValueError(f"cannot split {n_patterns} patterns over {n_shards} shards")
    tiles = (n_patterns + align - 1) // align
    if tiles >= n_shards:
        base, extra = divmod(tiles, n_shards)
        t0 = shard * base + min(shard, extra)
        t1 = t0 + base + (1 if shard < extra else 0)
        start, end = t0 * align, min(t1 * align, n_patterns)
    else:
        base, extra = divmod(n_patterns, n_shards)
        start = shard * base + min(shard, extra)
        end = start + base + (1 if shard < extra else 0)
    return start, end - start


def batch_parallel_rows(n_rows: int, world: int, rank: int) -> Tuple[int, int]:
    """Rows [start, start+count) of rank `rank` in batch-parallel mode (even split, remainder to the first ranks)."""
    base, extra = divmod(n_rows, world)
    start = rank * base + min(rank, extra)
    return start, base + (1 if rank < extra else 0)


class PeerExchange:
    """Peer-mapped exchange buffers of one rank + the mappings of every peer's (one node, one process per GPU).

    Layout of a rank's block (fp32 words): o (N*W) | lse (N) | rowmax (N) | argmax (N) | 2 x [out (N*W) | lse (N)
    | argmax (N)] | flags (HOP_PEER_FLAG_WORDS).  Result buffers alternate with the call parity, so a returned
    tensor stays valid until the next-but-one exchange.  The blocks are shared through CUDA IPC handles
    (`hop_sm100_ipc_export` / `_import`), exchanged with `all_gather_object`."""

    def __init__(self, n_rows: int, width: int, group=None, device: Optional[torch.device] = None):
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        if self.world > cabi.MAX_PEERS:
            raise ValueError(f"peer exchange supports at most {cabi.MAX_PEERS} ranks")
        self.n, self.w = int(n_rows), int(width)
        dev = device if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.device = dev
        nw = self.n * self.w
        self.off = dict(o=0, lse=nw, rowmax=nw + self.n, argmax=nw + 2 * self.n)
        base = nw + 3 * self.n
        base = (base + 63) // 64 * 64
        self.res = []
        for _ in range(2):
            self.res.append(dict(out=base, lse=base + nw, argmax=base + nw + self.n))
            base = (base + nw + 2 * self.n + 63) // 64 * 64
        self.off["flags"] = base
        words = base + cabi.PEER_FLAG_WORDS
        # a dedicated allocation (not a slice of a cached block), zeroed: the flag words start at epoch 0
        self.block = torch.zeros((words,), device=dev, dtype=torch.float32)
        torch.cuda.synchronize(dev)
        import ctypes as C
        lib = cabi.load()
        handle = C.create_string_buffer(64)
        offset = C.c_uint64(0)
        with torch.cuda.device(dev):
            cabi.check(lib.hop_sm100_ipc_export(self.block.data_ptr(), handle, C.byref(offset)), "hop_sm100_ipc_export")
        handles = [None] * self.world
        dist.all_gather_object(handles, (int(dev.index), handle.raw, int(offset.value)), group=group)
        self.base_ptr: List[int] = []
        self._imported: List[Tuple[int, int]] = []
        for g, (dev_g, raw, off_g) in enumerate(handles):
            if g == self.rank:
                self.base_ptr.append(self.block.data_ptr())
                continue
            p = C.c_void_p(0)
            with torch.cuda.device(dev):     # mapped on THIS rank's device: its kernels load / store it over NVLink
                cabi.check(lib.hop_sm100_enable_peer_access(int(dev_g)), "hop_sm100_enable_peer_access")
                cabi.check(lib.hop_sm100_ipc_import(raw, off_g, C.byref(p)), "hop_sm100_ipc_import")
            self.base_ptr.append(int(p.value))
            self._imported.append((int(p.value), off_g))
        self.epoch = 0
        dist.barrier(group=group)          # every rank has mapped every block before the first kernel runs

    def close(self) -> None:
        """Unmap the peers' blocks (call on every rank after a barrier; the block itself is a torch tensor)."""
        lib = cabi.load()
        for p, off in self._imported:
            lib.hop_sm100_ipc_close(p, off)
        self._imported = []

    def _view(self, off: int, shape, dtype=torch.float32) -> Tensor:
        n = 1
        for d in shape:
            n *= d
        t = self.block[off:off + n]
        if dtype != torch.float32:
            t = t.view(dtype)
        return t.view(*shape)

    def local_slots(self) -> Tuple[Tensor, Tensor, Tensor]:
        """(out (N,W), stats (2,N), argmax (N) i32) views of this rank's partial slots (`retrieve(into=...)`)."""
        return (self._view(self.off["o"], (self.n, self.w)), self._view(self.off["lse"], (2, self.n)),
                self._view(self.off["argmax"], (self.n,), torch.int32))

    def merge(self, row_begin: int, row_count: int, push_all: bool) -> Tuple[Tensor, Tensor, Tensor]:
        """Launch the fused exchange on the current stream; returns views (O (N,W), lse (N), argmax (N)) of this
        rank's result buffers (rows outside [row_begin, +row_count) are valid only with push_all)."""
        import ctypes as C
        from .ops import _stream, _on_device
        self.epoch += 1
        r = self.res[self.epoch & 1]
        m = cabi.HopPeerMerge()
        m.n_rows, m.d_v, m.n_ranks, m.rank = self.n, self.w, self.world, self.rank
        m.row_begin, m.row_count, m.push_all, m.epoch = row_begin, row_count, 1 if push_all else 0, self.epoch
        for g, b in enumerate(self.base_ptr):
            m.o[g] = b + 4 * self.off["o"]
            m.lse[g] = b + 4 * self.off["lse"]
            m.rowmax[g] = b + 4 * self.off["rowmax"]
            m.argmax[g] = b + 4 * self.off["argmax"]
            m.out[g] = b + 4 * r["out"]
            m.lse_out[g] = b + 4 * r["lse"]
            m.argmax_out[g] = b + 4 * r["argmax"]
            m.flags[g] = b + 4 * self.off["flags"]
        with _on_device(self.device):
            cabi.check(cabi.load().hop_sm100_peer_merge(C.byref(m), _stream(self.device)), "hop_sm100_peer_merge")
        return (self._view(r["out"], (self.n, self.w)), self._view(r["lse"], (self.n,)),
                self._view(r["argmax"], (self.n,), torch.int32))


class CudaBackend:
    """The product path: sm_100a kernels through the C-ABI custom ops."""
    peer_capable = True

    def prepare(self, pack: Dict[str, Optional[Tensor]], flags: int, path: int, d_assoc: int, d_v: int,
                update_steps: int, ln_eps: float) -> List[Tensor]:
        from .ops import prepare_patterns_op, _PATTERN_SIDE
        args = [None if pack.get(k) is None else pack[k].detach().contiguous() for k in _PATTERN_SIDE]
        return prepare_patterns_op(*args, flags, path, d_assoc, d_v, update_steps, ln_eps)

    def retrieve(self, x: Tensor, pats: List[Tensor], ln_w: Optional[Tensor], ln_b: Optional[Tensor], flags: int,
                 path: int, d_in: int, beta: float, ln_eps: float, pattern_offset: int, over_keys: bool,
                 fold: Optional[Tuple[Tensor, Optional[Tensor]]] = None, into=None):
        """One association pass on this shard.  `over_keys` reads out over K^ (a retrieve update).  With
        `fold=(wo, bo)` and a folded operand in `pats`, the pass returns O_g Wo^T + bo (the shard weights sum
        to 1, so the bias survives the merge unchanged); the second return value says whether it did."""
        from .ops import retrieve
        khat, vhat, k16, kt16, vt16, vot16 = pats[:6]
        kstats = pats[9]
        if over_keys:
            vhat, vt16 = khat, kt16
        empty = vot16.new_empty((0,))
        # one T = 0 pass per call: no update operands; the nearest-pattern index of a shard is merged across shards
        # on its 16-bit row maximum, so the per-shard fp32 repair (HOP_F_EXACT_ARGMAX) stays off here
        flags = flags & ~cabi.F_EXACT_ARGMAX
        one = lambda vo: [khat, vhat, k16, empty, vt16, vo, empty, empty, empty, kstats]
        if fold is not None and not over_keys and vot16.numel() > 0 and path != cabi.PATH_FP32_SIMT:
            wo, bo = fold
            res = retrieve(x, one(vot16), ln_w, ln_b, None, None, wo, bo,
                           flags | cabi.F_OUT_PROJ, path, d_in, wo.shape[0], 0, beta, ln_eps, pattern_offset, False,
                           False, into)
            return (res[0], res[3], res[5], res[4]), True
        res = retrieve(x, one(empty), ln_w, ln_b, None, None, None, None,
                       flags & ~cabi.F_OUT_PROJ, path, d_in, 1, 0, beta, ln_eps, pattern_offset, True, False, into)
        # no out_proj: out == read-out.  (O_g, lse_g, rowmax_g, argmax_g)
        return (res[0], res[3], res[5], res[4]), False

    def merge(self, retrieved: Tensor, lse_all: Tensor, rowmax_all: Tensor, argmax_all: Tensor, shard: int):
        from .ops import merge_shards
        return merge_shards(retrieved, lse_all, rowmax_all, argmax_all, shard)

    def out_projection(self, retrieved: Tensor, wo: Tensor, bo: Optional[Tensor]) -> Tensor:
        from .ops import out_projection
        return out_projection(retrieved, wo, bo)

    def resolve_path(self, spec, n_rows: int, n_patterns: int) -> int:
        from .ops import resolve_path
        return resolve_path(spec, n_rows, n_patterns)


class PatternShardedHopfield:
    """Pattern-sharded evaluation of one `HopfieldLayer` over a process group (inference).

    Built on every rank from the *full* layer (or any object with `spec()` / `param_pack()` / `quantity`);
    each rank keeps only its contiguous slice of the pattern-side tensors.
    """

    def __init__(self, layer, group: Optional[dist.ProcessGroup] = None, backend=None, scatter_rows: bool = False,
                 exchange: str = "auto"):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.backend = backend if backend is not None else CudaBackend()
        self.scatter_rows = scatter_rows
        if exchange not in ("auto", "peer", "collective"):
            raise ValueError("exchange must be 'auto', 'peer' or 'collective'")
        peer_ok = bool(getattr(self.backend, "peer_capable", False)) and self.world > 1
        if exchange == "peer" and not peer_ok:
            raise ValueError("exchange='peer' needs the CUDA backend and more than one rank")
        # "auto": the fused peer-memory exchange on a CUDA backend (one node), NCCL collectives otherwise
        self.use_peer = peer_ok and exchange != "collective"
        self._peer: Dict[Tuple[int, int], PeerExchange] = {}
        self.spec = layer.spec()
        if self.spec.flags & (cabi.F_PROJ_STATE | cabi.F_PROJ_STORED):
            raise NotImplementedError("pattern-sharded mode is built for static state/stored patterns "
                                      "(what Hop-VAE uses, HopVAE.py:149-170)")
        self.n_patterns = int(layer.quantity)
        self.start, self.count = shard_range(self.n_patterns, self.world, self.rank)
        pack = layer.param_pack()
        sl = slice(self.start, self.start + self.count)
        self.pack = dict(pack)
        self.pack["lookup"] = pack["lookup"].detach()[sl].contiguous()
        if pack.get("lookup_targets") is not None:
            self.pack["lookup_targets"] = pack["lookup_targets"].detach()[sl].contiguous()
        self._pats = None
        self._path = None
        self.last_argmax: Optional[Tensor] = None
        self.last_lse: Optional[Tensor] = None

    # ------------------------------------------------------------------ one (local pass + exchange) round
    def _exchange(self, o: Tensor, lse: Tensor, rowmax: Tensor, argmax: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
        """Shard-local (O_g, lse_g, max_g, arg_g) -> global (O, lse, argmax) on every rank."""
        if self.world == 1:
            return o, lse, argmax
        n = lse.shape[0]
        stats = torch.stack([lse, rowmax, argmax.view(torch.float32)])          # (3, N), 12 bytes per row
        gathered = torch.empty((self.world * 3, n), device=stats.device, dtype=torch.float32)
        dist.all_gather_into_tensor(gathered, stats, group=self.group)
        gathered = gathered.view(self.world, 3, n)
        lse_all = gathered[:, 0].contiguous()
        max_all = gathered[:, 1].contiguous()
        arg_all = gathered[:, 2].contiguous().view(torch.int32)
        g_lse, g_arg = self.backend.merge(o, lse_all, max_all, arg_all, self.rank)   # o *= exp(lse_g - lse)
        if self.scatter_rows:
            if n % self.world != 0:
                raise ValueError("scatter_rows needs N divisible by the world size")
            rows = n // self.world
            mine = torch.empty((rows, o.shape[1]), device=o.device, dtype=o.dtype)
            dist.reduce_scatter_tensor(mine, o, op=dist.ReduceOp.SUM, group=self.group)
            return mine, g_lse, g_arg
        dist.all_reduce(o, op=dist.ReduceOp.SUM, group=self.group)
        return o, g_lse, g_arg

    def _pass(self, state: Tensor, ln_w, ln_b, st_flags: int, over_keys: bool, fold):
        """One local association pass + exchange -> (O, global lse, global argmax, folded?)."""
        spec = self.spec
        n = state.shape[0]
        args = (state, self._pats, ln_w, ln_b, st_flags, self._path, spec.d_in, spec.beta, spec.ln_eps, self.start,
                over_keys)
        if not self.use_peer:
            (o, lse, rowmax, argmax), folded = self.backend.retrieve(*args, fold)
            o, g_lse, g_arg = self._exchange(o, lse, rowmax, argmax)
            return o, g_lse, g_arg, folded
        will_fold = (fold is not None and not over_keys and self._pats[5].numel() > 0
                     and self._path != cabi.PATH_FP32_SIMT)
        width = spec.d_assoc if over_keys else (fold[0].shape[0] if will_fold else spec.d_v)
        ex = self._peer.get((n, width))
        if ex is None:
            if len(self._peer) >= 4:          # variable batch sizes: keep the IPC mappings bounded (collective call:
                self.close()                  # every rank sees the same sequence of (N, width))
            ex = self._peer[(n, width)] = PeerExchange(n, width, self.group, state.device)
        _, folded = self.backend.retrieve(*args, fold, ex.local_slots())
        start, count = batch_parallel_rows(n, self.world, self.rank)
        scatter = self.scatter_rows and not over_keys
        o, g_lse, g_arg = ex.merge(start, count, push_all=not scatter)
        if scatter:
            o, g_lse, g_arg = o[start:start + count], g_lse[start:start + count], g_arg[start:start + count]
        return o, g_lse, g_arg, folded

    def forward(self, x: Tensor) -> Tensor:
        spec = self.spec
        lead = x.shape[:-1]
        x2 = x.reshape(-1, spec.d_in).contiguous()
        n = x2.shape[0]
        if self._pats is None:
            # path chosen on the GLOBAL problem so that every rank runs the same kernel family
            self._path = self.backend.resolve_path(spec, n, self.n_patterns)
            # (every pass is its own T = 0 call on this shard: plain 16-bit update operands, kt16 when T > 0)
            self._pats = self.backend.prepare(self.pack, spec.flags, self._path, spec.d_assoc, spec.d_v,
                                              min(spec.update_steps, 1), spec.ln_eps)
        flags = spec.flags & ~cabi.F_OUT_PROJ
        state, st_flags = x2, flags
        ln_w, ln_b = self.pack.get("ln_state_w"), self.pack.get("ln_state_b")
        if self.scatter_rows and spec.update_steps > 0:
            raise NotImplementedError("scatter_rows with retrieve updates (every pass needs all rows' state)")
        for t in range(spec.update_steps):
            state, _, _, _ = self._pass(state, ln_w, ln_b, st_flags, True, None)   # q <- A K^ (beta: next pass)
            st_flags = flags & ~cabi.F_NORM_STATE                        # upstream does not re-normalise the state
            ln_w = ln_b = None
        fold = None
        if spec.flags & cabi.F_OUT_PROJ:
            fold = (self.pack["wo"].detach().contiguous(),
                    None if self.pack.get("bo") is None else self.pack["bo"].detach().contiguous())
        o, g_lse, g_arg, folded = self._pass(state, ln_w, ln_b, st_flags, False, fold)
        self.last_lse, self.last_argmax = g_lse, g_arg
        if fold is not None and not folded:
            o = self.backend.out_projection(o, fold[0], fold[1])
        if self.scatter_rows:
            return o
        return o.reshape(*lead, o.shape[-1])

    __call__ = forward

    def close(self) -> None:
        """Unmap the peers' exchange blocks (collective: every rank calls it; a barrier first, so that no peer still
        reads this rank's block).  The object can be used again afterwards (the exchanges are rebuilt on demand)."""
        if self._peer:
            torch.cuda.synchronize()
            if dist.is_initialized():
                dist.barrier(group=self.group)
            for ex in self._peer.values():
                ex.close()
            self._peer.clear()
