"""CPU (`-m "not gpu"`): the multi-GPU host logic under `gloo`, world_size 2 and 3.

The CUDA kernels cannot run here, so the arithmetic backend of `PatternShardedHopfield` is replaced by an
oracle-backed stand-in (tests may use the oracle); what is exercised is the product's sharding, offsets,
exchange (all_gather of per-row statistics, LSE merge weights, all_reduce / reduce_scatter) and the per-pass
merge for retrieve updates.
"""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


class OracleBackend:
    """CPU stand-in with the same interface as sharded.CudaBackend (checker arithmetic, fp32 torch)."""

    def resolve_path(self, spec, n, p):
        return 1

    def prepare(self, pack, flags, path, d_assoc, d_v, update_steps, ln_eps):
        import torch.nn.functional as F
        w = pack["lookup"]
        wt = pack.get("lookup_targets")
        wt = w if wt is None else wt
        k = F.layer_norm(w, (w.shape[1],), pack["ln_stored_w"], pack["ln_stored_b"], ln_eps) if flags & 2 else w
        v = F.layer_norm(wt, (wt.shape[1],), pack["ln_proj_w"], pack["ln_proj_b"], ln_eps) if flags & 4 else wt
        if flags & 32:
            v = F.linear(v, pack["wv"], pack["bv"])
        e = torch.empty(0)
        return [k, v, e, e, e, e]

    fold_ok = False

    def retrieve(self, x, pats, ln_w, ln_b, flags, path, d_in, beta, ln_eps, pattern_offset, over_keys, fold=None):
        import torch.nn.functional as F
        k, v = pats[0], pats[1]
        if over_keys:
            v = k
        q = F.layer_norm(x, (d_in,), ln_w, ln_b, ln_eps) if flags & 1 else x
        s = (beta * q) @ k.t()
        lse = torch.logsumexp(s, -1)
        a = torch.softmax(s, -1)
        rowmax, arg = s.max(-1)
        o = a @ v
        folded = fold is not None and self.fold_ok and not over_keys
        if folded:
            o = torch.nn.functional.linear(o, fold[0], fold[1])
        return (o, lse, rowmax, (arg + pattern_offset).to(torch.int32)), folded

    def merge(self, retrieved, lse_all, rowmax_all, argmax_all, shard):
        lse = torch.logsumexp(lse_all, 0)
        retrieved.mul_(torch.exp(lse_all[shard] - lse)[:, None])
        best = rowmax_all.max(0).values
        cand = torch.where(rowmax_all == best[None], argmax_all.long(), torch.full_like(argmax_all.long(), 2 ** 31 - 1))
        return lse, cand.min(0).values.to(torch.int32)

    def out_projection(self, retrieved, wo, bo):
        return torch.nn.functional.linear(retrieved, wo, bo)


def _worker(rank, world, port, T, scatter, fold, ret):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import hflayers
        from oracle import hflayers_oracle as O
        from util import randomize_
        impl = hflayers._impl
        kw = dict(input_size=32, quantity=300, output_size=20, update_steps_max=T, update_steps_eps=0.0,
                  stored_pattern_as_static=True, state_pattern_as_static=True)
        ref = randomize_(O.HopfieldLayer(**kw), 7, 1.0)
        full = impl.HopfieldLayer(**kw)                      # parameters only; never run on the CPU
        full.load_state_dict(ref.state_dict())
        be = OracleBackend()
        be.fold_ok = fold
        sh = impl.PatternShardedHopfield(full, backend=be, scatter_rows=scatter)
        assert (sh.start, sh.count) == impl.shard_range(300, world, rank)
        x = torch.randn(12 * world, 32, generator=torch.Generator().manual_seed(3))
        with torch.no_grad():
            want, _, scores, _ = ref(x[None], return_details=True)
            got = sh(x)
        want = want[0]
        if scatter:
            r0, cnt = impl.batch_parallel_rows(x.shape[0], world, rank)
            want = want[r0:r0 + cnt]
        err = float((got - want).norm() / want.norm())
        ok_arg = bool(torch.equal(sh.last_argmax.long(), scores.reshape(x.shape[0], -1).argmax(-1))) if T == 0 else True
        ret[rank] = (err, ok_arg)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,T,scatter,fold", [(2, 0, False, False), (2, 0, True, True), (3, 0, False, True),
                                                  (2, 2, False, False)])
def test_pattern_sharded_exchange_gloo(world, T, scatter, fold):
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, T, scatter, fold, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    for r in range(world):
        err, ok_arg = ret[r]
        assert err < 1e-5, (r, err)
        assert ok_arg


def test_shard_ranges_cover_and_align():
    sys.path.insert(0, ROOT)
    import hflayers
    impl = hflayers._impl
    for p, g in [(65536, 8), (1000, 3), (300, 2), (130, 4), (7, 7), (1 << 20, 8)]:
        got = [impl.shard_range(p, g, i) for i in range(g)]
        assert got[0][0] == 0 and sum(c for _, c in got) == p
        for (s0, c0), (s1, _) in zip(got, got[1:]):
            assert s0 + c0 == s1 and c0 > 0
        if p >= 128 * g:
            assert all(s % 128 == 0 for s, _ in got)
    rows = [impl.batch_parallel_rows(10, 4, r) for r in range(4)]
    assert rows == [(0, 3), (3, 3), (6, 2), (8, 2)]
