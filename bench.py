#!/usr/bin/env python
"""bench.py -- retrieved latents/sec of the fused Hopfield retrieval on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

One "step" = one forward pass of the Hopfield association over one batch of synthetic queries
(SURVEY.md section 8d).  Default workload = the smallest point of BASELINE.json config 5 (the one the
"% of bf16 tensor peak" target is quoted on): N = 16,384 query rows per GPU, P = 65,536 stored
patterns, D = Dv = 256, T = 0 extra update steps, beta = 1/sqrt(D).  With --gpus N the latent batch
is data-parallel (every rank holds the patterns and its own N rows; no data-path collective; weak
scaling).  Rank 0 prints ONE JSON line.  That line also carries
  parity     rel-L2 and nearest-pattern mismatches of the device result against the CPU checker (oracle/chunked.py)
             on a row subsample, computed outside the timed regions;
  roofline   the fused kernel alone through the C-ABI: burst (L2 flushed, one launch per event pair) and `sustained`
             (>= 2 s of back-to-back launches, clocks recorded);
  prep       the one-off pattern-side preparation (K^ / V^ packs) that `value` and `e2e` do NOT include -- it is
             cached per parameter version (inference), see DESIGN.md 6;
  sharded    (N > 1 only) BASELINE config 5's pattern-sharded mode on the same ranks: P = 256k / 1M split P/N per
             GPU, fused peer-memory and NCCL exchanges, strong scaling against the one-GPU time, parity.

`--impl reference` times the reference's own CPU path for the same metric on a bounded sample of
the same workload: PyTorch-eager fp32 through the restated `hflayers` (oracle/; upstream
hopfield-layers cannot be installed offline -- see DESIGN.md), all host threads.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (N rows per GPU, P patterns, D = Dv, T update steps)
    "sweep_p64k_d256": (16384, 65536, 256, 0),
    "sweep_p64k_d256_t1": (16384, 65536, 256, 1),
    "sweep_p64k_d128": (16384, 65536, 128, 0),
    "sweep_p64k_d64": (16384, 65536, 64, 0),
    "sweep_p256k_d256": (16384, 262144, 256, 0),
    "sweep_p64k_d512": (16384, 65536, 512, 0),
    "sweep_p64k_d512_t1": (16384, 65536, 512, 1),
    "sweep_p256k_d512": (16384, 262144, 512, 0),
    "sweep_p64k_d1024": (16384, 65536, 1024, 0),
    "sweep_p64k_d1024_t1": (16384, 65536, 1024, 1),
    "ffhq32_layer": (2592, 512, 64, 0),       # one `hopfield` layer of FFHQ-32 (launch-bound; context only)
    "tiny": (1024, 4096, 64, 0),              # plumbing check
}
DEFAULT_WORKLOAD = "sweep_p64k_d256"
METRIC = "retrieved latents/sec (Hopfield fwd, iters=cfg)"
UNIT = "latents/s"


def fwd_flops(n, p, d, dv, t):
    """Algorithmic FLOPs of one forward: T+1 score GEMMs, T state updates, 1 read-out."""
    return 2.0 * n * p * ((2 * t + 1) * d + dv)


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            j = json.load(f)
        return dict(bf16_tflops=float(j["bf16_tflops"]), bf16_tflops_sustained=float(j.get("bf16_tflops_sustained", 0)),
                    hbm_gbs=float(j["hbm_gbs"]), source="measured (MEASURED_PEAKS.json)")
    return dict(bf16_tflops=1590.0, bf16_tflops_sustained=1400.0, hbm_gbs=6650.0,
                source="fallback (B200_PROFILING.md)")


def make_state(p, d, t, seed=1265):
    """Synthetic config-5 layer: W ~ N(0,1), LN affine identity, Wv ~ N(0, 1/D)."""
    import torch
    g = torch.Generator().manual_seed(seed)
    w = torch.randn(p, d, generator=g)
    g2 = torch.Generator().manual_seed(seed + 1)
    wv = torch.randn(d, d, generator=g2) / math.sqrt(d)
    wo = torch.randn(d, d, generator=g2) / math.sqrt(d)
    return w, wv, wo


def make_queries(w, n, seed=1267):
    import torch
    g = torch.Generator().manual_seed(seed)
    idx = torch.randint(0, w.shape[0], (n,), generator=g)
    return w[idx] + 0.5 * torch.randn(n, w.shape[1], generator=g), idx


def build_layer(cls, p, d, t, w, wv, wo):
    import torch
    layer = cls(input_size=d, quantity=p, stored_pattern_as_static=True, state_pattern_as_static=True,
                update_steps_max=t)
    with torch.no_grad():
        layer.lookup_weights.copy_(w.unsqueeze(0))
        core = layer.hopfield.association_core
        core.in_proj_weight.copy_(wv)
        core.in_proj_bias.zero_()
        core.out_proj.weight.copy_(wo)
        core.out_proj.bias.zero_()
    return layer.eval()


# ------------------------------------------------------------------------------------------ CPU arms
def cpu_reference_run(n, p, d, t, steps, warmup, budget_s=12.0):
    """Reference CPU path (restated hflayers, PyTorch eager fp32, all host threads) on a bounded sample
    of the workload's query rows; returns (latents/s, sample description, cores, ms per step, rows)."""
    import torch
    from oracle import hflayers_oracle as O          # checker / CPU baseline only
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    w, wv, wo = make_state(p, d, t)
    layer = build_layer(O.HopfieldLayer, p, d, t, w, wv, wo)
    # size the sample: probe with 64 rows, then scale to about budget_s of total CPU work
    xq, _ = make_queries(w, min(n, 4096))
    with torch.no_grad():
        t0 = time.perf_counter(); layer(xq[:64]); probe = time.perf_counter() - t0
        t0 = time.perf_counter(); layer(xq[:64]); probe = min(probe, time.perf_counter() - t0)
    per_row = probe / 64
    total_steps = max(1, steps + warmup)
    rows = int(max(64, min(xq.shape[0], budget_s / total_steps / max(per_row, 1e-9))))
    x = xq[:rows].contiguous()
    with torch.no_grad():
        for _ in range(warmup):
            layer(x)
        t0 = time.perf_counter()
        for _ in range(steps):
            layer(x)
        dt = time.perf_counter() - t0
    ms = dt / steps * 1e3
    sample = (f"{rows} of {n} query rows per step, all P={p} patterns, D={d}, T={t}; oracle restatement of hflayers "
              f"(upstream hopfield-layers not installable offline), torch eager fp32, {cores} threads")
    return rows * steps / dt, sample, cores, ms, rows


class ClockSampler:
    """Samples nvidia-smi SM clocks / throttle reasons while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._read, daemon=True)
            self.thr.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [s.strip() for s in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------ GPU arm
class RawForward:
    """Direct C-ABI call (include/hop_sm100.h) of the association without out_proj on preallocated
    buffers: isolates the fused retrieval kernel for the roofline timing."""

    def __init__(self, impl, layer, x_dev, path):
        import ctypes as C
        import torch
        cabi = impl.cabi
        self.C, self.cabi, self.lib = C, cabi, cabi.load()
        spec = layer.spec()
        n, p = x_dev.shape[0], layer.quantity
        flags = spec.flags & ~cabi.F_OUT_PROJ
        self.desc = cabi.make_desc(flags=flags, path=path, n_rows=n, n_patterns=p, d_in=spec.d_in,
                                   d_assoc=spec.d_assoc, d_v=spec.d_v, d_out=1, update_steps=spec.update_steps,
                                   beta=spec.beta, ln_eps=spec.ln_eps, tune_splits=spec.tune_splits,
                                   tune_flags=spec.tune_flags)
        pack = layer.param_pack()
        self.keep = pack
        self.pats_t = list(layer._prepared(pack, spec, path))
        self.pats_t[5] = self.pats_t[5].new_empty((0,))          # no out-projection here: no folded operand
        dev = x_dev.device
        self.x = x_dev
        self.out = torch.empty((n, spec.d_v), device=dev)
        self.lse = torch.empty((n,), device=dev)
        self.rowmax = torch.empty((n,), device=dev)
        self.argmax = torch.empty((n,), device=dev, dtype=torch.int32)
        need = C.c_size_t(0)
        cabi.check(self.lib.hop_sm100_workspace_bytes(C.byref(self.desc), 0, C.byref(need)), "workspace_bytes")
        self.ws = torch.empty((max(int(need.value), 16),), device=dev, dtype=torch.uint8)
        self.ws_bytes = int(need.value)
        self.params = cabi.HopParams(ln_state_w=cabi.ptr(pack["ln_state_w"]), ln_state_b=cabi.ptr(pack["ln_state_b"]))
        self.pats = cabi.patterns_struct(self.pats_t)

    def __call__(self):
        import torch
        C, cabi = self.C, self.cabi
        st = torch.cuda.current_stream().cuda_stream
        cabi.check(self.lib.hop_sm100_forward(C.byref(self.desc), C.byref(self.params), C.byref(self.pats),
                                              cabi.ptr(self.x), cabi.ptr(self.out), None, None, cabi.ptr(self.lse),
                                              cabi.ptr(self.argmax), cabi.ptr(self.rowmax), cabi.ptr(self.ws),
                                              self.ws_bytes, st), "hop_sm100_forward")


def timed_steps(fn, steps, flush, torch):
    """Runs fn() `steps` times, L2 flushed (outside the timed events) before each; returns per-step ms."""
    evs = []
    for _ in range(steps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        evs.append((e0, e1))
    torch.cuda.synchronize()
    return [a.elapsed_time(b) for a, b in evs]


def traffic_of(workload, world, sharded):
    """DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture of this workload
    (profiles/traffic.json names the capture), or None when none was taken (other workloads, sharded runs)."""
    if sharded:
        return None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            e = json.load(f).get(workload)
        return None if e is None else {"dram_bytes_per_launch": int(e["bytes_per_launch"]), "unit": "B",
                                       "source": e["source"]}
    except (OSError, ValueError, KeyError):
        return None


def bind_to_gpu_numa_node(index):
    """Multi-rank runs: pin this process to the CPUs next to its GPU (NVML's ideal affinity) BEFORE the pinned
    host buffers are allocated, so that they land on that NUMA node and 8 ranks do not all pull their host-side
    copies through one socket.  Best effort: returns the CPU count bound to, or None."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        n_cpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (n_cpu + 63) // 64)
        cpus = {64 * w + b for w, m in enumerate(words) for b in range(64) if (m >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:  # noqa: BLE001  (no NVML / not permitted: keep the default placement)
        pass
    return None


def parity_block(y_dev, arg_dev, xh, w, wv, wo, t, d, sub, world_note=""):
    """Device result vs the CPU checker on the first `sub` rows (outside every timed region): rel-L2 of the layer
    output and nearest-pattern mismatches.  A row is excluded from the index comparison only where the fp32 checker
    itself cannot decide it (gap below its rounding; with updates: below the 16-bit rounding the state carries)."""
    import torch
    from oracle.chunked import chunked_oracle          # the checker, never the product
    got, got_arg = y_dev[:sub].float().cpu(), arg_dev[:sub].cpu().long()
    want, arg, gap = chunked_oracle(xh[:sub], w, wv, wo, t, 1.0 / math.sqrt(d))
    decidable = gap > (1e-4 if t == 0 else 2e-2)
    rel = float((got.double() - want.double()).norm() / want.double().norm())
    return {"rel_l2": rel, "bar": 1e-3, "rows_checked": int(sub), "argmax_rows_compared": int(decidable.sum()),
            "argmax_mismatches": int(((got_arg != arg) & decidable).sum()),
            "against": "oracle/chunked.py (P-chunked fp32/fp64 restatement of hflayers, CPU)" + world_note}


def sustained_kernel(raw, flush, fl, min_seconds, torch):
    """>= `min_seconds` of back-to-back launches of the fused kernel (C-ABI), the L2 flush inside the loop and its own
    cost (timed the same way) subtracted: the figure under the board's power cap, next to the burst one."""
    def loop(fn, reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1)
    probe = loop(lambda: (flush.zero_(), raw()), 5) / 5
    reps = max(10, int(min_seconds * 1e3 / max(probe, 1e-3)))
    total = loop(lambda: (flush.zero_(), raw()), reps)
    only_flush = loop(lambda: flush.zero_(), reps)
    k_ms = (total - only_flush) / reps
    return {"seconds": total * 1e-3, "launches": reps, "kernel_ms": k_ms, "achieved": fl / (k_ms * 1e-3) / 1e12}


def backward_record(layer_cls, impl, dev, n, p, d, w, wv, wo, pk, torch, steps=5):
    """Training step of the same layer (forward + `backward()`, main.py:35) on this GPU: device ms of the backward and
    its algorithmic rate F_bwd = 2 N P (3 D + 2 Dv) against the measured peak.  Everything the backward
    enqueues is inside: stats / operand packs, the three tcgen05 core launches, the fp32 LayerNorm / projection /
    weight-gradient kernels.  Gradient parity of this path: tests/test_backward_tc_gpu.py."""
    layer = build_layer(layer_cls, p, d, 0, w, wv, wo).to(dev).train()
    layer.kernel_path = "f16"
    xh, _ = make_queries(w, n)
    x = xh.to(dev).requires_grad_(True)
    gout = torch.randn(n, d, device=dev)
    path = impl.ops.resolve_train_path(layer.spec(), n, p)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    t_f, t_b = [], []
    for it in range(steps + 2):
        layer.zero_grad(set_to_none=True)
        x.grad = None
        flush.zero_()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record(); y = layer(x); e[1].record(); y.backward(gout); e[2].record()
        torch.cuda.synchronize()
        if it >= 2:
            t_f.append(e[0].elapsed_time(e[1])); t_b.append(e[1].elapsed_time(e[2]))
    ms_f, ms_b = sorted(t_f)[len(t_f) // 2], sorted(t_b)[len(t_b) // 2]
    f_bwd = 2.0 * n * p * 5 * d
    ach = f_bwd / (ms_b * 1e-3) / 1e12
    return {"train_path": {1: "fp32_simt", 2: "tc_f16", 3: "tc_bf16"}[path], "ms_backward": ms_b,
            "ms_forward_incl_prep": ms_f, "algorithmic_flops": f_bwd, "achieved": ach, "unit": "TFLOP/s",
            "peak": pk["bf16_tflops"], "frac": ach / pk["bf16_tflops"],
            "executed_over_algorithmic": 8.0 / 5.0 if path != 1 else 7.0 / 5.0,
            "note": "whole backward of the layer (core + fp32 tails), median of %d steps, L2 flushed" % steps}


SHARDED_POINTS = [(262144, 256), (1048576, 256), (1048576, 512)]      # (P, D) of BASELINE config 5, T = 0


def sharded_record(impl, hflayers, dev, rank, world, n, steps, torch, dist):
    """BASELINE config 5's pattern-sharded mode on the ranks of this run: P split P/world per GPU, all ranks see the
    same n rows; per point the step time with the fused peer-memory exchange and with the NCCL form, the local
    kernel alone, the one-GPU time of the unsharded layer (every rank runs it; rank 0's number), parity."""
    out = []
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def timed(fn, reps):
        ts = []
        for _ in range(reps):
            flush.zero_()
            torch.cuda.synchronize(); dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        tt = torch.tensor([sorted(ts)[len(ts) // 2]], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return float(tt.item())

    for (p, d) in SHARDED_POINTS:
        w, wv, wo = make_state(p, d, 0)
        layer = build_layer(hflayers.HopfieldLayer, p, d, 0, w, wv, wo).to(dev)
        layer.kernel_path = "f16"
        xh, _ = make_queries(w, n, seed=1267)
        x = xh.to(dev)
        rec = {"patterns": p, "dim": d, "update_steps": 0, "rows": n, "patterns_per_gpu": None}
        with torch.no_grad():
            for _ in range(2):
                layer(x)
            single_ms = timed(lambda: layer(x), max(3, steps // 2))
            layer.invalidate()                       # drop the full-size packs before the shards are prepared
            torch.cuda.empty_cache()
            y = None
            for exchange in ("peer", "collective"):
                runner = impl.PatternShardedHopfield(layer, exchange=exchange)
                rec["patterns_per_gpu"] = runner.count
                for _ in range(3):
                    y = runner(x)
                ms = timed(lambda: runner(x), steps)
                rec[f"ms_per_step_{exchange}"] = ms
                if exchange == "peer":
                    y_keep, arg_keep = y.clone(), runner.last_argmax.clone()
                    # the local fused kernel alone on this rank's shard (same operands, no exchange)
                    spec = runner.spec
                    flags = spec.flags & ~impl.cabi.F_EXACT_ARGMAX

                    def local():
                        runner.backend.retrieve(x, runner._pats, runner.pack.get("ln_state_w"),
                                                runner.pack.get("ln_state_b"), flags & ~impl.cabi.F_OUT_PROJ,
                                                runner._path, spec.d_in, spec.beta, spec.ln_eps, runner.start, False,
                                                (runner.pack["wo"].detach(), runner.pack["bo"].detach()))
                    for _ in range(2):
                        local()
                    rec["ms_local_kernel"] = timed(local, steps)
                    # the fused exchange kernel alone (the partials of the last step are still in the peer buffers)
                    ex = next(iter(runner._peer.values()), None)
                    if ex is not None:
                        r0, rc = impl.batch_parallel_rows(n, world, rank)
                        rec["exchange_us_peer_kernel"] = timed(lambda: ex.merge(r0, rc, push_all=True), steps) * 1e3
                runner.close()
                del runner
            rec["ms_one_gpu"] = single_ms
            # (step minus local kernel: meaningful while the kernel is short; at P = 1M the two timings differ by more than
            # the exchange costs -- the board's power cap moves the clocks between them -- hence the direct figure above)
            rec["exchange_us_peer_by_difference"] = (rec["ms_per_step_peer"] - rec["ms_local_kernel"]) * 1e3
            rec["exchange_us_collective_by_difference"] = (rec["ms_per_step_collective"] - rec["ms_local_kernel"]) * 1e3
            rec["speedup_vs_one_gpu"] = single_ms / rec["ms_per_step_peer"]
            rec["strong_scaling_efficiency"] = single_ms / rec["ms_per_step_peer"] / world
            # bytes a rank moves over NVLink per step with the fused exchange: it pulls its n/world rows of
            # (O, lse, max, arg) from world-1 peers and pushes the merged rows (O, lse, arg) to world-1 peers
            rows_mine = n // world
            rec["nvlink_bytes_per_rank_per_step"] = (world - 1) * rows_mine * ((d + 3) * 4 + (d + 2) * 4)
            rec["partial_bytes_per_shard"] = 4 * n * (d + 3)
            rec["latents_per_s"] = n / (rec["ms_per_step_peer"] * 1e-3)
            rec["algorithmic_tflops_all_gpus"] = fwd_flops(n, p, d, d, 0) / (rec["ms_per_step_peer"] * 1e-3) / 1e12
            if rank == 0:
                sub = 32 if p * d > (1 << 28) else 64
                rec["parity"] = parity_block(y_keep.reshape(n, d), arg_keep, xh, w, wv, wo, 0, d, sub,
                                             f"; result of the {world}-way sharded run, peer exchange")
        out.append(rec)
        del layer, x, w, wv, wo, y
        torch.cuda.empty_cache()
        dist.barrier()
    return out


def gpu_run(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the Hopfield retrieval op is sm_100a only; no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = bind_to_gpu_numa_node(local) if world > 1 else None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import hflayers
    impl = hflayers._impl
    cabi = impl.cabi
    cabi.load()
    n, p, d, t = WORKLOADS[args.workload]
    w, wv, wo = make_state(p, d, t)
    layer = build_layer(hflayers.HopfieldLayer, p, d, t, w, wv, wo).to(dev)
    layer.kernel_path = args.path
    x_host, _ = make_queries(w, n, seed=1267 + rank)
    x_host = x_host.contiguous().pin_memory()
    out_host = torch.empty((n, d), dtype=torch.float32).pin_memory()
    x_dev = x_host.to(dev)
    x_stage = torch.empty_like(x_dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    path = impl.resolve_path(layer.spec(), n, p)
    sharded = args.mode == "sharded"
    if sharded:
        # pattern-sharded: every rank holds P/world patterns and sees the SAME n query rows (strong scaling in P)
        x_host, _ = make_queries(w, n, seed=1267)
        x_host = x_host.contiguous().pin_memory()
        x_dev = x_host.to(dev)
        runner = impl.PatternShardedHopfield(layer, exchange=args.exchange)
        p_local = runner.count
        sub = build_layer(hflayers.HopfieldLayer, p_local, d, t, w[runner.start:runner.start + p_local], wv, wo).to(dev)
        sub.kernel_path = {1: "fp32", 2: "f16", 3: "bf16"}[path]
        raw = RawForward(impl, sub, x_dev, path)
    else:
        runner = layer
        raw = RawForward(impl, layer, x_dev, path)

    e2e_chunks = args.e2e_chunks
    if isinstance(e2e_chunks, str):
        e2e_chunks = tuple(float(v) for v in e2e_chunks.split(",")) if "," in e2e_chunks else int(e2e_chunks)

    def step_dev():
        with torch.no_grad():
            return runner(x_dev)

    def step_e2e():
        # the reference-facing call with HOST buffers: pinned x in, pinned result out, copies inside the timed region
        with torch.no_grad():
            if sharded:
                x_stage.copy_(x_host, non_blocking=True)
                y = runner(x_stage)
                out_host.copy_(y.reshape(n, d), non_blocking=True)
            else:
                layer.forward_host(x_host, out_host, chunks=e2e_chunks)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def max_over_ranks(v):
        if world == 1:
            return v
        tt = torch.tensor([v], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return float(tt.item())

    # one-off pattern-side preparation (cached per parameter version; outside `value` and `e2e`): the SECOND call of the
    # process (the first one also pays the lazy loading of its kernels), caches cold (L2 flushed)
    prep_ms = None
    if not sharded:
        with torch.no_grad():
            layer.invalidate()
            layer._prepared(layer.param_pack(), layer.spec(), path)
            layer.invalidate()
            torch.empty(256 << 20, dtype=torch.uint8, device=dev).zero_()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); layer._prepared(layer.param_pack(), layer.spec(), path); e1.record()
        torch.cuda.synchronize()
        prep_ms = e0.elapsed_time(e1)
        raw = RawForward(impl, layer, x_dev, path)

    for _ in range(max(args.warmup, 3)):
        step_dev(); raw(); step_e2e()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = cabi.launch_count()
    barrier()
    ms_steps = timed_steps(step_dev, args.steps, flush, torch)
    barrier()
    launches = cabi.launch_count() - l0
    ms_total = max_over_ranks(sum(ms_steps))
    ms_kernel = timed_steps(raw, args.steps, flush, torch)
    barrier()
    ms_e2e_total = max_over_ranks(sum(timed_steps(step_e2e, args.steps, flush, torch)))
    barrier()
    clocks = sampler.stop() if rank == 0 else None

    fl = fwd_flops(n, runner.count if sharded else p, d, d, t)
    sustained = None
    if not args.no_sustained:
        sus_sampler = ClockSampler(local)
        if rank == 0:
            sus_sampler.start()
        sustained = sustained_kernel(raw, flush, fl, args.sustained_seconds, torch)
        barrier()
        if rank == 0:
            sustained["clocks"] = sus_sampler.stop()

    parity = None
    if rank == 0 and not args.no_parity:
        with torch.no_grad():
            y = step_dev()
        arg = runner.last_argmax
        parity = parity_block(y.reshape(n, -1), arg, x_host, w, wv, wo, t, d, 64 if p * d <= (1 << 26) else 32)

    bwd_rec = None
    if rank == 0 and world == 1 and not sharded and t == 0 and not args.no_backward:
        try:
            bwd_rec = backward_record(hflayers.HopfieldLayer, impl, dev, n, p, d, w, wv, wo, peaks(), torch)
        except Exception as e:  # noqa: BLE001
            bwd_rec = {"error": f"{type(e).__name__}: {e}"[:400]}

    shard_rec = None
    if world > 1 and not sharded and not args.no_sharded_record:
        del raw
        layer.invalidate()
        torch.cuda.empty_cache()
        try:
            shard_rec = sharded_record(impl, hflayers, dev, rank, world, n, max(3, args.steps // 2), torch, dist)
        except Exception as e:  # noqa: BLE001  (the headline line must survive a failure of the side record)
            shard_rec = [{"error": f"{type(e).__name__}: {e}"[:500]}]

    if rank == 0:
        pk = peaks()
        ms_step = ms_total / args.steps
        rows_total = n if sharded else world * n
        value = rows_total * args.steps / (ms_total * 1e-3)
        e2e = rows_total * args.steps / (ms_e2e_total * 1e-3)
        k_ms = sum(ms_kernel) / len(ms_kernel)
        achieved = fl / (k_ms * 1e-3) / 1e12
        dtype_name = {1: "f32", 2: "f16", 3: "bf16"}[path]
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "strong" if sharded else "weak",
            "vs_baseline": None, "dtype": dtype_name + " operands, f32 accumulate",
            "dtype_note": ("fp16-encoded operands replace the north star's bf16: same 16-bit width and tensor-pipe rate, "
                           "11 instead of 8 significand bits -- plain bf16 operands miss the 1e-3 parity bar "
                           "(SURVEY.md B.3, DESIGN.md 4.1)") if path == 2 else None,
            "data": "synthetic (config-5 recipe, SURVEY.md 8d; seeds 1265/1266/1267+rank)",
            "config": {"workload": args.workload, "rows_per_gpu": n, "patterns": p, "dim": d, "update_steps": t,
                       "beta": 1.0 / math.sqrt(d),
                       "parallelism": (f"pattern-sharded{world} (P/{world} per GPU, " +
                                       ("fused peer-memory LSE merge over NVLink)" if getattr(runner, "use_peer", False)
                                        else "NCCL LSE merge)") if sharded else f"batch-dp{world}"), "kernel_path":
                       {1: "fp32_simt", 2: "tc_f16", 3: "tc_bf16"}[path],
                       "host_numa_cpus": numa, "l2": "flushed between timed steps (256 MiB memset outside the events)"},
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": x_host.numel() * 4,
                    "d2h_bytes_per_step": out_host.numel() * 4,
                    "api": ("PatternShardedHopfield.forward on a staged copy" if sharded else
                            "HopfieldLayer.forward_host (pinned host in/out, H2D / kernel / D2H pipelined in row chunks)")},
            "gpu_launches": int(launches),
            "roofline": {"bound": "tensor", "achieved": achieved, "peak": pk["bf16_tflops"], "unit": "TFLOP/s",
                         "frac": achieved / pk["bf16_tflops"], "traffic": traffic_of(args.workload, world, sharded),
                         "kernel_ms": k_ms, "algorithmic_flops": fl, "peak_source": pk["source"],
                         "timing": "burst: one C-ABI forward per CUDA-event pair, L2 flushed before each",
                         "includes": "fused forward + stream-K / split merge + exact-argmax certificate (and repair "
                                     "launch) -- everything hop_sm100_forward enqueues"},
            "clocks": clocks,
        }
        if sustained is not None:
            sp = pk["bf16_tflops_sustained"]
            sustained.update({"unit": "TFLOP/s", "peak_sustained": sp,
                              "frac_of_sustained_peak": (sustained["achieved"] / sp) if sp else None,
                              "frac_of_burst_peak": sustained["achieved"] / pk["bf16_tflops"],
                              "method": "back-to-back C-ABI forwards for the stated seconds, L2 flush inside the loop, "
                                        "the flush loop's own time subtracted"})
            line["roofline"]["sustained"] = sustained
        if prep_ms is not None:
            ppad = (p + 127) // 128 * 128
            prep_bytes = 4 * p * d * 2 + 4 * p * d * 2 + 2 * ppad * d * (3 if t == 0 else 4) + 4 * p * d
            line["prep"] = {"ms": prep_ms, "in_value": False, "in_e2e": False,
                            "what": "hop_sm100_prepare_patterns: LN_k / LN_v / V-projection / out_proj fold -> fp32 K^, V^ "
                                    "and the 16-bit packs; once per parameter version (cached for inference)",
                            "algorithmic_bytes": prep_bytes, "gbs": prep_bytes / (prep_ms * 1e-3) / 1e9,
                            "hbm_peak_gbs": pk["hbm_gbs"]}
        if parity is not None:
            line["parity"] = parity
        if bwd_rec is not None:
            line["backward"] = bwd_rec
        if shard_rec is not None:
            line["sharded"] = {"mode": f"pattern-sharded over {world} GPUs (BASELINE config 5), T = 0, "
                                       "f16 operands; times = median of max-over-ranks device times, L2 flushed",
                               "points": shard_rec}
        if world == 1 and not args.no_cpu_baseline:
            v, sample, cores, ms, rows = cpu_reference_run(n, p, d, t, steps=2, warmup=1, budget_s=15.0)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def reference_run(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    n, p, d, t = WORKLOADS[args.workload]
    v, sample, cores, ms, rows = cpu_reference_run(n, p, d, t, steps=args.steps, warmup=args.warmup, budget_s=60.0)
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic (config-5 recipe, SURVEY.md 8d)",
        "config": {"workload": args.workload, "rows_per_gpu": n, "patterns": p, "dim": d, "update_steps": t,
                   "beta": 1.0 / math.sqrt(d), "parallelism": "host cores", "sample_rows_per_step": rows},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--exchange", default="auto", choices=["auto", "peer", "collective"],
                    help="--mode sharded: fused peer-memory exchange kernel (default on CUDA) or NCCL collectives")
    ap.add_argument("--path", default="auto", choices=["auto", "fp32", "f16", "bf16"])
    ap.add_argument("--mode", default="batch", choices=["batch", "sharded"],
                    help="multi-GPU mode: batch-parallel rows (default, weak scaling) or pattern-sharded with an "
                         "NCCL log-sum-exp merge (strong scaling in P)")
    ap.add_argument("--e2e-chunks", default=None,
                    help="row chunks of the pipelined host-buffer call: a count (equal pieces) or weights like 1,3,3,1")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-sustained", action="store_true")
    ap.add_argument("--no-backward", action="store_true")
    ap.add_argument("--sustained-seconds", type=float, default=2.0)
    ap.add_argument("--no-sharded-record", action="store_true",
                    help="multi-GPU batch-parallel runs: skip the pattern-sharded sub-record (config 5 points)")
    args = ap.parse_args()
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if args.impl == "reference":
        reference_run(args)
    else:
        gpu_run(args)


if __name__ == "__main__":
    main()
