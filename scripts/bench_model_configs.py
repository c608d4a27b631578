"""Latency of the three-layer Hopfield bottleneck at the reference's own model shapes (BASELINE.json
configs 1-3: MNIST-28, FFHQ-32, FFHQ-64; B = 32, P = 512, D = 64/64/3) on one B200:

  ours      hflayers.HopfieldLayer (sm_100a library, exact fp32 kernels; HOP_PATH_AUTO)
  eager     the oracle restatement run on the SAME GPU with PyTorch's stock kernels (ATen LayerNorm /
            softmax + cuBLAS, TF32 off) -- what the reference itself does on a CUDA device, both in the
            de-duplicated form and in the upstream-faithful form (patterns expanded over the batch)

These sizes are launch-latency bound (<= 2.5 GFLOP): the numbers reported are microseconds per bottleneck
(forward, and forward+backward), library launches per call, and retrieved latents/s -- not a roofline.

    python scripts/bench_model_configs.py            # on the GPU box
"""
import copy
import json
import os
import statistics
import sys
import time

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import hflayers  # noqa: E402
import hopvae_harness as H  # noqa: E402
from oracle import hflayers_oracle as O  # noqa: E402  (baseline / checker only)

torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = False
impl = hflayers._impl
dev = torch.device("cuda", 0)


def timed(fn, iters=50, warm=10):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3   # microseconds


def timed_cpu(fn, iters=10, warm=3):
    """Median wall time (us) of a CPU callable: the reference's CPU path timed beside the GPU numbers."""
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return statistics.median(ts) * 1e6


torch.set_num_threads(os.cpu_count() or 1)


def graphed(fn):
    """Capture fn() (static shapes, no host sync inside) into a CUDA graph; returns a replay callable."""
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(3):
            fn()
    torch.cuda.current_stream().wait_stream(s)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    return g.replay


rows = []
for name, cfg in (("mnist28", H.MNIST_28), ("ffhq32", H.FFHQ_32), ("ffhq64", H.FFHQ_64)):
    torch.manual_seed(1265)
    ref = H.HopVAEHarness(cfg, O.HopfieldLayer).to(dev).eval()
    dut = H.HopVAEHarness(cfg, hflayers.HopfieldLayer)
    dut.load_state_dict(ref.state_dict())
    dut = dut.to(dev).eval()
    R = cfg["representation_dim"]
    n = 32 * R * R
    z = torch.randn(32, R * R, cfg["embedding_dim"], device=dev)
    with torch.no_grad():
        a, b = dut.bottleneck(z), ref.bottleneck(z)
        rel = float((a["emb_recon"] - b["emb_recon"]).norm() / b["emb_recon"].norm())
    res = {"config": name, "rows": n, "parity_rel_l2": rel}

    def fwd(m):
        def f():
            with torch.no_grad():
                m.bottleneck(z)
        return f

    l0 = impl.cabi.launch_count()
    fwd(dut)()
    res["launches_fwd_ours"] = impl.cabi.launch_count() - l0
    res["us_fwd_ours"] = timed(fwd(dut))
    for layer in (ref.hopfield, ref.embedding_to_index, ref.index_to_embedding):
        layer.faithful = False
    res["us_fwd_eager"] = timed(fwd(ref))
    try:
        res["us_fwd_ours_graph"] = timed(graphed(fwd(dut)))
    except Exception as e:  # noqa: BLE001
        res["us_fwd_ours_graph"] = f"capture failed: {type(e).__name__}"
    # the same bottleneck as ONE launch (hop_chain.cu)
    def fused():
        with torch.no_grad():
            hflayers.bottleneck_forward(dut.hopfield, dut.embedding_to_index, dut.index_to_embedding, z, cfg["num_levels"])
    with torch.no_grad():
        c = dut.bottleneck_fused(z)
    res["fused_parity_rel_l2_emb"] = float((c["emb"] - b["emb"]).norm() / b["emb"].norm())
    res["fused_loss_vs_eager"] = [float(c["emb_recon_loss"]), float(b["emb_recon_loss"])]
    l0 = impl.cabi.launch_count()
    fused()
    res["launches_fwd_fused"] = impl.cabi.launch_count() - l0
    res["us_fwd_fused"] = timed(fused)
    try:
        res["us_fwd_fused_graph"] = timed(graphed(fused))
    except Exception as e:  # noqa: BLE001
        res["us_fwd_fused_graph"] = f"capture failed: {type(e).__name__}"
    for layer in (ref.hopfield, ref.embedding_to_index, ref.index_to_embedding):
        layer.faithful = True
    res["us_fwd_eager_upstream_shape"] = timed(fwd(ref))
    ref_cpu, z_cpu = copy.deepcopy(ref).cpu(), z.cpu()

    def fwd_cpu():
        with torch.no_grad():
            ref_cpu.bottleneck(z_cpu)
    res["us_fwd_cpu_upstream_shape"] = timed_cpu(fwd_cpu)
    res["cpu_threads"] = torch.get_num_threads()
    del ref_cpu

    def fb(m):
        zz = z.clone().requires_grad_(True)

        def f():
            m.zero_grad(set_to_none=True)
            out = m.bottleneck(zz)
            (out["emb_recon"].square().mean() + out["emb_recon_loss"]).backward()
        return f

    dut.train(); ref.train()
    res["us_fwdbwd_ours"] = timed(fb(dut), iters=30)
    res["us_fwdbwd_eager_upstream_shape"] = timed(fb(ref), iters=30)
    # the same training step with the host out of the way (whole fwd+bwd captured into a CUDA graph, both arms)
    for key, m in (("us_fwdbwd_ours_graph", dut), ("us_fwdbwd_eager_upstream_shape_graph", ref)):
        try:
            res[key] = timed(graphed(fb(m)), iters=30)
        except Exception as e:  # noqa: BLE001
            res[key] = f"capture failed: {type(e).__name__}: {str(e)[:120]}"
    res["latents_per_s_fwd_ours"] = n / (res["us_fwd_ours"] * 1e-6)
    res["latents_per_s_fwd_eager"] = n / (res["us_fwd_eager"] * 1e-6)
    rows.append(res)
    print(json.dumps(res), flush=True)


# BASELINE config 4 (HopVAE.sample after the prior, HopVAE.py:188-200): index tensor -> index_to_embedding (D = 3
# Hopfield layer) -> decoder.  S = 1 as the reference samples, S = 1024 as the synthetic batch of SURVEY.md 8d.
torch.manual_seed(1265)
ref = H.HopVAEHarness(H.MNIST_28, O.HopfieldLayer).to(dev).eval()
dut = H.HopVAEHarness(H.MNIST_28, hflayers.HopfieldLayer)
dut.load_state_dict(ref.state_dict())
dut = dut.to(dev).eval()
for S in (1, 1024):
    idx = torch.randint(0, 512, (S, 3, 8, 8), generator=torch.Generator().manual_seed(1265)).to(dev)
    with torch.no_grad():
        ea, xa = dut.sample_from_indices(idx)
        eb, xb = ref.sample_from_indices(idx)
    res = {"config": f"pixelcnn_sample_S{S}", "rows": 64 * S,
           "parity_rel_l2_embeddings": float((ea - eb).norm() / eb.norm()),
           "parity_rel_l2_image": float((xa - xb).norm() / xb.norm())}

    def lookup(m):
        r = m.cfg.representation_dim
        z = (idx.type(torch.int64) / (m.cfg.num_levels - 1)).permute(0, 2, 3, 1).contiguous().view(-1, r * r, 3)

        def f():
            with torch.no_grad():
                m.index_to_embedding(z)
        return f

    def whole(m):
        def f():
            with torch.no_grad():
                m.sample_from_indices(idx)
        return f

    for layer in (ref.index_to_embedding,):
        layer.faithful = True
    res["us_index_to_embedding_ours"] = timed(lookup(dut))
    res["us_index_to_embedding_eager_upstream_shape"] = timed(lookup(ref))
    res["us_lookup_plus_decoder_ours"] = timed(whole(dut))
    res["us_lookup_plus_decoder_eager_upstream_shape"] = timed(whole(ref))
    res["us_index_to_embedding_ours_graph"] = timed(graphed(lookup(dut)))
    ref_cpu, idx_cpu = copy.deepcopy(ref).cpu(), idx.cpu()

    def whole_cpu():
        with torch.no_grad():
            ref_cpu.sample_from_indices(idx_cpu)
    res["us_lookup_plus_decoder_cpu_upstream_shape"] = timed_cpu(whole_cpu, iters=5 if S > 1 else 10, warm=1 if S > 1 else 3)
    res["cpu_threads"] = torch.get_num_threads()
    res["latents_per_s_ours"] = 64 * S / (res["us_index_to_embedding_ours"] * 1e-6)
    print(json.dumps(res), flush=True)

