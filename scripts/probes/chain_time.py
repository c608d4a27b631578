"""Device time of the fused bottleneck (CUDA events around eager calls; the call is GPU-bound) at configs 1 / 2 / 3,
per cluster split count (tune splits of the first layer; 0 = the library's choice)."""
import os, sys, json
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import hflayers
import hopvae_harness as H
dev = torch.device("cuda", 0)
only = sys.argv[1] if len(sys.argv) > 1 else None
sweep = [int(v) for v in sys.argv[2].split(",")] if len(sys.argv) > 2 else [0]
for name, cfg in (("mnist28", H.MNIST_28), ("ffhq32", H.FFHQ_32), ("ffhq64", H.FFHQ_64)):
    if only and only != "all" and name != only:
        continue
    torch.manual_seed(1265)
    dut = H.HopVAEHarness(cfg, hflayers.HopfieldLayer).to(dev).eval()
    R = cfg["representation_dim"]
    z = torch.randn(32, R * R, 64, device=dev)
    for sp in sweep:
        dut.hopfield.tune = dict(splits=sp) if sp else {}
        def fused():
            with torch.no_grad():
                hflayers.bottleneck_forward(dut.hopfield, dut.embedding_to_index, dut.index_to_embedding, z, 512)
        for _ in range(5):
            fused()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fused()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g.replay(); torch.cuda.synchronize()
        e0.record()
        for _ in range(50):
            g.replay()
        e1.record(); torch.cuda.synchronize()
        print(json.dumps({"config": name, "rows": 32 * R * R, "splits": sp, "us_fused_graph": e0.elapsed_time(e1) * 1e3 / 50}), flush=True)
