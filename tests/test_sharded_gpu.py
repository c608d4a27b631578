"""GPU (-m gpu), needs >= 2 B200s: pattern-sharded retrieval over NCCL vs the unsharded oracle."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
sys.path.insert(0, os.environ["HOP_ROOT"]); sys.path.insert(0, os.path.join(os.environ["HOP_ROOT"], "tests"))
import torch, torch.distributed as dist
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
import hflayers
from oracle import hflayers_oracle as O
from util import randomize_, rel_l2
impl = hflayers._impl
for (D, P, N, T, path, bar) in [(64, 512, 300, 0, "fp32", 1e-5), (64, 515, 77, 2, "fp32", 1e-4),
                                (128, 4096, 1024, 0, "f16", 1e-3), (256, 2048, 512, 1, "f16", 1e-3),
                                (512, 4096, 256, 0, "f16", 1e-3)]:     # the 64-rows-per-CTA pair kernel per shard
    kw = dict(input_size=D, quantity=P, update_steps_max=T, update_steps_eps=0.0,
              stored_pattern_as_static=True, state_pattern_as_static=True)
    ref = randomize_(O.HopfieldLayer(**kw), 5, 1.0)
    g = torch.Generator().manual_seed(6)
    with torch.no_grad():
        ref.lookup_weights.copy_(torch.randn(1, P, D, generator=g))
    full = impl.HopfieldLayer(**kw); full.load_state_dict(ref.state_dict()); full = full.cuda(); full.kernel_path = path
    idx = torch.randint(0, P, (N,), generator=g)
    x = ref.lookup_weights[0, idx].detach() + 0.5 * torch.randn(N, D, generator=g)
    with torch.no_grad():
        want, _, scores, _ = ref(x[None], return_details=True)
    s = scores.reshape(N, P); top2 = s.topk(2, -1).values
    decided = (top2[:, 0] - top2[:, 1]) > (2e-2 if path == "f16" else 1e-5)
    outs = {}
    # "peer": one fused kernel per rank over NVLink peer memory; "collective": all_gather + merge + all_reduce
    for exchange in ("peer", "collective"):
        sh = impl.PatternShardedHopfield(full, exchange=exchange)
        assert sh.use_peer == (exchange == "peer")
        with torch.no_grad():
            for rep in range(3):                       # repeated calls: flag epochs and buffer parity
                got = sh(x.cuda()).clone()
        r = rel_l2(got, want[0])
        if T == 0:
            assert torch.equal(sh.last_argmax.cpu().long()[decided], s.argmax(-1)[decided])
        if T == 0:
            assert rel_l2(sh.last_lse, torch.logsumexp(s, -1)) < (1e-3 if path == "f16" else 1e-5)
        print(f"rank {rank}: D={D} P={P} T={T} {path} {exchange}: rel-L2 {r:.3e}", flush=True)
        assert r < bar, r
        outs[exchange] = got
    assert rel_l2(outs["peer"], outs["collective"]) < 1e-5
    if T == 0 and N % world == 0:
        sh = impl.PatternShardedHopfield(full, exchange="peer", scatter_rows=True)
        with torch.no_grad():
            mine = sh(x.cuda())
        rows = N // world
        assert rel_l2(mine, want[0][rank * rows:(rank + 1) * rows]) < bar
dist.destroy_process_group()
'''


@pytest.mark.parametrize("world", [2, 4, 8])
def test_pattern_sharded_nccl_two_gpus(tmp_path, world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    env = dict(os.environ, HOP_ROOT=ROOT)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)],
                       env=env, capture_output=True, text=True, timeout=600)
    print(r.stdout[-3000:], r.stderr[-3000:])
    assert r.returncode == 0


DDP_WORKER = r'''
import os, sys
sys.path.insert(0, os.environ["HOP_ROOT"]); sys.path.insert(0, os.path.join(os.environ["HOP_ROOT"], "tests"))
import torch, torch.distributed as dist, torch.nn.functional as F
from torch.nn.parallel import DistributedDataParallel as DDP
torch.backends.cuda.matmul.allow_tf32 = False; torch.backends.cudnn.allow_tf32 = False
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
import hflayers, hopvae_harness as H
from oracle import hflayers_oracle as O
from util import rel_l2
B = 4
torch.manual_seed(1265)
ref = H.HopVAEHarness(H.FFHQ_64, O.HopfieldLayer)                       # single-process oracle (CPU, fp32)
with torch.no_grad():
    for m in (ref.hopfield, ref.embedding_to_index, ref.index_to_embedding):
        m.lookup_weights.normal_(0, 0.5)
dut = H.HopVAEHarness(H.FFHQ_64, hflayers.HopfieldLayer)
dut.load_state_dict(ref.state_dict())
ddp = DDP(dut.cuda(), device_ids=[rank], find_unused_parameters=True)   # post_vq_conv is never called (as upstream)
xs = [torch.rand(B, 3, 64, 64, generator=torch.Generator().manual_seed(1265 + r)) - 0.5 for r in range(world)]
def loss_of(model, x):
    recon, aux = model(x)
    return F.mse_loss(recon, x) + aux                                   # main.py:32-33
loss = loss_of(ddp, xs[rank].cuda())
loss.backward()                                                         # DDP all-reduces (averages) the grads
lt = loss.detach().clone(); dist.all_reduce(lt); lt /= world
want = loss_of(ref, torch.cat(xs)); want.backward()
rl = abs(float(lt) - float(want)) / abs(float(want))
msgs = [f"rank {rank}: FFHQ-64 batch-parallel fwd+bwd, world {world}: loss {float(lt):.6f} vs oracle {float(want):.6f} (rel {rl:.2e})"]
assert rl < 1e-4, rl
for name in ("hopfield", "embedding_to_index", "index_to_embedding"):
    got = getattr(ddp.module, name).lookup_weights.grad
    exp = getattr(ref, name).lookup_weights.grad
    r = rel_l2(got, exp)
    msgs.append(f"  d {name}.lookup_weights rel-L2 {r:.3e}")
    assert r < 1e-3, (name, r)
print(chr(10).join(msgs), flush=True)
dist.destroy_process_group()
'''


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_batch_parallel_ddp_training_step_two_gpus(tmp_path):
    """BASELINE config 3 (FFHQ-64 fwd+bwd, batch-parallel): DistributedDataParallel around the unchanged model
    with the sm_100a Hopfield layers; loss and every `lookup_weights.grad` against the single-process oracle on
    the concatenated batch."""
    script = tmp_path / "ddp_worker.py"
    script.write_text(DDP_WORKER)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    env = dict(os.environ, HOP_ROOT=ROOT)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)],
                       env=env, capture_output=True, text=True, timeout=600)
    print(r.stdout[-3000:], r.stderr[-3000:])
    assert r.returncode == 0
