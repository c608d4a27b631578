"""GPU (-m gpu): the tcgen05 backward (hop_bwd_tc.cu) against the oracle's autograd.

What `loss.backward()` at /root/reference/main.py:35 computes through one Hopfield layer, T = 0: gradients of the
input and of EVERY parameter.  The association core (dq, dK^, dV^ -- five dense contractions) runs on the tensor pipe
with fp16-encoded operands and fp32 accumulation; LayerNorm / projection backwards stay fp32.  Bar, written here: every
gradient within rel-L2 <= 5e-3 of the fp32 oracle (16-bit operand rounding, 2^-11 relative per operand, enters the five
contractions once each; measured values are printed), the forward of the same step within the 1e-3 bar.

Queries are planted near stored patterns with a noise level that puts the retrieval at its working point -- nearest
pattern weight w of order 0.3 .. 0.9, score of the planted pattern ~ ln P + 1 -- where the gradient is not degenerate.
On a saturated softmax (w -> 1) the true gradient vanishes like 1 - w while every 16-bit operand keeps its 2^-11 rounding:
there `dS = A (dA - delta)` is a difference of nearly equal numbers, which the kernel's consistency fix
(bwd_fix_dq_kernel) holds to ~6e-4 for d_x down to 1 - w ~ 1e-3 (case `d256_saturated`), but beyond that even the fp32
log-sum-exp the weights are normalised with (ulp 2e-6 at lse ~ 20) is coarser than 1 - w, and the LayerNorm backward of
the pattern side cancels the component of dK^ along K^: gradients that are ~1e-5 of the usual scale are then only
accurate to a few per cent of themselves (bar 5e-2 there, stated in the case).
"""
import math

import pytest
import torch

from oracle import hflayers_oracle as O
from util import REL_L2_BAR, rel_l2, randomize_

pytestmark = pytest.mark.gpu

GRAD_BAR = 5e-3
STATIC = dict(stored_pattern_as_static=True, state_pattern_as_static=True)


@pytest.fixture(scope="module")
def hf():
    import hflayers
    hflayers._impl.cabi.load()
    return hflayers


def _pair(hf, D, P, seed, **extra):
    kw = dict(input_size=D, quantity=P, **STATIC, **extra)
    ref = randomize_(O.HopfieldLayer(**kw), seed, 1.0)
    g = torch.Generator().manual_seed(seed + 1)
    with torch.no_grad():
        ref.lookup_weights.copy_(torch.randn(1, P, D, generator=g))
    ref.faithful = False                      # de-duplicated evaluation of the same arithmetic (one flat matmul)
    dut = hf.HopfieldLayer(**kw)
    dut.load_state_dict(ref.state_dict())
    return ref, dut.cuda(), g


CASES = {
    # name: (D, P, N, noise, bar)   -- N x P >= 2^25 selects the tensor-core training path; noise None = working point
    "d256_p4k": (256, 4096, 8192, None, GRAD_BAR),
    "d128_p8k": (128, 8192, 4096, None, GRAD_BAR),
    "d512_p4k": (512, 4096, 8192, None, GRAD_BAR),
    "d256_ragged": (256, 5000, 7001, None, GRAD_BAR),      # ragged rows and patterns: padded tiles on both axes
    "d256_p64k": (256, 65536, 2048, None, GRAD_BAR),       # BASELINE config 5's pattern count
    "d512_p64k": (512, 65536, 1024, None, GRAD_BAR),
    "d256_recipe": (256, 65536, 2048, 0.5, GRAD_BAR),      # config 5's own recipe (w ~ 0.94)
    "d256_saturated": (256, 4096, 8192, 0.5, 5e-2),        # w ~ 0.999: vanishing gradients, see the module docstring
}


def working_point_noise(D, P):
    """Noise std for which the planted pattern scores ~ ln P + 1 (beta D cos = sqrt(D / (1 + s^2)))."""
    return math.sqrt(max(0.05, D / (math.log(P) + 1.0) ** 2 - 1.0))


@pytest.mark.parametrize("name", list(CASES))
def test_backward_parity_tensor_core(hf, name):
    D, P, N, noise, bar = CASES[name]
    noise = working_point_noise(D, P) if noise is None else noise
    impl = hf._impl
    ref, dut, g = _pair(hf, D, P, seed=91)
    dut.kernel_path = "f16"
    assert impl.ops.resolve_train_path(dut.spec(), N, P) == impl.cabi.PATH_TC_F16
    idx = torch.randint(0, P, (N,), generator=g)
    x = ref.lookup_weights[0, idx].detach() + noise * torch.randn(N, D, generator=g)
    w = torch.randn(N, D, generator=g)
    xr = x.clone().requires_grad_(True)
    xd = x.cuda().requires_grad_(True)
    yr = ref(xr)
    (yr * w).sum().backward()
    l0 = impl.cabi.launch_count()
    yd = dut(xd)
    (yd * w.cuda()).sum().backward()
    torch.cuda.synchronize()
    print(f"{name}: forward rel-L2 {rel_l2(yd.detach(), yr.detach()):.3e}, launches fwd+bwd {impl.cabi.launch_count() - l0}")
    assert rel_l2(yd.detach(), yr.detach()) < REL_L2_BAR
    worst = rel_l2(xd.grad, xr.grad)
    with torch.no_grad():
        _, a_ref, _, _ = ref(x, return_details=True)
        wtop = float(a_ref.reshape(N, P).max(-1).values.mean())
    print(f"{name}: noise {noise:.2f}, mean nearest-pattern weight {wtop:.4f}, d_x rel-L2 {worst:.3e}")
    assert worst < bar
    rp, dp = dict(ref.named_parameters()), dict(dut.named_parameters())
    bad = []
    for k in rp:
        assert dp[k].grad is not None, k
        if k.endswith("norm_stored_pattern.bias"):
            # exactly 0 in exact arithmetic (softmax is shift invariant): both sides hold rounding noise only
            scale = float(rp[k.replace(".bias", ".weight")].grad.norm())
            assert float(dp[k].grad.norm()) < 2e-2 * scale, k
            continue
        r = rel_l2(dp[k].grad, rp[k].grad)
        print(f"{name}: {k} rel-L2 {r:.3e}")
        if not r < bar:
            bad.append((k, r))
    assert not bad, bad


@pytest.mark.parametrize("D,P,N", [(128, 8192, 4096), (256, 5000, 7001)])
def test_backward_tensor_core_matches_fp32_kernels(hf, D, P, N):
    """Same layer, same inputs: the tensor-core step against the library's own fp32 step (tune bwd_simt keeps the whole
    step on the exact kernels) -- the two cores must agree to the operand rounding, and tiny gradients (1e-6 of the
    usual scale) must survive the fp16 operand range through the global rescale.  In between: the tensor-core core with
    the projection / weight-gradient products on the fp32 kernels (tune bwd_tails_simt), which isolates the tcgen05 GEMM
    tails (hi + lo operand pairs, range-scaled gradients, split contraction): they must match the fp32 tails to the
    level at which the core re-quantises their output to fp16 (a dR that differs in the 7th digit rounds a few operand
    words the other way: 4e-5 measured; bar 2e-4)."""
    ref, dut, g = _pair(hf, D, P, seed=95)
    dut.kernel_path = "f16"
    x = torch.randn(N, D, generator=g).cuda()
    w = (1e-6 * torch.randn(N, D, generator=g)).cuda()
    grads = {}
    modes = ("tc", "tc_tails_fp32") + (("fp32",) if D <= 128 else ())       # (the fp32 core stops at width 128)
    for mode in modes:
        dut.tune = dict(bwd_simt=(mode == "fp32"), bwd_tails_simt=(mode == "tc_tails_fp32"))
        dut.zero_grad(set_to_none=True)
        xd = x.clone().requires_grad_(True)
        (dut(xd) * w).sum().backward()
        torch.cuda.synchronize()
        grads[mode] = dict(d_x=xd.grad.clone(), **{k: v.grad.clone() for k, v in dut.named_parameters() if v.grad is not None})
    for what in ("d_x", "lookup_weights") if "fp32" in grads else ():
        r = rel_l2(grads["tc"][what], grads["fp32"][what])
        print(f"tensor-core vs fp32 core, {what}: rel-L2 {r:.3e}")
        assert r < GRAD_BAR
    bad = []
    for what, a in grads["tc"].items():
        b = grads["tc_tails_fp32"][what]
        if what.endswith("norm_stored_pattern.bias"):
            continue                                      # rounding noise around an exact zero on both sides
        r = rel_l2(a, b)
        print(f"GEMM tails vs fp32 tails, {what}: rel-L2 {r:.3e}")
        bad += [(what, r)] if not r < 2e-4 else []
    assert not bad, bad
