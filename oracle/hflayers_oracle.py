"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py) -- PyTorch-eager fp32
restatement of ``hflayers.HopfieldLayer`` as the reference uses it.

Follows (reference side, all citable):
  * construction kwargs            /root/reference/HopVAE.py:149-170
  * call sites / tensor layouts    /root/reference/HopVAE.py:193,215,216,234,255,257,271
  * state-dict contract            /root/reference/utils/__init__.py:36-49 and the
                                   key dump of checkpoints/MNIST-28.ckpt
Follows (third-party, restated from the published ml-jku/hopfield-layers
package, version un-pinned by the reference -- SURVEY.md Appendix A):
  * hflayers/__init__.py  HopfieldLayer / Hopfield   (module tree, _associate)
  * hflayers/activation.py  HopfieldCore             (in_proj slices, out_proj)
  * hflayers/functional.py  hopfield_core_forward    (scale, bmm, softmax,
                                                      retrieve loop, bmm, out_proj)

Two evaluation forms of the SAME arithmetic are provided:
  ``faithful=True``  upstream-shaped: the lookup matrix is expanded over the
                     image batch, LayerNorm/projection run on the expanded copy
                     and the association is a per-image ``torch.bmm`` with the
                     score matrix materialised (this is *the* CPU baseline);
  ``faithful=False`` de-duplicated: patterns prepared once, one flat matmul.
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

__all__ = ["HopfieldLayer", "Hopfield", "HopfieldCore", "hopfield_association"]

_UNSUPPORTED_DEFAULTS = dict(
    num_heads=1,
    normalize_hopfield_space=False,
    normalize_hopfield_space_affine=False,
    pattern_projection_as_connected=False,
    association_activation=None,
    dropout=0.0,
    concat_bias_pattern=False,
    add_zero_association=False,
    batch_first=True,
)


def hopfield_association(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, beta: float,
                         update_steps_max: int = 0, update_steps_eps: float = 1e-4,
                         return_details: bool = False):
    """Batched association on already-normalised / projected operands.

    q: (B, L, H)   k: (B, P, H)   v: (B, P, Dv)   ->  (B, L, Dv)

    Restates the body of ``hopfield_core_forward`` between the input and output
    projections:  A = softmax(beta q k^T); up to
    ``update_steps_max`` retrieve updates q <- beta (A k) while the per-image
    Frobenius norm of the change of A, max-reduced over the batch, exceeds
    ``update_steps_eps``; read-out A v.
    """
    q = q * beta
    xi_old = None
    step = 0
    while True:
        scores = torch.bmm(q, k.transpose(1, 2))            # (B, L, P) materialised, as upstream
        xi = torch.softmax(scores, dim=-1)
        with torch.no_grad():
            active = (step < update_steps_max) or (update_steps_max < 0)
            if xi_old is not None and active:
                change = (xi_old - xi).norm(p=2, dim=(1, 2)).max()
                active = bool(change > update_steps_eps)
            xi_old = xi
        step += 1
        if not active:
            break
        q = torch.bmm(xi, k) * beta                          # retrieve update, no re-normalisation
    out = torch.bmm(xi, v)
    if return_details:
        return out, xi, scores, step - 1
    return out


class HopfieldCore(nn.Module):
    """Parameter holder mirroring upstream ``HopfieldCore`` (names pinned by the ckpt)."""

    def __init__(self, embed_dim: int, head_dim: int, pattern_dim: int, out_dim: int,
                 q_static: bool, k_static: bool, v_static: bool, bias: bool,
                 disable_out_projection: bool):
        super().__init__()
        self.embed_dim, self.head_dim, self.pattern_dim = embed_dim, head_dim, pattern_dim
        self.q_static, self.k_static, self.v_static = q_static, k_static, v_static
        rows = (0 if q_static else head_dim) + (0 if k_static else head_dim) + (0 if v_static else pattern_dim)
        if rows > 0:
            self.in_proj_weight = nn.Parameter(torch.empty(rows, embed_dim))
            self.in_proj_bias = nn.Parameter(torch.empty(rows)) if bias else None
        else:
            self.register_parameter("in_proj_weight", None)
            self.register_parameter("in_proj_bias", None)
        self.out_proj = None if disable_out_projection else nn.Linear(pattern_dim, out_dim, bias=bias)
        self.reset_parameters()

    def reset_parameters(self):
        if self.in_proj_weight is not None:
            nn.init.normal_(self.in_proj_weight, mean=0.0, std=0.02)
        if self.in_proj_bias is not None:
            nn.init.constant_(self.in_proj_bias, 0.0)
        if self.out_proj is not None:
            nn.init.normal_(self.out_proj.weight, mean=0.0, std=0.02)
            if self.out_proj.bias is not None:
                nn.init.constant_(self.out_proj.bias, 0.0)

    def slices(self) -> Tuple[Optional[slice], Optional[slice], Optional[slice]]:
        """Row ranges of the q / k / v projections inside in_proj_weight (order q, k, v)."""
        o = 0
        sq = sk = sv = None
        if not self.q_static:
            sq = slice(o, o + self.head_dim); o += self.head_dim
        if not self.k_static:
            sk = slice(o, o + self.head_dim); o += self.head_dim
        if not self.v_static:
            sv = slice(o, o + self.pattern_dim); o += self.pattern_dim
        return sq, sk, sv

    def project(self, x: torch.Tensor, sl: Optional[slice]) -> torch.Tensor:
        if sl is None:
            return x
        b = None if self.in_proj_bias is None else self.in_proj_bias[sl]
        return F.linear(x, self.in_proj_weight[sl], b)


class Hopfield(nn.Module):
    """Mirrors upstream ``Hopfield``: three LayerNorms + the association core."""

    def __init__(self, input_size: int, hidden_size: Optional[int], output_size: Optional[int],
                 pattern_size: Optional[int], scaling: Optional[float], update_steps_max: int,
                 update_steps_eps: float, norm_flags, static_flags, input_bias: bool,
                 disable_out_projection: bool):
        super().__init__()
        (n_k, n_k_aff, n_q, n_q_aff, n_v, n_v_aff) = norm_flags
        (k_static, q_static, v_static) = static_flags
        head_dim = input_size if hidden_size is None else hidden_size
        pattern_dim = head_dim if pattern_size is None else pattern_size
        if q_static or k_static:
            assert head_dim == input_size, "static patterns live in the input space: hidden_size must equal input_size"
        if v_static:
            assert pattern_dim == input_size
        out_dim = input_size if output_size is None else output_size
        self.association_core = HopfieldCore(input_size, head_dim, pattern_dim, out_dim,
                                             q_static, k_static, v_static, input_bias,
                                             disable_out_projection)
        self.norm_stored_pattern = nn.LayerNorm(input_size, elementwise_affine=n_k_aff) if n_k else None
        self.norm_state_pattern = nn.LayerNorm(input_size, elementwise_affine=n_q_aff) if n_q else None
        self.norm_pattern_projection = nn.LayerNorm(input_size, elementwise_affine=n_v_aff) if n_v else None
        self.scaling = scaling
        self.update_steps_max = update_steps_max
        self.update_steps_eps = update_steps_eps

    @property
    def beta(self) -> float:
        if self.scaling is not None:
            return float(self.scaling)
        return 1.0 / math.sqrt(self.association_core.head_dim)

    def forward(self, stored: torch.Tensor, state: torch.Tensor, projection: torch.Tensor,
                return_details: bool = False):
        core = self.association_core
        if self.norm_stored_pattern is not None:
            stored = self.norm_stored_pattern(stored)
        if self.norm_state_pattern is not None:
            state = self.norm_state_pattern(state)
        if self.norm_pattern_projection is not None:
            projection = self.norm_pattern_projection(projection)
        sq, sk, sv = core.slices()
        q = core.project(state, sq)
        k = core.project(stored, sk)
        v = core.project(projection, sv)
        res = hopfield_association(q, k, v, self.beta, self.update_steps_max, self.update_steps_eps,
                                   return_details=return_details)
        out = res[0] if return_details else res
        if core.out_proj is not None:
            out = core.out_proj(out)
        if return_details:
            return (out,) + tuple(res[1:])
        return out


class HopfieldLayer(nn.Module):
    """Restated ``hflayers.HopfieldLayer`` (constructor surface of SURVEY.md A.1)."""

    def __init__(self, input_size: int, hidden_size: Optional[int] = None, output_size: Optional[int] = None,
                 pattern_size: Optional[int] = None, num_heads: int = 1, scaling: Optional[float] = None,
                 update_steps_max: int = 0, update_steps_eps: float = 1e-4,
                 lookup_weights_as_separated: bool = False, lookup_targets_as_trainable: bool = True,
                 normalize_stored_pattern: bool = True, normalize_stored_pattern_affine: bool = True,
                 normalize_state_pattern: bool = True, normalize_state_pattern_affine: bool = True,
                 normalize_pattern_projection: bool = True, normalize_pattern_projection_affine: bool = True,
                 normalize_hopfield_space: bool = False, normalize_hopfield_space_affine: bool = False,
                 stored_pattern_as_static: bool = False, state_pattern_as_static: bool = False,
                 pattern_projection_as_static: bool = False, pattern_projection_as_connected: bool = False,
                 stored_pattern_size: Optional[int] = None, pattern_projection_size: Optional[int] = None,
                 batch_first: bool = True, association_activation: Optional[str] = None, dropout: float = 0.0,
                 input_bias: bool = True, concat_bias_pattern: bool = False, add_zero_association: bool = False,
                 disable_out_projection: bool = False, quantity: int = 1, trainable: bool = True):
        super().__init__()
        given = dict(num_heads=num_heads, normalize_hopfield_space=normalize_hopfield_space,
                     normalize_hopfield_space_affine=normalize_hopfield_space_affine,
                     pattern_projection_as_connected=pattern_projection_as_connected,
                     association_activation=association_activation, dropout=dropout,
                     concat_bias_pattern=concat_bias_pattern, add_zero_association=add_zero_association,
                     batch_first=batch_first)
        for name, default in _UNSUPPORTED_DEFAULTS.items():
            if given[name] != default:
                raise NotImplementedError(f"HopfieldLayer option {name}={given[name]!r} is not used by Hop-VAE "
                                          f"(HopVAE.py:149-170) and is not restated")
        if stored_pattern_size not in (None, input_size) or pattern_projection_size not in (None, input_size):
            raise NotImplementedError("stored_pattern_size / pattern_projection_size different from input_size")
        self.hopfield = Hopfield(
            input_size, hidden_size, output_size, pattern_size, scaling, update_steps_max, update_steps_eps,
            (normalize_stored_pattern, normalize_stored_pattern_affine, normalize_state_pattern,
             normalize_state_pattern_affine, normalize_pattern_projection, normalize_pattern_projection_affine),
            (stored_pattern_as_static, state_pattern_as_static, pattern_projection_as_static),
            input_bias, disable_out_projection)
        self.lookup_weights = nn.Parameter(torch.empty(1, quantity, input_size), requires_grad=trainable)
        if lookup_weights_as_separated:
            self.lookup_targets = nn.Parameter(torch.empty(1, quantity, input_size),
                                               requires_grad=lookup_targets_as_trainable)
        else:
            self.register_parameter("lookup_targets", None)
        self.faithful = True
        self.reset_parameters()

    def reset_parameters(self):
        nn.init.normal_(self.lookup_weights, mean=0.0, std=0.02)
        if self.lookup_targets is not None:
            nn.init.normal_(self.lookup_targets, mean=0.0, std=0.02)

    def forward(self, input: torch.Tensor, stored_pattern_padding_mask=None, association_mask=None,
                return_details: bool = False):
        if stored_pattern_padding_mask is not None or association_mask is not None:
            raise NotImplementedError("masks are not used by Hop-VAE and are not restated")
        squeeze = input.dim() == 2
        x = input.unsqueeze(0) if squeeze else input
        bsz = x.shape[0] if self.faithful else 1
        if not self.faithful:
            shape = x.shape
            x = x.reshape(1, -1, shape[-1])
        stored = self.lookup_weights.expand(bsz, -1, -1)
        proj = (self.lookup_weights if self.lookup_targets is None else self.lookup_targets).expand(bsz, -1, -1)
        res = self.hopfield(stored, x, proj, return_details=return_details)
        out = res[0] if return_details else res
        if not self.faithful:
            out = out.reshape(*shape[:-1], out.shape[-1])
        if squeeze:
            out = out.squeeze(0)
        if return_details:
            return (out,) + tuple(res[1:])
        return out
