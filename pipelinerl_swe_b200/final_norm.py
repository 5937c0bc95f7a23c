"""Final RMSNorm and the value head in one sweep over the last hidden state.

Inside ``model(**model_inputs)`` of pipelinerl/finetune/rl/__init__.py:183-195 the last op of the trunk is the
final RMSNorm (transformers ``Qwen2RMSNorm`` / ``LlamaRMSNorm``: upcast, pow, mean, rsqrt, two multiplies and a
downcast -- seven elementwise launches over [tokens, hidden] in eager, and as many again in backward), and the
reference's value-head wrapper (pipelinerl/finetune/value_model.py:40-52,100-106) then runs ``Linear(H, 1)`` over
the same normalised tensor.  ``install_fused_final_norm`` re-classes that norm module IN PLACE (same object, same
``weight`` Parameter, state dict and hooks untouched -- the trick ``install_fused_lm_head`` uses) so that its forward
is one kernel launch that also produces the value head's output; the re-classed ``ValueHead`` picks the values up
when it is handed the very tensor the norm returned, and falls back to its own ``Linear`` otherwise.

Opt-in (``rl_step`` does not install it): norm variants differ between model families -- only modules with a plain
``weight`` and ``variance_epsilon`` (Qwen2 / Llama / Mistral) are accepted; Gemma's ``(1 + weight)`` is refused.
"""
from __future__ import annotations

import threading

import torch
from torch import nn

from . import ops


class FinalNormValueFunction(torch.autograd.Function):
    """(x [T,H] bf16, gamma [H], value_w [H] | None, value_b [1] | None) -> (y [T,H] bf16, value [T] f32 | None)."""

    @staticmethod
    def forward(ctx, x, gamma, value_w, value_b, eps: float):
        vw = value_w.reshape(-1).contiguous() if value_w is not None else None
        vb = value_b.reshape(-1).contiguous() if value_b is not None else None
        y, rstd, value = ops.final_norm_fwd(x, gamma, eps, vw, vb)
        ctx.save_for_backward(x, gamma, rstd, vw if vw is not None else gamma)
        ctx.has_value, ctx.has_bias = vw is not None, vb is not None
        ctx.w_shape = value_w.shape if value_w is not None else None
        ctx.b_shape = value_b.shape if value_b is not None else None
        if value is None:
            ctx.mark_non_differentiable(rstd)
            return y, rstd  # second output unused; keeps the arity fixed
        return y, value

    @staticmethod
    def backward(ctx, dy, dvalue):
        x, gamma, rstd, vw = ctx.saved_tensors
        if dy is None:
            dy = torch.zeros_like(x)
        use_v = ctx.has_value and dvalue is not None
        dx, dgamma, dvw, dvb = ops.final_norm_bwd(x, gamma, rstd, dy.contiguous().to(torch.bfloat16),
                                                  dvalue.float().contiguous() if use_v else None, vw if use_v else None)
        gw = dvw.to(vw.dtype).reshape(ctx.w_shape) if use_v else None
        gb = dvb.to(vw.dtype).reshape(ctx.b_shape) if (use_v and ctx.has_bias) else None
        return dx, dgamma.to(gamma.dtype), gw, gb, None


_tls = threading.local()


def final_norm_value(x: torch.Tensor, gamma: torch.Tensor, eps: float, value_linear: nn.Linear | None = None):
    """x [..., H] -> (y [..., H] in x's dtype, value [...] f32 | None).  bf16 compute: other dtypes are cast."""
    shape = x.shape
    x2 = x.reshape(-1, shape[-1])
    xb = x2 if x2.dtype == torch.bfloat16 else x2.to(torch.bfloat16)
    g = gamma if gamma.dtype == torch.bfloat16 else gamma.to(torch.bfloat16)
    vw = vb = None
    if value_linear is not None:
        vw = value_linear.weight if value_linear.weight.dtype == torch.bfloat16 else value_linear.weight.to(torch.bfloat16)
        if value_linear.bias is not None:
            vb = value_linear.bias if value_linear.bias.dtype == torch.bfloat16 else value_linear.bias.to(torch.bfloat16)
    y, second = FinalNormValueFunction.apply(xb.contiguous(), g.contiguous(), vw, vb, float(eps))
    y = y.reshape(shape)
    if y.dtype != x.dtype:
        y = y.to(x.dtype)
    return y, (second.reshape(shape[:-1]) if value_linear is not None else None)


class FusedFinalNorm(nn.Module):
    """Mixin given to the model's final RMSNorm by ``install_fused_final_norm``."""

    def forward(self, hidden_states: torch.Tensor) -> torch.Tensor:
        if not hidden_states.is_cuda or hidden_states.dtype != torch.bfloat16 or self.weight.dtype != torch.bfloat16:
            # CPU tensors (a meta / init-time call) and non-bf16 models keep the module's own forward: the kernel computes
            # what the transformers module computes on bf16 and would round an fp32 model's hidden state
            _tls.last = None
            return super().forward(hidden_states)
        linear = self.__dict__.get("_b200_value_linear")
        y, value = final_norm_value(hidden_states, self.weight, self.variance_epsilon, linear)
        _tls.last = (y, value) if value is not None else None
        return y


class FusedValueHead(nn.Module):
    """Mixin given to the reference's ``ValueHead`` (value_model.py:40-52): returns the values the fused final norm
    already produced when called on that norm's output, else runs the original ``Linear``."""

    def forward(self, hidden_states: torch.Tensor) -> torch.Tensor:
        last = getattr(_tls, "last", None)
        if last is not None:
            y, value = last
            if y.data_ptr() == hidden_states.data_ptr() and y.shape == hidden_states.shape:
                _tls.last = None
                return value.to(hidden_states.dtype)  # the reference's Linear returns the model dtype
        return super().forward(hidden_states)


def _find_final_norm(model: nn.Module) -> nn.Module:
    owner = model
    for _ in range(8):
        norm = getattr(owner, "norm", None)
        if isinstance(norm, nn.Module) and hasattr(owner, "layers"):
            return norm
        nxt = None
        for attr in ("module", "pretrained_model", "base_model", "model", "language_model"):
            cand = getattr(owner, attr, None)
            if isinstance(cand, nn.Module) and cand is not owner:
                nxt = cand
                break
        if nxt is None:
            break
        owner = nxt
    raise AttributeError("install_fused_final_norm: no `<decoder>.norm` next to `<decoder>.layers` found in this model")


def _find_value_head(model: nn.Module):
    owner = model
    for _ in range(4):
        if hasattr(owner, "value_head"):
            return owner.value_head
        owner = getattr(owner, "module", None)
        if not isinstance(owner, nn.Module):
            break
    return None


def install_fused_final_norm(model: nn.Module) -> nn.Module:
    """Idempotently fuse the model's final RMSNorm (and, if the model is the reference's value-head wrapper, the value
    head) into one kernel.  Returns the norm module."""
    norm = _find_final_norm(model)
    if not isinstance(norm, FusedFinalNorm):
        if not (isinstance(getattr(norm, "weight", None), torch.Tensor) and norm.weight.dim() == 1 and hasattr(norm, "variance_epsilon")):
            raise ValueError("final norm must be an RMSNorm with a 1-D `weight` and `variance_epsilon` (Qwen2 / Llama / Mistral)")
        if "gemma" in type(norm).__name__.lower():
            raise ValueError("Gemma's RMSNorm scales by (1 + weight): not the op this kernel computes")
        base = type(norm)
        norm.__class__ = type("Fused" + base.__name__, (FusedFinalNorm, base), {"__module__": __name__})
    head = _find_value_head(model)
    if head is not None:
        linear = getattr(head, "output", None)
        if isinstance(linear, nn.Linear) and linear.out_features == 1 and linear.in_features == norm.weight.numel():
            object.__setattr__(norm, "_b200_value_linear", linear)  # a plain attribute: NOT a submodule, the state dict is unchanged
            if not isinstance(head, FusedValueHead):
                hb = type(head)
                head.__class__ = type("Fused" + hb.__name__, (FusedValueHead, hb), {"__module__": __name__})
    return norm
