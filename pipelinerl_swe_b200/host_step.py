"""Host-buffer GRPO step: the end-to-end C entry point ``b200grpo_rl_step_host`` behind a small class.

One packed micro-batch whose inputs live in PINNED HOST memory (what the trainer's
``run_data_loader`` thread holds before ``to_device``, pipelinerl/finetune_loop.py:92-115)
goes through H2D -> K1 -> finalize -> K3 -> K2 -> D2H(metrics[, grad_hidden]) on the
current stream in one library call.  ``lm_head.weight`` and its gradient stay on the device
(model state).  This is what ``bench.py`` times for its ``e2e`` number.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _abi
from ._abi import HostBatch, M_COUNT, check, lib
from .rl import make_loss_cfg

_F32_FIELDS = ("rewards", "advantages", "ref_logprobs", "old_logprobs", "group_tokens", "num_labels", "overflow")
_I64_FIELDS = ("input_ids", "labels", "position_ids")


class HostRLStep:
    def __init__(self, L: int, H: int, V: int, weight: torch.Tensor, config, current_step: int = 0, max_step: int = 1,
                 grad_weight_dtype: torch.dtype = torch.bfloat16):
        if not weight.is_cuda or weight.dtype != torch.bfloat16 or not weight.is_contiguous():
            raise _abi.B200GrpoError("HostRLStep: weight must be a contiguous CUDA bf16 tensor (no CPU path)")
        assert weight.shape == (V, H)
        self.L, self.H, self.V = L, H, V
        self.weight = weight
        self.temperature = float(config.temperature)
        self.cfg = make_loss_cfg(config, current_step, max_step, True, False)
        dev = weight.device
        self.grad_weight = torch.zeros(V, H, dtype=grad_weight_dtype, device=dev)
        pin = lambda *s, dt: torch.empty(*s, dtype=dt).pin_memory()  # noqa: E731
        self.hidden_host = pin(L, H, dt=torch.bfloat16)
        self.i64_host = {k: pin(L, dt=torch.int64) for k in _I64_FIELDS}
        self.f32_host = {k: pin(L, dt=torch.float32) for k in _F32_FIELDS}
        self.grad_hidden_host = pin(L, H, dt=torch.bfloat16)
        self.metrics_host = pin(M_COUNT, dt=torch.float64)
        n = lib().b200grpo_rl_step_host_workspace(L, H, V)
        self.workspace = torch.empty(n, dtype=torch.uint8, device=dev)
        self.hb = HostBatch(hidden_bf16=self.hidden_host.data_ptr(),
                            **{k: t.data_ptr() for k, t in self.i64_host.items()},
                            **{k: t.data_ptr() for k, t in self.f32_host.items()})
        self.h2d_bytes = L * H * 2 + 3 * L * 8 + 7 * L * 4
        self.d2h_bytes = M_COUNT * 8

    def load_batch(self, batch: dict, hidden: torch.Tensor) -> None:
        """Copy a CPU batch (PipelineBatchEncoding fields, [1,L]) and hidden [L,H] into the pinned buffers."""
        self.hidden_host.copy_(hidden.reshape(self.L, self.H).to(torch.bfloat16))
        for k, t in self.i64_host.items():
            t.copy_(batch[k].reshape(-1))
        for k, t in self.f32_host.items():
            t.copy_(batch[k].reshape(-1))

    def run(self, want_grad_hidden: bool = False, accumulate: bool = False, backward: bool = True) -> torch.Tensor:
        """One fwd+bwd; returns the pinned f64 metrics vector (valid after a stream sync)."""
        gw = self.grad_weight.data_ptr() if backward else None
        gh = self.grad_hidden_host.data_ptr() if (want_grad_hidden and backward) else None
        check(lib().b200grpo_rl_step_host(
            C.byref(self.hb), self.L, self.H, self.V, self.weight.data_ptr(), self.H, self.temperature,
            C.byref(self.cfg), gh, gw, int(self.grad_weight.dtype == torch.float32), int(accumulate),
            self.metrics_host.data_ptr(), self.workspace.data_ptr(), self.workspace.numel(),
            torch.cuda.current_stream().cuda_stream))
        return self.metrics_host
