"""The allocation-free entry point ``b200grpo_segment_bounds_static`` (scratch in a ``static __device__`` array inside the
library; no counterpart in the reference -- added for callers that cannot call
cudaMalloc) against the existing ``b200grpo_segment_bounds``, through the C ABI.  Index work: bit-exact.

Collected last on purpose (a failure here must not hide another test)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_T = 1 << 22   # B200GRPO_SEGMENT_BOUNDS_STATIC_MAX_T (include/b200grpo.h)


def _positions(T, p_zero, first, seed):
    g = torch.Generator().manual_seed(seed)
    pos = torch.randint(1, 100, (T,), generator=g)
    pos[torch.rand(T, generator=g) < p_zero] = 0
    pos[0] = first
    return pos.to("cuda:0")


def _static_call(pos, bounds, n):
    """One call of the new entry point on torch's current stream: inputs and outputs are the caller's, nothing else is passed."""
    from pipelinerl_swe_b200 import _abi
    stream = torch.cuda.current_stream(pos.device).cuda_stream
    return _abi.lib().b200grpo_segment_bounds_static(pos.data_ptr(), pos.numel(), bounds.data_ptr(), n.data_ptr(), stream)


def test_header_limit_matches():
    import os
    import re
    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "b200grpo.h")).read()
    m = re.search(r"#define B200GRPO_SEGMENT_BOUNDS_STATIC_MAX_T \(1 << (\d+)\)", hdr)
    assert m and (1 << int(m.group(1))) == MAX_T


def test_static_scratch_entry_point_matches_the_workspace_one_on_consecutive_calls():
    from pipelinerl_swe_b200 import _abi, ops
    # consecutive calls share the one scratch array and it keeps the previous call's contents: a large problem first, then
    # smaller ones (stale counts / mask words beyond their extent), then the maximum size, then the first one again
    cases = [(1 << 20, 0.001, 0), (5000, 0.3, 7), (1, 0.0, 5), (8192, 0.002, 3), (4097, 1.0, 0), (MAX_T, 0.0005, 0), (1 << 20, 0.001, 0)]
    inputs = [_positions(T, p, first, seed=T + k) for k, (T, p, first) in enumerate(cases)]
    want = [ops.segment_bounds(pos).clone() for pos in inputs]            # the existing entry point (caller-provided workspace)
    bounds = torch.empty(MAX_T + 1, dtype=torch.int32, device="cuda:0")   # outputs allocated by the caller, up front
    n = torch.zeros(1, dtype=torch.int32, device="cuda:0")
    assert _static_call(inputs[1], bounds, n) == 0, _abi.lib().b200grpo_last_error()   # set-up call (symbol address, module load)
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    allocated_before = torch.cuda.memory_allocated()
    for pos, w in zip(inputs, want):
        bounds.fill_(-7)
        n.zero_()
        assert _static_call(pos, bounds, n) == 0, _abi.lib().b200grpo_last_error()
        torch.cuda.synchronize()
        S1 = int(n.item())
        assert S1 == w.numel()
        assert torch.equal(bounds[:S1], w)                 # bit-exact
        assert int(bounds[S1].item()) == -7 if S1 < bounds.numel() else True   # nothing written past the S+1 boundaries
    free_after, _ = torch.cuda.mem_get_info()
    assert torch.cuda.memory_allocated() == allocated_before
    assert free_after >= free_before                       # the calls took no device memory


def test_static_scratch_entry_point_rejects_sizes_beyond_its_limit():
    from pipelinerl_swe_b200 import _abi
    pos = _positions(64, 0.1, 0, seed=1)
    bounds = torch.full((65,), -7, dtype=torch.int32, device="cuda:0")
    n = torch.full((1,), -7, dtype=torch.int32, device="cuda:0")
    stream = torch.cuda.current_stream().cuda_stream
    rc = _abi.lib().b200grpo_segment_bounds_static(pos.data_ptr(), MAX_T + 1, bounds.data_ptr(), n.data_ptr(), stream)
    assert rc == -3                                        # B200GRPO_EWORKSPACE
    msg = _abi.lib().b200grpo_last_error().decode()
    assert "exceeds" in msg and str(MAX_T) in msg
    torch.cuda.synchronize()
    assert int(n.item()) == -7 and bool((bounds == -7).all())   # refused before any launch
    with pytest.raises(_abi.B200GrpoError):
        _abi.check(_abi.lib().b200grpo_segment_bounds_static(None, 8, bounds.data_ptr(), n.data_ptr(), stream))
    # and the entry point still works after a refusal
    assert _abi.lib().b200grpo_segment_bounds_static(pos.data_ptr(), 64, bounds.data_ptr(), n.data_ptr(), stream) == 0
    torch.cuda.synchronize()
    assert int(bounds[int(n.item()) - 1].item()) == 64
