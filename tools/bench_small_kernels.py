#!/usr/bin/env python
"""Achieved HBM bandwidth of the byte-bound kernels around the GEMMs (north star: "achieved HBM GB/s for the
reductions ... against B200 peak").

    python tools/bench_small_kernels.py [--iters 20] [--once]

One JSON line per (kernel, size): algorithmic bytes (the per-unit figures of DESIGN.md 4 x units), CUDA-event
time per launch (after 3 warm-ups; a 512 MB buffer is rewritten between launches so nothing is served from
L2), GB/s and the fraction of the measured HBM peak (MEASURED_PEAKS.json, else 6540 GB/s).  ``--once`` runs
each kernel once without timing: the form used under ``ncu --set full`` for the DRAM-traffic column.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def hbm_peak():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        for k in ("hbm_gbps", "hbm_GBps", "hbm_gb_s", "hbm_copy_gbps"):
            if k in p:
                return float(p[k]), "measured"
        for k, v in p.items():
            if "hbm" in k.lower() and isinstance(v, (int, float)):
                return float(v), "measured"
    except Exception:  # noqa: BLE001
        pass
    return 6540.0, "fallback (earlier MEASURED_PEAKS.json of this pool)"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--once", action="store_true")
    ap.add_argument("--only", default="", help="comma list of: k3,segments,k4,ingest,final_norm,k6,dlogits (default: all)")
    args = ap.parse_args()
    from helpers import StubTokenizer, synthetic_packed_batch, synthetic_rollout_examples
    from pipelinerl_swe_b200 import RLConfig, ops, rl
    from pipelinerl_swe_b200.ingest import pack_rollouts
    dev = torch.device("cuda:0")
    peak, peak_kind = hbm_peak()
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)

    def timed(fn):
        if args.once:
            fn(); torch.cuda.synchronize()
            return None
        for _ in range(3):
            fn()
        ms = 0.0
        for _ in range(args.iters):
            flush.fill_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); torch.cuda.synchronize()
            ms += e0.elapsed_time(e1)
        return ms / args.iters

    def emit(kernel, size, nbytes, ms, note=""):
        line = {"kernel": kernel, "size": size, "algorithmic_bytes": int(nbytes)}
        if ms is not None:
            line.update(ms=round(ms, 5), GBps=round(nbytes / ms / 1e6, 1), frac_of_hbm_peak=round(nbytes / ms / 1e6 / peak, 4),
                        peak_GBps=peak, peak_kind=peak_kind)
        if note:
            line["note"] = note
        print(json.dumps(line), flush=True)

    g = torch.Generator(device=dev).manual_seed(0)
    only = set(filter(None, args.only.split(",")))
    want = lambda name: not only or name in only  # noqa: E731
    # ---- K3 token loss: 60 B/token in (11 fields incl. logprob/entropy) + 8 B/token coefficients out -----------
    for T, n_seq in ((8192, 16), (1 << 20, 2048), (1 << 23, 16384)) if want("k3") else ():
        lens = [T // n_seq] * n_seq
        b = synthetic_packed_batch(lens, 152064, seed=1)
        bd = {k: (v.to(dev).contiguous() if isinstance(v, torch.Tensor) else v) for k, v in b.items()}
        lp = -torch.rand(1, T - 1, device=dev, generator=g)
        ent = torch.rand(1, T - 1, device=dev, generator=g)
        cfg = rl.make_loss_cfg(RLConfig(batch_size=n_seq, kl_coef=0.1, final_kl_coef=0.1), 0, 1, True, False)
        ms = timed(lambda: ops.token_loss(lp, ent, bd, cfg))
        emit("token_loss (K3)", f"T={T} segments={n_seq}", 60 * T, ms, "8 B labels + 8 B position_ids + 7x4 B columns + 2x4 B logprob/entropy in, 2x4 B coefficients out")
    # ---- segment bounds: 8 B/token ---------------------------------------------------------------------------
    for T in (1 << 20, 1 << 24) if want("segments") else ():
        pos = (torch.arange(T, device=dev) % 977).contiguous()
        ms = timed(lambda: ops.segment_bounds(pos))
        emit("segment_bounds", f"T={T}", 8 * T, ms, "position_ids read once (flags kept as per-block counts + a 1-bit mask)")
    # ---- K4 group advantages: 16 B/token out ------------------------------------------------------------------
    for n_rows, n_tok in ((4096, 2048), (65536, 512)) if want("k4") else ():
        reward = torch.rand(n_rows, device=dev, dtype=torch.float64, generator=g)
        gkey = (torch.arange(n_rows, device=dev, dtype=torch.int32) // 16).contiguous()
        offs = (torch.arange(n_rows + 1, device=dev, dtype=torch.int64) * n_tok).contiguous()
        nlab = torch.full((n_rows,), n_tok - 8, device=dev, dtype=torch.int32)
        eos = torch.ones(n_rows, device=dev, dtype=torch.uint8)
        order, starts = ops.group_order(gkey, n_rows // 16)
        ms = timed(lambda: ops.group_advantage(reward, gkey, offs, nlab, eos, n_rows // 16, True, None, order, starts))
        emit("group_advantage (K4: stats + per-token expand)", f"rollouts={n_rows} x {n_tok} tokens", 16 * n_rows * n_tok + 25 * n_rows, ms)
        ms = timed(lambda: ops.group_advantage_rows(reward, gkey, offs, n_rows // 16, True, order, starts))
        emit("group_advantage_rows (K4: stats + per-rollout f64)", f"rollouts={n_rows}", 41 * n_rows, ms,
             "launch-bound: 25 B/rollout in (reward, key, order, two offsets), 16 B/rollout out")
    # ---- packed ingest expansion: 16 B/token in, 60 B/token out -------------------------------------------------
    for n_seq, n_tok in ((16, 512), (2048, 512)) if want("ingest") else ():
        ex = synthetic_rollout_examples([n_tok] * n_seq, 152064, seed=2)
        wire = pack_rollouts(ex, StubTokenizer(), 1)
        T = wire.T
        ms = timed(lambda: wire.to_device(dev))
        emit("ingest: H2D + expand_packed", f"T={T} rollouts={n_seq}", 76 * T, ms,
             f"includes the {wire.nbytes} B pinned H2D copy and the output allocation; device traffic 16 B in + 60 B out per token")
    # ---- final RMSNorm + value head: 4 B/element forward, 6 B/element backward -------------------
    for T, H in ((8192, 3584), (65536, 3584), (32768, 5120)) if want("final_norm") else ():
        x = torch.randn(T, H, device=dev, generator=g).bfloat16()
        gamma = torch.ones(H, device=dev).bfloat16()
        vw = (0.01 * torch.randn(H, device=dev, generator=g)).bfloat16()
        vb = torch.zeros(1, device=dev).bfloat16()
        dy = torch.randn(T, H, device=dev, generator=g).bfloat16()
        dv = torch.randn(T, device=dev, generator=g)
        y, rstd, _ = ops.final_norm_fwd(x, gamma, 1e-6, vw, vb)
        ms = timed(lambda: ops.final_norm_fwd(x, gamma, 1e-6, vw, vb))
        emit("final_norm_fwd (+ value head)", f"T={T} H={H}", 4 * T * H, ms, "bf16 in + bf16 out; includes the output allocation")
        ms = timed(lambda: ops.final_norm_bwd(x, gamma, rstd, dy, dv, vw))
        emit("final_norm_bwd (+ value head)", f"T={T} H={H}", 6 * T * H, ms, "x, dy in; dx out; column partials + reduce kernel included")
        del x, dy, y
    # ---- K6 bucket pack / unpack: 4 B/element (bf16 in + out) --------------------------------------------------
    if want("k6"):
        shapes = [(3584, 3584), (3584,), (512, 3584), (512,), (18944, 3584), (18944, 3584), (3584, 18944), (3584,), (3584,)] * 4
        params = [torch.randn(*s, device=dev).bfloat16() for s in shapes]
        n = ops.bucket_numel(params)
        bucket = torch.empty(n, dtype=torch.bfloat16, device=dev)
        ms = timed(lambda: ops.bucket_pack(params, bucket))
        emit("bucket_pack (K6)", f"{len(params)} tensors, {2 * n / 1e9:.2f} GB", 4 * n, ms)
        ms = timed(lambda: ops.bucket_unpack(bucket, params))
        emit("bucket_unpack (K6)", f"{len(params)} tensors, {2 * n / 1e9:.2f} GB", 4 * n, ms)
        masters = [p.float() for p in params[:18]]
        n2 = ops.bucket_numel(masters)
        b2 = torch.empty(n2, dtype=torch.bfloat16, device=dev)
        ms = timed(lambda: ops.bucket_pack(masters, b2))
        emit("bucket_pack (K6, fp32 masters)", f"{len(masters)} tensors, {2 * n2 / 1e9:.2f} GB out", 6 * n2, ms)
    # ---- dlogits pass + merge, at the C3 shape -----------------------------------------------------------------
    if want("dlogits"):
        M, H, V = 8191, 3584, 152064
        h = torch.randn(M, H, device=dev, generator=g).bfloat16()
        w = (torch.randn(V, H, device=dev, generator=g) * 0.02).bfloat16()
        tgt = torch.randint(0, V, (M,), device=dev, generator=g)
        saved = ops.lmhead_saved_buffer(M, H, V, dev)
        stats = ops.lmhead_fwd_stats(h, w, tgt, 1.0, saved=saved)
        lp, ent, lse = ops.lmhead_finalize(stats)
        c = torch.randn(2, M, device=dev, generator=g) * 0.1
        from pipelinerl_swe_b200 import _abi
        _abi.profile_enable(4096)
        for _ in range(1 if args.once else 4):
            ops.lmhead_bwd(h, w, tgt, lse, ent, c[0].contiguous(), c[1].contiguous(), 1.0, need_grad_hidden=False, need_grad_weight=False,
                           saved=saved)
        torch.cuda.synchronize()
        if not args.once:
            t = [ms for tag, ms in _abi.profile_read() if tag == "dlogits_pass"]
            emit("dlogits_from_saved (K2a)", f"M={M} V={V}", 4 * M * V, sum(t[1:]) / len(t[1:]), "bf16 read + bf16 write in place; exp + 10 ALU per element")


if __name__ == "__main__":
    main()
