#!/usr/bin/env python
"""SASS evidence for the tcgen05 / TMEM / TMA pipeline: `cuobjdump -sass` of the in-tree library, reduced to the
instructions that prove what the GEMM kernels are made of.

    python tools/sass_summary.py [out.txt]          (default profiles/r02/sass_gemm.txt; needs only cuobjdump, no GPU)

Per kernel: mnemonic counts of the tensor-core / TMA / TMEM / barrier / reduction instructions (B200_PROFILING.md names
them: UTCHMMA = tcgen05.mma, UTMALDG = cp.async.bulk.tensor load, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit, REDG = red.global,
LDGMC / multimem = NVSwitch multicast) and, for the three production GEMM instantiations, every such instruction line.
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "pipelinerl_swe_b200", "libb200grpo.so")
KEY = ("UTCHMMA", "UTCQMMA", "UTMALDG", "UTMASTG", "UTMAPF", "LDTM", "STTM", "UTCBAR", "UTCATOMSWS", "SYNCS", "REDG", "LDGMC", "STGMC",
       "REDGMC", "MUFU.EX2", "HMMA", "UCGABAR", "ACQBULK", "LDGSTS")
# template args: <epi, a_major, b_major, ctas, block_n, stages, save, a_panel, m_rep[, ...]>
PRODUCTION = {
    "forward + saved logits (EPI_FWD_STATS, 2-CTA)": re.compile(r"gemm_sm100_kernelILi1ELi0ELi0ELi2ELi256ELi\d+ELb1ELb0E"),
    "dH = dlogits . W (EPI_STORE, A panels K-major, B MN-major, 256-row CTA tiles)": re.compile(r"gemm_sm100_kernelILi0ELi0ELi1ELi2ELi256ELi\d+ELb0ELb1ELi2E"),
    "dW = dlogits^T . hidden (EPI_STORE, both MN-major, f32 red.add epilogue)": re.compile(r"gemm_sm100_kernelILi0ELi1ELi1ELi2ELi256ELi\d+ELb0ELb1ELi1E"),
}


def main():
    out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "r02", "sass_gemm.txt")
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
    usage = {}
    cur = None
    for ln in res.splitlines():
        m = re.search(r"Function (\S+):", ln)
        if m:
            cur = m.group(1)
        elif cur and "REG:" in ln:
            usage[cur] = ln.strip()
            cur = None
    funcs, name = collections.OrderedDict(), None
    for ln in sass.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            name = m.group(1)
            funcs[name] = []
        elif name and re.search(r"/\*[0-9a-f]{4}\*/", ln):
            funcs[name].append(ln.strip())
    lines = [f"# {os.path.relpath(LIB, ROOT)}: cuobjdump -sass, sm_100a, {len(funcs)} kernels, {sum(map(len, funcs.values()))} instructions", ""]
    lines.append("## mnemonic counts per kernel (kernels with tensor-core, TMA, TMEM or multimem instructions)")
    for fn, body in funcs.items():
        cnt = collections.Counter()
        for ln in body:
            for k in KEY:
                if re.search(r"\b" + re.escape(k), ln):
                    cnt[k] += 1
        if any(cnt[k] for k in ("UTCHMMA", "UTMALDG", "LDTM", "LDGMC", "STGMC", "REDGMC", "UTMASTG")):
            demangled = subprocess.run(["c++filt", fn], capture_output=True, text=True).stdout.strip()[:160]
            lines.append(f"{demangled}\n    {len(body)} instr; " + ", ".join(f"{k} {v}" for k, v in sorted(cnt.items())) + f"\n    {usage.get(fn, '')}")
    for title, pat in PRODUCTION.items():
        for fn, body in funcs.items():
            if pat.search(fn):
                lines += ["", f"## {title}", f"## {fn}", f"## {usage.get(fn, '')}"]
                lines += [ln for ln in body if any(re.search(r"\b" + re.escape(k), ln) for k in KEY if k != "MUFU.EX2")]
                break
    os.makedirs(os.path.dirname(out), exist_ok=True)
    open(out, "w").write("\n".join(lines) + "\n")
    print(f"wrote {out}: {len(lines)} lines")


if __name__ == "__main__":
    main()
