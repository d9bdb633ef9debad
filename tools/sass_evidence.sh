#!/bin/bash
# SASS evidence that the hot kernels are Blackwell-native (B200_PROFILING.md "What proves a
# Blackwell-native kernel"): counts of UTC*MMA / LDTM / UTMALDG / multimem per kernel.
# Runs on a CPU-only box (cuobjdump only needs the .so).
set -e
LIB=huggingface_sagemaker_tensorflow_distributed_b200/lib/libhstd_b200.so
OUT=profiles/sass
mkdir -p $OUT
cuobjdump -sass $LIB > /tmp/hstd_all.sass
python - <<'PY'
import re, collections
txt = open("/tmp/hstd_all.sass").read()
funcs = re.split(r"\n\s*Function : ", txt)[1:]
rows = []
for f in funcs:
    name = f.split("\n", 1)[0].strip()
    c = collections.Counter()
    for mn in ("UTCHMMA", "UTCHMMA.2CTA", "UTCBAR", "LDTM", "UTMALDG", "UTMALDG.4D.2CTA", "UTMASTG", "SYNCS", "HMMA", "MUFU.EX2",
               "FFMA2", "FADD2", "FMUL2", "LDGMC", "STG.E.128.STRONG.SYS", "LDG.E.128.STRONG.SYS", "ACQBULK", "ELECT",
               "REDG.E.ADD.F32", "MUFU.RCP", "MUFU.SQRT"):
        c[mn] = len(re.findall(r"(?<![A-Z0-9.])" + re.escape(mn), f))
    rows.append((name, c))
import subprocess
dem = {}
try:
    names = "\n".join(r[0] for r in rows)
    out = subprocess.run(["c++filt"], input=names, capture_output=True, text=True).stdout.splitlines()
    dem = dict(zip([r[0] for r in rows], out))
except Exception:
    pass
with open("profiles/sass/summary.txt", "w") as w:
    w.write("SASS mnemonic counts per kernel of libhstd_b200.so (sm_100a).  UTCHMMA = tcgen05.mma (.2CTA = cta_group::2),\n"
            "LDTM = tcgen05.ld, UTMALDG / UTMASTG = TMA tensor load / store, UTCBAR = tcgen05.commit, SYNCS = mbarrier ops,\n"
            "FFMA2 / FADD2 / FMUL2 = packed fp32x2 arithmetic (sm_100), ELECT = elect.sync, ACQBULK = griddepcontrol.wait,\n"
            "LDGMC = multimem.ld_reduce (in-switch NVLS reduction), *.STRONG.SYS = peer loads/stores over NVLink,\n"
            "REDG.E.ADD.F32 = red.global.add.v4.f32 (column sums of the fused LayerNorm backward).\n"
            "HMMA (legacy mma.sync) must be 0 everywhere.\n\n")
    for name, c in sorted(rows, key=lambda r: -r[1]["UTCHMMA"]):
        d = dem.get(name, name)
        d = re.sub(r"\(.*", "", d)[:80]
        w.write(f"{d:<82s} " + " ".join(f"{k}={v}" for k, v in c.items() if v) + "\n")
print(open("profiles/sass/summary.txt").read()[:3000])
PY
# full listings: flagship tensor-core kernels + the fused row / head / optimizer kernels
rm -f $OUT/*.sass
python - <<'PY'
import re
txt = open("/tmp/hstd_all.sass").read()
funcs = re.split(r"\n(?=\s*Function : )", txt)
want = {"gemm_256x256_2sm_TN": "gemm_bf16_kernelILi256ELb0ELb0ELi2E", "gemm_wgrad_128_1sm": "gemm_bf16_kernelILi128ELb1ELb1ELi1E",
        "gemm_192_1sm_TN": "gemm_bf16_kernelILi192ELb0ELb0ELi1E",
        "attn_fwd_S512": "attn_fwd_kernelILi4E", "attn_bwd_S512": "attn_bwd_kernelILi4E",
        "fused_allreduce_nvls_adamw_w8_coresident": "fused_allreduce_kernelILi8ELi2ELb1ELi0ELi128E",
        "fused_allreduce_nvls_adamw_w8_tail_rows_both": "fused_allreduce_kernelILi8ELi2ELb1ELi3ELi512E",
        "fused_allreduce_two_shot_adamw_w2_coresident": "fused_allreduce_kernelILi2ELi0ELb1ELi0ELi128E",
        "sparse_rows_mark": "mark_rows_kernelENS_14ArKernelParamsE",
        "ln_fwd_residual_dropout_d768": "ln_fwd_kernelILi3ELb0E", "ln_fwd_embedding_d768": "ln_fwd_kernelILi3ELb1E",
        "ln_bwd_fused_d768": "ln_bwd_fused_kernelILi3ELi16E", "head_logits_xent": "head_logits_loss_kernel",
        "adamw_flat": "adamw_flat_kernel"}
for out, key in want.items():
    for f in funcs:
        if key in f.split("\n", 2)[0] + f.split("\n", 2)[1 if len(f.split("\n")) > 1 else 0]:
            open(f"profiles/sass/{out}.sass", "w").write(f)
            break
PY
ls -la $OUT
