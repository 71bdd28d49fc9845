"""Ablation builds of the tcgen05 prefill attention: variants of libdli_b200.so whose softmax loop leaves one piece of
work out (results are WRONG on purpose; only the time is of interest).  Built here, run on the GPU by abl_attn_time.py."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "distributed-llm-inferencing_b200")
OUT = os.path.join(ROOT, "scratch", "_abl")
src = open(os.path.join(PKG, "csrc", "attn_prefill_tc.cu")).read()

EDITS = {
    "base": [],
    "nostore": [("st_16B(dst + gq * 8, pack8(v));", "if (inv == 123.f) st_16B(dst + gq * 8, pack8(v));"),
                ("st_16B(dst + 32 + gq * 8, pack8(v));", "if (inv == 123.f) st_16B(dst + 32 + gq * 8, pack8(v));")],
    "nomask": [("if (need_mask) {", "if (need_mask && scale_log2 == 123.f) {")],
    "noepi": [("if (qrel < T) {", "if (qrel < T && inv == 123.f) {")],
}
EDITS["nomask_noepi"] = EDITS["nomask"] + EDITS["noepi"]
os.makedirs(OUT, exist_ok=True)
objs = [os.path.join(PKG, "build", f) for f in os.listdir(os.path.join(PKG, "build")) if f.endswith(".o") and f != "attn_prefill_tc.o"]
for name, edits in EDITS.items():
    s = src
    for a, b in edits:
        assert a in s, (name, a)
        s = s.replace(a, b)
    cu = os.path.join(OUT, f"attn_{name}.cu")
    open(cu, "w").write(s)
    o = cu.replace(".cu", ".o")
    subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC",
                           "-I", os.path.join(PKG, "csrc"), "-c", cu, "-o", o])
    subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", os.path.join(OUT, f"libdli_{name}.so"), o, *objs])
    print("built", name)
