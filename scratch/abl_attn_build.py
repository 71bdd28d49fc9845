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
    "noqk": [("""          umma_bf16(tS + st * 128, make_smem_desc(sQ + off), make_smem_desc(sK + st * kTileBytes2 + off), idesc_s,
                    kk > 0 ? 1u : 0u);""", "          if (off == 0xffffffu) umma_bf16(tS + st * 128, make_smem_desc(sQ + off), make_smem_desc(sK + st * kTileBytes2 + off), idesc_s, kk > 0 ? 1u : 0u);")],
    "nopv": [("          umma_bf16(tO, adesc, bdesc, idesc_o, (it > 0 || kk > 0) ? 1u : 0u);", "          if (it == 0x7fffff) umma_bf16(tO, adesc, bdesc, idesc_o, (it > 0 || kk > 0) ? 1u : 0u);")],
    "notma": [("""        mbar_expect_tx(k_full + 8 * st, kTileBytes2);
        tma_load_2d(sK + st * kTileBytes2, &tm, k_full + 8 * st, (H + kvh) * kTcHd, krow);
        tma_load_2d(sK + st * kTileBytes2 + kSpanBytes, &tm, k_full + 8 * st, (H + kvh) * kTcHd + 64, krow);""",
               "        mbar_expect_tx(k_full + 8 * st, 0);"),
              ("""        mbar_expect_tx(v_full + 8 * st, kTileBytes2);
        tma_load_2d(sV + st * kTileBytes2, &tm, v_full + 8 * st, (H + KV + kvh) * kTcHd, krow);
        tma_load_2d(sV + st * kTileBytes2 + kSpanBytes, &tm, v_full + 8 * st, (H + KV + kvh) * kTcHd + 64, krow);""",
               "        mbar_expect_tx(v_full + 8 * st, 0);")],
}
EDITS["nomma"] = EDITS["noqk"] + EDITS["nopv"]
EDITS["nomma_notma"] = EDITS["nomma"] + EDITS["notma"]
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
