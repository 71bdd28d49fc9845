"""Phase trace of the tcgen05 prefill attention: a variant of libdli_b200.so whose CTA 0 stamps %clock64 at the hand-off
points of its first item (softmax warp 2 / MMA thread / producer thread) into a __device__ array; scratch/trace_attn_run.py
prints the per-tile differences.  Instrumentation only: nothing here is part of the product."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "distributed-llm-inferencing_b200")
OUT = os.path.join(ROOT, "scratch", "_abl")
s = open(os.path.join(PKG, "csrc", "attn_prefill_tc.cu")).read()


def rep(a, b):
    global s
    assert s.count(a) == 1, (s.count(a), a)
    s = s.replace(a, b)


rep("namespace dli {\n\nconstexpr int kRowThreads", """__device__ long long g_trace[3 * 64 * 8];
#define TR(role, tile, pt) do { if (blockIdx.x == 0 && n == 0 && (tile) < 64) g_trace[((role) * 64 + (tile)) * 8 + (pt)] = clock64(); } while (0)
extern "C" int dli_dbg_trace(long long* host) { return (int)cudaMemcpyFromSymbol(host, g_trace, sizeof(g_trace)); }
namespace dli {

constexpr int kRowThreads""")
# softmax warp 2 lane 0
rep("        mbar_wait(s_full + 8 * st, (g >> 1) & 1);\n        tcgen05_fence_after();\n        float s[kSmCols];",
    "        mbar_wait(s_full + 8 * st, (g >> 1) & 1);\n        if (threadIdx.x == 64) TR(0, g, 0);\n        tcgen05_fence_after();\n        float s[kSmCols];")
rep("        tcgen05_fence_before();\n        __syncwarp();\n        if (lane == 0) mbar_arrive(s_empty + 8 * st);",
    "        if (threadIdx.x == 64) TR(0, g, 1);\n        tcgen05_fence_before();\n        __syncwarp();\n        if (lane == 0) mbar_arrive(s_empty + 8 * st);")
rep("        for (int q = 0; q < kRowThreads; ++q) mx = fmaxf(mx, xrow[q * 128]);",
    "        for (int q = 0; q < kRowThreads; ++q) mx = fmaxf(mx, xrow[q * 128]);\n        if (threadIdx.x == 64) TR(0, g, 2);")
rep("        l_run = l_run * corr + rs;", "        l_run = l_run * corr + rs;\n        if (threadIdx.x == 64 && l_run != 12345.f) TR(0, g, 3);")
rep("        tmem_st_wait();            // P (and O) are in tensor memory before the MMA warp is told",
    "        if (threadIdx.x == 64) TR(0, g, 4);\n        tmem_st_wait();\n        if (threadIdx.x == 64) TR(0, g, 5);")
rep("        if (lane == 0) mbar_arrive(p_full);\n      }", "        if (lane == 0) mbar_arrive(p_full);\n        if (threadIdx.x == 64) TR(0, g, 6);\n      }")
# MMA thread
rep("          if (it + 1 < nt) issue_s(g + 1, it + 2 == nt);", "          TR(1, g, 0);\n          if (it + 1 < nt) issue_s(g + 1, it + 2 == nt);\n          TR(1, g, 1);")
rep("          mbar_wait(v_full + 8 * st, (g >> 1) & 1);\n          tcgen05_fence_after();",
    "          TR(1, g, 2);\n          mbar_wait(v_full + 8 * st, (g >> 1) & 1);\n          TR(1, g, 3);\n          tcgen05_fence_after();")
rep("            umma_commit(v_empty + 8 * st);\n          }\n          __syncwarp();", "            umma_commit(v_empty + 8 * st);\n          }\n          __syncwarp();\n          TR(1, g, 4);")
# producer (unpaged)
rep("          mbar_wait(k_empty + 8 * st, ph ^ 1);\n          mbar_expect_tx(k_full + 8 * st, kTileBytes2);\n          tma_load_2d(sK + st * kTileBytes2, &tm,",
    "          mbar_wait(k_empty + 8 * st, ph ^ 1);\n          TR(2, g, 0);\n          mbar_expect_tx(k_full + 8 * st, kTileBytes2);\n          tma_load_2d(sK + st * kTileBytes2, &tm,")
rep("          mbar_wait(v_empty + 8 * st, ph ^ 1);\n          mbar_expect_tx(v_full + 8 * st, kTileBytes2);\n          tma_load_2d(sV + st * kTileBytes2, &tm,",
    "          mbar_wait(v_empty + 8 * st, ph ^ 1);\n          TR(2, g, 1);\n          mbar_expect_tx(v_full + 8 * st, kTileBytes2);\n          tma_load_2d(sV + st * kTileBytes2, &tm,")
os.makedirs(OUT, exist_ok=True)
cu = os.path.join(OUT, "attn_trace.cu")
open(cu, "w").write(s)
objs = [os.path.join(PKG, "build", f) for f in os.listdir(os.path.join(PKG, "build")) if f.endswith(".o") and f != "attn_prefill_tc.o"]
o = cu.replace(".cu", ".o")
subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC",
                       "-I", os.path.join(PKG, "csrc"), "-c", cu, "-o", o])
subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", os.path.join(OUT, "libdli_trace.so"), o, *objs])
print("built trace")
