"""Timeline of one decode layer at batch 1: variants of gemv.cu and attn_decode.cu whose first and last CTA stamp
%globaltimer at entry, after the dependency wait and at exit.  Instrumentation only (scratch/_abl, not the product)."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "distributed-llm-inferencing_b200")
OUT = os.path.join(ROOT, "scratch", "_abl")
os.makedirs(OUT, exist_ok=True)
HDR = '''
__device__ unsigned long long g_tl_%(n)s[1 << 16];
__device__ unsigned int g_tl_n_%(n)s;
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %%0, %%%%globaltimer;" : "=l"(t)); return t; }
#define TL_BEGIN(kind) unsigned long long tl0__ = 0, tl1__ = 0; const bool tl__ = threadIdx.x == 0 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1) && blockIdx.y == 0 && blockIdx.z == 0; const int tlk__ = (kind); if (tl__) tl0__ = gtime();
#define TL_WAITED() if (tl__) tl1__ = gtime();
#define TL_END() if (tl__) { unsigned int i = atomicAdd(&g_tl_n_%(n)s, 1u); if (i < (1u << 14)) { g_tl_%(n)s[4 * i] = ((unsigned long long)tlk__ << 32) | blockIdx.x; g_tl_%(n)s[4 * i + 1] = tl0__; g_tl_%(n)s[4 * i + 2] = tl1__; g_tl_%(n)s[4 * i + 3] = gtime(); } }
extern "C" int dli_dbg_tl_%(n)s(unsigned long long* host, unsigned int* n) { cudaMemcpyFromSymbol(n, g_tl_n_%(n)s, 4); unsigned int z = 0; cudaMemcpyToSymbol(g_tl_n_%(n)s, &z, 4); return (int)cudaMemcpyFromSymbol(host, g_tl_%(n)s, sizeof(unsigned long long) << 16); }
'''


def patch(src, name, edits):
    s = open(os.path.join(PKG, "csrc", src)).read()
    s = s.replace("namespace dli {", (HDR % {"n": name}) + "\nnamespace dli {", 1)
    for a, b in edits:
        assert s.count(a) == 1, (src, s.count(a), a)
        s = s.replace(a, b)
    cu = os.path.join(OUT, f"tl_{src}")
    open(cu, "w").write(s)
    o = cu.replace(".cu", ".o")
    subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC",
                           "-I", os.path.join(PKG, "csrc"), "-I", os.path.join(ROOT, "include"), "-c", cu, "-o", o])
    return o


objs = [patch("gemv.cu", "gemv", [
    ("  const uint4* wrow[kGemvRows];\n  if (pf_l2 && ngroups > 0", "  TL_BEGIN(N)\n  const uint4* wrow[kGemvRows];\n  if (pf_l2 && ngroups > 0"),
    ("  pdl_wait();\n  pdl_trigger();  // the grid is resident at once", "  pdl_wait();\n  TL_WAITED()\n  pdl_trigger();  // the grid is resident at once"),
    ("  if (cand) {\n    // per-CTA candidate", "  TL_END()\n  if (cand) {\n    // per-CTA candidate"),
]), patch("attn_decode.cu", "attn", [
    ("  if (!fused) {\n    pdl_wait();\n    pdl_trigger();\n  }\n  const int seq_len = seq_lens[b];",
     "  TL_BEGIN(1)\n  unsigned long long tA__ = 0, tB__ = 0, tC__ = 0, tD__ = 0, tE__ = 0;\n  if (!fused) {\n    pdl_wait();\n    pdl_trigger();\n  }\n  const int seq_len = seq_lens[b];"),
    ("  if (fused) {\n    pdl_wait();  // q and the new token's k, v\n    pdl_trigger();\n  }",
     "  if (tl__) tA__ = gtime();\n  if (fused) {\n    pdl_wait();  // q and the new token's k, v\n    pdl_trigger();\n  }\n  TL_WAITED()"),
    ("    else cp_async_wait<NS - 1>();             // all but the newest NS-1 groups have landed: this tile is in\n    __syncthreads();",
     "    else cp_async_wait<NS - 1>();             // all but the newest NS-1 groups have landed: this tile is in\n    __syncthreads();\n    if (tl__ && tile == t_begin) tB__ = gtime();"),
    ("      if (!kStreamQ) {\n#pragma unroll\n        for (int kk = 0; kk < HD / 16; ++kk) {\n          const int mi = lane >> 3;\n          ldmatrix_x4(qf[kStreamQ ? 0 : kk],",
     "      if (tl__) tC__ = gtime();\n      if (!kStreamQ) {\n#pragma unroll\n        for (int kk = 0; kk < HD / 16; ++kk) {\n          const int mi = lane >> 3;\n          ldmatrix_x4(qf[kStreamQ ? 0 : kk],"),
    ("    __syncthreads();  // every warp is done with this stage: refill it", "    if (tl__ && tile == t_begin) tD__ = gtime();\n    __syncthreads();  // every warp is done with this stage: refill it"),
    ("  __syncthreads();\n  for (int idx = tid; idx < G * HD; idx += kAttnThreads) {\n    const int row = idx / HD, col = idx - row * HD;\n    float mmax = -INFINITY;",
     "  __syncthreads();\n  if (tl__) tE__ = gtime();\n  for (int idx = tid; idx < G * HD; idx += kAttnThreads) {\n    const int row = idx / HD, col = idx - row * HD;\n    float mmax = -INFINITY;"),
    ("  if (num_splits > 1 && cluster_merge) {",
     "  TL_END()\n  if (tl__) { unsigned int i = atomicAdd(&g_tl_n_attn, 1u); if (i < (1u << 14)) { g_tl_attn[4 * i] = (2ull << 32) | blockIdx.x; g_tl_attn[4 * i + 1] = tA__; g_tl_attn[4 * i + 2] = tB__; g_tl_attn[4 * i + 3] = gtime(); } i = atomicAdd(&g_tl_n_attn, 1u); if (i < (1u << 14)) { g_tl_attn[4 * i] = (4ull << 32) | blockIdx.x; g_tl_attn[4 * i + 1] = tC__; g_tl_attn[4 * i + 2] = tD__; g_tl_attn[4 * i + 3] = tE__; } }\n  if (num_splits > 1 && cluster_merge) {"),
    ("  const int lane = threadIdx.x & 31;\n  pdl_wait();\n  pdl_trigger();\n  if (bh >= n_bh) return;",
     "  const int lane = threadIdx.x & 31;\n  TL_BEGIN(3)\n  pdl_wait();\n  TL_WAITED()\n  pdl_trigger();\n  if (bh >= n_bh) return;"),
    ("    *reinterpret_cast<uint32_t*>(o + i) = pack_bf16(acc[i] * inv, acc[i + 1] * inv);\n}",
     "    *reinterpret_cast<uint32_t*>(o + i) = pack_bf16(acc[i] * inv, acc[i + 1] * inv);\n  TL_END()\n}"),
])]
HDR = HDR.replace("(blockIdx.x == 0 || blockIdx.x == gridDim.x - 1) && blockIdx.y == 0 && blockIdx.z == 0", "blockIdx.x == 0 && (blockIdx.y %% 16 == 0 || blockIdx.y == gridDim.y - 1) && (blockIdx.z == 0 || blockIdx.z == gridDim.z - 1)")
HDR = HDR.replace("| blockIdx.x;", "| (blockIdx.y * 16 + blockIdx.z);")
objs.append(patch("gemm_tcgen05.cu", "gemm", [
    ("  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;\n  // split-K (gridDim.z > 1, swap mode)",
     "  TL_BEGIN(100000 + N)\n  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;\n  // split-K (gridDim.z > 1, swap mode)"),
    ("      pdl_wait();\n      for (int kb = 0; kb < num_kb; ++kb) {\n        const int s = kb % Cfg::kStages;\n        const uint32_t ph = (kb / Cfg::kStages) & 1;\n        const uint32_t sA = smem_addr(smem + s * Cfg::kStageBytes);\n        const uint32_t sB = sA + Cfg::kABytes;\n        if (kb < pre) {",
     "      pdl_wait();\n      TL_WAITED()\n      for (int kb = 0; kb < num_kb; ++kb) {\n        const int s = kb % Cfg::kStages;\n        const uint32_t ph = (kb / Cfg::kStages) & 1;\n        const uint32_t sA = smem_addr(smem + s * Cfg::kStageBytes);\n        const uint32_t sB = sA + Cfg::kABytes;\n        if (kb < pre) {"),
    ("    tcgen05_fence_before();\n  }\n  __syncthreads();\n  if (warp == 1) {\n    tcgen05_fence_after();\n    tmem_dealloc(tmem_base, Cfg::kTmemCols);\n  }\n}\n\n\n// ====",
     "    tcgen05_fence_before();\n  }\n  __syncthreads();\n  TL_END()\n  if (warp == 1) {\n    tcgen05_fence_after();\n    tmem_dealloc(tmem_base, Cfg::kTmemCols);\n  }\n}\n\n\n// ===="),
]))
others = [os.path.join(PKG, "build", f) for f in os.listdir(os.path.join(PKG, "build")) if f.endswith(".o") and f not in ("gemv.o", "attn_decode.o", "gemm_tcgen05.o")]
subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", os.path.join(OUT, "libdli_tl.so"), *objs, *others])
print("built timeline library")
