"""Turns the r1c gpurun outputs (gpurun_out/f2, gpurun_out/*r1g*, ar_check) into profiles/r1c_*.md."""
import json, os, glob, shutil, subprocess, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(R, "gpurun_out"), os.path.join(R, "profiles")


def rows(path):
    out = []
    if os.path.exists(path):
        for l in open(path):
            l = l.strip()
            if l.startswith("{"):
                try:
                    out.append(json.loads(l))
                except Exception:
                    pass
    return out


def decode_table():
    md = ["# r1c: Llama-3-8B W4A16 (gptq_marlin g128) decode tokens/s, ctx 1024, CUDA-graph replay (benchmarks/decode_bench.py)",
          "", "Sampler/scheduler/tokenizer excluded; lm_head fp16 torch matmul + argmax. `glue` = who runs RMSNorm / RoPE+cache write / SiLU*mul;",
          "`PDL` = programmatic dependent launch across the chain; `all_reduce` = TP partial-sum reduction (messages above the",
          "world-size one-shot limit -- 2 MiB / 1 MiB / 512 KiB at world 2 / 4 / 8, i.e. batch 256 at TP4 -- go through NCCL in every configuration;",
          "the fused add/RMSNorm kernel is taken up to 144 rows).", "",
          "| TP | glue | PDL | all_reduce | batch | ms/step | tokens/s (job) | tokens/s/GPU | % HBM peak (weights+KV+lm_head) |", "|---|---|---|---|---|---|---|---|---|"]
    srcs = [("f2/decode_tp1_torchglue.jsonl", "torch ops", "off (torch)"), ("f2/decode_tp1_fused_nopdl.jsonl", "ours", "off"),
            ("f2/decode_tp1_fused_pdl.jsonl", "ours", "on")]
    for n in (2, 4, 8):
        srcs += [(f"decode_tp{n}_r1g_nccl.jsonl", "ours", "on"), (f"decode_tp{n}_r1g_oneshot.jsonl", "ours", "on"),
                 (f"decode_tp{n}_r1h_fusednorm.jsonl", "ours", "on")]
    for f, glue, pdl in srcs:
        for r in rows(os.path.join(G, f)):
            md.append(f"| {r['tp']} | {glue} | {pdl} | {(r.get('all_reduce') or '-') + (' + fused add/RMSNorm' if r.get('ar_norm_fused') else '')} | {r['batch']} | {r['ms_per_step']:.3f} | "
                      f"{r['tokens_per_s']:.0f} | {r['tokens_per_s_per_gpu']:.0f} | {100 * r['hbm_frac']:.1f} |")
    open(os.path.join(P, "r1c_decode_tokens.md"), "w").write("\n".join(md) + "\n")


def glue_table():
    md = ["# r1c: decode-side glue kernels on B200 (benchmarks/glue_bench.py; CUDA-graph replay, buffer sets rotated past L2 at prefill sizes)", "",
          "| op | tokens | us/call | algorithmic bytes | GB/s | % HBM peak | L2 exceeded |", "|---|---|---|---|---|---|---|"]
    for r in rows(os.path.join(G, "f2/glue_bench.jsonl")):
        md.append(f"| {r['op']} | {r['tokens']} | {r['us_per_call']} | {r['bytes']} | {r['GBps']} | {100 * r['hbm_frac']:.1f} | {r['l2_exceeded']} |")
    open(os.path.join(P, "r1c_glue_bench.md"), "w").write("\n".join(md) + "\n")


def ar_table():
    md = ["# r1c: one-shot NVLink all-reduce vs NCCL (benchmarks/ar_check.py; 64 calls per CUDA graph, max over ranks, fp16)", "",
          "Correctness line of the same run first (bit-identical across ranks, bit-exact vs rank-order fp32 sum, graph replay).", ""]
    for n in (2, 4, 8):
        rs = rows(os.path.join(G, f"ar_check_n{n}.jsonl"))
        if not rs:
            continue
        md += [f"## world {n}: {json.dumps(rs[0])}", "", "| bytes | one-shot us | NCCL us | path |", "|---|---|---|---|"]
        for r in rs[1:]:
            path = "flag-in-data push (LL)" if r["bytes"] <= 2 * 2**20 else "staged one-shot"
            md.append(f"| {r['bytes']} | {r['us_one_shot']} | {r['us_nccl']} | {path} |")
        md.append("")
    open(os.path.join(P, "r1c_all_reduce.md"), "w").write("\n".join(md) + "\n")


def gemm_table():
    rs = rows(os.path.join(G, "f2/gemm_all.jsonl"))
    if not rs:
        return
    md = ["# r1c: quantized-linear microbench on B200 after PDL (benchmarks/gemm_bench.py --op all; CUDA-graph replay, weights rotated past L2)", "",
          "| op | layer | M | KxN | us/call | GB/s (algorithmic) | % HBM peak | TFLOP/s | % tensor peak |", "|---|---|---|---|---|---|---|---|---|"]
    for r in rs:
        md.append(f"| {r['op']} | {r.get('layer','')} | {r['m']} | {r['k']}x{r['n']} | {r['us']:.1f} | {r['GBps']:.0f} | {100*r['hbm_frac']:.1f} | {r['TFLOPs']:.0f} | {100*r['tensor_frac']:.1f} |")
    open(os.path.join(P, "r1c_gemm_bench.md"), "w").write("\n".join(md) + "\n")


def copies():
    for src, dst in (("f2/pytest_gpu.log", "r1c_pytest_gpu.log"), ("f2/smoke.log", "r1c_smoke.log"),
                     ("f2/bench_fp16.json", "r1c_bench_attention_fp16.json"), ("f2/bench_fp8.json", "r1c_bench_attention_fp8.json"),
                     ("f2/bench_ref.json", "r1c_bench_reference_cpu.json"), ("f2/launches.csv", "r1c_bench_launches.csv")):
        s = os.path.join(G, src)
        if os.path.exists(s):
            shutil.copy(s, os.path.join(P, dst))
    for rep, title, dst in (("f2/glue_add_rms.ncu-rep", "r1c: ncu --set full, rms_norm fp16 (the FUSED=0 instantiation was the one captured), 16384 tokens x 4096", "r1c_glue_add_rms_ncu.md"),
                            ("f2/glue_silu.ncu-rep", "r1c: ncu --set full, silu_and_mul fp16, 16384 tokens x 2x14336", "r1c_glue_silu_ncu.md"),
                            ("f2/glue_rotary_cache.ncu-rep", "r1c: ncu --set full, rotary_embedding_and_cache fp16, 16384 tokens", "r1c_glue_rotary_cache_ncu.md"),
                            ("f2/marlin_m16_gateup.ncu-rep", "r1c: ncu --set full, gptq_marlin_gemm W4A16 g128 decode kernel with PDL producer, gate_up M=16", "r1c_marlin_decode_ncu.md")):
        s = os.path.join(G, rep)
        if os.path.exists(s):
            txt = subprocess.run([sys.executable, os.path.join(P, "summarize_ncu.py"), s, title], capture_output=True, text=True).stdout
            open(os.path.join(P, dst), "w").write(txt)


if __name__ == "__main__":
    decode_table(); glue_table(); ar_table(); gemm_table(); copies()
    print(sorted(f for f in os.listdir(P) if f.startswith("r1c")))
