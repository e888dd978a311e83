#!/usr/bin/env python
"""Instruction mix of the episode kernels from the built library (cuobjdump -sass): per kernel the opcode histogram, the counts of
the mnemonics that prove the code path (UTCHMMA / LDTM / STTM / UBLKCP / HMMA / MUFU.TANH / SYNCS) and the listing around the first
tensor instruction.  Writes profiles/r02_sass_<tag>.txt.   python scripts/sass_summary.py"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "lunar_lander_b200", "libll_gpu.so")
WANT = [("tc05_32_128_32", r"ll_episode_tc05_kernel.*Li5ELi32ELi128ELi32ELi4EEEELi0E"), ("mma_resume_32_128_32", r"ll_episode_mma_resume_kernel.*Li5ELi32ELi128ELi32ELi4E"),
        ("mmat_resume_32_128_32", r"ll_episode_mmat_resume_kernel.*Li5ELi32ELi128ELi32ELi4EEEELi2E"), ("exactw_f64_16_16", r"ll_episode_exactw_kernelId.*Li5ELi16ELi16ELi4E"),
        ("mma_32_128_32", r"ll_episode_mma_kernel.*Li5ELi32ELi128ELi32ELi4E.*Li0E"), ("mmat_16_64_32", r"ll_episode_mmat_kernel.*Li5ELi16ELi64ELi32ELi4E.*Li1E"),
        ("warp_8_16_8", r"ll_episode_warp_kernel.*Li5ELi8ELi16ELi8ELi4E"), ("exact_f64", r"ll_episode_kernel.*IdLi0E"),
        ("create_population_f32_fast", r"ll_create_population_f32_kernelILb1E"), ("create_population_f32_exact", r"ll_create_population_f32_kernelILb0E")]
KEY = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UBLKCP", "UTMALDG", "HMMA", "MOVM", "MUFU.TANH", "MUFU", "SYNCS", "BAR.SYNC", "FFMA", "DFMA", "DMUL", "DADD", "SHFL", "LDS", "STS"]

sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s*Function : ", sass)[1:]
os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
for tag, pat in WANT:
    hit = [f for f in funcs if re.search(pat, f.split("\n", 1)[0])]
    if not hit:
        print(tag, "not found")
        continue
    f = hit[0]
    name = f.split("\n", 1)[0]
    ins = re.findall(r"^\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", f, flags=re.M)
    hist = collections.Counter(i.split(".")[0] for i in ins)
    full = collections.Counter(ins)
    lines = [l for l in f.split("\n") if re.match(r"^\s+/\*[0-9a-f]{4,}\*/", l)]
    first = next((i for i, l in enumerate(lines) if re.search(r"UTCHMMA|HMMA|DFMA|DMUL|SHFL", l)), 0)
    with open(os.path.join(ROOT, "profiles", f"r02_sass_{tag}.txt"), "w") as out:
        out.write(f"# {name}\n# cuobjdump -sass lunar_lander_b200/libll_gpu.so (sm_100a), {len(ins)} instructions in the kernel\n")
        out.write("# proof mnemonics: " + ", ".join(f"{k} {sum(v for n, v in full.items() if n.startswith(k))}" for k in KEY) + "\n")
        out.write("# opcode histogram: " + ", ".join(f"{k} {v}" for k, v in hist.most_common(40)) + "\n")
        out.write(f"# listing: 120 instructions from the first tensor / FMA-chain instruction (instruction {first})\n")
        for l in lines[max(0, first - 10): first + 110]:
            out.write(re.sub(r"\s+/\* 0x[0-9a-f]+ \*/\s*$", "", l).rstrip() + "\n")
    print(tag, len(ins), {k: sum(v for n, v in full.items() if n.startswith(k)) for k in ("UTCHMMA", "LDTM", "STTM", "HMMA", "MUFU")})
