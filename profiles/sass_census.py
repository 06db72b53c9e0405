"""SASS instruction census of the library's objects: which kernels use which tensor / copy path.

    python profiles/sass_census.py > profiles/sass_census.txt      (after <pkg>/build.py; no GPU needed)
"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "distributed-real-time-chat-and-collaboration-tool_b200", "build")
OBJECTS = ["gemm_tcgen05", "decode_fused", "attention_tc", "attention", "ops", "engine"]
COLS = ["UTCHMMA", "UTCHMMA.2CTA", "UTMALDG", "UTMASTG", "UTMAREDG", "LDTM", "HMMA", "LDGSTS", "SYNCS", "REDUX"]
KERNELS = re.compile(r"attn|ln_gemm|gemm_redadd")


def sass(obj):
    return subprocess.run(["cuobjdump", "-sass", os.path.join(BUILD, obj + ".o")], capture_output=True, text=True).stdout


def count(text, mnemonic):
    if mnemonic == "UTCHMMA.2CTA":
        return len(re.findall(r"\bUTCHMMA\.2CTA", text))
    return len(re.findall(r"\b" + re.escape(mnemonic) + r"\b", text))


print("# SASS instruction census of libb200llm.so's objects (cuobjdump -sass <obj>, occurrences of a mnemonic), sm_100a")
print("# UTCHMMA = tcgen05.mma (kind::f16), .2CTA = cta_group::2; UTMALDG/UTMASTG/UTMAREDG = TMA load / store / reduce-add;")
print("# LDTM = tcgen05.ld; HMMA = mma.sync (legacy tensor path); LDGSTS = cp.async; SYNCS = mbarrier ops; REDUX = redux.sync")
print("%-16s" % "object" + "".join("%14s" % c for c in COLS))
per_kernel = []
for obj in OBJECTS:
    text = sass(obj)
    print("%-16s" % obj + "".join("%14d" % count(text, c) for c in COLS))
    for block in re.split(r"\n\s*Function : ", text)[1:]:
        name = block.split("\n", 1)[0].strip()
        if KERNELS.search(name):
            c = collections.OrderedDict((m, count(block, m)) for m in COLS)
            per_kernel.append((name, c))
print()
print("# per kernel (attention and fused decode kernels): non-zero counts")
for name, c in per_kernel:
    short = re.sub(r"^_ZN\w*?\d+_GLOBAL__N__\w+?\d+", "", name)[-96:]
    print("  %-96s %s" % (short, " ".join(f"{k}={v}" for k, v in c.items() if v)))
