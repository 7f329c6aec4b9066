"""SASS evidence: per kernel, count the mnemonics that prove which hardware path it uses
(tcgen05 -> UTC*MMA, TMEM ld/st -> LDTM/STTM, TMA -> UTMALDG/UTMASTG, legacy mma.sync -> HMMA,
cp.async -> LDGSTS, vector RED, peer/system-scope memory ops).  Runs on the CPU box (cuobjdump)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "pretraining_llm_b200", "ops", "libb200llm.so")
PAT = ["UTCHMMA", "UTCBAR", "UTMALDG", "UTMASTG", "UTMAPF", "LDTM", "STTM", "UTCATOMSWS", "HMMA", "LDSM", "LDGSTS",
       "SYNCS", "RED.E", "MUFU.EX2", "UCGABAR", "MEMBAR.SC.SYS", "LD.E.STRONG.SYS", "ST.E.STRONG.SYS", "LDG.E.STRONG.SYS",
       "STG.E.STRONG.SYS", "ERRBAR"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True, check=True).stdout
    kernels = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = re.sub(r"\(anonymous namespace\)::", "", name)
            name = re.sub(r"\(.*", "", name)
            cur = kernels.setdefault(name, collections.Counter())
            continue
        if cur is None:
            continue
        for p in PAT:
            if re.search(r"\b" + re.escape(p), line):
                cur[p] += 1
        cur["_instr"] += 1 if re.search(r"/\*[0-9a-f]{4}\*/", line) else 0
    path = os.path.join(ROOT, "profiles", "sass_evidence.md")
    with open(path, "w") as f:
        f.write("# SASS evidence (cuobjdump -sass libb200llm.so, sm_100a)\n\n"
                "UTCHMMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st (TMEM), UTMALDG/UTMASTG = TMA load/store, UTCBAR = "
                "tcgen05.commit, SYNCS = mbarrier, HMMA = legacy mma.sync, LDGSTS = cp.async, RED = vector reductions.\n\n"
                "| kernel | SASS instrs | " + " | ".join(PAT) + " |\n|---|---|" + "---|" * len(PAT) + "\n")
        for k, c in kernels.items():
            if c["_instr"] == 0:
                continue
            f.write(f"| `{k[:70]}` | {c['_instr']} | " + " | ".join(str(c[p]) if c[p] else "" for p in PAT) + " |\n")
    print("wrote", path, len(kernels), "kernels")


if __name__ == "__main__":
    main()
