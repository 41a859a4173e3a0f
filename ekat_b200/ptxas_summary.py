"""Summarise ekat_b200/lib/ptxas.log: registers / stack / spills per kernel (python -m ekat_b200.ptxas_summary)."""
import os
import re
import subprocess


def summary():
    log = open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", "ptxas.log")).read()
    ents = re.findall(
        r"Compiling entry function '([^']+)' for 'sm_100a'\n.*?Function properties for \1\n\s+(\d+) bytes stack frame, "
        r"(\d+) bytes spill stores, (\d+) bytes spill loads\n.*?Used (\d+) registers", log, re.S)
    out = {}
    for name, stack, ss, sl, regs in ents:
        d = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
        d = re.sub(r"\(.*", "", d).replace("void ekb::", "")
        out[d] = (int(regs), int(stack), int(ss), int(sl))
    return out


if __name__ == "__main__":
    for k, v in sorted(summary().items()):
        print(f"{k:70s} regs={v[0]:3d} stack={v[1]} spill={v[2]}/{v[3]}")
