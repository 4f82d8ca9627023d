"""Instruction mix of a kernel's hottest loop from the built library's SASS.
Usage: python tools/sass_loop_mix.py <substring of the mangled kernel name> [lib]
Finds the innermost backward branch with the largest body and counts opcodes in it (per loop iteration)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "n-body-simulation_b200", "libnbody_b200.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
blocks = sass.split("Function : ")[1:]
pick = [b for b in blocks if sys.argv[1] in b.splitlines()[0]]
assert pick, "kernel not found"
for b in pick:
    name = b.splitlines()[0]
    ins = []   # (addr, opcode, text)
    for line in b.splitlines():
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m:
            text = m.group(2)
            text = re.sub(r"^@!?U?P\d+\s+", "", text)
            ins.append((int(m.group(1), 16), text.split()[0], text))
    loops = []
    for a, op, text in ins:
        if op.startswith("BRA"):
            m = re.search(r"0x([0-9a-f]+)", text)
            if m and int(m.group(1), 16) < a:
                loops.append((int(m.group(1), 16), a))
    print("==", name)
    for lo, hi in sorted(loops, key=lambda x: x[0] - x[1])[:int(os.environ.get("NLOOPS", "2"))]:
        body = [i for i in ins if lo <= i[0] <= hi]
        c = collections.Counter(re.sub(r"\..*", "", i[1]) if not i[1].startswith("MUFU") else i[1] for i in body)
        print(f"loop 0x{lo:x}..0x{hi:x}: {len(body)} instructions")
        print("   ", ", ".join(f"{k} {v}" for k, v in c.most_common()))
