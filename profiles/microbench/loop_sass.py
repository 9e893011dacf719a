#!/usr/bin/env python3
"""Lists, for every kernel of a binary, the opcode histogram of its innermost backward-branch
loop (cuobjdump -sass).  Usage: loop_sass.py <binary> [name-filter]"""
import collections, re, subprocess, sys
sass = subprocess.run(["cuobjdump", "-sass", sys.argv[1]], capture_output=True, text=True).stdout
flt = sys.argv[2] if len(sys.argv) > 2 else ""
name, lines = None, []
def flush():
    if not name or flt not in name: return
    ins = [(int(m.group(1), 16), m.group(2)) for m in (re.search(r"/\*([0-9a-f]{4,})\*/\s+(.*?);", l) for l in lines) if m]
    best = None
    for addr, text in ins:
        m = re.search(r"\bBRA\b.*?(0x[0-9a-f]+)", text)
        if m and int(m.group(1), 16) < addr:
            tgt = int(m.group(1), 16)
            if best is None or addr - tgt < best[1] - best[0]: best = (tgt, addr)
    if best is None: print(name, "no loop"); return
    body = [t for a, t in ins if best[0] <= a <= best[1]]
    ops = collections.Counter(re.sub(r"^@!?U?P\d+\s+", "", t).split()[0].split(".")[0] for t in body)
    print(name[:110]); print("   ", len(body), dict(ops.most_common()))
for l in sass.splitlines():
    m = re.search(r"Function : (\S+)", l)
    if m: flush(); name, lines = m.group(1), []
    else: lines.append(l)
flush()
