"""Count the instructions of a kernel's hot loops in the SHIPPED libgpuhash.so, by issue pipe.

The hashing kernels are bound by the INT32 ALU pipe (LOP3 / SHF / IADD3 / PRMT / LEA: 64 lanes per clock
per SM), so the figure every roofline in this repo rests on is "ALU-pipe warp-instructions per 64-byte
block". This tool reads it off the SASS instead of trusting a hand count: for every backward branch of
a function it lists the loop body's instruction mix. bench.py imports loop_table() at run time and
records the count together with the library's hash.

usage: python tests/tools/sass_count.py [substring of the kernel name ...]
"""
import collections
import hashlib
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
LIB = os.path.join(ROOT, "downloader-go_b200", "libgpuhash.so")

ALU = ("LOP3", "SHF", "IADD3", "IADD", "PRMT", "LEA", "ISETP", "SEL", "IMNMX", "VIMNMX", "MOV", "POPC", "FLO",
       "BMSK", "SGXT", "PLOP3", "LOP", "IABS", "VABSDIFF", "FMNMX", "R2P", "P2R")
FMA = ("IMAD", "FFMA", "FMUL", "FADD", "HFMA2", "VIADD")  # MOV may be emitted as IMAD.MOV: counted under IMAD
LSU = ("LDS", "STS", "LDG", "STG", "LD", "ST", "LDC", "LDL", "STL", "ATOM", "RED", "ATOMS", "ATOMG", "LDSM")


def pipe_of(op):
    base = op.split(".")[0]
    if base in ("IMAD", "VIADD") or base in FMA:
        return "fma"
    if base in ALU:
        return "alu"
    if base in LSU or base.startswith("UTMA") or base.startswith("SYNCS") or base.startswith("UBLK"):
        return "lsu"
    if base in ("BRA", "BSSY", "BSYNC", "EXIT", "WARPSYNC", "NANOSLEEP", "YIELD", "BAR", "CALL", "RET", "BREAK"):
        return "ctl"
    if base.startswith("U") or base in ("S2R", "S2UR", "R2UR", "CS2R", "ELECT", "VOTE", "VOTEU", "REDUX", "SHFL", "MATCH"):
        return "uni"
    return "other"


def disassemble(lib=LIB):
    out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    funcs, name = {}, None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = m.group(1)
            funcs[name] = []
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m and name:
            addr = int(m.group(1), 16)
            text = m.group(2).strip()
            text = re.sub(r"^@!?U?P\d+\s+", "", text)
            funcs[name].append((addr, text))
    return funcs


def demangle(names):
    try:
        out = subprocess.run(["cu++filt"] + list(names), capture_output=True, text=True, check=True).stdout.split("\n")
        return dict(zip(names, out))
    except Exception:
        return {n: n for n in names}


def loops(instrs):
    """[(start_addr, end_addr, Counter(pipe), Counter(opcode))] for every backward branch, largest first."""
    res = []
    for addr, text in instrs:
        m = re.match(r"BRA(?:\.\S+)?\s+(?:!?U?P\d+,\s*)?`?\(?\.?L?_?x?_?\d*\)?", text)
        tgt = re.search(r"0x([0-9a-f]+)", text) if text.startswith("BRA") else None
        if not tgt:
            continue
        t = int(tgt.group(1), 16)
        if t >= addr:
            continue
        body = [x for a, x in instrs if t <= a <= addr]
        pipes, ops = collections.Counter(), collections.Counter()
        for x in body:
            op = x.split()[0]
            pipes[pipe_of(op)] += 1
            ops[op.split(".")[0]] += 1
        res.append((t, addr, pipes, ops))
    res.sort(key=lambda r: -(r[1] - r[0]))
    return res


def loop_table(lib=LIB, min_instrs=100):
    """{demangled kernel: [{"instrs", "alu", "fma", "lsu", ...}]} for loops of at least min_instrs."""
    funcs = disassemble(lib)
    names = demangle(list(funcs))
    table = {}
    for mangled, instrs in funcs.items():
        rows = []
        for t, a, pipes, ops in loops(instrs):
            n = sum(pipes.values())
            if n < min_instrs:
                continue
            rows.append({"instrs": n, **{k: pipes.get(k, 0) for k in ("alu", "fma", "lsu", "uni", "ctl", "other")},
                         "from": hex(t), "to": hex(a)})
        short = re.sub(r"\(anonymous namespace\)::", "", names.get(mangled, mangled))
        short = re.sub(r"\(CUtensorMap_st.*", "", short)
        short = re.sub(r"^void ", "", short)
        table[short] = rows
    return table


def lib_sha256(lib=LIB):
    h = hashlib.sha256()
    with open(lib, "rb") as f:
        for blk in iter(lambda: f.read(1 << 20), b""):
            h.update(blk)
    return h.hexdigest()


if __name__ == "__main__":
    want = sys.argv[1:]
    tab = loop_table()
    for k in sorted(tab):
        if want and not any(w in k for w in want):
            continue
        print(k)
        for r in tab[k]:
            print("   ", json.dumps(r))
