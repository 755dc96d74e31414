#!/usr/bin/env python
"""Static model of an emitted kernel from its SASS (no GPU needed).

  python tools/sass_model.py <lib.so> <kernel-name-substring> [--buckets 512]

Reads `cuobjdump -sass`, decodes the control word of every instruction (sm_70+ layout, /opt/skills/guides/
B300_MICROARCH.md "Terminology": stall = bits 105-108, yield = 109, write/read barrier = 110-112 / 113-115,
wait mask = 116-121, operand reuse = 122-125) and prints
  * the opcode histogram (FFMA / FMUL / FADD / ... -> executed FP32 operations per lane),
  * sum of stall fields = issue-to-issue lower bound for ONE warp running alone (fixed-latency dependencies
    are encoded there: the hardware does not interlock them), i.e. the static CPI of the schedule,
  * the same per bucket of N instructions (to see where the schedule is dependency-bound),
  * evidence mnemonics for the Blackwell paths (UBLKCP = TMA bulk, SYNCS = mbarrier, ...).
Used to compare emitter variants on the CPU box before spending GPU minutes, and to produce the committed
opcode summaries under profiles/.
"""
import collections
import re
import subprocess
import sys


def dump(lib: str, pattern: str):
    txt = subprocess.run(["cuobjdump", "-sass", lib], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    funcs = {}
    name = None
    pend = None
    for line in txt.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = m.group(1)
            funcs[name] = []
            continue
        if name is None:
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);\s*/\* (0x[0-9a-f]{16}) \*/", line)
        if m:
            pend = [int(m.group(1), 16), m.group(2).strip(), int(m.group(3), 16)]
            continue
        m = re.match(r"\s+/\* (0x[0-9a-f]{16}) \*/", line)
        if m and pend is not None:
            pend.append(int(m.group(1), 16))
            funcs[name].append(tuple(pend))
            pend = None
    sel = {k: v for k, v in funcs.items() if pattern in k}
    return sel


def opcode(text: str) -> str:
    t = text
    if t.startswith("@"):
        t = t.split(None, 1)[1]
    return t.split()[0].split(".")[0]


def bank_cycles(text: str) -> int:
    """Dispatch cycles of one instruction under the register-file port model of B300_MICROARCH.md ("RF banking":
    two banks, even / odd register index, one read port each per cycle; operands taken from the operand-reuse
    cache (.reuse) or given as immediates / constant-bank / uniform registers cost no port):
    max(1, #distinct even registers read, #distinct odd registers read).  64-bit operands (F32x2 pairs, .64) read
    both registers of the aligned pair."""
    t = text
    if t.startswith("@"):
        t = t.split(None, 1)[1]
    parts = t.split(None, 1)
    if len(parts) < 2:
        return 1
    op = parts[0]
    args = [a.strip() for a in parts[1].rstrip(" ;").split(",")]
    srcs = args[1:]
    if op.split(".")[0] in ("STS", "STG", "STL", "ST"):
        srcs = args
    ev, od = set(), set()
    for s_ in srcs:
        if ".reuse" in s_:
            continue
        for m in re.finditer(r"(?<![A-Za-z])R(\d+)((?:\.[A-Za-z0-9_]+)*)", s_):
            r = int(m.group(1))
            wide = "F32x2" in m.group(2) or ".64" in m.group(2)
            for q in ((r, r + 1) if wide else (r,)):
                (ev if q % 2 == 0 else od).add(q)
    return max(1, len(ev), len(od))


def analyse(ins, buckets=512, out=sys.stdout):
    n = len(ins)
    hist = collections.Counter(opcode(t) for _, t, _, _ in ins)
    full = collections.Counter(t.split()[1].rstrip(";") if t.startswith("@") else t.split()[0] for _, t, _, _ in ins)
    stalls = [(hi >> 41) & 0xF for _, _, _, hi in ins]
    waits = [(hi >> 52) & 0x3F for _, _, _, hi in ins]
    reuse = [(hi >> 58) & 0xF for _, _, _, hi in ins]
    fp = {k: hist.get(k, 0) for k in ("FFMA", "FMUL", "FADD", "FFMA2", "FMUL2", "FADD2", "DFMA", "DMUL", "DADD", "MUFU")}
    flops = 2 * fp["FFMA"] + fp["FMUL"] + fp["FADD"] + 4 * fp["FFMA2"] + 2 * fp["FMUL2"] + 2 * fp["FADD2"] \
        + 2 * fp["DFMA"] + fp["DMUL"] + fp["DADD"]
    print(f"instructions {n}  ({n * 16 / 1024:.1f} KB)", file=out)
    print("fp:", {k: v for k, v in fp.items() if v}, f" flop/lane (static) {flops}", file=out)
    print("top opcodes:", hist.most_common(24), file=out)
    ev = {k: sum(v for kk, v in full.items() if kk.startswith(k)) for k in
          ("UBLKCP", "UBLKPF", "SYNCS", "ACQBULK", "BAR", "LDL", "STL", "LDS", "STS", "LDG", "STG", "CALL", "BRA", "SHFL", "MOV")}
    print("evidence/overhead mnemonics:", {k: v for k, v in ev.items() if v}, file=out)
    tot = sum(stalls)
    print(f"sum(stall) {tot}  static CPI (one warp alone, no scoreboard waits) {tot / max(n, 1):.3f}", file=out)
    print(f"instructions with a scoreboard wait mask {sum(1 for w in waits if w)}; with an operand .reuse flag "
          f"{sum(1 for r in reuse if r)}", file=out)
    FPOPS = ("FFMA", "FMUL", "FADD", "FFMA2", "FMUL2", "FADD2")
    bc = [bank_cycles(t) if opcode(t) in FPOPS else 1 for _, t, _, _ in ins]
    print(f"register-port model: dispatch cycles {sum(bc)} ({sum(bc) / max(n, 1):.3f} per instruction; "
          f"FP instructions needing 2 / 3+ cycles: {sum(1 for c in bc if c == 2)} / {sum(1 for c in bc if c > 2)})", file=out)
    sh = collections.Counter(stalls)
    print("stall histogram:", sorted(sh.items()), file=out)
    if buckets:
        print(f"# per {buckets}-instruction bucket: first_instr  sum(stall)  CPI  FP-instr  waits", file=out)
        for b in range(0, n, buckets):
            s = stalls[b:b + buckets]
            f = sum(1 for _, t, _, _ in ins[b:b + buckets] if opcode(t) in ("FFMA", "FMUL", "FADD", "FFMA2", "FMUL2", "FADD2"))
            w = sum(1 for x in waits[b:b + buckets] if x)
            print(f"{b:7d} {sum(s):7d} {sum(s) / len(s):6.2f} {f:6d} {w:5d}", file=out)
    return {"n": n, "flops": flops, "stall": tot, "fp": fp, "dispatch": sum(bc)}


def main():
    lib, pat = sys.argv[1], sys.argv[2]
    buckets = 512
    if "--buckets" in sys.argv:
        buckets = int(sys.argv[sys.argv.index("--buckets") + 1])
    sel = dump(lib, pat)
    if not sel:
        print("no function matches", pat)
        return 1
    for name, ins in sel.items():
        print("==", name)
        analyse(ins, buckets)
    return 0


if __name__ == "__main__":
    sys.exit(main())
