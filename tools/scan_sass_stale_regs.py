"""Heuristic scan for the miscompile seen in the grouped generic epilogue: after a 16-byte global load into Rn..Rn+3,
uniform-register moves (R2UR) within the next instructions read a register that is neither part of the loaded quad nor
written since the load, while at least one R2UR reads the quad (i.e. the sequence is unpacking that load).
Validated on the defective build (commit 2ac0960: four hits per layout in gemm_kernel<128, 5, 4, ., ., 1, 8>, R9 / R11
and R17 / R19 after `LDG.E.128 R4` / `R12`); the fixed libraries (bf16 and fp32 check) give 0 hits.
Usage: cuobjdump -sass k3m_b200/libk3m_sm100a.so > /tmp/all.sass; python tools/scan_sass_stale_regs.py /tmp/all.sass"""
import re, sys
ins = re.compile(r"^\s*/\*[0-9a-f]+\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)\s+(.*?);")
def regs(s): return [int(x) for x in re.findall(r"\bR(\d+)\b", s)]
def scan(lines, name, out):
    code = []
    for l in lines:
        m = ins.match(l)
        if m: code.append((m.group(2), m.group(3)))
    for i, (op, args) in enumerate(code):
        if not op.startswith("LDG.E.128") : continue
        dst = regs(args.split(",")[0])
        if not dst: continue
        n = dst[0]; quad = set(range(n, n + 4)); written = set(quad); hit_quad = False; bad = []
        for op2, a2 in code[i + 1:i + 40]:
            if op2.startswith(("BRA", "EXIT", "BSYNC", "CALL", "RET", "BSSY", "WARPSYNC")): break
            parts = a2.split(",")
            d = regs(parts[0]) if parts else []
            srcs = regs(",".join(parts[1:]))
            if op2 == "R2UR":
                for r in srcs:
                    if r in quad: hit_quad = True
                    elif r not in written: bad.append((op2, a2))
            if op2.startswith(("ST", "RED", "ATOM")):   # no register destination
                continue
            for r in d:
                written.add(r)
                if ".64" in op2 or "WIDE" in op2: written.add(r + 1)
                if ".128" in op2: written.update(range(r, r + 4))
        if hit_quad and bad:
            out.append((name, i, op + " " + args, bad[:3]))
cur, buf, out = None, [], []
for l in open(sys.argv[1], errors="replace"):
    m = re.match(r"\s*Function : (\S+)", l)
    if m:
        if cur: scan(buf, cur, out)
        cur, buf = m.group(1), []
    else:
        buf.append(l)
if cur: scan(buf, cur, out)
print(len(out), "suspicious sequences")
for o in out[:40]: print(o)
