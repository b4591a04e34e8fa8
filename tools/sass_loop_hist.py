"""Static opcode histogram of the hottest loop of one kernel, from `cuobjdump -sass`.
usage: python tools/sass_loop_hist.py lib.so <substring of the mangled kernel name> [--nth k] [--dump]
The hottest loop is taken to be the longest backward branch span in the function (for gcm_record_kernel: the
per-round loop over the record's 16-byte slots).  A static count: every instruction of the span once, both sides
of its branches included, so it is an upper bound of what one iteration executes."""
import collections
import re
import subprocess
import sys

lib, pat = sys.argv[1], sys.argv[2]
txt = subprocess.run(["/usr/local/cuda/bin/cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s*Function : ", txt)
body = next((f for f in funcs[1:] if pat in f.split("\n", 1)[0]), None)
if body is None:
    sys.exit("kernel not found; have: " + ", ".join(f.split("\n", 1)[0] for f in funcs[1:])[:2000])
ins = []
for m in re.finditer(r"/\*([0-9a-f]{4,})\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)\s*([^;]*);", body):
    ins.append((int(m.group(1), 16), m.group(3), m.group(4)))
spans = []
for a, op, args in ins:
    if op.startswith("BRA"):
        t = re.search(r"0x([0-9a-f]+)", args)
        if t:
            tgt = int(t.group(1), 16)
            if tgt < a:
                spans.append((a - tgt, tgt, a))
spans.sort(reverse=True)
nth = int(sys.argv[sys.argv.index("--nth") + 1]) if "--nth" in sys.argv else 0      # 0 = longest, 1 = next (the nested loop) ...
print("backward branches (span bytes, from, to):", [(hex(s), hex(l), hex(h)) for s, l, h in spans[:6]])
_, lo, hi = spans[nth]
loop = [(a, op, args) for a, op, args in ins if lo <= a <= hi]


def key(op):
    return ".".join(op.split(".")[:2]) if op.startswith(("LD", "ST", "IMAD", "SHFL")) else op.split(".")[0]


h = collections.Counter(key(op) for _, op, _ in loop)
print(f"{body.splitlines()[0]}\nfunction: {len(ins)} instructions; loop 0x{lo:x}..0x{hi:x}: {len(loop)} instructions")
alu = sum(v for k, v in h.items() if k in ("LOP3", "PRMT", "SHF", "IADD3", "ISETP", "VIADD", "SEL", "LEA", "IADD", "VIMNMX", "PLOP3", "P2R", "R2P", "BMSK", "SGXT", "FLO", "POPC"))
fma = sum(v for k, v in h.items() if k.startswith("IMAD") or k in ("IMUL", "FFMA"))
print(f"ALU-pipe {alu}, FMA-pipe {fma}, LDS {sum(v for k, v in h.items() if k.startswith('LDS'))}, local {sum(v for k, v in h.items() if k in ('LDL', 'STL'))}")
for k, v in h.most_common(40):
    print(f"  {k:12s} {v}")
if "--dump" in sys.argv:
    for a, op, args in loop:
        print(f"{a:06x} {op} {args}")
