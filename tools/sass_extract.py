# dev: SASS evidence without an .ncu-rep.  Dumps one kernel of lib/libvcrt.so with cuobjdump and prints (a) the opcode histogram of the
# whole kernel and (b) its hot loop body: the longest backward-branch loop in which `marker` opcodes make up the largest share.
# usage: sass_extract.py <kernel name substring> <marker opcode, e.g. FFMA2> [dense]
import re, subprocess, sys
name, marker = sys.argv[1], sys.argv[2]
lib = "vulkancomputeraytracing_b200/lib/libvcrt.so"
dense = len(sys.argv) > 3 and sys.argv[3] == "dense"   # pick the loop with the highest SHARE of marker opcodes (>= 4 of them) instead of the highest count
text = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s*Function : ", text)
body = next(f for f in funcs if name in f.split("\n", 1)[0])
print("Function :", body.split("\n", 1)[0])
ins = []
for line in body.split("\n"):
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if m:
        ins.append((int(m.group(1), 16), m.group(2).strip()))
def opcode(t):
    w = t.split()
    return (w[1] if w[0].startswith("@") else w[0]).split(".")[0]
hist = {}
for _, t in ins:
    hist[opcode(t)] = hist.get(opcode(t), 0) + 1
print(f"{len(ins)} instructions; opcodes: " + " ".join(f"{k}x{v}" for k, v in sorted(hist.items(), key=lambda kv: -kv[1])[:24]))
addr = {a: i for i, (a, _) in enumerate(ins)}
best = None
for i, (a, t) in enumerate(ins):
    m = re.search(r"BRA\S*\s+(?:\S+,\s*)?`?\(?0x([0-9a-f]+)", t)
    if m and int(m.group(1), 16) in addr and addr[int(m.group(1), 16)] < i:
        j = addr[int(m.group(1), 16)]
        n = sum(1 for _, x in ins[j:i + 1] if opcode(x) == marker)
        score = n / (i + 1 - j) if dense else n
        if n >= (4 if dense else 1) and (best is None or score > best[3]):
            best = (n, j, i, score)
if best:
    n, j, i, _ = best
    loop = ins[j:i + 1]
    h = {}
    for _, t in loop:
        h[opcode(t)] = h.get(opcode(t), 0) + 1
    print(f"\nhot loop: {len(loop)} instructions at 0x{loop[0][0]:04x}-0x{loop[-1][0]:04x}, {n} {marker}; opcodes: " + " ".join(f"{k}x{v}" for k, v in sorted(h.items(), key=lambda kv: -kv[1])))
    for a, t in loop:
        print(f"  /*{a:04x}*/ {t}")
