# dev: basic-block table of an `ncu --set full --import-source on` capture: share of executed instructions, active lanes and
# stall samples per run of SASS instructions that execute equally often.  usage: ncu_blocks.py rep.ncu-rep [min_share_percent]
import csv, io, subprocess, sys
rep = sys.argv[1]; floor = float(sys.argv[2]) if len(sys.argv) > 2 else 0.3
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
allrows = list(csv.reader(io.StringIO(raw)))
# one section per profiled launch: a "Kernel Name" row, a header row, then one row per SASS instruction
starts = [i for i, r in enumerate(allrows) if r and r[0] == "Kernel Name"] + [len(allrows)]
pick = int(sys.argv[3]) if len(sys.argv) > 3 else None     # optional third argument: which launch of the capture (0-based); default all
for si in range(len(starts) - 1):
    if pick is not None and si != pick: continue
    rows = allrows[starts[si]:starts[si + 1]]
    print(f"---- launch {si}: {rows[0][1]}")
    h = rows[1]; data = [r for r in rows[2:] if len(r) == len(h)]
    iS, iN, iE, iT = h.index("Source"), h.index("# Samples"), h.index("Instructions Executed"), h.index("Thread Instructions Executed")
    tot = sum(int(r[iE]) for r in data); smp = sum(int(r[iN]) for r in data)
    print(f"{len(data)} SASS instructions, {tot} warp instructions executed, {smp} stall samples")
    print("first-last  count  share of executed instructions  active lanes  share of samples  opcodes")
    blocks = []
    for k, r in enumerate(data):
        e, t = int(r[iE]), int(r[iT])
        if blocks and blocks[-1][-1][1] == e and abs(blocks[-1][-1][2] - t / max(e, 1)) < 0.05: blocks[-1].append((k, e, t / max(e, 1), int(r[iN]), r[iS]))
        else: blocks.append([(k, e, t / max(e, 1), int(r[iN]), r[iS])])
    for b in blocks:
        e = sum(x[1] for x in b) / tot * 100; s = sum(x[3] for x in b) / smp * 100
        if e < floor and s < floor: continue
        ops = {}
        for x in b:
            w = x[4].split(); op = (w[1] if w[0].startswith("@") else w[0]).split(".")[0]; ops[op] = ops.get(op, 0) + 1
        print(f"{b[0][0]:5d}-{b[-1][0]:5d} {len(b):4d}  {e:6.2f} %  {b[0][2]:5.1f}  {s:6.2f} %  " + " ".join(f"{k}x{v}" for k, v in sorted(ops.items(), key=lambda kv: -kv[1])[:6]))
