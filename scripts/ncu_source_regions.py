"""Where the warp samples of an `ncu --set full --import-source on` capture fall: reads the SASS source page exported with
`ncu -i X.ncu-rep --page source --csv` and sums samples / executed instructions over code regions of the tile kernel:
the main body up to the hot loop (prologue, re-seed, tile staging), the hot loop (densest FMA loop), the rest of the main
body, and every out-of-line function after the kernel's EXIT (drain_hits, flush_queue / replay_block), classified by content.
usage: python scripts/ncu_source_regions.py gpurun_out/<tag>_ncu_source.csv"""
import csv, io, re, sys
txt = open(sys.argv[1]).read().split("\n")
h = next(i for i, l in enumerate(txt) if l.startswith('"Address"'))
rows = list(csv.DictReader(io.StringIO("\n".join(txt[h:]))))
base = int(rows[0]["Address"], 16)
def op(r): return re.sub(r"^@!?U?P\d+\s+", "", r["Source"].strip()).split()[0]
S = [int(r["# Samples"] or 0) for r in rows]; E = [int(r["Instructions Executed"] or 0) for r in rows]
tot, inst = sum(S), sum(E)
addr = [int(r["Address"], 16) - base for r in rows]
# loops: backward branches
loops = []
for k, r in enumerate(rows):
    if op(r).startswith("BRA"):
        mm = re.search(r"0x([0-9a-f]+)", r["Source"].split("BRA")[1])
        if mm:
            t = int(mm.group(1), 16) - base
            if 0 <= t <= addr[k]:
                k0 = addr.index(t) if t in addr else None
                if k0 is not None:
                    loops.append((sum(1 for x in rows[k0:k + 1] if op(x).startswith(("DFMA", "FFMA"))), k0, k))
_, h0, h1 = max(loops)
ends = [k for k, r in enumerate(rows) if re.match(r"(EXIT|RET)", r["Source"].strip())]   # unpredicated only
main_end = ends[0]
regions = [("main: prologue + re-seed + tile staging", 0, h0), ("main: hot loop", h0, h1 + 1), ("main: after the hot loop (push / tile end / epilogue)", h1 + 1, main_end + 1)]
prev = main_end + 1
for e in ends[1:] + [len(rows) - 1]:
    if e + 1 <= prev: continue
    seg = rows[prev:e + 1]
    ops = [op(x) for x in seg]
    name = "function: drain_hits + apply_winners (ATOM/CAS)" if any(o.startswith(("ATOM", "ATOMG")) and "CAS" in x["Source"] for o, x in zip(ops, seg)) else \
           ("function: replay (flush_queue / replay_block)" if sum(o.startswith(("DMUL", "FMUL")) for o in ops) > 32 else "function: other")
    regions.append((name, prev, e + 1)); prev = e + 1
print(f"{'region':<58}{'SASS range':>20}{'samples %':>11}{'instr %':>9}")
for name, a, b in regions:
    print(f"{name:<58}{hex(addr[a]):>10}-{hex(addr[b - 1]):<9}{100 * sum(S[a:b]) / tot:>10.1f}{100 * sum(E[a:b]) / inst:>9.1f}")
print(f"total samples {tot}, warp instructions executed {inst:.4g}")
# stall reasons of the whole kernel
keys = [k for k in rows[0].keys() if k.startswith("stall_") and "Not Issued" not in k]
st = {k: sum(int(r[k] or 0) for r in rows) for k in keys}
print("stall reasons (all samples): " + ", ".join(f"{k[6:]} {100 * v / tot:.1f}%" for k, v in sorted(st.items(), key=lambda x: -x[1])[:8]))
