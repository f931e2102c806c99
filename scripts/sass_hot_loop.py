"""SASS listing + hot-loop instruction histogram of the tile kernel (make the per-block instruction mix re-readable).
Usage: python scripts/sass_hot_loop.py [object] [kernel-substring] > profiles/<name>.txt
Finds every backward branch of the kernel, takes the loop body with the most DFMA/FFMA and prints its opcode histogram per
loop iteration and per filter block (the fp64 loop is unrolled over MPB_HOT_UNROLL_F64 * BPG blocks), then the loop's SASS."""
import collections, re, subprocess, sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
obj = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "matrix-profile-hls_b200", "csrc", "build", "scamp_tile.o")
want = sys.argv[2] if len(sys.argv) > 2 else "tile_kernelIdLb0ELb1E"
blocks_per_iter = int(sys.argv[3]) if len(sys.argv) > 3 else 2
txt = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True, check=True).stdout
funcs = re.split(r"\n\s*Function : ", txt)
body = next(f for f in funcs if f.startswith("_ZN3mpb11" + want) or want in f.split("\n", 1)[0])
name = body.split("\n", 1)[0]
ins = []   # (addr, opcode, text)
for ln in body.split("\n"):
    mm = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
    if mm:
        t = mm.group(2).strip()
        t2 = re.sub(r"^@!?U?P\d+\s+", "", t)
        ins.append((int(mm.group(1), 16), t2.split()[0], t))
loops = []
for k, (a, op, t) in enumerate(ins):
    if op.startswith("BRA"):
        mm = re.search(r"0x([0-9a-f]+)", t)
        if mm and int(mm.group(1), 16) <= a:
            tgt = int(mm.group(1), 16)
            k0 = next(i for i, x in enumerate(ins) if x[0] == tgt)
            fma = sum((2 if x[1].startswith("FFMA2") else 1) for x in ins[k0:k + 1] if x[1].startswith(("DFMA", "FFMA")))
            loops.append((fma, k0, k))
fma, k0, k1 = max(loops)
seg = ins[k0:k1 + 1]
hist = collections.Counter(re.sub(r"\..*", "", x[1]) for x in seg)
full = collections.Counter(x[1] for x in seg)
print(f"kernel: {name}")
print(f"instructions in the kernel: {len(ins)}; backward branches (loops): {len(loops)}")
print(f"hot loop: SASS 0x{seg[0][0]:x} .. 0x{seg[-1][0]:x}, {len(seg)} instructions per iteration = {blocks_per_iter} filter blocks of 8 rows x 8 diagonals per lane")
print(f"{'opcode':<12}{'per iteration':>14}{'per block':>12}")
for op, c in hist.most_common():
    print(f"{op:<12}{c:>14}{c / blocks_per_iter:>12.1f}")
d = sum(c for op, c in hist.items() if op in ("DFMA", "DMUL", "DADD"))
print(f"issue slots per block (FP64-pipe instructions cost two): {(len(seg) + d) / blocks_per_iter:.1f}; of them DFMA: {2 * hist.get('DFMA', 0) / blocks_per_iter:.0f}")
print("\nfull opcode variants:")
for op, c in sorted(full.items(), key=lambda x: -x[1]):
    print(f"  {op:<28}{c:>6}")
print("\nSASS of the loop body:")
for a, op, t in seg:
    print(f"  /*{a:05x}*/ {t}")
