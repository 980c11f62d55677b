"""Opcode mix of one kernel from `ncu -i rep --page source --csv --print-source cuda,sass -k regex:NAME` (stdin or file):
warp-level instructions executed per SASS opcode, sorted, with shares."""
import csv, sys, collections
def main(path):
    rows = csv.reader(open(path, newline=""))
    seen = set(); ops = collections.Counter()
    for r in rows:
        if len(r) < 8 or r[0] != "" or not r[2].startswith("0x"):
            continue
        if r[2] in seen:
            continue
        seen.add(r[2])
        try:
            n = int(r[7])
        except ValueError:
            continue
        toks = r[3].split()
        op = toks[1] if toks and toks[0].startswith("@") else (toks[0] if toks else "?")
        ops[op.split(".")[0] + ("." + op.split(".")[1] if op.startswith(("SHFL", "LDS", "STS", "ATOM", "RED", "MUFU", "LDG", "STG")) and "." in op else "")] += n
    tot = sum(ops.values())
    print("total warp instructions %d" % tot)
    for op, n in ops.most_common(40):
        print("%-14s %12d  %5.1f%%" % (op, n, 100.0 * n / tot))
if __name__ == "__main__":
    main(sys.argv[1])
