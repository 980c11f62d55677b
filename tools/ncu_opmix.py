"""Opcode mix of one kernel from `ncu -i rep --page source --csv --print-source sass [-k regex:NAME]` (file argument):
warp-level instructions executed and stall samples per SASS opcode, sorted, with shares.  `--list` prints every instruction."""
import collections
import csv
import sys


def main(path, listing=False):
    rows = list(csv.reader(open(path, newline="")))
    hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    h = rows[hdr]
    ci, cs = h.index("Instructions Executed"), h.index("# Samples")
    ops, samples = collections.Counter(), collections.Counter()
    for r in rows[hdr + 1:]:
        if len(r) <= ci or not r[0].startswith("0x"):
            continue
        n, smp = int(r[ci]), int(r[cs])
        toks = r[1].split()
        op = toks[1] if toks and toks[0].startswith("@") else (toks[0] if toks else "?")
        parts = op.rstrip(";").split(".")
        key = parts[0] + ("." + parts[1] if parts[0] in ("SHFL", "LDS", "STS", "ATOMS", "ATOM", "RED", "MUFU", "LDG", "STG") and len(parts) > 1 else "")
        ops[key] += n
        samples[key] += smp
        if listing:
            print("%10d %6d  %s" % (n, smp, r[1].strip()))
    tot, stot = sum(ops.values()), max(sum(samples.values()), 1)
    print("total warp instructions %d, stall samples %d" % (tot, stot))
    for op, n in ops.most_common(45):
        print("%-14s %12d  %5.1f%%   samples %5.1f%%" % (op, n, 100.0 * n / tot, 100.0 * samples[op] / stot))


if __name__ == "__main__":
    main(sys.argv[1], "--list" in sys.argv)
