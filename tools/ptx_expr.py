#!/usr/bin/env python3
"""Symbolic evaluator for the float dataflow of one PTX kernel: prints, for every st.global.f32 (or a selection),
the expression tree of the stored value in terms of loads/params, with mul/add/fma kept exactly as emitted.
Used to compare FMA contraction between the reference build and this repo's kernels (bit-parity work).
usage: ptx_expr.py file.ptx kernel_substring [max_depth]"""
import re
import sys


def main():
    path, kname = sys.argv[1], sys.argv[2]
    text = open(path).read()
    m = re.search(r"\.entry\s+(\S*%s\S*)\(" % re.escape(kname), text)
    start = m.start()
    end = text.find("\n}", start)
    body = text[start:end]
    env = {}
    stores = []
    n_ld = [0]

    def val(tok):
        tok = tok.strip()
        if tok in env:
            return env[tok]
        if tok.startswith("0f"):
            import struct
            return repr(struct.unpack(">f", bytes.fromhex(tok[2:]))[0])
        return tok

    for line in body.split("\n"):
        line = line.strip().rstrip(";")
        if not line or line.startswith("//") or line.startswith("."):
            continue
        pm = re.match(r"(@!?%p\d+\s+)?(\S+)\s+(.*)", line)
        if not pm:
            continue
        pred, op, rest = pm.groups()
        args = [a.strip() for a in re.split(r",\s*(?![^\[]*\])", rest)]
        p = ("[%s] " % pred.strip()) if pred else ""
        if op.startswith("ld.param"):
            env[args[0]] = "P(" + args[1].strip("[]").split("_param_")[-1] + ")"
        elif op.startswith("ld.global") or op.startswith("ld.const") or op.startswith("ld.shared"):
            a = args[1].strip("[]")
            base, _, off = a.partition("+")
            env[args[0]] = "LD(%s+%s)" % (val(base), off or "0")
        elif op.startswith("st.global") or op.startswith("st.shared"):
            a = args[0].strip("[]")
            base, _, off = a.partition("+")
            stores.append((op, "%s+%s" % (val(base), off or "0"), p + val(args[1])))
        elif op in ("mul.f32", "mul.rn.f32", "mul.ftz.f32"):
            env[args[0]] = p + "(%s * %s)" % (val(args[1]), val(args[2]))
        elif op in ("add.f32", "add.rn.f32"):
            env[args[0]] = p + "(%s + %s)" % (val(args[1]), val(args[2]))
        elif op in ("sub.f32", "sub.rn.f32"):
            env[args[0]] = p + "(%s - %s)" % (val(args[1]), val(args[2]))
        elif op in ("fma.rn.f32",):
            env[args[0]] = p + "fma(%s, %s, %s)" % (val(args[1]), val(args[2]), val(args[3]))
        elif op == "neg.f32":
            env[args[0]] = p + "(-%s)" % val(args[1])
        elif op in ("mov.f32", "mov.b32", "mov.u32", "mov.u64", "mov.b64", "cvta.to.global.u64"):
            env[args[0]] = p + val(args[1]) if p else val(args[1])
        elif op.startswith("mul.wide") or op.startswith("mad.wide") or op.startswith("add.s64") or op.startswith("add.s32") or \
                op.startswith("mad.lo") or op.startswith("mul.lo") or op.startswith("shl") or op.startswith("cvt"):
            env[args[0]] = "%s(%s)" % (op.split(".")[0], ",".join(val(a) for a in args[1:]))
        elif len(args) >= 2 and args[0].startswith("%"):
            env[args[0]] = p + "%s(%s)" % (op, ",".join(val(a) for a in args[1:]))
    sel = sys.argv[3] if len(sys.argv) > 3 else None
    for op, addr, v in stores:
        pm = re.search(r"P\((\d+)\)", addr)
        tag = "P%s" % (pm.group(1) if pm else "?")
        if sel and tag != sel:
            continue
        print(op, "@", tag, addr[-40:])
        print_dag(v)


def tokenize(v):
    return re.findall(r"fma\(|LD\(|P\(|[A-Za-z_.]+[\w.]*\(|\(|\)|,|\*|\+|-|[^(),*+\- ]+", v)


def print_dag(v):
    """Hash-cons the fully expanded string into named temporaries."""
    names = {}
    out = []

    def intern(expr):
        if expr not in names:
            names[expr] = "t%d" % len(names)
            out.append("      %s = %s" % (names[expr], expr))
        return names[expr]

    # parse recursively
    pos = [0]
    toks = tokenize(v)

    def parse():
        t = toks[pos[0]]
        if t in ("fma(",) or (t.endswith("(") and t not in ("(",)):
            fn = t[:-1]
            pos[0] += 1
            args = []
            depth_guard = 0
            while toks[pos[0]] != ")":
                if toks[pos[0]] == ",":
                    pos[0] += 1
                    continue
                args.append(parse_sum())
                depth_guard += 1
            pos[0] += 1
            if fn in ("LD", "P"):
                import hashlib
                raw = ",".join(args)
                return "%s<%s>" % (fn, hashlib.md5(raw.encode()).hexdigest()[:6] if fn == "LD" else raw)
            return intern("%s(%s)" % (fn, ", ".join(args)))
        if t == "(":
            pos[0] += 1
            e = parse_sum()
            assert toks[pos[0]] == ")", toks[pos[0]]
            pos[0] += 1
            return e
        if t == "-":
            pos[0] += 1
            return intern("neg(%s)" % parse())
        pos[0] += 1
        return t

    def parse_sum():
        a = parse()
        while pos[0] < len(toks) and toks[pos[0]] in ("*", "+", "-"):
            o = toks[pos[0]]
            pos[0] += 1
            b = parse()
            a = intern("%s %s %s" % (a, o, b))
        return a

    r = parse_sum()
    for l in out:
        print(l)
    print("      => ", r)


if __name__ == "__main__":
    main()
