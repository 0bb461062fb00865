#!/usr/bin/env python3
"""Symbolic tracer for straight-line FP dataflow in cuobjdump SASS (sm_100a).

Purpose: bit-exact parity of the per-Gaussian preprocess needs the *same FP32
rounding sequence* as the reference binary (nvcc fuses mul+add into FFMA in
NVVM and again in ptxas, and which multiply gets fused is not visible in the
C++ source).  This tool walks the fast path of one kernel's SASS, builds a
hash-consed expression DAG for every register, and prints the expression that
reaches every global store.  Run on the reference object and on our kernel it
gives two canonical forms that can be diffed offline (no GPU needed).

Usage:
  cuobjdump -sass foo.o > foo.sass
  python tools/sass_trace.py foo.sass <kernel-substring> [--force ADDR=take|skip ...]
                             [--leafmap file.json] [--canon]

Control flow: predicated EXIT is assumed not taken; a conditional BRA is taken
iff falling through would reach a CALL (slow-path subroutine of div/sqrt/rcp,
printf) before the branch target; `--force` overrides per address.
"""
import json
import re
import sys
from collections import OrderedDict


class DAG:
    def __init__(self):
        self.nodes = {}      # key -> id
        self.defs = []       # id -> key

    def mk(self, op, *args):
        key = (op,) + tuple(args)
        i = self.nodes.get(key)
        if i is None:
            i = len(self.defs)
            self.nodes[key] = i
            self.defs.append(key)
        return i


COMM = {"fmul", "fadd", "fmul.ftz", "fadd.ftz", "dmul", "dadd", "fmin", "fmax"}


def parse_sass(path, kernel_sub):
    """Return list of (addr:int, pred:str|None, opcode:str, operands:[str])."""
    lines = open(path).read().splitlines()
    out = []
    active = False
    for ln in lines:
        if "Function :" in ln:
            active = kernel_sub in ln
            continue
        if not active:
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if not m:
            continue
        addr = int(m.group(1), 16)
        body = m.group(2).strip()
        pred = None
        pm = re.match(r"@(!?U?P\d+|!?PT)\s+(.*)", body)
        if pm:
            pred, body = pm.group(1), pm.group(2)
        parts = body.split(None, 1)
        opc = parts[0]
        ops = []
        if len(parts) > 1:
            # split on commas not inside brackets
            depth = 0
            cur = ""
            for ch in parts[1]:
                if ch in "[(":
                    depth += 1
                elif ch in "])":
                    depth -= 1
                if ch == "," and depth == 0:
                    ops.append(cur.strip())
                    cur = ""
                else:
                    cur += ch
            if cur.strip():
                ops.append(cur.strip())
        out.append((addr, pred, opc, ops))
    return out


class Tracer:
    def __init__(self, instrs, force=None, verbose=False):
        self.ins = instrs
        self.by_addr = {a: i for i, (a, _, _, _) in enumerate(instrs)}
        self.d = DAG()
        self.reg = {}
        self.force = force or {}
        self.stores = []
        self.verbose = verbose
        self.nld = 0

    # ---- operand handling
    def rd(self, tok, wide=False):
        t = tok.replace(".reuse", "").strip()
        neg = False
        ab = False
        if t.startswith("-"):
            neg, t = True, t[1:]
        if t.startswith("|") and t.endswith("|"):
            ab, t = True, t[1:-1]
        if t.startswith("-"):
            neg, t = (not neg), t[1:]
        if t in ("RZ", "URZ"):
            v = self.d.mk("const", "0")
        elif re.fullmatch(r"U?R\d+", t):
            if wide:
                v = self.reg.get(t + ":w")
                if v is None:
                    lo = self.reg.get(t)
                    nxt = t[: -len(re.search(r"\d+$", t).group())] + str(int(re.search(r"\d+$", t).group()) + 1)
                    hi = self.reg.get(nxt)
                    v = self.d.mk("pair", lo if lo is not None else self.d.mk("undef", t),
                                  hi if hi is not None else self.d.mk("undef", nxt))
            else:
                v = self.reg.get(t)
                if v is None:
                    v = self.d.mk("undef", t)
        elif t.startswith("c["):
            v = self.d.mk("param", t)
        elif re.fullmatch(r"U?P\d+|PT", t.lstrip("!")):
            v = self.d.mk("pred", t)
        else:
            v = self.d.mk("const", t)
        if ab:
            v = self.d.mk("abs", v)
        if neg:
            v = self.d.mk("neg", v)
        return v

    def wr(self, tok, val, wide=False):
        t = tok.strip()
        if t in ("RZ", "URZ", "PT"):
            return
        if wide:
            self.reg[t + ":w"] = val
            pre = t[: -len(re.search(r"\d+$", t).group())]
            n = int(re.search(r"\d+$", t).group())
            self.reg[t] = self.d.mk("lo", val)
            self.reg[pre + str(n + 1)] = self.d.mk("hi", val)
        else:
            self.reg[t] = val
            self.reg.pop(t + ":w", None)
            # writing the high half of a wide pair invalidates it
            pre = t[: -len(re.search(r"\d+$", t).group())]
            n = int(re.search(r"\d+$", t).group())
            if n > 0:
                self.reg.pop(pre + str(n - 1) + ":w", None)

    def comm(self, op, a, b):
        if a > b:
            a, b = b, a
        return self.d.mk(op, a, b)

    # ---- execution
    def run(self, start=0, max_steps=20000):
        pc = start
        steps = 0
        while pc < len(self.ins) and steps < max_steps:
            steps += 1
            addr, pred, opc, ops = self.ins[pc]
            base = opc.split(".")[0]
            nxt = pc + 1
            if base == "EXIT":
                if pred is None:
                    break
                pc = nxt
                continue
            if base in ("BRA",):
                tgt_tok = ops[-1]
                tgt = int(tgt_tok, 16)
                take = None
                key = "%04x" % addr
                if key in self.force:
                    take = self.force[key] == "take"
                elif pred is None and len(ops) == 1:
                    take = True
                else:
                    # conditional: take iff a CALL lies between fallthrough and target
                    take = False
                    if tgt > addr:
                        j = nxt
                        while j < len(self.ins) and self.ins[j][0] < tgt:
                            if self.ins[j][2].startswith("CALL"):
                                take = True
                                break
                            j += 1
                if self.verbose:
                    print("  [%04x] BRA %s -> %s" % (addr, ops, "taken" if take else "fallthrough"), file=sys.stderr)
                pc = self.by_addr[tgt] if take else nxt
                continue
            if base in ("BSSY", "BSYNC", "NOP", "BAR", "WARPSYNC", "DEPBAR", "CALL", "RET", "BMOV", "ERRBAR", "MEMBAR", "CCTL"):
                pc = nxt
                continue
            self.exec_one(addr, pred, opc, ops)
            pc = nxt
        return self

    def exec_one(self, addr, pred, opc, ops):
        d = self.d
        base = opc.split(".")[0]
        ftz = ".FTZ" in opc
        sat = ".SAT" in opc

        def fin(dst, val, wide=False):
            if sat:
                val = d.mk("sat", val)
            if pred is not None and pred != "PT":
                old = self.rd(dst, wide)
                val = d.mk("select", d.mk("pred", pred), val, old)
            self.wr(dst, val, wide)

        if base == "FMUL":
            fin(ops[0], self.comm("fmul.ftz" if ftz else "fmul", self.rd(ops[1]), self.rd(ops[2])))
        elif base == "FADD":
            fin(ops[0], self.comm("fadd.ftz" if ftz else "fadd", self.rd(ops[1]), self.rd(ops[2])))
        elif base == "FFMA":
            a, b, c = self.rd(ops[1]), self.rd(ops[2]), self.rd(ops[3])
            if a > b:
                a, b = b, a
            fin(ops[0], d.mk("ffma.ftz" if ftz else "ffma", a, b, c))
        elif base == "FMNMX":
            a, b = self.rd(ops[1]), self.rd(ops[2])
            p = ops[3].strip()
            if p == "PT":
                fin(ops[0], self.comm("fmin", a, b))
            elif p == "!PT":
                fin(ops[0], self.comm("fmax", a, b))
            else:
                fin(ops[0], d.mk("fmnmx", a, b, self.rd(p)))
        elif base == "MUFU":
            fin(ops[0], d.mk(opc.lower(), self.rd(ops[1])))
        elif base in ("FCHK", "FSETP", "ISETP", "PLOP3", "DSETP", "UISETP", "UPLOP3"):
            # predicates: record symbolic
            srcs = tuple(self.rd(o) for o in ops[2:] if o)
            dst = ops[0]
            self.reg[dst] = d.mk(opc.lower(), *srcs)
            if len(ops) > 1 and re.fullmatch(r"U?P\d+", ops[1].strip()):
                self.reg[ops[1].strip()] = d.mk(opc.lower() + ".2", *srcs)
        elif base == "F2F":
            if opc == "F2F.F64.F32":
                fin(ops[0], d.mk("f2d", self.rd(ops[1])), wide=True)
            elif opc == "F2F.F32.F64":
                fin(ops[0], d.mk("d2f", self.rd(ops[1], wide=True)))
            else:
                fin(ops[0], d.mk(opc.lower(), self.rd(ops[1])))
        elif base in ("DADD", "DMUL"):
            a = self.rd(ops[1], wide=True)
            b = self.rd(ops[2], wide=True) if re.fullmatch(r"-?\|?U?R\d+.*", ops[2].strip()) else self.rd(ops[2])
            fin(ops[0], self.comm(base.lower(), a, b), wide=True)
        elif base == "DFMA":
            def w(o):
                return self.rd(o, wide=True) if re.fullmatch(r"-?\|?U?R\d+.*", o.strip()) else self.rd(o)
            a, b, c = w(ops[1]), w(ops[2]), w(ops[3])
            if a > b:
                a, b = b, a
            fin(ops[0], d.mk("dfma", a, b, c), wide=True)
        elif base == "I2F":
            if ".F64" in opc:
                fin(ops[0], d.mk("i2d", self.rd(ops[1])), wide=True)
            else:
                fin(ops[0], d.mk(opc.lower(), self.rd(ops[1])))
        elif base in ("I2FP", "F2I", "FRND", "F2FP"):
            fin(ops[0], d.mk(opc.lower(), self.rd(ops[1])))
        elif base == "MOV" or base == "UMOV":
            fin(ops[0], self.rd(ops[1]))
        elif base in ("IMAD", "UIMAD") and opc.startswith(base + ".MOV"):
            fin(ops[0], self.rd(ops[3]))
        elif base == "HFMA2" and ops[1].replace(".reuse", "") == "-RZ" and ops[2].replace(".reuse", "") == "RZ":
            fin(ops[0], d.mk("const", "h2(%s,%s)" % (ops[3], ops[4])))
        elif base in ("FSEL", "SEL", "USEL"):
            fin(ops[0], d.mk(base.lower(), self.rd(ops[1]), self.rd(ops[2]), self.rd(ops[3])))
        elif base in ("LDG", "LD", "LDS", "LDL"):
            width = 1
            if ".64" in opc:
                width = 2
            if ".128" in opc:
                width = 4
            inner = re.findall(r"\[([^\[\]]*)\]", ops[1])[-1]
            am = re.match(r"(U?R\d+)(?:\.64)?(?:\.U32)?(?:\+(U?R\d+))?(?:\+(-?0x[0-9a-f]+))?", inner)
            regname = am.group(1) if am else inner
            off = am.group(3) if (am and am.group(3)) else "0x0"
            abase = self.rd(regname, wide=True)
            dreg = ops[0].strip()
            pre = dreg[: -len(re.search(r"\d+$", dreg).group())]
            n = int(re.search(r"\d+$", dreg).group())
            for k in range(width):
                v = d.mk("ld", base.lower(), abase, "%s+%d" % (off, 4 * k))
                if pred is not None and pred != "PT":
                    v = d.mk("select", d.mk("pred", pred), v, self.rd(pre + str(n + k)))
                self.wr(pre + str(n + k), v)
        elif base in ("STG", "ST", "STS", "STL"):
            width = 1
            if ".64" in opc:
                width = 2
            if ".128" in opc:
                width = 4
            inner = re.findall(r"\[([^\[\]]*)\]", ops[0])[-1]
            am = re.match(r"(U?R\d+)(?:\.64)?(?:\+(-?0x[0-9a-f]+))?", inner)
            regname = am.group(1) if am else inner
            off = am.group(2) if (am and am.group(2)) else "0x0"
            abase = self.rd(regname, wide=True)
            sreg = ops[1].strip()
            pre = sreg[: -len(re.search(r"\d+$", sreg).group())] if sreg != "RZ" else "R"
            n = int(re.search(r"\d+$", sreg).group()) if sreg != "RZ" else None
            for k in range(width):
                v = self.rd(pre + str(n + k)) if n is not None else d.mk("const", "0")
                self.stores.append((addr, opc, abase, "%s+%d" % (off, 4 * k), v, pred))
        elif base in ("LDC", "LDCU", "ULDC"):
            width = 2 if ".64" in opc else 1
            m = re.search(r"c\[(0x[0-9a-f]+)\]\[(0x[0-9a-f]+)\]", ops[1])
            if m is None:
                v = d.mk("param", ops[1])
                self.wr(ops[0], v)
            else:
                boff = int(m.group(2), 16)
                if width == 2:
                    self.wr(ops[0], d.mk("param", "c[%s]" % hex(boff)), wide=True)
                    # individual halves keep distinct names
                    dreg = ops[0].strip()
                    pre = dreg[: -len(re.search(r"\d+$", dreg).group())]
                    n = int(re.search(r"\d+$", dreg).group())
                    self.reg[pre + str(n)] = d.mk("param", "c[%s]" % hex(boff))
                    self.reg[pre + str(n + 1)] = d.mk("param", "c[%s]" % hex(boff + 4))
                else:
                    self.wr(ops[0], d.mk("param", "c[%s]" % hex(boff)))
        elif base in ("S2R", "S2UR", "CS2R"):
            self.wr(ops[0], d.mk("sreg", ops[1]))
        else:
            # generic integer / misc op: all operands after the first are sources
            dsts = [ops[0]] if ops else []
            srcs = []
            for o in ops[1:]:
                o2 = o.strip()
                if re.fullmatch(r"!?U?P\d+|!?PT", o2):
                    srcs.append(d.mk("pred", o2))
                else:
                    srcs.append(self.rd(o2))
            wide = ".WIDE" in opc or (".64" in opc and base not in ("SHF",))
            if dsts:
                # second operand may be a predicate destination (IADD3 R, P0, ...)
                val = d.mk(opc.lower(), *srcs)
                if re.fullmatch(r"U?R\d+", dsts[0].strip()):
                    if pred is not None and pred != "PT":
                        val = d.mk("select", d.mk("pred", pred), val, self.rd(dsts[0], wide))
                    self.wr(dsts[0], val, wide)
                else:
                    self.reg[dsts[0].strip()] = val

    # ---- printing
    def find_params(self, i, acc=None, seen=None):
        if acc is None:
            acc, seen = [], set()
        if i in seen:
            return acc
        seen.add(i)
        k = self.d.defs[i]
        if k[0] == "param":
            if k[1] not in acc:
                acc.append(k[1])
        elif k[0] in ("const", "undef", "sreg", "pred"):
            pass
        else:
            for a in k[1:]:
                if isinstance(a, int):
                    self.find_params(a, acc, seen)
        return acc

    def leafname(self, k, leafmap):
        if k[0] == "ld":
            ps = [p for p in self.find_params(k[2]) if int(p[2:-1], 16) >= 0x380]
            nm = "ld(%s%s)" % ("|".join(ps) if ps else "?", "" if k[3] == "0x0+0" else "," + k[3])
            return leafmap.get(nm, nm)
        if k[0] == "param":
            return leafmap.get(k[1], k[1])
        if k[0] == "const":
            return k[1]
        if k[0] == "undef":
            return "undef:" + k[1]
        if k[0] == "sreg":
            return k[1]
        if k[0] == "pred":
            return k[1]
        return None

    FPOPS = ("fmul", "fadd", "ffma", "fmin", "fmax", "mufu", "f2d", "d2f", "dadd", "dmul", "dfma",
             "neg", "abs", "i2d", "i2fp", "f2i", "sat", "fsel", "frnd", "select", "lo", "hi", "pair",
             "fmnmx", "i2f")

    def dump(self, roots, leafmap=None, out=sys.stdout):
        """Print expressions for roots with shared temporaries."""
        leafmap = leafmap or {}
        # count uses among fp nodes
        uses = {}
        order = []
        seen = set()

        def isfp(k):
            return any(k[0].startswith(p) for p in self.FPOPS)

        def visit(i):
            uses[i] = uses.get(i, 0) + 1
            if i in seen:
                return
            seen.add(i)
            k = self.d.defs[i]
            if self.leafname(k, leafmap) is not None:
                return
            if not isfp(k):
                return
            for a in k[1:]:
                if isinstance(a, int):
                    visit(a)
            order.append(i)

        for r in roots:
            visit(r)
        names = {}

        def expr(i, top=False):
            k = self.d.defs[i]
            ln = self.leafname(k, leafmap)
            if ln is not None:
                return ln
            if not isfp(k):
                ps = self.find_params(i)
                return "int{%s:%s}" % (k[0], "|".join(ps))
            if not top and i in names:
                return names[i]
            if k[0] == "neg":
                return "-" + expr(k[1])
            if k[0] == "abs":
                return "|" + expr(k[1]) + "|"
            args = ", ".join(expr(a) if isinstance(a, int) else str(a) for a in k[1:])
            return "%s(%s)" % (k[0], args)

        tcount = 0
        for i in order:
            k = self.d.defs[i]
            if k[0] in ("neg", "abs", "lo", "hi"):
                continue
            if uses.get(i, 0) > 1 or True:
                nm = "t%d" % tcount
                tcount += 1
                print("%s = %s" % (nm, expr(i, top=True)), file=out)
                names[i] = nm
        return names, expr


def main():
    args = sys.argv[1:]
    if len(args) < 2:
        print(__doc__)
        sys.exit(1)
    path, ksub = args[0], args[1]
    force = {}
    leafmap = {}
    verbose = False
    i = 2
    while i < len(args):
        if args[i] == "--force":
            a, v = args[i + 1].split("=")
            force[a.lower().rjust(4, "0")] = v
            i += 2
        elif args[i] == "--leafmap":
            leafmap = json.load(open(args[i + 1]))
            i += 2
        elif args[i] == "-v":
            verbose = True
            i += 1
        else:
            i += 1
    ins = parse_sass(path, ksub)
    if not ins:
        print("kernel not found", file=sys.stderr)
        sys.exit(2)
    tr = Tracer(ins, force, verbose).run()
    roots = [s[4] for s in tr.stores]
    names, expr = tr.dump(roots, leafmap)
    print("---- stores")
    for addr, opc, abase, off, v, pred in tr.stores:
        ps = tr.find_params(abase)
        print("[%04x] %s%s -> *(%s %s) = %s" % (addr, ("@%s " % pred) if pred else "", opc, "|".join(ps), off, expr(v)))


if __name__ == "__main__":
    main()
