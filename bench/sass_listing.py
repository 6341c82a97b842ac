#!/usr/bin/env python
"""Per-kernel SASS evidence from the built library (no GPU needed): for every kernel of
``libdtf_kernels.so`` one file ``profiles/sass/<short name>.sass`` with its resource usage
(registers, shared memory, spill: ``cuobjdump -res-usage``), an opcode histogram, and the instruction
windows around what the design claims -- tcgen05.mma (``UTCHMMA``), TMA loads / stores / reduce-add
(``UTMALDG`` / ``UTMASTG`` / ``UTMAREDG``), TMEM loads (``LDTM``), mbarrier waits (``SYNCS``), the
gather's 128-bit peer loads, ``multimem`` (``LDGMC`` / ``.SYS`` stores), barrier polls.  The MMA issue
loop, the TMA producer loop and the gather loop of every instantiation are therefore in the listing
in full; ``profiles/sass/INDEX.md`` maps the step's launch list to the files.

    python bench/sass_listing.py            # rewrites profiles/sass/
"""
import collections
import os
import re
import subprocess
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(REPO, "distributed_tensorflow_b200", "_lib", "libdtf_kernels.so")
OUT = os.path.join(REPO, "profiles", "sass")
KEY = re.compile(r"UTCHMMA|UTMALDG|UTMASTG|UTMAREDG|UTCBAR|UTCATOMSWS|LDTM|LDGMC|LDG\.E\.128|"
                 r"STG\.E\.(?:64|128)\.STRONG\.SYS|LD\.E\.STRONG\.SYS|LDG\.E\.STRONG\.SYS|ST\.E\.STRONG\.SYS|REDG|RED\.E|UCGABAR|ATOM")
CTX = 3


def demangle(names):
    p = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True)
    return p.stdout.splitlines() if p.returncode == 0 else names


def short(name):
    m = re.search(r"gemm_sm100_kernel<(\d+), *\(?(?:bool\))?(\w+), *\(?(?:bool\))?(\w+), *(\d+), *\(?(?:bool\))?(\w+), *(\d+)>", name)
    if m:
        b = lambda x: "1" if x in ("true", "1") else "0"
        return "gemm_bn%s_amn%s_bmn%s_epi%s_gather%s_ctas%s" % (m.group(1), b(m.group(2)), b(m.group(3)), m.group(4), b(m.group(5)), m.group(6))
    base = re.sub(r"\(.*", "", name)
    base = re.sub(r"^void\s+", "", base)
    return re.sub(r"[^A-Za-z0-9_]+", "_", base.replace("dtf::", "")).strip("_")[:90]


def main():
    os.makedirs(OUT, exist_ok=True)
    for f in os.listdir(OUT):
        os.unlink(os.path.join(OUT, f))
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout.splitlines()
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout.splitlines()
    usage = {}
    for i, l in enumerate(res):
        m = re.match(r"\s*Function (\S+):", l)
        if m and i + 1 < len(res):
            usage[m.group(1)] = res[i + 1].strip()
    funcs, cur = collections.OrderedDict(), None
    for l in sass:
        m = re.match(r"\s*Function : (\S+)", l)
        if m:
            cur = m.group(1)
            funcs[cur] = []
        elif cur is not None:
            funcs[cur].append(l)
    mangled = list(funcs)
    names = demangle(mangled)
    index = []
    seen = collections.Counter()
    for mg, nm in zip(mangled, names):
        body = [l for l in funcs[mg] if re.search(r"/\*[0-9a-f]{4}\*/", l)]
        ops = collections.Counter()
        for l in body:
            m = re.search(r"/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", l)
            if m:
                ops[m.group(1).split(".")[0] if not m.group(1).startswith(("UTC", "UTMA", "LDTM", "LDGMC", "SYNCS")) else m.group(1)] += 1
        hits = [i for i, l in enumerate(body) if KEY.search(l)]
        keep = set()
        for i in hits:
            keep.update(range(max(0, i - CTX), min(len(body), i + CTX + 1)))
        sn = short(nm)
        seen[sn] += 1
        if seen[sn] > 1:
            sn += "_%d" % seen[sn]
        with open(os.path.join(OUT, sn + ".sass"), "w") as f:
            f.write("// %s\n// %s\n// %d SASS instructions; resource usage: %s\n" % (nm, mg, len(body), usage.get(mg, "?")))
            f.write("// opcode histogram: %s\n\n" % ", ".join("%s x%d" % kv for kv in ops.most_common(40)))
            last = -2
            for i in sorted(keep):
                if i != last + 1:
                    f.write("        ...\n")
                f.write(body[i].rstrip() + "\n")
                last = i
        tc = sum(v for k, v in ops.items() if k.startswith("UTCHMMA"))
        tma = sum(v for k, v in ops.items() if k.startswith("UTMA"))
        index.append((sn, len(body), usage.get(mg, "?"), tc, tma, sum(v for k, v in ops.items() if k.startswith("LDTM")),
                      ops.get("HMMA", 0)))
    with open(os.path.join(OUT, "INDEX.md"), "w") as f:
        f.write("# SASS listings per kernel (`python bench/sass_listing.py`, from `cuobjdump -sass` / `-res-usage` of libdtf_kernels.so)\n\n")
        f.write("gemm names: bn = N tile, amn/bmn = operand is MN-major, epi = epilogue (2 bias+ReLU+mask, 3 dReLU, 5 fp32 TMA reduce-add, "
                "6 push rows), gather = fused pull, ctas 2 = CTA pairs (`cta_group::2`).\n\n")
        f.write("| file | SASS instr. | resources | UTCHMMA | UTMA* | LDTM | HMMA |\n|---|---|---|---|---|---|---|\n")
        for row in index:
            f.write("| `%s.sass` | %d | %s | %d | %d | %d | %d |\n" % row)
    print("wrote %d listings to %s" % (len(index), OUT))


if __name__ == "__main__":
    sys.exit(main())
