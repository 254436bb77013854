"""Writes one trimmed SASS listing per kernel of sputnik_b200/libbsmm.so under profiles/sass/
(cuobjdump -sass): every tensor-core / TMA / TMEM / mbarrier instruction with two lines of context,
gaps marked, plus an index with the mnemonic counts per kernel.  Run in the build container:

    python tools/sass_dump.py            (after __graft_entry__.build())
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "sputnik_b200", "libbsmm.so")
OUT = os.path.join(ROOT, "profiles", "sass")
KEY = re.compile(r"\b(UTC[A-Z]*MMA|UTMALDG|UTMASTG|UTMAREDG|UTMAPF|UTMACCTL|UTMACMDFLUSH|UBLKCP|LDTM|STTM|"
                 r"UTCBAR|UTCATOMSWS|SYNCS|HMMA|ACQBULK|UGETNEXTWORKID|ATOMG|ATOM|RED|STG|LDG)\b")
COUNTED = ["UTCHMMA", "UTCQMMA", "UTMALDG", "UTMASTG", "UTMAREDG", "UTMAPF", "LDTM", "STTM",
           "SYNCS", "HMMA", "STG", "LDG"]


def short(name):
    """bsmm::bsmm_gemm_kernel<1, false, true, 256, false, 1>(...) -> bsmm_gemm_kernel_1_0_1_256_0_1"""
    m = re.match(r"(?:void )?(?:bsmm::)?(\w+)(?:<(.*?)>)?\(", name)
    if not m:
        return re.sub(r"\W+", "_", name)[:80]
    args = m.group(2) or ""
    args = args.replace("true", "1").replace("false", "0").replace("(int)", "").replace("(bool)", "")
    args = re.sub(r"[^0-9,]", "", args).replace(",", "_")
    return m.group(1) + ("_" + args if args else "")


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
    regs = dict(re.findall(r"Function (\S+):\s*\n\s*REG:(\d+)", res))
    os.makedirs(OUT, exist_ok=True)
    for f in os.listdir(OUT):
        os.remove(os.path.join(OUT, f))
    funcs = re.split(r"\n\s*Function : ", sass)[1:]
    index = []
    for blk in funcs:
        mangled, _, body = blk.partition("\n")
        mangled = mangled.strip()
        dem = subprocess.run(["c++filt", mangled], capture_output=True, text=True).stdout.strip()
        lines = [l for l in body.splitlines() if re.search(r"/\*[0-9a-f]{4,}\*/", l)]
        ins = [re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", l).rstrip() for l in lines
               if not re.match(r"\s*/\* 0x", l)]
        keep = set()
        counts = collections.Counter()
        for i, l in enumerate(ins):
            m = KEY.search(l)
            if m:
                op = m.group(1)
                for c in COUNTED:
                    if op.startswith(c):
                        counts[c] += 1
                if op in ("STG", "LDG", "ATOM", "RED", "ATOMG") and not dem.startswith("bsmm::bsmm_"):
                    pass
                if op not in ("STG", "LDG") or "bsmm_" not in dem or True:
                    if op in ("STG", "LDG", "SYNCS") and "kernel" in dem and "bsmm_" in dem and \
                            not any(k in l for k in ("UTC", "UTMA")):
                        if op != "SYNCS":
                            continue   # global ld/st are counted, not listed, for the GEMM kernels
                    keep.update(range(max(0, i - 2), min(len(ins), i + 3)))
        name = short(dem)
        path = os.path.join(OUT, name + ".sass")
        with open(path, "w") as f:
            f.write(f"// {dem}\n// {mangled}\n// registers/thread: {regs.get(mangled, '?')}, "
                    f"{len(ins)} SASS instructions; listed: tensor-core (UTC*MMA), TMA (UTMA*), TMEM "
                    f"(LDTM/STTM) and mbarrier (SYNCS) instructions with 2 lines of context\n")
            f.write("// counts: " + ", ".join(f"{k} {counts[k]}" for k in COUNTED if counts[k]) + "\n")
            last = -2
            for i in sorted(keep):
                if i != last + 1:
                    f.write("        ...\n")
                f.write(ins[i] + "\n")
                last = i
        index.append((name, regs.get(mangled, "?"), len(ins), counts))
    with open(os.path.join(OUT, "INDEX.txt"), "w") as f:
        f.write("# cuobjdump -sass sputnik_b200/libbsmm.so, per kernel: registers, instructions, mnemonic counts\n")
        f.write(f"# {'kernel':58s} regs  instr  " + " ".join(f"{c:>8s}" for c in COUNTED) + "\n")
        for name, r, n, c in sorted(index):
            f.write(f"{name:60s} {r:>4s} {n:6d}  " + " ".join(f"{c[k]:8d}" for k in COUNTED) + "\n")
    print(f"wrote {len(index)} listings to {OUT}")


if __name__ == "__main__":
    main()
