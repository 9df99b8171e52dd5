"""Writes profiles/r02_sass_summary.txt: per kernel of lib/libggml_b200.so the SASS instruction counts that prove which
hardware paths it uses (cuobjdump -sass / -res-usage; no GPU needed).  usage: python tools/sass_summary.py [out.txt]"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "llama.cpp-q_4_0_b200", "lib", "libggml_b200.so")
COLS = ["UTCIMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UBLKCP", "UTMAPF", "IDP.4A", "FFMA2", "SYNCS", "LDGSTS", "ELECT", "USETMAXREG", "STL", "LDL"]


def main():
    out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "r02_sass_summary.txt")
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True, check=True).stdout
    usage = {}
    name = None
    for ln in res.splitlines():
        m = re.match(r"\s*Function (\S+):", ln)
        if m:
            name = m.group(1)
            continue
        m = re.search(r"REG:(\d+) STACK:(\d+)", ln)
        if m and name:
            usage[name] = (int(m.group(1)), int(m.group(2)))
            name = None
    kernels, cur = {}, None
    for ln in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", ln)
        if m:
            cur = m.group(1)
            kernels[cur] = {"instrs": 0, **{c: 0 for c in COLS}}
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
        if m and cur:
            op = m.group(1)
            kernels[cur]["instrs"] += 1
            for c in COLS:
                if op == c or op.startswith(c + "."):
                    kernels[cur][c] += 1
    names = list(kernels)
    dem = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    with open(out, "w") as f:
        f.write("SASS instruction counts per kernel of libggml_b200.so (cuobjdump -sass, sm_100a), round 2; written by tools/sass_summary.py.\n")
        f.write("UTCIMMA = tcgen05.mma, UTCBAR = tcgen05.commit, LDTM/STTM = tcgen05.ld/st, UTMALDG = TMA tensor load, UBLKCP = cp.async.bulk,\n")
        f.write("IDP.4A = dp4a, FFMA2 = fma.rn.f32x2, SYNCS = mbarrier ops, STL/LDL = local-memory spills.  regs/stack from -res-usage.\n\n")
        f.write(f"{'instrs':8s} {'regs':6s} {'stack':6s} " + " ".join(f"{c:8s}" for c in COLS) + " kernel\n")
        for n, d in zip(names, dem):
            k = kernels[n]
            r, s = usage.get(n, (-1, -1))
            short = re.sub(r"\(.*$", "", d)
            f.write(f"{k['instrs']:<8d} {r:<6d} {s:<6d} " + " ".join(f"{k[c]:<8d}" for c in COLS) + f" {short}\n")
    print(out, len(names), "kernels")


if __name__ == "__main__":
    main()
