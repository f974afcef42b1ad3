"""Opcode histogram per kernel of libssl4gie_b200.so (cuobjdump -sass): the evidence of which kernels are Blackwell-native
(UTCHMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UTMALDG / UTMASTG = TMA, UTCBAR = tcgen05.commit) and which still use the
legacy tensor path (HMMA = mma.sync).  Usage: python scripts/sass_summary.py > profiles/rNN_sass_summary.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "ssl4gie_b200", "libssl4gie_b200.so")
KEYS = ["UTCHMMA.2CTA", "UTCHMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTCBAR", "UTCCP", "HMMA", "MUFU.EX2", "SYNCS", "BAR", "LDGSTS", "FLOAT_ATOMIC", "STL", "LDL"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True, check=True).stdout
    kernels = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip() or m.group(1)
            name = re.sub(r"^void ", "", name)
            name = name.replace("(anonymous namespace)::", "")
            name = re.sub(r"\((?!anonymous).*", "", name)
            cur = kernels.setdefault(name, collections.Counter())
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and cur is not None:
            op = m.group(1)
            cur["total"] += 1
            if re.match(r"(ATOM|ATOMG|ATOMS|RED|REDG)\..*\.(F32|F64|F16|BF16)", op) and ".ADD" in op:
                cur["FLOAT_ATOMIC"] += 1   # floating-point atomic add: the only non-reproducible instruction class
                continue
            for k in KEYS:
                if op.startswith(k):
                    cur[k] += 1
                    break
    print("# cuobjdump -sass ssl4gie_b200/libssl4gie_b200.so (sm_100a): instruction counts per kernel (static; zero columns omitted)")
    print("# tcgen05.mma -> UTCHMMA(.2CTA), tcgen05.ld/st -> LDTM/STTM, TMA -> UTMALDG/UTMASTG, tcgen05.commit -> UTCBAR, mma.sync -> HMMA,")
    print("# FLOAT_ATOMIC = ATOM* / RED* floating-point adds (only the fp32 SIMT validation wgrad has them), register spills as STL / LDL")
    for name, c in kernels.items():
        items = " ".join(f"{k}={c[k]}" for k in KEYS if c[k])
        print(f"{name[:110]:110s} total={c['total']:5d} {items}")
    native = [n for n, c in kernels.items() if c["UTCHMMA"] or c["UTCHMMA.2CTA"]]
    legacy = [n for n, c in kernels.items() if c["HMMA"]]
    print("#\n# tcgen05 kernels: " + "; ".join(native))
    print("# mma.sync kernels (kept for A/B via S4G_ATTN=mma, not on the default path): " + "; ".join(legacy))
    atom = [n for n, c in kernels.items() if c["FLOAT_ATOMIC"]]
    print("# kernels with floating-point atomic adds: " + ("; ".join(atom) if atom else "none") +
          "  (the fp32 / SIMT validation back end's split-K weight gradient; no kernel of the bf16 tensor-core product path)")


if __name__ == "__main__":
    main()
