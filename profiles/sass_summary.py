"""Writes profiles/sass_summary.txt: per kernel of libtmdn.so, the SASS instruction count and the mnemonics that show which
hardware paths it uses (tcgen05, TMA, mbarriers, cp.async, clusters, atomics).  `python profiles/sass_summary.py`, no GPU needed."""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "torchmd-net_b200", "torchmdnet", "lib", "libtmdn.so")
KEYS = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTCBAR", "UTCCP", "UTMALDG", "UTMASTG", "UTMAREDG", "SYNCS", "LDGSTS", "UCGABAR",
        "CCTL", "ATOMG", "ATOMS", "REDG", "RED.", "BAR.SYNC", "LDG.E.128", "LDS.128", "SHFL", "MUFU", "HMMA", "FFMA"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    arch = sorted(set(re.findall(r"arch = (sm_\w+)", sass)))
    kernels, name = collections.OrderedDict(), None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = m.group(1)
            while name in kernels:          # the same kernel compiled into two objects: keep both listings
                name += "'"
            kernels[name] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(.*?);", line)
        if m and name:
            ins = re.sub(r"^@!?U?P(\d+|T)\s+", "", m.group(1))
            kernels[name]["insts"] += 1
            for k in KEYS:
                if ins.startswith(k) or (k.endswith(".") and ins.startswith(k)):
                    kernels[name][k] += 1
    names = list(kernels)
    dem = subprocess.run(["c++filt"], input="\n".join(n.rstrip("'") for n in names), capture_output=True, text=True).stdout.splitlines()
    with open(os.path.join(ROOT, "profiles", "sass_summary.txt"), "w") as f:
        f.write("# python profiles/sass_summary.py -- cuobjdump -sass of torchmd-net_b200/torchmdnet/lib/libtmdn.so, per kernel: SASS\n"
                "# instruction count and the mnemonics that show the hardware path: tcgen05 (UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld,\n"
                "# UTCBAR = tcgen05.commit), TMA (UTMALDG / UTMASTG / UTMAREDG = cp.async.bulk.tensor load / store / reduce-add),\n"
                "# mbarriers (SYNCS), cp.async (LDGSTS), cluster barrier (UCGABAR), atomics (ATOMG / REDG).  arch of every cubin: %s\n"
                % ", ".join(arch))
        for n, d in zip(names, dem):
            d = re.sub(r"\(.*", "", d.replace("(anonymous namespace)::", ""))
            c = kernels[n]
            f.write("%-92s insts=%-6d %s\n" % (d[:92], c["insts"], " ".join("%s=%d" % (k, c[k]) for k in KEYS if c[k] and k != "FFMA")
                                             + (" FFMA=%d" % c["FFMA"] if c["FFMA"] else "")))
    print(len(names), "kernels;", arch)


if __name__ == "__main__":
    main()
