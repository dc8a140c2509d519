#!/usr/bin/env python
"""SASS listings of the hot kernels for profiles/ (BASELINE north_star: "committed SASS listings").

  python tools/dump_sass.py [round_tag]

For each kernel: a histogram of the mnemonics (the lines that prove which pipes the kernel uses: UTCQMMA /
UTCIMMA = tcgen05 MMA, UTMALDG / UBLKCP = TMA bulk copies, POPC, LOP3, SHFL, REDUX ...) followed by the listing
with the encoding words stripped.  Runs on the container: cuobjdump reads the sm_100a cubin inside the built
library, no GPU needed."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "racing-slam_b200", "libslamgpu.so")
KERNELS = {
    "bf_tc4_scan_kernel": "K7b block-scaled FP4 tensor-core scan (headline, C5)",
    "bf_tc_scan_kernel": "K7a int8 tensor-core scan (exact kNN-2)",
    "bf_scan_kernelILi2ELb0EE": "K1 integer XOR+POPC scan, two query rows per thread",
    "proj_gate_kernel": "K4 pass 1: projection + gates, thread per point",
    "proj_search_kernelILi8EE": "K4 pass 2: grid search + candidate x observation Hamming, 8 lanes per point",
    "proj_match_group_kernelILi8EE": "K4 fused form (small visit lists)",
    "voc_transform_kernelILi10EE": "V1 vocabulary descent, 10 lanes per descriptor",
}


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
    text = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    parts = re.split(r"\n\s*Function : ", text)
    for part in parts[1:]:
        name = part.split("\n", 1)[0].strip()
        hit = [k for k in KERNELS if k in name]
        if not hit:
            continue
        key = hit[0]
        lines, hist = [], collections.Counter()
        for ln in part.split("\n")[1:]:
            m = re.match(r"\s*/\*([0-9a-f]{4})\*/\s+(.*?)\s*/\* 0x[0-9a-f]+ \*/", ln)
            if m:
                ins = m.group(2).rstrip(" ;")
                lines.append(f"/*{m.group(1)}*/ {ins}")
                op = ins.split()[1] if ins.startswith("@") and len(ins.split()) > 1 else ins.split()[0]
                hist[op.split(".")[0]] += 1
            elif ln.strip().startswith(".L_"):
                lines.append(ln.strip())
        short = re.sub(r"[^A-Za-z0-9_]", "", key)
        out = os.path.join(ROOT, "profiles", f"{tag}_sass_{short}.txt")
        with open(out, "w") as f:
            f.write(f"# {KERNELS[key]}\n# {name}\n# cuobjdump -sass racing-slam_b200/libslamgpu.so (sm_100a), encoding words stripped; "
                    f"{len([l for l in lines if l.startswith('/*')])} instructions\n#\n# mnemonic histogram (static count):\n")
            for op, n in hist.most_common():
                f.write(f"#   {op:12s} {n}\n")
            f.write("#\n")
            f.write("\n".join(lines) + "\n")
        print(out, len(lines), "lines;", ", ".join(f"{o} {n}" for o, n in hist.most_common(8)))


if __name__ == "__main__":
    main()
