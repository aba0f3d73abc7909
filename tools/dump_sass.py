#!/usr/bin/env python
"""Regenerate profiles/sass/*.sass and profiles/sass_tcgen05_evidence.txt from
the built library (cuobjdump -sass; runs without a GPU)."""
import collections
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "gru_mpi_cuda_b200" / "lib" / "libb200gru.so"
OUT = ROOT / "profiles" / "sass"
KEEP = ["gru_layer_pair_kernel", "gru_layer_kernel", "fc_sample_kernel", "gru_persistent_kernel",
        "compact_plan_kernel", "compact_move_kernel", "compact_commit_kernel",
        "sgemm_tn_kernel", "exact_gates_kernel", "exact_sample_kernel"]
PAT = re.compile(r"\b(UTC[A-Z]*MMA(?:\.2CTA)?|UTMALDG\.\dD(?:\.2CTA)?(?:\.MULTICAST)?|UTMASTG\.\dD|LDTM\.x\d+|STTM\.x\d+|"
                 r"UTCBAR(?:\.2CTA)?(?:\.MULTICAST)?|MUFU\.(?:TANH|EX2)|ELECT|LDG\.E\.ENL2\.LTC128B\.256|"
                 r"SYNCS\.ARRIVE\.TRANS64(?:\.RED)?(?:\.A1T0)?|REDG\.E\.ADD\.STRONG\.(?:GPU|SYS)|STG\.E\.128\.STRONG\.SYS|"
                 r"LDG\.E\.STRONG\.(?:GPU|SYS)|UCGABAR_ARV|HMMA)\b")


def main():
    txt = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True, check=True).stdout
    OUT.mkdir(parents=True, exist_ok=True)
    for f in OUT.glob("*.sass"):
        f.unlink()
    funcs = re.split(r"\n\s*Function : ", txt)
    counts = {}
    for blk in funcs[1:]:
        mangled = blk.split("\n", 1)[0].strip()
        name = next((k for k in KEEP if k in mangled), None)
        if not name:
            continue
        tag = name
        if "ILb1E" in mangled:
            tag += "_layer0"
        elif "ILb0E" in mangled:
            tag += "_layerN"
        (OUT / f"{tag}.sass").write_text("Function : " + blk)
        c = collections.Counter(m.group(1) for m in PAT.finditer(blk))
        counts[tag] = c
    lines = []
    for tag in sorted(counts):
        for op, k in sorted(counts[tag].items(), key=lambda kv: (-kv[1], kv[0])):
            lines.append(f"{k:5d} {tag} {op}")
    (ROOT / "profiles" / "sass_tcgen05_evidence.txt").write_text("\n".join(lines) + "\n")
    print("wrote", len(counts), "listings:", ", ".join(sorted(counts)))


if __name__ == "__main__":
    sys.exit(main())
