#!/usr/bin/env python
"""Sum of NVLink data counters of one GPU (`nvidia-smi nvlink -gt d`), for a
before/after difference around a run:  python tools/nvlink_delta.py [gpu] -> "tx_kib rx_kib"."""
import re
import subprocess
import sys

gpu = sys.argv[1] if len(sys.argv) > 1 else "0"
out = subprocess.run(["nvidia-smi", "nvlink", "-gt", "d", "-i", gpu], capture_output=True, text=True).stdout
tx = sum(int(x) for x in re.findall(r"Data Tx:\s*(\d+)\s*KiB", out))
rx = sum(int(x) for x in re.findall(r"Data Rx:\s*(\d+)\s*KiB", out))
print(tx, rx)
