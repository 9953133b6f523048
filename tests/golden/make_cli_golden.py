"""Writes tests/golden/cli_config1.json: digest and header of the bytes the REFERENCE's own command-line tool
(oracle/_ref/vkt_ref_cli = /root/reference/src/cli/main.cpp compiled unmodified by oracle/Makefile) produces for BASELINE
config 1.  Run where /root/reference exists:  python tests/golden/make_cli_golden.py"""
import hashlib
import json
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
CLI = os.path.join(ROOT, "oracle", "_ref", "vkt_ref_cli")
STAGES = ["declare-sv --dims 64 64 64 --data-format uint8", "fill --value 0.1",
          "fill-range --value 1 --first 0 0 0 --last 8 8 8", "resample --dims 32 32 32 --data-format uint16"]

subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "ref"], check=True)
data = b""
for s in STAGES:
    data = subprocess.run([CLI] + s.split(), input=data, capture_output=True, check=True).stdout
out = {"pipeline": " | ".join("vkt " + s for s in STAGES), "bytes": len(data), "header_hex": data[:44].hex(),
       "sha256": hashlib.sha256(data).hexdigest(), "tool": "oracle/_ref/vkt_ref_cli (reference src/cli/main.cpp, unmodified)"}
json.dump(out, open(os.path.join(ROOT, "tests", "golden", "cli_config1.json"), "w"), indent=1)
print(out)
