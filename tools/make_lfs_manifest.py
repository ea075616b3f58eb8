#!/usr/bin/env python3
"""Record the Git-LFS pointer (sha256 oid + size) of every consensus-spec-tests KZG data.yaml that kateth's
conformance tests read (/root/reference/src/kzg/setup.rs:305-317) into tests/golden/lfs_oids.json, and copy the
public trusted setup next to it.  Runs only in the build container (where /root/reference is mounted); the
outputs are committed so that nothing at test/bench time needs /root/reference.
"""
import json, os, re, shutil, sys
REF = "/root/reference"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
base = os.path.join(REF, "consensus-spec-tests/tests/general/deneb/kzg")
out = {}
for handler in sorted(os.listdir(base)):
    d = os.path.join(base, handler, "kzg-mainnet")
    for case in sorted(os.listdir(d)):
        txt = open(os.path.join(d, case, "data.yaml")).read()
        m = re.search(r"oid sha256:([0-9a-f]{64})\s+size (\d+)", txt)
        assert m, (handler, case)
        out.setdefault(handler, {})[case] = {"oid": m.group(1), "size": int(m.group(2))}
os.makedirs(os.path.join(ROOT, "tests/golden"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "tests/golden/lfs_oids.json"), "w"), indent=0, sort_keys=True)
shutil.copy(os.path.join(REF, "trusted_setup_4096.json"), os.path.join(ROOT, "tests/golden/trusted_setup_4096.json"))
print({k: len(v) for k, v in out.items()}, sum(len(v) for v in out.values()))
