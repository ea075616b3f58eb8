#!/usr/bin/env python3
"""tests/golden/g1_monomial_4096.json: the monomial-basis G1 points [tau^k]G1 of the mainnet KZG ceremony, copied from the
consensus-spec-tests presets file the reference's submodule carries (public ceremony output; kateth's own trusted_setup_4096.json has
only g1_lagrange and g2_monomial).  Used by the g1_monomial cross-check tests (kb200_setup_check_g1_monomial).  Run in the build
container (reads /root/reference, which does not exist on the GPU box)."""
import json, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src = "/root/reference/consensus-spec-tests/presets/mainnet/trusted_setups/trusted_setup_4096.json"
d = json.load(open(src))
ours = json.load(open(os.path.join(ROOT, "tests/golden/trusted_setup_4096.json")))
assert d["g1_lagrange"] == ours["g1_lagrange"] and d["g2_monomial"] == ours["g2_monomial"], "presets file and kateth's setup differ"
assert len(d["g1_monomial"]) == 4096
json.dump({"g1_monomial": d["g1_monomial"]}, open(os.path.join(ROOT, "tests/golden/g1_monomial_4096.json"), "w"))
print("wrote tests/golden/g1_monomial_4096.json")
