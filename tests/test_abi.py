"""CPU tests of the drop-in boundary: the C-ABI library builds, loads, exports every symbol include/kateth_b200.h
declares, and refuses to run without a GPU (no CPU fallback)."""
import os, re, subprocess, sys
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    txt = open(os.path.join(ROOT, "include", "kateth_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(kb200_[a-z0-9_]+)\s*\(", txt)))


def test_header_and_binding_agree():
    from kateth_b200.setup import SYMBOLS
    assert _header_symbols() == sorted(SYMBOLS)


def test_library_exports_every_declared_symbol():
    import kateth_b200 as K
    assert os.path.exists(K.lib_path()), "build first: python -c 'import __graft_entry__ as g; g.build()'"
    out = subprocess.check_output(["nm", "-D", "--defined-only", K.lib_path()]).decode()
    exported = set(re.findall(r"\bT (kb200_[a-z0-9_]+)", out))
    assert set(_header_symbols()) <= exported
    l = K.lib()
    for s in _header_symbols():
        assert hasattr(l, s)


def test_library_is_sm100a_sass():
    import kateth_b200 as K
    out = subprocess.run(["cuobjdump", "-lelf", K.lib_path()], capture_output=True, text=True).stdout
    assert "sm_100a" in out


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import kateth_b200 as K
    with pytest.raises(K.KzgError) as e:
        K.Setup.load_json(os.path.join(ROOT, "tests/golden/trusted_setup_4096.json"))
    assert e.value.status == -1


def test_engine_does_not_reference_oracle():
    # the product path must never import, link or call anything under oracle/
    for dirpath, _, files in os.walk(os.path.join(ROOT, "kateth_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h", "Makefile")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "kzg_oracle" not in txt and "from oracle" not in txt and "import oracle" not in txt, f
    out = subprocess.check_output(["nm", "-D", os.path.join(ROOT, "kateth_b200", "libkateth_b200.so")]).decode()
    assert "kzo_" not in out


def _write_txt(path, mutate=None):
    import json
    d = json.load(open(os.path.join(ROOT, "tests/golden/trusted_setup_4096.json")))
    g1 = [x[2:] for x in d["g1_lagrange"]]; g2 = [x[2:] for x in d["g2_monomial"]]
    lines = [str(len(g1)), str(len(g2))] + g1 + g2 + g1[:3]          # trailing G1 monomial lines are ignored
    if mutate:
        lines = mutate(lines)
    open(path, "w").write("\n".join(lines) + "\n")


def test_txt_setup_parser(tmp_path):
    """c-kzg text setups: parse errors are reported before the device is touched; a good file gets as far as the GPU."""
    import torch
    import kateth_b200 as K
    good = tmp_path / "trusted_setup.txt"
    _write_txt(good)
    if torch.cuda.is_available():
        s = K.Setup.load_txt(str(good)); s.close()
    else:
        with pytest.raises(K.KzgError) as e:
            K.Setup.load_txt(str(good))
        assert e.value.status == -1                                   # parsed; no CUDA device (no CPU fallback)
    cases = [(lambda l: ["4095"] + l[1:], -10), (lambda l: [l[0], "64"] + l[2:], -11), (lambda l: l[:100], -3),
             (lambda l: l[:2] + ["zz" + l[2][2:]] + l[3:], -3), (lambda l: l[:2] + [l[2][:-2]] + l[3:], 3)]
    for k, (mut, want) in enumerate(cases):
        bad = tmp_path / ("bad%d.txt" % k)
        _write_txt(bad, mut)
        with pytest.raises(K.KzgError) as e:
            K.Setup.load_txt(str(bad))
        assert e.value.status == want, k
    with pytest.raises(K.KzgError) as e:
        K.Setup.load_txt(str(tmp_path / "missing.txt"))
    assert e.value.status == -2
