"""CPU: the offline helpers either side of the path and repository hygiene."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_convert_glove_vec_truncates_and_lowercases(tmp_path):
    """util/convert_glove_vec.py:1-23 of the reference: first K dims, lower-cased words, same JSON schema -
    and the result loads through the loader's word-vector path."""
    import induction_b200  # noqa: F401
    from induction_b200.util.convert_glove_vec import convert, main
    src = tmp_path / "glove.6B.50d.json"
    vecs = [{"word": "The", "vec": [0.1 * i for i in range(50)]}, {"word": "hello", "vec": [1.0] * 50}]
    src.write_text(json.dumps(vecs))
    out = tmp_path / "glove.6B.5d.json"
    assert convert(str(src), str(out), 5) == 2
    got = json.loads(out.read_text())
    assert got == [{"word": "the", "vec": vecs[0]["vec"][:5]}, {"word": "hello", "vec": [1.0] * 5}]
    assert main(["--k", "3", "--input", str(src)]) == 0
    assert len(json.loads((tmp_path / "glove.6B.3d.json").read_text())[0]["vec"]) == 3


def test_pytest_collects_only_tests_dir():
    """Scratch scripts elsewhere in the tree must never be collected (one used to run CUDA at import)."""
    r = subprocess.run([sys.executable, "-m", "pytest", "--collect-only", "-q", "-m", "not gpu"], cwd=ROOT,
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "tools/" not in r.stdout


def test_tools_compile():
    import py_compile
    for name in sorted(os.listdir(os.path.join(ROOT, "tools"))):
        if name.endswith(".py"):
            py_compile.compile(os.path.join(ROOT, "tools", name), doraise=True)


def test_every_exported_entry_point_is_declared_in_the_header():
    """include/indnet.h is the boundary: an `extern "C"` entry in csrc/ that the header does not declare is invisible to
    the ctypes binding (which is generated from the header) - only the timeline-build debug readers are exempt."""
    import re
    import induction_b200  # noqa: F401
    from induction_b200._lib import parse_header
    declared = set(parse_header())
    defined = set()
    csrc = os.path.join(ROOT, "induction-network-on-fewrel_b200", "csrc")
    for name in os.listdir(csrc):
        if name.endswith(".cu"):
            src = open(os.path.join(csrc, name)).read()
            defined |= set(re.findall(r'extern\s+"C"\s+(?:const\s+char\s*\*|long long|int)\s+(indnet_\w+)\s*\(', src))
    extra = {n for n in defined - declared if not n.startswith("indnet_debug")}
    assert not extra, sorted(extra)
    assert not (declared - defined), sorted(declared - defined)
