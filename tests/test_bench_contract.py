"""Driver-facing contracts that can be checked without a GPU: the reference arm of bench.py reports
that the TensorFlow-1.x reference cannot run here (one JSON line, exit 0), and __graft_entry__.build()
compiles both native libraries in-tree."""
import json
import os
import subprocess
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_reports_unavailable_and_exits_zero():
    p = subprocess.run([sys.executable, os.path.join(REPO, "bench.py"), "--impl", "reference", "--gpus", "1", "--steps", "2",
                        "--warmup", "1"], capture_output=True, text=True, timeout=300, cwd=REPO)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.strip().startswith("{")]
    assert len(lines) == 1
    out = json.loads(lines[0])
    assert out["impl"] == "reference" and isinstance(out["unavailable"], str) and out["unavailable"]


def test_graft_entry_build_produces_in_tree_libraries():
    sys.path.insert(0, REPO)
    import __graft_entry__ as g
    g.build()
    lib = os.path.join(REPO, "distributed_tensorflow_b200", "_lib")
    for name in ("libdtf_kernels.so", "libdtf_native.so"):
        assert os.path.getsize(os.path.join(lib, name)) > 100_000
    assert callable(g.smoke)


def test_packaging_stub_lists_the_three_packages():
    p = subprocess.run([sys.executable, "setup.py", "--name", "--version"], capture_output=True, text=True, timeout=120, cwd=REPO)
    assert p.returncode == 0, p.stderr[-1000:]
    assert p.stdout.split()[-2:] == ["distributed_tensorflow_b200", "1.0"]
    for pkg in ("trainer", "auto_stop_ps", "distributed_tensorflow_b200"):
        assert os.path.isfile(os.path.join(REPO, pkg, "__init__.py"))
