"""CPU, build container only: tests/golden/ref_layers.npz is exactly what the reference's own unmodified files return
under tests/golden/tf_shim (runs make_ref_golden.py --check in a subprocess so the stand-in ``tensorflow`` never enters
this interpreter).  Skipped where /root/reference is not mounted (the GPU box)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/algorithm/deep_learning"


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference not mounted")
def test_committed_fixture_is_the_reference_output():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "golden", "make_ref_golden.py"), "--check"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "reproduced bit for bit" in r.stdout


def test_shim_is_test_infrastructure_only():
    for dp, _, fs in os.walk(os.path.join(ROOT, "recommender_system_b200")):
        for f in fs:
            if f.endswith(".py"):
                src = open(os.path.join(dp, f)).read()
                assert "tf_shim" not in src and "ref_layers" not in src, os.path.join(dp, f)
    assert "tf_shim" not in open(os.path.join(ROOT, "bench.py")).read()
