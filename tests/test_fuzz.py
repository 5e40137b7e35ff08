"""Bounded runs of the randomised sweeps (tools/fuzz_*.py): random image sizes, extractor parameters and image families chosen
to produce ties, empty levels and saturation; random frame / local-map / vocabulary scenarios. The long runs (tens of thousands
of configurations) found the one descriptor bit that led to the glibc sincosf restatement; these keep a fixed
sample of them in the suites. The comparison is always against the reference's own object code (oracle/_ref)."""
import os
import subprocess
import sys

import pytest

from oracle import ref as R

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
needs_ref = pytest.mark.skipif(not R.available(), reason="oracle/_ref not built and /root/reference absent")


def run(tool, *args):
    p = subprocess.run([sys.executable, os.path.join(ROOT, "tools", tool), *map(str, args)], capture_output=True, text=True, cwd=ROOT,
                       timeout=600)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    return p.stdout


@needs_ref
def test_oracle_equals_reference_on_random_configurations():
    out = run("fuzz_oracle_cpu.py", 300, 101, 60)  # 60 cases, seed 101
    assert " 0 mismatches" in out


@pytest.mark.gpu
def test_cuda_path_equals_reference_on_random_configurations():
    out = run("fuzz_vs_reference.py", 300, 102, 80)
    assert "80 random configurations" in out and " 0 mismatches" in out


@pytest.mark.gpu
def test_cuda_rows_equal_reference_on_random_scenarios():
    out = run("fuzz_rows.py", 300, 103, 60)
    assert " 0 mismatches" in out
