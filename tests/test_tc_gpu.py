"""tcgen05 kernels (fp16 path) against matmul references, each case in a subprocess (a kernel trap would poison
the CUDA context of the whole pytest process)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def probe(*args):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "tc_probe.py")] + [str(a) for a in args],
                       capture_output=True, text=True, timeout=300)
    line = [l for l in r.stdout.splitlines() if l.startswith("RESULT")]
    assert r.returncode == 0 and line, (r.stdout[-2000:], r.stderr[-2000:])
    assert line[0].endswith("ok=1"), line[0]
    return line[0]


@pytest.mark.parametrize("H", [128, 256, 512])
@pytest.mark.parametrize("mode", [0, 1])
def test_layer_gemm(H, mode):
    probe("layer", H, mode, 0, 128 * 3 + 5)          # a few tiles, ragged tail
    probe("layer", H, mode, 0, 148 * 128 * 3 + 77)   # several tiles per CTA: ring + TMEM double-buffer phases


@pytest.mark.parametrize("H", [128, 256, 512])
def test_dw_gemm(H):
    probe("dw", H, 0, 128 * 2 + 9)
    probe("dw", H, 0, 148 * 128 * 2 + 300)
