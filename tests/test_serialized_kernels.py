"""The two persistent kernels of the scheduler need each other and must run side by side. A profiler (ncu), a debugger or
CUDA_LAUNCH_BLOCKING=1 makes the runtime run kernels one after the other: the first kernel of the pair then waits for a partner
that cannot start before it ends. The library notices (one second without progress and without the partner's start flag: device
code 901), switches to the one-kernel scheduler and launches the unfinished packets again -- the timestep completes and the
caller sees nothing but the delay. Checked here in a child process with CUDA_LAUNCH_BLOCKING=1."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)

CHILD = r"""
import sys
sys.path.insert(0, %(repo)r); sys.path.insert(0, %(here)r)
import numpy as np
import goldenlib
from artis_b200 import abi
from artis_b200.transport import Transport
fx = goldenlib.load_fixture("classic_nogrey_ts12.npz")
T = Transport(%(lib)r)
T.init(fx["static"], seed=99)
T.upload_cellstate(fx["cell"])
T.zero_estimators()
before = T.scheduler_geometry()
pk = fx["pre"].copy()
T.update_packets(pk)
after = T.scheduler_geometry()
st = fx["static"]
t2 = st["ts_start"][fx["nts"]] + st["ts_width"][fx["nts"]]
assert ((pk["type"] == abi.TYPE_ESCAPE) | (pk["prop_time"] == t2)).all()
c = T.download_estimators()["counters"]
assert c[45] == pk["interactions"].sum()
newly = (pk["type"] == abi.TYPE_ESCAPE) & (fx["pre"]["type"] != abi.TYPE_ESCAPE)
assert c[33] == newly.sum()
pk2 = fx["pre"].copy()
T.zero_estimators()
T.update_packets(pk2)            # the one-kernel scheduler from now on: no second delay
assert ((pk2["type"] == abi.TYPE_ESCAPE) | (pk2["prop_time"] == t2)).all()
n1, n2 = int(newly.sum()), int(((pk2["type"] == abi.TYPE_ESCAPE) & (fx["pre"]["type"] != abi.TYPE_ESCAPE)).sum())
assert abs(n1 - n2) <= 0.2 * max(n1, n2) + 20, (n1, n2)   # packets in flight at the hand-back redraw a pending action: statistics, not histories
print("GEOMETRY", list(before), list(after))
T.finalize()
"""


def test_kernel_pair_falls_back_when_kernels_are_serialised(gpu_lib):
    env = dict(os.environ, CUDA_LAUNCH_BLOCKING="1")
    out = subprocess.run([sys.executable, "-c", CHILD % dict(repo=REPO, here=HERE, lib=gpu_lib)], env=env, capture_output=True,
                         text=True, timeout=300)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    line = [ln for ln in out.stdout.splitlines() if ln.startswith("GEOMETRY")][-1]
    before, after = eval(line[len("GEOMETRY"):].replace("] [", "],["))
    assert before[0] == 2 and after[0] == 1, line  # the pair was planned, one kernel ran
