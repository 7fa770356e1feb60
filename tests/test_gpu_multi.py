"""Multi-GPU tests (need >= 2 GPUs): launched through torchrun, one rank per GPU."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(nproc, mode, extra_env=None):
    env = dict(os.environ)
    env.update(extra_env or {})
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", "29617", os.path.join(ROOT, "tests", "_mp_worker.py"), mode]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env, cwd=ROOT)
    line = [ln for ln in out.stdout.splitlines() if ln.startswith("MPRESULT ")]
    assert line, out.stdout[-3000:] + out.stderr[-3000:]
    return json.loads(line[-1][len("MPRESULT "):])


@pytest.mark.parametrize("mode", ["auto", "ipc", "torch"])
def test_multi_gpu_matches_oracle(mode):
    ng = torch.cuda.device_count()
    if ng < 2:
        pytest.skip("needs >= 2 GPUs")
    res = _run(min(ng, 8), mode)
    bad = [r for r in res if not r.get("ok", True)]
    assert not bad, bad
    assert any(r.get("swaps", 0) > 0 for r in res)  # the P2P block-row swap path was exercised


def test_flag_protocol_litmus_two_gpus():
    """Race detection for the release/acquire flag protocol used by every cross-GPU exchange."""
    import ctypes as C

    from mpi_jordan_crazy_acceleration_b200 import _native

    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    errs = C.c_ulonglong(123)
    for a, b in ((0, 1), (1, 0)):
        _native.check(_native.cuda().jb_flag_litmus(a, b, 200000, C.byref(errs)))
        assert errs.value == 0
