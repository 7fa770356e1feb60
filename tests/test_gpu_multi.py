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


@pytest.mark.parametrize("mode", ["auto", "ipc"])
def test_multi_gpu_matches_oracle(mode):
    ng = torch.cuda.device_count()
    if ng < 2:
        pytest.skip("needs >= 2 GPUs")
    res = _run(min(ng, 8), mode)
    bad = [r for r in res if not r.get("ok", True)]
    assert not bad, bad
    assert any(r.get("swaps", 0) > 0 for r in res)  # the P2P block-row swap path was exercised
