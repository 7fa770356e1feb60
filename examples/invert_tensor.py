"""Single-GPU convenience API: invert a dense fp64 torch tensor."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

from mpi_jordan_crazy_acceleration_b200 import invert  # noqa: E402

a = torch.randn(1024, 1024, dtype=torch.float64)
x = invert(a, m=128)
print("max |A X - I| =", float((a.cuda() @ x - torch.eye(1024, dtype=torch.float64, device="cuda")).abs().max()))
