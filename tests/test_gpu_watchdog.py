"""A device-side flag wait that nobody satisfies must fail, not hang (in the
reference a lost MPI message hangs the job).  The trap poisons the CUDA context, so the
waiting kernel runs in a child process."""
import os
import subprocess
import sys
import time

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r"""
import sys, torch
sys.path.insert(0, %r)
from tenzing_b200 import _C
k = _C.kernels
k.set_wait_timeout_ms(300)
flag = torch.zeros(4, dtype=torch.int64, device="cuda")
torch.cuda.synchronize()
# epoch 7 is never published: a schedule that waits before anybody put
k.flags_wait(flag.data_ptr(), 4, 7, 0)
try:
    torch.cuda.synchronize()
except Exception as e:
    print("WATCHDOG:", type(e).__name__, str(e)[:200])
    sys.exit(3)
print("the wait returned although nobody published the flag")
sys.exit(0)
""" % ROOT


def test_unsatisfied_flag_wait_traps_instead_of_hanging():
    t0 = time.time()
    p = subprocess.run([sys.executable, "-c", CHILD], capture_output=True, text=True, timeout=120,
                       env=dict(os.environ, TENZING_ACK_NOTICE="1"))
    dt = time.time() - t0
    assert p.returncode == 3, (p.returncode, p.stdout[-500:], p.stderr[-500:])
    assert "WATCHDOG" in p.stdout
    assert dt < 100


def test_satisfied_wait_is_unaffected_by_the_watchdog():
    import torch
    from tenzing_b200 import _C
    k = _C.kernels
    flag = torch.zeros(2, dtype=torch.int64, device="cuda")
    k.flags_signal([flag.data_ptr(), flag.data_ptr() + 8], 5, 0)
    k.flags_wait(flag.data_ptr(), 2, 5, 0)
    torch.cuda.synchronize()
    assert flag.tolist() == [5, 5]
