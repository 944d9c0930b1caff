"""The ``.share_memory()`` / CUDA-IPC contract of the reference's multi-process loop (src/NICE_SLAM.py:117-148,
357-376): ``shared_c`` and ``shared_decoders`` are created once, handed to spawned processes, the MAPPER process
trains them in place and the TRACKER process copies them with ``update_para_from_mapping`` (src/Tracker.py:137-149).
tests/ipc_driver.py plays the main + tracker process in its own interpreter and spawns the mapper on the objects it
created; afterwards it must see the trained grids and colour decoder through its own handles -- with
``shared.prepare_shared`` (zero-copy) and without it (the loop re-allocates, keeps the old storages and refreshes
them in ``MappingLoop.finish``)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode", ["prepared", "unprepared"])
def test_tracker_process_sees_what_the_mapper_process_trained(mode):
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    driver = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ipc_driver.py")
    r = subprocess.run([sys.executable, driver, mode], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "IPC_OK" in r.stdout, (r.returncode, r.stdout[-2000:], r.stderr[-4000:])
