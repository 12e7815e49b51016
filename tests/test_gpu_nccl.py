"""Halo exchange inside the library over NCCL (rx_comm_init / rx_halo_exchange / rx_comm_allreduce_rms; csrc/rx_comm.cuh): a run
partitioned over the GPUs of the box — strips and graph partitions with RCM numbering, laminar and RANS — equals the single-partition
run to 1e-12, with rx_iterate doing exchanges and global norms without the host in the loop.  Mode "turb": the SST transport step on
every partition as well (rx_turb_iterate with a communicator; criteria in nccl_worker.turb_mode).  Needs at least two GPUs."""
import os
import subprocess
import sys

import pytest

from rxcase import ROOT

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode", ["strips", "graph", "turb", "unsteady"])
def test_partitioned_run_with_library_side_exchange_equals_single(mode):
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs two GPUs (one rank per GPU); run with gpurun --gpus 2")
    n = min(n, 4)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr", "127.0.0.1", "--master-port", "29731",
           os.path.join(ROOT, "tests", "nccl_worker.py"), mode]
    log = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    assert log.returncode == 0, log.stdout[-3000:]
    line = [l for l in log.stdout.splitlines() if l.startswith("NCCL_WORKER")]
    assert line and line[-1].endswith("OK"), log.stdout[-3000:]
    print(line[-1])
