"""Data-parallel learner on 2 GPUs of one box, with both histogram exchanges (peer memory over NVLink, NCCL all-reduce) (skipped on a single-GPU box; the N > 1 host logic is covered on CPU by
tests/test_capi_cpu.py::test_world_size_2_rendezvous_and_unique_id_handoff and the oracle's shard-count test)."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("exchange", ["peer", "nccl"])
def test_two_gpu_data_parallel_matches_oracle(native, exchange):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", "multi_gpu_worker.py")]
    env = dict(os.environ, LGB_B200_PEER_EXCHANGE="1" if exchange == "peer" else "0", LGB_B200_PEER_DEBUG="1")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "quantized dp2 model identical to oracle: True" in r.stdout
    assert "quantized sampled dp2 model identical to oracle: True" in r.stdout
    assert ("exchange over peer memory" in r.stderr) == (exchange == "peer"), r.stderr[-2000:]
