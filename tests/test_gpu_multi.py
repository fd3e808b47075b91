"""Data-parallel learner on 2 / 4 / 8 GPUs of one box with each histogram exchange: peer memory all-to-all, peer memory feature-sliced
(reduce-scatter by feature + all-gather of best splits) and the NCCL all-reduce fallback.  Skipped where the box has fewer GPUs; the
N > 1 host logic is covered on CPU by tests/test_capi_cpu.py::test_world_size_2_rendezvous_and_unique_id_handoff and the oracle's
shard-count test, and bench.py repeats a parity check at whatever N it is launched with."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("exchange", ["peer", "sliced", "nccl"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_data_parallel_matches_oracle(native, world, exchange):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=%d" % world, "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", "multi_gpu_worker.py")]
    env = dict(os.environ, LGB_B200_PEER_EXCHANGE="0" if exchange == "nccl" else "1", LGB_B200_SLICED="1" if exchange == "sliced" else "0",
               LGB_B200_PEER_DEBUG="1")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "quantized dp%d model identical to oracle: True" % world in r.stdout
    assert "quantized sampled dp%d model identical to oracle: True" % world in r.stdout
    assert "quantized categorical dp%d model identical to oracle: True" % world in r.stdout
    assert "quantized wide dp%d model identical to oracle: True" % world in r.stdout
    assert "quantized constrained dp%d model identical to oracle: True" % world in r.stdout
    expect = {"peer": "all-to-all", "sliced": "feature-sliced", "nccl": "NCCL all-reduce exchange"}[exchange]
    assert expect in r.stderr, r.stderr[-2000:]
    log = os.environ.get("LGB_B200_MULTI_LOG")      # profiles/r02_multi_gpu_parity.log is written this way
    if log:
        with open(log, "a") as f:
            f.write("---- world=%d exchange=%s rc=%d\n%s\n" % (world, exchange, r.returncode, r.stdout))
            f.write("\n".join(l for l in r.stderr.splitlines() if l.startswith("[peer]")) + "\n")
