"""The C-ABI library must load without a GPU and export every symbol include/*.h declares (no compute calls here)."""
import ctypes as C
import os
import re
import socket
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "lightgbm_b200_c_api.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(LGBM_\w+)\s*\(", text)))


def test_every_declared_symbol_is_exported(native):
    syms = declared_symbols()
    assert len(syms) >= 40
    missing = [s for s in syms if not hasattr(native, s)]
    assert not missing, missing


def test_reference_training_subset_is_present(native):
    # the 21 training-critical functions of SURVEY.md Appendix B (marked T)
    for s in ["LGBM_DatasetCreateFromSampledColumn", "LGBM_DatasetSerializeReferenceToBinary", "LGBM_ByteBufferGetAt", "LGBM_ByteBufferFree",
              "LGBM_DatasetCreateFromSerializedReference", "LGBM_DatasetInitStreaming", "LGBM_DatasetPushRowsWithMetadata",
              "LGBM_DatasetSetWaitForManualFinish", "LGBM_DatasetMarkFinished", "LGBM_DatasetGetNumData", "LGBM_DatasetGetNumFeature",
              "LGBM_DatasetSetFeatureNames", "LGBM_DatasetFree", "LGBM_BoosterCreate", "LGBM_BoosterUpdateOneIter", "LGBM_BoosterResetParameter",
              "LGBM_BoosterSaveModelToString", "LGBM_BoosterFree", "LGBM_NetworkInit", "LGBM_NetworkFree", "LGBM_GetLastError"]:
        assert hasattr(native, s)


def test_errors_do_not_cross_the_boundary(native):
    """Bad input returns -1 and leaves a message; nothing throws.  (Creating a dataset needs the GPU, so without one the call
    must fail loudly rather than fall back to a CPU path.)"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("CPU-only check")
    from synapseml_b200.lightgbm.native import LightGBMNativeError, validate
    col = np.array([1.0, 2.0, 3.0])
    idx = np.array([0, 1, 2], dtype=np.int32)
    cnt = np.array([3], dtype=np.int32)
    dpp = (C.POINTER(C.c_double) * 1)(col.ctypes.data_as(C.POINTER(C.c_double)))
    ipp = (C.POINTER(C.c_int) * 1)(idx.ctypes.data_as(C.POINTER(C.c_int)))
    out = C.c_void_p()
    rc = native.LGBM_DatasetCreateFromSampledColumn(dpp, ipp, 1, cnt.ctypes.data_as(C.POINTER(C.c_int)), 3, 1, C.c_int64(3), b"max_bin=255", C.byref(out))
    assert rc == -1
    with pytest.raises(LightGBMNativeError):
        validate(rc, "Dataset create")
    assert native.LGBM_NetworkInit(b"", 12400, 120, 1) == 0      # one machine: no-op
    assert native.LGBM_NetworkFree() == 0


_WORKER = r"""
import ctypes as C, os, sys
sys.path.insert(0, %(root)r)
import torch.distributed as dist
from synapseml_b200.lightgbm.native import lightgbmlib, validate
from synapseml_b200.lightgbm import network_manager as nm
dist.init_process_group(backend="gloo")
info = nm.get_global_network_info(default_listen_port=%(port)d)
assert info.world == 2 and len(info.lightgbm_network_string.split(",")) == 2
blob = (C.c_uint8 * 128)(*([(7 * i + 3) %% 256 for i in range(128)] if info.rank == 0 else [0] * 128))
r = C.c_int()
validate(lightgbmlib().LGBM_B200_NetworkBootstrapBlob(info.lightgbm_network_string.encode(), info.local_listen_port, 30, blob, 128, C.byref(r)))
assert r.value == dist.get_rank(), (r.value, dist.get_rank())
assert list(blob) == [(7 * i + 3) %% 256 for i in range(128)]
dist.barrier()
# LGBM_NetworkInit opens the TCP star that carries the fit's host-side values (no GPU, no NCCL); values come back in rank order
nm.init_lightgbm_network(info, timeout=30)
for trial in range(3):
    out = (C.c_int64 * 2)()
    rk, wd = C.c_int(), C.c_int()
    validate(lightgbmlib().LGBM_B200_NetworkAllGatherInt64(C.c_int64(100 * trial + info.rank), out, C.byref(rk), C.byref(wd)))
    assert (rk.value, wd.value) == (info.rank, 2) and list(out) == [100 * trial, 100 * trial + 1], list(out)
nm.free_lightgbm_network()
# row sharding bookkeeping the estimator does under torch.distributed
counts = [None, None]
dist.all_gather_object(counts, 1000 + dist.get_rank())
assert counts == [1000, 1001]
# the bin sample is taken from the head of the GLOBAL row order (df.limit in the reference), whichever ranks hold those rows
import numpy as np
from synapseml_b200.lightgbm.estimators import LightGBMClassifier, LocalDataFrame
rk = dist.get_rank()
Xl = np.arange(rk * 30 * 3, (rk + 1) * 30 * 3, dtype=np.float64).reshape(30, 3)
est = LightGBMClassifier(samplingMode="fixed", binSampleCount=40)
got = est._sample_across_ranks(LocalDataFrame.from_arrays(Xl, np.zeros(30, dtype=np.float32), num_partitions=2), 60, [30, 30], rk, dist)
if rk == 0:
    assert got.shape == (40, 3) and np.array_equal(got, np.arange(120, dtype=np.float64).reshape(40, 3)), got
else:
    assert got is None
dist.barrier()
open(os.path.join(%(out)r, "rank%%d.ok" %% dist.get_rank()), "w").write("ok")
"""


def test_world_size_2_rendezvous_and_unique_id_handoff(native, tmp_path):
    """N > 1 host path on CPU: gloo rendezvous builds the machine list (NetworkManager), then the library's own TCP bootstrap
    moves rank 0's blob to rank 1, and LGBM_NetworkInit's TCP star all-gathers host values in rank order."""
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    master_port = s.getsockname()[1]
    s.close()
    script = tmp_path / "worker.py"
    script.write_text(_WORKER % dict(root=ROOT, port=23400 + (os.getpid() % 500) * 4, out=str(tmp_path)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(master_port), str(script)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=180)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert (tmp_path / "rank0.ok").exists() and (tmp_path / "rank1.ok").exists()


def test_hot_kernels_are_built_for_sm_100a_without_spills(native):
    """Build evidence that needs no GPU: the library carries sm_100a code only, and the kernels of the boosting loop's hot path keep
    their working set in registers — no local-memory frame on the histogram kernels' main variants or on the plain split-search
    kernel (the categorical / wide / constrained search is compiled into its own instantiation so that it costs the plain one nothing)."""
    import shutil
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    lib = os.path.join(ROOT, "synapseml_b200", "lib", "lib_lightgbm.so")
    out = subprocess.run([cuobjdump, "--dump-resource-usage", lib], capture_output=True, text=True, check=True).stdout
    archs = set(re.findall(r"arch = (sm_\w+)", out))
    assert archs == {"sm_100a"}, archs
    usage = {}
    lines = out.splitlines()
    for i, line in enumerate(lines):
        m = re.match(r"\s*Function (\S+):", line)
        if m and i + 1 < len(lines):
            usage[m.group(1)] = {k: int(v) for k, v in re.findall(r"(REG|STACK|SHARED|LOCAL):(\d+)", lines[i + 1])}
    find_plain = [u for n, u in usage.items() if re.match(r"_ZN3lgb9k_bt_findILb[01]ELb0EEE", n)]
    find_ext = [u for n, u in usage.items() if re.match(r"_ZN3lgb9k_bt_findILb[01]ELb1EEE", n)]
    hist = [u for n, u in usage.items() if n.startswith("_ZN3lgb9k_bt_histI")]
    assert len(find_plain) == 2 and len(find_ext) == 2 and len(hist) >= 12
    assert all(u["STACK"] == 0 and u["LOCAL"] == 0 for u in find_plain), find_plain
    assert all(u["SHARED"] < min(e["SHARED"] for e in find_ext) for u in find_plain)
    assert all(u["REG"] <= 64 and u["LOCAL"] == 0 and u["STACK"] <= 32 for u in hist), hist      # 1024 threads per CTA need <= 64 registers
    assert all(u["LOCAL"] == 0 for u in usage.values())
