"""LGBM_B200_RootHistogramNoAlloc: the histogram entry point for callers that cannot allocate GPU memory (a plugin running
inside an external memory pool or a fixed per-task budget, where cudaMalloc / cudaFree are not allowed).  Its scratch is a
static array inside the library, shared between calls, so this test calls it twice in a row — the second call starts from
the first call's leftovers and must still be right — and compares both results with the existing, allocating entry point
(LGBM_B200_BoosterRootHistogram).  Integer histograms: the comparison is exact.

Collected last on purpose (file name), so a failure here cannot hide another test's result under `pytest -x`.
"""
import ctypes as C

import numpy as np
import pytest

from helpers import NativeBooster, NativeDataset
from synapseml_b200.data import synthetic as S

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,f", [(70000, 28), (9000, 130)])
def test_no_alloc_histogram_matches_allocating_entry_point_twice(native, n, f):
    import torch
    params = "objective=binary num_leaves=4 min_data_in_leaf=1 max_bin=255 use_quantized_grad=true stochastic_rounding=false verbosity=-1"
    X, y = S.higgs_shape(n, f=f)
    nd = NativeDataset(X, y, params, micro_batches=[8192])
    nb = NativeBooster(nd, params)
    nb.update()
    expected = nb.root_histogram()                      # existing entry point (allocates its scratch with cudaMalloc)
    qg, qh, _ = nb.quant_gradients()
    rows = nd.row_major()
    used = nd.num_used()
    gh = ((qg.astype(np.uint8).astype(np.uint16) << 8) | qh.astype(np.uint8).astype(np.uint16)).view(np.int16)
    d_rows = torch.from_numpy(rows).cuda()
    d_gh = torch.from_numpy(gh).cuda()
    d_out = torch.full((used, 256, 2), -7, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    for call in range(2):                               # the second call must not see the first call's sums
        d_out.fill_(-7)
        torch.cuda.synchronize()
        rc = native.LGBM_B200_RootHistogramNoAlloc(C.c_void_p(d_rows.data_ptr()), C.c_int32(rows.shape[1]), C.c_void_p(d_gh.data_ptr()),
                                                   C.c_int32(n), C.c_int32(used), C.c_int32(4), C.c_void_p(d_out.data_ptr()), None)
        assert rc == 0, native.LGBM_GetLastError()
        torch.cuda.synchronize()
        assert np.array_equal(d_out.cpu().numpy(), expected), "call %d" % call
    # beyond the fixed capacity: refused with a clear message, nothing written
    rc = native.LGBM_B200_RootHistogramNoAlloc(C.c_void_p(d_rows.data_ptr()), C.c_int32(384), C.c_void_p(d_gh.data_ptr()), C.c_int32(n),
                                               C.c_int32(300), C.c_int32(4), C.c_void_p(d_out.data_ptr()), None)
    assert rc == -1 and b"capacity" in native.LGBM_GetLastError()
    nb.free()
    nd.free()
