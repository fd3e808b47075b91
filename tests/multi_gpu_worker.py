"""Worker for the multi-GPU parity check (launched by torchrun, one rank per GPU): each rank trains on its contiguous row
shard through the estimator API; the resulting model must be byte-identical to the oracle's simulated data-parallel learner
with the same number of machines (quantized mode: integer histograms make the all-reduce order irrelevant)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from helpers import strip_parameters  # noqa: E402
from synapseml_b200.data import synthetic as S  # noqa: E402
from synapseml_b200.lightgbm.estimators import LightGBMClassifier, LocalDataFrame  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group(backend="cpu:gloo,cuda:nccl", rank=rank, world_size=world)
    n = int(os.environ.get("MG_ROWS", "200000"))
    iters = int(os.environ.get("MG_ITERS", "8"))
    lo, hi = n * rank // world, n * (rank + 1) // world
    X, y = S.higgs_shape(hi - lo, start=lo)
    ok = True
    sampling = dict(baggingFraction=0.7, baggingFreq=1, featureFraction=0.7)      # every rank bags its own rows, all draw the same columns
    cats = dict(categoricalSlotIndexes=[2, 3], minDataPerGroup=50)      # two signal columns as 63 and 3 categories

    def with_cats(A):
        A = A.copy()
        A[:, 2] = np.floor((A[:, 2] + np.pi) * 10.0)      # the label depends on sin(x2): not monotone in the category
        return A                                          # (column 3 has three values already: one category against the rest)

    wide = dict(maxBin=1023)      # features with more than 256 bins: several uint8 columns each, searched as one feature
    # monotone constraints and per-leaf column samples: both are decided from sums every rank sees alike
    shaped = dict(monotoneConstraints=[1, -1] + [0] * 26, monotonePenalty=0.5, passThroughArgs="feature_fraction_bynode=0.6")
    for case, (quant, extra) in enumerate(((True, {}), (False, {}), (True, sampling), (True, cats), (True, wide), (True, shaped))):
        is_cat = "categoricalSlotIndexes" in extra
        Xc = with_cats(X) if is_cat else X
        kw = dict(maxBin=63)
        kw.update(extra)
        est = LightGBMClassifier(numIterations=iters, numLeaves=31, useQuantizedGrad=quant, stochasticRounding=False,
                                 samplingMode="fixed", microBatchSize=8192, binSampleCount=50000,
                                 defaultListenPort=12400 + (os.getpid() % 1000) * 12 + case * 4, **kw)
        model = est.fit(LocalDataFrame.from_arrays(Xc, y, num_partitions=2))
        if rank == 0:
            from oracle import oracle as O
            Xa, ya = S.higgs_shape(n)
            if is_cat:
                Xa = with_cats(Xa)
            params = est.get_train_params(world, 2).to_string()
            dsp = "max_bin=%d bin_construct_sample_cnt=50000 min_data_in_leaf=20" % kw["maxBin"] + (" categorical_feature=2,3" if is_cat else "")
            bst, _ = O.fit(Xa, ya, params, dataset_params=dsp, n_shards=world, num_iterations=iters, sample_rows=50000)
            ours, theirs = strip_parameters(model.model_str), strip_parameters(bst.save_model_to_string())
            if quant:
                same = ours == theirs
                kind = " categorical" if is_cat else " wide" if "maxBin" in extra else " constrained" if "monotoneConstraints" in extra else " sampled" if extra else ""
                if is_cat:
                    same = same and "num_cat=0" not in ours.split("Tree=1")[0]      # the first tree has categorical splits
                print("quantized%s dp%d model identical to oracle: %s" % (kind, world, same), flush=True)
                ok &= same
            else:
                from synapseml_b200.lightgbm.model_text import TextModel
                a, b = TextModel(model.model_str), TextModel(bst.save_model_to_string())
                same3 = a.structure()[:3] == b.structure()[:3]
                pa, pb = a.predict_raw(Xa[:50000]), b.predict_raw(Xa[:50000])
                rel = np.abs(pa - pb).max() / np.abs(pb).max()
                print("float dp%d first trees identical: %s, raw score rel diff %.2e" % (world, same3, rel), flush=True)
                ok &= same3 and rel < 1e-5
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, src=0)
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 1 else 1)


if __name__ == "__main__":
    main()
