"""Synthetic inputs of the BASELINE.json shapes; float64 features, float32 labels (the cast the
reference applies at lightgbm/.../StreamingPartitionTask.scala:332).

Rows are generated in fixed 1M-row chunks keyed by (seed, chunk index), so the first N rows are the same whatever the
total size and a rank can generate just its own row shard.
"""
import numpy as np

CHUNK = 1 << 20


def _chunks(n, start=0):
    c0, c1 = start // CHUNK, (start + n + CHUNK - 1) // CHUNK
    for c in range(c0, c1):
        lo, hi = max(start, c * CHUNK), min(start + n, (c + 1) * CHUNK)
        yield c, lo - c * CHUNK, hi - c * CHUNK


def _assemble(gen_chunk, n, start):
    # chunks are independent streams: generate them on a thread pool (numpy's generators release the GIL)
    from concurrent.futures import ThreadPoolExecutor
    import os
    spans = list(_chunks(n, start))
    f = None
    x_out = y_out = None
    pos = [0]
    for _, lo, hi in spans:
        pos.append(pos[-1] + hi - lo)

    def work(i):
        c, lo, hi = spans[i]
        x, y = gen_chunk(c)
        return i, x[lo:hi], y[lo:hi]

    with ThreadPoolExecutor(max_workers=min(len(spans), max(1, (os.cpu_count() or 2) - 1))) as ex:
        for i, x, y in ex.map(work, range(len(spans))):
            if x_out is None:
                x_out = np.empty((n, x.shape[1]), dtype=x.dtype)
                y_out = np.empty(n, dtype=y.dtype)
            x_out[pos[i]:pos[i + 1]] = x
            y_out[pos[i]:pos[i + 1]] = y
    del f
    return x_out, y_out


def higgs_shape(n, start=0, seed=20240601, f=28):
    """Higgs-shape (C1/C2): N(0,1) / Exp(1) / U(-pi,pi) / 3-valued columns; Bernoulli labels."""
    def gen(c):
        rng = np.random.Generator(np.random.PCG64([seed, c]))
        x = np.empty((CHUNK, f), dtype=np.float64)
        for j in range(f):
            k = j % 4
            if k == 0:
                x[:, j] = rng.standard_normal(CHUNK)
            elif k == 1:
                x[:, j] = rng.exponential(1.0, CHUNK)
            elif k == 2:
                x[:, j] = rng.uniform(-np.pi, np.pi, CHUNK)
            else:
                u = rng.random(CHUNK)
                x[:, j] = np.where(u < 0.5, 0.0, np.where(u < 0.8, 1.0865, 2.1731))
        z = 0.8 * x[:, 0] - 0.6 * x[:, 1 % f] + 0.5 * np.sin(x[:, 2 % f]) + 0.7 * (x[:, 3 % f] > 1) + 0.3 * x[:, 4 % f] * x[:, 8 % f] - 0.2
        y = (rng.random(CHUNK) < 1.0 / (1.0 + np.exp(-z))).astype(np.float32)
        return x, y
    return _assemble(gen, n, start)


def dense100(n, start=0, seed=7, f=100):
    """Dense-100 (C3): x_j ~ N(0,(1+j/50)^2); y = sum_{j<10} x_j/(1+j) + 2 sin(x_10) + N(0,0.5^2)."""
    def gen(c):
        rng = np.random.Generator(np.random.PCG64([seed, c]))
        x = rng.standard_normal((CHUNK, f)) * (1.0 + np.arange(f) / 50.0)
        y = sum(x[:, j] / (1.0 + j) for j in range(min(10, f))) + 2.0 * np.sin(x[:, 10 % f]) + 0.5 * rng.standard_normal(CHUNK)
        return x, y.astype(np.float32)
    return _assemble(gen, n, start)


def epsilon_shape(n, start=0, seed=13, f=2000):
    """Epsilon-shape (C5): unit-norm gaussian rows; label = [w.x + 0.02 N(0,1) > 0]."""
    w = np.random.Generator(np.random.PCG64([seed, 1 << 30])).standard_normal(f)
    small = 1 << 16

    def gen_small(c):
        rng = np.random.Generator(np.random.PCG64([seed, c]))
        x = rng.standard_normal((small, f))
        x /= np.linalg.norm(x, axis=1, keepdims=True)
        y = (x @ w + 0.02 * rng.standard_normal(small) > 0).astype(np.float32)
        return x, y
    xs, ys = [], []
    c0, c1 = start // small, (start + n + small - 1) // small
    for c in range(c0, c1):
        lo, hi = max(start, c * small) - c * small, min(start + n, (c + 1) * small) - c * small
        x, y = gen_small(c)
        xs.append(x[lo:hi])
        ys.append(y[lo:hi])
    return np.concatenate(xs), np.concatenate(ys)


def mslr_shape(n_target, seed=11, f=136):
    """MSLR-shape (C4): lognormal query sizes clipped to [1,1251], rows contiguous per query; returns (X, relevance, query_id).
    40% zero-heavy Exp(1) columns, 30% Poisson(3), 30% N(0,1); relevance in {0..4}."""
    rng = np.random.Generator(np.random.PCG64([seed, 0]))
    sizes = []
    tot = 0
    while tot < n_target:
        s = int(np.clip(np.round(rng.lognormal(4.6, 0.6)), 1, 1251))
        s = min(s, n_target - tot)
        sizes.append(s)
        tot += s
    n = tot
    qid = np.repeat(np.arange(len(sizes), dtype=np.int32), sizes)
    x = np.empty((n, f), dtype=np.float64)
    for j in range(f):
        k = j % 10
        if k < 4:
            col = rng.exponential(1.0, n)
            col[rng.random(n) < 0.6] = 0.0
            x[:, j] = col
        elif k < 7:
            x[:, j] = rng.poisson(3.0, n)
        else:
            x[:, j] = rng.standard_normal(n)
    latent = 0.9 + 0.5 * x[:, 0] + 0.4 * x[:, 50 % f] + 0.8 * rng.standard_normal(n)
    rel = np.clip(np.floor(latent), 0, 4).astype(np.float32)
    return x, rel, qid
