"""Synthetic generators of BASELINE.json's configs (SURVEY.md 8(d) "Synthetic inputs" 2-4), shared by bench.py, the
at-size parity tests and the record scripts under tools/.  All float32 row-major, numpy default_rng(seed)."""
import numpy as np

HIGGS_PARAMS = ("objective=binary max_bin=255 num_leaves=255 learning_rate=0.1 min_data_in_leaf=0 "
                "min_sum_hessian_in_leaf=100 verbosity=-1")
EPSILON_PARAMS = HIGGS_PARAMS
MSLR_PARAMS = ("objective=lambdarank max_bin=255 num_leaves=255 learning_rate=0.1 min_data_in_leaf=0 "
               "min_sum_hessian_in_leaf=100 verbosity=-1")


def higgs_shaped(n, seed=0, F=28):
    """config #2: 21 gaussian columns, 4 half-zero columns (max(0, N(0,1))), 3 three-level columns; label =
    1[x.w + 0.5 x0 x1 + eps > 0]."""
    rng = np.random.default_rng(seed)
    X = np.empty((n, F), dtype=np.float32)
    chunk = 1 << 20
    w = np.random.default_rng(seed + 1).standard_normal(F).astype(np.float32)
    y = np.empty(n, dtype=np.float32)
    levels = np.array([0.0, 1.087, 2.173], dtype=np.float32)
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        b = rng.standard_normal((e - s, F), dtype=np.float32)
        b[:, 21:25] = np.maximum(0, b[:, 21:25])
        b[:, 25:28] = levels[rng.integers(0, 3, size=(e - s, 3))]
        X[s:e] = b
        y[s:e] = ((b @ w + 0.5 * b[:, 0] * b[:, 1] + rng.standard_normal(e - s, dtype=np.float32)) > 0)
    return X, y


def epsilon_shaped(n=400_000, F=2000, seed=2):
    """config #3: rows = unit-normalised N(0,1) vectors, label = sign of a random hyperplane + noise."""
    rng = np.random.default_rng(seed)
    X = np.empty((n, F), dtype=np.float32)
    w = np.random.default_rng(seed + 1).standard_normal(F).astype(np.float32)
    y = np.empty(n, dtype=np.float32)
    chunk = 1 << 15
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        b = rng.standard_normal((e - s, F), dtype=np.float32)
        b /= np.linalg.norm(b, axis=1, keepdims=True)
        X[s:e] = b
        y[s:e] = ((b @ w + 0.5 * rng.standard_normal(e - s, dtype=np.float32)) > 0)
    return X, y


def mslr_shaped(n=2_270_000, F=137, seed=3):
    """config #4: query sizes ~ clip(Poisson(120), 1, 1000) summing to n; features a mix of N(0,1) and integer counts;
    labels 0-4 with P ~ {.52, .32, .13, .02, .01}, correlated with a latent score.  Returns X, y, query sizes."""
    rng = np.random.default_rng(seed)
    sizes = []
    tot = 0
    while tot < n:
        s = int(np.clip(rng.poisson(120), 1, 1000))
        s = min(s, n - tot)
        sizes.append(s); tot += s
    sizes = np.asarray(sizes, dtype=np.int32)
    X = np.empty((n, F), dtype=np.float32)
    latent = np.zeros(n, dtype=np.float32)
    w = np.random.default_rng(seed + 1).standard_normal(F).astype(np.float32) / np.sqrt(F)
    chunk = 1 << 18
    half = F // 2
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        b = rng.standard_normal((e - s, F), dtype=np.float32)
        b[:, half:] = np.round(np.abs(b[:, half:]) * 3.0)   # integer counts
        X[s:e] = b
        latent[s:e] = b @ w + 0.7 * rng.standard_normal(e - s, dtype=np.float32)
    cuts = np.quantile(latent[:min(n, 1 << 20)], [0.52, 0.84, 0.97, 0.99])
    y = np.digitize(latent, cuts).astype(np.float32)
    return X, y, sizes
