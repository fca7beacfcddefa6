"""Oracle (test infrastructure): CompressAI 1.2.4 entropy-model arithmetic used by the codec,
restated in numpy (compressai is unvendored: sender/encoder/requirements.txt:10).
Call sites followed: codec_pipeline.py:305-306, 397-437; codec_parallel.py:291-318, 382-419.
"""
import math

import numpy as np
from scipy.special import erfc, ndtri

PRECISION = 16
SCALE_MIN, SCALE_MAX, SCALE_LEVELS = 0.11, 256.0, 64
SCALE_BOUND = 0.11
TAIL_MASS = 1e-9


# ---- _CXX.pmf_to_quantized_cdf ------------------------------------------------------------------------
def pmf_to_quantized_cdf(pmf, precision=PRECISION):
    pmf = np.asarray(pmf, dtype=np.float32)
    scale = np.float32(1 << precision)
    cdf = [0]
    for p in pmf:
        v = float(np.float32(p) * scale)
        cdf.append(int(math.floor(v + 0.5)) if v >= 0 else -int(math.floor(-v + 0.5)))  # std::round
    total = sum(cdf)
    if total == 0:
        raise ValueError("zero pmf")
    cdf = [((1 << precision) * c) // total for c in cdf]
    for i in range(1, len(cdf)):
        cdf[i] += cdf[i - 1]
    cdf[-1] = 1 << precision
    n = len(cdf) - 1
    for i in range(n):
        if cdf[i] == cdf[i + 1]:
            best_freq, best_steal = None, -1
            for j in range(n):
                f = cdf[j + 1] - cdf[j]
                if f > 1 and (best_freq is None or f < best_freq):
                    best_freq, best_steal = f, j
            if best_steal < 0:
                raise ValueError("cannot quantise pmf")
            if best_steal < i:
                for j in range(best_steal + 1, i + 1):
                    cdf[j] -= 1
            else:
                for j in range(i + 1, best_steal + 1):
                    cdf[j] += 1
    return np.array(cdf, dtype=np.int32)


def _phi(x):
    return 0.5 * erfc(-x / math.sqrt(2.0))


# ---- GaussianConditional ---------------------------------------------------------------------------------
def scale_table():
    t = np.exp(np.linspace(math.log(SCALE_MIN), math.log(SCALE_MAX), SCALE_LEVELS))
    return t.astype(np.float32)


def gaussian_pmf(table=None, dtype=np.float32):
    """The pmf rows (tail entry last) GaussianConditional.update() quantises, and pmf_center.

    CompressAI 1.2.4 evaluates this on float32 tensors (the scale table is a float32 tensor; `samples.float()`,
    `samples_scale.float()`, torch.erfc in float32); only the quantile multiplier is a Python double.  dtype =
    np.float32 restates exactly that with numpy float32 arithmetic and scipy's erfc (an independent erfc: the
    product uses torch's); dtype = np.float64 is the "mathematically exact" variant, kept to MEASURE what the
    float32 convention changes (tests/test_cpu_oracle.py::test_gaussian_tables_float32_vs_float64)."""
    table = scale_table() if table is None else np.asarray(table, dtype=np.float32)
    t = table.astype(dtype)
    multiplier = dtype(-ndtri(TAIL_MASS / 2))
    pmf_center = np.ceil(t * multiplier).astype(np.int64)
    pmf_length = 2 * pmf_center + 1
    half, const = dtype(0.5), dtype(-(2 ** -0.5))
    rows = []
    for i, s in enumerate(t):
        samples = np.abs(np.arange(pmf_length[i]) - pmf_center[i]).astype(dtype)
        upper = half * erfc(const * ((half - samples) / s))
        lower = half * erfc(const * ((-half - samples) / s))
        assert upper.dtype == dtype
        pmf = upper - lower
        tail = dtype(2) * lower[:1]
        rows.append(np.concatenate([pmf, tail]).astype(np.float32))
    return rows, pmf_center


def gaussian_tables(table=None, dtype=np.float32):
    """GaussianConditional.update(): list of cdf rows (ragged), cdf_length, offset."""
    prows, pmf_center = gaussian_pmf(table, dtype)
    rows = [pmf_to_quantized_cdf(r) for r in prows]
    return rows, (2 * pmf_center + 3).astype(np.int32), (-pmf_center).astype(np.int32)


def lower_bound_scale(s):
    return np.maximum(np.asarray(s, dtype=np.float32), np.float32(SCALE_BOUND))


def build_indexes(scales, table=None):
    """GaussianConditional.build_indexes: idx = L-1 - #(s <= table[:-1]) on the lower-bounded scale."""
    table = scale_table() if table is None else np.asarray(table, dtype=np.float32)
    s = lower_bound_scale(scales)
    idx = np.full(s.shape, len(table) - 1, dtype=np.int32)
    for t in table[:-1]:
        idx -= (s <= t).astype(np.int32)
    return idx


def quantize_symbols(x, means):
    """torch.round(x - means).int(): fp32 subtract, round half to even."""
    d = np.asarray(x, dtype=np.float32) - np.asarray(means, dtype=np.float32)
    r = np.rint(d)
    r = np.clip(r, -(2.0 ** 30), 2.0 ** 30)
    return r.astype(np.int32)


def gaussian_encode_symbols(y, scales_hat, means_hat, scale, table=None):
    """codec_pipeline.py:425-430 for one quality: y, scales_hat, means_hat [C, N]; scale [C]."""
    s = np.asarray(scale, dtype=np.float32).reshape(-1, 1)
    idx = build_indexes(np.asarray(scales_hat, np.float32) * s, table)
    sym = quantize_symbols(np.asarray(y, np.float32) * s, np.asarray(means_hat, np.float32) * s)
    return sym, idx


def bin_offset(sd):
    """get_offsets core: centroid shift toward zero of a unit bin under a Laplacian of std `sd`
    (the real `entropy_model.get_offsets` is in the missing submodule; this is this repo's own
    definition)."""
    b = np.asarray(sd, dtype=np.float32) * np.float32(0.70710678118654752)
    with np.errstate(over="ignore", divide="ignore", invalid="ignore"):
        ib = np.float32(1.0) / b
        far = ib * np.float32(1.0 / 12.0) - ib * ib * ib * np.float32(1.0 / 720.0)
        near = np.float32(0.5) / np.tanh(np.float32(0.5) / b) - b
    d = np.where(b > 4.0, far, near).astype(np.float32)
    return np.clip(d, 0.0, 0.5).astype(np.float32)


def gaussian_dequant(sym, scales_hat, means_hat, scale, offset_gain=None):
    """codec_parallel.py:394-409: sym, scales_hat, means_hat [C, N]; scale [C] -> y_hat [C, N]."""
    s = np.asarray(scale, dtype=np.float32).reshape(-1, 1)
    rescale = np.float32(1.0) / s
    q = np.asarray(sym).astype(np.float32)
    q_abs, signs = np.abs(q), np.sign(q)
    sd = lower_bound_scale(np.asarray(scales_hat, np.float32) * s)
    g = np.ones_like(s) if offset_gain is None else np.asarray(offset_gain, np.float32).reshape(-1, 1)
    q_off = (-1.0 * (g * bin_offset(sd))).astype(np.float32)
    q_off[q_abs < 0.0001] = 0
    y = signs * (q_abs + q_off)
    return (y * rescale + np.asarray(means_hat, np.float32)).astype(np.float32)


def scale_nn(q, w1, b1, w2, b2):
    """entropy_model.scale_nn (own definition): softplus(W2 relu(W1 q + b1) + b2), evaluated as
    sequential fp32 accumulation so encoder and decoder agree bit for bit on any host."""
    q = np.asarray(q, dtype=np.float32).reshape(-1)
    h = np.asarray(b1, np.float32).copy()
    for k in range(q.shape[0]):
        h = h + np.asarray(w1, np.float32)[k] * q[k]
    h = np.maximum(h, np.float32(0))
    o = np.asarray(b2, np.float32).copy()
    for k in range(h.shape[0]):
        o = o + np.asarray(w2, np.float32)[k] * h[k]
    # softplus in fp32
    return (np.maximum(o, 0) + np.log1p(np.exp(-np.abs(o)))).astype(np.float32)


# ---- EntropyBottleneck --------------------------------------------------------------------------------------
def _softplus(x):
    return np.maximum(x, 0) + np.log1p(np.exp(-np.abs(x)))


def _sigmoid(v):
    with np.errstate(over="ignore"):
        return (1 / (1 + np.exp(-v))).astype(v.dtype)


def eb_logits_cumulative(x, matrices, biases, factors, dtype=np.float32):
    """x [C,1,N]; EntropyBottleneck._logits_cumulative in `dtype` arithmetic (CompressAI: float32)."""
    logits = x.astype(dtype)
    for i in range(len(matrices)):
        logits = np.matmul(_softplus(matrices[i].astype(dtype)), logits) + biases[i].astype(dtype)
        if i < len(factors):
            logits = logits + np.tanh(factors[i].astype(dtype)) * np.tanh(logits)
    return logits


def eb_pmf(matrices, biases, factors, quantiles, dtype=np.float32):
    """The pmf rows (tail entry last) EntropyBottleneck.update(force=True) quantises.  float32 like CompressAI
    (parameters and samples are float32 tensors); float64 kept as the measuring stick."""
    q = np.asarray(quantiles, dtype=np.float32).astype(dtype)
    medians = q[:, 0, 1]
    minima = np.maximum(np.ceil(medians - q[:, 0, 0]), 0).astype(np.int64)
    maxima = np.maximum(np.ceil(q[:, 0, 2] - medians), 0).astype(np.int64)
    pmf_start = medians - minima.astype(dtype)
    pmf_length = maxima + minima + 1
    max_length = int(pmf_length.max())
    samples = np.arange(max_length).astype(dtype)[None, None, :] + pmf_start[:, None, None]
    half = dtype(0.5)
    lower = eb_logits_cumulative(samples - half, matrices, biases, factors, dtype)
    upper = eb_logits_cumulative(samples + half, matrices, biases, factors, dtype)
    sign = -np.sign(lower + upper)
    pmf = np.abs(_sigmoid(sign * upper) - _sigmoid(sign * lower))[:, 0, :]
    tail = _sigmoid(lower[:, 0, :1]) + _sigmoid(-upper[:, 0, -1:])
    rows = [np.concatenate([pmf[c, :int(pmf_length[c])], tail[c]]).astype(np.float32) for c in range(pmf.shape[0])]
    return rows, minima, pmf_length, medians.astype(np.float32)


def eb_tables(matrices, biases, factors, quantiles, dtype=np.float32):
    """EntropyBottleneck.update(force=True) -> (cdf rows, cdf_length, offset, medians)."""
    prows, minima, pmf_length, medians = eb_pmf(matrices, biases, factors, quantiles, dtype)
    rows = [pmf_to_quantized_cdf(r) for r in prows]
    return rows, (pmf_length + 2).astype(np.int32), (-minima).astype(np.int32), medians


def factorized_quantize(z, medians):
    """EntropyBottleneck.compress symbols + decompress values: z [C, N]."""
    m = np.asarray(medians, np.float32).reshape(-1, 1)
    sym = quantize_symbols(z, m)
    return sym, (sym.astype(np.float32) + m).astype(np.float32)
