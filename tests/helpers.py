"""Shared test helpers: move between the product's Morton row order and the oracle's canonical order."""
import numpy as np

from oracle import sparse_ref as S


def lex_rank_of_rows(C):
    """For coordinates C (any row order): order[r] = row index holding canonical rank r."""
    return np.argsort(S.lex_key(C), kind="stable")


def to_canonical(C, *arrays):
    order = lex_rank_of_rows(C)
    return (C[order],) + tuple(a[order] for a in arrays)


def stack_frames(frames):
    """frames -> (coords int32 [N,4], feats float32 [N,4] = (1,r,g,b)) like codec_pipeline.unpack_batch."""
    cs, fs = [], []
    for b, f in enumerate(frames):
        p = np.asarray(f["points"]).astype(np.int32)
        cs.append(np.concatenate([np.full((len(p), 1), b, np.int32), p], axis=1))
        fs.append(np.concatenate([np.ones((len(p), 1)), np.asarray(f["colors"], np.float64)], axis=1))
    return np.concatenate(cs), np.concatenate(fs).astype(np.float32)


def report(name, **values):
    """Append one line of measured parity numbers to gpurun_out/parity_report.jsonl (best effort; the
    summaries quoted in DESIGN.md section 2 are copied from there into profiles/)."""
    import json
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    line = {"test": name}
    for k, v in values.items():
        if isinstance(v, (np.floating, np.integer)):
            v = v.item()
        elif isinstance(v, np.ndarray):
            v = v.tolist()
        line[k] = v
    print("PARITY " + json.dumps(line))
    try:
        os.makedirs(os.path.join(root, "gpurun_out"), exist_ok=True)
        with open(os.path.join(root, "gpurun_out", "parity_report.jsonl"), "a") as f:
            f.write(json.dumps(line) + "\n")
    except OSError:
        pass
