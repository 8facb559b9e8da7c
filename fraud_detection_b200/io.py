"""Input path: files -> pinned host rows -> device.

The reference reads its rows with ``np.load('data/preprocessed_data.npz')['X_test']``
(explain_model.py:12-13) and ``pd.read_csv(DATA_CSV)`` (train_model.py:22, preprocess.py:15,
columns ``Time, V1..V28, Amount, Class``).  At 25 G rows/s on the device the file reader is the
end-to-end bound, so rows are parsed with pyarrow's multi-threaded CSV reader straight into
page-locked memory (DMA-able, no second host copy) and streamed through the chunked host
pipeline of ``fdb_linear_pass_host``.

    rows = read_rows("data/creditcard.csv")                  # PinnedRows: .array (N, 30) + .labels
    out  = score_file("data/creditcard.csv", scorer)         # proba / phi / mean|SHAP| for the whole file
"""

from __future__ import annotations

import os
from dataclasses import dataclass

import numpy as np

from ._lib import N_FEATURES, PinnedBuffer

FEATURE_NAMES = ["Time"] + [f"V{i}" for i in range(1, 29)] + ["Amount"]


class _HostBuffer:
    """Pageable stand-in with PinnedBuffer's shape (``pinned=False``: parsing without a CUDA driver)."""

    def __init__(self, shape, dtype):
        self.array = np.empty(shape, dtype=dtype)


def _alloc(shape, dtype, pinned: bool):
    return PinnedBuffer(shape, dtype) if pinned else _HostBuffer(shape, dtype)


@dataclass
class PinnedRows:
    buffer: object                    # PinnedBuffer (owns the page-locked memory) or _HostBuffer
    array: np.ndarray                 # (N, 30) view of it
    labels: np.ndarray | None = None  # the Class column when the file has one


def _to_pinned(cols: list[np.ndarray], dtype, pinned: bool = True) -> PinnedRows:
    n = int(cols[0].shape[0])
    buf = _alloc((n, N_FEATURES), dtype, pinned)
    for k, c in enumerate(cols):
        buf.array[:, k] = c           # column-wise scatter into the row-major pinned matrix
    return PinnedRows(buf, buf.array)


def read_csv_rows(path: str, feature_names=None, dtype=np.float32, label: str | None = "Class",
                  pinned: bool = True) -> PinnedRows:
    """CSV with a header row -> pinned (N, 30) matrix in the model's feature order."""
    import pyarrow.csv as pacsv
    names = list(feature_names or FEATURE_NAMES)
    table = pacsv.read_csv(path, read_options=pacsv.ReadOptions(use_threads=True))
    missing = [c for c in names if c not in table.column_names]
    if missing:
        raise ValueError(f"{path}: missing feature columns {missing[:5]}{'...' if len(missing) > 5 else ''}")
    cols = [table.column(c).to_numpy(zero_copy_only=False) for c in names]
    rows = _to_pinned(cols, dtype, pinned)
    if label and label in table.column_names:
        rows.labels = table.column(label).to_numpy(zero_copy_only=False)
    return rows


def read_npz_rows(path: str, key: str = "X_test", dtype=None, label_key: str | None = "y_test",
                  pinned: bool = True) -> PinnedRows:
    """``np.load(path)[key]`` (explain_model.py:12-13) -> pinned (N, 30) matrix."""
    with np.load(path) as data:
        X = data[key]
        if X.ndim != 2 or X.shape[1] != N_FEATURES:
            raise ValueError(f"{path}[{key}] has shape {X.shape}, expected (N, {N_FEATURES})")
        dt = np.dtype(dtype or (np.float32 if X.dtype == np.float32 else np.float64))
        buf = _alloc(X.shape, dt, pinned)
        buf.array[...] = X
        rows = PinnedRows(buf, buf.array)
        if label_key and label_key in data.files:
            rows.labels = data[label_key]
    return rows


def read_rows(path: str, **kw) -> PinnedRows:
    ext = os.path.splitext(path)[1].lower()
    if ext == ".npz":
        return read_npz_rows(path, **kw)
    if ext in (".csv", ".txt", ".gz"):
        return read_csv_rows(path, **kw)
    raise ValueError(f"unsupported file type {ext!r} (use .csv or .npz)")


def score_file(path: str, scorer, phi: bool = True, device=None, **read_kw) -> dict:
    """Read a file and run the fused raw-row pass (``device.FraudScorer.score_host``) over it.

    Returns ``{"rows", "proba", "phi", "mean_abs_shap", "top3", "labels"}``; proba/phi are views of pinned
    buffers owned by the returned dict.
    """
    rows = read_rows(path, **read_kw)
    n = rows.array.shape[0]
    p_buf = PinnedBuffer((n,), rows.array.dtype)
    out = {"proba": p_buf.array}
    keep = [rows.buffer, p_buf]
    if phi:
        phi_buf = PinnedBuffer((n, N_FEATURES), rows.array.dtype)
        out["phi"] = phi_buf.array
        keep.append(phi_buf)
    res = scorer.score_host(rows.array, phi=phi, abs_sum=True, out=out, device=device)
    mean_abs = res["abs_sum"][:N_FEATURES] / max(n, 1)
    return {"rows": n, "proba": res["proba"], "phi": res.get("phi"), "mean_abs_shap": mean_abs,
            "top3": mean_abs.argsort()[-3:][::-1], "labels": rows.labels, "_buffers": keep}
