"""File readers of the input path (CPU only, pageable buffers): same rows as pandas / np.load."""

import numpy as np
import pandas as pd
import pytest

from fraud_detection_b200 import io as fio
from oracle import synth


def test_csv_reader_matches_pandas(tmp_path):
    X = synth.reference_rows(1000)
    df = pd.DataFrame(np.column_stack([X, (np.arange(1000) % 7 == 0).astype(float)]), columns=synth.FEATURE_NAMES + ["Class"])
    df = df[["Class"] + synth.FEATURE_NAMES[::-1]]                    # any column order in the file
    path = tmp_path / "creditcard.csv"
    df.to_csv(path, index=False)
    rows = fio.read_rows(str(path), dtype=np.float64, pinned=False)
    ref = pd.read_csv(path)[synth.FEATURE_NAMES].to_numpy()
    assert rows.array.shape == (1000, 30) and rows.array.flags["C_CONTIGUOUS"]
    np.testing.assert_array_equal(rows.array, X)                       # correctly rounded parse: exact round trip
    np.testing.assert_allclose(rows.array, ref, rtol=1e-12)            # pandas' default fast parser is off by an ulp
    np.testing.assert_array_equal(rows.labels, df["Class"].to_numpy())
    r32 = fio.read_csv_rows(str(path), pinned=False)
    assert r32.array.dtype == np.float32
    np.testing.assert_array_equal(r32.array, X.astype(np.float32))


def test_csv_reader_rejects_missing_columns(tmp_path):
    path = tmp_path / "approval.csv"                                   # like the reference's shipped data/credit_card.csv
    pd.DataFrame({"Gender": [1, 0], "Age": [30.0, 41.5]}).to_csv(path, index=False)
    with pytest.raises(ValueError, match="missing feature columns"):
        fio.read_csv_rows(str(path), pinned=False)
    with pytest.raises(ValueError, match="unsupported file type"):
        fio.read_rows(str(tmp_path / "x.parquet"))


def test_npz_reader(tmp_path):
    X = synth.rows(64, seed=1)
    path = tmp_path / "preprocessed_data.npz"
    np.savez(path, X_test=X, y_test=np.arange(64) % 2)
    rows = fio.read_rows(str(path), pinned=False)
    np.testing.assert_array_equal(rows.array, X)
    assert rows.labels.shape == (64,)
    np.savez(path, X_test=X[:, :10])
    with pytest.raises(ValueError, match="expected"):
        fio.read_npz_rows(str(path), pinned=False)
