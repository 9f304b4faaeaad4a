"""The PCM16 restatement (oracle/pcm16.py) against outputs of the reference's own convert_to_16_bit_wav
(tests/golden/pcm16_*.npz, made by oracle/gen_golden_pcm16.py).  CPU only; bit-exact."""
import glob
import os

import numpy as np
import pytest

from conftest import GOLDEN, load_golden
from oracle.pcm16 import make_wave, pcm16_row, pcm16_rows

CASES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "pcm16_*.npz")))


def test_fixtures_present():
    assert len(CASES) >= 5


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_output(name):
    z, meta = load_golden(name)
    x = make_wave(meta["B"], meta["L"], meta["seed"], meta["scale"], meta["bf16"])
    got, peaks = pcm16_rows(x)
    assert got.dtype == np.int16 and got.shape == z["pcm"].shape
    assert np.array_equal(got, z["pcm"])
    assert np.array_equal(peaks, np.abs(x).max(axis=1))


def test_branches_and_truncation():
    # peak == 1 exactly: no normalisation (the reference tests "> 1.0", main.py:2973); truncation is toward zero
    x = np.array([1.0, -1.0, 0.5 / 32767, -0.5 / 32767, 1.5 / 32767, -1.5 / 32767, 0.0], np.float32)
    assert pcm16_row(x).tolist() == [32767, -32767, 0, 0, 1, -1, 0]
    # peak > 1: everything is divided by the peak first
    y = np.array([2.0, -1.0, 0.5], np.float32)
    assert pcm16_row(y).tolist() == [32767, -16383, 8191]
    assert pcm16_row(np.zeros((0,), np.float32)).shape == (0,)
