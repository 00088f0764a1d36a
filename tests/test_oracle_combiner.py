"""Combiner oracle against AverageCombinerTest.java's known answers and the MnistCombiner fixture."""
import os

import numpy as np
import pytest

from oracle.combiner import CombinerError, mean_combiner, standardise

from conftest import GOLDEN


def test_average_combiner_kats():
    # AverageCombinerTest.java:41-75 / 77-111
    out = mean_combiner([np.array([[1.0, 1.0]]), np.array([[1.0, 0.5]]), np.array([[2.2, 0.9]])])
    assert abs(out[0, 0] - (1.0 + 1.0 + 2.2) / 3) < 1e-7 and abs(out[0, 1] - (1.0 + 0.5 + 0.9) / 3) < 1e-7
    # L113-141 two 1x1
    assert mean_combiner([np.array([[1.0]]), np.array([[1.0]])])[0, 0] == 1.0
    # L143-165 single child is the identity
    assert np.array_equal(mean_combiner([np.array([[1.0, 5.0, 0.3]])]), np.array([[1.0, 5.0, 0.3]]))


def test_average_combiner_errors():
    with pytest.raises(CombinerError):  # L167-175
        mean_combiner([])
    with pytest.raises(CombinerError):  # L177-193 no values -> no 2-D shape
        mean_combiner([np.array([]), np.array([])])
    with pytest.raises(CombinerError):  # L195-225
        mean_combiner([np.zeros((1, 2)), np.zeros((1, 2)), np.zeros((1, 3))])


def test_mnist_combiner_fixture():
    g = np.load(os.path.join(GOLDEN, "combiner_mnist.npz"))
    assert np.array_equal(mean_combiner([g["a"], g["b"]]), g["out"])


def test_standardise():
    X = np.array([[1.0, 2.0], [3.0, 4.0]])
    out = standardise(X, np.array([1.0, 2.0]), np.array([2.0, 0.0]))
    assert np.allclose(out[:, 0], [0.0, 1.0]) and out[1, 1] == 2.0 / 1e-10
