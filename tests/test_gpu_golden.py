"""GPU tier: the CUDA path against the committed golden vectors of the unmodified reference."""
from pathlib import Path

import numpy as np
import pytest

import golden_cases
from helpers import assert_parity

pytestmark = pytest.mark.gpu
GOLDEN = np.load(Path(__file__).parent / "golden" / "golden_v1.npz")


@pytest.mark.parametrize("name", sorted(golden_cases.CASES))
def test_product_matches_golden(cuda, product, name):
    got = golden_cases.CASES[name](product)
    assert_parity(got, GOLDEN[name], what=name)


GOLDEN_NEXT = np.load(Path(__file__).parent / "golden" / "golden_v2.npz")


@pytest.mark.parametrize("name", sorted(golden_cases.NEXT_CASES))
def test_product_matches_golden_next_rows(cuda, product, name):
    got = golden_cases.NEXT_CASES[name](product)
    assert_parity(got, GOLDEN_NEXT[name], what=name)
