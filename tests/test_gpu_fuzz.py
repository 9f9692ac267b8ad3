"""-m gpu: a few seeded seconds of the differential fuzzer (tests/fuzz_parity.py) -- random range-proof shapes, secrets,
mutations and kernel paths, random ElGamal vectors with damaged points -- against the oracle through the C ABI."""
import pytest

import fuzz_parity

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('seed', [11, 12])
def test_fuzz_rangeproofs(backend, oracle, seed):
    stats = fuzz_parity.fuzz_rangeproofs(backend, oracle, seed, seconds=60, max_cases=6)
    assert stats['cases'] == 6 and stats['rejected'] > 0 and stats['accepted'] > 0


def test_fuzz_elgamal(oracle):
    from drynx_b200 import Backend
    stats = fuzz_parity.fuzz_elgamal(None, oracle, 21, seconds=60, max_cases=10, make_backend=lambda: Backend(0))
    assert stats['cases'] == 10


def test_fuzz_mixed_batches(backend, oracle):
    stats = fuzz_parity.fuzz_mixed_batches(backend, oracle, 31, seconds=60, max_cases=5)
    assert stats['cases'] == 5 and stats['rejected'] > 0 and stats['accepted'] > 0


def test_fuzz_group_operations(backend, oracle):
    stats = fuzz_parity.fuzz_group_ops(backend, oracle, 41, seconds=60, max_cases=8)
    assert stats['cases'] == 8


def test_fuzz_other_entry_points(oracle):
    from drynx_b200 import Backend
    stats = fuzz_parity.fuzz_misc(None, oracle, 51, seconds=60, max_cases=8, make_backend=lambda: Backend(0))
    assert stats['cases'] == 8
