"""The C++ host mirror (host/drynx_host.hpp) run through its reference-style test program (host/test_host.cpp)."""
import subprocess

import pytest


def _binary():
    from drynx_b200 import build
    return build.build_host_test()


def test_plaintext_encoders_match_reference_kats():
    """CPU: logistic_regression_test.go:20-72 and :544-558 known answers, ToBase, vector layouts."""
    p = subprocess.run([_binary(), '--cpu'], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0 and 'HOST_TESTS_OK' in p.stdout, p.stdout + p.stderr


@pytest.mark.gpu
def test_reference_style_round_trips_on_gpu():
    """sum / mean / variance / frequency count (with proofs) / lin-reg / log-reg round trips, as in lib/encoding/*_test.go."""
    p = subprocess.run([_binary()], capture_output=True, text=True, timeout=600)
    assert p.returncode == 0 and 'HOST_TESTS_OK' in p.stdout, p.stdout + p.stderr
