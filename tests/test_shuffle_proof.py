"""The shuffle proof of SURVEY 8(f)-2 (libunlynxshuffle.ShuffleProofCreation / ShuffleProofVerification over kyber's
shuffle.PairShuffle, verified from lib/proof/structs_proofs.go:389-420): tests/shuffle_proof.py restates Neff's argument
over rep_commit / rep_verify. CPU: the oracle engine accepts an honest proof and refuses every damaged part, a wrong
permutation, a wrong re-randomisation. GPU: the same proof through drx_rep_commit / drx_rep_verify is byte-identical
and gets the same verdicts. kyber's own test (shuffle/shuffle_test.go as recalled: shuffle k pairs, prove, verify) is the
first case."""
import pytest

import bn256_ref as R
import rep_proofs as RP
import shuffle_proof as SP
from conftest import SplitMix


def _pt(k):
    return R.g1_to_bytes(R.g1_mul(k, R.G1_GEN))


def _case(rng, k):
    g = R.g1_to_bytes(R.G1_GEN)
    h = _pt(rng.scalar())                                   # the collective public key
    X = [_pt(rng.scalar()) for _ in range(k)]               # compressed lines: (K, C) of one ciphertext each
    Y = [_pt(rng.scalar()) for _ in range(k)]
    beta = [rng.scalar() for _ in range(k)]
    pi = list(range(k))
    for i in range(k - 1, 0, -1):                           # Fisher-Yates over the test's generator
        j = rng.randint(0, i)
        pi[i], pi[j] = pi[j], pi[i]
    secrets = dict(u=[rng.scalar() for _ in range(k)], w=[rng.scalar() for _ in range(k)], a=[rng.scalar() for _ in range(k)],
                   tau0=rng.scalar(), nu=rng.scalar(), gamma=rng.scalar(), theta=[rng.scalar() for _ in range(2 * k - 1)])
    return g, h, X, Y, beta, pi, secrets


def _run(engine, rng, k):
    g, h, X, Y, beta, pi, secrets = _case(rng, k)
    Xbar, Ybar = SP.shuffle_pairs(engine, pi, g, h, beta, X, Y)
    proof = SP.pair_shuffle_prove(engine, pi, g, h, beta, X, Y, secrets)
    return g, h, X, Y, Xbar, Ybar, proof, (beta, pi, secrets)


def _verdicts(engine, g, h, X, Y, Xbar, Ybar, proof, k):
    """Verdicts on the honest proof and on a list of damaged variants (same list for every engine)."""
    out = [SP.pair_shuffle_verify(engine, g, h, X, Y, Xbar, Ybar, proof),
           SP.pair_shuffle_verify(engine, g, h, X, Y, Xbar, Ybar, proof, link_simple_shuffle=True)]
    # one flipped bit in every message of the transcript: Gamma, A, C, U, W, Lambda1, Lambda2, D, sigma, tau,
    # the simple shuffle's X, Y, Theta, alpha (scalars: low bit of the last byte; points: replaced by another point)
    other = _pt(12345)
    offs = {'Gamma': 0, 'A0': 64, 'C1': 64 * (1 + k + 1), 'U0': 64 * (1 + 2 * k), 'W1': 64 * (1 + 3 * k + 1),
            'Lambda1': 64 * (4 * k + 1), 'Lambda2': 64 * (4 * k + 2), 'D0': 64 * (4 * k + 3)}
    for name, o in offs.items():
        bad = bytearray(proof)
        bad[o:o + 64] = other
        out.append(SP.pair_shuffle_verify(engine, g, h, X, Y, Xbar, Ybar, bytes(bad)))
    s0 = 64 * (4 * k + 3) + 64 * k
    for o in (s0 + 31, s0 + 32 * k + 31,                                  # sigma_0, tau
              s0 + 32 * (k + 1) + 128 * k + 128 * k + 31,                  # alpha_0
              len(proof) - 1):                                             # alpha_last
        bad = bytearray(proof)
        bad[o] ^= 1
        out.append(SP.pair_shuffle_verify(engine, g, h, X, Y, Xbar, Ybar, bytes(bad)))
    bad = bytearray(proof)
    o = s0 + 32 * (k + 1) + 128 * k                                        # Theta_0
    bad[o:o + 64] = other
    out.append(SP.pair_shuffle_verify(engine, g, h, X, Y, Xbar, Ybar, bytes(bad)))
    bad = bytearray(proof)
    bad[5] ^= 1                                                            # Gamma leaves the curve
    out.append(SP.pair_shuffle_verify(engine, g, h, X, Y, Xbar, Ybar, bytes(bad)))
    out.append(SP.pair_shuffle_verify(engine, g, h, X, Y, Xbar, Ybar, proof[:-32]))      # truncated
    # a proof for other data: two outputs swapped, one output not re-randomised as claimed, another public key
    Xs, Ys = list(Xbar), list(Ybar)
    Xs[0], Xs[1] = Xs[1], Xs[0]
    out.append(SP.pair_shuffle_verify(engine, g, h, X, Y, Xs, Ybar, proof))
    Ys[1] = other
    out.append(SP.pair_shuffle_verify(engine, g, h, X, Y, Xbar, Ys, proof))
    out.append(SP.pair_shuffle_verify(engine, g, other, X, Y, Xbar, Ybar, proof))
    Xo = list(X)
    Xo[k - 1] = other
    out.append(SP.pair_shuffle_verify(engine, g, h, Xo, Y, Xbar, Ybar, proof))
    return out


EXPECT = [True, True] + [False] * 19


@pytest.mark.parametrize('k', [2, 5])
def test_oracle_pair_shuffle_round_trip_and_damage(oracle, k):
    rng = SplitMix(700 + k)
    eng = RP.OracleEngine(oracle, threads=4)
    g, h, X, Y, Xbar, Ybar, proof, (beta, pi, secrets) = _run(eng, rng, k)
    assert len(proof) == SP.proof_size(k)
    # the shuffle is what it claims: output i is input pi[i] plus beta[pi[i]] * (g, h)
    for i in range(k):
        want = R.g1_add(R.g1_from_bytes(X[pi[i]]), R.g1_mul(beta[pi[i]], R.G1_GEN))
        assert Xbar[i] == R.g1_to_bytes(want)
    assert _verdicts(eng, g, h, X, Y, Xbar, Ybar, proof, k) == EXPECT
    # a prover that lies about the permutation it used
    pi2 = list(pi)
    pi2[0], pi2[1] = pi2[1], pi2[0]
    lie = SP.pair_shuffle_prove(eng, pi2, g, h, beta, X, Y, secrets)
    assert not SP.pair_shuffle_verify(eng, g, h, X, Y, Xbar, Ybar, lie)
    # ... or about one re-randomisation value
    beta2 = list(beta)
    beta2[pi[0]] = (beta2[pi[0]] + 1) % R.n
    lie = SP.pair_shuffle_prove(eng, pi, g, h, beta2, X, Y, secrets)
    assert not SP.pair_shuffle_verify(eng, g, h, X, Y, Xbar, Ybar, lie)


def test_identity_shuffle_and_zero_beta(oracle):
    """pi = identity with beta = 0 (outputs equal inputs) is a valid shuffle; so is any pi with beta = 0."""
    rng = SplitMix(77)
    eng = RP.OracleEngine(oracle, threads=4)
    k = 3
    g, h, X, Y, beta, pi, secrets = _case(rng, k)
    for perm in (list(range(k)), pi):
        zero = [0] * k
        Xbar, Ybar = SP.shuffle_pairs(eng, perm, g, h, zero, X, Y)
        assert Xbar == [X[perm[i]] for i in range(k)]
        proof = SP.pair_shuffle_prove(eng, perm, g, h, zero, X, Y, secrets)
        assert SP.pair_shuffle_verify(eng, g, h, X, Y, Xbar, Ybar, proof)


@pytest.mark.gpu
@pytest.mark.parametrize('k', [2, 7, 64])
def test_gpu_pair_shuffle_matches_oracle(backend, oracle, k):
    rng = SplitMix(900 + k)
    cpu, gpu = RP.OracleEngine(oracle, threads=8), RP.BackendEngine(backend)
    g, h, X, Y, Xbar, Ybar, proof, _ = _run(gpu, SplitMix(900 + k), k)
    g2, h2, X2, Y2, Xbar2, Ybar2, proof2, _ = _run(cpu, SplitMix(900 + k), k)
    assert (Xbar, Ybar) == (Xbar2, Ybar2) and proof == proof2          # every group element and scalar of the proof
    got = _verdicts(gpu, g, h, X, Y, Xbar, Ybar, proof, k)
    assert got == EXPECT
    if k <= 7:
        assert got == _verdicts(cpu, g, h, X, Y, Xbar, Ybar, proof, k)
