"""The pin to the reference's own bytes. go/vectors/dump_test.go, run once inside a checkout of github.com/ldsec/drynx
on a machine with Go and the module cache, writes tests/golden/kyber_vectors.json: what kyber v3.0.8 pairing/bn256,
unlynx v1.4.0 and drynx lib/range produce for fixed inputs. While that file is absent every test here SKIPS with the
reason "parity unpinned" (the BN256 oracle is then only pinned to its own mathematical definition);
once it is present the tests are strict: the Python definition, the C++ oracle and (-m gpu) the CUDA path must all
reproduce every vector byte for byte, which settles the open conventions (generator signs, G2 generator, GT coefficient
order, G2 coordinate order, what UnmarshalBinary refuses)."""
import json
import os

import numpy as np
import pytest

import bn256_ref as R

PATH = os.environ.get('DRYNX_KYBER_VECTORS', os.path.join(os.path.dirname(__file__), 'golden', 'kyber_vectors.json'))  # the override exists for tools/make_mock_kyber_vectors.py
REASON = ('parity unpinned: tests/golden/kyber_vectors.json is absent (generate it with go/vectors/dump_test.go on a '
          'machine that has Go and the reference\'s module cache; see go/README.md)')


@pytest.fixture(scope='module')
def vec():
    if not os.path.exists(PATH):
        pytest.skip(REASON)
    with open(PATH) as f:
        return json.load(f)


def H(s):
    return bytes.fromhex(s)


def replay_picks(stream, count):
    """kyber Scalar.Pick over a cipher.Stream: 32 bytes per attempt, accepted when 0 < value < n."""
    out, pos = [], 0
    while len(out) < count:
        v = int.from_bytes(stream[pos:pos + 32], 'big')
        pos += 32
        if 0 < v < R.n:
            out.append(v)
    assert pos == len(stream), 'the prover consumed a different number of bytes than the replay'
    return out


# ---------------------------------------------------------------------------------------------- CPU: definition + oracle
def test_generators_and_group_operations(vec, oracle):
    g1, g2 = oracle.generators()
    assert g1 == H(vec['g1_base']) == R.g1_to_bytes(R.G1_GEN)
    assert g2 == H(vec['g2_base']) == R.g2_to_bytes(R.G2_GEN)
    assert H(vec['g1_null']) == bytes(64)
    for e in vec['g1_mul_base']:
        k = int.from_bytes(H(e['k']), 'big')
        assert R.g1_to_bytes(R.g1_mul(k, R.G1_GEN)) == H(e['out'])
        assert oracle.g1_mul(H(e['k']), g1) == H(e['out'])
    for e in vec['g2_mul_base']:
        k = int.from_bytes(H(e['k']), 'big')
        assert R.g2_to_bytes(R.g2_mul(k, R.G2_GEN)) == H(e['out'])
        assert oracle.g2_mul(H(e['k']), g2) == H(e['out'])
    assert int.from_bytes(H(vec['scalar_setint64_minus5']), 'big') == R.n - 5
    sb = vec['scalar_setbytes_64']
    assert int.from_bytes(H(sb['in']), 'big') % R.n == int.from_bytes(H(sb['out']), 'big')


def test_pairings(vec, oracle):
    for e in vec['pair']:
        P, Q = R.g1_from_bytes(H(e['g1'])), R.g2_from_bytes(H(e['g2']))
        assert R.f12_to_bytes(R.pair(P, Q)) == H(e['out'])
        assert oracle.pair(H(e['g1']), H(e['g2'])) == H(e['out'])
    gm = vec['gt_mul']
    e = R.pair(R.G1_GEN, R.G2_GEN)
    assert R.f12_to_bytes(R.f12pow(e, int.from_bytes(H(gm['k']), 'big'))) == H(gm['out'])
    assert R.f12_to_bytes(R.f12mul(e, e)) == H(gm['add_e_e'])


def test_elgamal_and_aggregation(vec, oracle):
    eg = vec['elgamal']
    pub = H(eg['pub'])
    assert R.g1_to_bytes(R.g1_mul(int.from_bytes(H(eg['sk']), 'big'), R.G1_GEN)) == pub
    for e in eg['enc']:
        assert oracle.encrypt(pub, [e['m']], H(e['r'])) == H(e['ct'])
    a = vec['aggregate']
    rc, out = oracle.aggregate([H(a['a']), H(a['b'])])
    assert rc == 0 and out == H(a['sum'])


def test_range_proof_bytes_and_signatures(vec, oracle):
    sd = vec['sig_deterministic_u16']
    Y, A = oracle.make_sigs([12], 16)
    assert Y == H(sd['Y']) and A == H(sd['A'])
    rp = vec['range_proof']
    ns, u, l, secret = rp['ns'], rp['u'], rp['l'], rp['secret']
    Yb, Ab = b''.join(H(y) for y in rp['Y']), b''.join(H(a) for a in rp['A'])
    nd = len(R.to_base(secret, u, l))
    rnd = replay_picks(H(rp['stream']), nd * (3 + ns))
    rndb = b''.join(x.to_bytes(32, 'big') for x in rnd)
    proofs, off = oracle.rp_create(ns, u, l, Yb, Ab, H(rp['pub']), [secret], H(rp['r']), H(rp['ct']), [rndb])
    assert proofs == H(rp['proof'])
    assert list(oracle.rp_verify(ns, u, l, Yb, H(rp['pub']), H(rp['proof']), off)) == [1]


def test_schnorr_and_unmarshal_rules(vec, oracle):
    s = vec['schnorr']
    assert list(oracle.schnorr_verify([H(s['pub'])], [H(s['sig'])], [H(s['msg'])])) == [1]
    out = np.zeros(128, np.uint8)
    for e in vec['g2_outside_subgroup']:
        accepted_here = oracle.lib.orc_g2_mul((1).to_bytes(32, 'big'), H(e['point']), out.ctypes.data) == 0
        assert accepted_here == e['accepted'], 'G2 subgroup rule differs from kyber: ' + e['error']
    over = vec['g1_coordinate_above_p']
    o64 = np.zeros(64, np.uint8)
    accepted_here = oracle.lib.orc_g1_mul((1).to_bytes(32, 'big'), H(over['point']), o64.ctypes.data) == 0
    if accepted_here != over['accepted']:
        pytest.xfail('documented deviation (DESIGN.md section 2): coordinates >= p are rejected here, kyber v3.0.8 reduces them')


# ---------------------------------------------------------------------------------------------- GPU, through the C ABI
@pytest.mark.gpu
def test_gpu_reproduces_the_reference_bytes(vec):
    from drynx_b200 import Backend
    from drynx_b200._lib import DrynxGpuError
    be = Backend(0, H(vec['g1_base']), H(vec['g2_base']))
    try:
        g1, g2 = H(vec['g1_base']), H(vec['g2_base'])
        for e in vec['g1_mul_base']:
            assert be.g1_mul(H(e['k']), g1).tobytes() == H(e['out'])
        for e in vec['g2_mul_base']:
            assert be.g2_mul(H(e['k']), g2).tobytes() == H(e['out'])
        for e in vec['pair']:
            assert be.pair(H(e['g1']), H(e['g2'])).tobytes() == H(e['out'])
        eg = vec['elgamal']
        be.set_pubkey(H(eg['pub']))
        be.create_decryption_table(10000)
        for e in eg['enc']:
            ct = be.encrypt_int_vector([e['m']], H(e['r']))
            assert ct.tobytes() == H(e['ct'])
            assert int(be.decrypt_int_vector(H(eg['sk']), ct)[0]) == e['dec']
        a = vec['aggregate']
        assert be.aggregate([H(a['a']), H(a['b'])]).tobytes() == H(a['sum'])
        sd = vec['sig_deterministic_u16']
        Y, A = be.init_range_proof_signature([12], 16)
        assert Y.tobytes() == H(sd['Y']) and A.tobytes() == H(sd['A'])
        rp = vec['range_proof']
        ns, u, l, secret = rp['ns'], rp['u'], rp['l'], rp['secret']
        be.set_pubkey(H(rp['pub']))
        be.set_signatures(0, ns, u, b''.join(H(y) for y in rp['Y']), b''.join(H(x) for x in rp['A']))
        nd = be.ndigits(secret, u, l)
        rnd = b''.join(x.to_bytes(32, 'big') for x in replay_picks(H(rp['stream']), nd * (3 + ns)))
        proofs, off = be.create_range_proofs([0], {0: ns}, {0: u}, [l], [secret], H(rp['r']), H(rp['ct']), [rnd])
        assert proofs.tobytes() == H(rp['proof'])
        for path in (1, 2):
            be.set_option('rp_verify_path', path)
            assert list(be.verify_range_proofs([0], [l], H(rp['proof']), off)) == [1]
        s = vec['schnorr']
        assert list(be.schnorr_verify([H(s['pub'])], [H(s['sig'])], [H(s['msg'])])) == [1]
        for e in vec['g2_outside_subgroup']:
            try:
                be.g2_mul((1).to_bytes(32, 'big'), H(e['point']))
                accepted_here = True
            except DrynxGpuError:
                accepted_here = False
            assert accepted_here == e['accepted']
    finally:
        be.close()
