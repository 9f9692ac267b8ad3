package libdrynxrange_test

// Dumps the bytes the reference (kyber v3.0.8 pairing/bn256, unlynx v1.4.0, drynx lib/range) produces for fixed
// inputs, as JSON, for tests/test_kyber_vectors.py of the B200 backend. Pure Go: no cgo, no GPU. Copy this file to
// <drynx>/lib/range/kyber_vectors_test.go and run
//
//	DRYNX_VECTORS_OUT=/path/to/tests/golden/kyber_vectors.json go test ./lib/range/ -run TestDumpKyberVectors
//
// NOT COMPILED in the repository that carries it (no Go toolchain there); see ../README.md.

import (
	"crypto/cipher"
	"encoding/hex"
	"encoding/json"
	"io/ioutil"
	"os"
	"testing"

	libdrynx "github.com/ldsec/drynx/lib"
	libdrynxrange "github.com/ldsec/drynx/lib/range"
	libunlynx "github.com/ldsec/unlynx/lib"
	"github.com/stretchr/testify/require"
	"go.dedis.ch/kyber/v3"
	"go.dedis.ch/kyber/v3/pairing/bn256"
	"go.dedis.ch/kyber/v3/sign/schnorr"
	"go.dedis.ch/kyber/v3/suites"
	"go.dedis.ch/kyber/v3/xof/blake2xb"
)

// recordingStream hands out the bytes of a seeded XOF and keeps a copy of everything that was consumed, so the other
// side can replay the prover's draws (Scalar.Pick = 32 bytes per attempt, accepted when 0 < value < n).
type recordingStream struct {
	inner cipher.Stream
	taken []byte
}

func (r *recordingStream) XORKeyStream(dst, src []byte) {
	r.inner.XORKeyStream(dst, src)
	// Pick passes a zeroed src, so dst now holds the raw key-stream bytes
	r.taken = append(r.taken, dst[:len(src)]...)
}

// seededSuite is the suite Drynx uses with its random stream replaced.
type seededSuite struct {
	suites.Suite
	s cipher.Stream
}

func (d seededSuite) RandomStream() cipher.Stream { return d.s }

func hx(b []byte) string { return hex.EncodeToString(b) }

func mb(t *testing.T, m interface{ MarshalBinary() ([]byte, error) }) []byte {
	b, err := m.MarshalBinary()
	require.NoError(t, err)
	return b
}

func scalarFromUint(s kyber.Scalar, v int64) kyber.Scalar { return s.SetInt64(v) }

func TestDumpKyberVectors(t *testing.T) {
	out := os.Getenv("DRYNX_VECTORS_OUT")
	if out == "" {
		t.Skip("set DRYNX_VECTORS_OUT to the path of tests/golden/kyber_vectors.json")
	}
	libunlynx.SuiTe = libdrynx.Suite
	require.True(t, libdrynx.CurvePairingTest())
	pair := bn256.NewSuite()
	g1, g2, gt := libunlynx.SuiTe, pair.G2(), pair.GT()
	v := map[string]interface{}{
		"kyber": "go.dedis.ch/kyber/v3 v3.0.8", "unlynx": "github.com/ldsec/unlynx v1.4.0", "suite": libunlynx.SuiTe.String(),
	}

	// ---- generators, scalar multiples, pairings
	v["g1_base"] = hx(mb(t, g1.Point().Base()))
	v["g2_base"] = hx(mb(t, g2.Point().Base()))
	v["g1_null"] = hx(mb(t, g1.Point().Null()))
	ks := []int64{1, 2, 3, 12, 65537, 1234567891011}
	var g1m, g2m, pairs []map[string]string
	for _, k := range ks {
		s := g1.Scalar().SetInt64(k)
		g1m = append(g1m, map[string]string{"k": hx(mb(t, s)), "out": hx(mb(t, g1.Point().Mul(s, nil)))})
		s2 := g2.Scalar().SetInt64(k)
		g2m = append(g2m, map[string]string{"k": hx(mb(t, s2)), "out": hx(mb(t, g2.Point().Mul(s2, nil)))})
	}
	big := g1.Scalar().SetBytes([]byte("a fixed scalar well above 2^64 for the vector file, reduced mod n by SetBytes"))
	g1m = append(g1m, map[string]string{"k": hx(mb(t, big)), "out": hx(mb(t, g1.Point().Mul(big, nil)))})
	g2m = append(g2m, map[string]string{"k": hx(mb(t, big)), "out": hx(mb(t, g2.Point().Mul(big, nil)))})
	neg := g1.Scalar().Neg(g1.Scalar().SetInt64(5))
	g1m = append(g1m, map[string]string{"k": hx(mb(t, neg)), "out": hx(mb(t, g1.Point().Mul(neg, nil)))})
	v["g1_mul_base"], v["g2_mul_base"] = g1m, g2m
	v["scalar_setint64_minus5"] = hx(mb(t, neg))
	v["scalar_setbytes_64"] = map[string]string{
		"in":  hx([]byte("0123456789abcdef0123456789abcdef0123456789abcdef0123456789abcdef")),
		"out": hx(mb(t, g1.Scalar().SetBytes([]byte("0123456789abcdef0123456789abcdef0123456789abcdef0123456789abcdef")))),
	}
	for _, ab := range [][2]int64{{1, 1}, {2, 3}, {65537, 12}} {
		a, b := g1.Scalar().SetInt64(ab[0]), g2.Scalar().SetInt64(ab[1])
		P, Q := g1.Point().Mul(a, nil), g2.Point().Mul(b, nil)
		pairs = append(pairs, map[string]string{"g1": hx(mb(t, P)), "g2": hx(mb(t, Q)), "out": hx(mb(t, pair.Pair(P, Q)))})
	}
	v["pair"] = pairs
	e := pair.Pair(g1.Point().Base(), g2.Point().Base())
	v["gt_mul"] = map[string]string{ // GT "Mul" by a scalar = exponentiation; GT "Add" = product
		"k": hx(mb(t, big)), "out": hx(mb(t, gt.Point().Mul(big, e))),
		"add_e_e": hx(mb(t, gt.Point().Add(e, e))),
	}

	// ---- ElGamal (libunlynx.EncryptIntGetR: K = rB, C = mB + rP), aggregation, decryption
	sk := g1.Scalar().SetBytes([]byte("collective secret key of the vector file"))
	P := g1.Point().Mul(sk, nil)
	var encs []map[string]interface{}
	for _, m := range []int64{0, 1, 25, -11955, 99999} {
		ct, r := libunlynx.EncryptIntGetR(P, m)
		b, err := ct.ToBytes()
		require.NoError(t, err)
		encs = append(encs, map[string]interface{}{"m": m, "r": hx(mb(t, r)), "ct": hx(b),
			"dec": libunlynx.DecryptIntWithNeg(sk, *ct)})
	}
	v["elgamal"] = map[string]interface{}{"sk": hx(mb(t, sk)), "pub": hx(mb(t, P)), "enc": encs}
	c1, _ := libunlynx.EncryptIntGetR(P, 40)
	c2, _ := libunlynx.EncryptIntGetR(P, 2)
	sum := libunlynx.CipherText{K: g1.Point().Add(c1.K, c2.K), C: g1.Point().Add(c1.C, c2.C)}
	b1, _ := c1.ToBytes()
	b2, _ := c2.ToBytes()
	bs, _ := sum.ToBytes()
	v["aggregate"] = map[string]string{"a": hx(b1), "b": hx(b2), "sum": hx(bs)}

	// ---- range-proof signatures and one full proof with recorded randomness
	det := libdrynxrange.InitRangeProofSignatureDeterministic(16)
	v["sig_deterministic_u16"] = map[string]string{"Y": hx(mb(t, det.Public)), "A": hx(det.Signature)}

	const ns = 3
	u, l, secret := int64(16), int64(4), int64(40000)
	sigBytes := make([]libdrynx.PublishSignatureBytes, ns)
	sigs := make([]libdrynx.PublishSignature, ns)
	ys := make([]kyber.Point, ns)
	var ysHex, asHex []string
	for i := range sigs {
		sigBytes[i] = libdrynxrange.InitRangeProofSignature(u)
		sigs[i] = libdrynxrange.PublishSignatureBytesToPublishSignatures(sigBytes[i])
		ys[i] = sigs[i].Public
		ysHex = append(ysHex, hx(mb(t, ys[i])))
		asHex = append(asHex, hx(sigBytes[i].Signature))
	}
	cipherText, r := libunlynx.EncryptIntGetR(P, secret)
	rec := &recordingStream{inner: blake2xb.New([]byte("drynx-b200 range proof vector"))}
	libunlynx.SuiTe = seededSuite{libdrynx.Suite, rec}
	proof := libdrynxrange.CreatePredicateRangeProofForAllServ(libdrynxrange.CreateProof{
		Sigs: sigs, U: u, L: l, Secret: secret, R: r, CaPub: P, Cipher: *cipherText})
	libunlynx.SuiTe = libdrynx.Suite
	require.True(t, libdrynxrange.RangeProofVerification(proof, u, l, ys, P))
	pb := proof.ToBytes()
	flat := append([]byte{}, pb.Commit...)
	flat = append(flat, *pb.RP.Challenge...)
	flat = append(flat, *pb.RP.Zr...)
	flat = append(flat, *pb.RP.D...)
	flat = append(flat, *pb.RP.Zphi...)
	for i := 0; i < ns; i++ {
		flat = append(flat, (*pb.RP.V)[i]...)
		flat = append(flat, (*pb.RP.A)[i]...)
		flat = append(flat, (*pb.RP.Zv)[i]...)
	}
	ctb, _ := cipherText.ToBytes()
	v["range_proof"] = map[string]interface{}{
		"ns": ns, "u": u, "l": l, "secret": secret, "pub": hx(mb(t, P)), "r": hx(mb(t, r)), "ct": hx(ctb),
		"Y": ysHex, "A": asHex,
		// every byte the prover took from the suite's random stream, in order: s_j, t_j, m_j, v_0j, v_1j, v_2j per digit,
		// 32 bytes per Pick attempt (an attempt is rejected when the value is 0 or >= n)
		"stream": hx(rec.taken),
		"proof":  hx(flat), "verifies": true,
	}

	// ---- Schnorr (kyber sign/schnorr over G1, as lib/proof/structs_proofs.go:117,500 uses it)
	priv := g1.Scalar().SetBytes([]byte("schnorr private key of the vector file"))
	pubS := g1.Point().Mul(priv, nil)
	msg := []byte("a serialised proof request")
	sig, err := schnorr.Sign(libunlynx.SuiTe, priv, msg)
	require.NoError(t, err)
	require.NoError(t, schnorr.Verify(pair.G1(), pubS, msg, sig))
	v["schnorr"] = map[string]string{"priv": hx(mb(t, priv)), "pub": hx(mb(t, pubS)), "msg": hx(msg), "sig": hx(sig)}

	// ---- twist points outside the order-n subgroup: what does UnmarshalBinary say?
	outside := []string{
		"7d99a696dca9e563182cfb0276396fa6f4027d33da3e8a49f7bfc971cf6267cb080e305082d5ab28784ad392340c5201357ade6da03a40442baabf70a0ee925b4c6464e4187a89a1d99781679bc1900df4c17b8a6edeadf393c78f008030de9f2f3ef0e8c6ce8244b79d25449c9720e3cdcd86cb25b967ef071825d4ef901058",
		"533608e2f8be1a21bc067aabb04583a7a07b158c8f30fe6fa504f6df288f144b82e191ac3e8557b3640b9819f7a179ed4b6f7074ac504e1abac0cf5f4af1d09a5ece4ae05460a2dd9aa66a9c9ab92a92a8cd79a74df6bb5f2d64ff27c2d3ec7f577160825466d0b85bf684aa2339c18334a6e5840ca3900af139080e94ddb6a1",
	}
	var outRes []map[string]interface{}
	for _, h := range outside {
		raw, _ := hex.DecodeString(h)
		err := g2.Point().UnmarshalBinary(raw)
		msgErr := ""
		if err != nil {
			msgErr = err.Error()
		}
		outRes = append(outRes, map[string]interface{}{"point": h, "accepted": err == nil, "error": msgErr})
	}
	v["g2_outside_subgroup"] = outRes
	// a coordinate >= p: kyber v3.0.8 reduces it silently or rejects it?
	over, _ := hex.DecodeString(v["g1_base"].(string))
	over[0] |= 0x80 // x + 2^255 > p for the generator's x
	errOver := g1.Point().UnmarshalBinary(over)
	v["g1_coordinate_above_p"] = map[string]interface{}{"point": hx(over), "accepted": errOver == nil}

	js, err := json.MarshalIndent(v, "", " ")
	require.NoError(t, err)
	require.NoError(t, ioutil.WriteFile(out, js, 0644))
}
