//go:build gpu
// +build gpu

package gpubackend

// Round trips in the style of the reference's own tests (lib/encoding/sum_test.go:15-31,
// lib/range/range_proof_test.go:43-77), through the GPU backend and checked against the reference's CPU functions.
// Needs a B200 and libdrynx_gpu.so; not compiled in the repository that carries it.

import (
	"testing"

	libdrynx "github.com/ldsec/drynx/lib"
	libdrynxrange "github.com/ldsec/drynx/lib/range"
	libunlynx "github.com/ldsec/unlynx/lib"
	"github.com/stretchr/testify/require"
	"go.dedis.ch/kyber/v3"
)

func TestEncryptAggregateDecrypt(t *testing.T) {
	x, err := New(0)
	require.NoError(t, err)
	defer x.Close()
	sk, pk := libunlynx.GenKey()
	require.NoError(t, x.SetPubKey(pk))
	require.NoError(t, x.CreateDecryptionTable(10000))
	a, _, err := x.EncryptIntVectorGetRs([]int64{1, 2, -3, 400})
	require.NoError(t, err)
	b, _, err := x.EncryptIntVectorGetRs([]int64{10, 20, 30, -500})
	require.NoError(t, err)
	// the CPU path decrypts what the GPU encrypted
	require.Equal(t, []int64{1, 2, -3, 400}, libunlynx.DecryptIntVectorWithNeg(sk, &a))
	ab, _, _ := a.ToBytes()
	bb, _, _ := b.ToBytes()
	sum, err := x.Aggregate([][]byte{ab, bb}, 4)
	require.NoError(t, err)
	cv := libunlynx.NewCipherVector(4)
	cv.FromBytes(sum, 4)
	got, err := x.DecryptIntVectorWithNeg(sk, *cv)
	require.NoError(t, err)
	require.Equal(t, []int64{11, 22, 27, -100}, got)
}

func TestRangeProofRoundTrip(t *testing.T) {
	if !libdrynx.CurvePairingTest() {
		t.Skip("no pairing suite")
	}
	x, err := New(0)
	require.NoError(t, err)
	defer x.Close()
	const u, l, ns = int64(16), int64(4), 3
	sk, pk := libunlynx.GenKey()
	_ = sk
	require.NoError(t, x.SetPubKey(pk))
	sigBytes := make([]libdrynx.PublishSignatureBytes, ns)
	sigs := make([]libdrynx.PublishSignature, ns)
	for i := range sigBytes {
		sigBytes[i] = libdrynxrange.InitRangeProofSignature(u)
		sigs[i] = libdrynxrange.PublishSignatureBytesToPublishSignatures(sigBytes[i])
	}
	require.NoError(t, x.SetSignatures(0, u, sigBytes))
	secret := int64(25)
	cipher, r := libunlynx.EncryptIntGetR(pk, secret)
	cp := libdrynxrange.CreateProof{Sigs: sigs, U: u, L: l, Secret: secret, R: r, CaPub: pk, Cipher: *cipher}
	proofs, err := x.CreatePredicateRangeProofListForAllServers([]libdrynxrange.CreateProof{cp}, []int32{0})
	require.NoError(t, err)
	// the reference's CPU verifier accepts the GPU's proof ...
	ys := make([]kyber.Point, ns)
	for i := range ys {
		ys[i] = sigs[i].Public
	}
	require.True(t, libdrynxrange.RangeProofVerification(proofs[0], u, l, ys, pk))
	// ... and the GPU verifier accepts the reference's proof and rejects a tampered one
	ref := libdrynxrange.CreatePredicateRangeProofForAllServ(cp)
	data := []libdrynxrange.RangeProofBytes{ref.ToBytes(), proofs[0].ToBytes()}
	ok, err := x.RangeProofListVerification(libdrynxrange.RangeProofListBytes{Data: &data}, []int32{0, 0}, []int64{l, l}, pk)
	require.NoError(t, err)
	require.True(t, ok)
	(*data[0].RP.Zr)[5] ^= 1
	ok, err = x.RangeProofListVerification(libdrynxrange.RangeProofListBytes{Data: &data}, []int32{0, 0}, []int64{l, l}, pk)
	require.NoError(t, err)
	require.False(t, ok)
}
