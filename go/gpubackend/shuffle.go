//go:build gpu
// +build gpu

package gpubackend

/*
#include "drynx_gpu.h"
*/
import "C"

import (
	"bytes"

	libunlynx "github.com/ldsec/unlynx/lib"
	"go.dedis.ch/kyber/v3"
)

// PairShuffleVerify does the verifier's side of kyber's shuffle.PairShuffle (Neff's verifiable shuffle of ElGamal pairs;
// kyber v3.0.8 shuffle/pair.go and shuffle/simple.go) with every group operation on the device: what
// libunlynxshuffle.ShuffleProofVerification runs after it has compressed the lines (the verifying node calls it at
// lib/proof/structs_proofs.go:412). The proof bytes are the messages in the order the prover Put them:
//
//	Gamma | A[k] | C[k] | U[k] | W[k] | Lambda1 | Lambda2 | D[k] | Zsigma[k] | Ztau | X[k] | Y[k] | Theta[2k] | Zalpha[2k-1]
//
// and the public random values (Zrho[k], Zlambda, Zt, Zc) come from the same transcript proof.HashVerify keeps (the
// `transcript` type of proofs.go). g == nil means the suite's base point, as in shuffle.Verifier.
//
// The message flow and the checks (31)-(35) are written from memory of kyber's source (it is not part of the drynx
// tree) and mirror tests/shuffle_proof.py, which runs the same flow through the CUDA library and the CPU oracle;
// NOT COMPILED here (no Go toolchain), like the rest of this package.
func (x *Ctx) PairShuffleVerify(g, h kyber.Point, X, Y, Xbar, Ybar []kyber.Point, prf []byte) (bool, error) {
	k := len(X)
	want := G1Len*(4*k+3) + G1Len*k + ScalarLen*(k+1) + 2*G1Len*k + 2*G1Len*k + ScalarLen*(2*k-1)
	if k < 2 || len(Y) != k || len(Xbar) != k || len(Ybar) != k || len(prf) != want {
		return false, nil
	}
	suite := libunlynx.SuiTe
	if g == nil {
		g = suite.Point().Base()
	}
	gb, hb := make([]byte, G1Len), make([]byte, G1Len)
	if err := marshalPoints(gb, g); err != nil {
		return false, err
	}
	if err := marshalPoints(hb, h); err != nil {
		return false, err
	}
	flat := func(pts []kyber.Point) ([]byte, error) {
		out := make([]byte, G1Len*len(pts))
		return out, marshalPoints(out, pts...)
	}
	xb, err := flat(X)
	if err != nil {
		return false, err
	}
	yb, err := flat(Y)
	if err != nil {
		return false, err
	}
	xbarb, err := flat(Xbar)
	if err != nil {
		return false, err
	}
	ybarb, err := flat(Ybar)
	if err != nil {
		return false, err
	}

	// the transcript: messages are absorbed in the order they were Put, public randomness is read in between
	tr := newTranscript("PairShuffle")
	pos := 0
	take := func(n int) []byte { b := prf[pos : pos+n]; pos += n; return b }
	scalarsOf := func(msg []byte, n int) ([][]byte, error) {
		out := make([][]byte, n)
		for i := 0; i < n; i++ {
			var m []byte
			if i == 0 {
				m = msg
			}
			c, err := tr.challenge(m)
			if err != nil {
				return nil, err
			}
			out[i], _ = c.MarshalBinary()
		}
		return out, nil
	}
	neg := func(s []byte) []byte {
		v := suite.Scalar()
		_ = v.UnmarshalBinary(s)
		b, _ := v.Neg(v).MarshalBinary()
		return b
	}
	one, _ := suite.Scalar().One().MarshalBinary()
	minusOne := neg(one)
	zero, _ := suite.Scalar().Zero().MarshalBinary()

	p1 := take(G1Len * (4*k + 3))
	gamma := p1[:G1Len]
	w := p1[G1Len*(1+3*k) : G1Len*(1+4*k)]
	lambdas := p1[G1Len*(4*k+1):]
	rho, err := scalarsOf(p1, k)
	if err != nil {
		return false, err
	}
	d := take(G1Len * k)
	if _, err = scalarsOf(d, 1); err != nil { // Zlambda: only the strict variant of the check uses it
		return false, err
	}
	p5 := take(ScalarLen * (k + 1))
	sigma, tau := p5[:ScalarLen*k], p5[ScalarLen*k:]
	p0 := take(2 * G1Len * k)
	ts, err := scalarsOf(append(append([]byte{}, p5...), p0...), 1)
	if err != nil {
		return false, err
	}
	theta := take(2 * G1Len * k)
	cs, err := scalarsOf(theta, 1)
	if err != nil {
		return false, err
	}
	alpha := take(ScalarLen * (2*k - 1))
	t, c := ts[0], cs[0]

	// simple shuffle, V step 5: Xhat_i = X_i - t g, Yhat_i = Y_i - t Gamma (two-term sums), then 2k checks
	// a A - b B == T as one-term relations (resp = a, bases = A, chal = -b, P = B, V = T)
	sc := bytes.Repeat(append(append([]byte{}, one...), neg(t)...), k)
	pts := make([]byte, 0, 2*G1Len*k)
	for i := 0; i < k; i++ {
		pts = append(append(pts, p0[G1Len*i:G1Len*(i+1)]...), gb...)
	}
	xhat := make([]byte, G1Len*k)
	if C.drx_rep_commit(x.c, C.size_t(k), 2, u8(sc), u8(pts), u8(xhat)) != 0 {
		return false, nil // a point of the proof does not unmarshal
	}
	pts = pts[:0]
	for i := 0; i < k; i++ {
		pts = append(append(pts, p0[G1Len*(k+i):G1Len*(k+i+1)]...), gamma...)
	}
	yhat := make([]byte, G1Len*k)
	if C.drx_rep_commit(x.c, C.size_t(k), 2, u8(sc), u8(pts), u8(yhat)) != 0 {
		return false, nil
	}
	thlen := 2*k - 1
	al := func(i int) []byte { return alpha[ScalarLen*i : ScalarLen*(i+1)] }
	var resp, chal, A, B []byte
	add := func(a, b, pa, pb []byte) {
		resp, chal, A, B = append(resp, a...), append(chal, neg(b)...), append(A, pa...), append(B, pb...)
	}
	add(c, al(0), xhat[:G1Len], yhat[:G1Len])
	for i := 1; i < k; i++ {
		add(al(i-1), al(i), xhat[G1Len*i:G1Len*(i+1)], yhat[G1Len*i:G1Len*(i+1)])
	}
	for i := k; i < thlen; i++ {
		add(al(i-1), al(i), gamma, gb)
	}
	add(al(thlen-1), c, gamma, gb)
	ok := make([]byte, 2*k)
	if C.drx_rep_verify(x.c, C.size_t(2*k), 1, u8(resp), u8(chal), u8(A), u8(B), u8(theta), u8(ok)) != 0 {
		return false, lastErr()
	}
	if bytes.IndexByte(ok, 0) >= 0 {
		return false, nil
	}
	// (33): Zsigma_i Gamma == W_i + D_i
	ok = ok[:k]
	if C.drx_rep_verify(x.c, C.size_t(k), 1, u8(sigma), u8(bytes.Repeat(minusOne, k)), u8(bytes.Repeat(gamma, k)), u8(d), u8(w), u8(ok)) != 0 {
		return false, lastErr()
	}
	if bytes.IndexByte(ok, 0) >= 0 {
		return false, nil
	}
	// (31), (32), (34), (35): Lambda == sum Zsigma_i bar_i - sum Zrho_i orig_i - Ztau base, sums of 2k+1 terms
	r := append([]byte{}, sigma...)
	for i := 0; i < k; i++ {
		r = append(r, neg(rho[i])...)
	}
	r = append(r, neg(tau)...)
	phi1, err := x.wideSum(r, append(append(append([]byte{}, xbarb...), xb...), gb...), one, zero, gb)
	if err != nil {
		return false, nil
	}
	phi2, err := x.wideSum(r, append(append(append([]byte{}, ybarb...), yb...), hb...), one, zero, gb)
	if err != nil {
		return false, nil
	}
	return bytes.Equal(phi1, lambdas[:G1Len]) && bytes.Equal(phi2, lambdas[G1Len:]), nil
}

// wideSum returns sum_j scalars[j] * points[j]. One device thread walks the terms of a relation, so a wide sum goes up
// a tree: 16 terms per relation (all relations of a level in one drx_rep_commit), then the partial sums with scalar 1,
// until one point is left (k = 1024 pairs: 94 ms for the whole verification instead of 2.6 s with two 2049-term
// relations). pad is any valid point; it fills the last chunk with zero terms.
func (x *Ctx) wideSum(scalars, points, one, zero, pad []byte) ([]byte, error) {
	const chunk = 16
	scalars, points = append([]byte{}, scalars...), append([]byte{}, points...)
	n := len(scalars) / ScalarLen
	for n > chunk {
		for n%chunk != 0 {
			scalars, points = append(scalars, zero...), append(points, pad...)
			n++
		}
		part := make([]byte, G1Len*(n/chunk))
		if C.drx_rep_commit(x.c, C.size_t(n/chunk), chunk, u8(scalars), u8(points), u8(part)) != 0 {
			return nil, lastErr()
		}
		n /= chunk
		points, scalars = part, bytes.Repeat(one, n)
	}
	out := make([]byte, G1Len)
	if C.drx_rep_commit(x.c, 1, C.int(n), u8(scalars), u8(points), u8(out)) != 0 {
		return nil, lastErr()
	}
	return out, nil
}
