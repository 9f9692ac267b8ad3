//go:build gpu
// +build gpu

package gpubackend

/*
#include "drynx_gpu.h"
*/
import "C"

import (
	"crypto/sha512"
	"unsafe"

	libunlynx "github.com/ldsec/unlynx/lib"
	"go.dedis.ch/kyber/v3"
	"go.dedis.ch/kyber/v3/util/random"
)

// KeySwitchSequence replaces libunlynxkeyswitch.KeySwitchSequence (unlynx v1.4.0 lib/key_switch; called at
// data/data.go:80 and by the key-switching protocol, services/service.go:565-616): per element
// cv[i] = (v_i B, v_i Q - sk rB_i), plus -rB_i and v_i for the proof.
func (x *Ctx) KeySwitchSequence(targetPubKey kyber.Point, rBs []kyber.Point, secretKey kyber.Scalar) (libunlynx.CipherVector, []kyber.Point, []kyber.Point, []kyber.Scalar, error) {
	n := len(rBs)
	q, err := targetPubKey.MarshalBinary()
	if err != nil {
		return nil, nil, nil, nil, err
	}
	if C.drx_set_target_pubkey(x.c, u8(q)) != 0 {
		return nil, nil, nil, nil, lastErr()
	}
	rb, v := make([]byte, G1Len*n), make([]byte, ScalarLen*n)
	vis := make([]kyber.Scalar, n)
	for i := range rBs {
		b, err := rBs[i].MarshalBinary()
		if err != nil {
			return nil, nil, nil, nil, err
		}
		copy(rb[G1Len*i:], b)
		vis[i] = libunlynx.SuiTe.Scalar().Pick(random.New())
		s, _ := vis[i].MarshalBinary()
		copy(v[ScalarLen*i:], s)
	}
	sk, err := secretKey.MarshalBinary()
	if err != nil {
		return nil, nil, nil, nil, err
	}
	cvb, neg := make([]byte, CtLen*n), make([]byte, G1Len*n)
	if C.drx_keyswitch(x.c, C.size_t(n), u8(sk), u8(rb), u8(v), u8(cvb), u8(neg)) != 0 {
		return nil, nil, nil, nil, lastErr()
	}
	cv := libunlynx.NewCipherVector(n)
	cv.FromBytes(cvb, n)
	ks2s, rBNegs := make([]kyber.Point, n), make([]kyber.Point, n)
	for i := 0; i < n; i++ {
		ks2s[i] = (*cv)[i].C
		rBNegs[i] = libunlynx.SuiTe.Point()
		if err := rBNegs[i].UnmarshalBinary(neg[G1Len*i : G1Len*(i+1)]); err != nil {
			return nil, nil, nil, nil, err
		}
	}
	return *cv, ks2s, rBNegs, vis, nil
}

// VerifyProofSignatures replaces VerifyProofSignature (lib/proof/structs_proofs.go:498-505) for all the proof
// requests a verifying node received in one round: kyber schnorr.Verify over G1, h = SHA-512(R || P || msg).
func (x *Ctx) VerifyProofSignatures(pubKeys []kyber.Point, datas [][]byte, sigs [][]byte) ([]bool, error) {
	n := len(datas)
	res := make([]bool, n)
	if n == 0 {
		return res, nil
	}
	pub, sig, dig, ok := make([]byte, G1Len*n), make([]byte, 96*n), make([]byte, 64*n), make([]byte, n)
	wellFormed := make([]bool, n)
	for i := range datas {
		if len(sigs[i]) != 96 { // schnorr.Verify: "schnorr: signature of invalid length"
			continue
		}
		p, err := pubKeys[i].MarshalBinary()
		if err != nil {
			return nil, err
		}
		wellFormed[i] = true
		copy(pub[G1Len*i:], p)
		copy(sig[96*i:], sigs[i])
		h := sha512.New()
		h.Write(sigs[i][:64])
		h.Write(p)
		h.Write(datas[i])
		copy(dig[64*i:], h.Sum(nil))
	}
	if C.drx_schnorr_verify(x.c, C.size_t(n), u8(pub), u8(sig), u8(dig), u8(ok)) != 0 {
		return nil, lastErr()
	}
	for i := range res {
		res[i] = wellFormed[i] && ok[i] == 1
	}
	return res, nil
}

// MulCipherVectorByScalars replaces the MulCipherTextbyScalar loop of the obfuscation protocol
// (protocols/obfuscation_protocol.go:236-253): out[i] = (s_i K_i, s_i C_i).
func (x *Ctx) MulCipherVectorByScalars(cv libunlynx.CipherVector, ss []kyber.Scalar) (libunlynx.CipherVector, error) {
	n := len(cv)
	if n == 0 {
		return libunlynx.CipherVector{}, nil
	}
	ct, _, _ := cv.ToBytes()
	sb := make([]byte, ScalarLen*n)
	for i := range ss {
		b, err := ss[i].MarshalBinary()
		if err != nil {
			return nil, err
		}
		copy(sb[ScalarLen*i:], b)
	}
	out := make([]byte, CtLen*n)
	if C.drx_ct_scalar_mul(x.c, C.size_t(n), u8(sb), u8(ct), u8(out)) != 0 {
		return nil, lastErr()
	}
	res := libunlynx.NewCipherVector(n)
	res.FromBytes(out, n)
	return *res, nil
}

// ShuffleRerandomize replaces the per-ciphertext work of libunlynxshuffle.ShuffleSequence (data/data.go:68): output
// line i is input line pi[i], every ciphertext re-randomised with beta[pi[i]][j] (beta and pi stay drawn in Go).
func (x *Ctx) ShuffleRerandomize(lines []libunlynx.CipherVector, beta [][]kyber.Scalar, pi []int32) ([]libunlynx.CipherVector, error) {
	k := len(lines)
	if k == 0 {
		return nil, nil
	}
	nq := len(lines[0])
	in, bb := make([]byte, 0, CtLen*k*nq), make([]byte, 0, ScalarLen*k*nq)
	for i := range lines {
		ct, _, _ := lines[i].ToBytes()
		in = append(in, ct...)
		for j := 0; j < nq; j++ {
			b, err := beta[i][j].MarshalBinary()
			if err != nil {
				return nil, err
			}
			bb = append(bb, b...)
		}
	}
	out := make([]byte, CtLen*k*nq)
	if C.drx_shuffle_rerandomize(x.c, C.size_t(k), C.size_t(nq), (*C.int32_t)(unsafe.Pointer(&pi[0])), u8(in), u8(bb), u8(out)) != 0 {
		return nil, lastErr()
	}
	res := make([]libunlynx.CipherVector, k)
	for i := range res {
		cv := libunlynx.NewCipherVector(nq)
		cv.FromBytes(out[CtLen*nq*i:CtLen*nq*(i+1)], nq)
		res[i] = *cv
	}
	return res, nil
}
