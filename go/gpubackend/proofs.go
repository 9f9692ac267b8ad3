//go:build gpu
// +build gpu

package gpubackend

/*
#include "drynx_gpu.h"
*/
import "C"

import (
	"bytes"
	"math"

	libdrynxobfuscation "github.com/ldsec/drynx/lib/obfuscation"
	libunlynx "github.com/ldsec/unlynx/lib"
	libunlynxkeyswitch "github.com/ldsec/unlynx/lib/key_switch"
	"go.dedis.ch/kyber/v3"
	"go.dedis.ch/kyber/v3/util/random"
)

// The two proof.Rep list proofs on the verifying-node path with their group arithmetic on the device:
// libdrynxobfuscation.ObfuscationListProofCreation / ObfuscationListProofVerification
// (lib/obfuscation/obfuscation_proof.go:62-111) here; the key-switch list proof has the same shape with three
// relations (drx_rep_commit / drx_rep_verify with nterms 1 and 2, see tests/rep_proofs.py for the layout).
//
// What stays in Go is the Fiat-Shamir transcript: kyber's proof.HashProve keeps an XOF seeded with the protocol name,
// reseeds it with every commitment the prover Put, and reads the challenge from it (kyber v3.0.8 proof/hash.go). The
// hashProver type is unexported, so `transcript` below repeats its few lines over kyber's public XOF interface; the
// bytes of the proof (commitments, then one response per secret) are laid out in the order kyber's Rep and And
// predicates Put them, so a proof made here verifies with proof.HashVerify and the other way round.
// Not compiled or run in the build container (no Go toolchain): written against kyber v3.0.8.
type transcript struct {
	pubrand kyber.XOF
}

func newTranscript(protocolName string) *transcript {
	return &transcript{pubrand: libunlynx.SuiTe.XOF([]byte(protocolName))}
}

// challenge consumes the message bytes Put so far (hashProver.consumeMsg) and reads one scalar (hashProver.PubRand).
func (t *transcript) challenge(msg []byte) (kyber.Scalar, error) {
	if len(msg) > 0 {
		t.pubrand.Reseed()
		if _, err := t.pubrand.Write(msg); err != nil {
			return nil, err
		}
	}
	c := libunlynx.SuiTe.Scalar().Pick(t.pubrand)
	return c, nil
}

const obfProofLen = 2*G1Len + ScalarLen

// ObfuscationListProofCreation replaces libdrynxobfuscation.ObfuscationListProofCreation (called at
// protocols/obfuscation_protocol.go:259 and data/data.go:57): one commitment call and one response call for the list
// where the Go code starts a goroutine per proof.
func (x *Ctx) ObfuscationListProofCreation(c, co []libunlynx.CipherText, s []kyber.Scalar) (libdrynxobfuscation.PublishedListObfuscationProof, error) {
	n := len(s)
	plop := libdrynxobfuscation.PublishedListObfuscationProof{Prs: make([]libdrynxobfuscation.PublishedObfuscationProof, n)}
	if n == 0 {
		return plop, nil
	}
	cb, vb, sb := make([]byte, CtLen*n), make([]byte, ScalarLen*n), make([]byte, ScalarLen*n)
	for i := 0; i < n; i++ {
		b, err := c[i].ToBytes()
		if err != nil {
			return plop, err
		}
		copy(cb[CtLen*i:], b)
		v, _ := libunlynx.SuiTe.Scalar().Pick(random.New()).MarshalBinary()
		copy(vb[ScalarLen*i:], v)
		sv, err := s[i].MarshalBinary()
		if err != nil {
			return plop, err
		}
		copy(sb[ScalarLen*i:], sv)
	}
	// V1 || V2 = v*c1 || v*c2: the commitments of the two Rep predicates, which share the blinding value of "s"
	V := make([]byte, CtLen*n)
	if C.drx_ct_scalar_mul(x.c, C.size_t(n), u8(vb), u8(cb), u8(V)) != 0 {
		return plop, lastErr()
	}
	chal := make([]byte, ScalarLen*n)
	for i := 0; i < n; i++ {
		ch, err := newTranscript("proofTest").challenge(V[CtLen*i : CtLen*(i+1)])
		if err != nil {
			return plop, err
		}
		b, _ := ch.MarshalBinary()
		copy(chal[ScalarLen*i:], b)
	}
	r := make([]byte, ScalarLen*n)
	if C.drx_rep_respond(x.c, C.size_t(n), 1, u8(vb), u8(chal), u8(sb), u8(r)) != 0 {
		return plop, lastErr()
	}
	for i := 0; i < n; i++ {
		pr := make([]byte, 0, obfProofLen)
		pr = append(pr, V[CtLen*i:CtLen*(i+1)]...)
		pr = append(pr, r[ScalarLen*i:ScalarLen*(i+1)]...)
		plop.Prs[i] = libdrynxobfuscation.PublishedObfuscationProof{C: c[i], Co: co[i], Proof: pr}
	}
	return plop, nil
}

// ObfuscationListProofVerification replaces libdrynxobfuscation.ObfuscationListProofVerification
// (lib/obfuscation/obfuscation_proof.go:94-111; the verifying node calls it at lib/proof/structs_proofs.go:328): the
// first ceil(percent*len) proofs, 2 relations each, in one device call.
func (x *Ctx) ObfuscationListProofVerification(plop libdrynxobfuscation.PublishedListObfuscationProof, percent float64) (bool, error) {
	n := int(math.Ceil(percent * float64(len(plop.Prs))))
	if n == 0 {
		return true, nil
	}
	resp, chal := make([]byte, 2*ScalarLen*n), make([]byte, 2*ScalarLen*n)
	bases, P, V := make([]byte, CtLen*n), make([]byte, CtLen*n), make([]byte, CtLen*n)
	for i := 0; i < n; i++ {
		pr := plop.Prs[i]
		if len(pr.Proof) != obfProofLen {
			return false, nil
		}
		ch, err := newTranscript("proofTest").challenge(pr.Proof[:CtLen])
		if err != nil {
			return false, err
		}
		cbin, _ := ch.MarshalBinary()
		for h := 0; h < 2; h++ {
			copy(resp[ScalarLen*(2*i+h):], pr.Proof[CtLen:])
			copy(chal[ScalarLen*(2*i+h):], cbin)
		}
		cb, err := pr.C.ToBytes()
		if err != nil {
			return false, err
		}
		cob, err := pr.Co.ToBytes()
		if err != nil {
			return false, err
		}
		copy(bases[CtLen*i:], cb)
		copy(P[CtLen*i:], cob)
		copy(V[CtLen*i:], pr.Proof[:CtLen])
	}
	ok := make([]byte, 2*n)
	if C.drx_rep_verify(x.c, C.size_t(2*n), 1, u8(resp), u8(chal), u8(bases), u8(P), u8(V), u8(ok)) != 0 {
		return false, lastErr()
	}
	if bytes.IndexByte(ok, 0) >= 0 {
		return false, nil
	}
	return true, nil
}

const ksProofLen = 3*G1Len + 2*ScalarLen

func marshalPoints(dst []byte, pts ...kyber.Point) error {
	for i, p := range pts {
		b, err := p.MarshalBinary()
		if err != nil {
			return err
		}
		copy(dst[G1Len*i:], b)
	}
	return nil
}

// KeySwitchListProofVerification replaces libunlynxkeyswitch.KeySwitchListProofVerification (unlynx v1.4.0
// lib/key_switch; the verifying node calls it at lib/proof/structs_proofs.go:476): predicate
// And(Rep(viB = vi*B), Rep(K = k*B), Rep(ks2 = vi*Q + k*rBNeg)), proof bytes V1 || V2 || V3 || r_vi || r_k. Two device
// calls for the first ceil(percent*len) proofs: 2n relations with one term, n relations with two.
func (x *Ctx) KeySwitchListProofVerification(pkslp libunlynxkeyswitch.PublishedKSListProof, percent float64) (bool, error) {
	n := int(math.Ceil(percent * float64(len(pkslp.List))))
	if n == 0 {
		return true, nil
	}
	B := libunlynx.SuiTe.Point().Base()
	r1, c1 := make([]byte, 2*ScalarLen*n), make([]byte, 2*ScalarLen*n)
	b1, p1, v1 := make([]byte, 2*G1Len*n), make([]byte, 2*G1Len*n), make([]byte, 2*G1Len*n)
	c2, b2, p2, v2 := make([]byte, ScalarLen*n), make([]byte, 2*G1Len*n), make([]byte, G1Len*n), make([]byte, G1Len*n)
	for i := 0; i < n; i++ {
		pr := pkslp.List[i]
		if len(pr.Proof) != ksProofLen {
			return false, nil
		}
		ch, err := newTranscript("proofTest").challenge(pr.Proof[:3*G1Len])
		if err != nil {
			return false, err
		}
		cbin, _ := ch.MarshalBinary()
		copy(r1[2*ScalarLen*i:], pr.Proof[3*G1Len:]) // r_vi, r_k: relation 1 uses r_vi, relation 2 uses r_k
		copy(c1[2*ScalarLen*i:], cbin)
		copy(c1[2*ScalarLen*i+ScalarLen:], cbin)
		copy(c2[ScalarLen*i:], cbin)
		if err := marshalPoints(b1[2*G1Len*i:], B, B); err != nil {
			return false, err
		}
		if err := marshalPoints(p1[2*G1Len*i:], pr.ViB, pr.K); err != nil {
			return false, err
		}
		copy(v1[2*G1Len*i:], pr.Proof[:2*G1Len])
		if err := marshalPoints(b2[2*G1Len*i:], pr.Q, pr.RbNeg); err != nil {
			return false, err
		}
		if err := marshalPoints(p2[G1Len*i:], pr.Ks2); err != nil {
			return false, err
		}
		copy(v2[G1Len*i:], pr.Proof[2*G1Len:3*G1Len])
	}
	ok1, ok2 := make([]byte, 2*n), make([]byte, n)
	if C.drx_rep_verify(x.c, C.size_t(2*n), 1, u8(r1), u8(c1), u8(b1), u8(p1), u8(v1), u8(ok1)) != 0 {
		return false, lastErr()
	}
	// the two-term relation reads (r_vi, r_k) in the order they stand in the proof
	if C.drx_rep_verify(x.c, C.size_t(n), 2, u8(r1), u8(c2), u8(b2), u8(p2), u8(v2), u8(ok2)) != 0 {
		return false, lastErr()
	}
	if bytes.IndexByte(ok1, 0) >= 0 || bytes.IndexByte(ok2, 0) >= 0 {
		return false, nil
	}
	return true, nil
}
