//go:build gpu
// +build gpu

package gpubackend

/*
#include "drynx_gpu.h"
*/
import "C"

import (
	"errors"
	"unsafe"

	libdrynx "github.com/ldsec/drynx/lib"
	libdrynxrange "github.com/ldsec/drynx/lib/range"
	libunlynx "github.com/ldsec/unlynx/lib"
	"go.dedis.ch/kyber/v3"
)

// ProofSize is the length of one serialised range proof in the flat layout the device reads and writes:
//   Commit(128) | Challenge(32) | Zr(32) | D(64) | Zphi(32 l) | for each server i: V_i(128 l) | A_i(384 l) | Zv_i(32 l)
// i.e. the fields of libdrynxrange.RangeProofBytes (lib/range/range_proof.go:42-58) back to back.
func ProofSize(ns int, nd int64) int { return int(C.drx_rp_proof_size(C.int(ns), C.int64_t(nd))) }

// SetSignatures registers the servers' signatures for one output column under `setID`
// (libdrynx.PublishSignatureBytes per server, as produced by InitRangeProofSignature, lib/range/range_proof.go:270-288
// and read by ReadColumn, :598-621). The device builds the fixed-base tables of Y_i, A_{i,k} and e(B, A_{i,k}).
func (x *Ctx) SetSignatures(setID int, u int64, sigs []libdrynx.PublishSignatureBytes) error {
	ns := len(sigs)
	Y := make([]byte, 0, G1Len*ns)
	A := make([]byte, 0, G2Len*ns*int(u))
	for _, s := range sigs {
		yb, err := s.Public.MarshalBinary()
		if err != nil {
			return err
		}
		if len(s.Signature) != G2Len*int(u) {
			return errors.New("gpubackend: signature bytes do not hold u points of G2")
		}
		Y = append(Y, yb...)
		A = append(A, s.Signature...)
	}
	if C.drx_rp_set_sigs(x.c, C.int(setID), C.int(ns), C.int64_t(u), u8(Y), u8(A)) != 0 {
		return lastErr()
	}
	return nil
}

// InitRangeProofSignature replaces libdrynxrange.InitRangeProofSignature (lib/range/range_proof.go:270-288; called per
// server and range at simul/drynx_simul.go:293-295 and by the range-proof tests): the key pair is still drawn in Go (key.NewKeyPair),
// Y = x B and the u signatures A_k = (x + k)^-1 g2 come from one device call. InitRangeProofSignatureDeterministic
// (:249-267) is the same with x = 12.
func (x *Ctx) InitRangeProofSignature(u int64) (libdrynx.PublishSignatureBytes, error) {
	priv := libunlynx.SuiTe.Scalar().Pick(libunlynx.SuiTe.RandomStream())
	return x.initRangeProofSignature(u, priv)
}

func (x *Ctx) InitRangeProofSignatureDeterministic(u int64) (libdrynx.PublishSignatureBytes, error) {
	return x.initRangeProofSignature(u, libunlynx.SuiTe.Scalar().SetInt64(12))
}

func (x *Ctx) initRangeProofSignature(u int64, priv kyber.Scalar) (libdrynx.PublishSignatureBytes, error) {
	var out libdrynx.PublishSignatureBytes
	xb, err := priv.MarshalBinary()
	if err != nil {
		return out, err
	}
	Y, A := make([]byte, G1Len), make([]byte, G2Len*int(u))
	if C.drx_rp_make_sigs(x.c, 1, C.int64_t(u), u8(xb), u8(Y), u8(A)) != 0 {
		return out, lastErr()
	}
	pub := libunlynx.SuiTe.Point()
	if err := pub.UnmarshalBinary(Y); err != nil {
		return out, err
	}
	return libdrynx.PublishSignatureBytes{Signature: A, Public: pub}, nil
}

// drawProofRandomness draws, for one proof with nd digits and ns servers, the scalars
// s_j, t_j, m_j, v_0j .. v_(ns-1)j for j = 0 .. nd-1 in the order CreatePredicateRangeProofForAllServ draws them
// (lib/range/range_proof.go:377-390) from the suite's random stream.
func drawProofRandomness(nd, ns int) ([]byte, error) {
	out := make([]byte, 0, ScalarLen*nd*(3+ns))
	for j := 0; j < nd; j++ {
		for k := 0; k < 3+ns; k++ {
			b, err := libunlynx.SuiTe.Scalar().Pick(libunlynx.SuiTe.RandomStream()).MarshalBinary()
			if err != nil {
				return nil, err
			}
			out = append(out, b...)
		}
	}
	return out, nil
}

// CreatePredicateRangeProofListForAllServers replaces the function of the same name
// (lib/range/range_proof.go:305-317; caller protocols/data_collection_protocol.go:278-292). setIDs[i] names the
// signature set registered for cps[i].Sigs (SetSignatures); a proof with U = L = 0 passes setID -1 and is returned as
// the bare commitment, as the reference does (:322-324).
func (x *Ctx) CreatePredicateRangeProofListForAllServers(cps []libdrynxrange.CreateProof, setIDs []int32) ([]libdrynxrange.RangeProof, error) {
	n := len(cps)
	if n == 0 {
		return nil, nil
	}
	if len(setIDs) != n {
		return nil, errors.New("gpubackend: one set id per proof")
	}
	ls, secrets := make([]int64, n), make([]int64, n)
	rb, ct := make([]byte, ScalarLen*n), make([]byte, CtLen*n)
	proofOff, rndOff := make([]C.size_t, n+1), make([]C.size_t, n+1)
	nds := make([]int64, n)
	var rnd []byte
	for i, cp := range cps {
		ls[i], secrets[i] = cp.L, cp.Secret
		b, err := cp.R.MarshalBinary()
		if err != nil {
			return nil, err
		}
		copy(rb[ScalarLen*i:], b)
		cb, err := cp.Cipher.ToBytes()
		if err != nil {
			return nil, err
		}
		copy(ct[CtLen*i:], cb)
		size := 128
		if setIDs[i] >= 0 {
			nds[i] = int64(C.drx_rp_ndigits(C.int64_t(cp.Secret), C.int64_t(cp.U), C.int64_t(cp.L)))
			size = ProofSize(len(cp.Sigs), nds[i])
			r, err := drawProofRandomness(int(nds[i]), len(cp.Sigs))
			if err != nil {
				return nil, err
			}
			rnd = append(rnd, r...)
		}
		proofOff[i+1] = proofOff[i] + C.size_t(size)
		rndOff[i+1] = C.size_t(len(rnd))
	}
	proofs := make([]byte, int(proofOff[n]))
	if C.drx_rp_create(x.c, C.size_t(n), (*C.int32_t)(unsafe.Pointer(&setIDs[0])), (*C.int64_t)(unsafe.Pointer(&ls[0])),
		(*C.int64_t)(unsafe.Pointer(&secrets[0])), u8(rb), u8(ct), u8(rnd), &rndOff[0], u8(proofs), &proofOff[0]) != 0 {
		return nil, lastErr()
	}
	out := make([]libdrynxrange.RangeProof, n)
	for i := range cps {
		flat := proofs[int(proofOff[i]):int(proofOff[i+1])]
		if setIDs[i] < 0 {
			out[i] = libdrynxrange.RangeProof{Commit: cps[i].Cipher}
			continue
		}
		out[i].FromBytes(unflatten(flat, len(cps[i].Sigs), int(nds[i])))
	}
	return out, nil
}

// unflatten slices the flat layout into the reference's RangeProofBytes (no copies of the field bytes).
func unflatten(flat []byte, ns, nd int) libdrynxrange.RangeProofBytes {
	commit := flat[0:128]
	challenge, zr, d := flat[128:160], flat[160:192], flat[192:256]
	zphi := flat[256 : 256+32*nd]
	off := 256 + 32*nd
	v, a, zv := make([][]byte, ns), make([][]byte, ns), make([][]byte, ns)
	for i := 0; i < ns; i++ {
		v[i] = flat[off : off+128*nd]
		off += 128 * nd
		a[i] = flat[off : off+384*nd]
		off += 384 * nd
		zv[i] = flat[off : off+32*nd]
		off += 32 * nd
	}
	return libdrynxrange.RangeProofBytes{Commit: commit, RP: &libdrynxrange.RangeProofDataBytes{
		Challenge: &challenge, Zr: &zr, D: &d, Zphi: &zphi, V: &v, A: &a, Zv: &zv}}
}

// Flatten is the inverse: the bytes of RangeProof.ToBytes() (lib/range/range_proof.go:89-155) in the flat layout.
func Flatten(pb libdrynxrange.RangeProofBytes) []byte {
	out := append([]byte{}, pb.Commit...)
	if pb.RP == nil || pb.RP.Challenge == nil {
		return out
	}
	out = append(out, *pb.RP.Challenge...)
	out = append(out, *pb.RP.Zr...)
	out = append(out, *pb.RP.D...)
	out = append(out, *pb.RP.Zphi...)
	for i := range *pb.RP.V {
		out = append(out, (*pb.RP.V)[i]...)
		out = append(out, (*pb.RP.A)[i]...)
		out = append(out, (*pb.RP.Zv)[i]...)
	}
	return out
}

// RangeProofListVerification replaces the function of the same name (lib/range/range_proof.go:484-501; caller
// lib/proof/structs_proofs.go:171): one batched call, then the AND over the list (:497-499). setIDs[i] is the signature
// set of proof i (-1 for U = L = 0, which the reference accepts, :508-510); ls[i] its digit count.
func (x *Ctx) RangeProofListVerification(list libdrynxrange.RangeProofListBytes, setIDs []int32, ls []int64, P kyber.Point) (bool, error) {
	n := len(*list.Data)
	if n == 0 {
		return true, nil
	}
	if err := x.SetPubKey(P); err != nil {
		return false, err
	}
	off := make([]C.size_t, n+1)
	var flat []byte
	for i, pb := range *list.Data {
		flat = append(flat, Flatten(pb)...)
		off[i+1] = C.size_t(len(flat))
	}
	ok := make([]byte, n)
	if C.drx_rp_verify(x.c, C.size_t(n), (*C.int32_t)(unsafe.Pointer(&setIDs[0])), (*C.int64_t)(unsafe.Pointer(&ls[0])),
		u8(flat), &off[0], u8(ok)) != 0 {
		return false, lastErr()
	}
	for _, b := range ok {
		if b == 0 {
			return false, nil
		}
	}
	return true, nil
}
