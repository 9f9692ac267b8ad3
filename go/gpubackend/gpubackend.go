//go:build gpu
// +build gpu

// Package gpubackend binds libdrynx_gpu.so (include/drynx_gpu.h) for the Drynx hot path: batched ElGamal over
// bn256.G1, homomorphic aggregation, table-based decryption. Range proofs are in rangeproof.go, the other
// verifying-node operations in extras.go. Every method replaces one goroutine fan-out of the reference by one
// batched call; inputs and outputs are the reference's own types.
//
// NOT COMPILED in the repository that carries it (no Go toolchain there); see ../README.md.
package gpubackend

/*
#cgo LDFLAGS: -ldrynx_gpu -lcudart
#include <stdlib.h>
#include "drynx_gpu.h"
*/
import "C"

import (
	"errors"
	"runtime"
	"unsafe"

	libunlynx "github.com/ldsec/unlynx/lib"
	"go.dedis.ch/kyber/v3"
	"go.dedis.ch/kyber/v3/pairing/bn256"
	"go.dedis.ch/kyber/v3/util/random"
)

// Sizes of the wire forms (kyber pairing/bn256 MarshalBinary).
const (
	ScalarLen = 32
	G1Len     = 64
	G2Len     = 128
	GTLen     = 384
	CtLen     = 2 * G1Len
)

// Ctx is one device context. Calls on one Ctx are serialised inside the library; use one Ctx per GPU (or several
// per GPU for goroutines that should overlap).
type Ctx struct{ c *C.drx_ctx }

func lastErr() error { return errors.New(C.GoString(C.drx_last_error())) }

func u8(b []byte) *C.uint8_t {
	if len(b) == 0 {
		return nil
	}
	return (*C.uint8_t)(unsafe.Pointer(&b[0]))
}

// New creates a context on CUDA device `device` and hands it the suite's generators, so that no generator
// convention is hard-coded on the device (replaces the package init of lib/suite.go:12-15 for the GPU side).
func New(device int) (*Ctx, error) {
	g1, err := libunlynx.SuiTe.Point().Base().MarshalBinary()
	if err != nil {
		return nil, err
	}
	g2, err := bn256.NewSuite().G2().Point().Base().MarshalBinary()
	if err != nil {
		return nil, err
	}
	var c *C.drx_ctx
	if rc := C.drx_init(&c, C.int(device), u8(g1), u8(g2)); rc != 0 {
		return nil, lastErr()
	}
	x := &Ctx{c}
	runtime.SetFinalizer(x, func(x *Ctx) { x.Close() })
	return x, nil
}

// Close frees the device context.
func (x *Ctx) Close() {
	if x.c != nil {
		C.drx_free(x.c)
		x.c = nil
	}
}

// SetPubKey registers the collective key P (builds its fixed-base table).
func (x *Ctx) SetPubKey(p kyber.Point) error {
	b, err := p.MarshalBinary()
	if err != nil {
		return err
	}
	if C.drx_set_pubkey(x.c, u8(b)) != 0 {
		return lastErr()
	}
	return nil
}

// EncryptIntVectorGetRs replaces the goroutine fan-out over libunlynx.EncryptIntGetR
// (lib/encoding/mean.go:29-39, variance.go:36, frequency_count.go:32-41, logistic_regression.go:420).
// The host keeps drawing the randomness, exactly as libunlynx does (Pick over random.New()).
func (x *Ctx) EncryptIntVectorGetRs(ms []int64) (libunlynx.CipherVector, []kyber.Scalar, error) {
	n := len(ms)
	if n == 0 {
		return libunlynx.CipherVector{}, nil, nil
	}
	rs := make([]kyber.Scalar, n)
	rb := make([]byte, ScalarLen*n)
	for i := range rs {
		rs[i] = libunlynx.SuiTe.Scalar().Pick(random.New())
		b, err := rs[i].MarshalBinary()
		if err != nil {
			return nil, nil, err
		}
		copy(rb[ScalarLen*i:], b)
	}
	ct := make([]byte, CtLen*n)
	if C.drx_encrypt(x.c, C.size_t(n), (*C.int64_t)(unsafe.Pointer(&ms[0])), u8(rb), u8(ct)) != 0 {
		return nil, nil, lastErr()
	}
	cv := libunlynx.NewCipherVector(n)
	cv.FromBytes(ct, n)
	return *cv, rs, nil
}

// Aggregate replaces the FromBytes + CipherVector.Add loop of the root node
// (protocols/data_collection_protocol.go:145-166); dps[d] is ResponseDPBytes.Data[group] of provider d, v the
// number of ciphertexts in each. The pointer table lives in C memory and the Go slices are pinned for the call
// (cgo forbids passing Go memory that holds Go pointers).
func (x *Ctx) Aggregate(dps [][]byte, v int) ([]byte, error) {
	if len(dps) == 0 || v == 0 {
		return make([]byte, CtLen*v), nil
	}
	ptrs := C.malloc(C.size_t(len(dps)) * C.size_t(unsafe.Sizeof(uintptr(0))))
	if ptrs == nil {
		return nil, errors.New("gpubackend: malloc failed")
	}
	defer C.free(ptrs)
	table := unsafe.Slice((**C.uint8_t)(ptrs), len(dps))
	var pinner runtime.Pinner
	defer pinner.Unpin()
	for d := range dps {
		if len(dps[d]) != CtLen*v {
			return nil, errors.New("gpubackend: a provider's vector has the wrong length")
		}
		pinner.Pin(&dps[d][0])
		table[d] = u8(dps[d])
	}
	out := make([]byte, CtLen*v)
	if C.drx_aggregate(x.c, C.size_t(len(dps)), C.size_t(v), (**C.uint8_t)(ptrs), u8(out)) != 0 {
		return nil, lastErr()
	}
	return out, nil
}

// CreateDecryptionTable replaces libunlynx.CreateDecryptionTable (services/api.go:50).
func (x *Ctx) CreateDecryptionTable(limit int64) error {
	if C.drx_dlog_table(x.c, C.int64_t(limit)) != 0 {
		return lastErr()
	}
	return nil
}

// DecryptIntVectorWithNeg replaces libunlynx.DecryptIntVectorWithNeg (lib/encoding/encode_decode.go:224,
// logistic_regression.go:753) and, element-wise, DecryptIntWithNeg (sum.go:38, mean.go:66, variance.go:68 ...).
func (x *Ctx) DecryptIntVectorWithNeg(sk kyber.Scalar, cv libunlynx.CipherVector) ([]int64, error) {
	if len(cv) == 0 {
		return nil, nil
	}
	ct, _, _ := cv.ToBytes()
	skb, err := sk.MarshalBinary()
	if err != nil {
		return nil, err
	}
	out := make([]int64, len(cv))
	if C.drx_decrypt(x.c, C.size_t(len(cv)), u8(skb), u8(ct), 1, (*C.int64_t)(unsafe.Pointer(&out[0])), nil) != 0 {
		return nil, lastErr()
	}
	return out, nil
}

// DecryptCheckZero replaces libunlynx.DecryptCheckZero over a vector (lib/encoding/OR_AND.go:61,114).
func (x *Ctx) DecryptCheckZero(sk kyber.Scalar, cv libunlynx.CipherVector) ([]int64, error) {
	if len(cv) == 0 {
		return nil, nil
	}
	ct, _, _ := cv.ToBytes()
	skb, err := sk.MarshalBinary()
	if err != nil {
		return nil, err
	}
	out := make([]int64, len(cv))
	if C.drx_decrypt_check_zero(x.c, C.size_t(len(cv)), u8(skb), u8(ct), (*C.int64_t)(unsafe.Pointer(&out[0]))) != 0 {
		return nil, lastErr()
	}
	return out, nil
}
