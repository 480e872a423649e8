package guda

/*
#include "guda_b200.h"
*/
import "C"

import "unsafe"

// Multi-GPU forms (new, additive; the reference has no multi-device notion: guda.go:227-229).
// One process per GPU.  Rank 0 calls CommUniqueID and distributes the 128 bytes out of band
// (MPI, a file, a socket); every rank then calls CommInitRank.

const CommIDBytes = 128

type Comm struct{ h C.guda_comm_t }

func CommUniqueID() ([CommIDBytes]byte, error) {
	var id [CommIDBytes]byte
	err := status(C.guda_comm_unique_id((*C.uint8_t)(unsafe.Pointer(&id[0]))), "CommUniqueID")
	return id, err
}

func CommInitRank(nranks, rank int, id [CommIDBytes]byte) (*Comm, error) {
	c := &Comm{}
	err := status(C.guda_comm_init_rank(&c.h, C.int(nranks), C.int(rank), (*C.uint8_t)(unsafe.Pointer(&id[0]))), "CommInitRank")
	return c, err
}

func (c *Comm) Destroy() error { return status(C.guda_comm_destroy(c.h), "CommDestroy") }

// GEMMSharded: 2-D block-partitioned SGEMM over a pr x pc process grid (rank = r*pc + c).  This rank holds the A row
// panel (m/pr x k) and the B column panel (k x n/pc); it computes its C block with full K locally (same per-element
// k-order as one GPU: bit-identical) and the blocks are all-gathered so that cFull is complete on every rank.
func (c *Comm) GEMMSharded(pr, pc, m, n, k int, alpha float32, aPanel DevicePtr, lda int, bPanel DevicePtr, ldb int,
	cBlock DevicePtr, ldcBlock int, cFull DevicePtr, ldcFull int) error {
	return status(C.guda_sgemm_sharded(c.h, C.int(pr), C.int(pc), C.int(m), C.int(n), C.int(k), C.float(alpha),
		(*C.float)(aPanel.ptr), C.int(lda), (*C.float)(bPanel.ptr), C.int(ldb), (*C.float)(cBlock.ptr), C.int(ldcBlock),
		(*C.float)(cFull.ptr), C.int(ldcFull), nil), "GEMMSharded")
}

// DOTSharded: per-GPU partial dot products combined by an all-reduce; every rank gets the sum.
func (c *Comm) DOTSharded(x, y DevicePtr, nLocal int) (float32, error) {
	var r C.float
	err := status(C.guda_sdot_sharded(c.h, C.int(nLocal), (*C.float)(x.ptr), (*C.float)(y.ptr), &r, nil), "DOTSharded")
	return float32(r), err
}

// DgemmSharded: the fp64 twin of GEMMSharded (compute/dgemm.go:25 partitioned 2-D with full K per rank; the FP64 tensor-pipe
// kernel's epilogue stores each finished tile into every rank's cFull).
func (c *Comm) DgemmSharded(pr, pc, m, n, k int, alpha float64, aPanel DevicePtr, lda int, bPanel DevicePtr, ldb int,
	cBlock DevicePtr, ldcBlock int, cFull DevicePtr, ldcFull int) error {
	return status(C.guda_dgemm_sharded(c.h, C.int(pr), C.int(pc), C.int(m), C.int(n), C.int(k), C.double(alpha),
		(*C.double)(aPanel.ptr), C.int(lda), (*C.double)(bPanel.ptr), C.int(ldb), (*C.double)(cBlock.ptr), C.int(ldcBlock),
		(*C.double)(cFull.ptr), C.int(ldcFull), nil), "DgemmSharded")
}

// AXPYSharded: y += alpha * x on this rank's contiguous chunk of a sharded vector (math.go:39); elementwise, no collective.
func (c *Comm) AXPYSharded(alpha float32, x, y DevicePtr, nLocal int) error {
	return status(C.guda_saxpy_sharded(c.h, C.int(nLocal), C.float(alpha), (*C.float)(x.ptr), (*C.float)(y.ptr), nil), "AXPYSharded")
}
