package guda

/*
#include "guda_b200.h"
*/
import "C"

import (
	"fmt"
	"unsafe"
)

// MemcpyKind mirrors memory.go:13-21 (same iota order as guda_memcpy_kind_t).
type MemcpyKind int

const (
	MemcpyHostToHost MemcpyKind = iota
	MemcpyHostToDevice
	MemcpyDeviceToHost
	MemcpyDeviceToDevice
	MemcpyDefault
)

// DevicePtr mirrors guda.go:95-99.  ptr is now a real device address (HBM), so the typed views
// Float32()/Float64()/Int32()/Byte() are device-to-host SNAPSHOTS: reading works exactly as in
// the reference, but a write through the returned slice no longer reaches the device -- use
// Memcpy or CopyFrom (SURVEY.md H1).
type DevicePtr struct {
	ptr    unsafe.Pointer
	size   int
	offset int
}

// Malloc mirrors memory.go:57-59 / guda.go:140: real device allocation behind the reference's
// pool semantics (64-byte accounting, first-fit reuse, stats).
func Malloc(size int) (DevicePtr, error) { return defaultContext.Malloc(size) }

// Free mirrors memory.go:64-66: unknown pointer -> Memory error, second Free -> ErrDoubleFree.
func Free(ptr DevicePtr) error { return defaultContext.Free(ptr) }

// Memcpy mirrors memory.go:81-145 (same accepted dst/src types); kind is now honoured.
func Memcpy(dst, src interface{}, size int, kind MemcpyKind) error {
	return defaultContext.Memcpy(dst, src, size, kind)
}

func (ctx *Context) Malloc(size int) (DevicePtr, error) {
	if size < 0 {
		return DevicePtr{}, ErrInvalidSize
	}
	var p unsafe.Pointer
	if err := status(C.guda_malloc(&p, C.size_t(size)), "Malloc"); err != nil {
		return DevicePtr{}, err
	}
	return DevicePtr{ptr: p, size: size}, nil
}

func (ctx *Context) Free(ptr DevicePtr) error {
	return status(C.guda_free(ptr.ptr), "Free")
}

// GetStats mirrors MemoryPool.GetStats (memory.go:230-234).
func (ctx *Context) GetStats() (allocated, peak int64) {
	var a, p C.int64_t
	C.guda_pool_stats(&a, &p)
	return int64(a), int64(p)
}

func hostPointer(v interface{}, op, which string) (unsafe.Pointer, bool, error) {
	switch d := v.(type) {
	case DevicePtr:
		return d.ptr, true, nil
	case unsafe.Pointer:
		return d, false, nil
	case []byte:
		if len(d) > 0 {
			return unsafe.Pointer(&d[0]), false, nil
		}
	case []float32:
		if len(d) > 0 {
			return unsafe.Pointer(&d[0]), false, nil
		}
	case []float64:
		if len(d) > 0 {
			return unsafe.Pointer(&d[0]), false, nil
		}
	case []int32:
		if len(d) > 0 {
			return unsafe.Pointer(&d[0]), false, nil
		}
	case []uint16:
		if len(d) > 0 {
			return unsafe.Pointer(&d[0]), false, nil
		}
	default:
		return nil, false, NewInvalidArgError(op, fmt.Sprintf("unsupported %s type: %T", which, v))
	}
	return nil, false, nil
}

func (ctx *Context) Memcpy(dst, src interface{}, size int, kind MemcpyKind) error {
	dp, ddev, err := hostPointer(dst, "Memcpy", "dst")
	if err != nil {
		return err
	}
	sp, sdev, err := hostPointer(src, "Memcpy", "src")
	if err != nil {
		return err
	}
	if dp == nil || sp == nil || size <= 0 {
		return nil // memory.go:140-142
	}
	// The reference ignores kind (memory.go:81-84) and its own callers pass kinds that do not always match the
	// operands, so for typed operands the direction is derived from the operand types; a raw unsafe.Pointer
	// operand is taken to be what kind says it is.
	_, dRaw := dst.(unsafe.Pointer)
	_, sRaw := src.(unsafe.Pointer)
	if !(dRaw || sRaw) || kind == MemcpyDefault {
		switch {
		case ddev && sdev:
			kind = MemcpyDeviceToDevice
		case ddev:
			kind = MemcpyHostToDevice
		case sdev:
			kind = MemcpyDeviceToHost
		default:
			kind = MemcpyHostToHost
		}
	}
	return status(C.guda_memcpy(dp, sp, C.size_t(size), C.int(kind), nil), "Memcpy")
}

// MemcpyAsync (docs/manual/07-runtime-api.md:179-192): stream-ordered copy; host memory should be pinned.
func MemcpyAsync(dst, src interface{}, size int, kind MemcpyKind, stream *Stream) error {
	dp, _, err := hostPointer(dst, "MemcpyAsync", "dst")
	if err != nil {
		return err
	}
	sp, _, err := hostPointer(src, "MemcpyAsync", "src")
	if err != nil {
		return err
	}
	if dp == nil || sp == nil || size <= 0 {
		return nil
	}
	return status(C.guda_memcpy_async(dp, sp, C.size_t(size), C.int(kind), stream.handle()), "MemcpyAsync")
}

// MallocHost / FreeHost: pinned host staging for MemcpyAsync (C memory, so it may outlive the cgo call).
func MallocHost(size int) (unsafe.Pointer, error) {
	var p unsafe.Pointer
	err := status(C.guda_malloc_host(&p, C.size_t(size)), "MallocHost")
	return p, err
}
func FreeHost(p unsafe.Pointer) error { return status(C.guda_free_host(p), "FreeHost") }

func Memset(ptr DevicePtr, value int, size int) error {
	return status(C.guda_memset(ptr.ptr, C.int(value), C.size_t(size), nil), "Memset")
}

func MemGetInfo() (free, total uint64, err error) {
	var f, t C.size_t
	err = status(C.guda_mem_get_info(&f, &t), "MemGetInfo")
	return uint64(f), uint64(t), err
}

func snapshot[T any](d DevicePtr, elem int) []T {
	if d.ptr == nil {
		return nil
	}
	n := d.size / elem
	out := make([]T, n)
	if n > 0 {
		if err := status(C.guda_memcpy(unsafe.Pointer(&out[0]), d.ptr, C.size_t(n*elem), C.GUDA_MEMCPY_D2H, nil), "Memcpy"); err != nil {
			panic(err)
		}
	}
	return out
}

func (d DevicePtr) Float32() []float32 { return snapshot[float32](d, 4) } // memory.go:246
func (d DevicePtr) Float64() []float64 { return snapshot[float64](d, 8) } // memory.go:261
func (d DevicePtr) Int32() []int32     { return snapshot[int32](d, 4) }   // memory.go:276
func (d DevicePtr) Byte() []byte       { return snapshot[byte](d, 1) }    // memory.go:291

// CopyFrom is the write half the live views used to give: host slice -> device.
func (d DevicePtr) CopyFrom(src interface{}) error {
	sp, _, err := hostPointer(src, "CopyFrom", "src")
	if err != nil || sp == nil {
		return err
	}
	var n int
	switch s := src.(type) {
	case []byte:
		n = len(s)
	case []float32:
		n = 4 * len(s)
	case []float64:
		n = 8 * len(s)
	case []int32:
		n = 4 * len(s)
	case []uint16:
		n = 2 * len(s)
	}
	if n > d.size {
		return NewInvalidArgError("CopyFrom", "source larger than the device buffer")
	}
	return status(C.guda_memcpy(d.ptr, sp, C.size_t(n), C.GUDA_MEMCPY_H2D, nil), "CopyFrom")
}

// Offset mirrors memory.go:303-309: an alias into the parent block; must not be freed.
func (d DevicePtr) Offset(bytes int) DevicePtr {
	return DevicePtr{ptr: unsafe.Add(d.ptr, bytes), size: d.size - bytes, offset: d.offset + bytes}
}

func (d DevicePtr) Size() int { return d.size } // memory.go:312
