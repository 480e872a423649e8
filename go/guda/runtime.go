package guda

/*
#include "guda_b200.h"
*/
import "C"

import (
	"fmt"
	"sync"
)

// Device mirrors guda.go:34-40; NumCores is the SM count, MaxThreads the resident-thread capacity.
type Device struct {
	ID         int
	Name       string
	TotalMem   uint64
	NumCores   int
	MaxThreads int
}

// Context mirrors guda.go:43-49.  The memory pool and the streams live in the C library.
type Context struct {
	device        *Device
	mu            sync.Mutex
	streams       map[*Stream]struct{}
	defaultStream *Stream
}

// Stream mirrors guda.go:57-62: an in-order asynchronous queue -- here a CUDA stream.
type Stream struct {
	h   C.guda_stream_t
	ctx *Context
}

// Event is new (the manual documents it, docs/manual/07-runtime-api.md:122; the reference never implemented it).
type Event struct{ h C.guda_event_t }

var (
	defaultDevice  *Device
	defaultContext *Context
	initOnce       sync.Once
	initErr        error
)

func init() {
	initOnce.Do(func() {
		defaultContext = &Context{streams: make(map[*Stream]struct{})}
		defaultContext.defaultStream = &Stream{h: nil, ctx: defaultContext} // NULL handle = the library's default stream
		if initErr = status(C.guda_init(0), "Init"); initErr != nil {
			return // no CUDA device: every later call fails loudly with a Device error (there is no CPU fallback)
		}
		var p C.guda_device_props_t
		if initErr = status(C.guda_device_properties(0, &p), "GetDeviceProperties"); initErr != nil {
			return
		}
		defaultDevice = &Device{ID: int(p.id), Name: C.GoString(&p.name[0]), TotalMem: uint64(p.total_mem),
			NumCores: int(p.sm_count), MaxThreads: int(p.max_threads)}
		defaultContext.device = defaultDevice
	})
}

func GetDevice() *Device { return defaultDevice } // guda.go:209

// SetDevice mirrors guda.go:216-221: the single-device API accepts only device 0.
func SetDevice(id int) error {
	if id != 0 {
		return ErrInvalidDevice
	}
	return nil
}

// GetDeviceCount stays 1 (guda.go:227-229, pinned by guda_test.go:332-335); the multi-GPU forms are additive
// (sharded.go) and run one process per GPU.
func GetDeviceCount() int { return 1 }

// PhysicalDeviceCount is new: the number of CUDA devices visible to this process.
func PhysicalDeviceCount() (int, error) {
	var n C.int
	err := status(C.guda_device_count(&n), "PhysicalDeviceCount")
	return int(n), err
}

func GetDeviceProperties(id int) (*Device, error) {
	if id != 0 {
		return nil, NewInvalidArgError("GetDeviceProperties", fmt.Sprintf("invalid device ID: %d", id))
	}
	return defaultDevice, nil
}

func Synchronize() error { return defaultContext.Synchronize() } // guda.go:200

func (ctx *Context) Synchronize() error { // guda.go:278-283
	return status(C.guda_device_synchronize(), "Synchronize")
}

func (ctx *Context) CreateStream() *Stream { // guda.go:242-256
	s := &Stream{ctx: ctx}
	if err := status(C.guda_stream_create(&s.h), "CreateStream"); err != nil {
		panic(err)
	}
	ctx.mu.Lock()
	ctx.streams[s] = struct{}{}
	ctx.mu.Unlock()
	return s
}

func (s *Stream) Synchronize() { // guda.go:297-299
	if err := status(C.guda_stream_synchronize(s.handle()), "StreamSynchronize"); err != nil {
		panic(err)
	}
}

// Submit mirrors guda.go:302-305 for host closures: the closure runs after all work already queued on the
// stream has finished (the reference ran it on the stream's worker goroutine).
func (s *Stream) Submit(task func()) {
	s.Synchronize()
	task()
}

func (s *Stream) Destroy() error {
	if s.h == nil {
		return nil
	}
	s.ctx.mu.Lock()
	delete(s.ctx.streams, s)
	s.ctx.mu.Unlock()
	return status(C.guda_stream_destroy(s.h), "StreamDestroy")
}

func (s *Stream) WaitEvent(e *Event) error {
	return status(C.guda_stream_wait_event(s.handle(), e.h), "StreamWaitEvent")
}

// CUDA-runtime-style aliases the manual documents (docs/manual/07-runtime-api.md:122,179-192).
func StreamCreate() (*Stream, error)      { return defaultContext.CreateStream(), nil }
func StreamDestroy(s *Stream) error       { return s.Destroy() }
func StreamSynchronize(s *Stream) error   { s.Synchronize(); return nil }
func StreamWaitEvent(s *Stream, e *Event) error { return s.WaitEvent(e) }

func EventCreate() (*Event, error) {
	e := &Event{}
	return e, status(C.guda_event_create(&e.h), "EventCreate")
}
func (e *Event) Record(s *Stream) error { return status(C.guda_event_record(e.h, s.handle()), "EventRecord") }
func (e *Event) Synchronize() error     { return status(C.guda_event_synchronize(e.h), "EventSynchronize") }
func (e *Event) Query() (bool, error) {
	var done C.int
	err := status(C.guda_event_query(e.h, &done), "EventQuery")
	return done != 0, err
}
func (e *Event) Destroy() error { return status(C.guda_event_destroy(e.h), "EventDestroy") }

// ElapsedTime returns milliseconds between two recorded events.
func ElapsedTime(start, stop *Event) (float32, error) {
	var ms C.float
	err := status(C.guda_event_elapsed_ms(start.h, stop.h, &ms), "EventElapsedTime")
	return float32(ms), err
}
