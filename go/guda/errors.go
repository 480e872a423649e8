package guda

import (
	"errors"
	"fmt"
)

// ErrorType mirrors errors.go:11-24 (same iota order; the C ABI uses value+1 so that 0 is success).
type ErrorType int

const (
	ErrTypeMemory ErrorType = iota
	ErrTypeInvalidArg
	ErrTypeExecution
	ErrTypeNumerical
	ErrTypeDevice
	ErrTypeNotImplemented
)

var errorTypeNames = [...]string{"Memory", "InvalidArgument", "Execution", "Numerical", "Device", "NotImplemented"}

func (t ErrorType) String() string {
	if int(t) >= 0 && int(t) < len(errorTypeNames) {
		return errorTypeNames[t]
	}
	return "Unknown"
}

// GUDAError mirrors errors.go:27-44.
type GUDAError struct {
	Type    ErrorType
	Op      string
	Message string
	Err     error
	Context interface{}
}

func (e *GUDAError) Error() string {
	if e.Err != nil {
		return fmt.Sprintf("GUDA %s error in %s: %s (caused by: %v)", e.Type, e.Op, e.Message, e.Err)
	}
	return fmt.Sprintf("GUDA %s error in %s: %s", e.Type, e.Op, e.Message)
}

func (e *GUDAError) Unwrap() error { return e.Err }

func NewMemoryError(op, message string, err error) error {
	return &GUDAError{Type: ErrTypeMemory, Op: op, Message: message, Err: err}
}
func NewInvalidArgError(op, message string) error {
	return &GUDAError{Type: ErrTypeInvalidArg, Op: op, Message: message}
}
func NewExecutionError(op, message string, err error) error {
	return &GUDAError{Type: ErrTypeExecution, Op: op, Message: message, Err: err}
}

// Sentinels (errors.go:113-131, fused.go:11-17).
var (
	ErrOutOfMemory   = NewMemoryError("Malloc", "out of memory", nil)
	ErrInvalidSize   = NewInvalidArgError("Malloc", "size must be positive")
	ErrNullPointer   = NewInvalidArgError("Memory", "null pointer")
	ErrDoubleFree    = NewMemoryError("Free", "double free detected", nil)
	ErrInvalidDevice = NewInvalidArgError("SetDevice", "invalid device ID")
	ErrNoDevice      = &GUDAError{Type: ErrTypeDevice, Op: "Device", Message: "no compute device available"}
	ErrNotSupported  = errors.New("operation not supported")
	// fused.go:12-16
	ErrInvalidShape     = NewInvalidArgError("FusedOp", "invalid shape")
	ErrInvalidParameter = NewInvalidArgError("FusedOp", "invalid parameter")
	ErrInvalidAxis      = NewInvalidArgError("FusedOp", "invalid axis")
)

func isType(err error, t ErrorType) bool {
	var g *GUDAError
	return errors.As(err, &g) && g.Type == t
}

func IsMemoryError(err error) bool         { return isType(err, ErrTypeMemory) }
func IsInvalidArgError(err error) bool     { return isType(err, ErrTypeInvalidArg) }
func IsExecutionError(err error) bool      { return isType(err, ErrTypeExecution) }
func IsNumericalError(err error) bool      { return isType(err, ErrTypeNumerical) }
func IsDeviceError(err error) bool         { return isType(err, ErrTypeDevice) }
func IsNotImplementedError(err error) bool { return isType(err, ErrTypeNotImplemented) }
