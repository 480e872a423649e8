package guda

// The reference's kernel-launch surface (guda.go:66-89,184-196,258-276,310-339; execution.go:9-86): Dim3, ThreadID, Kernel,
// KernelFunc, Launch, LaunchFunc, LaunchStream, LaunchFuncStream.
//
// In the reference a "kernel" is a Go closure that a pool of goroutines runs once per (block, thread) over slices that ARE the
// device buffers (unified host memory).  Here device buffers are HBM allocations and the device cannot call back into Go, so
// the launch entry points exist -- a program written against the reference compiles unchanged -- and report
// ErrTypeNotImplemented, naming the device-kernel entry points that cover the closures the reference itself launches:
//
//	reference closure                              device entry point
//	Add/Multiply/ReLU/Sigmoid (math.go:133-270)    Add, Multiply, ReLU, Sigmoid            (elementwise.go)
//	FusedKernel.Execute        (fused.go:123)      (*FusedKernel).Execute, LinearReLU, GELU (elementwise.go)
//	GEMMBFloat16 / float16 conversions             GEMMBFloat16, GEMMFloat16                (blas.go)
//	ForEach / Map / Reduce     (execution.go:198+) the reductions and FusedKernel ops above
//
// The index helpers (Global, GlobalX/Y/Z, Size) are pure integer functions and behave exactly as in the reference.

// Dim3 is guda.go:66-68.
type Dim3 struct {
	X, Y, Z int
}

// Size is guda.go:330-332.
func (d Dim3) Size() int { return d.X * d.Y * d.Z }

// ThreadID is guda.go:73-78.
type ThreadID struct {
	BlockIdx  Dim3
	ThreadIdx Dim3
	BlockDim  Dim3
	GridDim   Dim3
}

func (tid ThreadID) axis(b, d, t int) int { return b*d + t }

// Global, GlobalX, GlobalY, GlobalZ are guda.go:310-327.
func (tid ThreadID) Global() int  { return tid.axis(tid.BlockIdx.X, tid.BlockDim.X, tid.ThreadIdx.X) }
func (tid ThreadID) GlobalX() int { return tid.axis(tid.BlockIdx.X, tid.BlockDim.X, tid.ThreadIdx.X) }
func (tid ThreadID) GlobalY() int { return tid.axis(tid.BlockIdx.Y, tid.BlockDim.Y, tid.ThreadIdx.Y) }
func (tid ThreadID) GlobalZ() int { return tid.axis(tid.BlockIdx.Z, tid.BlockDim.Z, tid.ThreadIdx.Z) }

// Kernel and KernelFunc are guda.go:83-89,335-337.
type Kernel interface {
	Execute(tid ThreadID, args ...interface{})
}
type KernelFunc func(tid ThreadID, args ...interface{})

func (fn KernelFunc) Execute(tid ThreadID, args ...interface{}) { fn(tid, args...) }

const launchUnsupported = "host closures cannot run on the device: use the device-kernel entry points " +
	"(Add/Multiply/ReLU/Sigmoid, FusedKernel.Execute, LinearReLU, GELU, GEMMBFloat16/GEMMFloat16, the reductions)"

func launchRefused(op string, grid, block Dim3) error {
	if grid.Size() <= 0 || block.Size() <= 0 {
		return &GUDAError{Type: ErrTypeInvalidArg, Op: op, Message: "grid and block dimensions must be positive"}
	}
	return &GUDAError{Type: ErrTypeNotImplemented, Op: op, Message: launchUnsupported}
}

// Launch, LaunchFunc (guda.go:184-196) and their Context / stream forms (guda.go:258-276).
func Launch(kernel Kernel, grid, block Dim3, args ...interface{}) error {
	return defaultContext.Launch(kernel, grid, block, args...)
}
func LaunchFunc(fn KernelFunc, grid, block Dim3, args ...interface{}) error {
	return defaultContext.LaunchFunc(fn, grid, block, args...)
}
func (ctx *Context) Launch(kernel Kernel, grid, block Dim3, args ...interface{}) error {
	return launchRefused("Launch", grid, block)
}
func (ctx *Context) LaunchFunc(fn KernelFunc, grid, block Dim3, args ...interface{}) error {
	return launchRefused("LaunchFunc", grid, block)
}
func (ctx *Context) LaunchStream(kernel Kernel, grid, block Dim3, stream *Stream, args ...interface{}) error {
	return launchRefused("LaunchStream", grid, block)
}
func (ctx *Context) LaunchFuncStream(fn KernelFunc, grid, block Dim3, stream *Stream, args ...interface{}) error {
	return launchRefused("LaunchFuncStream", grid, block)
}
