package guda

/*
#include "guda_b200.h"
*/
import "C"

// ConvParams mirrors conv.go:9-31 field for field.
type ConvParams struct {
	BatchSize, InChannels, InHeight, InWidth int
	OutChannels, KernelHeight, KernelWidth   int
	StrideH, StrideW, PadH, PadW             int
	DilationH, DilationW                     int
	UseBias                                  bool
}

// Validate mirrors conv.go:34-51 (the C library repeats these checks with the same messages).
func (p *ConvParams) Validate() error {
	if p.BatchSize <= 0 || p.InChannels <= 0 || p.InHeight <= 0 || p.InWidth <= 0 {
		return NewInvalidArgError("Conv2D", "invalid input dimensions: batch, channels, height and width must be positive")
	}
	if p.OutChannels <= 0 || p.KernelHeight <= 0 || p.KernelWidth <= 0 {
		return NewInvalidArgError("Conv2D", "invalid kernel dimensions: output channels and kernel size must be positive")
	}
	if p.StrideH <= 0 || p.StrideW <= 0 {
		return NewInvalidArgError("Conv2D", "invalid stride: stride must be positive")
	}
	if p.DilationH <= 0 || p.DilationW <= 0 {
		return NewInvalidArgError("Conv2D", "invalid dilation: dilation must be positive")
	}
	if p.PadH < 0 || p.PadW < 0 {
		return NewInvalidArgError("Conv2D", "invalid padding: padding cannot be negative")
	}
	return nil
}

func (p *ConvParams) OutputHeight() int { // conv.go:54-57
	return (p.InHeight+2*p.PadH-((p.KernelHeight-1)*p.DilationH+1))/p.StrideH + 1
}

func (p *ConvParams) OutputWidth() int { // conv.go:60-63
	return (p.InWidth+2*p.PadW-((p.KernelWidth-1)*p.DilationW+1))/p.StrideW + 1
}

// Conv2D mirrors conv.go:69-81: NCHW input, [OC][IC][KH][KW] kernel, optional bias (a zero DevicePtr means none).
// An implicit GEMM on the GPU (im2col gathered in the tile loader, bias + NCHW placement in the epilogue) with the reference's
// FMA chain per output element; bit-identical to the reference for BatchSize == 1.
func Conv2D(input, kernel, bias, output DevicePtr, params *ConvParams) error {
	if err := params.Validate(); err != nil {
		return err
	}
	cp := C.guda_conv_params_t{
		batch_size: C.int(params.BatchSize), in_channels: C.int(params.InChannels), in_height: C.int(params.InHeight), in_width: C.int(params.InWidth),
		out_channels: C.int(params.OutChannels), kernel_height: C.int(params.KernelHeight), kernel_width: C.int(params.KernelWidth),
		stride_h: C.int(params.StrideH), stride_w: C.int(params.StrideW), pad_h: C.int(params.PadH), pad_w: C.int(params.PadW),
		dilation_h: C.int(params.DilationH), dilation_w: C.int(params.DilationW), use_bias: cbool(params.UseBias),
	}
	return status(C.guda_conv2d((*C.float)(input.ptr), (*C.float)(kernel.ptr), (*C.float)(bias.ptr), (*C.float)(output.ptr), &cp, nil), "Conv2D")
}

// Conv2DDirect mirrors conv.go:231-300: the direct loops (unfused sum over the in-bounds taps in (ic, kh, kw) order, then + bias),
// bit-identical to the reference.  It is the cross-check API; Conv2D is the fast route.
func Conv2DDirect(input, kernel, bias, output DevicePtr, params *ConvParams) error {
	if err := params.Validate(); err != nil {
		return err
	}
	cp := C.guda_conv_params_t{
		batch_size: C.int(params.BatchSize), in_channels: C.int(params.InChannels), in_height: C.int(params.InHeight), in_width: C.int(params.InWidth),
		out_channels: C.int(params.OutChannels), kernel_height: C.int(params.KernelHeight), kernel_width: C.int(params.KernelWidth),
		stride_h: C.int(params.StrideH), stride_w: C.int(params.StrideW), pad_h: C.int(params.PadH), pad_w: C.int(params.PadW),
		dilation_h: C.int(params.DilationH), dilation_w: C.int(params.DilationW), use_bias: cbool(params.UseBias),
	}
	return status(C.guda_conv2d_direct((*C.float)(input.ptr), (*C.float)(kernel.ptr), (*C.float)(bias.ptr), (*C.float)(output.ptr), &cp, nil), "Conv2DDirect")
}
