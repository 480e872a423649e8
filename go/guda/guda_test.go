package guda

// Parity tests in the shape of the reference's own (guda_test.go:16-133,312-375; reference_test.go:11-342;
// fused_gelu_test.go:81-170).  They need a B200, the built libguda_b200.so and a Go toolchain.

import (
	"math"
	"testing"
)

func TestMallocMemcpyRoundTrip(t *testing.T) { // guda_test.go:83-133
	const n = 1024
	h := make([]float32, n)
	for i := range h {
		h[i] = float32(i)
	}
	d1, err := Malloc(n * 4)
	if err != nil {
		t.Fatal(err)
	}
	d2, _ := Malloc(n * 4)
	if err := Memcpy(d1, h, n*4, MemcpyHostToDevice); err != nil {
		t.Fatal(err)
	}
	if err := Memcpy(d2, d1, n*4, MemcpyDeviceToDevice); err != nil {
		t.Fatal(err)
	}
	out := make([]float32, n)
	if err := Memcpy(out, d2, n*4, MemcpyDeviceToHost); err != nil {
		t.Fatal(err)
	}
	for i := range out {
		if out[i] != h[i] {
			t.Fatalf("element %d: got %v want %v", i, out[i], h[i])
		}
	}
	if err := Free(d1); err != nil {
		t.Fatal(err)
	}
	if err := Free(d1); err != ErrDoubleFree { // guda_test.go:314-324
		t.Fatalf("second Free: got %v, want ErrDoubleFree", err)
	}
	Free(d2)
}

func TestDeviceAPI(t *testing.T) { // guda_test.go:326-337
	if GetDeviceCount() != 1 {
		t.Fatal("GetDeviceCount must stay 1")
	}
	if err := SetDevice(1); err == nil {
		t.Fatal("SetDevice(1) must fail")
	}
}

func TestAXPYDOTBaselineKAT(t *testing.T) { // baseline_test.go:34-87 -> baseline_20250807_221713.json
	const n = 1000
	x, y := make([]float32, n), make([]float32, n)
	for i := range x {
		x[i] = float32(i) * 0.001
		y[i] = float32(i) * 0.002
	}
	dx, _ := Malloc(n * 4)
	dy, _ := Malloc(n * 4)
	dx.CopyFrom(x)
	dy.CopyFrom(y)
	AXPY(2.5, dx, dy, n)
	Synchronize()
	got := dy.Float32()
	for i := range got {
		want := float32(math.FMA(2.5, float64(x[i]), float64(y[i]))) // one rounding per element
		if got[i] != want {
			t.Fatalf("AXPY[%d] = %v, want %v", i, got[i], want)
		}
	}
	Free(dx)
	Free(dy)
}

func TestFusedGELUClosedForm(t *testing.T) { // fused_gelu_test.go:95-160
	const m, n, k = 8, 8, 4
	a, b, bias := make([]float32, m*k), make([]float32, k*n), make([]float32, n)
	for i := 0; i < m; i++ {
		for j := 0; j < k; j++ {
			a[i*k+j] = 0.1
			if i == j {
				a[i*k+j] = 1.0
			}
		}
	}
	for i := 0; i < k; i++ {
		for j := 0; j < n; j++ {
			b[i*n+j] = float32(j) * 0.1
		}
	}
	for j := range bias {
		bias[j] = float32(j) * 0.05
	}
	da, _ := Malloc(len(a) * 4)
	db, _ := Malloc(len(b) * 4)
	dc, _ := Malloc(m * n * 4)
	dbias, _ := Malloc(n * 4)
	da.CopyFrom(a)
	db.CopyFrom(b)
	dbias.CopyFrom(bias)
	if err := FusedGEMMBiasGELU(false, false, m, n, k, 1, da, k, db, n, 0, dc, n, dbias); err != nil {
		t.Fatal(err)
	}
	Synchronize()
	c := dc.Float32()
	for i := 0; i < m; i++ {
		for j := 0; j < n; j++ {
			var s float32
			for l := 0; l < k; l++ {
				s += a[i*k+l] * b[l*n+j]
			}
			x := float64(s + bias[j])
			want := 0.5 * x * (1 + math.Tanh(math.Sqrt(2/math.Pi)*(x+0.044715*x*x*x)))
			if math.Abs(float64(c[i*n+j])-want) > 0.015 {
				t.Fatalf("C[%d,%d] = %v, want %v", i, j, c[i*n+j], want)
			}
		}
	}
	if err := FusedGEMMBiasGELU(true, false, m, n, k, 1, da, k, db, n, 0, dc, n, dbias); err != ErrNotSupported {
		t.Fatalf("transA: got %v, want ErrNotSupported", err)
	}
}

func TestGEMMPanics(t *testing.T) { // compute/sgemm.go:30-71
	defer func() {
		if r := recover(); r != "blas: m < 0" {
			t.Fatalf("got %v", r)
		}
	}()
	var z DevicePtr
	GEMM(false, false, -1, 1, 1, 1, z, 1, z, 1, 0, z, 1)
}
