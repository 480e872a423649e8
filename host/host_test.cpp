// host_test.cpp -- the reference's own test shapes through the C++ host mirror (host/guda.hpp), checked against the
// CPU oracle (TEST INFRASTRUCTURE: oracle/guda_oracle.h).  Mirrors guda_test.go:83-133,312-375, reference_test.go:11-120,
// fused_gelu_test.go:95-160.  Exit code 0 = all parity checks passed; 3 = no CUDA device (the product path never falls back).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "guda.hpp"
#include "../oracle/guda_oracle.h"

static int failures = 0;
#define EXPECT(cond, ...) do { if (!(cond)) { failures++; std::printf("FAIL %s:%d: ", __FILE__, __LINE__); std::printf(__VA_ARGS__); std::printf("\n"); } } while (0)

static std::vector<float> uniform(size_t n, uint64_t seed) { std::vector<float> v(n); orc_generate_f32_masked(v.data(), n, seed, -1.0f, 1.0f); return v; }

int main() {
  using namespace guda;
  try {
    check(guda_init(0), "Init");
  } catch (const GUDAError& e) {
    std::printf("no usable CUDA device: %s\n", e.what());
    return e.type == ErrorType::Device ? 3 : 4;
  }
  // ---- guda_test.go:83-133 H2D -> D2D -> D2H round trip; :312-375 double free, stats
  {
    const int n = 1024;
    std::vector<float> h(n); for (int i = 0; i < n; i++) h[i] = (float)i;
    int64_t a0, p0, a1, p1;
    GetStats(&a0, &p0);
    DevicePtr d1 = Malloc(n * 4), d2 = Malloc(n * 4);
    GetStats(&a1, &p1);
    EXPECT(a1 - a0 == 2 * n * 4 && p1 >= a1, "pool stats after two allocations: %lld", (long long)(a1 - a0));
    Memcpy(d1, h.data(), n * 4); Memcpy(d2, d1, n * 4);
    std::vector<float> out(n); Memcpy(out.data(), d2, n * 4);
    EXPECT(out == h, "round trip");
    Free(d1);
    bool dbl = false; try { Free(d1); } catch (const ErrDoubleFree&) { dbl = true; }
    EXPECT(dbl, "second Free must raise ErrDoubleFree");
    Free(d2);
    bool bad = false; try { SetDevice(1); } catch (const GUDAError&) { bad = true; }
    EXPECT(bad && GetDeviceCount() == 1, "device API");
  }
  // ---- reference_test.go:11-60 AXPY n=1024 alpha=2.5, DOT n=1024
  {
    const int n = 1024;
    auto x = uniform(n, 1), y = uniform(n, 2);
    DevicePtr dx = Malloc(n * 4), dy = Malloc(n * 4);
    dx.CopyFrom(x); dy.CopyFrom(y);
    AXPY(2.5f, dx, dy, n);
    float dot = DOT(dx, dy, n);
    orc_saxpy(n, 2.5f, x.data(), x.size(), 1, y.data(), y.size(), 1);
    EXPECT(dy.Float32() == y, "AXPY bit-exact");
    float want = 0; orc_sdot(n, x.data(), x.size(), 1, y.data(), y.size(), 1, &want);
    EXPECT(std::fabs(dot - want) <= 1e-4f * std::fmax(1.0f, std::fabs(want)), "DOT %g vs %g", dot, want);
    Free(dx); Free(dy);
  }
  // ---- reference_test.go:62-120 GEMM 32^3 alpha=1.5 beta=0.5; arch_aware_gemm_test.go shapes
  for (auto shp : std::vector<std::vector<int>>{{32, 32, 32}, {63, 65, 64}, {128, 128, 128}}) {
    int m = shp[0], n = shp[1], k = shp[2];
    auto a = uniform((size_t)m * k, 3), b = uniform((size_t)k * n, 4), c = uniform((size_t)m * n, 5);
    DevicePtr da = Malloc(a.size() * 4), db = Malloc(b.size() * 4), dc = Malloc(c.size() * 4);
    da.CopyFrom(a); db.CopyFrom(b); dc.CopyFrom(c);
    GEMM(false, false, m, n, k, 1.5f, da, k, db, n, 0.5f, dc, n);
    Synchronize();
    orc_sgemm(0, 0, m, n, k, 1.5f, a.data(), a.size(), k, b.data(), b.size(), n, 0.5f, c.data(), c.size(), n, 0, 1);
    EXPECT(dc.Float32() == c, "GEMM %dx%dx%d bit-exact", m, n, k);
    Free(da); Free(db); Free(dc);
  }
  // ---- fused_gelu_test.go:95-160 closed-form fused GELU; fused.go:324-329 ErrNotSupported
  {
    const int m = 8, n = 8, k = 4;
    std::vector<float> a(m * k, 0.1f), b(k * n), bias(n), c(m * n, 0.0f);
    for (int i = 0; i < k; i++) a[i * k + i] = 1.0f;
    for (int i = 0; i < k; i++) for (int j = 0; j < n; j++) b[i * n + j] = 0.1f * j;
    for (int j = 0; j < n; j++) bias[j] = 0.05f * j;
    DevicePtr da = Malloc(a.size() * 4), db = Malloc(b.size() * 4), dc = Malloc(c.size() * 4), dbias = Malloc(n * 4);
    da.CopyFrom(a); db.CopyFrom(b); dbias.CopyFrom(bias);
    FusedGEMMBiasGELU(false, false, m, n, k, 1.0f, da, k, db, n, 0.0f, dc, n, dbias);
    Synchronize();
    orc_fused_gemm_bias_gelu(0, 0, m, n, k, 1.0f, a.data(), k, b.data(), n, 0.0f, c.data(), n, bias.data(), 0, 1);
    EXPECT(dc.Float32() == c, "fused GEMM+bias+GELU bit-exact");
    bool ns = false; try { FusedGEMMBiasGELU(true, false, m, n, k, 1.0f, da, k, db, n, 0.0f, dc, n, dbias); } catch (const ErrNotSupported&) { ns = true; }
    EXPECT(ns, "transA must raise ErrNotSupported");
    bool pan = false; try { GEMM(false, false, -1, n, k, 1.0f, da, k, db, n, 0.0f, dc, n); } catch (const BlasPanic& e) { pan = std::string(e.what()) == "blas: m < 0"; }
    EXPECT(pan, "m < 0 must panic with the reference's string");
    Free(da); Free(db); Free(dc); Free(dbias);
  }
  // ---- conv_test.go:8-76 TestConv2DBasic (3x3 edge detector, pad 1)
  {
    ConvParams p; p.BatchSize = 1; p.InChannels = 1; p.InHeight = 4; p.InWidth = 4; p.OutChannels = 1; p.KernelHeight = 3; p.KernelWidth = 3;
    p.PadH = p.PadW = 1;
    std::vector<float> in(16), ker{-1, -1, -1, -1, 8, -1, -1, -1, -1};
    for (int i = 0; i < 16; i++) in[i] = (float)(i + 1);
    const std::vector<float> expected{-5, -6, -3, 14, 12, 0, 0, 27, 24, 0, 0, 39, 71, 54, 57, 90};
    DevicePtr di = Malloc(64), dk = Malloc(36), dout = Malloc(64);
    di.CopyFrom(in); dk.CopyFrom(ker);
    Conv2D(di, dk, DevicePtr{}, dout, p);
    EXPECT(dout.Float32() == expected, "Conv2D basic");
    Free(di); Free(dk); Free(dout);
  }
  // ---- Sgemv N / T and Sger (compute/level2float32.go:244-308, :24-65): bit-exact against the oracle
  {
    const int m = 37, n = 75;
    std::vector<float> a = uniform(m * n, 71), xn = uniform(n, 72), xt = uniform(m, 73), yn = uniform(m, 74), yt = uniform(n, 75);
    DevicePtr da = Malloc(a.size() * 4), dxn = Malloc(n * 4), dxt = Malloc(m * 4), dyn = Malloc(m * 4), dyt = Malloc(n * 4);
    da.CopyFrom(a); dxn.CopyFrom(xn); dxt.CopyFrom(xt); dyn.CopyFrom(yn); dyt.CopyFrom(yt);
    Sgemv(false, m, n, 1.5f, da, n, dxn, 1, 0.25f, dyn, 1);
    Sgemv(true, m, n, 1.5f, da, n, dxt, 1, 0.25f, dyt, 1);
    Synchronize();
    orc_sgemv(0, m, n, 1.5f, a.data(), a.size(), n, xn.data(), n, 1, 0.25f, yn.data(), m, 1);
    orc_sgemv(1, m, n, 1.5f, a.data(), a.size(), n, xt.data(), m, 1, 0.25f, yt.data(), n, 1);
    EXPECT(dyn.Float32() == yn, "Sgemv NoTrans bit-exact");
    EXPECT(dyt.Float32() == yt, "Sgemv Trans bit-exact");
    Sger(m, n, -0.75f, dxt, 1, dxn, 1, da, n);
    Synchronize();
    orc_sger(m, n, -0.75f, xt.data(), m, 1, xn.data(), n, 1, a.data(), a.size(), n);
    EXPECT(da.Float32() == a, "Sger bit-exact");
    bool pan = false; try { Sgemv(false, m, n, 1.0f, da, n - 1, dxn, 1, 0.0f, dyn, 1); } catch (const BlasPanic& e) { pan = std::string(e.what()) == "blas: bad leading dimension of A"; }
    EXPECT(pan, "bad lda must panic with the reference's string");
    pan = false; try { Sgemv(false, m, n, 1.0f, da, n, dxt, 1, 0.0f, dyn, 1); } catch (const BlasPanic& e) { pan = std::string(e.what()) == "blas: insufficient length of x"; }
    EXPECT(pan, "short x must panic with the reference's string");
    Free(da); Free(dxn); Free(dxt); Free(dyn); Free(dyt);
  }
  // ---- GEMMHost: host operands, pipelined copies; bit-identical to the oracle
  {
    const int m = 150, n = 70, k = 90;
    std::vector<float> a = uniform(m * k, 91), b = uniform(k * n, 92), c = uniform(m * n, 93), want = c;
    GEMMHost(false, false, m, n, k, 1.5f, a.data(), a.size(), k, b.data(), b.size(), n, 0.5f, c.data(), c.size(), n);
    orc_sgemm(0, 0, m, n, k, 1.5f, a.data(), a.size(), k, b.data(), b.size(), n, 0.5f, want.data(), want.size(), n, 0, 1);
    EXPECT(c == want, "GEMMHost bit-exact");
  }
  // ---- reductions (reduce.go:13-56): index / max / min outputs bit-exact, NaN and signed zero included
  {
    std::vector<float> v = uniform(100003, 81);
    v[777] = 5.0f; v[99000] = 5.0f; v[31] = -7.0f;
    DevicePtr d = Malloc(v.size() * 4); d.CopyFrom(v);
    EXPECT(ArgMax(d, v.size()) == orc_sargmax(v.data(), v.size()) && ArgMax(d, v.size()) == 777, "ArgMax first of equal maxima");
    EXPECT(ArgMin(d, v.size()) == 31, "ArgMin");
    long long want = -1; orc_isamax((int)v.size(), v.data(), v.size(), 1, &want);
    EXPECT(Isamax((int)v.size(), d, 1) == want && want == 31, "Isamax");
    float mx = Max(d, v.size()), omx = orc_smax(v.data(), v.size());
    EXPECT(std::memcmp(&mx, &omx, 4) == 0, "Max bit-exact");
    v[5] = std::nanf(""); v[99999] = std::nanf(""); d.CopyFrom(v);
    mx = Max(d, v.size()); omx = orc_smax(v.data(), v.size());
    EXPECT(std::memcmp(&mx, &omx, 4) == 0, "Max with NaNs follows VMAXPS");
    float mn = Min(d, v.size()), omn = orc_smin(v.data(), v.size());
    EXPECT(std::memcmp(&mn, &omn, 4) == 0, "Min with NaNs follows VMINPS");
    EXPECT(ArgMax(d, 0) == -1 && Max(d, 0) == -INFINITY && Min(d, 0) == INFINITY && Sum(d, 0) == 0.0f, "empty reductions");
    Free(d);
  }
  // ---- tolerance framework (tolerance_test.go:8-240, tolerance_arch_test.go:27-57) through the C++ mirror, device verification vs the oracle
  {
    ToleranceConfig d = DefaultTolerance(), g = GetOperationTolerance("gemm"), ga = GetOperationTolerance("gemm", "arm64");
    EXPECT(d.AbsTol == 1e-7f && d.RelTol == 1e-5f && d.ULPTol == 4 && d.CheckNaN && d.CheckInf, "DefaultTolerance");
    EXPECT(g.AbsTol == 1e-6f && g.ULPTol == 4 && g.KScale == 2.0f && ga.ULPTol == 16 && ga.KScale == 0.0f, "GetOperationTolerance(gemm) sm_100a / arm64");
    ToleranceConfig base; base.AbsTol = 1e-7f; base.RelTol = 1e-6f; base.ULPTol = 2; base.CheckNaN = true;
    ToleranceConfig over; over.RelTol = 1e-4f; over.ULPTol = 16;
    ToleranceConfig m = mergeTolerances(base, over);
    EXPECT(m.AbsTol == 1e-7f && m.RelTol == 1e-4f && m.ULPTol == 16 && m.CheckNaN && !m.CheckInf, "merge keeps non-overridden fields");
    EXPECT(Float32NearEqual(1e-8f, 2e-8f, d) && !Float32NearEqual(1e-6f, 2e-6f, d) && Float32NearEqual(0.0f, -0.0f, d), "Float32NearEqual");
    EXPECT(Float32ULPDiff(1.0f, -1.0f) == 0x7FFFFFFF && Float32ULPDiff(1.0f, std::nextafterf(1.0f, 2.0f)) == 1, "Float32ULPDiff");
    const size_t n = 100003;
    auto e = uniform(n, 77), a = e;
    for (size_t i = 0; i < n; i += 97) a[i] *= 1.0f + 3e-4f;
    a[5] = -a[5]; e[9] = std::nanf(""); a[9] = std::nanf("");
    DevicePtr de = Malloc(n * 4), da = Malloc(n * 4);
    de.CopyFrom(e); da.CopyFrom(a);
    VerificationResult v = VerifyFloat32Array(de, da, (long long)n, g);
    orc_tol_t ot{g.AbsTol, g.RelTol, g.ULPTol, 1, 1};
    orc_verify_t ov;
    orc_verify_f32(e.data(), a.data(), n, &ot, &ov);
    EXPECT(v.NumErrors == ov.num_errors && v.FirstError == ov.first_error && v.MaxULPError == ov.max_ulp && v.TotalItems == (long long)n &&
           std::memcmp(&v.MaxAbsError, &ov.max_abs, 4) == 0 && std::memcmp(&v.MaxRelError, &ov.max_rel, 4) == 0 && v.NumErrors > 0,
           "VerifyFloat32Array on the device == oracle: %lld errors", v.NumErrors);
    EXPECT(!v.IsAcceptable(g), "IsAcceptable");
    Free(de); Free(da);
    bool refused = false;
    try { Launch([](ThreadID) {}, Dim3{4, 1, 1}, Dim3{64, 1, 1}); } catch (const GUDAError& err) { refused = err.type == ErrorType::NotImplemented; }
    ThreadID t{Dim3{3, 2, 1}, Dim3{5, 1, 0}, Dim3{64, 4, 1}, Dim3{8, 8, 1}};
    EXPECT(refused && t.Global() == 197 && t.GlobalY() == 9 && (Dim3{4, 2, 3}).Size() == 24, "Launch surface");
  }
  std::printf(failures ? "host_test: %d FAILED\n" : "host_test: all checks passed\n", failures);
  return failures ? 1 : 0;
}
