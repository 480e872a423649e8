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
  std::printf(failures ? "host_test: %d FAILED\n" : "host_test: all checks passed\n", failures);
  return failures ? 1 : 0;
}
