#!/usr/bin/env python
"""bench.py -- GUDA BLAS hot path on B200: throughput, roofline fraction and the CPU baseline beside it.

  python bench.py --gpus N --steps K --warmup W            our arm (N>1: launched by torchrun, one rank per GPU)
  python bench.py --impl reference --gpus N --steps K ...  the reference's CPU path (oracle port) on the host cores

Headline workload (BASELINE.json configs[4], the large-GEMM config the north star scales over 1/2/4/8 GPUs):
SGEMM 65536^3 fp32, 2-D block-partitioned, C all-gathered over NVLink at N>1 -- STRONG scaling (total work fixed).
A "step" is one full C = A*B.  The other BASELINE configs (C1 SGEMM 1024^3, C2 SAXPY/SDOT 16M, C3 fused FFN,
C4 DGEMM 16384^3) run once per invocation at N=1 and are reported in the same JSON line under "suite", each with
its own roofline.  Prints ONE JSON line on rank 0.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--size", type=int, default=int(os.environ.get("GUDA_BENCH_SIZE", "65536")), help="N of the headline N^3 SGEMM")
    ap.add_argument("--mode", default=os.environ.get("GUDA_BENCH_SGEMM_MODE", "exact"), choices=["exact", "3xtf32"])
    ap.add_argument("--no-suite", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    return ap.parse_args()


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return {"hbm_gbs": p["hbm_gbs"], "bf16_tflops": p["bf16_tflops"], "bf16_tflops_sustained": p.get("bf16_tflops_sustained"),
                "source": "measured"}
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


# ---------------------------------------------------------------- reference arm / CPU baseline (oracle port)
CPU_SAMPLE_N = 4096


def cpu_sgemm_sample(steps, warmup, threads):
    """The reference's CPU SGEMM path (oracle port of compute/sgemm.go on the AVX2+FMA route, one task per 64x64
    C block over all host threads) on a bounded sample of the headline workload: SGEMM CPU_SAMPLE_N^3."""
    from oracle import oracle as O
    n = CPU_SAMPLE_N
    a = O.uniform_f32(n * n, 1); b = O.uniform_f32(n * n, 2); c = np.zeros(n * n, dtype=np.float32)
    for _ in range(warmup):
        O.sgemm(0, 0, n, n, n, 1.0, a, n, b, n, 0.0, c, n, threads=threads)
    t0 = time.perf_counter()
    for _ in range(steps):
        O.sgemm(0, 0, n, n, n, 1.0, a, n, b, n, 0.0, c, n, threads=threads)
    dt = (time.perf_counter() - t0) / max(steps, 1)
    return 2.0 * n ** 3 / dt / 1e12, dt


def _timed(fn, reps=2):
    fn()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    return (time.perf_counter() - t0) / reps


def cpu_suite_baselines(cores):
    """The reference's CPU path (oracle port) beside each suite config, threaded the way the reference threads it
    (BASELINE.md section 3): SGEMM over all cores (compute/sgemm.go:165-264), AXPY / DOT / DGEMM on ONE thread
    (compute/level1float32.go:247-249, compute/dgemm.go:122).  Bounded: about 10 s in total."""
    from concurrent.futures import ThreadPoolExecutor
    from oracle import oracle as O
    out = {}
    n = 1024
    a, b, c = O.uniform_f32(n * n, 1), O.uniform_f32(n * n, 2), np.zeros(n * n, np.float32)
    dt = _timed(lambda: O.sgemm(0, 0, n, n, n, 1.0, a, n, b, n, 0.0, c, n, threads=cores), 3)
    out["sgemm_1024"] = {"value": 2.0 * n ** 3 / dt / 1e12, "unit": "TFLOP/s", "cores": cores, "kind": "port",
                         "sample": "SGEMM 1024^3 fp32 alpha=1 beta=0, full size, one task per 64x64 C block over all cores, %.1f ms" % (dt * 1e3)}
    nv = 1 << 24
    x, y = O.uniform_f32(nv, 3), O.uniform_f32(nv, 4)
    dt = _timed(lambda: O.saxpy(nv, 2.5, x, 1, y, 1), 3)
    chunks = [(i * (nv // cores), (i + 1) * (nv // cores) if i < cores - 1 else nv) for i in range(cores)]
    with ThreadPoolExecutor(cores) as ex:
        dta = _timed(lambda: list(ex.map(lambda ch: O.saxpy(ch[1] - ch[0], 2.5, x[ch[0]:ch[1]], 1, y[ch[0]:ch[1]], 1), chunks)), 3)
        dtd_all = _timed(lambda: sum(ex.map(lambda ch: float(O.sdot(ch[1] - ch[0], x[ch[0]:ch[1]], 1, y[ch[0]:ch[1]], 1)), chunks)), 3)
    out["saxpy_16M"] = {"value": 12.0 * nv / dt / 1e9, "unit": "GB/s", "cores": 1, "kind": "port",
                        "sample": "SAXPY n=2^24 alpha=2.5 full size, ONE thread as the reference (compute/level1float32.go:247-249), %.1f ms" % (dt * 1e3),
                        "all_cores_not_the_reference": {"value": 12.0 * nv / dta / 1e9, "unit": "GB/s", "cores": cores}}
    dt = _timed(lambda: O.sdot(nv, x, 1, y, 1), 3)
    out["sdot_16M"] = {"value": 8.0 * nv / dt / 1e9, "unit": "GB/s", "cores": 1, "kind": "port",
                       "sample": "SDOT n=2^24 full size, ONE thread as the reference (compute/level1float32_sdot.go:18), %.1f ms" % (dt * 1e3),
                       "all_cores_not_the_reference": {"value": 8.0 * nv / dtd_all / 1e9, "unit": "GB/s", "cores": cores}}
    M, K, N = 1024, 4096, 16384
    a, b, bias, c = O.uniform_f32(M * K, 5), O.uniform_f32(K * N, 6, -0.05, 0.05), O.uniform_f32(N, 7), np.zeros(M * N, np.float32)
    dt = _timed(lambda: O.fused_gemm_bias_gelu(0, 0, M, N, K, 1.0, a, K, b, N, 0.0, c, N, bias, threads=cores), 1)
    out["fused_gemm_bias_gelu_f32_8192x4096x16384"] = {
        "value": 2.0 * M * N * K / dt / 1e12, "unit": "TFLOP/s", "cores": cores, "kind": "port",
        "sample": "M reduced 8192 -> 1024 (K=4096, N=16384 kept): GEMM over all cores, then the reference's second bias+GELU pass "
                  "(fused_gelu.go:96-137), %.2f s" % dt}
    n = 2048
    ad, bd, cd = O.uniform_f64(n * n, 8), O.uniform_f64(n * n, 9), np.zeros(n * n)
    dt = _timed(lambda: O.dgemm(0, 0, n, n, n, 1.0, ad, n, bd, n, 0.0, cd, n), 1)
    out["dgemm_16384"] = {"value": 2.0 * n ** 3 / dt / 1e12, "unit": "TFLOP/s", "cores": 1, "kind": "port",
                          "sample": "N reduced 16384 -> 2048, ONE thread as the reference (compute/dgemm.go:122 -> dgemm_fix.go:7-11), %.2f s" % dt}
    return out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = host_cores()
    tf, dt = cpu_sgemm_sample(args.steps, min(args.warmup, 1), cores)
    sample = "SGEMM %d^3 fp32 per step (bounded sample of the %d^3 headline; oracle port of compute/sgemm.go, %d threads)" % (
        CPU_SAMPLE_N, args.size, cores)
    line = {"impl": "reference", "metric": "sgemm_tflops", "value": tf, "unit": "TFLOP/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "SGEMM %d^3 fp32 NN alpha=1 beta=0 (BASELINE configs[4])" % args.size, "sample": sample},
            "cpu_baseline": {"value": tf, "unit": "TFLOP/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": tf, "unit": "TFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# ---------------------------------------------------------------- device helpers
class Dev:
    def __init__(self, local_rank):
        import guda_b200 as g
        self.g = g
        g.Init(local_rank)
        self.L = g.load()
        self.e0, self.e1 = g.Event(), g.Event()

    def time_ms(self, fn, iters=1):
        """CUDA events on the library's stream around `iters` calls of fn()."""
        self.e0.Record()
        for _ in range(iters):
            fn()
        self.e1.Record()
        self.e1.Synchronize()
        return self.g.ElapsedTime(self.e0, self.e1) / iters

    def peak(self, kind):
        v = C.c_double(0)
        self.g.capi.check(self.L.guda_microbench_peak(kind, 3, C.byref(v)), "microbench")
        return v.value


def ncu_traffic(key):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the named kernel, from the committed ncu capture
    (profiles/traffic.json, written by profiles/summarize.py); None when no capture covers it."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(key)
    except Exception:
        return None


def suite(dev, peaks, ffma_peak, dfma_peak, dmma_peak):
    """BASELINE configs C1..C4 at one GPU; inputs resident in HBM; rotating buffer sets larger than L2."""
    g = dev.g
    out = {}
    # ---- C2: SAXPY / SDOT n = 2^24, alpha = 2.5 (benchmark_test.go:47); 4 rotating pairs = 512 MiB > 126 MB L2
    n = 1 << 24
    pairs = []
    for i in range(4):
        x, y = g.Malloc(n * 4), g.Malloc(n * 4)
        g.FillUniform(x, 1, n, n, 100 + i); g.FillUniform(y, 1, n, n, 200 + i)
        pairs.append((x, y))
    it = [0]

    def axpy():
        x, y = pairs[it[0] % 4]; it[0] += 1
        g.AXPY(2.5, x, y, n)

    def dot():
        x, y = pairs[it[0] % 4]; it[0] += 1
        r = C.c_float(0)
        dev.L.guda_sdot_device(n, x.ptr, 1, y.ptr, 1, scal.ptr, None)

    scal = g.Malloc(64)
    for _ in range(8):
        axpy(); dot()
    ms = dev.time_ms(axpy, 40)
    gbs = 12.0 * n / (ms * 1e-3) / 1e9
    out["saxpy_16M"] = {"value": gbs, "unit": "GB/s", "ms": ms, "roofline": {"bound": "hbm", "achieved": gbs, "peak": peaks["hbm_gbs"],
                        "unit": "GB/s", "frac": gbs / peaks["hbm_gbs"], "frac_of_nominal_8TBs": gbs / 8000.0, "peak_source": peaks["source"],
                        "algorithmic_bytes": 12.0 * n, "traffic": ncu_traffic("saxpy_16M")}}
    ms = dev.time_ms(dot, 40)
    gbs = 8.0 * n / (ms * 1e-3) / 1e9
    out["sdot_16M"] = {"value": gbs, "unit": "GB/s", "ms": ms, "roofline": {"bound": "hbm", "achieved": gbs, "peak": peaks["hbm_gbs"],
                       "unit": "GB/s", "frac": gbs / peaks["hbm_gbs"], "frac_of_nominal_8TBs": gbs / 8000.0, "peak_source": peaks["source"],
                       "algorithmic_bytes": 8.0 * n, "traffic": ncu_traffic("sdot_16M")}}
    # blocking DOT through the public call (host scalar back every call)
    t0 = time.perf_counter()
    for _ in range(40):
        x, y = pairs[it[0] % 4]; it[0] += 1
        g.DOT(x, y, n)
    out["sdot_16M"]["blocking_call_gbs"] = 8.0 * n / ((time.perf_counter() - t0) / 40) / 1e9
    for x, y in pairs:
        g.Free(x); g.Free(y)
    g.Free(scal)

    # ---- level 2: SGEMV N / T and SGER on a 16384 x 16384 matrix (1 GiB, 2 rotating copies > L2)
    m2 = n2 = 16384
    mats = []
    for i in range(2):
        a = g.Malloc(m2 * n2 * 4); g.FillUniform(a, m2, n2, n2, 250 + i); mats.append(a)
    vx, vy = g.Malloc(n2 * 4), g.Malloc(n2 * 4)
    g.FillUniform(vx, 1, n2, n2, 252); g.FillUniform(vy, 1, n2, n2, 253)

    def l2(fn, nbytes, key):
        def call():
            a = mats[it[0] % 2]; it[0] += 1
            fn(a)
        for _ in range(3):
            call()
        ms = dev.time_ms(call, 20)
        gbs = nbytes / (ms * 1e-3) / 1e9
        out[key] = {"value": gbs, "unit": "GB/s", "ms": ms, "roofline": {"bound": "hbm", "achieved": gbs, "peak": peaks["hbm_gbs"],
                    "unit": "GB/s", "frac": gbs / peaks["hbm_gbs"], "frac_of_nominal_8TBs": gbs / 8000.0, "peak_source": peaks["source"],
                    "algorithmic_bytes": nbytes, "traffic": ncu_traffic(key)}}

    l2(lambda a: dev.L.guda_sgemv(0, m2, n2, 1.0, a.ptr, n2, vx.ptr, 1, 0.5, vy.ptr, 1, None), 4.0 * m2 * n2, "sgemv_n_16384")
    l2(lambda a: dev.L.guda_sgemv(1, m2, n2, 1.0, a.ptr, n2, vx.ptr, 1, 0.5, vy.ptr, 1, None), 4.0 * m2 * n2, "sgemv_t_16384")
    l2(lambda a: dev.L.guda_sger(m2, n2, 1e-6, vx.ptr, 1, vy.ptr, 1, a.ptr, n2, None), 8.0 * m2 * n2, "sger_16384")
    for a in mats:
        g.Free(a)
    g.Free(vx); g.Free(vy)

    # ---- reductions: blocking scalar-returning calls over 2^28 floats (1 GiB, 2 rotating buffers)
    nr = 1 << 28
    rb = []
    for i in range(2):
        x = g.Malloc(nr * 4); g.FillUniform(x, 1, nr, nr, 260 + i); rb.append(x)
    rf, ri = C.c_float(0), C.c_longlong(0)
    for key, fn, res in (("ssum_256M", dev.L.guda_ssum, rf), ("smax_256M", dev.L.guda_smax, rf), ("sargmax_256M", dev.L.guda_sargmax, ri)):
        def call():
            x = rb[it[0] % 2]; it[0] += 1
            fn(nr, x.ptr, C.byref(res), None)
        for _ in range(3):
            call()
        ms = dev.time_ms(call, 10)
        gbs = 4.0 * nr / (ms * 1e-3) / 1e9
        out[key] = {"value": gbs, "unit": "GB/s", "ms": ms, "includes": "launch + stream sync + host read of the scalar, every call",
                    "roofline": {"bound": "hbm", "achieved": gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": gbs / peaks["hbm_gbs"],
                                 "peak_source": peaks["source"], "algorithmic_bytes": 4.0 * nr, "traffic": ncu_traffic(key)}}
    for x in rb:
        g.Free(x)

    # ---- ReduceSum / ReduceMax along an axis (reduce.go:105-178): 16384 x 16384 (1 GiB, beyond L2)
    ra, ro = g.Malloc(16384 * 16384 * 4), g.Malloc(16384 * 4)
    g.FillUniform(ra, 16384, 16384, 16384, 650)
    for key, fn, axis in (("reduce_sum_axis0_16384", g.ReduceSum, 0), ("reduce_sum_axis1_16384", g.ReduceSum, 1),
                          ("reduce_max_axis0_16384", g.ReduceMax, 0), ("reduce_max_axis1_16384", g.ReduceMax, 1)):
        for _ in range(2):
            fn(ra, [16384, 16384], axis, ro)
        ms = dev.time_ms(lambda: fn(ra, [16384, 16384], axis, ro), 10)
        gbs = 4.0 * 16384 * 16384 / (ms * 1e-3) / 1e9
        out[key] = {"value": gbs, "unit": "GB/s", "ms": ms,
                    "roofline": {"bound": "hbm", "achieved": gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": gbs / peaks["hbm_gbs"],
                                 "peak_source": peaks["source"], "algorithmic_bytes": 4.0 * 16384 * 16384, "traffic": ncu_traffic(key)}}
    g.Free(ra); g.Free(ro)

    # ---- elementwise device ops (math.go:133-169, fused.go:252): n = 2^26 (256 MiB per vector, beyond L2)
    ne = 1 << 26
    ea, eb, ec = g.Malloc(ne * 4), g.Malloc(ne * 4), g.Malloc(ne * 4)
    g.FillUniform(ea, 1, ne, ne, 700); g.FillUniform(eb, 1, ne, ne, 701)
    for key, fn, bpe in (("add_64M", lambda: g.Add(ea, eb, ec, ne), 12.0), ("linear_relu_64M", lambda: g.LinearReLU(ec, 1.01, 0.1, ne), 8.0),
                         ("gelu_64M", lambda: g.GELU(ec, ne), 8.0), ("sigmoid_64M", lambda: g.Sigmoid(ec, ne), 8.0)):
        for _ in range(3):
            fn()
        ms = dev.time_ms(fn, 20)
        gbs = bpe * ne / (ms * 1e-3) / 1e9
        out[key] = {"value": gbs, "unit": "GB/s", "ms": ms,
                    "roofline": {"bound": "hbm", "achieved": gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": gbs / peaks["hbm_gbs"],
                                 "peak_source": peaks["source"], "algorithmic_bytes": bpe * ne, "traffic": ncu_traffic(key)}}
    for p in (ea, eb, ec):
        g.Free(p)

    # ---- C1: SGEMM 1024^3 (README headline), alpha=1 beta=0; 16 rotating triples (192 MiB)
    n = 1024
    sets = []
    for i in range(16):
        a, b, c = g.Malloc(n * n * 4), g.Malloc(n * n * 4), g.Malloc(n * n * 4)
        g.FillUniform(a, n, n, n, 300 + i); g.FillUniform(b, n, n, n, 400 + i)
        sets.append((a, b, c))

    def sg():
        a, b, c = sets[it[0] % 16]; it[0] += 1
        dev.L.guda_sgemm(0, 0, n, n, n, 1.0, a.ptr, n, b.ptr, n, 0.0, c.ptr, n, None)

    for _ in range(5):
        sg()
    ms = dev.time_ms(sg, 50)
    tf = 2.0 * n ** 3 / (ms * 1e-3) / 1e12
    out["sgemm_1024"] = {"value": tf, "unit": "TFLOP/s", "ms": ms, "roofline": {"bound": "fp32_simt", "achieved": tf, "peak": ffma_peak,
                         "unit": "TFLOP/s", "frac": tf / ffma_peak, "peak_source": "FFMA microbenchmark, this run"}}
    # the same call in the opt-in 3xTF32 math mode (stated tolerance, not the reference's bits: tests/test_gpu_tensor.py), for scale
    g.SetSgemmMode(1)
    try:
        for _ in range(5):
            sg()
        ms3 = dev.time_ms(sg, 50)
    finally:
        g.SetSgemmMode(0)
    tf3 = 2.0 * n ** 3 / (ms3 * 1e-3) / 1e12
    tpeak3 = peaks["bf16_tflops"] / 2.0 / 3.0
    out["sgemm_1024_3xtf32"] = {"value": tf3, "unit": "TFLOP/s", "ms": ms3, "includes": "hi/lo split pre-pass + pool traffic",
                                "math": "3xTF32 split on tcgen05 (opt-in, abs(delta) <= 8*K*u*sum|a||b|); NOT bit-identical to the reference",
                                "roofline": {"bound": "tensor", "achieved": tf3, "peak": tpeak3, "unit": "TFLOP/s", "frac": tf3 / tpeak3,
                                             "peak_source": "MEASURED_PEAKS.json bf16 burst / 2 (tf32 rate) / 3 (three MMAs per product)"}}
    for a, b, c in sets:
        g.Free(a); g.Free(b); g.Free(c)

    # ---- Conv2D: conv_test.go:262-300's 64x56x56 -> 64, 3x3 pad 1 shape at batch 32; implicit GEMM vs materialised im2col
    cp = g.ConvParams(BatchSize=32, InChannels=64, InHeight=56, InWidth=56, OutChannels=64, KernelHeight=3, KernelWidth=3,
                      StrideH=1, StrideW=1, PadH=1, PadW=1, DilationH=1, DilationW=1, UseBias=True)
    n_in, n_w, n_out = 32 * 64 * 56 * 56, 64 * 64 * 9, 32 * 64 * 56 * 56
    cx, cw, cb, co = g.Malloc(n_in * 4), g.Malloc(n_w * 4), g.Malloc(64 * 4), g.Malloc(n_out * 4)
    g.FillUniform(cx, 1, n_in, n_in, 600); g.FillUniform(cw, 1, n_w, n_w, 601); g.FillUniform(cb, 1, 64, 64, 602)
    conv_flops = 2.0 * 64 * (64 * 9) * (32 * 56 * 56)
    conv_ms = {}
    for route, name in ((0, "implicit"), (1, "im2col")):
        g.SetConvRoute(route)
        for _ in range(2):
            g.Conv2D(cx, cw, cb, co, cp)
        conv_ms[name] = dev.time_ms(lambda: g.Conv2D(cx, cw, cb, co, cp), 10)
    g.SetConvRoute(-1)
    tf = conv_flops / (conv_ms["implicit"] * 1e-3) / 1e12
    out["conv2d_b32_64x56x56_3x3"] = {
        "value": tf, "unit": "TFLOP/s", "ms": conv_ms["implicit"], "ms_im2col_route": conv_ms["im2col"],
        "includes": "stream-ordered calls (one implicit-GEMM kernel each; the offset table is cached per geometry)",
        "roofline": {"bound": "fp32_simt", "achieved": tf, "peak": ffma_peak, "unit": "TFLOP/s", "frac": tf / ffma_peak,
                     "peak_source": "FFMA microbenchmark, this run"}}
    for p in (cx, cw, cb, co):
        g.Free(p)

    # ---- C3: fused GEMM + bias + GELU, FFN shape M=8192 K=4096 N=16384, fp32
    M, K, N = 8192, 4096, 16384
    a, b, c, bias = g.Malloc(M * K * 4), g.Malloc(K * N * 4), g.Malloc(M * N * 4), g.Malloc(N * 4)
    s = 1.0 / np.sqrt(K)
    g.FillUniform(a, M, K, K, 500); g.FillUniform(b, K, N, N, 501, lo=-3 * s, hi=3 * s); g.FillUniform(bias, 1, N, N, 502)

    def fused():
        dev.L.guda_sgemm_bias_gelu(0, 0, M, N, K, 1.0, a.ptr, K, b.ptr, N, 0.0, c.ptr, N, bias.ptr, None)

    def plain():
        dev.L.guda_sgemm(0, 0, M, N, K, 1.0, a.ptr, K, b.ptr, N, 0.0, c.ptr, N, None)

    fused()
    ms = dev.time_ms(fused, 3)
    ms_plain = dev.time_ms(plain, 3)
    tf = 2.0 * M * N * K / (ms * 1e-3) / 1e12
    out["fused_gemm_bias_gelu_f32_8192x4096x16384"] = {
        "value": tf, "unit": "TFLOP/s", "ms": ms, "ms_plain_sgemm_same_shape": ms_plain,
        "roofline": {"bound": "fp32_simt", "achieved": tf, "peak": ffma_peak, "unit": "TFLOP/s", "frac": tf / ffma_peak,
                     "peak_source": "FFMA microbenchmark, this run"}}
    for p in (a, b, c, bias):
        g.Free(p)

    # ---- C3 bf16: same FFN shape, bf16 storage / fp32 TMEM accumulation, tcgen05 + fused bias+GELU epilogue
    a, b, c, bias = g.Malloc(M * K * 2), g.Malloc(K * N * 2), g.Malloc(M * N * 2), g.Malloc(N * 4)
    tmp = g.Malloc(K * N * 4)
    g.FillUniform(tmp, M, K, K, 510); dev.L.guda_f32_to_bf16(tmp.ptr, a.ptr, M * K, None)
    g.FillUniform(tmp, K, N, N, 511, lo=-3 * s, hi=3 * s); dev.L.guda_f32_to_bf16(tmp.ptr, b.ptr, K * N, None)
    g.FillUniform(bias, 1, N, N, 512)
    g.Synchronize(); g.Free(tmp)

    def fused16():
        dev.L.guda_gemm_bf16_bias_gelu(M, N, K, 1.0, a.ptr, K, b.ptr, N, c.ptr, N, bias.ptr, None)

    def plain16():
        dev.L.guda_gemm_bf16(0, 0, M, N, K, 1.0, a.ptr, K, b.ptr, N, 0.0, c.ptr, N, None)

    fused16(); plain16()
    ms = dev.time_ms(fused16, 10)
    ms_plain = dev.time_ms(plain16, 10)
    tf = 2.0 * M * N * K / (ms * 1e-3) / 1e12
    out["fused_gemm_bias_gelu_bf16_8192x4096x16384"] = {
        "value": tf, "unit": "TFLOP/s", "ms": ms, "ms_plain_gemm_same_shape": ms_plain, "tflops_plain": 2.0 * M * N * K / (ms_plain * 1e-3) / 1e12,
        "roofline": {"bound": "tensor", "achieved": tf, "peak": peaks["bf16_tflops"], "unit": "TFLOP/s", "frac": tf / peaks["bf16_tflops"],
                     "peak_source": "MEASURED_PEAKS.json bf16 burst (%s)" % peaks["source"]}}
    for p in (a, b, c, bias):
        g.Free(p)

    # ---- SGEMM 8192^3 on the 3xTF32 tensor-core route (opt-in math mode; stated tolerance, see tests/test_gpu_tensor.py)
    n = 8192
    a, b, c = g.Malloc(n * n * 4), g.Malloc(n * n * 4), g.Malloc(n * n * 4)
    g.FillUniform(a, n, n, n, 520); g.FillUniform(b, n, n, n, 521)
    g.SetSgemmMode(1)

    def sg3():
        dev.L.guda_sgemm(0, 0, n, n, n, 1.0, a.ptr, n, b.ptr, n, 0.0, c.ptr, n, None)

    try:
        sg3()
        ms = dev.time_ms(sg3, 3)
    finally:
        g.SetSgemmMode(0)
    tf = 2.0 * n ** 3 / (ms * 1e-3) / 1e12
    tpeak = peaks["bf16_tflops"] / 2.0 / 3.0
    out["sgemm_3xtf32_8192"] = {"value": tf, "unit": "TFLOP/s", "ms": ms, "includes": "hi/lo split pre-pass + pool traffic",
                                "roofline": {"bound": "tensor", "achieved": tf, "peak": tpeak, "unit": "TFLOP/s", "frac": tf / tpeak,
                                             "peak_source": "MEASURED_PEAKS.json bf16 burst / 2 (tf32 rate) / 3 (three MMAs per product)"}}
    for p in (a, b, c):
        g.Free(p)

    # ---- C4: DGEMM 16384^3, alpha=1 beta=0
    n = 16384
    a, b, c = g.Malloc(n * n * 8), g.Malloc(n * n * 8), g.Malloc(n * n * 8)
    g.FillUniform(a, n, n, n, 600, dtype=np.float64); g.FillUniform(b, n, n, n, 601, dtype=np.float64)

    def dg():
        dev.L.guda_dgemm(0, 0, n, n, n, 1.0, a.ptr, n, b.ptr, n, 0.0, c.ptr, n, None)

    dg()
    ms = dev.time_ms(dg, 2)
    tf = 2.0 * n ** 3 / (ms * 1e-3) / 1e12
    fp64_peak = max(dmma_peak, dfma_peak)
    out["dgemm_16384"] = {"value": tf, "unit": "TFLOP/s", "ms": ms, "route": "FP64 tensor pipe (DMMA m8n8k4), bit-exact vs the oracle's FMA chain",
                          "roofline": {"bound": "tensor", "pipe": "fp64 tensor (DMMA)", "achieved": tf, "peak": fp64_peak,
                                       "unit": "TFLOP/s", "frac": tf / fp64_peak,
                                       "peak_source": "max(DMMA, DFMA) microbenchmarks, this run: DMMA %.1f, DFMA %.1f TFLOP/s" % (dmma_peak, dfma_peak),
                                       "traffic": ncu_traffic("dgemm_dmma_16384")}}
    g.SetDgemmMode(1)
    try:
        dg()
        ms_simt = dev.time_ms(dg, 1)
    finally:
        g.SetDgemmMode(0)
    out["dgemm_16384"]["simt_route_tflops"] = 2.0 * n ** 3 / (ms_simt * 1e-3) / 1e12
    for p in (a, b, c):
        g.Free(p)
    dev.L.guda_pool_trim()
    return out


# ---------------------------------------------------------------- headline: SGEMM size^3, strong scaling over a pr x pc grid
from guda_b200.sharding import grid_for, block_extents  # noqa: E402  (integer geometry only; no device code)


def _panel_host(g, rows, cols, seed, row0, col0, full_cols, dtype=np.float32):
    """rows x cols window at (row0, col0) of the seeded full matrix: regenerated on the device (counter-based RNG) and copied down."""
    it = np.dtype(dtype).itemsize
    p = g.Malloc(rows * cols * it)
    g.FillUniform(p, rows, cols, cols, seed, row0=row0, col0=col0, full_cols=full_cols, dtype=dtype)
    g.Synchronize()
    h = (p.Float32() if dtype == np.float32 else p.Float64())[:rows * cols].copy()
    g.Free(p)
    return h


def parity_blocks(g, read_block, n, blocks, seed_a, seed_b, dtype=np.float32):
    """SURVEY.md 8(d) parity at scale: a C element is ONE sequential-k chain over a row of A and a column of B, and the inputs come
    from a counter-based generator, so the 4 x n / n x 4 panels behind any 4 x 4 block of C can be regenerated and handed to the
    oracle (outside every timed region).  Returns the number of blocks that are bit-identical to the oracle's chains."""
    from oracle import oracle as O
    good = 0
    for i0, j0 in blocks:
        ar = _panel_host(g, 4, n, seed_a, i0, 0, n, dtype)
        bc = _panel_host(g, n, 4, seed_b, 0, j0, n, dtype)
        if dtype == np.float32:
            want = O.sgemm(0, 0, 4, 4, n, 1.0, ar, n, bc, 4, 0.0, np.zeros(16, np.float32), 4).reshape(4, 4)
        else:
            want = O.dgemm(0, 0, 4, 4, n, 1.0, ar, n, bc, 4, 0.0, np.zeros(16), 4).reshape(4, 4)
        got = read_block(i0, j0)
        good += int(np.array_equal(got.view(np.uint32 if dtype == np.float32 else np.uint64), want.view(np.uint32 if dtype == np.float32 else np.uint64))
                    and bool(np.all(np.isfinite(got))))
    return good


def _device_block_reader(g, c, ldc, row_off, col_off, dtype=np.float32):
    it = np.dtype(dtype).itemsize

    def read(i0, j0):
        out = np.empty((4, 4), dtype)
        for i in range(4):
            row = g.DevicePtr(c.ptr + ((i0 - row_off + i) * ldc + (j0 - col_off)) * it, 4 * it)
            out[i] = (row.Float32() if dtype == np.float32 else row.Float64())[:4]
        return out
    return read


def sample_blocks(n, pr, pc, rank, world):
    """>= 6 blocks of the gathered C: the two far corners, a block owned by ANOTHER rank, one straddling the seam of four ranks'
    blocks (or the centre), one straddling a 64 x 128 tile corner, one at an unaligned position."""
    m_loc, n_loc = n // pr, n // pc
    other = (rank + 1) % world
    orow, ocol = (other // pc) * m_loc, (other % pc) * n_loc
    seam = (m_loc - 2 if pr > 1 else n // 2 - 2, n_loc - 2 if pc > 1 else n // 2 - 2)
    return [(0, 0), (n - 4, n - 4), (orow + m_loc // 2 + 1, ocol + n_loc // 2 + 3), seam, (62, 126), ((40001 * 7) % (n - 4), (50003 * 3) % (n - 4))]


def sharded_secondary(dev, comm, dist, world, rank, pr, pc, peaks, dmma_peak, barrier, max_over_ranks, min_over_ranks):
    """N > 1 only: the other sharded forms of SURVEY.md 8(e), each timed on the device (max over ranks) and checked.
      dgemm_16384     DGEMM 16384^3 (BASELINE configs[3]) 2-D partitioned, C gathered by the DMMA kernel's peer-store epilogue
      sdot_sharded    per-rank DOT over 2^24 resident elements + one-float all-reduce (weak scaling: n = 2^24 * N)
      saxpy_sharded   per-rank AXPY over 2^24 resident elements, no collective"""
    import torch
    from oracle import oracle as O
    from guda_b200 import tolerance as T
    g, L = dev.g, dev.L
    out = {}
    ok = True
    # ---- DGEMM 16384^3
    n = 16384
    row0, m_loc, col0, n_loc = block_extents(rank, pr, pc, n, n)
    a_p, b_p, c_b, c_f = g.Malloc(m_loc * n * 8), g.Malloc(n * n_loc * 8), g.Malloc(m_loc * n_loc * 8), g.Malloc(n * n * 8)
    g.FillUniform(a_p, m_loc, n, n, seed=21, row0=row0, col0=0, full_cols=n, dtype=np.float64)
    g.FillUniform(b_p, n, n_loc, n_loc, seed=22, row0=0, col0=col0, full_cols=n, dtype=np.float64)
    g.Memset(c_f, 0xFF, n * n * 8)

    def dstep():
        g.capi.check(L.guda_dgemm_sharded(comm, pr, pc, n, n, n, 1.0, a_p.ptr, n, b_p.ptr, n_loc, c_b.ptr, n_loc, c_f.ptr, n, None), "DgemmSharded")
    dstep(); dstep()
    barrier()
    ms = max_over_ranks(dev.time_ms(dstep, 3))
    barrier()
    good = parity_blocks(g, _device_block_reader(g, c_f, n, 0, 0, np.float64), n, sample_blocks(n, pr, pc, rank, world), 21, 22, np.float64)
    good = min_over_ranks(good)
    kms = dev.time_ms(lambda: L.guda_dgemm(0, 0, m_loc, n_loc, n, 1.0, a_p.ptr, n, b_p.ptr, n_loc, 0.0, c_b.ptr, n_loc, None), 2)
    ktf = 2.0 * m_loc * n_loc * n / (kms * 1e-3) / 1e12
    out["dgemm_16384"] = {"value": 2.0 * n ** 3 / (ms * 1e-3) / 1e12, "unit": "TFLOP/s", "ms_per_step": ms, "scaling": "strong",
                          "workload": "DGEMM 16384^3 fp64 NN alpha=1 beta=0 (BASELINE configs[3]), %dx%d block partition, C gathered on every rank "
                                      "by the FP64 tensor-pipe kernel's peer-store epilogue" % (pr, pc),
                          "roofline": {"bound": "tensor", "pipe": "fp64 tensor (DMMA)", "achieved": ktf, "peak": dmma_peak, "unit": "TFLOP/s",
                                       "frac": ktf / dmma_peak if dmma_peak else None, "what": "this rank's block GEMM alone, per GPU"},
                          "parity": {"blocks": 6, "bit_identical": bool(good == 6), "where": "gathered C on every rank vs the oracle's chains"}}
    ok = ok and good == 6
    for p in (a_p, b_p, c_b, c_f):
        g.Free(p)
    L.guda_pool_trim()
    # ---- sharded Level-1: n_local = 2^24 per rank, 4 rotating pairs (512 MiB > L2)
    nl = 1 << 24
    pairs = []
    for i in range(4):
        x, y = g.Malloc(nl * 4), g.Malloc(nl * 4)
        g.FillUniform(x, 1, nl, nl, 300 + i, row0=0, col0=rank * nl, full_cols=nl * world)
        g.FillUniform(y, 1, nl, nl, 400 + i, row0=0, col0=rank * nl, full_cols=nl * world)
        pairs.append((x, y))
    it = [0]
    res = C.c_float(0)

    def sdot():
        x, y = pairs[it[0] % 4]; it[0] += 1
        g.capi.check(L.guda_sdot_sharded(comm, nl, x.ptr, y.ptr, C.byref(res), None), "SdotSharded")

    def saxpy():
        x, y = pairs[it[0] % 4]; it[0] += 1
        g.capi.check(L.guda_saxpy_sharded(comm, nl, 1e-3, x.ptr, y.ptr, None), "SaxpySharded")

    dres = g.Malloc(64)

    def sdot_dev():                                        # stream-ordered form: the scalar stays on the device, no host read
        x, y = pairs[it[0] % 4]; it[0] += 1
        g.capi.check(L.guda_sdot_sharded_device(comm, nl, x.ptr, y.ptr, dres.ptr, None), "SdotShardedDevice")

    def tiny_dot():                                        # the collective + host read alone: the latency floor of the blocking call
        x, y = pairs[0]
        g.capi.check(L.guda_sdot_sharded(comm, 32, x.ptr, y.ptr, C.byref(res), None), "SdotSharded")
    for _ in range(4):
        sdot()
    barrier()
    t0 = time.perf_counter()
    for _ in range(40):
        sdot()
    ms = max_over_ranks((time.perf_counter() - t0) / 40 * 1e3)
    for _ in range(4):
        tiny_dot()
    barrier()
    t0 = time.perf_counter()
    for _ in range(40):
        tiny_dot()
    floor_ms = max_over_ranks((time.perf_counter() - t0) / 40 * 1e3)
    # parity: the sharded result of pair 0 against the one-GPU DOT of the whole vector (rank 0 regenerates it), ReduceArchTolerance
    it[0] = 0
    sdot()
    got = float(res.value)
    dot_ok = 1.0
    if rank == 0:
        xf, yf = g.Malloc(nl * world * 4), g.Malloc(nl * world * 4)
        g.FillUniform(xf, 1, nl * world, nl * world, 300); g.FillUniform(yf, 1, nl * world, nl * world, 400)
        whole = float(g.DOT(xf, yf, nl * world))
        dot_ok = 1.0 if T.Float32NearEqual(whole, got, T.GetOperationTolerance("dot")) else 0.0
        g.Free(xf); g.Free(yf)
    dot_ok = min_over_ranks(dot_ok)
    for _ in range(4):
        sdot_dev()
    barrier()
    ms_dev = max_over_ranks(dev.time_ms(sdot_dev, 40))
    barrier()
    g.Free(dres)
    gbs_dev = 8.0 * nl * world / (ms_dev * 1e-3) / 1e9
    out["sdot_sharded_stream_ordered"] = {
        "value": gbs_dev, "unit": "GB/s", "ms_per_call": ms_dev, "scaling": "weak", "n_total": nl * world,
        "what": "guda_sdot_sharded_device: per-rank DOT + one-float ncclAllReduce, result left on the device; CUDA events over 40 back-to-back calls, max over ranks",
        "roofline": {"bound": "hbm", "achieved": gbs_dev, "peak": peaks["hbm_gbs"] * world, "unit": "GB/s", "frac": gbs_dev / (peaks["hbm_gbs"] * world)}}
    gbs = 8.0 * nl * world / (ms * 1e-3) / 1e9
    out["sdot_sharded"] = {"value": gbs, "unit": "GB/s", "ms_per_call": ms, "scaling": "weak", "n_total": nl * world,
                           "what": "blocking call: per-rank DOT + one-float ncclAllReduce + host read of the scalar, wall clock, max over ranks",
                           "allreduce_and_host_read_floor_ms": floor_ms,
                           "roofline": {"bound": "hbm", "achieved": gbs, "peak": peaks["hbm_gbs"] * world, "unit": "GB/s",
                                        "frac": gbs / (peaks["hbm_gbs"] * world), "frac_excluding_the_floor": 8.0 * nl * world / (max(ms - floor_ms, 1e-6) * 1e-3) / 1e9 / (peaks["hbm_gbs"] * world)},
                           "parity": {"within_ReduceArchTolerance_of_one_gpu": bool(dot_ok > 0.5)}}
    ok = ok and dot_ok > 0.5
    for _ in range(4):
        saxpy()
    barrier()
    ms = max_over_ranks(dev.time_ms(saxpy, 40))
    barrier()
    # parity: a 4096-element window of this rank's chunk replayed by the oracle (one FMA per element: every bit)
    x, y = pairs[0]
    win = 4096
    hx = g.DevicePtr(x.ptr + 4 * 12345, win * 4).Float32()[:win].copy()
    hy = g.DevicePtr(y.ptr + 4 * 12345, win * 4).Float32()[:win].copy()
    g.capi.check(L.guda_saxpy_sharded(comm, win, 2.5, x.ptr + 4 * 12345, y.ptr + 4 * 12345, None), "SaxpySharded")
    g.Synchronize()
    ax_ok = float(np.array_equal(g.DevicePtr(y.ptr + 4 * 12345, win * 4).Float32()[:win], O.saxpy(win, 2.5, hx, 1, hy.copy(), 1)))
    ax_ok = min_over_ranks(ax_ok)
    gbs = 12.0 * nl * world / (ms * 1e-3) / 1e9
    out["saxpy_sharded"] = {"value": gbs, "unit": "GB/s", "ms_per_call": ms, "scaling": "weak", "n_total": nl * world,
                            "what": "stream-ordered per-rank AXPY on its resident chunk, no collective; CUDA events, max over ranks",
                            "roofline": {"bound": "hbm", "achieved": gbs, "peak": peaks["hbm_gbs"] * world, "unit": "GB/s", "frac": gbs / (peaks["hbm_gbs"] * world)},
                            "parity": {"bit_identical_window": bool(ax_ok > 0.5)}}
    ok = ok and ax_ok > 0.5
    for x, y in pairs:
        g.Free(x); g.Free(y)
    L.guda_pool_trim()
    out["all_parity_ok"] = bool(ok)
    return out


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if rank == 0:
            print("bench.py: --gpus %d but WORLD_SIZE=%d (launch N>1 with torchrun)" % (args.gpus, world), file=sys.stderr)
        sys.exit(2)
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    dev = Dev(local_rank)
    g, L = dev.g, dev.L
    peaks = measured_peaks()
    if args.mode == "3xtf32":
        g.SetSgemmMode(1)

    def barrier():
        if dist is not None:
            dist.barrier()
        g.Synchronize()

    def min_over_ranks(x):
        if dist is None:
            return x
        import torch
        t = torch.tensor([float(x)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return float(t.item())

    def max_over_ranks(x):
        if dist is None:
            return x
        import torch
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    ffma_peak = max(dev.peak(0), dev.peak(4))          # scalar FFMA and packed FFMA2: the roofline uses the higher
    dfma_peak = dev.peak(1) if rank == 0 else 0.0
    dmma_peak = dev.peak(3) if rank == 0 else 0.0
    copy_peak = dev.peak(2) if rank == 0 else 0.0

    n = args.size
    pr, pc = grid_for(world)
    row0, m_loc, col0, n_loc = block_extents(rank, pr, pc, n, n)
    comm = None
    if world > 1:
        ident = (C.c_uint8 * 128)()
        if rank == 0:
            g.capi.check(L.guda_comm_unique_id(ident), "CommUniqueId")
        obj = [bytes(ident)]
        dist.broadcast_object_list(obj, src=0)
        ident = (C.c_uint8 * 128).from_buffer_copy(obj[0])
        h = C.c_void_p()
        g.capi.check(L.guda_comm_init_rank(C.byref(h), world, rank, ident), "CommInitRank")
        comm = h

    # A row panel (m_loc x n), B column panel (n x n_loc), generated on the device from the same counter-based
    # stream as the full matrices: U[-1,1), alpha = 1, beta = 0 (benchmark_test.go:124 form)
    a_panel = g.Malloc(m_loc * n * 4); b_panel = g.Malloc(n * n_loc * 4); c_block = g.Malloc(m_loc * n_loc * 4)
    c_full = g.Malloc(n * n * 4) if world > 1 else None
    g.FillUniform(a_panel, m_loc, n, n, seed=11, row0=row0, col0=0, full_cols=n)
    g.FillUniform(b_panel, n, n_loc, n_loc, seed=12, row0=0, col0=col0, full_cols=n)
    g.Synchronize()

    def step():
        if world == 1:
            g.capi.check(L.guda_sgemm(0, 0, n, n, n, 1.0, a_panel.ptr, n, b_panel.ptr, n, 0.0, c_block.ptr, n, None), "Sgemm")
        else:
            g.capi.check(L.guda_sgemm_sharded(comm, pr, pc, n, n, n, 1.0, a_panel.ptr, n, b_panel.ptr, n_loc, c_block.ptr, n_loc,
                                              c_full.ptr, n, None), "SgemmSharded")

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    ms = dev.time_ms(step, args.steps)            # CUDA events on the launching stream, K steps back to back
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms = max_over_ranks(ms)
    flops = 2.0 * n ** 3
    value = flops / (ms * 1e-3) / 1e12

    # ---- parity, outside the timed region: sampled blocks of the (gathered) C against the oracle's chains, on EVERY rank
    c_check, ld_check = (c_full, n) if world > 1 else (c_block, n)
    blocks = sample_blocks(n, pr, pc, rank, world)
    good = parity_blocks(g, _device_block_reader(g, c_check, ld_check, 0, 0), n, blocks, 11, 12) if args.mode == "exact" else len(blocks)
    all_good = min_over_ranks(good)

    # dominant kernel alone (the local block GEMM), for the roofline
    def local_gemm():
        L.guda_sgemm(0, 0, m_loc, n_loc, n, 1.0, a_panel.ptr, n, b_panel.ptr, n_loc, 0.0, c_block.ptr, n_loc, None)
    kern_ms = dev.time_ms(local_gemm, 1) if world > 1 else ms
    kern_tf = 2.0 * m_loc * n_loc * n / (kern_ms * 1e-3) / 1e12
    if args.mode == "exact":
        roof = {"bound": "fp32_simt", "pipe": "fp32 FFMA (exact-order SIMT; tensor cores cannot reproduce the reference's per-step fp32 rounding)",
                "achieved": kern_tf, "peak": ffma_peak, "unit": "TFLOP/s", "frac": kern_tf / ffma_peak,
                "peak_source": "max(FFMA, FFMA2) microbenchmarks measured in this run (MEASURED_PEAKS.json has no fp32 SIMT figure)",
                "frac_of_measured_bf16_tensor_peak": kern_tf / peaks["bf16_tflops_sustained"],
                "traffic": ncu_traffic("sgemm_simt_%d_p%d" % (n, world))}
    else:
        tpeak = peaks["bf16_tflops_sustained"] / 2.0 / 3.0     # TF32 = half the bf16 rate, three MMAs per product
        roof = {"bound": "tensor", "achieved": kern_tf, "peak": tpeak, "unit": "TFLOP/s", "frac": kern_tf / tpeak,
                "peak_source": "MEASURED_PEAKS.json bf16 sustained / 2 (tf32) / 3 (3xTF32 split), %s" % peaks["source"], "traffic": None}

    # ---- e2e: the public call with HOST buffers; H2D of this rank's panels and D2H of its C block inside the timed region
    e2e = None
    if not args.no_e2e:
        import psutil
        need = (m_loc * n + n * n_loc + m_loc * n_loc) * 4
        avail = psutil.virtual_memory().available // max(world, 1)
        if need * 1.2 < avail:
            ha, hb, hc = C.c_void_p(), C.c_void_p(), C.c_void_p()
            g.capi.check(L.guda_malloc_host(C.byref(ha), m_loc * n * 4), "MallocHost")
            g.capi.check(L.guda_malloc_host(C.byref(hb), n * n_loc * 4), "MallocHost")
            g.capi.check(L.guda_malloc_host(C.byref(hc), m_loc * n_loc * 4), "MallocHost")
            L.guda_memcpy(ha, a_panel.ptr, m_loc * n * 4, 2, None); L.guda_memcpy(hb, b_panel.ptr, n * n_loc * 4, 2, None)

            def e2e_step():
                if world == 1:
                    # the host-operand entry point: uploads pipelined under the GEMM by K chunks, C streamed back by row panels
                    g.capi.check(L.guda_sgemm_host(0, 0, n, n, n, 1.0, ha, n, hb, n, 0.0, hc, n, None), "SgemmHost")
                    return
                g.capi.check(L.guda_sgemm_sharded_host(comm, pr, pc, n, n, n, 1.0, ha, n, hb, n_loc, hc, n_loc, a_panel.ptr, b_panel.ptr,
                                                       c_block.ptr, c_full.ptr, n, None), "SgemmShardedHost")
            e2e_step()
            barrier()
            e_steps = max(1, min(args.steps, 5))
            e_ms = max_over_ranks(dev.time_ms(e2e_step, e_steps))
            barrier()
            # parity of the HOST result block (this rank's m_loc x n_loc block of C), outside the timed region
            hc_arr = np.ctypeslib.as_array(C.cast(hc, C.POINTER(C.c_float)), shape=(m_loc, n_loc))
            eb = [(row0, col0), (row0 + m_loc - 4, col0 + n_loc - 4), (row0 + (m_loc // 2 | 1), col0 + (n_loc // 3 | 2))]
            e_good = 0
            if args.mode == "exact":
                e_good = parity_blocks(g, lambda i0, j0: hc_arr[i0 - row0:i0 - row0 + 4, j0 - col0:j0 - col0 + 4].copy(), n, eb, 11, 12)
            e_all = min_over_ranks(e_good)
            e2e = {"value": flops / (e_ms * 1e-3) / 1e12, "unit": "TFLOP/s", "ms_per_step": e_ms, "steps": e_steps,
                   "parity": {"blocks": len(eb), "bit_identical": bool(e_all == len(eb)), "where": "host C block of every rank vs the oracle's chains"}
                             if args.mode == "exact" else None,
                   # N > 1: every unique panel chunk crosses PCIe once (its owner uploads it, ncclBroadcast fans it out over NVLink)
                   "h2d_bytes_per_step": 2 * n * n * 4, "d2h_bytes_per_step": m_loc * n_loc * 4 * world,
                   "nvlink_broadcast_bytes_per_step": ((pc - 1) * pr * m_loc * n + (pr - 1) * pc * n * n_loc) * 4,
                   "host_memory": "pinned",
                   "call": "guda_sgemm_host (K-chunk pipelined H2D, row-panel D2H)" if world == 1 else
                           "guda_sgemm_sharded_host (K-chunk pipelined: one H2D per unique panel chunk + NVLink broadcast to the ranks sharing it, under the "
                           "GEMM; fused gather on the last chunk; D2H of the C block)"}
            for hp in (ha, hb, hc):
                L.guda_free_host(hp)
        else:
            e2e = {"value": None, "unit": "TFLOP/s", "skipped": "host RAM too small for %d bytes of pinned staging" % need}

    for p in (a_panel, b_panel, c_block, c_full):
        if p is not None:
            g.Free(p)
    L.guda_pool_trim()

    secondary = sharded_secondary(dev, comm, dist, world, rank, pr, pc, peaks, dmma_peak, barrier, max_over_ranks, min_over_ranks) \
        if (comm is not None and not args.no_suite) else None

    suite_res, cpu = None, None
    if rank == 0 and world == 1 and not args.no_suite:
        suite_res = suite(dev, peaks, ffma_peak, dfma_peak, dmma_peak)
        if not args.no_cpu:
            for key, cb in cpu_suite_baselines(host_cores()).items():
                if key in suite_res:
                    suite_res[key]["cpu_baseline"] = cb
    if rank == 0 and world == 1 and not args.no_cpu:
        cores = host_cores()
        tf, dt = cpu_sgemm_sample(2, 1, cores)
        cpu = {"value": tf, "unit": "TFLOP/s", "cores": cores, "kind": "port",
               "sample": "SGEMM %d^3 fp32 (bounded sample of the %d^3 headline), oracle port of compute/sgemm.go (AVX2+FMA route, "
                         "one task per 64x64 C block), %.2f s per pass" % (CPU_SAMPLE_N, n, dt)}
    gather = None
    if comm is not None:
        eff = C.c_int(-1)
        L.guda_comm_get_gather_mode(comm, C.byref(eff))
        gather = {0: "fused: GEMM epilogue stores tiles into every rank's C over NVLink peer pointers (CUDA IPC), one kernel",
                  1: "local GEMM, then ncclAllGather (+ unpack kernel)"}.get(eff.value, "?")
        barrier()
        L.guda_comm_destroy(comm)
    if rank == 0:
        # exact mode: K >= 16384 runs as 8192-deep launches resuming the same fp32 chains (gemm.cu launch_simt)
        # exact mode, per step: every 4096-deep K chunk is one pre-pass (t = alpha*A transposed k-major + zero flags) and one
        # pipeline launch (sgemm_pipe.cu); the NCCL gather route adds an unpack kernel when pc > 1
        kchunks = (n + 4095) // 4096 if (args.mode == "exact" and n >= 8192) else 1          # sgemm_pipe.cu: 4096-deep K chunks from k = 8192 up
        launches = args.steps * (2 * kchunks + (0 if (world == 1 or gather.startswith("fused") or pc == 1) else 1))
        line = {"metric": "sgemm_tflops", "value": value, "unit": "TFLOP/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": "SGEMM %d^3 fp32 NN alpha=1 beta=0 (BASELINE configs[4]); %dx%d block partition%s" % (
                               n, pr, pc, ", C gathered on every rank inside the step (%s)" % gather if world > 1 else ""),
                           "math": "exact fp32 SIMT, reference k-order (bit-exact vs oracle)" if args.mode == "exact" else "3xTF32 tcgen05",
                           "l2": "operands %.1f GiB per rank, far larger than the 126 MB L2" % ((m_loc * n + n * n_loc) * 4 / 2 ** 30)},
                "parity": {"blocks": len(blocks), "bit_identical": bool(all_good == len(blocks)),
                           "route": ("p2p" if gather.startswith("fused") else "nccl") if world > 1 else "single",
                           "where": "sampled 4x4 blocks of the %s C on every rank (corners, another rank's block, block seam, tile corner, "
                                    "unaligned) vs the oracle's FMA chains over regenerated panels" % ("gathered" if world > 1 else "full")}
                          if args.mode == "exact" else None,
                "roofline": roof, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks, "sharded": secondary,
                "measured_here": {"ffma_tflops": ffma_peak, "dfma_tflops": dfma_peak, "dmma_tflops": dmma_peak, "copy_gbs": copy_peak},
                "suite": suite_res}
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()
    bad = args.mode == "exact" and (all_good != len(blocks) or (e2e is not None and e2e.get("parity") and not e2e["parity"]["bit_identical"])
                                    or (secondary is not None and not secondary.get("all_parity_ok", True)))
    if bad:
        if rank == 0:
            print("bench.py: PARITY MISMATCH (see the parity objects of the JSON line)", file=sys.stderr)
        sys.exit(1)


if __name__ == "__main__":
    main()
