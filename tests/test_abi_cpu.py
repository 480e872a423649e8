"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol include/guda_b200.h declares,
the ctypes table covers the header, and -- with no GPU -- every compute path FAILS LOUDLY (there is no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    src = open(os.path.join(ROOT, "include", "guda_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(guda_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_hot_path():
    syms = _header_symbols()
    for need in ("guda_malloc", "guda_free", "guda_memcpy", "guda_stream_create", "guda_event_record", "guda_saxpy", "guda_sdot",
                 "guda_sscal", "guda_sgemm", "guda_dgemm", "guda_sgemm_bias_relu", "guda_sgemm_bias_gelu", "guda_gemm_bf16",
                 "guda_gemm_f16", "guda_sgemm_sharded", "guda_sdot_sharded"):
        assert need in syms


def test_library_exports_every_declared_symbol():
    import guda_b200
    lib = C.CDLL(guda_b200.LIB_PATH)
    missing = [s for s in _header_symbols() if not hasattr(lib, s)]
    assert not missing, missing


def test_ctypes_table_matches_header():
    from guda_b200 import capi
    assert sorted(capi.SIGNATURES) == _header_symbols()
    capi.load()


def test_no_torch_types_in_the_abi():
    src = open(os.path.join(ROOT, "include", "guda_b200.h")).read()
    assert 'extern "C"' in src
    code = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    assert "torch" not in code and "at::" not in code and "#include <torch" not in src


def _have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_have_gpu(), reason="only meaningful without a CUDA device")
def test_compute_fails_loudly_without_a_gpu():
    import guda_b200 as g
    from guda_b200 import capi
    L = g.load()
    assert L.guda_init(0) == capi.ERR_DEVICE
    assert b"no CPU fallback" in L.guda_last_error()
    x = (C.c_float * 4)()
    assert L.guda_saxpy(4, 1.0, x, 1, x, 1, None) == capi.ERR_DEVICE
    r = C.c_float()
    assert L.guda_sdot(4, x, 1, x, 1, C.byref(r), None) == capi.ERR_DEVICE
    assert L.guda_sgemm(0, 0, 2, 2, 1, 1.0, x, 1, x, 2, 0.0, x, 2, None) == capi.ERR_DEVICE
    with pytest.raises(g.GUDAError):
        g.Malloc(64)


def test_argument_panics_need_no_gpu():
    """compute/errors.go:9-35 strings come back through guda_last_error before any device work."""
    import guda_b200 as g
    from guda_b200 import capi
    L = g.load()
    x = (C.c_float * 4)()
    assert L.guda_sgemm(0, 0, -1, 2, 1, 1.0, x, 1, x, 2, 0.0, x, 2, None) == capi.ERR_INVALID_ARG
    assert L.guda_last_error() == b"blas: m < 0"
    assert L.guda_sgemm(0, 0, 2, 2, 2, 1.0, x, 1, x, 2, 0.0, x, 2, None) == capi.ERR_INVALID_ARG
    assert L.guda_last_error() == b"blas: bad leading dimension of A"
    assert L.guda_saxpy(4, 1.0, x, 0, x, 1, None) == capi.ERR_INVALID_ARG
    assert L.guda_last_error() == b"blas: zero x index increment"
    assert L.guda_sgemm_bias_gelu(1, 0, 2, 2, 1, 1.0, x, 1, x, 2, 0.0, x, 2, x, None) == capi.ERR_NOT_IMPLEMENTED


@pytest.mark.skipif(_have_gpu(), reason="only meaningful without a CUDA device")
def test_cpp_host_mirror_builds_and_refuses_to_run_without_a_gpu():
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "host")], stdout=subprocess.DEVNULL)
    p = subprocess.run([os.path.join(ROOT, "host", "host_test")], capture_output=True, text=True)
    assert p.returncode == 3 and "no CPU fallback" in p.stdout


def test_product_package_never_imports_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "guda_b200")):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, fn)).read()
                assert "oracle" not in text.replace("the oracle", "").replace("oracle's", "").replace("oracle checks", ""), os.path.join(dirpath, fn)


def _schedule(k, m=65536):
    from guda_b200 import capi
    L = capi.load()
    bounds, count, panels = (C.c_int * 64)(), C.c_int(0), C.c_int(0)
    capi.check(L.guda_host_gemm_schedule(k, m, bounds, 64, C.byref(count), C.byref(panels)), "HostGemmSchedule")
    return list(bounds[:count.value]), panels.value


@pytest.mark.parametrize("k", [1, 100, 8192, 8193, 9000, 16384, 20000, 24576, 32768, 65536, 131072])
def test_host_gemm_k_chunk_schedule(k):
    """Host logic of guda_sgemm_host / guda_sgemm_sharded_host (no device call): the chunk bounds cover [0, k) in order; up to one 8192
    block there is a single chunk; beyond it the depths start at 512 and at most double from chunk to chunk (chunk j+1 uploads under
    chunk j's GEMM), every interior bound keeps the 16-byte / TMA alignment, and from three blocks up the last chunk is more than one block (at most two) deep
    (the download of C hides under the last chunk only)."""
    b, _ = _schedule(k)
    assert b[0] == 0 and b[-1] == k and all(x < y for x, y in zip(b, b[1:]))
    depth = [y - x for x, y in zip(b, b[1:])]
    if k <= 8192:
        assert depth == [k]
        return
    assert depth[0] == 512
    assert all(x % 512 == 0 for x in b[1:-1])
    body = depth[:-1]
    assert all(d2 <= 2 * d1 for d1, d2 in zip(body, body[1:]))
    assert max(depth[:-1], default=0) <= 8192
    assert depth[-1] <= 2 * 8192
    if k >= 3 * 8192:
        assert depth[-1] > 8192


def test_host_gemm_row_panels():
    assert _schedule(100, 1024)[1] == 1 and _schedule(100, 2048)[1] == 8 and _schedule(100, 32768)[1] == 16
