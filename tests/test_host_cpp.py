"""The C++ host mirror (host/guda.hpp) on a B200: the reference's test shapes through the same C ABI the Go shim binds."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_cpp_host_mirror_parity_on_gpu():
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "host")], stdout=subprocess.DEVNULL)
    p = subprocess.run([os.path.join(ROOT, "host", "host_test")], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout + p.stderr
    assert "all checks passed" in p.stdout
