"""The C++ host layer (llm-slo-ebpf-toolkit_b200/host/llmslo.hpp): the reference's unit tests for the hot path,
transcribed in host_tests.cpp, run as a native binary against libllmslo_gpu.so."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "llm-slo-ebpf-toolkit_b200", "host")


def test_host_layer_builds_and_fails_loudly_without_gpu():
    subprocess.check_call(["make", "-s", "-C", HOST])
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    p = subprocess.run([os.path.join(HOST, "host_tests")], capture_output=True, text=True)
    assert p.returncode == 2 and "no usable CUDA device" in p.stdout      # no CPU path


@pytest.mark.gpu
def test_reference_unit_tests_through_cpp_host_layer():
    if not os.path.exists(os.path.join(HOST, "host_tests")):
        subprocess.check_call(["make", "-s", "-C", HOST])
    p = subprocess.run([os.path.join(HOST, "host_tests")], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stdout + p.stderr
    assert p.stdout.strip().startswith("ok:") and "0 failed" in p.stdout
