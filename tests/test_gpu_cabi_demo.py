"""The C ABI is usable without torch or Python: examples/cabi_demo.cu is compiled with nvcc against
include/b200sft.h and libb200sft.so, run as a plain executable (its own CUDA runtime, cudaMalloc
memory, its own stream) and checks the fused forward loss against a double-precision CPU loop."""
import os
import shutil
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_c_host_calls_the_library(cuda_device, tmp_path):
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available on this box")
    from llm_fine_tune_distributed_b200 import build

    build.build()
    pkg = os.path.join(ROOT, "llm_fine_tune_distributed_b200")
    exe = tmp_path / "cabi_demo"
    cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-O2", "-o", str(exe),
           os.path.join(ROOT, "examples", "cabi_demo.cu"), "-I" + os.path.join(ROOT, "include"), "-L" + pkg, "-lb200sft",
           "-Xlinker", "-rpath=" + pkg]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stderr[-3000:]
    run = subprocess.run([str(exe)], capture_output=True, text=True, timeout=120)
    print(run.stdout)
    assert run.returncode == 0, run.stdout + run.stderr
