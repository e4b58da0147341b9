"""The C++ host adapter (include/dindel_b200.hpp): the source-level drop-in for
DetInDel::computeLikelihoods (DInDel.cpp:1707-1739).

  * CPU: the header compiles as gnu++98 (the reference's dialect) and as C++17 against test doubles
    with the reference's member names, links the C-ABI library, and refuses to run without a CUDA
    device (exit code 2, "no usable CUDA device": there is no CPU path).
  * GPU: the same binary passes its alignment checks; and oracle/_ref/adapter_ref_test -- the header
    instantiated with the reference's OWN Haplotype / Read / MLAlignment / AlignedVariant, built where
    /root/reference exists -- reproduces the reference class's MLAlignment field by field.
"""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "adapter_mock_test.cpp")
LIBDIR = os.path.join(ROOT, "dindel_tgi_b200")
REF_TEST = os.path.join(ROOT, "oracle", "_ref", "adapter_ref_test")


def _build(tmp_path, std):
    from dindel_tgi_b200 import engine
    engine.load_library()   # fails loudly when the CUDA library has not been built
    exe = str(tmp_path / ("adapter_mock_" + std.replace("+", "p")))
    subprocess.check_call(["g++", "-std=" + std, "-O1", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), SRC,
                           "-o", exe, "-L" + LIBDIR, "-ldindel_b200", "-Wl,-rpath," + LIBDIR])
    return exe


@pytest.mark.parametrize("std", ["gnu++98", "c++17"])
def test_adapter_compiles_and_has_no_cpu_path(tmp_path, std):
    import torch
    exe = _build(tmp_path, std)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    if torch.cuda.is_available():
        assert r.returncode == 0, r.stdout + r.stderr
    else:
        assert r.returncode == 2 and "no usable CUDA device" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
def test_adapter_mock_types_on_gpu(tmp_path):
    exe = _build(tmp_path, "gnu++98")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "adapter_mock_test ok" in r.stdout, r.stdout + r.stderr
    # one operation slot per read to start with: every CIGAR longer than that goes through the adapter's second pass
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300, env=dict(os.environ, DINDEL_B200_CIGAR_CAP="1"))
    assert r.returncode == 0 and "adapter_mock_test ok" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
def test_adapter_with_reference_types_matches_reference_class():
    if not os.path.exists(REF_TEST):
        pytest.skip("oracle/_ref/adapter_ref_test not built (reference sources absent at build time)")
    r = subprocess.run([REF_TEST, "40"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "adapter_ref_test ok" in r.stdout, r.stdout + r.stderr
    r = subprocess.run([REF_TEST, "12"], capture_output=True, text=True, timeout=600, env=dict(os.environ, DINDEL_B200_CIGAR_CAP="2"))
    assert r.returncode == 0 and "adapter_ref_test ok" in r.stdout, r.stdout + r.stderr
