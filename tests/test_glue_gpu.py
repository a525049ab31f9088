"""The Arrow FunctionRegistry glue (arrow_glue/): the reference's own public API — CallFunction /
compute::Filter / SortIndices / internal::GroupBy / Grouper — run against the CUDA registry on
CudaBuffer-resident arrays and compared with the reference's CPU registry in the same process
(arrow_glue/test_glue.cc).  The binary is built where /root/reference exists and ships prebuilt."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "arrow_glue", "test_glue")
ARCHIVE = os.path.join(ROOT, "arrow_glue", "libarrow1_b200_arrow.a")


def test_glue_archive_defines_the_registration_entry_points():
    from arrow1_b200 import glue_build
    glue_build.build()
    if not os.path.exists(ARCHIVE):
        pytest.skip("arrow_glue not built (no reference headers on this box)")
    syms = subprocess.run(["nm", "-C", "--defined-only", ARCHIVE], capture_output=True, text=True).stdout
    for want in ("arrow::compute::cuda::RegisterCudaKernels(arrow::compute::FunctionRegistry*, int)",
                 "arrow::compute::cuda::MakeCudaRegistry(int)",
                 "arrow::compute::cuda::GroupBy(",
                 "arrow::compute::cuda::MakeGrouper("):
        assert want in syms, want
    # the glue computes nothing itself: every operation is an undefined reference into the C ABI
    undefined = subprocess.run(["nm", "-u", ARCHIVE], capture_output=True, text=True).stdout
    for want in ("b200_filter_apply", "b200_sort_indices", "b200_groupby_consume", "b200_grouper_consume", "b200_hash_agg_consume"):
        assert want in undefined, want


@pytest.mark.gpu
def test_glue_parity_against_reference_cpu_registry():
    assert os.path.exists(BIN), "arrow_glue/test_glue missing: run arrow_glue/build.sh where /root/reference exists"
    r = subprocess.run([BIN], capture_output=True, text=True, timeout=900)
    tail = "\n".join(r.stdout.splitlines()[-40:])
    assert r.returncode == 0, "test_glue failed (rc %d):\n%s\n%s" % (r.returncode, tail, r.stderr[-2000:])
    assert "GLUE OK" in r.stdout, tail
    assert "FAIL" not in r.stdout, tail
