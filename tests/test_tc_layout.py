"""CPU: the record layout of the tensor-core sweep path's sample table (csrc/tc_layout.cuh), compiled for the host.
The writer (sample_table_kernel) and the reader (accumulate_tc*_kernel) share tc_rec_index; what the reader relies on:
a bijection of (row d, sample k) onto the record's floats, in which lane (r8, kq) finds the eight samples
k = 8 (kq + 4 h) .. + 7 of row 8 rg + r8 as two float4 vectors p = 0, 1 at float4 index ((rg * 2 + h) * 2 + p) * 32 + lane."""
import os
import shutil
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = r'''
#include <cstdio>
#include "%s/ssvi_tf_b200/csrc/tc_layout.cuh"
int main() {
  for (int d = 0; d < 64; ++d)
    for (int k = 0; k < 64; ++k) printf("%%d\n", tc_rec_index(d, k));
  return 0;
}
'''


@pytest.mark.skipif(shutil.which("nvcc") is None, reason="nvcc not on PATH")
def test_tc_rec_index_is_the_lane_order_the_kernel_reads(tmp_path):
    src = tmp_path / "layout.cu"
    src.write_text(SRC % ROOT)
    exe = tmp_path / "layout"
    subprocess.check_call(["nvcc", "-std=c++17", "-o", str(exe), str(src)])
    idx = np.array(subprocess.check_output([str(exe)]).split(), dtype=np.int64).reshape(64, 64)
    for nt in range(1, 9):                                # a record of NT row groups is the first NT * 512 floats
        sub = idx[:8 * nt]
        assert sorted(sub.ravel().tolist()) == list(range(nt * 512))
    for d in (0, 5, 13, 31, 50, 63):
        for k in (0, 7, 8, 37, 63):
            rg, r8, k8 = d // 8, d % 8, k // 8
            h, kq, p, e = k8 // 4, k8 % 4, (k % 8) // 4, k % 4
            lane = r8 + 8 * kq
            assert idx[d, k] == ((((rg * 2 + h) * 2 + p) * 32 + lane) * 4 + e)
