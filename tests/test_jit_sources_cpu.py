"""The run-time specialised kernel bodies (csrc/jit/*.cuh) are compiled by NVRTC on the GPU box only; a syntax error there would
surface as a failing GPU test.  This CPU test compiles the many-groups body with nvcc for sm_100a for a representative plan in
every variant the host code can ask for (fused / update-only, with / without the range-chunked front region and the claim room)."""
import itertools
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
JIT = os.path.join(ROOT, "presto-1_b200", "csrc", "jit")

PLAN = ("#define PGW_AGGS(X) X(0, CNT, NONE, 0, 0, 0, C64, 3, 4, 0) X(1, SUMDEC, I64, 1, 1, 1, FLAG, 0, 4, 0) "
        "X(2, SUMF64, F64, 0, 0, 0, FLAG, 0, 6, 1) X(3, SUMDEC, DEC128, 0, 0, 1, C64, 7, 8, 0)\n")


@pytest.mark.parametrize("bidx,rng,quota,vec", [(b, r, q, v) for b, r, q, v in itertools.product((0, 1), (0, 1), (0, 1), (4, 1))
                                                if not (b == 1 and (q == 1 or r == 0 and v == 1))])
def test_wide_body_compiles_for_sm_100a(tmp_path, bidx, rng, quota, vec):
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available")
    src = tmp_path / "w.cu"
    src.write_text("#define PGW_BUCKET_WORDS 12\n#define PGW_HASKEYNULL 1\n#define PGW_FLAGWORD 2\n#define PGW_VEC %d\n#define PGW_RANGE %d\n"
                   "#define PGW_BIDX %d\n#define PGW_HASONLY %d\n#define PGW_QUOTA %d\n#define PGW_FLAGS_UNCONDITIONAL 0\n%s"
                   "#include \"%s/wide_args.h\"\n#include \"%s/wide_body.cuh\"\n" % (vec, rng, bidx, 0 if bidx else quota, quota, PLAN, JIT, JIT))
    r = subprocess.run([nvcc, "-arch=sm_100a", "-std=c++17", "-cubin", "-o", str(tmp_path / "w.cubin"), str(src)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
