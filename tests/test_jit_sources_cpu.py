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


# ---- the few-groups body (csrc/jit/lowcard_body.cuh): a plain plan (argument columns, a mask, a nullable argument) and the Q1 plan
# with its filter and derived long decimals fused in (the text jit_lowcard.cu + project_codegen generate for it)
LOWCARD_PLAIN = (
    "#define PG_KEYS(X) X(0, KVAR, 0, 0, 0, 1) X(1, K32, 0, 32, 1, 2)\n"
    "#define PG_LOADS(X) X(0, U8, 1) X(1, U64, 1) X(2, U128, 0) X(3, I32, 0)\n"
    "#define PG_SLOTS(X) X(0, CNT, 0) X(1, I64, 1) X(2, DEC, 3) X(3, MAX, 6) X(4, F64, 7)\n"
    "#define PG_ROW_FILTER  const bool pgx_pass = true;\n"
    "#define PG_ROW_UPDATE  const bool pgc0 = (ld0[v] != 0) && !ldn0[v] && !(ldn1[v] != 0); const bool pgc1 = true;"
    " if (pgc0) { rec[0 * PG_T] += 1; } if (pgc0) { upd_i64(rec + 1 * PG_T, ld1[v]); } if (pgc1) { upd_dec(rec + 3 * PG_T, ld2[v], ldh2[v]); }"
    " if (pgc1) { upd_max(rec + 6 * PG_T, ld3[v], 0); } if (pgc0) { upd_f64(rec + 7 * PG_T, ld1[v], 1); }\n")
LOWCARD_Q1_FUSED = (
    "#define PG_KEYS(X) X(0, KVAR, 0, 0, 0, 1) X(1, KVAR, 0, 32, 0, 2)\n"
    "#define PG_LOADS(X) X(0, I32, 0) X(1, U64, 0) X(2, U64, 0) X(3, U64, 0) X(4, U64, 1)\n"
    "#define PG_SLOTS(X) X(0, CNT, 0) X(1, I64, 1) X(2, I64, 3) X(3, DEC, 5) X(4, DEC, 8) X(5, I64, 11)\n"
    "#define PG_ROW_FILTER const pgx_u64 x2 = ((pgx_i64)(ld0[v]) <= (pgx_i64)(((pgx_u64)10471LL))) ? 1ULL : 0ULL;  const bool pgx_pass = (x2 != 0);\n"
    "#define PG_ROW_UPDATE  const pgx_u64 x7 = (pgx_u64)(((pgx_i64)(((pgx_u64)1LL)) * 100LL) - (pgx_i64)(ld3[v]));"
    " int x8e = 0; const Pgx2c x8 = pgx2c_mul_ss((pgx_i64)(ld2[v]), (pgx_i64)(x7));"
    " const pgx_u64 x10 = (pgx_u64)(((pgx_i64)(((pgx_u64)1LL)) * 100LL) + (pgx_i64)(ld4[v]));"
    " int x11e = 0; const Pgx2c x11 = pgx2c_mul_ls(x8, (pgx_i64)(x10), x11e); if (!(ldn4[v] != 0)) pgx_err |= x11e;"
    " const bool pgc0 = true; const bool pgc1 = !(ldn4[v] != 0);"
    " if (pgc0) { rec[0 * PG_T] += 1; } if (pgc0) { upd_i64(rec + 1 * PG_T, ld1[v]); } if (pgc0) { upd_i64(rec + 3 * PG_T, ld2[v]); }"
    " if (pgc0) { upd_dec2c(rec + 5 * PG_T, x8.w0, (i64)(x8.w1)); } if (pgc1) { upd_dec2c(rec + 8 * PG_T, x11.w0, (i64)(x11.w1)); }"
    " if (pgc0) { upd_i64(rec + 11 * PG_T, ld3[v]); }\n")


@pytest.mark.parametrize("plan,words", [(LOWCARD_PLAIN, 8), (LOWCARD_Q1_FUSED, 13)], ids=["plain", "q1_fused"])
@pytest.mark.parametrize("vec", [4, 1])
def test_lowcard_body_compiles_for_sm_100a(tmp_path, plan, words, vec):
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available")
    src = tmp_path / "l.cu"
    src.write_text("#define PG_T 512\n#define PG_VEC %d\n#define PG_MIN_BLOCKS 1\n#define PG_REC_WORDS %d\n#define PG_PREFETCH 1\n%s"
                   "#include \"%s/decimal_ops.cuh\"\n#include \"%s/lowcard_args.h\"\n#include \"%s/lowcard_body.cuh\"\n" % (vec, words, plan, JIT, JIT, JIT))
    r = subprocess.run([nvcc, "-arch=sm_100a", "-std=c++17", "-cubin", "-Xptxas", "-v", "-o", str(tmp_path / "l.cubin"), str(src)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    assert "bytes spill" not in r.stderr or " 0 bytes spill stores" in r.stderr, r.stderr[-1500:]
