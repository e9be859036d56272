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


@pytest.mark.parametrize("bidx,rng,quota,vec,rowids", [(b, r, q, v, 0) for b, r, q, v in itertools.product((0, 1), (0, 1), (0, 1), (4, 1))
                                                       if not (b == 1 and (q == 1 or r == 0 and v == 1))] + [(0, 0, 0, 4, 1), (0, 0, 1, 1, 1)])
def test_wide_body_compiles_for_sm_100a(tmp_path, bidx, rng, quota, vec, rowids):
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available")
    src = tmp_path / "w.cu"
    src.write_text("#define PGW_BUCKET_WORDS 12\n#define PGW_HASKEYNULL 1\n#define PGW_FLAGWORD 2\n#define PGW_VEC %d\n#define PGW_RANGE %d\n"
                   "#define PGW_BIDX %d\n#define PGW_HASONLY %d\n#define PGW_QUOTA %d\n#define PGW_FLAGS_UNCONDITIONAL 0\n#define PGW_ROWIDS %d\n#define PGW_DYNAMIC %d\n#define PGW_PEEK %d\n%s"
                   "#include \"%s/wide_args.h\"\n#include \"%s/wide_body.cuh\"\n" % (vec, rng, bidx, 0 if bidx else quota, quota, rowids, rowids, rowids, PLAN, JIT, JIT))
    r = subprocess.run([nvcc, "-arch=sm_100a", "-std=c++17", "-cubin", "-o", str(tmp_path / "w.cubin"), str(src)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]


# ---- the few-groups body (csrc/jit/lowcard_body.cuh): a plain plan (argument columns, a mask, a nullable argument) and the Q1 plan
# with its filter and derived long decimals fused in (the text jit_lowcard.cu + project_codegen generate for it)
LOWCARD_PLAIN = (
    "#define PG_NSUMS 2\n"
    "#define PG_KEYS(X) X(0, KVAR, 0, 0, 0, 1) X(1, K32, 0, 32, 1, 2)\n"
    "#define PG_LOADS(X) X(0, U8, 1) X(1, U64, 1) X(2, U128, 0) X(3, I32, 0)\n"
    "#define PG_SLOTS(X) X(0, CNT, 0) X(1, SUM, 1) X(2, SUM, 2) X(3, MAX, 3) X(4, F64, 4)\n"
    "#define PG_ROW_FILTER  const bool pgx_pass = true;\n"
    "#define PG_ROW_UPDATE  const bool pgc0 = (ld0[v] != 0) && !ldn0[v] && !(ldn1[v] != 0); const bool pgc1 = true;"
    " rec[0 * PG_T] += pgc0 ? 1ULL : 0ULL;"
    " const bool pgo1 = fits_fast(ld1[v]); rec[1 * PG_T] += (pgc0 && pgo1) ? (u64)(ld1[v]) : 0ULL; pg_slow |= (pgc0 && !pgo1) ? (1u << 1) : 0u;"
    " const Pgx2c pgv2 = pgx2c_from_long(ld2[v], ldh2[v]); const bool pgo2 = (fits_fast(pgv2.w0) && pgv2.w1 == ((i64)(pgv2.w0) >> 63));"
    " rec[2 * PG_T] += (pgc1 && pgo2) ? (u64)(pgv2.w0) : 0ULL; pg_slow |= (pgc1 && !pgo2) ? (1u << 2) : 0u;"
    " if (pgc1) upd_max(rec + 3 * PG_T, ld3[v], 0); if (pgc0) upd_f64(rec + 4 * PG_T, ld1[v], 1);"
    " if (pg_slow) { if (pg_slow & (1u << 1)) wide_add(wrec + 0, (u64)(ld1[v]), (u64)((i64)(ld1[v]) >> 63), (u64)((i64)(ld1[v]) >> 63));"
    " if (pg_slow & (1u << 2)) wide_add(wrec + 3, (u64)(pgv2.w0), (u64)(pgv2.w1), (u64)((i64)(pgv2.w1) >> 63)); }\n")
LOWCARD_Q1_FUSED = (
    "#define PG_NSUMS 5\n"
    "#define PG_KEYS(X) X(0, KVAR, 0, 0, 0, 1) X(1, KVAR, 0, 32, 0, 2)\n"
    "#define PG_LOADS(X) X(0, I32, 0) X(1, U64, 0) X(2, U64, 0) X(3, U64, 0) X(4, U64, 1)\n"
    "#define PG_SLOTS(X) X(0, CNT, 0) X(1, SUM, 1) X(2, SUM, 2) X(3, SUM, 3) X(4, SUM, 4) X(5, SUM, 5)\n"
    "#define PG_ROW_FILTER const pgx_u64 x2 = ((pgx_i64)(ld0[v]) <= (pgx_i64)(((pgx_u64)10471LL))) ? 1ULL : 0ULL;  const bool pgx_pass = (x2 != 0);\n"
    "#define PG_ROW_UPDATE  const pgx_u64 x7 = (pgx_u64)(((pgx_i64)(((pgx_u64)1LL)) * 100LL) - (pgx_i64)(ld3[v]));"
    " int x8e = 0; const Pgx2c x8 = pgx2c_mul_ss((pgx_i64)(ld2[v]), (pgx_i64)(x7));"
    " const pgx_u64 x10 = (pgx_u64)(((pgx_i64)(((pgx_u64)1LL)) * 100LL) + (pgx_i64)(ld4[v]));"
    " int x11e = 0; const Pgx2c x11 = pgx2c_mul_ls(x8, (pgx_i64)(x10), x11e); if (!(ldn4[v] != 0)) pgx_err |= x11e;"
    " const bool pgc0 = true; const bool pgc1 = !(ldn4[v] != 0);"
    " rec[0 * PG_T] += pgc0 ? 1ULL : 0ULL;"
    " const bool pgo1 = fits_fast(ld1[v]); rec[1 * PG_T] += (pgc0 && pgo1) ? (u64)(ld1[v]) : 0ULL; pg_slow |= (pgc0 && !pgo1) ? (1u << 1) : 0u;"
    " const bool pgo2 = fits_fast(ld2[v]); rec[2 * PG_T] += (pgc0 && pgo2) ? (u64)(ld2[v]) : 0ULL; pg_slow |= (pgc0 && !pgo2) ? (1u << 2) : 0u;"
    " const bool pgo3 = (fits_fast(x8.w0) && (i64)(x8.w1) == ((i64)(x8.w0) >> 63)); rec[3 * PG_T] += (pgc0 && pgo3) ? (u64)(x8.w0) : 0ULL;"
    " pg_slow |= (pgc0 && !pgo3) ? (1u << 3) : 0u;"
    " const bool pgo4 = (fits_fast(x11.w0) && (i64)(x11.w1) == ((i64)(x11.w0) >> 63)); rec[4 * PG_T] += (pgc1 && pgo4) ? (u64)(x11.w0) : 0ULL;"
    " pg_slow |= (pgc1 && !pgo4) ? (1u << 4) : 0u;"
    " const bool pgo5 = fits_fast(ld3[v]); rec[5 * PG_T] += (pgc0 && pgo5) ? (u64)(ld3[v]) : 0ULL; pg_slow |= (pgc0 && !pgo5) ? (1u << 5) : 0u;"
    " if (pg_slow) { if (pg_slow & (1u << 3)) wide_add(wrec + 6, (u64)(x8.w0), (u64)((i64)(x8.w1)), (u64)((i64)((i64)(x8.w1)) >> 63)); }\n")


@pytest.mark.parametrize("plan,words", [(LOWCARD_PLAIN, 5), (LOWCARD_Q1_FUSED, 6)], ids=["plain", "q1_fused"])
@pytest.mark.parametrize("vec,threads", [(4, 512), (2, 768), (1, 256)])
def test_lowcard_body_compiles_for_sm_100a(tmp_path, plan, words, vec, threads):
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available")
    src = tmp_path / "l.cu"
    src.write_text("#define PG_T %d\n#define PG_VEC %d\n#define PG_MIN_BLOCKS 1\n#define PG_REC_WORDS %d\n#define PG_PREFETCH 1\n#define PG_FAST_BITS 38\n%s"
                   "#include \"%s/decimal_ops.cuh\"\n#include \"%s/lowcard_args.h\"\n#include \"%s/lowcard_body.cuh\"\n" % (threads, vec, words, plan, JIT, JIT, JIT))
    r = subprocess.run([nvcc, "-arch=sm_100a", "-std=c++17", "-cubin", "-Xptxas", "-v", "-o", str(tmp_path / "l.cubin"), str(src)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    assert "bytes spill" not in r.stderr or " 0 bytes spill stores" in r.stderr, r.stderr[-1500:]
