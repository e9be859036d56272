"""Pins the oracle's filter + projection evaluator (oracle/oracle_project.h) against the reference's own known answers for decimal
arithmetic (tests/golden/decimal_operator_vectors.json <- TestDecimalOperators.java testAdd / testSubtract / testMultiply) and checks
the PageProcessor semantics the reference defines: filter first, projections over the selected positions only
(PageProcessor.java:139-180), NULL filter results reject the row, three-valued AND / OR."""
import numpy as np
import pytest

import oracle_lib as O
import presto_b200 as P
from presto_b200 import Expr as E
from project_cases import dec_block, dec_type, decimal_vectors, vector_case

VECTORS = decimal_vectors()


@pytest.mark.parametrize("case", VECTORS, ids=lambda c: "L%d" % c["line"])
def test_oracle_matches_reference_decimal_vectors(case):
    page, expr = vector_case(case)
    proc = O.OraclePageProcessor(None, [expr])
    rc, out = proc.process(page)
    exp = case["expected"]
    if exp == "NUMERIC_VALUE_OUT_OF_RANGE":
        assert rc == P.native.PG_E_NUMERIC_OVERFLOW
        return
    assert rc == 0
    t, p, s = proc.output_types[0]
    assert (p, s) == (exp["precision"], exp["scale"]) and t == dec_type(exp["precision"])
    assert out.blocks[0].to_list() == [exp["unscaled"]]


def test_vector_file_is_complete():
    ops = [c["op"] for c in VECTORS]
    assert len(VECTORS) == 84 and ops.count("+") == 32 and ops.count("-") == 31 and ops.count("*") == 21


def test_filter_runs_before_projections():
    # row 1 overflows in the projection but is rejected by the filter: no error (projections only see selected positions)
    big = 10 ** 37
    page = P.Page(dec_block([5, big * 9, 7], 38, 0), P.Block.long_array(np.array([1, 0, 1], dtype=np.int64)))
    x = E.input(0, P.LONG_DECIMAL, 38, 0)
    keep = E.input(1, P.BIGINT).eq(E.bigint(1))
    rc, out = O.OraclePageProcessor(keep, [x * E.decimal(9, 1, 0)]).process(page)
    assert rc == 0 and out.blocks[0].to_list() == [45, 63]
    rc, _ = O.OraclePageProcessor(None, [x * E.decimal(9, 1, 0)]).process(page)
    assert rc == P.native.PG_E_NUMERIC_OVERFLOW


def test_three_valued_logic_and_null_filter():
    t, f, n = 1, 0, None
    a = P.Block.byte_array(np.array([1, 1, 1, 0, 0, 0, 0, 0, 0], dtype=np.uint8), [0, 0, 0, 0, 0, 0, 1, 1, 1])
    b = P.Block.byte_array(np.array([1, 0, 0, 1, 0, 0, 1, 0, 0], dtype=np.uint8), [0, 0, 1, 0, 0, 1, 0, 0, 1])
    page = P.Page(a, b)
    x, y = E.input(0, P.BOOLEAN), E.input(1, P.BOOLEAN)
    rc, out = O.OraclePageProcessor(None, [x & y, x | y, ~x]).process(page)
    assert rc == 0
    assert out.blocks[0].to_list() == [t, f, n, f, f, f, n, f, n]        # AND: FALSE wins over NULL
    assert out.blocks[1].to_list() == [t, t, t, t, f, n, t, n, n]        # OR: TRUE wins over NULL
    assert out.blocks[2].to_list() == [f, f, f, t, t, t, n, n, n]
    rc, out = O.OraclePageProcessor(x & y, [x]).process(page)            # only TRUE passes; NULL rejects
    assert out.position_count == 1


def test_multiply_sign_of_zero_follows_the_reference():
    # UnscaledDecimal128Arithmetic.multiply packs (leftNegative != rightNegative) even for a zero product
    page = P.Page(dec_block([0], 20, 0), dec_block([-3], 5, 0))
    x, y = E.input(0, P.LONG_DECIMAL, 20, 0), E.input(1, P.SHORT_DECIMAL, 5, 0)
    rc, out = O.OraclePageProcessor(None, [x * y, x + x]).process(page)
    v = out.blocks[0].to_host().values
    assert int(v[0]) == 0 and int(v[1]) == 1 << 63
    w = out.blocks[1].to_host().values
    assert int(w[0]) == 0 and int(w[1]) == 0
