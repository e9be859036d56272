"""Writes tests/golden/reference_operator_vectors.json.

The reference is Java and cannot run in this image (no JDK), so the vectors are a transcription of the inputs and
expected rows that the reference's own operator tests define in their source — all of them synthetic sequences:

  OPT = core/trino-main/src/test/java/io/trino/operator
  OPT/TestHashJoinOperator.java           testInnerJoin :165-202, testInnerJoinWithNullProbe :547-585,
                                          testInnerJoinWithNullBuild :587-625, testInnerJoinWithNullOnBothSides :627-666,
                                          testProbeOuterJoin :668-712, testOuterJoinWithNullProbe :763-803,
                                          testOuterJoinWithNullBuild :850-889, testOuterJoinWithNullOnBothSides :936-976
  OPT/TestHashAggregationOperator.java    testHashAggregation :162-220 (max(varchar) left out: out of scope)
  RowPagesBuilder.addSequencePage(n, starts...) = row i of channel c holds (starts[c] + i) rendered in the channel's type.

Run:  python tests/golden/make_golden.py      (no inputs; deterministic)
"""
import hashlib
import json
import os

HERE = os.path.dirname(os.path.abspath(__file__))


def seq_rows(n, types, starts):
    rows = []
    for i in range(n):
        row = []
        for t, s in zip(types, starts):
            row.append(str(s + i) if t == "varchar" else s + i)
        rows.append(row)
    return rows


def join_cases():
    cases = []
    vbb = ["varchar", "bigint", "bigint"]
    build = seq_rows(10, vbb, (20, 30, 40))
    cases.append(dict(
        name="testInnerJoin", source="TestHashJoinOperator.java:165-202", join_type="inner",
        build_types=vbb, probe_types=vbb, build=build, probe=seq_rows(1000, vbb, (0, 1000, 2000)),
        expected=[[str(20 + i), 1020 + i, 2020 + i, str(20 + i), 30 + i, 40 + i] for i in range(10)]))
    cases.append(dict(
        name="testProbeOuterJoin", source="TestHashJoinOperator.java:668-712", join_type="probe_outer",
        build_types=vbb, probe_types=vbb, build=build, probe=seq_rows(15, vbb, (20, 1020, 2020)),
        expected=[[str(20 + i), 1020 + i, 2020 + i, str(20 + i), 30 + i, 40 + i] for i in range(10)]
        + [[str(30 + i), 1030 + i, 2030 + i, None, None, None] for i in range(5)]))
    v = ["varchar"]
    r = lambda *xs: [[x] for x in xs]
    abc = r("a", "b", "c")
    b_nulls = r("a", None, None, "a", "b")
    cases.append(dict(name="testInnerJoinWithNullProbe", source="TestHashJoinOperator.java:547-585", join_type="inner",
                      build_types=v, probe_types=v, build=abc, probe=r("a", None, None, "a", "b"),
                      expected=[["a", "a"], ["a", "a"], ["b", "b"]]))
    cases.append(dict(name="testInnerJoinWithNullBuild", source="TestHashJoinOperator.java:587-625", join_type="inner",
                      build_types=v, probe_types=v, build=b_nulls, probe=abc,
                      expected=[["a", "a"], ["a", "a"], ["b", "b"]]))
    cases.append(dict(name="testInnerJoinWithNullOnBothSides", source="TestHashJoinOperator.java:627-666", join_type="inner",
                      build_types=v, probe_types=v, build=b_nulls, probe=r("a", "b", None, "c"),
                      expected=[["a", "a"], ["a", "a"], ["b", "b"]]))
    cases.append(dict(name="testOuterJoinWithNullProbe", source="TestHashJoinOperator.java:763-803", join_type="probe_outer",
                      build_types=v, probe_types=v, build=abc, probe=r("a", None, None, "a", "b"),
                      expected=[["a", "a"], [None, None], [None, None], ["a", "a"], ["b", "b"]]))
    cases.append(dict(name="testOuterJoinWithNullBuild", source="TestHashJoinOperator.java:850-889", join_type="probe_outer",
                      build_types=v, probe_types=v, build=b_nulls, probe=abc,
                      expected=[["a", "a"], ["a", "a"], ["b", "b"], ["c", None]]))
    cases.append(dict(name="testOuterJoinWithNullOnBothSides", source="TestHashJoinOperator.java:936-976", join_type="probe_outer",
                      build_types=v, probe_types=v, build=b_nulls, probe=r("a", "b", None, "c"),
                      expected=[["a", "a"], ["a", "a"], ["b", "b"], [None, None], ["c", None]]))
    return cases


def agg_expected_rows(n):
    # (Integer.toString(i), 3L, 3L * i, (double) i, 3L, 3L) without the max(varchar) column
    return [[str(i), 3, 3 * i, float(i), 3, 3] for i in range(n)]


def aggregation_cases():
    n = 40_000
    rows = agg_expected_rows(n)
    digest = hashlib.sha256(json.dumps(sorted(rows), separators=(",", ":")).encode()).hexdigest()
    return [dict(
        name="testHashAggregation", source="TestHashAggregationOperator.java:162-220",
        types=["varchar", "varchar", "varchar", "bigint", "boolean"], group_by_channels=[1], step="single", expected_groups=100_000,
        sequence_pages=[[n, 100, 0, 100_000, 0, 500], [n, 100, 0, 200_000, 0, 500], [n, 100, 0, 300_000, 0, 500]],
        aggregates=[["count", None], ["sum", 3], ["avg", 3], ["count_col", 0], ["count_col", 4]],
        expected_rule="row i = [str(i), 3, 3*i, float(i), 3, 3] for i in range(40000), any order",
        expected_first_rows=rows[:5], expected_sha256_sorted=digest)]


def main():
    out = dict(comment=__doc__.strip().split("\n\n")[0], joins=join_cases(), aggregations=aggregation_cases())
    path = os.path.join(HERE, "reference_operator_vectors.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=None, separators=(",", ":"))
        f.write("\n")
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
