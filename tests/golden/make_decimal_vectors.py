"""Writes tests/golden/decimal_operator_vectors.json from the reference's own known-answer tests for decimal +, -, *:

  /root/reference/core/trino-main/src/test/java/io/trino/type/TestDecimalOperators.java   testAdd, testSubtract, testMultiply

Every line of the form   assertDecimalFunction("DECIMAL 'a' <op> DECIMAL 'b'", decimal("r"))   or
assertInvalidFunction("DECIMAL 'a' <op> DECIMAL 'b'", NUMERIC_VALUE_OUT_OF_RANGE)   is transcribed mechanically (regex over the
Java source, which exists in the build container only; the JSON is what travels).  A literal's type is decimal(total digits,
digits after the point), leading zeros included (Decimals.parseIncludeLeadingZerosInPrecision), as is the expected value's.

Run:  python tests/golden/make_decimal_vectors.py
"""
import json
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/core/trino-main/src/test/java/io/trino/type/TestDecimalOperators.java"


def literal(text):
    neg = text.startswith("-")
    body = text.lstrip("+-")
    whole, _, frac = body.partition(".")
    digits = whole + frac
    return {"unscaled": (-1 if neg else 1) * int(digits or "0"), "precision": len(digits), "scale": len(frac), "negative_zero_text": neg and int(digits or "0") == 0}


def main():
    ok = re.compile(r'assertDecimalFunction\("DECIMAL \'([^\']+)\' ([-+*]) DECIMAL \'([^\']+)\'", decimal\("([^"]+)"\)\)')
    bad = re.compile(r'assertInvalidFunction\("DECIMAL \'([^\']+)\' ([-+*]) DECIMAL \'([^\']+)\'", NUMERIC_VALUE_OUT_OF_RANGE\)')
    cases = []
    for ln, line in enumerate(open(SRC), 1):
        m = ok.search(line)
        if m:
            cases.append({"line": ln, "op": m.group(2), "a": literal(m.group(1)), "b": literal(m.group(3)), "expected": literal(m.group(4))})
            continue
        m = bad.search(line)
        if m:
            cases.append({"line": ln, "op": m.group(2), "a": literal(m.group(1)), "b": literal(m.group(3)), "expected": "NUMERIC_VALUE_OUT_OF_RANGE"})
    out = {"source": "core/trino-main/src/test/java/io/trino/type/TestDecimalOperators.java (testAdd :31-85, testSubtract :87-136, testMultiply :138-178)",
           "cases": cases}
    path = os.path.join(HERE, "decimal_operator_vectors.json")
    with open(path, "w") as f:
        json.dump(out, f, separators=(",", ":"))
        f.write("\n")
    print("wrote", path, len(cases), "cases")


if __name__ == "__main__":
    main()
