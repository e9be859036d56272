"""PagesSerde wire format (presto-1_b200/pages_serde.py) against what the reference pins for it: the serialized sizes asserted by
TestPagesSerde.java:63-109, the byte layouts stated in the block encodings (cited in the module), and round trips over every block
type of the path including nulls, RLE and DICTIONARY input."""
import importlib
import struct

import numpy as np
import pytest

P = importlib.import_module("presto_b200")
S = importlib.import_module("presto_b200.pages_serde")


def _name(s):
    return struct.pack("<i", len(s)) + s.encode()


def test_bigint_serialized_sizes_match_the_reference_test():
    """testBigintSerializedSize: 35 bytes of page overhead, 9 for the first value (value + the null flag), 8 for the second."""
    one = S.serialize_page(P.Page(P.Block.long_array([123])))
    two = S.serialize_page(P.Page(P.Block.long_array([123, 456])))
    assert len(one) == 35 + 9 and len(two) == 35 + 9 + 8
    assert one == struct.pack("<ibii", 1, 0, 31, 31) + struct.pack("<i", 1) + _name("LONG_ARRAY") + struct.pack("<ibq", 1, 0, 123)


def test_empty_bigint_page_as_the_reference_sends_it():
    """The reference's empty block is a RunLengthEncodedBlock of a NULL value: 52 bytes (TestPagesSerde.java:70-71).  Built here from
    the encodings' layouts; it has to have that size and to read back as an empty page."""
    value = _name("LONG_ARRAY") + struct.pack("<ibB", 1, 1, 0b1000_0000) + struct.pack("<i", 0)      # one position, null, no values
    body = struct.pack("<i", 1) + _name("RLE") + struct.pack("<i", 0) + value
    wire = struct.pack("<ibii", 0, 0, len(body), len(body)) + body
    assert len(wire) == 52
    page, end = S.deserialize_page(wire, [P.BIGINT])
    assert end == 52 and page.position_count == 0 and page.blocks[0].n == 0


def test_varchar_serialized_sizes_match_the_reference_test():
    """testVarcharSerializedSize: 44 bytes for the empty page, 4 + 5 for 'alice', 4 + 3 for 'bob'."""
    sizes = [len(S.serialize_page(P.Page(P.Block.variable_width(v), position_count=len(v)))) for v in ([], ["alice"], ["alice", "bob"])]
    assert sizes == [44, 44 + 9, 44 + 9 + 7]


def test_null_positions_carry_no_value_on_the_wire():
    """LongArrayBlockEncoding.writeBlock: nulls as bits (most significant first), then the count and the values of the non-null rows."""
    wire = S.serialize_page(P.Page(P.Block.long_array([7, 0, 9], [0, 1, 0])))
    body = wire[13:]
    assert body == struct.pack("<i", 1) + _name("LONG_ARRAY") + struct.pack("<ibB", 3, 1, 0b0100_0000) + struct.pack("<iqq", 2, 7, 9)
    back, _ = S.deserialize_page(wire, [P.BIGINT])
    assert back.blocks[0].to_list() == [7, None, 9]
    # variable width: end offsets per position, a null has length zero
    wire = S.serialize_page(P.Page(P.Block.variable_width(["ab", None, "c"])))
    assert wire[13:] == struct.pack("<i", 1) + _name("VARIABLE_WIDTH") + struct.pack("<iiii", 3, 2, 2, 3) + struct.pack("<bB", 1, 0b0100_0000) + struct.pack("<i", 3) + b"abc"


def _random_page(rng, n, null_rate):
    def nulls():
        return (rng.random(n) < null_rate).astype(np.uint8) if null_rate > 0 else None
    words = [None if rng.random() < null_rate else "w%d" % x * int(x % 3) for x in rng.integers(0, 50, n)]
    big = [None if rng.random() < null_rate else int(x) * 10 ** 20 * (-1 if x % 2 else 1) for x in rng.integers(1, 10 ** 9, n)]
    return P.Page(P.Block.long_array(rng.integers(-10 ** 15, 10 ** 15, n), nulls()), P.Block.double_array(rng.normal(size=n), nulls()),
                  P.Block.int_array(rng.integers(-1000, 1000, n).astype(np.int32), nulls(), P.DATE), P.Block.byte_array(rng.integers(0, 2, n).astype(np.uint8), nulls()),
                  P.Block.int128_array(big), P.Block.variable_width(words), P.Block.long_array(rng.integers(0, 10 ** 9, n), nulls(), P.SHORT_DECIMAL),
                  position_count=n)


TYPES = [P.BIGINT, P.DOUBLE, P.DATE, P.BOOLEAN, P.LONG_DECIMAL, P.VARCHAR, P.SHORT_DECIMAL]


@pytest.mark.parametrize("n,null_rate", [(0, 0.0), (1, 0.0), (7, 0.5), (8, 0.5), (9, 1.0), (1000, 0.1), (4097, 0.0)])
def test_round_trip_every_block_type(n, null_rate):
    rng = np.random.default_rng(n + 1)
    page = _random_page(rng, n, null_rate)
    back, end = S.deserialize_page(S.serialize_page(page), TYPES)
    assert back.position_count == n
    for a, b in zip(page.blocks, back.blocks):
        assert a.type == b.type and a.to_list() == b.to_list()


def test_streams_of_pages_and_framing_errors():
    rng = np.random.default_rng(3)
    pages = [_random_page(rng, n, 0.2) for n in (5, 0, 300)]
    wire = S.write_pages(pages)
    back = S.read_pages(wire, TYPES)
    assert [p.position_count for p in back] == [5, 0, 300]
    assert all(a.to_list() == b.to_list() for p, q in zip(pages, back) for a, b in zip(p.blocks, q.blocks))
    with pytest.raises(ValueError):
        S.deserialize_page(wire, TYPES[:-1])                                       # channel count
    with pytest.raises(ValueError):
        S.deserialize_page(wire, [P.INTEGER] + TYPES[1:])                          # an INT_ARRAY type for a LONG_ARRAY block
    compressed = bytearray(wire)
    compressed[4] = S.COMPRESSED
    with pytest.raises(ValueError):
        S.deserialize_page(bytes(compressed), TYPES)


def test_dictionary_and_rle_channels_are_flattened():
    """Pass-through channels of the reference arrive as DICTIONARY (ids into a dictionary block, then 24 bytes of source id) or RLE."""
    dictionary = _name("VARIABLE_WIDTH") + struct.pack("<iii", 2, 1, 3) + b"\x00" + struct.pack("<i", 3) + b"Nxy"
    dict_block = _name("DICTIONARY") + struct.pack("<i", 4) + dictionary + struct.pack("<iiii", 1, 0, 0, 1) + struct.pack("<qqq", 11, 22, 33)
    rle_block = _name("RLE") + struct.pack("<i", 4) + _name("INT_ARRAY") + struct.pack("<ibi", 1, 0, 42)
    body = struct.pack("<i", 2) + dict_block + rle_block
    wire = struct.pack("<ibii", 4, 0, len(body), len(body)) + body
    page, end = S.deserialize_page(wire, [P.VARCHAR, P.INTEGER])
    assert end == len(wire)
    assert page.blocks[0].to_list() == [b"xy", b"N", b"N", b"xy"] or page.blocks[0].to_list() == ["xy", "N", "N", "xy"]
    assert page.blocks[1].to_list() == [42, 42, 42, 42]
