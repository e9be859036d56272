"""The reference's page wire format (PagesSerde) for the blocks this path moves: what a PARTIAL aggregation on a CPU worker puts on the
exchange and a FINAL aggregation here takes off it, and the other way round.

Layout, all little-endian (io.airlift.slice):
  SerializedPage   int positionCount, byte markers, int uncompressedSize, int size, `size` bytes      (PagesSerdeUtil.java:66-74)
  raw page         int channelCount, then per channel a block                                         (PagesSerdeUtil.java:47-53)
  block            int nameLength, name (UTF-8), then the encoding's body                             (InternalBlockEncodingSerde.java:43-80)
  LONG_ARRAY / INT_ARRAY / BYTE_ARRAY / INT128_ARRAY
                   int positionCount, nulls, then either every value (no nulls) or int nonNullCount + the non-null values only
                                                                                                      (SPI/block/LongArrayBlockEncoding.java:37-61)
  VARIABLE_WIDTH   int positionCount, int endOffset[positionCount], nulls, int totalLength, bytes     (SPI/block/VariableWidthBlockEncoding.java:37-58)
  RLE              int positionCount, one-position value block                                        (SPI/block/RunLengthBlockEncoding.java:30-40)
  DICTIONARY       int positionCount, dictionary block, int ids[positionCount], 3 longs of source id  (SPI/block/DictionaryBlockEncoding.java:36-54)
  nulls            byte mayHaveNull, then (if set) one bit per position, most significant bit first   (SPI/block/EncoderUtil.java:33-66)

Flat blocks are written; RLE and DICTIONARY are also read (and flattened), since pass-through channels of the reference arrive that way.
Compressed or encrypted pages (markers 1 / 2, PageCodecMarker.java:27-29) are refused: the LZ4 codec and the spill cipher are outside this path.
Pinned by the serialized sizes the reference's own test asserts (TestPagesSerde.java:63-109: 35 bytes of overhead for a one-channel
BIGINT page + 9 for the first value + 8 for the next, 52 for the empty one that travels as RLE, 44 / 53 / 60 for VARCHAR):
tests/test_pages_serde_cpu.py.  Host-side code: the bytes go to and come from the network stack, not the GPU.
"""
import struct

import numpy as np

from . import _native as N
from .blocks import Block, Page

COMPRESSED, ENCRYPTED = 1, 2
_ENCODING = {N.BIGINT: "LONG_ARRAY", N.DOUBLE: "LONG_ARRAY", N.SHORT_DECIMAL: "LONG_ARRAY", N.INTEGER: "INT_ARRAY", N.DATE: "INT_ARRAY",
             N.BOOLEAN: "BYTE_ARRAY", N.LONG_DECIMAL: "INT128_ARRAY", N.VARCHAR: "VARIABLE_WIDTH", N.VARBINARY: "VARIABLE_WIDTH"}
_WIDTH = {"LONG_ARRAY": 8, "INT_ARRAY": 4, "BYTE_ARRAY": 1, "INT128_ARRAY": 16}
_TYPES_OF = {"LONG_ARRAY": (N.BIGINT, N.DOUBLE, N.SHORT_DECIMAL), "INT_ARRAY": (N.INTEGER, N.DATE), "BYTE_ARRAY": (N.BOOLEAN,),
             "INT128_ARRAY": (N.LONG_DECIMAL,), "VARIABLE_WIDTH": (N.VARCHAR, N.VARBINARY)}
_VALUE_DTYPE = {"LONG_ARRAY": np.int64, "INT_ARRAY": np.int32, "BYTE_ARRAY": np.uint8, "INT128_ARRAY": np.uint64}


# ------------------------------------------------------------------------------------------------ writing
def _null_mask(block):
    if block.nulls is None:
        return None
    m = np.asarray(block.nulls) != 0
    return m if m.any() else None


def _write_nulls(out, mask):
    if mask is None:
        out.append(b"\x00")
    else:
        out.append(b"\x01")
        out.append(np.packbits(mask).tobytes())          # most significant bit first, the last byte zero-padded


def _write_block(out, block):
    b = block.to_host()
    name = _ENCODING[b.type]
    out.append(struct.pack("<i", len(name)) + name.encode())
    n = b.n
    mask = _null_mask(b)
    if name == "VARIABLE_WIDTH":
        offs = np.asarray(b.offsets, dtype=np.int64)
        lens = offs[1:] - offs[:-1]
        data = np.asarray(b.values, dtype=np.uint8)
        if mask is not None and lens[mask].any():        # a null position has no bytes on the wire
            keep = np.repeat(~mask, lens)
            data = data[offs[0]:offs[-1]][keep]
            lens = np.where(mask, 0, lens)
        else:
            data = data[offs[0]:offs[-1]]
        ends = np.cumsum(lens).astype("<i4")
        out.append(struct.pack("<i", n))
        out.append(ends.tobytes())
        _write_nulls(out, mask)
        out.append(struct.pack("<i", int(ends[-1]) if n else 0))
        out.append(np.ascontiguousarray(data).tobytes())
        return
    per_row = 2 if name == "INT128_ARRAY" else 1
    values = np.ascontiguousarray(np.asarray(b.values)).view(_VALUE_DTYPE[name])[:n * per_row]
    out.append(struct.pack("<i", n))
    _write_nulls(out, mask)
    if mask is None:
        out.append(values.astype(values.dtype.newbyteorder("<"), copy=False).tobytes())
    else:
        kept = values.reshape(n, per_row)[~mask]
        out.append(struct.pack("<i", int(kept.shape[0])))
        out.append(np.ascontiguousarray(kept).tobytes())


def serialize_page(page):
    """PagesSerde.serialize + PagesSerdeUtil.writeSerializedPage for one page (no compression, no encryption): bytes."""
    body = [struct.pack("<i", len(page.blocks))]
    for b in page.blocks:
        _write_block(body, b)
    payload = b"".join(body)
    return struct.pack("<ibii", page.position_count, 0, len(payload), len(payload)) + payload


def write_pages(pages):
    """PagesSerdeUtil.writePages: the pages one after the other."""
    return b"".join(serialize_page(p) for p in pages)


# ------------------------------------------------------------------------------------------------ reading
class _Reader:
    def __init__(self, buf, at=0):
        self.buf, self.at = memoryview(buf), at

    def int32(self):
        v = struct.unpack_from("<i", self.buf, self.at)[0]
        self.at += 4
        return v

    def byte(self):
        v = self.buf[self.at]
        self.at += 1
        return v

    def array(self, dtype, count):
        a = np.frombuffer(self.buf, dtype=np.dtype(dtype).newbyteorder("<"), count=count, offset=self.at)
        self.at += a.nbytes
        return a

    def nulls(self, n):
        if self.byte() == 0:
            return None
        packed = self.array(np.uint8, (n + 7) // 8)
        return np.unpackbits(packed, count=n).astype(bool)


def _read_flat(r):
    """One block -> (encoding name, values array, null mask or None, offsets or None, positionCount), RLE / DICTIONARY flattened."""
    name = bytes(r.array(np.uint8, r.int32())).decode()
    n = r.int32()
    if name == "RLE":
        inner, values, mask, offs, one = _read_flat(r)
        if one != 1:
            raise ValueError("RLE value block has %d positions" % one)
        idx = np.zeros(n, dtype=np.int64)
        return (inner,) + _take(inner, values, mask, offs, idx) + (n,)
    if name == "DICTIONARY":
        inner, values, mask, offs, _ = _read_flat(r)
        ids = r.array(np.int32, n).astype(np.int64)
        r.at += 24                                   # DictionaryId: most / least significant bits, sequence id
        return (inner,) + _take(inner, values, mask, offs, ids) + (n,)
    if name == "VARIABLE_WIDTH":
        ends = r.array(np.int32, n)
        mask = r.nulls(n)
        total = r.int32()
        data = r.array(np.uint8, total).copy()
        offs = np.zeros(n + 1, dtype=np.int32)
        offs[1:] = ends
        return name, data, mask, offs, n
    if name not in _WIDTH:
        raise ValueError("block encoding %s is outside the hash aggregation / join path" % name)
    per_row = 2 if name == "INT128_ARRAY" else 1
    mask = r.nulls(n)
    if mask is None:
        return name, r.array(_VALUE_DTYPE[name], n * per_row).copy(), None, None, n
    kept = r.array(_VALUE_DTYPE[name], r.int32() * per_row).reshape(-1, per_row)
    values = np.zeros((n, per_row), dtype=kept.dtype)
    values[~mask] = kept                             # the wire carries the non-null values only, in order
    return name, values.reshape(-1), mask, None, n


def _take(name, values, mask, offs, idx):
    """Rows idx of a flat block (Block.getPositions)."""
    m = None if mask is None else mask[idx]
    if name == "VARIABLE_WIDTH":
        lens = (offs[1:] - offs[:-1])[idx]
        out_offs = np.zeros(len(idx) + 1, dtype=np.int32)
        np.cumsum(lens, out=out_offs[1:])
        pieces = [values[offs[i]:offs[i + 1]] for i in idx]
        data = np.concatenate(pieces) if pieces else np.zeros(0, dtype=np.uint8)
        return data, m, out_offs
    per_row = 2 if name == "INT128_ARRAY" else 1
    return values.reshape(-1, per_row)[idx].reshape(-1), m, None


def deserialize_page(buf, types, at=0):
    """PagesSerdeUtil.readSerializedPage + PagesSerde.deserialize: (Page of host blocks typed by `types`, offset behind the page).
    `types`: the plan's channel types — the encoding name tells the storage width, not BIGINT from DOUBLE or a short decimal."""
    r = _Reader(buf, at)
    positions, markers, uncompressed, size = r.int32(), r.byte(), r.int32(), r.int32()
    if markers & (COMPRESSED | ENCRYPTED):
        raise ValueError("compressed / encrypted pages are not supported (markers %d)" % markers)
    if uncompressed != size:
        raise ValueError("uncompressed size %d != size %d on an uncompressed page" % (uncompressed, size))
    end = r.at + size
    channels = r.int32()
    if channels != len(types):
        raise ValueError("the page has %d channels, the plan expects %d" % (channels, len(types)))
    blocks = []
    for k, t in enumerate(types):
        name, values, mask, offs, n = _read_flat(r)
        if t not in _TYPES_OF[name]:
            raise ValueError("channel %d: encoding %s cannot hold type %d" % (k, name, t))
        if n != positions:
            raise ValueError("channel %d has %d positions, the page %d" % (k, n, positions))
        nulls = None if mask is None else mask.astype(np.uint8)
        blocks.append(Block(t, np.ascontiguousarray(values), nulls, offs, n))
    if r.at != end:
        raise ValueError("page body is %d bytes, its header says %d" % (r.at - (end - size), size))
    return Page(*blocks, position_count=positions), end


def read_pages(buf, types):
    """PagesSerdeUtil.readPages: every page of the buffer."""
    pages, at = [], 0
    while at < len(buf):
        page, at = deserialize_page(buf, types, at)
        pages.append(page)
    return pages
