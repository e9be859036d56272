"""Host-side mirror of the reference's columnar containers (SPI/Page.java, SPI/block/LongArrayBlock.java:37-76,
IntArrayBlock.java, Int128ArrayBlock.java:37-46, VariableWidthBlock.java:38-48, ByteArrayBlock.java).

A Block holds either numpy arrays (host memory, what a JNI caller would hand over) or torch CUDA tensors
(device-resident column buffers).  Null masks are one byte per row like Java's boolean[] valueIsNull.
"""
import ctypes as C

import numpy as np

from . import _native as N

_NP_DTYPE = {
    N.BIGINT: np.int64, N.INTEGER: np.int32, N.DATE: np.int32, N.DOUBLE: np.int64, N.SHORT_DECIMAL: np.int64,
    N.LONG_DECIMAL: np.uint64, N.VARCHAR: np.uint8, N.VARBINARY: np.uint8, N.BOOLEAN: np.uint8,
}
TYPE_NAMES = {
    N.BIGINT: "bigint", N.INTEGER: "integer", N.DATE: "date", N.DOUBLE: "double", N.SHORT_DECIMAL: "decimal(short)",
    N.LONG_DECIMAL: "decimal(long)", N.VARCHAR: "varchar", N.VARBINARY: "varbinary", N.BOOLEAN: "boolean",
}


def _is_torch(x):
    return x is not None and type(x).__module__.startswith("torch")


def _ptr(x):
    if x is None:
        return None
    if _is_torch(x):
        return x.data_ptr()
    return x.ctypes.data


class Block:
    """One column.  `values` layout per type: int64 (BIGINT, DOUBLE as raw bits, SHORT_DECIMAL), int32 (INTEGER, DATE),
    uint64[2n] lo/hi sign-magnitude (LONG_DECIMAL), uint8 bytes + int32 offsets[n+1] (VARCHAR/VARBINARY), uint8 (BOOLEAN)."""

    def __init__(self, type_, values, nulls=None, offsets=None, n=None, precision=0, scale=0):
        self.type = type_
        self.values = values
        self.nulls = nulls
        self.offsets = offsets
        self.precision = precision
        self.scale = scale
        self.device = bool(_is_torch(values) and values.is_cuda)
        self.pinned = False      # host block in page-locked memory that the caller keeps alive until the operator consumed it (PG_MEM_HOST_PINNED)
        if n is None:
            if type_ in (N.VARCHAR, N.VARBINARY):
                n = len(offsets) - 1
            elif type_ == N.LONG_DECIMAL:
                n = len(values) // 2
            else:
                n = len(values)
        self.n = int(n)

    # ---- constructors named after the reference's block classes
    @staticmethod
    def long_array(values, nulls=None, type_=N.BIGINT):
        return Block(type_, _as_host(values, np.int64), _as_nulls(nulls))

    @staticmethod
    def double_array(values, nulls=None):
        v = np.ascontiguousarray(np.asarray(values, dtype=np.float64)).view(np.int64)
        return Block(N.DOUBLE, v, _as_nulls(nulls))

    @staticmethod
    def int_array(values, nulls=None, type_=N.INTEGER):
        return Block(type_, _as_host(values, np.int32), _as_nulls(nulls))

    @staticmethod
    def byte_array(values, nulls=None):
        return Block(N.BOOLEAN, _as_host(values, np.uint8), _as_nulls(nulls))

    @staticmethod
    def int128_array(py_ints, nulls=None):
        """Long decimals from python ints -> sign-magnitude lo/hi pairs (UnscaledDecimal128Arithmetic.java:165-176)."""
        out = np.zeros(2 * len(py_ints), dtype=np.uint64)
        for i, v in enumerate(py_ints):
            if v is None:
                continue
            lo, hi = encode_decimal128(int(v))
            out[2 * i] = lo
            out[2 * i + 1] = hi
        if nulls is None and any(v is None for v in py_ints):
            nulls = [v is None for v in py_ints]
        return Block(N.LONG_DECIMAL, out, _as_nulls(nulls))

    @staticmethod
    def variable_width(strings, type_=N.VARCHAR):
        """strings: list of bytes/str/None."""
        offs = np.zeros(len(strings) + 1, dtype=np.int32)
        chunks = []
        nulls = np.zeros(len(strings), dtype=np.uint8)
        pos = 0
        for i, s in enumerate(strings):
            if s is None:
                nulls[i] = 1
            else:
                b = s.encode() if isinstance(s, str) else bytes(s)
                chunks.append(b)
                pos += len(b)
            offs[i + 1] = pos
        data = np.frombuffer(b"".join(chunks), dtype=np.uint8).copy() if pos else np.zeros(0, dtype=np.uint8)
        return Block(type_, data, nulls if nulls.any() else None, offs)

    def slice(self, lo, hi):
        """Rows [lo, hi) as a new host block (Block.getRegion; var-width offsets are rebased)."""
        h = self.to_host()
        nulls = None if h.nulls is None else np.ascontiguousarray(h.nulls[lo:hi])
        if h.type in (N.VARCHAR, N.VARBINARY):
            o = np.asarray(h.offsets)
            vals = np.ascontiguousarray(np.asarray(h.values)[o[lo]:o[hi]])
            return Block(h.type, vals, nulls, (o[lo:hi + 1] - o[lo]).astype(np.int32), hi - lo, h.precision, h.scale)
        k = 2 if h.type == N.LONG_DECIMAL else 1
        return Block(h.type, np.ascontiguousarray(np.asarray(h.values)[k * lo:k * hi]), nulls, None, hi - lo, h.precision, h.scale)

    # ---- conversions
    def to_device(self, device="cuda"):
        import torch
        if self.device:
            return self
        t = lambda a: None if a is None else torch.from_numpy(np.ascontiguousarray(a).view(_sv(a))).to(device)
        return Block(self.type, t(self.values), t(self.nulls), t(self.offsets), self.n, self.precision, self.scale)

    def to_host(self):
        if not self.device and not _is_torch(self.values):
            return self
        h = lambda a, dt: None if a is None else a.cpu().numpy().view(dt)
        return Block(self.type, h(self.values, _NP_DTYPE[self.type]), h(self.nulls, np.uint8), h(self.offsets, np.int32), self.n,
                     self.precision, self.scale)

    def pg_col(self):
        c = N.PgCol()
        c.type = self.type
        c.mem = N.MEM_DEVICE if self.device else (N.MEM_HOST_PINNED if self.pinned else N.MEM_HOST)
        c.n = self.n
        c.values = _ptr(self.values)
        c.nulls = _ptr(self.nulls)
        c.offsets = _ptr(self.offsets)
        c.precision = self.precision
        c.scale = self.scale
        return c

    # ---- python-level accessors (host blocks only; used by tests)
    def is_null(self, i):
        return self.nulls is not None and bool(self.nulls[i])

    def get(self, i):
        b = self.to_host()
        if b.is_null(i):
            return None
        if b.type in (N.VARCHAR, N.VARBINARY):
            return bytes(b.values[b.offsets[i]:b.offsets[i + 1]])
        if b.type == N.LONG_DECIMAL:
            return decode_decimal128(int(b.values[2 * i]), int(b.values[2 * i + 1]))
        if b.type == N.DOUBLE:
            return float(np.int64(b.values[i]).view(np.float64))
        return int(b.values[i])

    def to_list(self):
        return [self.get(i) for i in range(self.n)]

    def __len__(self):
        return self.n


def _sv(a):
    # torch has no uint64 arithmetic but can carry the bits: view as int64
    return np.int64 if a.dtype == np.uint64 else a.dtype


def _as_host(values, dt):
    if _is_torch(values):
        return values
    return np.ascontiguousarray(np.asarray(values, dtype=dt))


def _as_nulls(nulls):
    if nulls is None:
        return None
    if _is_torch(nulls):
        return nulls
    return np.ascontiguousarray(np.asarray(nulls, dtype=np.uint8))


SIGN = 1 << 63


def encode_decimal128(v):
    m = -v if v < 0 else v
    assert m < (1 << 127)
    return m & ((1 << 64) - 1), (m >> 64) | (SIGN if v < 0 else 0)


def decode_decimal128(lo, hi):
    m = ((hi & (SIGN - 1)) << 64) | lo
    return -m if hi & SIGN else m


class Page:
    """SPI/Page.java: positionCount + blocks."""

    def __init__(self, *blocks, position_count=None):
        self.blocks = list(blocks)
        self.position_count = blocks[0].n if position_count is None else position_count
        for b in self.blocks:
            assert b.n == self.position_count, "all blocks of a page have the same positionCount"

    def get_block(self, channel):
        return self.blocks[channel]

    def get_channel_count(self):
        return len(self.blocks)

    def get_position_count(self):
        return self.position_count

    def to_device(self, device="cuda"):
        return Page(*[b.to_device(device) for b in self.blocks], position_count=self.position_count)


def col_array(blocks):
    """ctypes array of pg_col for a list of blocks (keeps nothing alive: callers hold the blocks)."""
    arr = (N.PgCol * max(1, len(blocks)))()
    for i, b in enumerate(blocks):
        arr[i] = b.pg_col()
    return arr


class OutBlock:
    """Caller-allocated output column (pg_col_out) backed by numpy (host) or torch (device)."""

    def __init__(self, type_, capacity, value_bytes=0, device=False, want_nulls=True, zero=True):
        """zero=False (device only): uninitialised memory, for calls that write every row of every array they are given."""
        self.type = type_
        self.capacity = int(capacity)
        self.device = device
        var = type_ in (N.VARCHAR, N.VARBINARY)
        cap = max(1, self.capacity)
        if var:
            nvals, dt = max(1, int(value_bytes)), np.uint8
        elif type_ == N.LONG_DECIMAL:
            nvals, dt = 2 * cap, np.uint64
        else:
            nvals, dt = cap, _NP_DTYPE[type_]
        if device:
            import torch
            tdt = {np.int64: torch.int64, np.uint64: torch.int64, np.int32: torch.int32, np.uint8: torch.uint8}[dt]
            new = torch.zeros if zero else torch.empty
            self.values = new(nvals, dtype=tdt, device="cuda")
            self.nulls = new(cap, dtype=torch.uint8, device="cuda") if want_nulls else None
            self.offsets = new(cap + 1, dtype=torch.int32, device="cuda") if var else None
        else:
            self.values = np.zeros(nvals, dtype=dt)
            self.nulls = np.zeros(cap, dtype=np.uint8) if want_nulls else None
            self.offsets = np.zeros(cap + 1, dtype=np.int32) if var else None
        self.value_bytes = int(value_bytes)

    def pg_col_out(self):
        o = N.PgColOut()
        o.type = self.type
        o.mem = N.MEM_DEVICE if self.device else N.MEM_HOST
        o.capacity = self.capacity
        o.values = _ptr(self.values)
        o.nulls = _ptr(self.nulls)
        o.offsets = _ptr(self.offsets)
        o.values_capacity_bytes = self.value_bytes
        return o

    def block(self, n):
        """View the first n rows as a Block."""
        var = self.type in (N.VARCHAR, N.VARBINARY)
        if var:
            return Block(self.type, self.values, None if self.nulls is None else self.nulls[:n], self.offsets[:n + 1], n)
        k = 2 if self.type == N.LONG_DECIMAL else 1
        return Block(self.type, self.values[:k * n], None if self.nulls is None else self.nulls[:n], None, n)
