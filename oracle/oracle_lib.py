"""ctypes wrapper of oracle/liboracle.so — TEST INFRASTRUCTURE ONLY (the CPU restatement of the reference).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this module.
It speaks the same pg_col structs as the product library so every parity test is "run both, compare".
"""
import ctypes as C
import os
import subprocess

import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import presto_b200 as P
from presto_b200 import _native as N
from presto_b200.blocks import Block, OutBlock, col_array

ORACLE_DIR = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(ORACLE_DIR)
LIB_PATH = os.path.join(ORACLE_DIR, "liboracle.so")

_lib = None
P_, I32, I64 = C.c_void_p, C.c_int32, C.c_int64
PCOL, PCOLOUT = C.POINTER(N.PgCol), C.POINTER(N.PgColOut)


def build():
    srcs = [os.path.join(ORACLE_DIR, f) for f in os.listdir(ORACLE_DIR) if f.endswith((".cpp", ".h"))]
    if os.path.exists(LIB_PATH) and all(os.path.getmtime(LIB_PATH) >= os.path.getmtime(s) for s in srcs):
        return LIB_PATH
    subprocess.run(["make", "-C", ORACLE_DIR, "liboracle.so"], check=True, capture_output=True)
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        l = C.CDLL(LIB_PATH)
        protos = {
            "orc_long_hash": (I64, [I64]), "orc_xxh64_long": (I64, [I64]), "orc_xxh64_bytes": (I64, [P_, I64]),
            "orc_fmix64": (C.c_uint64, [C.c_uint64]), "orc_array_size": (I64, [I64]), "orc_max_fill": (I64, [I64]),
            "orc_dec_add": (None, [P_, P_, P_]), "orc_dec_avg": (I32, [P_, I64, I64, P_]), "orc_dec_exceeds_1e38": (I32, [P_]),
            "orc_hash_rows": (None, [PCOL, I32, I64, P_]), "orc_partition": (None, [P_, I64, I32, I32, P_]),
            "orc_groupby_create": (P_, [P_, I32, I32, I64]), "orc_groupby_destroy": (None, [P_]),
            "orc_groupby_get_group_ids": (I32, [P_, PCOL, PCOL, I64, P_]), "orc_groupby_group_count": (I64, [P_]),
            "orc_groupby_capacity": (I64, [P_]), "orc_groupby_collisions": (I64, [P_]),
            "orc_groupby_contains": (None, [P_, PCOL, PCOL, I64, P_]), "orc_groupby_export": (I32, [P_, I64, I64, PCOLOUT, P_]),
            "orc_groupby_varwidth_bytes": (I64, [P_, I32]),
            "orc_agg_create": (P_, [I32, I32]), "orc_agg_destroy": (None, [P_]),
            "orc_agg_add_input": (I32, [P_, P_, I64, PCOL, PCOL, I64]), "orc_agg_add_intermediate": (I32, [P_, P_, I64, PCOL, I64]),
            "orc_agg_n_intermediate": (I32, [P_]), "orc_agg_evaluate": (I32, [P_, I64, I64, I32, PCOLOUT, P_]),
            "orc_join_create": (P_, [P_, I32, P_, I32, I32]), "orc_join_destroy": (None, [P_]),
            "orc_join_add_page": (None, [P_, PCOL, I64]), "orc_join_build": (None, [P_]), "orc_join_rows": (I64, [P_]),
            "orc_join_hash_size": (I64, [P_]), "orc_join_collisions": (I64, [P_]),
            "orc_join_probe": (I64, [P_, PCOL, I32, PCOL, I64, I32]), "orc_join_fetch": (None, [P_, P_, P_]),
            "orc_project_create": (P_, [P_, I32, I32, P_, I32]), "orc_project_destroy": (None, [P_]),
            "orc_project_types": (None, [P_, P_, P_, P_]), "orc_project_process": (I32, [P_, PCOL, I64, PCOLOUT, P_]),
        }
        for name, (res, args) in protos.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def _host(blocks):
    return [b.to_host() for b in blocks]


class OracleGroupBy:
    def __init__(self, types, expected_size, has_hash_channel=False):
        self.types = list(types)
        self.has_hash_channel = has_hash_channel
        arr = (C.c_int32 * len(types))(*types)
        self._h = lib().orc_groupby_create(arr, len(types), int(has_hash_channel), expected_size)

    def _split(self, blocks):
        blocks = _host(blocks.blocks if hasattr(blocks, "blocks") else blocks)
        nk = len(self.types)
        return blocks[:nk], (blocks[nk] if self.has_hash_channel and len(blocks) > nk else None)

    def get_group_ids(self, blocks):
        keys, hb = self._split(blocks)
        n = keys[0].n
        ids = np.empty(n, dtype=np.int32)
        hc = hb.pg_col() if hb is not None else None
        rc = lib().orc_groupby_get_group_ids(self._h, col_array(keys), C.byref(hc) if hc else None, n, ids.ctypes.data)
        assert rc == 0, rc
        return ids

    def add_page(self, blocks):
        keys, hb = self._split(blocks)
        hc = hb.pg_col() if hb is not None else None
        rc = lib().orc_groupby_get_group_ids(self._h, col_array(keys), C.byref(hc) if hc else None, keys[0].n, None)
        assert rc == 0, rc

    def contains(self, blocks):
        keys, hb = self._split(blocks)
        out = np.zeros(keys[0].n, dtype=np.uint8)
        hc = hb.pg_col() if hb is not None else None
        lib().orc_groupby_contains(self._h, col_array(keys), C.byref(hc) if hc else None, keys[0].n, out.ctypes.data)
        return out.astype(bool)

    def group_count(self):
        return lib().orc_groupby_group_count(self._h)

    def capacity(self):
        return lib().orc_groupby_capacity(self._h)

    def export(self, first=0, n=None, with_hash=True):
        if n is None:
            n = self.group_count() - first
        outs = [OutBlock(t, n, lib().orc_groupby_varwidth_bytes(self._h, k)) for k, t in enumerate(self.types)]
        arr = (N.PgColOut * len(outs))()
        for i, o in enumerate(outs):
            arr[i] = o.pg_col_out()
        hashes = np.zeros(n, dtype=np.int64)
        rc = lib().orc_groupby_export(self._h, first, n, arr, hashes.ctypes.data if with_hash else None)
        assert rc == 0, rc
        return [o.block(n) for o in outs], hashes

    def __del__(self):
        if lib and self._h:
            lib().orc_groupby_destroy(self._h)
            self._h = None


class OracleAccumulator:
    def __init__(self, kind, in_type=0):
        self.kind, self.in_type = kind, in_type
        self._h = lib().orc_agg_create(kind, in_type)

    def add_input(self, ids, group_count, arg=None, mask=None):
        ids = np.ascontiguousarray(np.asarray(ids, dtype=np.int32))
        a = arg.to_host().pg_col() if arg is not None else None
        m = mask.to_host().pg_col() if mask is not None else None
        return lib().orc_agg_add_input(self._h, ids.ctypes.data, group_count, C.byref(a) if a else None, C.byref(m) if m else None,
                                       len(ids))

    def add_intermediate(self, ids, group_count, state_blocks):
        ids = np.ascontiguousarray(np.asarray(ids, dtype=np.int32))
        blocks = _host(state_blocks)
        return lib().orc_agg_add_intermediate(self._h, ids.ctypes.data, group_count, col_array(blocks), len(ids))

    def evaluate(self, first, n, final=True):
        types = [P.final_type(self.kind, self.in_type)] if final else P.intermediate_types(self.kind, self.in_type)
        outs = [OutBlock(t, n, 32 * n if t == N.VARBINARY else 0) for t in types]
        arr = (N.PgColOut * len(outs))()
        for i, o in enumerate(outs):
            arr[i] = o.pg_col_out()
        eg = C.c_int64(-1)
        rc = lib().orc_agg_evaluate(self._h, first, n, 1 if final else 0, arr, C.byref(eg))
        return rc, [o.block(n) for o in outs], eg.value

    def __del__(self):
        if lib and self._h:
            lib().orc_agg_destroy(self._h)
            self._h = None


class OracleJoin:
    def __init__(self, types, join_channels, hash_channel=-1):
        t = (C.c_int32 * len(types))(*types)
        j = (C.c_int32 * len(join_channels))(*join_channels)
        self.types, self.join_channels = list(types), list(join_channels)
        self._h = lib().orc_join_create(t, len(types), j, len(join_channels), hash_channel)

    def add_page(self, blocks):
        blocks = _host(blocks.blocks if hasattr(blocks, "blocks") else blocks)
        lib().orc_join_add_page(self._h, col_array(blocks), blocks[0].n)

    def build(self):
        lib().orc_join_build(self._h)

    def rows(self):
        return lib().orc_join_rows(self._h)

    def probe(self, key_blocks, hash_block=None, join_type=0):
        keys = _host(key_blocks)
        hc = hash_block.to_host().pg_col() if hash_block is not None else None
        n = lib().orc_join_probe(self._h, col_array(keys), len(keys), C.byref(hc) if hc else None, keys[0].n, join_type)
        op = np.zeros(n, dtype=np.int32)
        ob = np.zeros(n, dtype=np.int64)
        lib().orc_join_fetch(self._h, op.ctypes.data, ob.ctypes.data)
        return op, ob

    def __del__(self):
        if lib and self._h:
            lib().orc_join_destroy(self._h)
            self._h = None


class OraclePageProcessor:
    """Filter + projections evaluated row by row (oracle_project.h).  Takes the same Expr trees as presto_b200.PageProcessor."""

    def __init__(self, filter=None, projections=()):
        from presto_b200.expressions import flatten
        self.projections = list(projections)
        roots = ([filter] if filter is not None else []) + self.projections
        arr, index, keep, n = flatten(roots)
        self._keep = (arr, keep)
        proj = (C.c_int32 * max(1, len(self.projections)))(*[index[id(p)] for p in self.projections])
        self._h = lib().orc_project_create(arr, n, index[id(filter)] if filter is not None else -1, proj, len(self.projections))
        k = len(self.projections)
        t, p, s = (C.c_int32 * max(1, k))(), (C.c_int32 * max(1, k))(), (C.c_int32 * max(1, k))()
        lib().orc_project_types(self._h, t, p, s)
        self.output_types = [(t[i], p[i], s[i]) for i in range(k)]

    def process(self, page):
        """Returns (status, projected host page)."""
        blocks = _host(page.blocks if hasattr(page, "blocks") else page)
        n = blocks[0].n
        outs = []
        for (t, p, s), e in zip(self.output_types, self.projections):
            vb = len(blocks[e.channel].values) if t in (N.VARCHAR, N.VARBINARY) else 0
            outs.append(OutBlock(t, n, vb))
        arr = (N.PgColOut * max(1, len(outs)))()
        for i, o in enumerate(outs):
            arr[i] = o.pg_col_out()
        n_out = C.c_int64()
        rc = lib().orc_project_process(self._h, col_array(blocks), n, arr, C.byref(n_out))
        m = n_out.value
        res = []
        for o, (t, p, s) in zip(outs, self.output_types):
            b = o.block(m)
            b.precision, b.scale = p, s
            res.append(b)
        return rc, P.Page(*res, position_count=m)

    def __del__(self):
        if lib and self._h:
            lib().orc_project_destroy(self._h)
            self._h = None


def hash_rows(blocks):
    blocks = _host(blocks)
    out = np.zeros(blocks[0].n, dtype=np.int64)
    lib().orc_hash_rows(col_array(blocks), len(blocks), blocks[0].n, out.ctypes.data)
    return out


def partition_ids(blocks, nparts, mode=0, hash_block=None):
    """HashGenerator.getPartition (mode 0) / LocalPartitionGenerator.getPartition (mode 1) for every row."""
    raw = hash_block.to_host().values if hash_block is not None else hash_rows(blocks)
    raw = np.ascontiguousarray(raw, dtype=np.int64)
    out = np.zeros(len(raw), dtype=np.int32)
    # oracle.cpp numbers the functions the other way round (0 = LocalPartitionGenerator, 1 = HashGenerator.getPartition)
    lib().orc_partition(raw.ctypes.data, len(raw), nparts, 1 - mode, out.ctypes.data)
    return out


class OraclePartitioner:
    """CPU stand-in for DevicePartitioner used ONLY by the gloo world_size-2 tests of the exchange's host logic."""

    def partition(self, page, key_channels, nparts, hash_channel=None, mode=0):
        blocks = _host(page.blocks)
        keys = [blocks[c] for c in key_channels]
        ids = partition_ids(keys, nparts, mode, blocks[hash_channel] if hash_channel is not None else None)
        order = np.argsort(ids, kind="stable")
        counts = [int((ids == p).sum()) for p in range(nparts)]
        out = []
        for b in blocks:
            k = 2 if b.type == N.LONG_DECIMAL else 1
            v = b.values.reshape(-1, k)[order].reshape(-1) if k == 2 else b.values[order]
            out.append(Block(b.type, np.ascontiguousarray(v), None if b.nulls is None else np.ascontiguousarray(b.nulls[order]), None, b.n))
        return P.Page(*out, position_count=page.position_count), counts
