"""Reader for TensorFlow "bundle" checkpoints (``model.ckpt.index`` + ``.data-*``).

The golden fixtures shipped with the reference
(``resnet/official/utils/testing/reference_data/resnet/*``) store the input tensor
and every block weight in this format; reading it without TensorFlow turns them
into a TF-free numerical oracle.  Format, as far as these files use
it: the index is an uncompressed LevelDB-style table — a 48-byte footer
(metaindex handle, index handle as varints, padding, magic
``0xdb4775248b80fb57``), an index block pointing at data blocks, every block being
``[shared | non_shared | value_len | key_delta | value]*`` + restart array +
``uint32 num_restarts``, followed by a 1-byte compression tag and a 4-byte crc.
Key ``""`` holds the bundle header; every other value is a ``BundleEntryProto``
(field 1 dtype, 2 shape{dim{size}}, 3 shard, 4 offset, 5 size).
"""
from __future__ import annotations

import os
import struct
from typing import Dict, List, Tuple

import numpy as np

TABLE_MAGIC = 0xDB4775248B80FB57
_DTYPES = {1: np.float32, 2: np.float64, 3: np.int32, 9: np.int64, 19: np.float16}


def _varint(buf: bytes, pos: int) -> Tuple[int, int]:
    shift = result = 0
    while True:
        b = buf[pos]
        pos += 1
        result |= (b & 0x7F) << shift
        if not b & 0x80:
            return result, pos
        shift += 7


def _block_entries(buf: bytes, offset: int, size: int) -> List[Tuple[bytes, bytes]]:
    block = buf[offset: offset + size]
    (num_restarts,) = struct.unpack_from("<I", block, len(block) - 4)
    limit = len(block) - 4 - 4 * num_restarts
    out, pos, key = [], 0, b""
    while pos < limit:
        shared, pos = _varint(block, pos)
        non_shared, pos = _varint(block, pos)
        vlen, pos = _varint(block, pos)
        key = key[:shared] + block[pos: pos + non_shared]
        pos += non_shared
        out.append((key, block[pos: pos + vlen]))
        pos += vlen
    return out


def _parse_proto(buf: bytes) -> Dict[int, list]:
    """Minimal protobuf wire decoder: {field: [values]} (varint and length-delimited)."""
    out: Dict[int, list] = {}
    pos = 0
    while pos < len(buf):
        tag, pos = _varint(buf, pos)
        fieldno, wt = tag >> 3, tag & 7
        if wt == 0:
            v, pos = _varint(buf, pos)
        elif wt == 2:
            n, pos = _varint(buf, pos)
            v = buf[pos: pos + n]
            pos += n
        elif wt == 5:
            v = buf[pos: pos + 4]
            pos += 4
        elif wt == 1:
            v = buf[pos: pos + 8]
            pos += 8
        else:
            raise ValueError(f"unsupported wire type {wt}")
        out.setdefault(fieldno, []).append(v)
    return out


def read_index(index_path: str) -> Dict[str, dict]:
    buf = open(index_path, "rb").read()
    footer = buf[-48:]
    (magic,) = struct.unpack_from("<Q", footer, 40)
    if magic != TABLE_MAGIC:
        raise ValueError("not a TF bundle index (bad magic)")
    pos = 0
    _, pos = _varint(footer, pos)      # metaindex offset
    _, pos = _varint(footer, pos)      # metaindex size
    idx_off, pos = _varint(footer, pos)
    idx_size, pos = _varint(footer, pos)
    entries: Dict[str, dict] = {}
    for _, handle in _block_entries(buf, idx_off, idx_size):
        off, p = _varint(handle, 0)
        size, _ = _varint(handle, p)
        for key, value in _block_entries(buf, off, size):
            if key == b"":
                continue
            f = _parse_proto(value)
            shape = []
            if 2 in f:
                for dim in _parse_proto(f[2][0]).get(2, []):
                    shape.append(_parse_proto(dim).get(1, [0])[0])
            entries[key.decode()] = {"dtype": f.get(1, [0])[0], "shape": shape,
                                     "shard": f.get(3, [0])[0], "offset": f.get(4, [0])[0],
                                     "size": f.get(5, [0])[0]}
    return entries


def read_bundle(prefix: str) -> Dict[str, np.ndarray]:
    """``prefix`` is e.g. ``.../model.ckpt``; returns {variable name: array}."""
    entries = read_index(prefix + ".index")
    shards: Dict[int, bytes] = {}
    out: Dict[str, np.ndarray] = {}
    for name, e in entries.items():
        if e["shard"] not in shards:
            n_shards = 1
            path = f"{prefix}.data-{e['shard']:05d}-of-{n_shards:05d}"
            if not os.path.isfile(path):
                cands = [p for p in os.listdir(os.path.dirname(prefix) or ".")
                         if p.startswith(os.path.basename(prefix) + f".data-{e['shard']:05d}-of-")]
                path = os.path.join(os.path.dirname(prefix), cands[0])
            shards[e["shard"]] = open(path, "rb").read()
        raw = shards[e["shard"]][e["offset"]: e["offset"] + e["size"]]
        out[name] = np.frombuffer(raw, dtype=_DTYPES[e["dtype"]]).reshape(e["shape"]).copy()
    return out
