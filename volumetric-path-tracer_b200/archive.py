""".vptscene / .vptarch named-array container — Python mirror of host/vpt_archive.h.

Layout (little endian): b"VPTARCH1", u32 count, then per array: u32 name_len,
name, u32 dtype (0 f32, 1 i32, 2 u8, 3 u64), u32 ndim, u64 dims[ndim],
u64 nbytes, raw data.
"""
from __future__ import annotations

import struct
from collections import OrderedDict

import numpy as np

_DTYPES = [np.float32, np.int32, np.uint8, np.uint64]


def load(path: str) -> "OrderedDict[str, np.ndarray]":
    out: "OrderedDict[str, np.ndarray]" = OrderedDict()
    with open(path, "rb") as f:
        if f.read(8) != b"VPTARCH1":
            raise ValueError(f"not a vptscene archive: {path}")
        (count,) = struct.unpack("<I", f.read(4))
        for _ in range(count):
            (name_len,) = struct.unpack("<I", f.read(4))
            name = f.read(name_len).decode()
            dtype, ndim = struct.unpack("<II", f.read(8))
            dims = struct.unpack(f"<{ndim}Q", f.read(8 * ndim))
            (nbytes,) = struct.unpack("<Q", f.read(8))
            raw = f.read(nbytes)
            if len(raw) != nbytes:
                raise ValueError(f"corrupt vptscene archive: {path}")
            out[name] = np.frombuffer(raw, dtype=_DTYPES[dtype]).reshape(dims).copy()
    return out


def save(path: str, arrays: "dict[str, np.ndarray]") -> None:
    with open(path, "wb") as f:
        f.write(b"VPTARCH1")
        f.write(struct.pack("<I", len(arrays)))
        for name, arr in arrays.items():
            arr = np.ascontiguousarray(arr)
            for code, dt in enumerate(_DTYPES):
                if arr.dtype == dt:
                    break
            else:
                raise TypeError(f"{name}: unsupported dtype {arr.dtype}")
            enc = name.encode()
            f.write(struct.pack("<I", len(enc)))
            f.write(enc)
            f.write(struct.pack("<II", code, arr.ndim))
            f.write(struct.pack(f"<{arr.ndim}Q", *arr.shape))
            f.write(struct.pack("<Q", arr.nbytes))
            f.write(arr.tobytes())
