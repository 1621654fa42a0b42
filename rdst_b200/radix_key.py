"""RadixKey — the key trait of the reference, lowered to a byte-extraction descriptor.

Reference: `pub trait RadixKey { const LEVELS: usize; fn get_level(&self, level) -> u8 }`
(reference src/radix_key.rs:1-5); level 0 is the LEAST significant byte
(reference README.md:37-38).  An arbitrary `get_level` cannot run on the device, so a key
is described as: for each level, which byte of the record holds the digit and which
transform applies (UINT / SINT / FLOAT word kinds, or a per-level byte table).  The
built-in impls mirror reference src/radix_key_impl.rs.
"""
from __future__ import annotations

import numpy as np

from . import _abi


class RadixKey:
    """Descriptor of how to read the LEVELS key bytes of one record type.

    kind        'uint' | 'sint' | 'float' | 'bytes'
    elem_size   size_of::<T>() — the record stride, payload included
    levels      T::LEVELS (may be smaller than key_width: partial keys, reference src/lib.rs:58-76)
    key_offset  byte offset of the little-endian key word inside the record (word kinds)
    key_width   bytes of the key word (word kinds)
    level_map   'bytes' kind: [(byte_offset, xor_mask)] per level, level 0 first
    """

    __slots__ = ("kind", "elem_size", "elem_align", "levels", "key_offset", "key_width", "level_map")

    _KINDS = {"uint": _abi.KEY_UINT, "sint": _abi.KEY_SINT, "float": _abi.KEY_FLOAT, "bytes": _abi.KEY_BYTES}

    def __init__(self, kind, elem_size, levels, key_offset=0, key_width=0, level_map=None, elem_align=1):
        if kind not in self._KINDS:
            raise ValueError(f"unknown key kind {kind!r}")
        self.kind = kind
        self.elem_size = int(elem_size)
        self.elem_align = int(elem_align)
        self.levels = int(levels)
        self.key_offset = int(key_offset)
        self.key_width = int(key_width)
        self.level_map = [(int(o), int(x)) for o, x in (level_map or [])]

    @property
    def LEVELS(self):
        return self.levels

    # ---- constructors mirroring reference src/radix_key_impl.rs ----
    @classmethod
    def for_dtype(cls, dtype, levels=None):
        """Built-in impls: u8..u64 (:3-37), i8..i64 (:87-121), f32/f64 (:162-185)."""
        dt = np.dtype(dtype)
        if dt.kind == "u":
            kind = "uint"
        elif dt.kind == "i":
            kind = "sint"
        elif dt.kind == "f" and dt.itemsize in (4, 8):
            kind = "float"
        elif dt.kind == "b":
            kind = "uint"
        else:
            raise TypeError(f"no built-in RadixKey for dtype {dt}")
        w = dt.itemsize
        return cls(kind, w, w if levels is None else levels, 0, w, elem_align=dt.alignment)

    @classmethod
    def u128(cls):
        return cls("uint", 16, 16, 0, 16, elem_align=16)  # :39-46

    @classmethod
    def i128(cls):
        return cls("sint", 16, 16, 0, 16, elem_align=16)  # :123-130

    @classmethod
    def byte_array(cls, n):
        """`[u8; N]`: level l reads self[N-1-l] — lexicographic order (:78-85)."""
        return cls("bytes", n, n, level_map=[(n - 1 - l, 0) for l in range(n)])

    @classmethod
    def field(cls, record_dtype, field, levels=None):
        """A struct sorted by one primitive field (the reference's struct_sort bench shape,
        reference benches/struct_sort.rs:11-27: `impl RadixKey for LargeStruct` delegating
        to the field's get_level)."""
        rdt = np.dtype(record_dtype)
        fdt, off = rdt.fields[field][:2]
        base = cls.for_dtype(fdt, levels)
        return cls(base.kind, rdt.itemsize, base.levels, off, base.key_width, elem_align=rdt.alignment)

    @classmethod
    def custom(cls, elem_size, level_map):
        """User keys (reference examples/impl_radix_key.rs): level l = record byte
        `offset` xor `mask`.  Accepts ints (offset) or (offset, mask) pairs."""
        lm = [(e, 0) if isinstance(e, int) else tuple(e) for e in level_map]
        return cls("bytes", elem_size, len(lm), level_map=lm)

    def with_levels(self, levels):
        return RadixKey(self.kind, self.elem_size, levels, self.key_offset, self.key_width,
                        self.level_map[:levels] if self.kind == "bytes" else None, self.elem_align)

    # ---- lowering ----
    def to_desc(self):
        d = _abi.KeyDesc()
        d.elem_size = self.elem_size
        d.elem_align = self.elem_align
        d.levels = self.levels
        d.kind = self._KINDS[self.kind]
        d.key_offset = self.key_offset
        d.key_width = self.key_width
        if self.kind == "bytes":
            if self.levels > _abi.B200_MAX_LEVELS:
                raise ValueError("too many levels")
            for l, (o, x) in enumerate(self.level_map[: self.levels]):
                d.level[l].byte_offset = o
                d.level[l].xor_mask = x
        return d

    def get_level(self, record_bytes, level):
        """RadixKey::get_level evaluated on the host THROUGH the C ABI (b200_get_level), so a
        binding can validate a user key against its lowering (reference src/radix_key.rs:13-21:
        level >= LEVELS is a debug_assert failure -> here an error)."""
        import ctypes as C

        buf = (C.c_uint8 * self.elem_size).from_buffer_copy(bytes(record_bytes))
        d = self.to_desc()
        r = _abi.lib().b200_get_level(C.cast(buf, C.c_void_p), C.byref(d), level)
        if r < 0:
            _abi.check(-r)
        return r

    def __repr__(self):
        return (f"RadixKey(kind={self.kind!r}, elem_size={self.elem_size}, levels={self.levels}, "
                f"key_offset={self.key_offset}, key_width={self.key_width})")


def key_for_array(arr, key=None):
    """Pick the RadixKey for a numpy array / torch tensor when the caller gave none."""
    if key is not None:
        return key
    dt = getattr(arr, "dtype", None)
    if isinstance(dt, np.dtype):
        if dt.fields is not None:
            raise TypeError("structured arrays need an explicit RadixKey (RadixKey.field(dtype, name))")
        if dt.kind == "V" or dt.kind == "S":
            return RadixKey.byte_array(dt.itemsize)
        return RadixKey.for_dtype(dt)
    # torch dtype
    import torch

    m = {
        torch.uint8: "u1", torch.int8: "i1", torch.int16: "i2", torch.int32: "i4", torch.int64: "i8",
        torch.float32: "f4", torch.float64: "f8", torch.uint16: "u2", torch.uint32: "u4", torch.uint64: "u8",
    }
    if dt in m:
        return RadixKey.for_dtype(m[dt])
    raise TypeError(f"no built-in RadixKey for {dt}; pass key=")
