//! How `RadixKey::get_level` is lowered for the device (include/b200sort.h: `b200_key_desc`).

use crate::radix_key::RadixKey;
use rdst_b200_sys as sys;

#[derive(Copy, Clone, Debug, PartialEq, Eq)]
#[repr(u32)]
pub enum KeyKind {
    /// `(self >> 8*level) as u8` (rdst src/radix_key_impl.rs:3-76)
    Uint = sys::B200_KEY_UINT,
    /// `((self ^ MIN) >> 8*level) as u8` (src/radix_key_impl.rs:87-160)
    Sint = sys::B200_KEY_SINT,
    /// IEEE total-order image, then as `Sint` (src/radix_key_impl.rs:162-185)
    Float = sys::B200_KEY_FLOAT,
    /// level `l` reads record byte `level[l].0` and xors it with `level[l].1`
    Bytes = sys::B200_KEY_BYTES,
}

/// A key descriptor: what the kernels need to extract every level's digit in registers.
#[derive(Copy, Clone)]
pub struct KeyDesc(pub(crate) sys::b200_key_desc);

impl KeyDesc {
    fn blank<T>(levels: usize, kind: KeyKind) -> sys::b200_key_desc {
        assert!(levels <= sys::B200_MAX_LEVELS, "at most {} levels", sys::B200_MAX_LEVELS);
        sys::b200_key_desc {
            elem_size: std::mem::size_of::<T>() as u32,
            elem_align: std::mem::align_of::<T>() as u32,
            levels: levels as u32,
            kind: kind as u32,
            key_offset: 0,
            key_width: 0,
            level: [sys::b200_level::default(); sys::B200_MAX_LEVELS],
        }
    }

    /// A little-endian integer or float word of `width` bytes at byte `offset` of `T`, of which the
    /// lowest `levels` bytes are the key (`levels < width`: a partial key, rdst src/lib.rs:58-76).
    /// `KeyDesc::word::<LargeStruct>(KeyKind::Float, offset_of!(LargeStruct, sort_key), 4, 4)` is the
    /// reference's benches/struct_sort.rs key.
    pub fn word<T>(kind: KeyKind, offset: usize, width: usize, levels: usize) -> Self {
        assert!(kind != KeyKind::Bytes);
        assert!(offset + width <= std::mem::size_of::<T>() && levels <= width);
        let mut d = Self::blank::<T>(levels, kind);
        d.key_offset = offset as u32;
        d.key_width = width as u32;
        KeyDesc(d)
    }

    /// A field of a struct: `KeyDesc::field::<Rec, u32>(offset_of!(Rec, key), 3)`.
    pub fn field<T, K: GpuRadixKey>(offset: usize, levels: usize) -> Self {
        let k = K::descriptor().0;
        assert!(k.kind != sys::B200_KEY_BYTES, "use KeyDesc::bytes for byte-string fields");
        let kind = match k.kind {
            sys::B200_KEY_UINT => KeyKind::Uint,
            sys::B200_KEY_SINT => KeyKind::Sint,
            _ => KeyKind::Float,
        };
        Self::word::<T>(kind, offset, k.key_width as usize, levels)
    }

    /// Any byte order: `levels[l] = (byte offset inside T, xor mask)`, level 0 first.
    /// rdst's examples/impl_radix_key.rs `PackedU8` is `&[(3, 0), (2, 0), (1, 0), (0, 0)]`.
    pub fn bytes<T>(levels: &[(usize, u8)]) -> Self {
        let mut d = Self::blank::<T>(levels.len(), KeyKind::Bytes);
        for (l, &(off, x)) in levels.iter().enumerate() {
            assert!(off < std::mem::size_of::<T>());
            d.level[l] = sys::b200_level { byte_offset: off as u16, xor_mask: x, _pad: 0 };
        }
        KeyDesc(d)
    }

    pub fn levels(&self) -> usize {
        self.0.levels as usize
    }

    /// `get_level` evaluated through the descriptor (what the device computes): `b200_get_level`.
    pub fn get_level<T>(&self, record: &T, level: usize) -> u8 {
        let r = unsafe { sys::b200_get_level(record as *const T as *const _, &self.0, level as u32) };
        if r < 0 {
            crate::engine_panic(-r);
        }
        r as u8
    }
}

/// A `RadixKey` the engine can sort: the key's levels as a static description.
/// `descriptor().levels()` must equal `LEVELS`, and `descriptor().get_level(x, l) == x.get_level(l)`
/// for every record and level — the validator checks both on sampled records.
pub trait GpuRadixKey: RadixKey {
    fn descriptor() -> KeyDesc;
}

macro_rules! word_key {
    ($kind:expr; $($t:ty),*) => {$(
        impl GpuRadixKey for $t {
            fn descriptor() -> KeyDesc {
                let w = std::mem::size_of::<$t>();
                KeyDesc::word::<$t>($kind, 0, w, w)
            }
        }
    )*};
}
word_key!(KeyKind::Uint; u8, u16, u32, u64, u128, usize);
word_key!(KeyKind::Sint; i8, i16, i32, i64, i128, isize);
word_key!(KeyKind::Float; f32, f64);

impl<const N: usize> GpuRadixKey for [u8; N] {
    fn descriptor() -> KeyDesc {
        let mut d = KeyDesc::blank::<[u8; N]>(N, KeyKind::Bytes);
        for l in 0..N {
            d.level[l] = sys::b200_level { byte_offset: (N - 1 - l) as u16, xor_mask: 0, _pad: 0 };
        }
        KeyDesc(d)
    }
}
