"""ctypes access to oracle/_build/librdst_oracle.so — the CPU restatement of the reference.
Test infrastructure: imported only from tests/, __graft_entry__.smoke() and bench.py's
CPU-baseline legs."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "_build", "librdst_oracle.so")

ALGO = {"MtOop": 0, "MtLsb": 1, "Scanning": 2, "Recombinating": 3, "Comparative": 4, "LrLsb": 5, "Lsb": 6,
        "Regions": 7, "Ska": 8}
TUNER = {"standard": 0, "low_memory": 1, "single_threaded": 2, "single_algo": 3}

_lib = None


def build():
    srcs = [os.path.join(ORACLE_DIR, f) for f in ("rdst_oracle.c", "rdst_algos.inc", "rdst_pool.h", "rdst_oracle.h")]
    if os.path.exists(ORACLE_SO) and all(os.path.getmtime(ORACLE_SO) >= os.path.getmtime(s) for s in srcs):
        return
    subprocess.run(["make", "-C", ORACLE_DIR], check=True, stdout=subprocess.DEVNULL)


def lib():
    global _lib
    if _lib is None:
        build()
        l = C.CDLL(ORACLE_SO)
        l.oracle_get_level.restype = C.c_int
        l.oracle_get_level.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
        l.oracle_get_level_builtin.restype = C.c_int
        l.oracle_get_level_builtin.argtypes = [C.c_char_p, C.c_void_p, C.c_uint32]
        l.oracle_get_level_bytes.restype = C.c_int
        l.oracle_get_level_bytes.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32]
        l.oracle_get_counts.restype = C.c_int
        l.oracle_get_counts.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_uint32, C.c_void_p,
                                        C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
        l.oracle_get_tile_counts_sorted.restype = C.c_int
        l.oracle_get_tile_counts_sorted.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_uint32,
                                                    C.c_void_p]
        l.oracle_sort.restype = C.c_int
        l.oracle_sort.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
        l.oracle_out_of_place_pass.restype = C.c_int
        l.oracle_out_of_place_pass.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_uint32, C.c_int]
        l.oracle_is_sorted.restype = C.c_int
        l.oracle_is_sorted.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
        l.oracle_hardware_threads.restype = C.c_int
        l.oracle_version.restype = C.c_char_p
        _lib = l
    return _lib


def _desc(key):
    d = key.to_desc()  # same struct layout as oracle_key_desc
    return d


def sort(arr: np.ndarray, key, tuner="standard", algo="Lsb", multi_threaded=True, threads=0):
    """In-place oracle sort of a C-contiguous numpy array of records."""
    assert arr.flags.c_contiguous
    d = _desc(key)
    n = arr.nbytes // key.elem_size
    rc = lib().oracle_sort(arr.ctypes.data, n, C.byref(d), TUNER[tuner], ALGO[algo], int(multi_threaded), threads)
    if rc:
        raise RuntimeError(f"oracle_sort rc={rc}")
    return arr


def sorted_copy(arr, key, **kw):
    out = arr.copy()
    sort(out, key, **kw)
    return out


def stable_sorted(arr, key):
    """rdst's stable order: the Lsb algorithm forced at every level."""
    return sorted_copy(arr, key, tuner="single_algo", algo="Lsb", multi_threaded=False, threads=1)


def get_level(rec_bytes: bytes, key, level):
    d = _desc(key)
    buf = C.create_string_buffer(bytes(rec_bytes), len(rec_bytes))
    return lib().oracle_get_level(buf, C.byref(d), level)


def get_level_builtin(tname, value_bytes, level):
    buf = C.create_string_buffer(bytes(value_bytes), len(value_bytes))
    return lib().oracle_get_level_builtin(tname.encode(), buf, level)


def get_counts(arr, key, level):
    d = _desc(key)
    counts = np.zeros(256, dtype=np.uint64)
    s, f, l = C.c_int(), C.c_int(), C.c_int()
    n = arr.nbytes // key.elem_size
    rc = lib().oracle_get_counts(arr.ctypes.data, n, C.byref(d), level, counts.ctypes.data, C.byref(s), C.byref(f),
                                 C.byref(l))
    assert rc == 0, rc
    return counts, bool(s.value), f.value, l.value


def tile_counts_sorted(arr, key, tile_size, level):
    d = _desc(key)
    n = arr.nbytes // key.elem_size
    counts = np.zeros(256, dtype=np.uint64)
    r = lib().oracle_get_tile_counts_sorted(arr.ctypes.data, n, C.byref(d), tile_size, level, counts.ctypes.data)
    assert r >= 0
    return bool(r), counts


def out_of_place_pass(arr, key, level, lr=False):
    d = _desc(key)
    n = arr.nbytes // key.elem_size
    dst = np.empty_like(arr)
    rc = lib().oracle_out_of_place_pass(arr.ctypes.data, dst.ctypes.data, n, C.byref(d), level, int(lr))
    assert rc == 0
    return dst


def is_sorted(arr, key):
    d = _desc(key)
    n = arr.nbytes // key.elem_size
    return bool(lib().oracle_is_sorted(arr.ctypes.data, n, C.byref(d)))


def hardware_threads():
    return lib().oracle_hardware_threads()


# ---- an INDEPENDENT model of the result definition (not the oracle): numpy stable argsort
# on the key bytes, level LEVELS-1 most significant (reference README.md:37-38) ----
def key_matrix(arr, key):
    """[n, LEVELS] uint8, column l = get_level(l), computed with numpy from the descriptor."""
    raw = arr.view(np.uint8).reshape(-1, key.elem_size)
    L = key.levels
    out = np.empty((raw.shape[0], L), dtype=np.uint8)
    if key.kind == "bytes":
        for l, (o, x) in enumerate(key.level_map[:L]):
            out[:, l] = raw[:, o] ^ np.uint8(x)
        return out
    w = key.key_width
    neg = (raw[:, key.key_offset + w - 1] & 0x80) != 0
    for l in range(L):
        b = raw[:, key.key_offset + l]
        if key.kind == "uint":
            out[:, l] = b
        elif key.kind == "sint":
            out[:, l] = b ^ np.uint8(0x80 if l == w - 1 else 0)
        else:  # float: negative -> flip all bits, else flip sign bit
            out[:, l] = np.where(neg, b ^ np.uint8(0xFF), b ^ np.uint8(0x80 if l == w - 1 else 0))
    return out


def numpy_stable_sorted(arr, key):
    km = key_matrix(arr, key)
    order = np.lexsort([km[:, l] for l in range(key.levels)])  # last key = most significant; stable
    raw = arr.view(np.uint8).reshape(-1, key.elem_size)
    return raw[order].reshape(-1).view(arr.dtype).reshape(arr.shape)
