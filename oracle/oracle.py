"""ctypes binding of the CPU oracle (oracle/liboracle.so) and, when built, of the reference's own
host field code (oracle/_ref/libulvt_ref.so).

TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke() and bench.py's CPU-baseline
legs; never from the product package.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_u32p = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")


def build(quiet=True):
    """(Re)build liboracle.so and, if /root/reference is present, oracle/_ref/*."""
    subprocess.run(["make", "-C", HERE, "all"], check=True, stdout=subprocess.DEVNULL if quiet else None)


def _load():
    path = os.path.join(HERE, "liboracle.so")
    if not os.path.exists(path):
        build()
    lib = C.CDLL(path)
    lib.orc_mul32.argtypes = [C.c_uint32, C.c_uint32]
    lib.orc_mul32.restype = C.c_uint32
    lib.orc_mul32_fast.argtypes = [C.c_uint32, C.c_uint32]
    lib.orc_mul32_fast.restype = C.c_uint32
    lib.orc_square32.argtypes = [C.c_uint32]
    lib.orc_square32.restype = C.c_uint32
    lib.orc_inverse32.argtypes = [C.c_uint32]
    lib.orc_inverse32.restype = C.c_uint32
    lib.orc_mul128_words.argtypes = [_u32p, _u32p, _u32p]
    lib.orc_mul_packed32.argtypes = [C.c_uint32, C.c_uint32, C.c_int]
    lib.orc_mul_packed32.restype = C.c_uint32
    lib.orc_interleave32.argtypes = [C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.c_int]
    lib.orc_transpose32.argtypes = [_u32p]
    lib.orc_bitslice.argtypes = [_u32p, C.c_int]
    lib.orc_unbitslice.argtypes = [_u32p, C.c_int]
    lib.orc_mul_bitsliced.argtypes = [_u32p, _u32p, _u32p, C.c_int]
    lib.orc_antt_constants.argtypes = [C.c_int, C.c_int, _u32p]
    lib.orc_antt_constants.restype = C.c_int
    lib.orc_antt_twiddle.argtypes = [_u32p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64]
    lib.orc_antt_twiddle.restype = C.c_uint32
    lib.orc_antt_forward.argtypes = [_u32p, _u32p, C.c_int, C.c_int, C.c_int, C.c_int]
    lib.orc_antt_forward.restype = C.c_int
    lib.orc_antt_forward_full128.argtypes = [_u32p, _u32p, C.c_int, C.c_int]
    lib.orc_antt_forward_full128.restype = C.c_int
    lib.orc_antt_inverse.argtypes = [_u32p, _u32p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
    lib.orc_antt_inverse.restype = C.c_int
    lib.orc_antt_set_plain.argtypes = [C.c_int]
    lib.orc_antt_general_constants.argtypes = [C.c_int, C.c_void_p, _u32p]
    lib.orc_antt_general_constants.restype = C.c_int
    lib.orc_antt_general.argtypes = [_u32p, _u32p, C.c_int, C.c_void_p, C.c_int]
    lib.orc_antt_general.restype = C.c_int
    lib.orc_antt_general_direct.argtypes = [_u32p, _u32p, C.c_int, C.c_void_p]
    lib.orc_antt_general_direct.restype = C.c_int
    lib.orc_md5_hex.argtypes = [C.c_void_p, C.c_size_t, C.c_char_p]
    lib.orc_bb31_mul.argtypes = [C.c_uint32, C.c_uint32]
    lib.orc_bb31_mul.restype = C.c_uint32
    lib.orc_bb31_pow.argtypes = [C.c_uint32, C.c_uint64]
    lib.orc_bb31_pow.restype = C.c_uint32
    lib.orc_bb31_ntt.argtypes = [_u32p, _u32p, C.c_int, C.c_uint32, C.c_int, C.c_int]
    lib.orc_bb31_ntt.restype = C.c_int
    lib.orc_mt19937_fill.argtypes = [C.c_uint32, _u32p, C.c_size_t]
    lib.orc_golden_md5.argtypes = [C.c_int, C.c_int]
    lib.orc_golden_md5.restype = C.c_char_p
    lib.orc_num_threads.restype = C.c_int
    lib.orc_sc_create.argtypes = [_u32p, C.c_int, C.c_int]
    lib.orc_sc_create.restype = C.c_void_p
    lib.orc_sc_destroy.argtypes = [C.c_void_p]
    lib.orc_sc_round_messages.argtypes = [C.c_void_p, _u32p, _u32p]
    lib.orc_sc_next_round.argtypes = [C.c_void_p, _u32p]
    lib.orc_sc_interpolate.argtypes = [_u32p, _u32p, C.c_int, _u32p]
    lib.orc_sc_eval_composition.argtypes = [_u32p, _u32p, C.c_int, C.c_int, _u32p]
    for name in ("orc_tower_mul", "orc_tower_square", "orc_tower_inverse", "orc_tower_mul_alpha"):
        pass  # 128-bit by-value arguments are not expressible in ctypes; use the word forms
    return lib


lib = _load()


def ref_lib():
    """The reference's own host field code, or None when oracle/_ref was not built."""
    path = os.path.join(HERE, "_ref", "libulvt_ref.so")
    if not os.path.exists(path):
        return None
    r = C.CDLL(path)
    for n in ("ref_mul", "ref_mul_packed"):
        getattr(r, n).argtypes = [C.c_uint32, C.c_uint32, C.c_int]
        getattr(r, n).restype = C.c_uint32
    for n in ("ref_square", "ref_inverse", "ref_mul_alpha"):
        getattr(r, n).argtypes = [C.c_uint32, C.c_int]
        getattr(r, n).restype = C.c_uint32
    r.ref_interleave.argtypes = [C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.c_int]
    r.ref_mul_unrolled.argtypes = [_u32p, _u32p, _u32p, C.c_int]
    r.ref_bitslice.argtypes = [_u32p, C.c_int]
    r.ref_unbitslice.argtypes = [_u32p, C.c_int]
    r.ref_mul128_words.argtypes = [_u32p, _u32p, _u32p]
    if hasattr(r, "ref_sc_compose"):
        r.ref_sc_compose.argtypes = [_u32p, _u32p, C.c_uint32, C.c_uint32]
        r.ref_sc_fold_batch.argtypes = [_u32p, _u32p, _u32p, _u32p, C.c_int]
        r.ref_sc_fold_small.argtypes = [_u32p, _u32p, _u32p, C.c_uint32]
        r.ref_sc_compute_sum.argtypes = [_u32p, _u32p, C.c_uint32]
        r.ref_sc_repeat_value.argtypes = [_u32p, _u32p]
    return r


# ---- numpy-level helpers -------------------------------------------------------------------------
def md5_hex(arr: np.ndarray) -> str:
    a = np.ascontiguousarray(arr)
    buf = C.create_string_buffer(33)
    lib.orc_md5_hex(a.ctypes.data, a.nbytes, buf)
    return buf.value.decode()


def mt19937(seed: int, n: int) -> np.ndarray:
    out = np.empty(n, dtype=np.uint32)
    lib.orc_mt19937_fill(seed & 0xFFFFFFFF, out, n)
    return out


def reference_input(log_h: int, log_rate: int) -> np.ndarray:
    """Input stream of ntt/tests/test_ntt.cu:192-199."""
    return mt19937(0xDEADBEEF + log_h + log_rate, 1 << log_h)


def golden_md5(log_rate: int, log_h: int):
    g = lib.orc_golden_md5(log_rate, log_h)
    return g.decode() if g else None


def antt_constants(log_h: int, log_rate: int) -> np.ndarray:
    pitch = log_h + log_rate - 1
    out = np.zeros((log_h, max(pitch, 1)), dtype=np.uint32)
    lib.orc_antt_constants(log_h, log_rate, out.reshape(-1))
    return out[:, :pitch] if pitch else out[:, :0]


def antt_forward(x: np.ndarray, log_h: int, log_rate: int, limbs: int = 1, threads: int = 0) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.uint32).reshape(-1)
    assert x.size == (1 << log_h) * limbs
    out = np.empty((1 << (log_h + log_rate)) * limbs, dtype=np.uint32)
    rc = lib.orc_antt_forward(x, out, log_h, log_rate, limbs, threads)
    if rc:
        raise ValueError("orc_antt_forward: bad arguments")
    return out


def antt_forward_full128(x, log_h, log_rate):
    x = np.ascontiguousarray(x, dtype=np.uint32).reshape(-1)
    out = np.empty((1 << (log_h + log_rate)) * 4, dtype=np.uint32)
    rc = lib.orc_antt_forward_full128(x, out, log_h, log_rate)
    if rc:
        raise ValueError("bad arguments")
    return out


def _basis_ptr(basis, log_h):
    if basis is None:
        return None, None
    b = np.ascontiguousarray(basis, dtype=np.uint32).reshape(-1)
    assert b.size == 4 * log_h
    return b, b.ctypes.data


def antt_general_constants(log_h, basis=None):
    """[log_h, log_h, 4] normalised constants W^_i(beta_{i+1+j}) (antt_general.c)."""
    keep, ptr = _basis_ptr(basis, log_h)
    out = np.zeros((log_h, log_h, 4), dtype=np.uint32)
    if lib.orc_antt_general_constants(log_h, ptr, out.reshape(-1)):
        raise ValueError("linearly dependent basis")
    return out


def antt_general(x, log_h, basis=None, inverse=False, direct=False):
    """GF(2^128) additive NTT on the basis `basis` (log_h x 4 words; None = 1 << i).  direct=True evaluates the
    definition instead of the butterfly network (log_h <= 8)."""
    x = np.ascontiguousarray(x, dtype=np.uint32).reshape(-1)
    assert x.size == 4 << log_h
    keep, ptr = _basis_ptr(basis, log_h)
    out = np.empty_like(x)
    rc = lib.orc_antt_general_direct(x, out, log_h, ptr) if direct else lib.orc_antt_general(x, out, log_h, ptr, int(inverse))
    if rc:
        raise ValueError("bad arguments or linearly dependent basis")
    return out


def antt_inverse(y: np.ndarray, log_h: int, log_rate: int = 0, coset: int = 0, limbs: int = 1, threads: int = 0):
    y = np.ascontiguousarray(y, dtype=np.uint32).reshape(-1)
    assert y.size == (1 << log_h) * limbs
    out = np.empty_like(y)
    rc = lib.orc_antt_inverse(y, out, log_h, log_rate, coset, limbs, threads)
    if rc:
        raise ValueError("bad arguments")
    return out


BB31_P = 15 * (1 << 27) + 1


def bb31_ntt(x, log_n, generator=137, log_group_order=27, input_bit_reversed=False):
    """BabyBear radix-2 NTT of ntt/gpuntt.cuh:126-209 on canonical residues (the reference stores Montgomery words)."""
    x = np.ascontiguousarray(x, dtype=np.uint32)
    out = np.empty(1 << log_n, dtype=np.uint32)
    rc = lib.orc_bb31_ntt(x, out, log_n, generator, log_group_order, 1 if input_bit_reversed else 0)
    if rc:
        raise ValueError(f"orc_bb31_ntt: {rc}")
    return out


def mul32(a, b):
    """Element-wise GF(2^32) tower product of two uint32 arrays (definitional, slow)."""
    a = np.asarray(a, dtype=np.uint32).reshape(-1)
    b = np.asarray(b, dtype=np.uint32).reshape(-1)
    return np.array([lib.orc_mul32_fast(int(x), int(y)) for x, y in zip(a, b)], dtype=np.uint32)


def mul128(a, b):
    """Element-wise GF(2^128) product, arrays of shape (n, 4) uint32."""
    a = np.ascontiguousarray(a, dtype=np.uint32).reshape(-1, 4)
    b = np.ascontiguousarray(b, dtype=np.uint32).reshape(-1, 4)
    out = np.empty_like(a)
    for i in range(a.shape[0]):
        lib.orc_mul128_words(a[i], b[i], out[i])
    return out


def bitslice(blocks: np.ndarray, width_bits: int) -> np.ndarray:
    """compact -> bit-sliced, blocks of 32 elements (width_bits words each), returns a copy."""
    x = np.array(blocks, dtype=np.uint32).reshape(-1, width_bits)
    for i in range(x.shape[0]):
        lib.orc_bitslice(x[i], width_bits)
    return x.reshape(-1)


def unbitslice(blocks: np.ndarray, width_bits: int) -> np.ndarray:
    x = np.array(blocks, dtype=np.uint32).reshape(-1, width_bits)
    for i in range(x.shape[0]):
        lib.orc_unbitslice(x[i], width_bits)
    return x.reshape(-1)


def mul_bitsliced(a, b, h):
    a = np.ascontiguousarray(a, dtype=np.uint32).reshape(-1, 1 << h)
    b = np.ascontiguousarray(b, dtype=np.uint32).reshape(-1, 1 << h)
    out = np.empty_like(a)
    for i in range(a.shape[0]):
        lib.orc_mul_bitsliced(a[i], b[i], out[i], h)
    return out.reshape(-1)


class Sumcheck:
    """Compact restatement of Sumcheck<NUM_VARS, COMPOSITION_SIZE, *> (sumcheck/sumcheck.cuh:10-300); evals: COMPACT
    GF(2^128) words, composition_size columns of 2^num_vars elements."""

    def __init__(self, evals, num_vars, composition_size):
        evals = np.ascontiguousarray(evals, dtype=np.uint32).reshape(-1)
        assert evals.size == 4 * composition_size << num_vars
        self.num_vars, self.comp = num_vars, composition_size
        self._h = lib.orc_sc_create(evals, num_vars, composition_size)

    def this_round_messages(self):
        s = np.zeros(4, dtype=np.uint32)
        p = np.zeros(4 * (self.comp + 1), dtype=np.uint32)
        lib.orc_sc_round_messages(self._h, s, p)
        return s, p.reshape(-1, 4)

    def move_to_next_round(self, challenge):
        lib.orc_sc_next_round(self._h, np.ascontiguousarray(challenge, dtype=np.uint32))

    def __del__(self):
        if getattr(self, "_h", None):
            lib.orc_sc_destroy(self._h)
            self._h = None


def sc_interpolate(challenge, points):
    """evaluate_univariate_given_points (sumcheck/test/verifier.cu:9-31)"""
    points = np.ascontiguousarray(points, dtype=np.uint32).reshape(-1, 4)
    out = np.zeros(4, dtype=np.uint32)
    lib.orc_sc_interpolate(np.ascontiguousarray(challenge, dtype=np.uint32), points.reshape(-1), points.shape[0], out)
    return out


def sc_eval_composition(evals, challenges, num_vars, composition_size):
    """evaluate_multilinear_composition (sumcheck/test/verifier.cu:92-107), brute force"""
    out = np.zeros(4, dtype=np.uint32)
    lib.orc_sc_eval_composition(np.ascontiguousarray(evals, dtype=np.uint32).reshape(-1), np.ascontiguousarray(challenges, dtype=np.uint32).reshape(-1), num_vars, composition_size, out)
    return out
