"""ctypes binding of the CPU oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY: imported by tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs; never by the product package."""
import ctypes as C
import os
import subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_SO = os.path.join(ROOT, "oracle", "liboracle.so")

R_MOD = 21888242871839275222246405745257275088548364400416034343698204186575808495617
Q_MOD = 21888242871839275222246405745257275088696311157297823662689037894645226208583
MONT_R = 1 << 256

u64p = C.POINTER(C.c_uint64)
u32p = C.POINTER(C.c_uint32)
u8p = C.POINTER(C.c_uint8)


def build():
    src = [os.path.join(ROOT, "oracle", f) for f in ("oracle.cpp", "bn254.hpp")]
    if not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in src):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "liboracle.so"])


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.orc_synth_field.restype = C.c_void_p
        _lib.orc_synth_binary.restype = C.c_void_p
        _lib.orc_circuit_new.restype = C.c_void_p
        _lib.orc_genkey.restype = C.c_void_p
        _lib.orc_key_serialize.restype = C.c_uint64
        _lib.orc_key_vk_json.restype = C.c_uint64
        _lib.orc_proof_json.restype = C.c_uint64
        _lib.orc_init()
    return _lib


def p64(a):
    assert a.dtype == np.uint64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(u64p)


def p32(a):
    assert a.dtype == np.uint32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(u32p)


def p8(a):
    assert a.dtype == np.uint8 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(u8p)


# ---------------------------------------------------------------- int <-> limb helpers
def int_to_limbs(x, n=4):
    return np.array([(x >> (64 * i)) & 0xFFFFFFFFFFFFFFFF for i in range(n)], dtype=np.uint64)


def limbs_to_int(a):
    a = np.asarray(a, dtype=np.uint64).reshape(-1)
    return sum(int(v) << (64 * i) for i, v in enumerate(a))


def ints_to_arr(xs, n=4):
    out = np.zeros((len(xs), n), dtype=np.uint64)
    for i, x in enumerate(xs):
        for j in range(n):
            out[i, j] = (x >> (64 * j)) & 0xFFFFFFFFFFFFFFFF
    return out


def arr_to_ints(a):
    a = np.asarray(a, dtype=np.uint64).reshape(-1, 4)
    return [limbs_to_int(r) for r in a]


def to_mont(x, p):
    return (x * MONT_R) % p


def from_mont(x, p):
    return (x * pow(MONT_R, -1, p)) % p


def rand_fr_mont(n, seed):
    """n uniform-ish Fr elements (< 2^253, Montgomery form is just another representation of some element)."""
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 1 << 63, size=(n, 4), dtype=np.uint64) * np.uint64(2) + rng.integers(0, 2, size=(n, 4), dtype=np.uint64)
    a[:, 3] &= np.uint64((1 << 60) - 1)
    return np.ascontiguousarray(a)


# ---------------------------------------------------------------- wrappers
def _b2(name, a, b):
    a = np.ascontiguousarray(a, dtype=np.uint64); b = np.ascontiguousarray(b, dtype=np.uint64)
    out = np.empty_like(a)
    getattr(lib(), name)(p64(a), p64(b), p64(out), C.c_size_t(a.size // 4))
    return out


def _b1(name, a):
    a = np.ascontiguousarray(a, dtype=np.uint64)
    out = np.empty_like(a)
    getattr(lib(), name)(p64(a), p64(out), C.c_size_t(a.size // 4))
    return out


def fr_mul(a, b): return _b2("orc_fr_mul", a, b)
def fr_add(a, b): return _b2("orc_fr_add", a, b)
def fr_sub(a, b): return _b2("orc_fr_sub", a, b)
def fq_mul(a, b): return _b2("orc_fq_mul", a, b)
def fq_add(a, b): return _b2("orc_fq_add", a, b)
def fq_sub(a, b): return _b2("orc_fq_sub", a, b)
def fr_from_mont(a): return _b1("orc_fr_from_mont", a)
def fr_to_mont(a): return _b1("orc_fr_to_mont", a)
def fq_from_mont(a): return _b1("orc_fq_from_mont", a)
def fq_to_mont(a): return _b1("orc_fq_to_mont", a)
def fr_inv(a): return _b1("orc_fr_inv", a)
def fq_inv(a): return _b1("orc_fq_inv", a)


def sha512_rng_fr(idx0, n):
    out = np.empty((n, 4), dtype=np.uint64)
    lib().orc_sha512_rng_fr(C.c_uint64(idx0), C.c_size_t(n), p64(out))
    return out


def constants():
    out = np.zeros(32, dtype=np.uint64)
    lib().orc_constants(p64(out))
    return out


def generators():
    g1 = np.zeros(8, dtype=np.uint64); g2 = np.zeros(16, dtype=np.uint64)
    lib().orc_generators(p64(g1), p64(g2))
    return g1, g2


def g1_fixed_base(scalars_mont):
    s = np.ascontiguousarray(scalars_mont, dtype=np.uint64); n = s.size // 4
    xy = np.zeros((n, 8), dtype=np.uint64); inf = np.zeros(n, dtype=np.uint8)
    lib().orc_g1_fixed_base(p64(s), C.c_size_t(n), p64(xy), p8(inf))
    return xy, inf


def g2_fixed_base(scalars_mont):
    s = np.ascontiguousarray(scalars_mont, dtype=np.uint64); n = s.size // 4
    xy = np.zeros((n, 16), dtype=np.uint64); inf = np.zeros(n, dtype=np.uint8)
    lib().orc_g2_fixed_base(p64(s), C.c_size_t(n), p64(xy), p8(inf))
    return xy, inf


def _add(name, w, a, ai, b, bi):
    out = np.zeros(w, dtype=np.uint64); oi = C.c_uint8(0)
    a = np.ascontiguousarray(a, dtype=np.uint64); b = np.ascontiguousarray(b, dtype=np.uint64)
    getattr(lib(), name)(p64(a), C.c_uint8(int(ai)), p64(b), C.c_uint8(int(bi)), p64(out), C.byref(oi))
    return out, oi.value


def g1_add(a, ai, b, bi): return _add("orc_g1_add", 8, a, ai, b, bi)
def g2_add(a, ai, b, bi): return _add("orc_g2_add", 16, a, ai, b, bi)


def _msm(name, w, bases, inf, scalars, c=0, prefilter=1, naive=False):
    bases = np.ascontiguousarray(bases, dtype=np.uint64); scalars = np.ascontiguousarray(scalars, dtype=np.uint64)
    n = scalars.size // 4
    assert bases.size == n * w
    inf = np.zeros(n, dtype=np.uint8) if inf is None else np.ascontiguousarray(inf, dtype=np.uint8)
    out = np.zeros(w, dtype=np.uint64); oi = C.c_uint8(0)
    if naive:
        getattr(lib(), name + "_naive")(p64(bases), p8(inf), p64(scalars), C.c_size_t(n), p64(out), C.byref(oi))
    else:
        getattr(lib(), name)(p64(bases), p8(inf), p64(scalars), C.c_size_t(n), C.c_uint(c), C.c_int(prefilter), p64(out), C.byref(oi))
    return out, oi.value


def msm_g1(bases, inf, scalars, **kw): return _msm("orc_msm_g1", 8, bases, inf, scalars, **kw)
def msm_g2(bases, inf, scalars, **kw): return _msm("orc_msm_g2", 16, bases, inf, scalars, **kw)


def fr_dot(a, b):
    a = np.ascontiguousarray(a, dtype=np.uint64); b = np.ascontiguousarray(b, dtype=np.uint64)
    out = np.zeros(4, dtype=np.uint64)
    lib().orc_fr_dot(p64(a), p64(b), C.c_size_t(a.size // 4), p64(out))
    return out


def ntt(data_mont, logm, mode):
    """mode: 0 FFT, 1 iFFT, 2 cosetFFT, 3 icosetFFT.  Returns a new array."""
    a = np.array(data_mont, dtype=np.uint64, copy=True)
    lib().orc_ntt(p64(a), C.c_uint(logm), C.c_int(mode))
    return a


def dft_naive(data_mont, logm):
    a = np.ascontiguousarray(data_mont, dtype=np.uint64); out = np.empty_like(a)
    lib().orc_dft_naive(p64(a), p64(out), C.c_uint(logm))
    return out


class Circuit:
    def __init__(self, handle):
        self.h = C.c_void_p(handle)
        d = np.zeros(8, dtype=np.uint64)
        lib().orc_circuit_dims(self.h, p64(d))
        self.nc, self.ni, self.nv, self.m = (int(x) for x in d[:4])
        self.nnz = [int(x) for x in d[4:7]]
        self.n_coeffs = int(d[7])
        self.logm = self.m.bit_length() - 1

    @staticmethod
    def synth_field(nc, ni, seed_a=0, seed_b=1):
        return Circuit(lib().orc_synth_field(C.c_uint64(nc), C.c_uint64(ni), C.c_uint64(seed_a), C.c_uint64(seed_b)))

    @staticmethod
    def from_csr(nc, ni, nv, mats, coeffs):
        """mats: three (row_ptr u64, col u32, coeff_idx u32); coeffs: (n, 4) Montgomery"""
        args = []
        keep = []
        for rp, col, ci in mats:
            rp = np.ascontiguousarray(rp, dtype=np.uint64); col = np.ascontiguousarray(col, dtype=np.uint32)
            ci = np.ascontiguousarray(ci, dtype=np.uint32)
            keep += [rp, col, ci]
            args += [p64(rp), p32(col), p32(ci)]
        coeffs = np.ascontiguousarray(coeffs, dtype=np.uint64)
        lib().orc_circuit_new.restype = C.c_void_p
        return Circuit(lib().orc_circuit_new(C.c_uint64(nc), C.c_uint64(ni), C.c_uint64(nv), *args, p64(coeffs),
                                              C.c_uint64(coeffs.shape[0])))

    @staticmethod
    def synth_binary(nc, ni, seed=1):
        return Circuit(lib().orc_synth_binary(C.c_uint64(nc), C.c_uint64(ni), C.c_uint64(seed)))

    def matrix(self, which):
        rp = np.zeros(self.nc + 1, dtype=np.uint64)
        col = np.zeros(max(self.nnz[which], 1), dtype=np.uint32); ci = np.zeros(max(self.nnz[which], 1), dtype=np.uint32)
        lib().orc_circuit_matrix(self.h, C.c_int(which), p64(rp), p32(col), p32(ci))
        return rp, col[: self.nnz[which]], ci[: self.nnz[which]]

    def coeffs(self):
        out = np.zeros((self.n_coeffs, 4), dtype=np.uint64)
        lib().orc_circuit_coeffs(self.h, p64(out))
        return out

    def witness(self):
        out = np.zeros((self.nv + 1, 4), dtype=np.uint64)
        assert lib().orc_circuit_witness(self.h, p64(out)) == 0
        return out

    def is_satisfied(self, w):
        return bool(lib().orc_is_satisfied(self.h, p64(np.ascontiguousarray(w, dtype=np.uint64))))

    def h_poly(self, w, evals=False):
        w = np.ascontiguousarray(w, dtype=np.uint64)
        out = np.zeros((self.m + 1, 4), dtype=np.uint64)
        if evals:
            ev = [np.zeros((self.m + 1, 4), dtype=np.uint64) for _ in range(3)]
            lib().orc_h(self.h, p64(w), p64(out), p64(ev[0]), p64(ev[1]), p64(ev[2]))
            return out, ev
        lib().orc_h(self.h, p64(w), p64(out), None, None, None)
        return out

    def __del__(self):
        try:
            lib().orc_circuit_free(self.h)
        except Exception:
            pass


FMT_MCL_NOZK, FMT_MCL_FULL, FMT_ALT_NOZK, FMT_ALT_FULL = 0, 1, 2, 3


class Key:
    def __init__(self, circuit, seed_base=1 << 32):
        self.circuit = circuit
        self.h = C.c_void_p(lib().orc_genkey(circuit.h, C.c_uint64(seed_base)))
        d = np.zeros(5, dtype=np.uint64)
        lib().orc_key_dims(self.h, p64(d))
        self.domain_size, self.nA, self.nB, self.nH, self.nL = (int(x) for x in d)

    def serialize(self, fmt=FMT_MCL_NOZK):
        n = lib().orc_key_serialize(self.h, C.c_int(fmt), None, C.c_uint64(0))
        buf = np.zeros(n, dtype=np.uint8)
        lib().orc_key_serialize(self.h, C.c_int(fmt), p8(buf), C.c_uint64(n))
        return buf.tobytes()

    def vk_json(self):
        n = lib().orc_key_vk_json(self.h, None, C.c_uint64(0))
        buf = C.create_string_buffer(n + 1)
        lib().orc_key_vk_json(self.h, buf, C.c_uint64(n + 1))
        return buf.value.decode()

    def prove(self, w, r=None, s=None, window=0):
        w = np.ascontiguousarray(w, dtype=np.uint64)
        A = np.zeros(8, dtype=np.uint64); B = np.zeros(16, dtype=np.uint64); Cc = np.zeros(8, dtype=np.uint64)
        inf = np.zeros(3, dtype=np.uint8)
        rp = p64(np.ascontiguousarray(r, dtype=np.uint64)) if r is not None else None
        sp = p64(np.ascontiguousarray(s, dtype=np.uint64)) if s is not None else None
        lib().orc_prove(self.h, self.circuit.h, p64(w), rp, sp, C.c_uint(window), p64(A), p64(B), p64(Cc), p8(inf))
        return A, B, Cc, inf

    def check_in_exponent(self, w, proof, r=None, s=None):
        A, B, Cc, inf = proof
        w = np.ascontiguousarray(w, dtype=np.uint64)
        rp = p64(np.ascontiguousarray(r, dtype=np.uint64)) if r is not None else None
        sp = p64(np.ascontiguousarray(s, dtype=np.uint64)) if s is not None else None
        return lib().orc_check_in_exponent(self.h, self.circuit.h, p64(w), rp, sp, p64(np.ascontiguousarray(A)),
                                           p64(np.ascontiguousarray(B)), p64(np.ascontiguousarray(Cc)),
                                           p8(np.ascontiguousarray(inf, dtype=np.uint8)))

    def __del__(self):
        try:
            lib().orc_key_free(self.h)
        except Exception:
            pass


def proof_json(proof, inputs_mont):
    A, B, Cc, inf = proof
    inputs = np.ascontiguousarray(inputs_mont, dtype=np.uint64)
    n_in = inputs.size // 4
    args = (p64(np.ascontiguousarray(A)), p64(np.ascontiguousarray(B)), p64(np.ascontiguousarray(Cc)),
            p8(np.ascontiguousarray(inf, dtype=np.uint8)), p64(inputs), C.c_uint64(n_in))
    n = lib().orc_proof_json(*args, None, C.c_uint64(0))
    buf = C.create_string_buffer(n + 1)
    lib().orc_proof_json(*args, buf, C.c_uint64(n + 1))
    return buf.value.decode()
