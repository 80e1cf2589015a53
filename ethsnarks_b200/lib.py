"""ctypes binding of libb200snark.so (the C ABI in include/b200snark.h).

This is the same kind of binding the reference already uses for its one native entry point
(`ethsnarks/verifier.py:199-212` loads `libethsnarks_verify` through ctypes).  There is no fallback: if the
shared library is missing, or no CUDA device is present, construction fails loudly.
"""
import ctypes as C
import os
import weakref
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get("B200SNARK_LIB") or os.path.join(_HERE, "libb200snark.so")   # the override is for kernel experiments

u64p = C.POINTER(C.c_uint64)
u32p = C.POINTER(C.c_uint32)
u8p = C.POINTER(C.c_uint8)

FMT_MCL_NOZK, FMT_MCL_FULL, FMT_ALT_NOZK, FMT_ALT_FULL, FMT_CACHE = 0, 1, 2, 3, 4
OP_MUL, OP_ADD, OP_SUB, OP_TO_MONT, OP_FROM_MONT, OP_INV, OP_SQR, OP_NEG, OP_INV_BINARY, OP_DOT2, OP_DOT4, OP_DIFF = range(12)
NTT_FFT, NTT_IFFT, NTT_COSET_FFT, NTT_ICOSET_FFT = range(4)

# every symbol include/b200snark.h declares (tests check the .so exports all of them)
SYMBOLS = [
    "b200_init", "b200_shutdown", "b200_strerror", "b200_last_error", "b200_stream", "b200_launch_count", "b200_set_option",
    "b200_timing_count", "b200_timing_get", "b200_timing_start", "b200_field_op", "b200_g1_add", "b200_g2_add", "b200_g1_fixed_base",
    "b200_g2_fixed_base", "b200_ntt", "b200_msm_g1", "b200_msm_g2", "b200_bases_upload", "b200_bases_generate",
    "b200_bases_download", "b200_bases_count", "b200_bases_free", "b200_msm", "b200_key_load", "b200_key_load_sharded",
    "b200_key_from_bases", "b200_key_free", "b200_key_info", "b200_circuit_load", "b200_prove", "b200_h",
    "b200_witness_eval", "b200_prove_partial", "b200_proof_combine", "b200_imad_peak", "b200_key_set_shard", "b200_bases_precompute", "b200_key_attach", "b200_key_export",
    "b200_generate_key", "b200_g1_sum", "b200_g2_sum", "b200_g1_sum_noalloc", "b200_device_count", "b200_key_convert", "b200_chain_owner", "b200_key_buffers", "b200_prove_partial_stage1", "b200_prove_partial_stage2", "b200_circuit_load_instanced", "b200_prove_partial_dev", "b200_proof_combine_dev",
]


class CSR(C.Structure):
    _fields_ = [("row_ptr", u64p), ("col", u32p), ("coeff_idx", u32p), ("n_rows", C.c_uint64)]


class INSTANCES(C.Structure):
    _fields_ = [("A", C.POINTER(CSR)), ("B", C.POINTER(CSR)), ("C", C.POINTER(CSR)), ("tables", u32p), ("table_len", C.c_uint32),
                ("n_instances", C.c_uint64)]


class B200Error(RuntimeError):
    pass


_lib = None


def key_convert(pk_bytes, fmt_in, fmt_out):
    """b200_key_convert: reference key bytes from one of the four layouts to another (pk_alt2mcl / pk_mcl2nozk, export.cpp:330-408);
    host only - works without a GPU"""
    lib = load()
    n = C.c_size_t()
    rc = lib.b200_key_convert(pk_bytes, C.c_size_t(len(pk_bytes)), fmt_in, fmt_out, None, C.c_size_t(0), C.byref(n))
    if rc:
        raise B200Error("key_convert: " + lib.b200_strerror(rc).decode())
    buf = C.create_string_buffer(n.value)
    rc = lib.b200_key_convert(pk_bytes, C.c_size_t(len(pk_bytes)), fmt_in, fmt_out, buf, n, C.byref(n))
    if rc:
        raise B200Error("key_convert: " + lib.b200_strerror(rc).decode())
    return buf.raw[:n.value]


def chain_owner(chain, world):
    """rank that takes evaluation vector `chain` (0 = A, 1 = B, 2 = C) through its first two transforms in a staged proof"""
    return int(load().b200_chain_owner(chain, world))


def load():
    """dlopen the in-tree shared library; raises if it has not been built (no CPU fallback exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise B200Error(f"{SO_PATH} is missing: build it with `make -C ethsnarks_b200/csrc -j` (or __graft_entry__.build())")
        lib = C.CDLL(SO_PATH)
        lib.b200_strerror.restype = C.c_char_p
        lib.b200_last_error.restype = C.c_char_p
        lib.b200_stream.restype = C.c_void_p
        lib.b200_launch_count.restype = C.c_ulonglong
        lib.b200_bases_count.restype = C.c_size_t
        for name in ("b200_shutdown", "b200_bases_free", "b200_key_free"):
            getattr(lib, name).restype = None
        _lib = lib
    return _lib


def _p64(a):
    return a.ctypes.data_as(u64p)


def _arr(a, dtype=np.uint64):
    return np.ascontiguousarray(a, dtype=dtype)


class Context:
    """One CUDA device, one stream (b200_init) - or, with a list of devices, ONE context spanning several GPUs of the box:
    keys loaded on it are split 1/N per device and prove() runs on all of them from this process."""

    def __init__(self, device=0):
        self.lib = load()
        self.h = C.c_void_p()
        devices = list(device) if isinstance(device, (list, tuple)) else [device]
        dev = (C.c_int * len(devices))(*devices)
        rc = self.lib.b200_init(dev, len(devices), C.byref(self.h))
        if rc:
            raise B200Error(f"b200_init(devices={devices}) failed: {self.lib.b200_strerror(rc).decode()}")
        self.device = devices[0]
        self.devices = devices
        self._children = weakref.WeakSet()    # live Bases / ProvingKey objects: freed before the context goes away

    def check(self, rc, what):
        if rc:
            raise B200Error(f"{what}: {self.lib.b200_strerror(rc).decode()} — {self.lib.b200_last_error(self.h).decode()}")

    def close(self):
        if self.h:
            for child in sorted(self._children, key=lambda c: not getattr(c, "_attached", False)):   # attached slots first
                child.free()
            self.lib.b200_shutdown(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def stream(self):
        return self.lib.b200_stream(self.h)

    @property
    def launches(self):
        return int(self.lib.b200_launch_count(self.h))

    def set_option(self, name, value):
        self.check(self.lib.b200_set_option(self.h, name.encode(), C.c_long(value)), "set_option")

    def timings(self):
        out = []
        n = self.lib.b200_timing_count(self.h)
        buf = C.create_string_buffer(64)
        ms = C.c_float()
        for i in range(n):
            self.lib.b200_timing_get(self.h, i, buf, C.c_size_t(64), C.byref(ms))
            out.append((buf.value.decode(), float(ms.value)))
        return out

    def timeline(self):
        """(name, start_ms, duration_ms) of every phase of the last call; start is relative to its first phase."""
        at = C.c_float()
        out = []
        for i, (name, ms) in enumerate(self.timings()):
            self.lib.b200_timing_start(self.h, i, C.byref(at))
            out.append((name, float(at.value), ms))
        return out

    def imad_peak(self, reps=5):
        v = C.c_double()
        self.check(self.lib.b200_imad_peak(self.h, reps, C.byref(v)), "imad_peak")
        return float(v.value)

    # ---- sub-operations
    def field_op(self, field, op, a, b=None):
        a = _arr(a)
        out = np.empty_like(a)
        bb = _arr(b) if b is not None else None
        self.check(self.lib.b200_field_op(self.h, field, op, _p64(a), _p64(bb) if bb is not None else None, _p64(out),
                                          C.c_size_t(a.size // 4)), "field_op")
        return out

    def _gadd(self, fn, w, a, ai, b, bi):
        a, b = _arr(a), _arr(b)
        n = a.size // w
        ai = _arr(ai, np.uint8) if ai is not None else np.zeros(n, np.uint8)
        bi = _arr(bi, np.uint8) if bi is not None else np.zeros(n, np.uint8)
        out = np.zeros((n, w), np.uint64)
        oi = np.zeros(n, np.uint8)
        self.check(fn(self.h, _p64(a), ai.ctypes.data_as(u8p), _p64(b), bi.ctypes.data_as(u8p), _p64(out), oi.ctypes.data_as(u8p),
                      C.c_size_t(n)), "group add")
        return out, oi

    def g1_add(self, a, ai, b, bi):
        return self._gadd(self.lib.b200_g1_add, 8, a, ai, b, bi)

    def g2_add(self, a, ai, b, bi):
        return self._gadd(self.lib.b200_g2_add, 16, a, ai, b, bi)

    def _gsum(self, fn, w, xy, inf):
        xy = _arr(xy)
        n = xy.size // w
        inf = _arr(inf, np.uint8) if inf is not None else np.zeros(n, np.uint8)
        out = np.zeros(w, np.uint64)
        oi = C.c_uint8(0)
        self.check(fn(self.h, _p64(xy), inf.ctypes.data_as(u8p), C.c_size_t(n), _p64(out), C.byref(oi)), "group sum")
        return out, oi.value

    def g1_sum(self, xy, inf=None):
        """sum of n affine G1 points (n x 8 Montgomery limbs) -> (8 limbs, infinity flag)"""
        return self._gsum(self.lib.b200_g1_sum, 8, xy, inf)

    def g2_sum(self, xy, inf=None):
        return self._gsum(self.lib.b200_g2_sum, 16, xy, inf)

    def g1_sum_noalloc(self, d_xy_ptr, n, d_out_ptr):
        """b200_g1_sum_noalloc: DEVICE pointers in and out (n x 8 limbs in; 8 limbs + u32 flag out), queued on the context's
        stream without synchronising; allocates nothing (static device scratch)."""
        self.check(self.lib.b200_g1_sum_noalloc(self.h, C.c_void_p(d_xy_ptr), C.c_size_t(n), C.c_void_p(d_out_ptr)), "g1_sum_noalloc")

    def _fixed(self, fn, w, scalars):
        s = _arr(scalars)
        n = s.size // 4
        out = np.zeros((n, w), np.uint64)
        oi = np.zeros(n, np.uint8)
        self.check(fn(self.h, _p64(s), C.c_size_t(n), _p64(out), oi.ctypes.data_as(u8p)), "fixed_base")
        return out, oi

    def g1_fixed_base(self, scalars):
        return self._fixed(self.lib.b200_g1_fixed_base, 8, scalars)

    def g2_fixed_base(self, scalars):
        return self._fixed(self.lib.b200_g2_fixed_base, 16, scalars)

    def ntt(self, data, logm, mode):
        a = np.array(data, dtype=np.uint64, copy=True)
        self.check(self.lib.b200_ntt(self.h, _p64(a), C.c_uint32(logm), mode), "ntt")
        return a

    def _msm(self, fn, w, bases, inf, scalars):
        bases, scalars = _arr(bases), _arr(scalars)
        n = scalars.size // 4
        infp = _arr(inf, np.uint8).ctypes.data_as(u8p) if inf is not None else None
        out = np.zeros(w, np.uint64)
        oi = C.c_uint8(0)
        self.check(fn(self.h, _p64(bases), infp, _p64(scalars), C.c_size_t(n), _p64(out), C.byref(oi)), "msm")
        return out, oi.value

    def msm_g1(self, bases, inf, scalars):
        return self._msm(self.lib.b200_msm_g1, 8, bases, inf, scalars)

    def msm_g2(self, bases, inf, scalars):
        return self._msm(self.lib.b200_msm_g2, 16, bases, inf, scalars)


class Bases:
    """A device-resident set of affine bases (G1 or G2)."""

    def __init__(self, ctx, group, xy=None, inf=None, scalars=None):
        self.ctx, self.group = ctx, group
        self.h = C.c_void_p()
        if scalars is not None:
            s = _arr(scalars)
            ctx.check(ctx.lib.b200_bases_generate(ctx.h, group, _p64(s), C.c_size_t(s.size // 4), C.byref(self.h)), "bases_generate")
        else:
            xy = _arr(xy)
            n = xy.size // (8 * group)
            infp = _arr(inf, np.uint8).ctypes.data_as(u8p) if inf is not None else None
            ctx.check(ctx.lib.b200_bases_upload(ctx.h, group, _p64(xy), infp, C.c_size_t(n), C.byref(self.h)), "bases_upload")
        self.n = int(ctx.lib.b200_bases_count(self.h))
        ctx._children.add(self)

    def precompute(self, window=0):
        self.ctx.check(self.ctx.lib.b200_bases_precompute(self.h, window), "bases_precompute")

    def download(self):
        xy = np.zeros((self.n, 8 * self.group), np.uint64)
        inf = np.zeros(self.n, np.uint8)
        self.ctx.check(self.ctx.lib.b200_bases_download(self.h, _p64(xy), inf.ctypes.data_as(u8p)), "bases_download")
        return xy, inf

    def msm(self, scalars=None, n=None, device_ptr=None):
        """scalars: host array (Montgomery) — copied inside the call; or device_ptr: int address of resident scalars."""
        out = np.zeros(8 * self.group, np.uint64)
        oi = C.c_uint8(0)
        if device_ptr is not None:
            sp = C.cast(C.c_void_p(device_ptr), u64p)
            on_dev = 1
        else:
            s = _arr(scalars)
            n = s.size // 4 if n is None else n
            sp = _p64(s)
            on_dev = 0
        self.ctx.check(self.ctx.lib.b200_msm(self.ctx.h, self.h, sp, C.c_size_t(n), on_dev, _p64(out), C.byref(oi)), "msm")
        return out, oi.value

    def free(self):
        if self.h and self.ctx.h:
            self.ctx.lib.b200_bases_free(self.h)
        self.h = C.c_void_p()

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def _csr_structs(mats):
    keep, structs = [], []
    for rp, col, ci in mats:
        rp, col, ci = _arr(rp), _arr(col, np.uint32), _arr(ci, np.uint32)
        if col.size == 0:
            col = np.zeros(1, np.uint32); ci = np.zeros(1, np.uint32)
        keep += [rp, col, ci]
        structs.append(CSR(_p64(rp), col.ctypes.data_as(u32p), ci.ctypes.data_as(u32p), C.c_uint64(rp.size - 1)))
    return keep, structs


class ProvingKey:
    """Device-resident proving key (+ circuit).  Mirrors load_proving_key / ProverContext of the reference."""

    def __init__(self, ctx, handle, keep=()):
        self.ctx, self.h, self._keep = ctx, handle, keep
        self._circuit_keep = None
        self._attached = False
        self._slots = weakref.WeakSet()       # prove slots attached to this key: freed before it
        ctx._children.add(self)

    @staticmethod
    def load(ctx, pk_bytes, fmt=FMT_MCL_NOZK, rank=0, world=1):
        h = C.c_void_p()
        if isinstance(pk_bytes, np.ndarray):
            buf, n = pk_bytes.ctypes.data_as(C.c_void_p), pk_bytes.nbytes
        else:
            buf, n = C.cast(C.c_char_p(bytes(pk_bytes) if not isinstance(pk_bytes, bytes) else pk_bytes), C.c_void_p), len(pk_bytes)
        ctx.check(ctx.lib.b200_key_load_sharded(ctx.h, buf, C.c_size_t(n), fmt, rank, world, C.byref(h)), "key_load")
        return ProvingKey(ctx, h)

    @staticmethod
    def from_bases(ctx, A, B_g2, H, L, g1_consts, g2_consts, num_variables, num_inputs, B_g1=None, A_idx=None, B_idx=None):
        h = C.c_void_p()
        g1c, g2c = _arr(g1_consts), _arr(g2_consts)
        ai = _arr(A_idx, np.uint32) if A_idx is not None else None
        bi = _arr(B_idx, np.uint32) if B_idx is not None else None
        ctx.check(ctx.lib.b200_key_from_bases(
            ctx.h, A.h, ai.ctypes.data_as(u32p) if ai is not None else None, B_g2.h, B_g1.h if B_g1 is not None else None,
            bi.ctypes.data_as(u32p) if bi is not None else None, H.h, L.h, _p64(g1c), _p64(g2c), C.c_uint64(num_variables),
            C.c_uint64(num_inputs), C.byref(h)), "key_from_bases")
        return ProvingKey(ctx, h, keep=(A, B_g2, B_g1, H, L))

    @staticmethod
    def generate(ctx, A, B, Cm, coeffs_mont, num_inputs, num_variables, toxic_mont, full=False):
        """Key generation on the device from caller-supplied toxic waste (t, alpha, beta, gamma, delta as 5 x 4 Montgomery
        limbs).  Returns (key, vk) with vk = dict of affine Montgomery limb arrays: alpha (8), beta/gamma/delta (16),
        gammaABC ((num_inputs + 1) x 8).  The key still needs load_circuit()."""
        keep, structs = _csr_structs((A, B, Cm))
        co, tox = _arr(coeffs_mont), _arr(toxic_mont).reshape(-1)
        if tox.size != 20:
            raise ValueError("toxic_mont must hold 5 field elements")
        vk = np.zeros(56 + 8 * (num_inputs + 1), np.uint64)
        h = C.c_void_p()
        ctx.check(ctx.lib.b200_generate_key(ctx.h, C.byref(structs[0]), C.byref(structs[1]), C.byref(structs[2]), _p64(co),
                                            C.c_size_t(co.size // 4), C.c_uint32(num_inputs), C.c_uint64(num_variables), _p64(tox),
                                            C.c_int(1 if full else 0), C.byref(h), _p64(vk)), "generate_key")
        del keep
        parts = dict(alpha=vk[0:8].copy(), beta=vk[8:24].copy(), gamma=vk[24:40].copy(), delta=vk[40:56].copy(),
                     gammaABC=vk[56:].reshape(-1, 8).copy())
        return ProvingKey(ctx, h), parts

    def export_cache(self):
        """The key as a device-native cache blob (FMT_CACHE); reload it with ProvingKey.load(ctx, blob, FMT_CACHE)."""
        n = C.c_size_t(0)
        self.ctx.check(self.ctx.lib.b200_key_export(self.h, None, C.c_size_t(0), C.byref(n)), "key_export")
        buf = np.zeros(n.value, np.uint8)
        self.ctx.check(self.ctx.lib.b200_key_export(self.h, buf.ctypes.data_as(C.c_void_p), n, C.byref(n)), "key_export")
        return buf

    def attach(self, ctx):
        """A second prove slot on this resident key, bound to another Context on the same device (pipelined proving)."""
        h = C.c_void_p()
        ctx.check(ctx.lib.b200_key_attach(self.h, ctx.h, C.byref(h)), "key_attach")
        slot = ProvingKey(ctx, h, keep=(self,))
        slot.m = self.m
        slot._attached = True
        self._slots.add(slot)
        return slot

    def set_shard(self, rank, world, L_lo):
        self.ctx.check(self.ctx.lib.b200_key_set_shard(self.h, rank, world, C.c_uint64(L_lo)), "key_set_shard")

    def info(self):
        out = np.zeros(8, np.uint64)
        self.ctx.check(self.ctx.lib.b200_key_info(self.h, _p64(out)), "key_info")
        names = ("domain_size", "nA", "nB", "nH", "nL", "full", "m", "num_inputs")
        return dict(zip(names, (int(x) for x in out)))

    def load_circuit(self, A, B, Cm, coeffs_mont, num_inputs, num_variables):
        """A, B, Cm: (row_ptr uint64, col uint32, coeff_idx uint32) triples."""
        keep, structs = _csr_structs((A, B, Cm))
        co = _arr(coeffs_mont)
        self.ctx.check(self.ctx.lib.b200_circuit_load(self.h, C.byref(structs[0]), C.byref(structs[1]), C.byref(structs[2]), _p64(co),
                                                      C.c_size_t(co.size // 4), C.c_uint32(num_inputs), C.c_uint64(num_variables)),
                       "circuit_load")
        self.num_inputs, self.num_variables = num_inputs, num_variables
        self.m = self.info()["m"]

    def load_circuit_instanced(self, A, B, Cm, master, tables, coeffs_mont, num_inputs, num_variables):
        """b200_circuit_load_instanced: plain rows A, B, Cm, then for every row j of `tables` (n_instances x table_len uint32) all
        rows of `master` = (mA, mB, mC) translated through it.  The device keeps the master rows and the tables."""
        keep, structs = _csr_structs((A, B, Cm))
        mkeep, mstructs = _csr_structs(master)
        tab = np.ascontiguousarray(tables, dtype=np.uint32).reshape(len(tables), -1)
        inst = INSTANCES(C.pointer(mstructs[0]), C.pointer(mstructs[1]), C.pointer(mstructs[2]), tab.ctypes.data_as(u32p),
                         C.c_uint32(tab.shape[1]), C.c_uint64(tab.shape[0]))
        co = _arr(coeffs_mont)
        self.ctx.check(self.ctx.lib.b200_circuit_load_instanced(self.h, C.byref(structs[0]), C.byref(structs[1]), C.byref(structs[2]),
                                                                C.byref(inst), _p64(co), C.c_size_t(co.size // 4), C.c_uint32(num_inputs),
                                                                C.c_uint64(num_variables)), "circuit_load_instanced")
        self.num_inputs, self.num_variables = num_inputs, num_variables
        self.m = self.info()["m"]

    def _wptr(self, witness):
        if isinstance(witness, int):
            return C.cast(C.c_void_p(witness), u64p), None
        w = _arr(witness)
        return _p64(w), w

    def prove(self, witness, r=None, s=None):
        """witness: numpy (nv+1, 4) uint64 Montgomery, or an int host address (e.g. a pinned torch tensor's data_ptr)."""
        wp, keep = self._wptr(witness)
        A = np.zeros(8, np.uint64); B = np.zeros(16, np.uint64); Cc = np.zeros(8, np.uint64); inf = np.zeros(3, np.uint8)
        rr = _arr(r) if r is not None else None
        ss = _arr(s) if s is not None else None
        self.ctx.check(self.ctx.lib.b200_prove(self.h, wp, _p64(rr) if rr is not None else None, _p64(ss) if ss is not None else None,
                                               _p64(A), _p64(B), _p64(Cc), inf.ctypes.data_as(u8p)), "prove")
        return A, B, Cc, inf

    def h_poly(self, witness):
        wp, keep = self._wptr(witness)
        out = np.zeros((self.m, 4), np.uint64)
        self.ctx.check(self.ctx.lib.b200_h(self.h, wp, _p64(out)), "h")
        return out

    def witness_eval(self, witness):
        wp, keep = self._wptr(witness)
        outs = [np.zeros((self.m, 4), np.uint64) for _ in range(3)]
        self.ctx.check(self.ctx.lib.b200_witness_eval(self.h, wp, _p64(outs[0]), _p64(outs[1]), _p64(outs[2])), "witness_eval")
        return outs

    def prove_partial(self, witness):
        wp, keep = self._wptr(witness)
        part = np.zeros(32, np.uint64); flags = np.zeros(3, np.uint8)
        self.ctx.check(self.ctx.lib.b200_prove_partial(self.h, wp, _p64(part), flags.ctypes.data_as(u8p)), "prove_partial")
        return part, flags

    def prove_partial_dev(self, witness, d_out_ptr):
        """b200_prove_partial_dev: this rank's 256-byte packed partial sums written to DEVICE memory at d_out_ptr, queued on the
        context's stream without synchronising"""
        wp, keep = self._wptr(witness)
        self.ctx.check(self.ctx.lib.b200_prove_partial_dev(self.h, wp, C.c_void_p(d_out_ptr)), "prove_partial_dev")

    def buffers(self):
        """device pointers of the three evaluation vectors (A, B, C) and their size in bytes"""
        bufs = (C.c_void_p * 3)()
        n = C.c_size_t()
        self.ctx.check(self.ctx.lib.b200_key_buffers(self.h, bufs, C.byref(n)), "key_buffers")
        return [int(b) for b in bufs], int(n.value)

    def prove_partial_stage1(self, witness, chains):
        wp, keep = self._wptr(witness)
        self.ctx.check(self.ctx.lib.b200_prove_partial_stage1(self.h, wp, C.c_int(chains)), "prove_partial_stage1")

    def prove_partial_stage2(self, d_out_ptr):
        self.ctx.check(self.ctx.lib.b200_prove_partial_stage2(self.h, C.c_void_p(d_out_ptr)), "prove_partial_stage2")

    def combine_dev(self, d_parts_ptr, world):
        """b200_proof_combine_dev: `world` packed partials already on this device -> the proof (standard form)"""
        A = np.zeros(8, np.uint64); B = np.zeros(16, np.uint64); Cc = np.zeros(8, np.uint64); inf = np.zeros(3, np.uint8)
        self.ctx.check(self.ctx.lib.b200_proof_combine_dev(self.h, C.c_void_p(d_parts_ptr), world, _p64(A), _p64(B), _p64(Cc),
                                                           inf.ctypes.data_as(u8p)), "proof_combine_dev")
        return A, B, Cc, inf

    def combine(self, partials, flags):
        partials, flags = _arr(partials), _arr(flags, np.uint8)
        world = partials.size // 32
        A = np.zeros(8, np.uint64); B = np.zeros(16, np.uint64); Cc = np.zeros(8, np.uint64); inf = np.zeros(3, np.uint8)
        self.ctx.check(self.ctx.lib.b200_proof_combine(self.h, _p64(partials), flags.ctypes.data_as(u8p), world, _p64(A), _p64(B),
                                                       _p64(Cc), inf.ctypes.data_as(u8p)), "proof_combine")
        return A, B, Cc, inf

    def free(self):
        for slot in list(self._slots):
            slot.free()
        if self.h and self.ctx.h:
            self.ctx.lib.b200_key_free(self.h)
        self.h = C.c_void_p()

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass
