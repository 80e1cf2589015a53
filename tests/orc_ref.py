"""Driver for oracle/_ref/ref_prove — the REFERENCE's own prover templates built from /root/reference by
oracle/ref/Makefile.  TEST INFRASTRUCTURE ONLY.  Writes a problem file from oracle-side data, runs the reference binary,
returns its proof points, its H coefficients and its prover time."""
import os
import struct
import subprocess
import tempfile
import numpy as np
import orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "ref_prove")


DUMP_BIN = os.path.join(ROOT, "oracle", "_ref", "circuit_dump")


def available():
    return os.path.exists(REF_BIN)


def dump_available():
    return os.path.exists(DUMP_BIN)


def reference_circuit(name):
    """Builds `name` ("sha256") with the reference's own gadget code (oracle/_ref/circuit_dump) and returns
    (oracle Circuit, witness (nv+1, 4) Montgomery, the dumper's stdout)."""
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "circuit.bin")
        res = subprocess.run([DUMP_BIN, name, path], check=True, capture_output=True, text=True)
        raw = open(path, "rb").read()
    nc, ni, nv, ncoef = struct.unpack("<QQQQ", raw[:32])
    off = 32
    coeffs = np.frombuffer(raw, np.uint64, ncoef * 4, off).reshape(ncoef, 4).copy(); off += 32 * ncoef
    mats = []
    for _ in range(3):
        nnz = struct.unpack("<Q", raw[off:off + 8])[0]; off += 8
        rp = np.frombuffer(raw, np.uint64, nc + 1, off).copy(); off += 8 * (nc + 1)
        col = np.frombuffer(raw, np.uint32, nnz, off).copy(); off += 4 * nnz
        ci = np.frombuffer(raw, np.uint32, nnz, off).copy(); off += 4 * nnz
        mats.append((rp, col, ci))
    w = np.frombuffer(raw, np.uint64, (nv + 1) * 4, off).reshape(nv + 1, 4).copy()
    return orc.Circuit.from_csr(nc, ni, nv, mats, coeffs), w, res.stdout


def write_problem(path, circ, w, pk_alt_nozk_bytes):
    with open(path, "wb") as f:
        coeffs = circ.coeffs()
        f.write(struct.pack("<QQQQ", circ.nc, circ.ni, circ.nv, coeffs.shape[0]))
        f.write(coeffs.tobytes())
        for which in range(3):
            rp, col, ci = circ.matrix(which)
            f.write(struct.pack("<Q", len(col)))
            f.write(np.ascontiguousarray(rp).tobytes()); f.write(np.ascontiguousarray(col).tobytes()); f.write(np.ascontiguousarray(ci).tobytes())
        f.write(np.ascontiguousarray(w, dtype=np.uint64).tobytes())
        f.write(struct.pack("<Q", len(pk_alt_nozk_bytes)))
        f.write(pk_alt_nozk_bytes)


def prove_phases(circ, w, key, repeats=2, threads=None):
    """mean milliseconds of the reference prover's own enter_block/leave_block blocks (…zok.tcc:454-544) over `repeats` proves"""
    import json
    with tempfile.TemporaryDirectory() as d:
        prob, out = os.path.join(d, "problem.bin"), os.path.join(d, "out.bin")
        write_problem(prob, circ, w, key.serialize(orc.FMT_ALT_NOZK))
        env = dict(os.environ, REF_PHASES="1")
        if threads:
            env["OMP_NUM_THREADS"] = str(threads)
        res = subprocess.run([REF_BIN, prob, out, str(repeats)], check=True, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True)
    return json.loads([l for l in res.stderr.splitlines() if l.startswith("{")][-1])


def prove(circ, w, key, repeats=1, threads=None, all_times=False, budget_s=None, min_repeats=None):
    """returns ((A, B, C, inf) in the oracle's layout, H coefficients (m, 4) Montgomery, seconds); seconds = the fastest
    repeat, or with all_times the list of every repeat's duration in call order"""
    with tempfile.TemporaryDirectory() as d:
        prob, out = os.path.join(d, "problem.bin"), os.path.join(d, "out.bin")
        write_problem(prob, circ, w, key.serialize(orc.FMT_ALT_NOZK))
        env = dict(os.environ)
        if threads:
            env["OMP_NUM_THREADS"] = str(threads)
        cmd = [REF_BIN, prob, out, str(repeats)]
        if budget_s:
            cmd += [str(float(budget_s)), str(int(min_repeats or 1))]
        subprocess.run(cmd, check=True, env=env, stdout=subprocess.DEVNULL)
        raw = open(out, "rb").read()
    pts = np.frombuffer(raw[:8 * 32], dtype=np.uint64).reshape(8, 4)
    A = pts[0:2].reshape(-1).copy(); B = pts[2:6].reshape(-1).copy(); Cc = pts[6:8].reshape(-1).copy()
    m = struct.unpack("<Q", raw[256:264])[0]
    h = np.frombuffer(raw[264:264 + 32 * m], dtype=np.uint64).reshape(m, 4).copy()
    sec = struct.unpack("<d", raw[264 + 32 * m:272 + 32 * m])[0]
    if all_times:
        nt = struct.unpack("<Q", raw[272 + 32 * m:280 + 32 * m])[0]
        sec = list(struct.unpack("<%dd" % nt, raw[280 + 32 * m:280 + 32 * m + 8 * nt]))
    return (A, B, Cc, np.zeros(3, np.uint8)), h, sec
