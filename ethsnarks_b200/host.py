"""Host-side mirror of the reference's prover entry points, on top of the C ABI (lib.py).

Names, argument meaning and error behaviour follow the reference (paths relative to the reference tree):
  Config                     src/prover_config.hpp:8-35        (CPU-only knobs are accepted and ignored)
  ProverContext              src/r1cs_gg_ppzksnark_zok/r1cs_gg_ppzksnark_zok.hpp:279-291
  load_proving_key           src/stubs.cpp:36-39
  get_domain                 src/stubs.cpp:49-75               (returns the domain size; the tables live on the device)
  prove                      src/stubs.cpp:42-47               (returns the proof JSON)
  proof_to_json              src/export.cpp:99-121
  Protoboard / R1CS          just enough of libsnark's protoboard (values, primary_input, constraint_system) to carry a
                             witness and a constraint list into prove(); gadget construction is out of scope
  generate_r1cs_example_*    depends/libsnark/libsnark/relations/constraint_satisfaction_problems/r1cs/examples/r1cs_examples.tcc:25-161
"""
import hashlib
import struct
import numpy as np
from . import lib as _lib

R_MOD = 21888242871839275222246405745257275088548364400416034343698204186575808495617
Q_MOD = 21888242871839275222246405745257275088696311157297823662689037894645226208583
MONT_R = 1 << 256


def ints_to_limbs(xs):
    """list of ints (< 2^256) -> (n, 4) uint64 little-endian limbs"""
    buf = b"".join(int(x).to_bytes(32, "little") for x in xs)
    return np.frombuffer(buf, dtype=np.uint64).reshape(-1, 4).copy()


def limbs_to_ints(a):
    a = np.ascontiguousarray(a, dtype=np.uint64).reshape(-1, 4)
    raw = a.tobytes()
    return [int.from_bytes(raw[i * 32:(i + 1) * 32], "little") for i in range(a.shape[0])]


def fr_to_mont(xs):
    return ints_to_limbs([(x * MONT_R) % R_MOD for x in xs])


def fr_from_mont(a):
    rinv = pow(MONT_R, -1, R_MOD)
    return [(x * rinv) % R_MOD for x in limbs_to_ints(a)]


def sha512_rng(idx, p=R_MOD):
    """libff::SHA512_rng<FieldT>(idx) (depends/libsnark/depends/libff/libff/common/rng.tcc:25-71)"""
    it = 0
    while True:
        h = hashlib.sha512(struct.pack("<QQ", idx, it)).digest()
        v = int.from_bytes(h[:32], "little") & ((1 << 254) - 1)
        if v < p:
            return v
        it += 1


class Config:
    """libsnark::Config (src/prover_config.hpp:8-35): same fields, same defaults."""

    def __init__(self):
        self.num_threads = 1
        self.smt = False
        self.fft = "recursive"
        self.radixes = []
        self.swapAB = True
        self.multi_exp_c = 0                 # 0 -> chosen by the backend's cost model (the CPU reference maps 0 -> 16)
        self.multi_exp_prefetch_locality = 0
        self.prefetch_stride = 128
        self.multi_exp_look_ahead = 1


class R1CS:
    """Constraint system as three CSR matrices over a shared coefficient table."""

    def __init__(self, num_inputs, num_variables, A, B, C, coeffs):
        self.num_inputs, self.num_variables = num_inputs, num_variables
        self.A, self.B, self.C = A, B, C          # (row_ptr u64, col u32, coeff_idx u32)
        self.coeffs = list(coeffs)                # ints, standard form
        self.num_constraints = len(A[0]) - 1

    def is_satisfied(self, values):
        def row(M, i):
            rp, col, ci = M
            return sum(self.coeffs[ci[k]] * values[col[k]] for k in range(int(rp[i]), int(rp[i + 1]))) % R_MOD
        return all((row(self.A, i) * row(self.B, i) - row(self.C, i)) % R_MOD == 0 for i in range(self.num_constraints))


class Protoboard:
    """values[0] is the constant ONE (LS/gadgetlib1/protoboard.tcc:26-33)."""

    def __init__(self, cs, assignment):
        self.cs = cs
        self.values = [1] + [int(v) % R_MOD for v in assignment]
        assert len(self.values) == cs.num_variables + 1

    def primary_input(self):
        return self.values[1:1 + self.cs.num_inputs]

    def get_constraint_system(self):
        return self.cs

    def num_constraints(self):
        return self.cs.num_constraints

    def num_inputs(self):
        return self.cs.num_inputs


class Translator:
    """libsnark::ITranslator (LS/relations/variable.hpp:279-284): maps a master gadget's variable indices to one instance's.
    `table[i]` is the instance's index of master variable i (table[0] is unused: index 0, the constant ONE, is never
    translated - `lt.index == 0 ? coeff : assignment[translate(lt.index)] * coeff`, variable.hpp:306-314)."""

    def __init__(self, table):
        self.table = np.asarray(table, dtype=np.uint32)

    def translate(self, index):
        return int(self.table[index])


def append_constraint_instances(cs, master, translators):
    """The reference's `r1cs_constraint_light_instance(master_constraint, translator)` (r1cs.hpp:145-211), as its gadgets use
    it (one instance constraint per master constraint per translator: poseidon.hpp:455-462, sha256_many.hpp:75-82): appends,
    for every translator in order, all rows of `master` (an R1CS whose variable indices are master-local) with their indices
    translated, to `cs`.  This is getA/getB/getC (r1cs.hpp:158-185) for all rows at once: one numpy gather per matrix
    instead of a virtual call and a vector copy per row, and the flattened CSR is what `b200_circuit_load` takes.  The
    coefficient tables must be the same list (the reference's global ConstantStorage, variable.hpp:47-86)."""
    assert cs.coeffs == master.coeffs, "master and instances share one coefficient table"
    out = []
    for M, N in ((cs.A, master.A), (cs.B, master.B), (cs.C, master.C)):
        rp, col, ci = (np.asarray(x) for x in M)
        mrp, mcol, mci = (np.asarray(x) for x in N)
        nnz = int(mrp[-1])
        cols, cis, rps = [col], [ci], [rp]
        base = int(rp[-1])
        for t in translators:
            tab = t.table if isinstance(t, Translator) else np.asarray(t, dtype=np.uint32)
            cols.append(np.where(mcol == 0, np.uint32(0), tab[mcol]).astype(np.uint32))
            cis.append(mci.astype(np.uint32))
            rps.append((mrp[1:] + np.uint64(base)).astype(np.uint64))
            base += nnz
        out.append((np.concatenate(rps), np.concatenate(cols), np.concatenate(cis)))
    cs.A, cs.B, cs.C = out
    cs.num_constraints = len(cs.A[0]) - 1
    return cs


def _csr(rows):
    rp = np.zeros(len(rows) + 1, dtype=np.uint64)
    cols, cis = [], []
    for i, r in enumerate(rows):
        for (v, c) in r:
            cols.append(v); cis.append(c)
        rp[i + 1] = len(cols)
    return rp, np.array(cols, dtype=np.uint32), np.array(cis, dtype=np.uint32)


def generate_r1cs_example_with_field_input(num_constraints, num_inputs, seed_a=0, seed_b=1):
    """r1cs_examples.tcc:25-93 with a, b = SHA512_rng(seed_a), SHA512_rng(seed_b) instead of random_element()."""
    assert num_inputs <= num_constraints + 2
    nv = num_constraints + 2
    a, b = sha512_rng(seed_a), sha512_rng(seed_b)
    fa = [a, b]
    # CSR built directly in numpy-friendly lists (coefficient table = [1])
    rpA = np.zeros(num_constraints + 1, np.uint64); rpB = np.zeros(num_constraints + 1, np.uint64); rpC = np.zeros(num_constraints + 1, np.uint64)
    colA, colB, colC = [], [], []
    for i in range(num_constraints - 1):
        if i % 2:
            colA.append(i + 1); colB.append(i + 2); colC.append(i + 3)
            tmp = a * b % R_MOD
        else:
            colB.append(0); colA.append(i + 1); colA.append(i + 2); colC.append(i + 3)
            tmp = (a + b) % R_MOD
        fa.append(tmp)
        a, b = b, tmp
        rpA[i + 1] = len(colA); rpB[i + 1] = len(colB); rpC[i + 1] = len(colC)
    fin = sum(fa[:nv - 1]) % R_MOD
    dense = list(range(1, nv))
    colA += dense; colB += dense; colC.append(nv)
    rpA[num_constraints] = len(colA); rpB[num_constraints] = len(colB); rpC[num_constraints] = len(colC)
    fa.append(fin * fin % R_MOD)
    mk = lambda rp, col: (rp, np.array(col, dtype=np.uint32), np.zeros(len(col), dtype=np.uint32))
    cs = R1CS(num_inputs, nv, mk(rpA, colA), mk(rpB, colB), mk(rpC, colC), [1])
    return Protoboard(cs, fa)


def generate_r1cs_example_with_binary_input(num_constraints, num_inputs, seed=1):
    """r1cs_examples.tcc:96-161 with std::rand replaced by a seeded 64-bit LCG (all witness values are 0/1)."""
    assert num_inputs >= 1
    mask = (1 << 64) - 1
    st = (seed * 6364136223846793005 + 1442695040888963407) & mask

    def rnd():
        nonlocal st
        st = (st * 6364136223846793005 + 1442695040888963407) & mask
        return st >> 33
    bits = [rnd() % 2 for _ in range(num_inputs)]
    rowsA, rowsB, rowsC = [], [], []
    lastvar = num_inputs - 1
    for i in range(num_constraints):
        lastvar += 1
        u = rnd() % num_inputs if i == 0 else rnd() % i
        v = rnd() % num_inputs if i == 0 else rnd() % i
        rowsA.append([(u + 1, 2)]); rowsB.append([(v + 1, 0)])
        rowsC.append(([(u + 1, 2)] if u == v else [(u + 1, 0), (v + 1, 0)]) + [(lastvar + 1, 1)])
        bits.append(bits[u] ^ bits[v])
    cs = R1CS(num_inputs, num_inputs + num_constraints, _csr(rowsA), _csr(rowsB), _csr(rowsC), [1, R_MOD - 1, 2, R_MOD - 2])
    return Protoboard(cs, bits)


# The reference prover's own timing blocks (libff::enter_block names, src/r1cs_gg_ppzksnark_zok/r1cs_gg_ppzksnark_zok.tcc:454-544) and
# the library phases (b200_timing_get names) each of them covers, so that CPU and GPU phase tables can be laid side by side.
REFERENCE_BLOCKS = {
    "Call to r1cs_gg_ppzksnark_zok_prover": ("prove.total",),
    "Compute the polynomial H": ("h2d.witness", "witness_eval", "h_chain", "h_chain.front", "h_chain.back"),
    "Compute evaluation to A-query": ("msm.A",),
    "Compute evaluation to B-query": ("msm.B_g2",),
    "Compute evaluation to H-query": ("msm.H",),
    "Compute evaluation to L-query": ("msm.L",),
    "Compute the proof": ("msm.A", "msm.B_g2", "msm.H", "msm.L", "finish", "combine"),
}


def phases_as_reference_blocks(phases_ms):
    """{library phase: ms} (serialised phases) -> {reference block name: ms}"""
    return {blk: float(sum(phases_ms.get(n, 0.0) for n in names)) for blk, names in REFERENCE_BLOCKS.items()}


def _hex(x):
    return "0x%x" % x


def proof_to_json(proof, primary_input, curve="alt"):
    """src/export.cpp:56-121: affine hex, G2 as [x.c1, x.c0], [y.c1, y.c0]; byte-for-byte the reference's layout.
    The one place where the string depends on the reference's curve build is a proof point at infinity (no real proof has
    one): the CURVE_ALT_BN128 branch prints libff's zero (0, 1) (alt_bn128_g1.cpp to_affine_coordinates), the default
    CURVE_MCL_BN128 branch prints the cleared mcl point, (0, 0) - `curve` selects which."""
    A, B, Cc, inf = proof
    a = limbs_to_ints(A); b = limbs_to_ints(B); c = limbs_to_ints(Cc)
    y0 = 1 if curve == "alt" else 0
    if inf[0]: a = [0, y0]
    if inf[1]: b = [0, 0, y0, 0]
    if inf[2]: c = [0, y0]
    s = "{\n"
    s += ' "A" :["%s", "%s"],\n' % (_hex(a[0]), _hex(a[1]))
    s += ' "B"  :[["%s", "%s"],\n ["%s", "%s"]],\n' % (_hex(b[1]), _hex(b[0]), _hex(b[3]), _hex(b[2]))
    s += ' "C"  :["%s", "%s"],\n' % (_hex(c[0]), _hex(c[1]))
    s += ' "input" :[' + ", ".join('"%s"' % _hex(v) for v in primary_input) + "]\n"
    s += "}"
    return s


def fq_from_mont(a):
    """Montgomery Fq limbs (n x 4) -> python ints (host-side; only used for the few points of a verification key)."""
    rinv = pow(1 << 256, -1, Q_MOD)
    return [x * rinv % Q_MOD for x in limbs_to_ints(a)]


def vk_to_json(vk):
    """vk2json (src/export.cpp:124-145) over the pieces ProvingKey.generate returns."""
    def g1(p):
        x, y = fq_from_mont(np.asarray(p).reshape(2, 4))
        return '"%s", "%s"' % (_hex(x), _hex(y)) if (x or y) else '"0x0", "0x1"'

    def g2(p):
        xa, xb, ya, yb = fq_from_mont(np.asarray(p).reshape(4, 4))
        if not (xa or xb or ya or yb):
            ya = 1
        return '["%s", "%s"],\n ["%s", "%s"]' % (_hex(xb), _hex(xa), _hex(yb), _hex(ya))

    s = "{\n"
    s += ' "alpha" :[' + g1(vk["alpha"]) + "],\n"
    s += ' "beta"  :[' + g2(vk["beta"]) + "],\n"
    s += ' "gamma" :[' + g2(vk["gamma"]) + "],\n"
    s += ' "delta" :[' + g2(vk["delta"]) + "],\n"
    s += '"gammaABC" :[' + ",".join("[" + g1(p) + "]" for p in vk["gammaABC"]) + "]}"
    return s


def pk_bellman2ethsnarks(bellman_json, alt=True):
    """pk_bellman2ethsnarks (src/export.cpp:267-328): a bellman-style proving key as JSON - "A", "B1", "B2", "C", "hExps" lists of
    Jacobian points ([x, y, z] decimal strings; G2 coordinates as [c0, c1] pairs) and vk_alfa_1, vk_beta_1/2, vk_delta_1/2 - to the
    bytes of the reference's stripped proving key (B200_FMT_ALT_NOZK, or MCL_NOZK with alt=False), which b200_key_load reads.
    As the reference does: B entries whose G1 half is zero are dropped, the first two "C" entries are skipped (:305), zero
    entries of A are dropped by the _nozk conversion (…zok.hpp:218-230), points are written normalised."""
    import json as _json
    d = _json.loads(bellman_json) if isinstance(bellman_json, (str, bytes)) else bellman_json
    q = Q_MOD
    one = MONT_R % q

    def fq(x):
        return ((int(x) % q) * MONT_R % q).to_bytes(32, "little")

    def g1(p):                                     # Jacobian (X, Y, Z) -> affine bytes, None for zero
        x, y, z = (int(v) % q for v in p)
        if z == 0:
            return None
        zi = pow(z, -1, q)
        return fq(x * zi * zi % q) + fq(y * zi * zi * zi % q)

    def g2(p):                                     # Fq2 = c0 + c1 u, u^2 = -1
        (x0, x1), (y0, y1), (z0, z1) = ((int(a) % q, int(b) % q) for a, b in p)
        if z0 == 0 and z1 == 0:
            return None
        mul = lambda a, b: ((a[0] * b[0] - a[1] * b[1]) % q, (a[0] * b[1] + a[1] * b[0]) % q)
        n = pow((z0 * z0 + z1 * z1) % q, -1, q)
        zi = (z0 * n % q, (-z1) * n % q)
        zi2 = mul(zi, zi)
        zi3 = mul(zi2, zi)
        X, Y = mul((x0, x1), zi2), mul((y0, y1), zi3)
        return fq(X[0]) + fq(X[1]) + fq(Y[0]) + fq(Y[1])

    def put(pt, nbytes):
        if alt:
            return (b"0" + pt) if pt is not None else (b"1" + bytes(nbytes // 2) + one.to_bytes(32, "little") + bytes(nbytes // 2 - 32))
        return (b"0" + pt + one.to_bytes(32, "little") + bytes(nbytes // 2 - 32)) if pt is not None else (b"1" + bytes(nbytes + nbytes // 2))

    out = [put(g1(d["vk_alfa_1"]), 64), put(g1(d["vk_beta_1"]), 64), put(g2(d["vk_beta_2"]), 128), put(g1(d["vk_delta_1"]), 64),
           put(g2(d["vk_delta_2"]), 128)]
    A = [g1(p) for p in d["A"]]
    keepA = [i for i, p in enumerate(A) if p is not None]
    out.append(("%d\n%d\n" % (len(A), len(keepA))).encode() + b"".join(b"%d\n" % i for i in keepA) + b"%d\n" % len(keepA))
    out += [put(A[i], 64) for i in keepA]
    keepB = [i for i, p in enumerate(d["B1"]) if g1(p) is not None]
    out.append(("%d\n%d\n" % (len(A), len(keepB))).encode() + b"".join(b"%d\n" % i for i in keepB) + b"%d\n" % len(keepB))
    out += [put(g2(d["B2"][i]), 128) for i in keepB]
    out.append(b"%d\n" % len(d["hExps"]))
    out += [put(g1(p), 64) for p in d["hExps"]]
    L = d["C"][2:]
    out.append(b"%d\n" % len(L))
    out += [put(g1(p), 64) for p in L]
    return b"".join(out)


def generate_keypair(ctx, pb, toxic, full=False):
    """libsnark's r1cs_gg_ppzksnark_zok_generator (…zok.tcc:277-449) on the device, with the toxic waste (t, alpha, beta,
    gamma, delta as python ints) passed in by the caller instead of drawn from FieldT::random_element().  Returns
    (ProvingKey without a circuit, vk dict); vk_to_json(vk) is the reference's vk.json."""
    cs = pb.get_constraint_system()
    return _lib.ProvingKey.generate(ctx, cs.A, cs.B, cs.C, fr_to_mont(cs.coeffs), cs.num_inputs, cs.num_variables,
                                    fr_to_mont(list(toxic)), full=full)


def load_proving_key(ctx, pk_file_or_bytes, fmt=_lib.FMT_MCL_NOZK, rank=0, world=1):
    """ethsnarks::load_proving_key (src/stubs.cpp:36-39); a missing file raises like the reference's assert."""
    if isinstance(pk_file_or_bytes, (bytes, bytearray, memoryview)):
        data = bytes(pk_file_or_bytes)
    else:
        with open(pk_file_or_bytes, "rb") as fh:
            data = fh.read()
    return _lib.ProvingKey.load(ctx, data, fmt, rank, world)


def get_domain(pb, pk=None, config=None):
    """src/stubs.cpp:49-75: the domain size roundUpToNearestPowerOf2(num_constraints + num_inputs + 1)."""
    need = pb.num_constraints() + pb.num_inputs() + 1
    m = 1
    while m < need:
        m *= 2
    return max(m, 2)


class ProverContext:
    """libsnark::ProverContext (…zok.hpp:279-291): the key, the constraint system, the domain and the scratch buffers
    that are reused across proofs — here they are device-resident behind the key handle."""

    def __init__(self, proving_key, pb, config=None):
        self.provingKey = proving_key
        self.config = config or Config()
        cs = pb.get_constraint_system()
        self.constraint_system = cs
        self.domain = get_domain(pb)
        if self.config.multi_exp_c:
            proving_key.ctx.set_option("window", self.config.multi_exp_c)
        proving_key.load_circuit(cs.A, cs.B, cs.C, fr_to_mont(cs.coeffs), cs.num_inputs, cs.num_variables)


def prove(context, pb, r=None, s=None):
    """ethsnarks::prove (src/stubs.cpp:42-47): witness in, proof JSON out."""
    w = fr_to_mont(pb.values)
    rr = fr_to_mont([r])[0] if r is not None else None
    ss = fr_to_mont([s])[0] if s is not None else None
    proof = context.provingKey.prove(w, rr, ss)
    return proof_to_json(proof, pb.primary_input())
