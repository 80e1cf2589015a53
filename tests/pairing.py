"""Pure-Python BN254 (alt_bn128) optimal-ate pairing and Groth16 verifier.  TEST INFRASTRUCTURE ONLY.

Restates the reference verifier r1cs_gg_ppzksnark_zok_verifier_strong_IC / online_verifier_weak_IC
(/root/reference/src/r1cs_gg_ppzksnark_zok/r1cs_gg_ppzksnark_zok.tcc:569-624) and the pairing it calls
(depends/libsnark/depends/libff/libff/algebra/curves/alt_bn128/alt_bn128_pairing.cpp), in the textbook form
Fq12 = Fq[w]/(w^12 - 18 w^6 + 82) with Fq2 embedded by u -> w^6 - 9.  It is pinned by the reference's own
known-answer vector (test/test_verify.py:10-12, copied in tests/kat.py): that vk/proof pair must verify, and any
flipped nibble must not.  Slow (seconds per check) and only used on a handful of proofs.
"""
P = 21888242871839275222246405745257275088696311157297823662689037894645226208583
R = 21888242871839275222246405745257275088548364400416034343698204186575808495617
ATE_LOOP = 29793968203157093288
LOG_ATE = 63
MODC = [82, 0, 0, 0, 0, 0, -18 % P, 0, 0, 0, 0, 0]   # w^12 = 18 w^6 - 82


def _deg(p):
    d = len(p) - 1
    while d and p[d] == 0:
        d -= 1
    return d


def _poly_div(a, b):
    """quotient of a / b over Fp (lists, low degree first)"""
    dega, degb = _deg(a), _deg(b)
    temp = list(a)
    o = [0] * len(a)
    inv_lead = pow(b[degb], -1, P)
    for i in range(dega - degb, -1, -1):
        q = temp[degb + i] * inv_lead % P
        o[i] = (o[i] + q) % P
        for c in range(degb + 1):
            temp[c + i] = (temp[c + i] - q * b[c]) % P
    return o[:_deg(o) + 1]


class F12:
    __slots__ = ("c",)

    def __init__(self, c):
        self.c = [x % P for x in c]

    @staticmethod
    def one():
        return F12([1] + [0] * 11)

    @staticmethod
    def zero():
        return F12([0] * 12)

    def __add__(self, o):
        return F12([a + b for a, b in zip(self.c, o.c)])

    def __sub__(self, o):
        return F12([a - b for a, b in zip(self.c, o.c)])

    def __neg__(self):
        return F12([-a for a in self.c])

    def __eq__(self, o):
        return self.c == o.c

    def scale(self, k):
        return F12([a * k for a in self.c])

    def __mul__(self, o):
        b = [0] * 23
        for i, x in enumerate(self.c):
            if x:
                for j, y in enumerate(o.c):
                    b[i + j] += x * y
        for e in range(22, 11, -1):   # reduce by w^12 = 18 w^6 - 82
            t = b[e] % P
            if t:
                b[e - 6] += 18 * t
                b[e - 12] -= 82 * t
        return F12(b[:12])

    def inv(self):
        lm, hm = [1] + [0] * 12, [0] * 13
        low, high = self.c + [0], MODC + [1]
        while _deg(low):
            r = _poly_div(high, low)
            r += [0] * (13 - len(r))
            nm, new = list(hm), list(high)
            for i in range(13):
                for j in range(13 - i):
                    nm[i + j] -= lm[i] * r[j]
                    new[i + j] -= low[i] * r[j]
            nm = [x % P for x in nm]
            new = [x % P for x in new]
            lm, low, hm, high = nm, new, lm, low
        k = pow(low[0], -1, P)
        return F12([x * k for x in lm[:12]])

    def __truediv__(self, o):
        return self * o.inv()

    def __pow__(self, e):
        r, b = F12.one(), self
        while e:
            if e & 1:
                r = r * b
            b = b * b
            e >>= 1
        return r


W = F12([0, 1] + [0] * 10)
W2, W3 = W * W, W * W * W


def twist(q):
    """q = ((xa, xb), (ya, yb)) with x = xa + xb*u on the twist -> point on E(Fq12)"""
    (xa, xb), (ya, yb) = q
    nx = F12([xa - 9 * xb] + [0] * 5 + [xb] + [0] * 5)
    ny = F12([ya - 9 * yb] + [0] * 5 + [yb] + [0] * 5)
    return (nx * W2, ny * W3)


def cast_g1(p):
    return (F12([p[0]] + [0] * 11), F12([p[1]] + [0] * 11))


def _double(pt):
    x, y = pt
    lam = (x * x).scale(3) / y.scale(2)
    nx = lam * lam - x.scale(2)
    return (nx, lam * (x - nx) - y)


def _add(p1, p2):
    if p1 is None or p2 is None:
        return p1 if p2 is None else p2
    x1, y1 = p1
    x2, y2 = p2
    if x1 == x2:
        return _double(p1) if y1 == y2 else None
    lam = (y2 - y1) / (x2 - x1)
    nx = lam * lam - x1 - x2
    return (nx, lam * (x1 - nx) - y1)


def _line(p1, p2, t):
    x1, y1 = p1
    x2, y2 = p2
    xt, yt = t
    if not (x1 == x2):
        lam = (y2 - y1) / (x2 - x1)
        return lam * (xt - x1) - (yt - y1)
    if y1 == y2:
        lam = (x1 * x1).scale(3) / y1.scale(2)
        return lam * (xt - x1) - (yt - y1)
    return xt - x1


def miller_loop(q12, p12):
    r, f = q12, F12.one()
    for i in range(LOG_ATE, -1, -1):
        f = f * f * _line(r, r, p12)
        r = _double(r)
        if ATE_LOOP & (1 << i):
            f = f * _line(r, q12, p12)
            r = _add(r, q12)
    q1 = (q12[0] ** P, q12[1] ** P)
    nq2 = (q1[0] ** P, -(q1[1] ** P))
    f = f * _line(r, q1, p12)
    r = _add(r, q1)
    f = f * _line(r, nq2, p12)
    return f


def final_exp(f):
    return f ** ((P ** 12 - 1) // R)


def g1_on_curve(p):
    return (p[1] * p[1] - p[0] ** 3 - 3) % P == 0


def g1_add(p1, p2):
    if p1 is None or p2 is None:
        return p1 if p2 is None else p2
    (x1, y1), (x2, y2) = p1, p2
    if x1 == x2:
        if (y1 + y2) % P == 0:
            return None
        lam = 3 * x1 * x1 * pow(2 * y1, -1, P) % P
    else:
        lam = (y2 - y1) * pow(x2 - x1, -1, P) % P
    nx = (lam * lam - x1 - x2) % P
    return (nx, (lam * (x1 - nx) - y1) % P)


def g1_mul(p, k):
    r = None
    while k:
        if k & 1:
            r = g1_add(r, p)
        p = g1_add(p, p)
        k >>= 1
    return r


def g1_neg(p):
    return None if p is None else (p[0], (-p[1]) % P)


def _h(s):
    return int(s, 16)


def parse_g1(j):
    return (_h(j[0]), _h(j[1]))


def parse_g2(j):
    """JSON order is [[x.c1, x.c0], [y.c1, y.c0]] (src/export.cpp:66-96) -> ((xa, xb), (ya, yb))"""
    return ((_h(j[0][1]), _h(j[0][0])), (_h(j[1][1]), _h(j[1][0])))


def groth16_verify(vk, proof):
    """vk / proof: dicts in the reference's JSON shape.  e(A,B) == e(alpha,beta) e(IC,gamma) e(C,delta)."""
    inputs = [_h(x) for x in proof["input"]]
    ic = [parse_g1(p) for p in vk["gammaABC"]]
    if len(inputs) + 1 != len(ic):
        return False
    acc = ic[0]
    for x, pt in zip(inputs, ic[1:]):
        acc = g1_add(acc, g1_mul(pt, x % R))
    A, C = parse_g1(proof["A"]), parse_g1(proof["C"])
    alpha = parse_g1(vk["alpha"])
    if not all(g1_on_curve(p) for p in (A, C, alpha, acc)):
        return False
    pairs = [(A, parse_g2(proof["B"])), (g1_neg(alpha), parse_g2(vk["beta"])), (g1_neg(acc), parse_g2(vk["gamma"])),
             (g1_neg(C), parse_g2(vk["delta"]))]
    f = F12.one()
    for g1p, g2p in pairs:
        f = f * miller_loop(twist(g2p), cast_g1(g1p))
    return final_exp(f) == F12.one()
