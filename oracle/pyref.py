"""Big-integer CPU oracle for the halo2 MSM + NTT hot path (BN254).

TEST INFRASTRUCTURE ONLY.  Nothing on the product path may import this file: only
`tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / --impl reference leg.

PARITY UNPINNED: the arithmetic of the reference lives in the third-party crate
`halo2curves = "0.6.0"` (halo2_backend/Cargo.toml:31) which is absent from /root/reference,
and the reference's own tests hold no BN254 known-answer vector for this path
(SURVEY.md §4).  What this oracle IS checked against (tests/test_oracle.py):
  * the published BN254 constants (moduli, generator, 2-adicity, ROOT_OF_UNITY = 7^((r-1)/2^28)),
  * the public EIP-196 alt_bn128 vector 2*G,
  * the algebraic identities the reference's tests assert:
      zal.rs:345-349            engine.msm == best_multiexp            (MSM == naive double-and-add)
      kzg/commitment.rs:381-408 commit(lagrange_to_coeff(a)) == commit_lagrange(a)
      domain.rs:478-517         rotation / NTT orientation
Everything here is plain Python ints: obviously-correct, slow, for sizes <= 2^12.

Each function cites the reference file:line it restates.
"""
from __future__ import annotations

# ---------------------------------------------------------------------------------------------
# BN254 constants (checked numerically in tests/test_oracle.py)
# ---------------------------------------------------------------------------------------------
R_MOD = 0x30644E72E131A029B85045B68181585D2833E84879B9709143E1F593F0000001  # Fr modulus r
Q_MOD = 0x30644E72E131A029B85045B68181585D97816A916871CA8D3C208C16D87CFD47  # Fq modulus q
FR_S = 28                                   # 2-adicity of r-1
FR_GENERATOR = 7                            # multiplicative generator used by halo2curves
ROOT_OF_UNITY = pow(FR_GENERATOR, (R_MOD - 1) >> FR_S, R_MOD)
# halo2curves Fr::ZETA (cube root of unity; SURVEY.md §A.1 "second" root)
FR_ZETA = 0x30644E72E131A029048B6E193FD84104CC37A73FEC2BC5E9B8CA0B2D36636F23
MONT_R = 1 << 256
CURVE_B = 3
G1_GEN = (1, 2)


def fr(x: int) -> int:
    return x % R_MOD


def fr_inv(x: int) -> int:
    return pow(x, R_MOD - 2, R_MOD)


def fq_inv(x: int) -> int:
    return pow(x, Q_MOD - 2, Q_MOD)


# ---------------------------------------------------------------------------------------------
# Byte layouts at the boundary (SURVEY.md §8 preamble / §A.2): 4x u64 LE limbs, Montgomery form.
# ---------------------------------------------------------------------------------------------
def to_mont_bytes(x: int, mod: int) -> bytes:
    return ((x * MONT_R) % mod).to_bytes(32, "little")


def from_mont_bytes(b: bytes, mod: int) -> int:
    v = int.from_bytes(b, "little")
    assert v < mod, "non-canonical Montgomery limb value"
    return (v * pow(MONT_R, -1, mod)) % mod


def fr_vec_to_bytes(v) -> bytes:
    return b"".join(to_mont_bytes(x, R_MOD) for x in v)


def fr_vec_from_bytes(b: bytes):
    return [from_mont_bytes(b[i:i + 32], R_MOD) for i in range(0, len(b), 32)]


def affine_to_bytes(P) -> bytes:
    """G1Affine{x,y} 64 B, identity = (0,0)."""
    if P is None:
        return bytes(64)
    return to_mont_bytes(P[0], Q_MOD) + to_mont_bytes(P[1], Q_MOD)


def affine_from_bytes(b: bytes):
    x = from_mont_bytes(b[:32], Q_MOD)
    y = from_mont_bytes(b[32:64], Q_MOD)
    if x == 0 and y == 0:
        return None
    return (x, y)


def jacobian_from_bytes(b: bytes):
    """G1 {x,y,z} 96 B Montgomery; z == 0 is the identity.  Returns affine or None."""
    X = from_mont_bytes(b[:32], Q_MOD)
    Y = from_mont_bytes(b[32:64], Q_MOD)
    Z = from_mont_bytes(b[64:96], Q_MOD)
    if Z == 0:
        return None
    zi = fq_inv(Z)
    return (X * zi * zi % Q_MOD, Y * zi * zi * zi % Q_MOD)


# ---------------------------------------------------------------------------------------------
# G1 arithmetic, affine with None = identity.  y^2 = x^3 + 3 over Fq.
# ---------------------------------------------------------------------------------------------
def g1_is_on_curve(P) -> bool:
    if P is None:
        return True
    x, y = P
    return (y * y - x * x * x - CURVE_B) % Q_MOD == 0


def g1_neg(P):
    if P is None:
        return None
    return (P[0], (-P[1]) % Q_MOD)


def g1_add(P, Q):
    if P is None:
        return Q
    if Q is None:
        return P
    x1, y1 = P
    x2, y2 = Q
    if x1 == x2:
        if (y1 + y2) % Q_MOD == 0:
            return None
        lam = 3 * x1 * x1 * fq_inv(2 * y1) % Q_MOD
    else:
        lam = (y2 - y1) * fq_inv((x2 - x1) % Q_MOD) % Q_MOD
    x3 = (lam * lam - x1 - x2) % Q_MOD
    y3 = (lam * (x1 - x3) - y1) % Q_MOD
    return (x3, y3)


# Jacobian versions (used for speed in the larger python checks; a = 0 curve)
def _jac_double(P):
    X, Y, Z = P
    if Z == 0:
        return P
    A = X * X % Q_MOD
    B = Y * Y % Q_MOD
    C = B * B % Q_MOD
    D = 2 * ((X + B) * (X + B) - A - C) % Q_MOD
    E = 3 * A % Q_MOD
    F = E * E % Q_MOD
    X3 = (F - 2 * D) % Q_MOD
    Y3 = (E * (D - X3) - 8 * C) % Q_MOD
    Z3 = 2 * Y * Z % Q_MOD
    return (X3, Y3, Z3)


def _jac_add_affine(P, Q):
    """P Jacobian + Q affine (or None)."""
    if Q is None:
        return P
    X1, Y1, Z1 = P
    x2, y2 = Q
    if Z1 == 0:
        return (x2, y2, 1)
    Z1Z1 = Z1 * Z1 % Q_MOD
    U2 = x2 * Z1Z1 % Q_MOD
    S2 = y2 * Z1 * Z1Z1 % Q_MOD
    if U2 == X1:
        if S2 == Y1:
            return _jac_double(P)
        return (1, 1, 0)
    H = (U2 - X1) % Q_MOD
    HH = H * H % Q_MOD
    HHH = H * HH % Q_MOD
    r = (S2 - Y1) % Q_MOD
    V = X1 * HH % Q_MOD
    X3 = (r * r - HHH - 2 * V) % Q_MOD
    Y3 = (r * (V - X3) - Y1 * HHH) % Q_MOD
    Z3 = Z1 * H % Q_MOD
    return (X3, Y3, Z3)


def _jac_to_affine(P):
    X, Y, Z = P
    if Z == 0:
        return None
    zi = fq_inv(Z)
    zi2 = zi * zi % Q_MOD
    return (X * zi2 % Q_MOD, Y * zi2 * zi % Q_MOD)


def g1_mul(P, k: int):
    """[k]P, k reduced mod r, plain left-to-right double-and-add."""
    k %= R_MOD
    if P is None or k == 0:
        return None
    acc = (1, 1, 0)
    for bit in bin(k)[2:]:
        acc = _jac_double(acc)
        if bit == "1":
            acc = _jac_add_affine(acc, P)
    return _jac_to_affine(acc)


def msm_naive(scalars, bases):
    """sum_i scalars[i]*bases[i]: what MsmAccel::msm (zal.rs:58) / best_multiexp computes."""
    assert len(scalars) == len(bases)
    acc = None
    for s, P in zip(scalars, bases):
        acc = g1_add(acc, g1_mul(P, s))
    return acc


# ---------------------------------------------------------------------------------------------
# NTT: best_fft contract: in place, natural order in and out,
#      a[i] <- sum_j a[j] * omega^(i*j).
# ---------------------------------------------------------------------------------------------
def ntt_naive(a, omega):
    n = len(a)
    return [sum(a[j] * pow(omega, i * j, R_MOD) for j in range(n)) % R_MOD for i in range(n)]


def best_fft(a, omega, log_n):
    """Radix-2 DIT after a bit-reversal permutation; same result as ntt_naive."""
    n = 1 << log_n
    assert len(a) == n
    a = list(a)
    for i in range(n):
        j = int(format(i, "0%db" % log_n)[::-1], 2) if log_n else 0
        if i < j:
            a[i], a[j] = a[j], a[i]
    m = 1
    for _ in range(log_n):
        w_m = pow(omega, n // (2 * m), R_MOD)
        for k in range(0, n, 2 * m):
            w = 1
            for j in range(m):
                t = a[k + j + m] * w % R_MOD
                a[k + j + m] = (a[k + j] - t) % R_MOD
                a[k + j] = (a[k + j] + t) % R_MOD
                w = w * w_m % R_MOD
        m *= 2
    return a


class EvaluationDomain:
    """Restates halo2_backend/src/poly/domain.rs:38-144 (constants) and :216-351 (transforms)."""

    def __init__(self, j: int, k: int):
        self.quotient_poly_degree = j - 1                       # domain.rs:40
        self.k = k
        self.n = 1 << k                                         # domain.rs:43
        extended_k = k
        while (1 << extended_k) < self.n * self.quotient_poly_degree:   # domain.rs:48-51
            extended_k += 1
        assert extended_k <= FR_S                                # domain.rs:54
        self.extended_k = extended_k
        ext_omega = ROOT_OF_UNITY
        for _ in range(extended_k, FR_S):                        # domain.rs:61-63
            ext_omega = ext_omega * ext_omega % R_MOD
        omega = ext_omega
        for _ in range(k, extended_k):                           # domain.rs:72-74
            omega = omega * omega % R_MOD
        self.extended_omega = ext_omega
        self.omega = omega
        self.g_coset = FR_ZETA                                   # domain.rs:83
        self.g_coset_inv = FR_ZETA * FR_ZETA % R_MOD             # domain.rs:84
        # t_evaluations: domain.rs:86-111
        orig = pow(FR_ZETA, self.n, R_MOD)
        step = pow(ext_omega, self.n, R_MOD)
        t = []
        cur = orig
        while True:
            t.append(cur)
            cur = cur * step % R_MOD
            if cur == orig:
                break
        assert len(t) == 1 << (extended_k - k)
        self.t_evaluations = [fr_inv((c - 1) % R_MOD) for c in t]
        self.ifft_divisor = fr_inv(1 << k)                       # domain.rs:113,129
        self.extended_ifft_divisor = fr_inv(1 << extended_k)     # domain.rs:114
        self.barycentric_weight = fr_inv(self.n)
        self.extended_omega_inv = fr_inv(ext_omega)
        self.omega_inv = fr_inv(omega)

    def extended_len(self):
        return 1 << self.extended_k

    # domain.rs:343-351
    @staticmethod
    def _ifft(a, omega_inv, log_n, divisor):
        a = best_fft(a, omega_inv, log_n)
        return [x * divisor % R_MOD for x in a]

    # domain.rs:325-341
    def _distribute_powers_zeta(self, a, into_coset: bool):
        powers = [self.g_coset, self.g_coset_inv] if into_coset else [self.g_coset_inv, self.g_coset]
        out = []
        for index, x in enumerate(a):
            i = index % 3
            out.append(x if i == 0 else x * powers[i - 1] % R_MOD)
        return out

    def lagrange_to_coeff(self, a):                              # domain.rs:216-226
        assert len(a) == 1 << self.k
        return self._ifft(a, self.omega_inv, self.k, self.ifft_divisor)

    def coeff_to_extended(self, a):                              # domain.rs:230-244
        assert len(a) == 1 << self.k
        a = self._distribute_powers_zeta(a, True)
        a = a + [0] * (self.extended_len() - len(a))
        return best_fft(a, self.extended_omega, self.extended_k)

    def extended_to_coeff(self, a):                              # domain.rs:271-293
        assert len(a) == self.extended_len()
        a = self._ifft(a, self.extended_omega_inv, self.extended_k, self.extended_ifft_divisor)
        a = self._distribute_powers_zeta(a, False)
        return a[: self.n * self.quotient_poly_degree]

    def divide_by_vanishing_poly(self, a):                       # domain.rs:297-316
        assert len(a) == self.extended_len()
        e = len(self.t_evaluations)
        return [x * self.t_evaluations[i % e] % R_MOD for i, x in enumerate(a)]

    def rotate_omega(self, value, rotation: int):                # domain.rs:380-390
        if rotation >= 0:
            return value * pow(self.omega, rotation, R_MOD) % R_MOD
        return value * pow(self.omega_inv, -rotation, R_MOD) % R_MOD


def eval_polynomial(poly, point):
    """arithmetic.rs:57-82 (Horner)."""
    acc = 0
    for c in reversed(poly):
        acc = (acc * point + c) % R_MOD
    return acc


def kate_division(a, b):
    """arithmetic.rs:99-120: a(X) / (X - b), remainder dropped (python ints)."""
    b = (-b) % R_MOD
    q = [0] * (len(a) - 1)
    tmp = 0
    for i in range(len(a) - 1, 0, -1):
        lead = (a[i] - tmp) % R_MOD
        q[i - 1] = lead
        tmp = lead * b % R_MOD
    return q


def batch_invert(a):
    """ff::BatchInvert semantics: zeros stay zero."""
    return [fr_inv(x) if x % R_MOD else 0 for x in a]


def grand_product(v, init, out_len):
    """z[0] = init, z[i] = z[i-1] * v[i-1]  (permutation/prover.rs:160-166)."""
    z, acc = [], init % R_MOD
    for i in range(out_len):
        z.append(acc)
        if i < len(v):
            acc = acc * v[i] % R_MOD
    return z


# ---------------------------------------------------------------------------------------------
# quotient evaluation: plonk/evaluation.rs restated on python ints (small domains only)
# ---------------------------------------------------------------------------------------------
def get_rotation_idx(idx: int, rot: int, rot_scale: int, isize: int) -> int:
    """evaluation.rs:21-24."""
    return (idx + rot * rot_scale) % isize


def _value_source(v, rot_idx, constants, inter, fixed, advice, instance, challenges, beta, gamma, theta, y, prev):
    """ValueSource::get, evaluation.rs:58-96.  v = (kind, ...) with the variant order of evaluation.rs:26-47."""
    kind = v[0]
    if kind == 0: return constants[v[1]]
    if kind == 1: return inter[v[1]]
    if kind == 2: return fixed[v[1]][rot_idx[v[2]]]
    if kind == 3: return advice[v[1]][rot_idx[v[2]]]
    if kind == 4: return instance[v[1]][rot_idx[v[2]]]
    if kind == 5: return challenges[v[1]]
    return (beta, gamma, theta, y, prev)[kind - 6]


def graph_evaluate(calculations, constants, rotations, fixed, advice, instance, challenges, beta, gamma, theta, y, previous,
                   size: int, rot_scale: int):
    """GraphEvaluator::evaluate (evaluation.rs:789-834) for every row; Calculation::evaluate is evaluation.rs:124-172.
    calculations: [(calc, target)], calc = ("add", a, b) | ("sub", a, b) | ("mul", a, b) | ("square", a) | ("double", a)
    | ("negate", a) | ("horner", start, parts, factor) | ("store", a).  previous: list or None (zero)."""
    n_inter = 1 + max([t for _, t in calculations], default=-1)
    out = []
    for idx in range(size):
        rot_idx = [get_rotation_idx(idx, r, rot_scale, size) for r in rotations]
        inter = [0] * n_inter
        prev = previous[idx] if previous is not None else 0
        get = lambda v: _value_source(v, rot_idx, constants, inter, fixed, advice, instance, challenges, beta, gamma, theta, y, prev)
        for calc, target in calculations:
            op = calc[0]
            if op == "add": r = get(calc[1]) + get(calc[2])
            elif op == "sub": r = get(calc[1]) - get(calc[2])
            elif op == "mul": r = get(calc[1]) * get(calc[2])
            elif op == "square": r = get(calc[1]) ** 2
            elif op == "double": r = 2 * get(calc[1])
            elif op == "negate": r = -get(calc[1])
            elif op == "horner":
                factor, r = get(calc[3]), get(calc[1])
                for part in calc[2]:
                    r = (r * factor + get(part)) % R_MOD
            else: r = get(calc[1])
            inter[target] = r % R_MOD
        out.append(inter[calculations[-1][1]] if calculations else 0)
    return out


def evaluate_expression(expr, idx, size, rot_scale, fixed, advice, instance, challenges):
    """The direct evaluator of an ExpressionBack (evaluation.rs:837-880), used to check the graph compiler."""
    tag = expr[0]
    if tag == "const": return expr[1] % R_MOD
    if tag == "query":
        cols = {"fixed": fixed, "advice": advice, "instance": instance}[expr[1]]
        return cols[expr[2]][get_rotation_idx(idx, expr[3], rot_scale, size)]
    if tag == "challenge": return challenges[expr[1]]
    ev = lambda e: evaluate_expression(e, idx, size, rot_scale, fixed, advice, instance, challenges)
    if tag == "neg": return (-ev(expr[1])) % R_MOD
    if tag == "sum": return (ev(expr[1]) + ev(expr[2])) % R_MOD
    if tag == "prod": return ev(expr[1]) * ev(expr[2]) % R_MOD
    raise ValueError(tag)


def permutation_fold(values, sets, col_values, col_perms, chunk_len, l0, l_last, l_active, beta, gamma, y, zeta, delta,
                     extended_omega, blinding_factors, size, rot_scale):
    """evaluation.rs:397-493."""
    if not sets:
        return list(values)
    out = []
    last_rotation = -(blinding_factors + 1)
    delta_start = beta * zeta % R_MOD
    beta_term = 1
    for idx in range(size):
        v = values[idx]
        r_next = get_rotation_idx(idx, 1, rot_scale, size)
        r_last = get_rotation_idx(idx, last_rotation, rot_scale, size)
        v = (v * y + (1 - sets[0][idx]) * l0[idx]) % R_MOD
        zl = sets[-1][idx]
        v = (v * y + (zl * zl - zl) * l_last[idx]) % R_MOD
        for s in range(1, len(sets)):
            v = (v * y + (sets[s][idx] - sets[s - 1][r_last]) * l0[idx]) % R_MOD
        current_delta = delta_start * beta_term % R_MOD
        for s, z in enumerate(sets):
            cols = range(s * chunk_len, min((s + 1) * chunk_len, len(col_values)))
            left = z[r_next]
            for c in cols:
                left = left * (col_values[c][idx] + beta * col_perms[c][idx] + gamma) % R_MOD
            right = z[idx]
            for c in cols:
                right = right * (col_values[c][idx] + current_delta + gamma) % R_MOD
                current_delta = current_delta * delta % R_MOD
            v = (v * y + (left - right) * l_active[idx]) % R_MOD
        out.append(v)
        beta_term = beta_term * extended_omega % R_MOD
    return out


def lookup_fold(values, table_values, product, permuted_input, permuted_table, l0, l_last, l_active, beta, gamma, y, size, rot_scale):
    """evaluation.rs:517-565."""
    out = []
    for idx in range(size):
        v = values[idx]
        r_next, r_prev = get_rotation_idx(idx, 1, rot_scale, size), get_rotation_idx(idx, -1, rot_scale, size)
        a_minus_s = permuted_input[idx] - permuted_table[idx]
        z = product[idx]
        v = (v * y + (1 - z) * l0[idx]) % R_MOD
        v = (v * y + (z * z - z) * l_last[idx]) % R_MOD
        v = (v * y + (product[r_next] * (permuted_input[idx] + beta) * (permuted_table[idx] + gamma) - z * table_values[idx])
             * l_active[idx]) % R_MOD
        v = (v * y + a_minus_s * l0[idx]) % R_MOD
        v = (v * y + a_minus_s * (permuted_input[idx] - permuted_input[r_prev]) * l_active[idx]) % R_MOD
        out.append(v)
    return out


def shuffle_fold(values, input_values, shuffle_values, product, l0, l_last, l_active, y, size, rot_scale):
    """evaluation.rs:573-620."""
    out = []
    for idx in range(size):
        v = values[idx]
        r_next = get_rotation_idx(idx, 1, rot_scale, size)
        z = product[idx]
        v = (v * y + (1 - z) * l0[idx]) % R_MOD
        v = (v * y + (z * z - z) * l_last[idx]) % R_MOD
        v = (v * y + l_active[idx] * (product[r_next] * shuffle_values[idx] - z * input_values[idx])) % R_MOD
        out.append(v)
    return out


# ---------------------------------------------------------------------------------------------
# KZG SRS with a known secret: kzg/commitment.rs:64-131
# ---------------------------------------------------------------------------------------------
class ParamsKZG:
    def __init__(self, k: int, s: int):
        self.k = k
        self.n = 1 << k
        self.s = s % R_MOD
        n = self.n
        # g[i] = [s^i] G1                                    commitment.rs:74-83
        self.g = []
        sp = 1
        for _ in range(n):
            self.g.append(g1_mul(G1_GEN, sp))
            sp = sp * self.s % R_MOD
        # g_lagrange[i] = [ (s^n - 1)/n * w^i / (s - w^i) ] G1  commitment.rs:95-112
        root = ROOT_OF_UNITY
        for _ in range(k, FR_S):
            root = root * root % R_MOD
        n_inv = fr_inv(n)
        multiplier = (pow(self.s, n, R_MOD) - 1) * n_inv % R_MOD
        self.g_lagrange = []
        for i in range(n):
            rp = pow(root, i, R_MOD)
            scalar = multiplier * rp % R_MOD * fr_inv((self.s - rp) % R_MOD) % R_MOD
            self.g_lagrange.append(g1_mul(G1_GEN, scalar))

    def commit(self, poly):                                      # commitment.rs:354-366
        assert len(self.g) >= len(poly)
        return msm_naive(poly, self.g[: len(poly)])

    def commit_lagrange(self, poly):                             # commitment.rs:305-317
        assert len(self.g_lagrange) >= len(poly)
        return msm_naive(poly, self.g_lagrange[: len(poly)])


# ---------------------------------------------------------------------------------------------
# Deterministic synthetic inputs shared by oracle, GPU tests and bench (splitmix64 stream).
# ---------------------------------------------------------------------------------------------
def splitmix64(state: int):
    state = (state + 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
    z = state
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
    return state, z ^ (z >> 31)


def random_fr(n: int, seed: int):
    """n Fr values: 4 splitmix64 words -> 256-bit, top 2 bits masked, rejection-free reduce."""
    out = []
    st = seed & 0xFFFFFFFFFFFFFFFF
    for _ in range(n):
        v = 0
        for limb in range(4):
            st, w = splitmix64(st)
            v |= w << (64 * limb)
        out.append(v % R_MOD)
    return out
