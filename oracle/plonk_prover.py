"""Independent CPU restatement of the reference's KZG prover -- TEST INFRASTRUCTURE ONLY.

This file is the oracle for SURVEY.md 8(a) row 15 ("prover call sites / RNG order"): a straight-line, python-int
restatement of what fixes the proof BYTES in the reference -- call order, transcript writes, RNG draws -- written from the
Rust sources alone.  It imports NOTHING from the product package (`yet-another-halo2-fork_b200` / `h2b200`): its own
transcript, its own RNGs, its own `parallelize` chunk rule, its own keygen, its own constraint-system query collection, its
own SHPLONK / GWC.  tests/test_abi.py::test_oracle_prover_is_independent checks that.  The product's proofs
(plonk.create_proof over the CUDA library) must equal this file's proofs byte for byte (tests/test_gpu_plonk.py, smoke()).

What each part follows (all paths under /root/reference):
    keygen_vk / keygen_pk            halo2_backend/src/plonk/keygen.rs:43-187, 192-368 (query collection)
    permutation keygen               halo2_backend/src/plonk/permutation/keygen.rs:30-280
    ConstraintSystemBack             halo2_backend/src/plonk/circuit.rs:47-189
    pinned vk Debug string + hash    halo2_backend/src/plonk.rs:165-244, circuit.rs:191-260, poly/domain.rs:469-476,
                                     halo2_middleware/src/{circuit,expression,lookup,shuffle,permutation,poly}.rs derives
    Prover::new / commit_phase / create_proof   halo2_backend/src/plonk/prover.rs:174-305, 309-494, 512-899
    lookup argument                  plonk/lookup/prover.rs:64-173, 182-324, 328-399, 410-494
    permutation argument             plonk/permutation/prover.rs:50-197, 214-332
    shuffle argument                 plonk/shuffle/prover.rs:43-255
    vanishing argument               plonk/vanishing/prover.rs:40-206
    evaluate_h                       plonk/evaluation.rs:317-623 (the constraint fold; expressions evaluated directly, which
                                     gives the same field elements as the GraphEvaluator), evaluate :836-871
    SHPLONK                          poly/kzg/multiopen/shplonk.rs:46-137, shplonk/prover.rs:27-305
    GWC                              poly/kzg/multiopen/gwc.rs:26-50, gwc/prover.rs:40-90
    arithmetic                       arithmetic.rs:57-82 (eval), 99-120 (kate), 124-172 (parallelize), 177-226 (lagrange)
    transcript                       transcript.rs:15-21, 313-322, 353-419, 509-540
    RNGs                             halo2_proofs/tests/frontend_backend_split.rs:477-489 (OneNg), rand_core 0.6 BlockRng,
                                     rand_chacha 0.3 ChaCha20Rng (64-bit counter, 64-word buffer)

PARITY UNPINNED where halo2curves 0.6 (absent from the tree) decides the bytes -- recalled, not verifiable here:
    * `Fr::random` = eight next_u64() as little-endian limbs of a 512-bit integer, reduced mod r;
    * `G1Affine::to_bytes` = x little-endian, bit 6 of byte 31 = y's canonical LSB, bit 7 = identity;
    * `{:?}` of Fr / Fq = "0x" + 64 lower-case hex digits (big-endian canonical), of G1Affine = "(x, y)" / "Infinity";
    * `Fr: Ord` = order of the canonical (non-Montgomery) integers -- it decides the sort in the lookup argument
      (lookup/prover.rs:418-432), the BTreeMap walk of the leftover table values (:434-466) and the BTreeSet point order
      of SHPLONK (shplonk.rs:46-137).
The `{:?}` forms are exercised by the reference's golden vk string (plonk_api.rs:664-1141: its field elements and points are
printed that way, and the oracle rebuilds the string, commitments included, from the circuit alone); the reference holds no
golden PROOF bytes, so the other three stay recalled -- and none of them changes whether a proof verifies (any total order
gives a valid lookup permutation, any injective point encoding a valid transcript), so the verifier says nothing about them.
The STRUCTURE of the pinned-vk Debug string is pinned: `render` reproduces the reference's own golden text
(halo2_proofs/tests/plonk_api.rs:664-1141) from its parse tree, see tests/test_oracle_prover.py.

Commitments use the SRS trapdoor: with the test secret s known, commit(p) = [p(s)]G and
commit_lagrange(v) = [sum_i v_i L_i(s)]G are the same group elements as the MSMs over g / g_lagrange
(kzg/commitment.rs:64-131, 305-317, 354-366).

A circuit is plain data (no classes from the package):
    circuit = {"k": k,
               "cs": {"num_fixed_columns", "num_advice_columns", "num_instance_columns", "num_challenges",
                      "unblinded_advice_columns": [...], "advice_column_phase": [...], "challenge_phase": [...],
                      "gates": [(name, expr)], "permutation_columns": [(kind, index)],
                      "lookups": [(name, [input exprs], [table exprs])], "shuffles": [(name, [input exprs], [shuffle exprs])],
                      "minimum_degree": None | int},
               "fixed": [[int] * n] * num_fixed_columns, "copies": [((kind, index), row, (kind, index), row)]}
    expr = ("const", v) | ("query", kind, column_index, rotation) | ("challenge", index) | ("neg", e) | ("sum", a, b) | ("prod", a, b)
    kind in {"advice", "fixed", "instance"}
"""
from __future__ import annotations

import hashlib
import re

from . import pyref as R

P = R.R_MOD
Q = R.Q_MOD
DELTA = pow(R.FR_GENERATOR, 1 << R.FR_S, P)              # PrimeField::DELTA = g^(2^S)  


def inv(x: int) -> int:
    return pow(x % P, -1, P)


# =====================================================================================================================
# RNGs
# =====================================================================================================================
class BlockRng:
    """rand_core 0.6 `BlockRng<R>`: a buffer of u32 results, refilled by `generate`."""

    def __init__(self, buf_words: int):
        self.results = [0] * buf_words
        self.index = buf_words

    def generate(self):                                   # -> list of len(self.results) u32
        raise NotImplementedError

    def _generate_and_set(self, index: int):
        self.results = self.generate()
        self.index = index

    def next_u32(self) -> int:
        if self.index >= len(self.results):
            self._generate_and_set(0)
        v = self.results[self.index]
        self.index += 1
        return v

    def next_u64(self) -> int:
        n, i = len(self.results), self.index
        if i < n - 1:
            self.index += 2
            return self.results[i] | (self.results[i + 1] << 32)
        if i >= n:
            self._generate_and_set(2)
            return self.results[0] | (self.results[1] << 32)
        x = self.results[n - 1]
        self._generate_and_set(1)
        return (self.results[0] << 32) | x

    def fill_bytes(self, length: int) -> bytes:
        out = bytearray()
        while len(out) < length:
            if self.index >= len(self.results):
                self._generate_and_set(0)
            # fill_via_u32_chunks: whole words are consumed, a partially used word is dropped
            need = length - len(out)
            words = min(len(self.results) - self.index, (need + 3) // 4)
            chunk = b"".join(w.to_bytes(4, "little") for w in self.results[self.index:self.index + words])
            out += chunk[:need]
            self.index += words
        return bytes(out)

    def random_fr(self) -> int:
        """`Fr::random(rng)` (halo2curves 0.6, recalled): from_u512 of eight next_u64 limbs, little-endian."""
        v = 0
        for i in range(8):
            v |= self.next_u64() << (64 * i)
        return v % P


class OneNg(BlockRng):
    """frontend_backend_split.rs:477-489: every generated u32 is 1."""

    def __init__(self):
        super().__init__(16)

    def generate(self):
        return [1] * 16


def _rotl32(v: int, c: int) -> int:
    return ((v << c) | (v >> (32 - c))) & 0xFFFFFFFF


def chacha20_block(key_words, counter: int):
    st = [0x61707865, 0x3320646E, 0x79622D32, 0x6B206574] + list(key_words) + [counter & 0xFFFFFFFF, (counter >> 32) & 0xFFFFFFFF, 0, 0]
    x = list(st)

    def qr(a, b, c, d):
        x[a] = (x[a] + x[b]) & 0xFFFFFFFF; x[d] = _rotl32(x[d] ^ x[a], 16)
        x[c] = (x[c] + x[d]) & 0xFFFFFFFF; x[b] = _rotl32(x[b] ^ x[c], 12)
        x[a] = (x[a] + x[b]) & 0xFFFFFFFF; x[d] = _rotl32(x[d] ^ x[a], 8)
        x[c] = (x[c] + x[d]) & 0xFFFFFFFF; x[b] = _rotl32(x[b] ^ x[c], 7)

    for _ in range(10):
        qr(0, 4, 8, 12); qr(1, 5, 9, 13); qr(2, 6, 10, 14); qr(3, 7, 11, 15)
        qr(0, 5, 10, 15); qr(1, 6, 11, 12); qr(2, 7, 8, 13); qr(3, 4, 9, 14)
    return [(a + b) & 0xFFFFFFFF for a, b in zip(x, st)]


class ChaCha20Rng(BlockRng):
    """rand_chacha 0.3 `ChaCha20Rng::from_seed`: 20 rounds, 64-bit block counter (words 12-13), stream 0, four blocks per refill."""

    def __init__(self, seed: bytes):
        assert len(seed) == 32
        super().__init__(64)
        self.key = [int.from_bytes(seed[4 * i:4 * i + 4], "little") for i in range(8)]
        self.counter = 0

    def generate(self):
        out = []
        for _ in range(4):
            out += chacha20_block(self.key, self.counter)
            self.counter += 1
        return out


# =====================================================================================================================
# curve / transcript
# =====================================================================================================================
def point_to_bytes(pt) -> bytes:
    """`G1Affine::to_bytes` (halo2curves 0.6, recalled): 32 bytes, x little-endian, bit 6 of the last byte = sign of y (its
    canonical LSB), bit 7 = identity."""
    if pt is None:
        b = bytearray(32)
        b[31] |= 0x80
        return bytes(b)
    x, y = pt
    b = bytearray(x.to_bytes(32, "little"))
    b[31] |= (y & 1) << 6
    return bytes(b)


class Blake2bWrite:
    """transcript.rs:313-322, 353-419: BLAKE2b-512 personalised "Halo2-Transcript"; prefixes 0 challenge, 1 point, 2 scalar."""

    def __init__(self):
        self.state = hashlib.blake2b(digest_size=64, person=b"Halo2-Transcript")
        self.proof = bytearray()

    def common_scalar(self, s: int):
        self.state.update(b"\x02")
        self.state.update((s % P).to_bytes(32, "little"))

    def common_point(self, pt):
        if pt is None:
            raise ValueError("cannot write points at infinity to the transcript")       # transcript.rs:401-406
        self.state.update(b"\x01")
        self.state.update(pt[0].to_bytes(32, "little"))
        self.state.update(pt[1].to_bytes(32, "little"))

    def write_point(self, pt):
        self.common_point(pt)
        self.proof += point_to_bytes(pt)

    def write_scalar(self, s: int):
        self.common_scalar(s)
        self.proof += (s % P).to_bytes(32, "little")

    def squeeze_challenge(self) -> int:
        self.state.update(b"\x00")
        digest = self.state.copy().digest()                 # the running state is NOT reset (transcript.rs:392-397)
        return int.from_bytes(digest, "little") % P         # Fr::from_uniform_bytes (transcript.rs:525-533)

    def finalize(self) -> bytes:
        return bytes(self.proof)


class Params:
    """KZG parameters with a known secret: commitments by the trapdoor (module docstring)."""

    def __init__(self, k: int, s: int):
        self.k, self.n, self.s = k, 1 << k, s % P
        omega = R.ROOT_OF_UNITY
        for _ in range(k, R.FR_S):
            omega = omega * omega % P
        mult = (pow(self.s, self.n, P) - 1) * inv(self.n) % P          # commitment.rs:95-112
        self.lagrange_at_s, w = [], 1
        for _ in range(self.n):
            self.lagrange_at_s.append(mult * w % P * inv(self.s - w) % P)
            w = w * omega % P

    def commit(self, coeffs):                               # commitment.rs:354-366 (the blind is ignored)
        assert len(coeffs) <= self.n
        return R.g1_mul(R.G1_GEN, R.eval_polynomial(coeffs, self.s))

    def commit_lagrange(self, values):                      # commitment.rs:305-317
        assert len(values) <= self.n
        return R.g1_mul(R.G1_GEN, sum(v * l for v, l in zip(values, self.lagrange_at_s)) % P)


# =====================================================================================================================
# arithmetic.rs
# =====================================================================================================================
def parallelize_chunks(total: int, num_threads: int):
    """arithmetic.rs:124-172: (offset, length) of the chunks `parallelize` hands out: the first total % T chunks have
    floor(total / T) + 1 items, the rest floor(total / T); nothing for empty chunks."""
    base = total // num_threads
    cutoff = total % num_threads
    split = cutoff * (base + 1)
    out = []
    if cutoff != 0:
        out += [(i * (base + 1), base + 1) for i in range(cutoff)]
    if base != 0:
        out += [(split + i * base, base) for i in range((total - split) // base)]
    return out


def lagrange_interpolate(points, evals):
    """arithmetic.rs:177-226: coefficients of the interpolant (the unique one, whatever the summation order)."""
    assert len(points) == len(evals)
    if len(points) == 1:
        return [evals[0] % P]
    final = [0] * len(points)
    for j, (xj, ev) in enumerate(zip(points, evals)):
        tmp = [1]
        for kk, xk in enumerate(points):
            if kk == j:
                continue
            d = inv(xj - xk)
            nxt = [0] * (len(tmp) + 1)
            for i, c in enumerate(tmp):                     # tmp * (X - xk) * d
                nxt[i] = (nxt[i] - c * xk % P * d) % P
                nxt[i + 1] = (nxt[i + 1] + c * d) % P
            tmp = nxt
        for i, c in enumerate(tmp):
            final[i] = (final[i] + c * ev) % P
    return final


def vanishing_eval(roots, z):                               # arithmetic.rs:228-247
    acc = 1
    for r_ in roots:
        acc = acc * (z - r_) % P
    return acc


# =====================================================================================================================
# constraint system (plonk/circuit.rs, keygen.rs:192-368)
# =====================================================================================================================
def expr_degree(e) -> int:
    t = e[0]
    if t in ("const", "challenge"):
        return 0
    if t == "query":
        return 1
    if t == "neg":
        return expr_degree(e[1])
    if t == "sum":
        return max(expr_degree(e[1]), expr_degree(e[2]))
    if t == "prod":
        return expr_degree(e[1]) + expr_degree(e[2])
    raise ValueError(t)


class ConstraintSystem:
    """`ConstraintSystemBack` with the query lists `collect_queries` builds (keygen.rs:311-341): gates, then lookups (inputs then
    tables, per lookup), then shuffles, then one Rotation::cur() query per permutation column; first occurrence wins."""

    def __init__(self, cs: dict):
        self.num_fixed_columns = cs["num_fixed_columns"]
        self.num_advice_columns = cs["num_advice_columns"]
        self.num_instance_columns = cs["num_instance_columns"]
        self.num_challenges = cs.get("num_challenges", 0)
        self.unblinded_advice_columns = list(cs.get("unblinded_advice_columns", []))
        self.advice_column_phase = list(cs.get("advice_column_phase") or [0] * self.num_advice_columns)
        self.challenge_phase = list(cs.get("challenge_phase", []))
        self.gates = [tuple(g) for g in cs.get("gates", [])]
        self.permutation_columns = [tuple(c) for c in cs.get("permutation_columns", [])]
        self.lookups = [tuple(l) for l in cs.get("lookups", [])]
        self.shuffles = [tuple(s) for s in cs.get("shuffles", [])]
        self.minimum_degree = cs.get("minimum_degree")
        self.queries = {"advice": [], "instance": [], "fixed": []}
        self.query_index = {}
        for _, poly in self.gates:
            self._collect(poly)
        for _, inputs, tables in self.lookups:
            for e in inputs:
                self._collect(e)
            for e in tables:
                self._collect(e)
        for _, inputs, shuffles in self.shuffles:
            for e in inputs:
                self._collect(e)
            for e in shuffles:
                self._collect(e)
        for kind, index in self.permutation_columns:
            self._add(kind, index, 0)
        self.num_advice_queries = [0] * self.num_advice_columns
        for col, _ in self.queries["advice"]:
            self.num_advice_queries[col] += 1

    def _add(self, kind, index, rot):
        key = (kind, index, rot)
        if key not in self.query_index:
            self.queries[kind].append((index, rot))
            self.query_index[key] = len(self.queries[kind]) - 1
        return self.query_index[key]

    def _collect(self, e):
        t = e[0]
        if t == "query":
            self._add(e[1], e[2], e[3])
        elif t == "neg":
            self._collect(e[1])
        elif t in ("sum", "prod"):
            self._collect(e[1])
            self._collect(e[2])

    def degree(self) -> int:                                # circuit.rs:80-117, 285-389
        deg = 3
        for _, inputs, tables in self.lookups:
            assert len(inputs) == len(tables)
            deg = max(deg, max(4, 2 + max([1] + [expr_degree(e) for e in inputs]) + max([1] + [expr_degree(e) for e in tables])))
        for _, inputs, shuffles in self.shuffles:
            assert len(inputs) == len(shuffles)
            deg = max(deg, max(2 + max([1] + [expr_degree(e) for e in shuffles]), 2 + max([1] + [expr_degree(e) for e in inputs])))
        for _, poly in self.gates:
            deg = max(deg, expr_degree(poly))
        return max(deg, self.minimum_degree if self.minimum_degree is not None else 1)

    def blinding_factors(self) -> int:                      # circuit.rs:121-152
        return max(3, max(self.num_advice_queries, default=1)) + 2

    def minimum_rows(self) -> int:
        return self.blinding_factors() + 3

    def phases(self):
        return list(range(max(self.advice_column_phase, default=0) + 1))


# =====================================================================================================================
# the pinned verifying key as Rust prints it with {:?}  (plonk.rs:189-244)
# =====================================================================================================================
# A tiny model of what `#[derive(Debug)]` / `Formatter::debug_*` emit:
#   ("struct", Name, [(field, node), ...])   Name { a: .., b: .. }     (no fields: just Name)
#   ("tuple", Name, [node, ...])             Name(a, b)                 (Name == "" : a plain tuple)
#   ("list", [node, ...])                    [a, b]
#   ("raw", text)                            numbers, "strings", 0x.. field elements, (0x.., 0x..) points
def render(node, pretty: bool = False, level: int = 0) -> str:
    kind = node[0]
    if kind == "raw":
        return node[1]
    pad, pad_in = "    " * level, "    " * (level + 1)
    if kind == "struct":
        _, name, fields = node
        if not fields:
            return name
        if pretty:
            body = "".join(f"{pad_in}{f}: {render(v, True, level + 1)},\n" for f, v in fields)
            return f"{name} {{\n{body}{pad}}}"
        return name + " { " + ", ".join(f"{f}: {render(v)}" for f, v in fields) + " }"
    if kind == "tuple":
        _, name, items = node
        if pretty:
            body = "".join(f"{pad_in}{render(v, True, level + 1)},\n" for v in items)
            return f"{name}(\n{body}{pad})"
        inner = ", ".join(render(v) for v in items)
        if name == "" and len(items) == 1:
            inner += ","                                     # Rust prints a 1-tuple as (a,)
        return f"{name}({inner})"
    if kind == "list":
        items = node[1]
        if not items:
            return "[]"
        if pretty:
            body = "".join(f"{pad_in}{render(v, True, level + 1)},\n" for v in items)
            return f"[\n{body}{pad}]"
        return "[" + ", ".join(render(v) for v in items) + "]"
    raise ValueError(kind)


_TOKEN = re.compile(r'\s*(\(0x[0-9a-f]+, 0x[0-9a-f]+\)|"(?:[^"\\]|\\.)*"|[A-Za-z_][A-Za-z0-9_]*|-?0x[0-9a-f]+|-?[0-9]+|[{}()\[\]:,])')


def parse_debug(text: str):
    """Parse Rust Debug output (either form) back into the node model above; used to pin `render` against the reference's own
    golden string."""
    tokens, pos = [], 0
    text = text.strip()
    while pos < len(text):
        m = _TOKEN.match(text, pos)
        if not m:
            raise ValueError(f"cannot tokenise at {pos}: {text[pos:pos + 40]!r}")
        tokens.append(m.group(1))
        pos = m.end()
    i = 0

    def peek():
        return tokens[i] if i < len(tokens) else None

    def take(expected=None):
        nonlocal i
        t = tokens[i]
        if expected is not None and t != expected:
            raise ValueError(f"expected {expected!r}, got {t!r}")
        i += 1
        return t

    def seq(close):
        items = []
        while peek() != close:
            items.append(value())
            if peek() == ",":
                take(",")
        take(close)
        return items

    def value():
        t = take()
        if t == "[":
            return ("list", seq("]"))
        if t == "(":
            return ("tuple", "", seq(")"))
        if re.fullmatch(r"[A-Za-z_][A-Za-z0-9_]*", t):
            if peek() == "{":
                take("{")
                fields = []
                while peek() != "}":
                    name = take()
                    take(":")
                    fields.append((name, value()))
                    if peek() == ",":
                        take(",")
                take("}")
                return ("struct", t, fields)
            if peek() == "(":
                take("(")
                return ("tuple", t, seq(")"))
            return ("struct", t, [])                          # unit: None, Advice, ...
        return ("raw", t)

    node = value()
    if i != len(tokens):
        raise ValueError("trailing tokens")
    return node


def _hex(v: int) -> str:
    return "0x%064x" % v


def _raw(v) -> tuple:
    return ("raw", str(v))


def _any(kind: str):
    return ("struct", {"advice": "Advice", "fixed": "Fixed", "instance": "Instance"}[kind], [])   # middleware circuit.rs:166-172


def _rotation(r: int):
    return ("tuple", "Rotation", [_raw(r)])


def _column(kind: str, index: int):
    return ("struct", "ColumnMid", [("column_type", _any(kind)), ("index", _raw(index))])        # middleware circuit.rs:201-207


def _point(pt):
    return ("raw", "Infinity" if pt is None else f"({_hex(pt[0])}, {_hex(pt[1])})")


def expr_node(cs: ConstraintSystem, e):
    """`ExpressionBack` Debug (middleware expression.rs:22-34; VarBack / QueryBack circuit.rs:8-21)."""
    t = e[0]
    if t == "const":
        return ("tuple", "Constant", [("raw", _hex(e[1] % P))])
    if t == "query":
        q = ("struct", "QueryBack", [("index", _raw(cs.query_index[(e[1], e[2], e[3])])), ("column_index", _raw(e[2])),
                                     ("column_type", _any(e[1])), ("rotation", _rotation(e[3]))])
        return ("tuple", "Var", [("tuple", "Query", [q])])
    if t == "challenge":
        ch = ("struct", "ChallengeMid", [("index", _raw(e[1])), ("phase", _raw(cs.challenge_phase[e[1]]))])
        return ("tuple", "Var", [("tuple", "Challenge", [ch])])
    if t == "neg":
        return ("tuple", "Negated", [expr_node(cs, e[1])])
    if t == "sum":
        return ("tuple", "Sum", [expr_node(cs, e[1]), expr_node(cs, e[2])])
    if t == "prod":
        return ("tuple", "Product", [expr_node(cs, e[1]), expr_node(cs, e[2])])
    raise ValueError(t)


def pinned_vk_node(cs: ConstraintSystem, domain, fixed_commitments, permutation_commitments, base_modulus=Q, scalar_modulus=P):
    """`domain` needs .k, .extended_k, .omega.  The moduli are parameters only so that the test can rebuild the reference's Pasta
    golden (plonk_api.rs:664-1141) through this very function."""
    qlist = lambda kind: ("list", [("tuple", "", [_column(kind, c), _rotation(r)]) for c, r in cs.queries[kind]])
    string = lambda s: ("raw", '"' + s.replace("\\", "\\\\").replace('"', '\\"') + '"')
    fields = [("num_fixed_columns", _raw(cs.num_fixed_columns)), ("num_advice_columns", _raw(cs.num_advice_columns)),
              ("num_instance_columns", _raw(cs.num_instance_columns))]
    if cs.num_challenges > 0:                               # circuit.rs:236-242: only shown when multi-phase is used
        fields += [("num_challenges", _raw(cs.num_challenges)), ("advice_column_phase", ("list", [_raw(p) for p in cs.advice_column_phase])),
                   ("challenge_phase", ("list", [_raw(p) for p in cs.challenge_phase]))]
    fields += [("gates", ("list", [expr_node(cs, poly) for _, poly in cs.gates])),     # PinnedGates: the polys only (circuit.rs:205-213)
               ("advice_queries", qlist("advice")), ("instance_queries", qlist("instance")), ("fixed_queries", qlist("fixed")),
               ("permutation", ("struct", "ArgumentMid", [("columns", ("list", [_column(k, i) for k, i in cs.permutation_columns]))])),
               ("lookups", ("list", [("struct", "Argument", [("name", string(name)),
                                                             ("input_expressions", ("list", [expr_node(cs, e) for e in ins])),
                                                             ("table_expressions", ("list", [expr_node(cs, e) for e in tabs]))])
                                     for name, ins, tabs in cs.lookups]))]
    if cs.shuffles:                                         # circuit.rs:250-252
        fields.append(("shuffles", ("list", [("struct", "Argument", [("name", string(name)),
                                                                     ("input_expressions", ("list", [expr_node(cs, e) for e in ins])),
                                                                     ("shuffle_expressions", ("list", [expr_node(cs, e) for e in shs]))])
                                             for name, ins, shs in cs.shuffles])))
    fields.append(("minimum_degree", ("struct", "None", []) if cs.minimum_degree is None else ("tuple", "Some", [_raw(cs.minimum_degree)])))
    return ("struct", "PinnedVerificationKey", [                                        # plonk.rs:238-246 field order
        ("base_modulus", ("raw", '"0x%064x"' % base_modulus)), ("scalar_modulus", ("raw", '"0x%064x"' % scalar_modulus)),
        ("domain", ("struct", "PinnedEvaluationDomain", [("k", _raw(domain.k)), ("extended_k", _raw(domain.extended_k)),
                                                         ("omega", ("raw", _hex(domain.omega)))])),
        ("cs", ("struct", "PinnedConstraintSystem", fields)),
        ("fixed_commitments", ("list", [_point(c) for c in fixed_commitments])),
        ("permutation", ("struct", "VerifyingKey", [("commitments", ("list", [_point(c) for c in permutation_commitments]))]))])


def transcript_repr_of(cs, domain, fixed_commitments, permutation_commitments) -> int:
    """plonk.rs:189-203: BLAKE2b-512 (personal "Halo2-Verify-Key") over len(s) as u64 LE, then s = format!("{:?}", vk.pinned())."""
    s = render(pinned_vk_node(cs, domain, fixed_commitments, permutation_commitments)).encode()
    h = hashlib.blake2b(digest_size=64, person=b"Halo2-Verify-Key")
    h.update(len(s).to_bytes(8, "little"))
    h.update(s)
    return int.from_bytes(h.digest(), "little") % P


# =====================================================================================================================
# keygen
# =====================================================================================================================
class Keys:
    pass


def permutation_mapping(n: int, columns, copies):
    """permutation/keygen.rs:30-117: the cycle structure after all `copy` calls."""
    m = len(columns)
    mapping = [[(i, j) for j in range(n)] for i in range(m)]
    aux = [[(i, j) for j in range(n)] for i in range(m)]
    sizes = [[1] * n for _ in range(m)]
    for lcol, lrow, rcol, rrow in copies:
        lc, rc = columns.index(tuple(lcol)), columns.index(tuple(rcol))
        if lrow >= n or rrow >= n:
            raise ValueError("BoundsFailure")
        left, right = aux[lc][lrow], aux[rc][rrow]
        if left == right:
            continue
        if sizes[left[0]][left[1]] < sizes[right[0]][right[1]]:
            left, right = right, left
        sizes[left[0]][left[1]] += sizes[right[0]][right[1]]
        i = right
        while True:
            aux[i[0]][i[1]] = left
            i = mapping[i[0]][i[1]]
            if i == right:
                break
        mapping[lc][lrow], mapping[rc][rrow] = mapping[rc][rrow], mapping[lc][lrow]
    return mapping


def keygen(circuit: dict, secret: int, transcript_repr=None) -> Keys:
    """keygen_vk + keygen_pk (keygen.rs:43-187).  `transcript_repr`: inject a value taken from a real Rust run instead of the
    restated Debug hash."""
    K = Keys()
    k = circuit["k"]
    K.cs = cs = ConstraintSystem(circuit["cs"])
    K.params = params = Params(k, secret)
    n = params.n
    K.cs_degree = cs.degree()
    K.domain = domain = R.EvaluationDomain(K.cs_degree, k)
    if n < cs.minimum_rows():
        raise ValueError("not enough rows available")
    K.fixed_values = [[v % P for v in col] for col in circuit["fixed"]]
    assert len(K.fixed_values) == cs.num_fixed_columns and all(len(c) == n for c in K.fixed_values)
    # permutation: sigma_i(omega^j) = delta^i' * omega^j' for (i', j') = mapping(i, j)        permutation/keygen.rs:138-175
    mapping = permutation_mapping(n, cs.permutation_columns, circuit["copies"])
    omega_powers = [pow(domain.omega, j, P) for j in range(n)]
    K.permutations = [[pow(DELTA, mapping[i][j][0], P) * omega_powers[mapping[i][j][1]] % P for j in range(n)]
                      for i in range(len(cs.permutation_columns))]
    K.permutation_commitments = [params.commit_lagrange(p) for p in K.permutations]            # build_vk :258-276
    K.fixed_commitments = [params.commit_lagrange(c) for c in K.fixed_values]                   # keygen.rs:66-83
    K.transcript_repr = (transcript_repr if transcript_repr is not None
                         else transcript_repr_of(cs, domain, K.fixed_commitments, K.permutation_commitments))
    # proving key                                                                                 keygen.rs:108-187
    K.fixed_polys = [domain.lagrange_to_coeff(c) for c in K.fixed_values]
    K.fixed_cosets = [domain.coeff_to_extended(p) for p in K.fixed_polys]
    bf = cs.blinding_factors()
    l0 = [0] * n; l0[0] = 1
    l_blind = [0] * n
    for i in range(bf):
        l_blind[n - 1 - i] = 1
    l_last = [0] * n; l_last[n - bf - 1] = 1
    ext = lambda v: domain.coeff_to_extended(domain.lagrange_to_coeff(v))
    K.l0, l_blind_c, K.l_last = ext(l0), ext(l_blind), ext(l_last)
    K.l_active_row = [(1 - (a + b)) % P for a, b in zip(K.l_last, l_blind_c)]
    K.permutation_polys = [domain.lagrange_to_coeff(p) for p in K.permutations]
    K.permutation_cosets = [domain.coeff_to_extended(p) for p in K.permutation_polys]
    return K


# =====================================================================================================================
# expression evaluation (evaluation.rs:836-871)
# =====================================================================================================================
def evaluate(e, size, rot_scale, fixed, advice, instance, challenges):
    out = []
    for idx in range(size):
        out.append(_eval_at(e, idx, size, rot_scale, fixed, advice, instance, challenges))
    return out


def _eval_at(e, idx, size, rot_scale, fixed, advice, instance, challenges):
    t = e[0]
    if t == "const":
        return e[1] % P
    if t == "query":
        col = {"fixed": fixed, "advice": advice, "instance": instance}[e[1]][e[2]]
        return col[(idx + e[3] * rot_scale) % size]         # get_rotation_idx, evaluation.rs:21-24
    if t == "challenge":
        return challenges[e[1]]
    if t == "neg":
        return (-_eval_at(e[1], idx, size, rot_scale, fixed, advice, instance, challenges)) % P
    a = _eval_at(e[1], idx, size, rot_scale, fixed, advice, instance, challenges)
    b = _eval_at(e[2], idx, size, rot_scale, fixed, advice, instance, challenges)
    return (a + b) % P if t == "sum" else a * b % P


def compress(exprs, theta, n, fixed, advice, instance, challenges):
    """lookup/prover.rs:93-111, shuffle/prover.rs:60-79: fold(empty, |acc, e| acc * theta + e)."""
    acc = [0] * n
    for e in exprs:
        vals = evaluate(e, n, 1, fixed, advice, instance, challenges)
        acc = [(a * theta + v) % P for a, v in zip(acc, vals)]
    return acc


def running_product(factors, n, bf, rng):
    """z = 1, f0, f0 f1, ... ; the first n - bf values are kept, then bf random blinding rows (lookup/prover.rs:254-265)."""
    z, acc = [], 1
    for i in range(n - bf):
        z.append(acc)
        acc = acc * factors[i] % P
    z += [rng.random_fr() for _ in range(bf)]
    return z


# =====================================================================================================================
# create_proof
# =====================================================================================================================
def create_proof(K: Keys, instances, witness_fn, rng, num_threads: int, multiopen: str = "shplonk") -> bytes:
    """`Prover::new` + `commit_phase` for every phase + `create_proof` for ONE circuit instance (prover.rs:174-899).
    instances: one list of canonical ints per instance column.  witness_fn(phase, challenges: dict) -> list with one entry per
    advice column: n ints for the columns of `phase`, None for the others."""
    return create_proof_batch(K, [instances], [witness_fn], rng, num_threads, multiopen)


def create_proof_batch(K: Keys, circuits_instances, witness_fns, rng, num_threads: int, multiopen: str = "shplonk") -> bytes:
    cs, domain, params = K.cs, K.domain, K.params
    n, bf = params.n, cs.blinding_factors()
    tr = Blake2bWrite()
    ncirc = len(circuits_instances)
    for inst in circuits_instances:
        if len(inst) != cs.num_instance_columns:
            raise ValueError("InvalidInstances")
    tr.common_scalar(K.transcript_repr)                                               # [TRANSCRIPT-1] prover.rs:194
    instance_values, instance_polys = [], []
    for inst in circuits_instances:                                                   # [TRANSCRIPT-2] prover.rs:203-271
        vals = []
        for column in inst:
            if len(column) > n - (bf + 1):
                raise ValueError("InstanceTooLarge")
            poly = [0] * n
            for i, v in enumerate(column):
                tr.common_scalar(v)                                                   # QUERY_INSTANCE == false for KZG
                poly[i] = v % P
            vals.append(poly)
        instance_values.append(vals)
        instance_polys.append([domain.lagrange_to_coeff(v) for v in vals])

    # ---- commit_phase per phase (prover.rs:309-494) ------------------------------------------------------------------
    advice_values = [[None] * cs.num_advice_columns for _ in range(ncirc)]
    advice_blinds = [[0] * cs.num_advice_columns for _ in range(ncirc)]
    challenges = {}
    unusable_rows_start = n - (bf + 1)
    for phase in cs.phases():
        column_indices = [i for i, p in enumerate(cs.advice_column_phase) if p == phase]
        for c in range(ncirc):
            witness = witness_fns[c](phase, dict(challenges))
            assert len(witness) == cs.num_advice_columns
            cols = []
            for i, w in enumerate(witness):
                if i in column_indices:
                    assert w is not None and len(w) == n, f"expected advice column {i} at phase {phase}"
                    cols.append([v % P for v in w])
                else:
                    assert w is None, f"expected no advice column {i} at phase {phase}"
            for i, col in zip(column_indices, cols):                                 # blinding rows first, column by column
                if i not in cs.unblinded_advice_columns:
                    for row in range(unusable_rows_start, n):
                        col[row] = rng.random_fr()
            blinds = [0 if i in cs.unblinded_advice_columns else rng.random_fr() for i in column_indices]   # then the blinds
            for col in cols:
                tr.write_point(params.commit_lagrange(col))                          # [TRANSCRIPT-3]
            for i, col, b in zip(column_indices, cols, blinds):
                advice_values[c][i], advice_blinds[c][i] = col, b
        for index, p in enumerate(cs.challenge_phase):                                # [TRANSCRIPT-4]
            if p == phase:
                assert index not in challenges
                challenges[index] = tr.squeeze_challenge()
    assert len(challenges) == cs.num_challenges
    challenges = [challenges[i] for i in range(cs.num_challenges)]

    # ---- create_proof (prover.rs:512-899) -------------------------------------------------------------------------------
    theta = tr.squeeze_challenge()                                                    # [TRANSCRIPT-5]
    permuted = []                                                                     # [TRANSCRIPT-6] lookup_commit_permuted
    for c in range(ncirc):
        per_circuit = []
        for _, ins, tabs in cs.lookups:
            L = {}
            L["compressed_input"] = compress(ins, theta, n, K.fixed_values, advice_values[c], instance_values[c], challenges)
            L["compressed_table"] = compress(tabs, theta, n, K.fixed_values, advice_values[c], instance_values[c], challenges)
            L["permuted_input"], L["permuted_table"] = permute_expression_pair(L["compressed_input"], L["compressed_table"], n, bf, rng)
            commits = []
            for name in ("permuted_input", "permuted_table"):                        # commit_values :130-143: iNTT, blind, commit
                L[name + "_poly"] = domain.lagrange_to_coeff(L[name])
                L[name + "_blind"] = rng.random_fr()
                commits.append(params.commit_lagrange(L[name]))
            for pt in commits:
                tr.write_point(pt)
            per_circuit.append(L)
        permuted.append(per_circuit)
    beta = tr.squeeze_challenge()                                                     # [TRANSCRIPT-7]
    gamma = tr.squeeze_challenge()                                                    # [TRANSCRIPT-8]

    # [TRANSCRIPT-9] permutation_commit (permutation/prover.rs:50-197)
    chunk_len = K.cs_degree - 2
    assert K.cs_degree >= 3
    perm_sets = []
    for c in range(ncirc):
        src = {"advice": advice_values[c], "fixed": K.fixed_values, "instance": instance_values[c]}
        sets, deltaomega, last_z = [], 1, 1
        cols_all = cs.permutation_columns
        for start in range(0, len(cols_all), chunk_len):
            columns = cols_all[start:start + chunk_len]
            perms = K.permutations[start:start + chunk_len]
            modified = [1] * n
            for (kind, index), perm in zip(columns, perms):
                vals = src[kind][index]
                modified = [m * (beta * pv + gamma + v) % P for m, v, pv in zip(modified, vals, perm)]
            modified = [inv(m) if m else 0 for m in modified]                       # batch_invert leaves zeros
            for kind, index in columns:
                vals = src[kind][index]
                cur = deltaomega
                for row in range(n):
                    modified[row] = modified[row] * (cur * beta + gamma + vals[row]) % P
                    cur = cur * domain.omega % P
                deltaomega = deltaomega * DELTA % P
            z = [last_z]
            for row in range(1, n):
                z.append(z[row - 1] * modified[row - 1] % P)
            for row in range(n - bf, n):
                z[row] = rng.random_fr()
            last_z = z[n - (bf + 1)]
            blind = rng.random_fr()
            commitment = params.commit_lagrange(z)
            poly = domain.lagrange_to_coeff(z)
            tr.write_point(commitment)
            sets.append({"poly": poly, "coset": domain.coeff_to_extended(poly), "blind": blind})
        perm_sets.append(sets)

    # [TRANSCRIPT-10] lookup products (lookup/prover.rs:182-324)
    for c in range(ncirc):
        for L in permuted[c]:
            denom = [(beta + a) * (gamma + s) % P for a, s in zip(L["permuted_input"], L["permuted_table"])]
            denom = [inv(d) if d else 0 for d in denom]
            prod = [d * (ci + beta) % P * (ct + gamma) % P for d, ci, ct in zip(denom, L["compressed_input"], L["compressed_table"])]
            z = running_product(prod, n, bf, rng)
            L["product_blind"] = rng.random_fr()
            commitment = params.commit_lagrange(z)
            L["product_poly"] = domain.lagrange_to_coeff(z)
            tr.write_point(commitment)

    # [TRANSCRIPT-11] shuffle products (shuffle/prover.rs:97-206)
    shuffles = []
    for c in range(ncirc):
        per_circuit = []
        for _, ins, shs in cs.shuffles:
            a = compress(ins, theta, n, K.fixed_values, advice_values[c], instance_values[c], challenges)
            s = compress(shs, theta, n, K.fixed_values, advice_values[c], instance_values[c], challenges)
            denom = [(gamma + v) % P for v in s]
            denom = [inv(d) if d else 0 for d in denom]
            prod = [d * (gamma + v) % P for d, v in zip(denom, a)]
            z = running_product(prod, n, bf, rng)
            S = {"product_blind": rng.random_fr()}
            commitment = params.commit_lagrange(z)
            S["product_poly"] = domain.lagrange_to_coeff(z)
            tr.write_point(commitment)
            per_circuit.append(S)
        shuffles.append(per_circuit)

    # [TRANSCRIPT-12] vanishing::Argument::commit (vanishing/prover.rs:40-99)
    chunk_size = n // num_threads
    offsets = [i * (chunk_size + 1) for i in range(n % num_threads)]
    if chunk_size != 0:
        first = (n % num_threads) * (chunk_size + 1)
        offsets += [first + i * chunk_size for i in range(num_threads - len(offsets))]
    offsets = offsets[:num_threads]
    thread_rngs = {off: ChaCha20Rng(rng.fill_bytes(32)) for off in offsets}           # seeds drawn in iteration order
    random_poly = [0] * n
    for off, length in parallelize_chunks(n, num_threads):
        r_ = thread_rngs[off]
        for i in range(off, off + length):
            random_poly[i] = r_.random_fr()
    random_blind = rng.random_fr()
    tr.write_point(params.commit(random_poly))

    # 6. advice polys (prover.rs:673-689)
    advice_polys = [[domain.lagrange_to_coeff(v) for v in advice_values[c]] for c in range(ncirc)]

    y = tr.squeeze_challenge()                                                        # [TRANSCRIPT-13]
    h_ext = evaluate_h(K, advice_polys, instance_polys, challenges, y, beta, gamma, theta, permuted, shuffles, perm_sets)

    # [TRANSCRIPT-14] vanishing construct (vanishing/prover.rs:101-155)
    h_coeffs = domain.extended_to_coeff(domain.divide_by_vanishing_poly(h_ext))
    h_pieces = [h_coeffs[i:i + n] for i in range(0, len(h_coeffs) - len(h_coeffs) % n, n)]        # chunks_exact(n)
    h_blinds = [rng.random_fr() for _ in h_pieces]
    for piece in h_pieces:
        tr.write_point(params.commit(piece))

    x = tr.squeeze_challenge()                                                        # [TRANSCRIPT-15]
    xn = pow(x, n, P)
    # [TRANSCRIPT-16] instance evals: skipped, QUERY_INSTANCE == false
    for c in range(ncirc):                                                            # [TRANSCRIPT-17]
        for col, rot in cs.queries["advice"]:
            tr.write_scalar(R.eval_polynomial(advice_polys[c][col], domain.rotate_omega(x, rot)))
    for col, rot in cs.queries["fixed"]:                                              # [TRANSCRIPT-18]
        tr.write_scalar(R.eval_polynomial(K.fixed_polys[col], domain.rotate_omega(x, rot)))
    # [TRANSCRIPT-19] vanishing evaluate (vanishing/prover.rs:157-186)
    h_poly = [0] * n
    h_blind = 0
    for piece, blind in zip(reversed(h_pieces), reversed(h_blinds)):
        h_poly = [(a * xn + b) % P for a, b in zip(h_poly, piece)]
        h_blind = (h_blind * xn + blind) % P
    tr.write_scalar(R.eval_polynomial(random_poly, x))
    for poly in K.permutation_polys:                                                  # [TRANSCRIPT-20] permutation/prover.rs:226-237
        tr.write_scalar(R.eval_polynomial(poly, x))
    x_next, x_prev = domain.rotate_omega(x, 1), domain.rotate_omega(x, -1)
    x_last = domain.rotate_omega(x, -(bf + 1))
    for c in range(ncirc):                                                            # [TRANSCRIPT-21] permutation/prover.rs:241-285
        for i, s in enumerate(perm_sets[c]):
            tr.write_scalar(R.eval_polynomial(s["poly"], x))
            tr.write_scalar(R.eval_polynomial(s["poly"], x_next))
            if i + 1 < len(perm_sets[c]):
                tr.write_scalar(R.eval_polynomial(s["poly"], x_last))
    for c in range(ncirc):                                                            # [TRANSCRIPT-22] lookup/prover.rs:328-356
        for L in permuted[c]:
            for poly, point in ((L["product_poly"], x), (L["product_poly"], x_next), (L["permuted_input_poly"], x),
                                (L["permuted_input_poly"], x_prev), (L["permuted_table_poly"], x)):
                tr.write_scalar(R.eval_polynomial(poly, point))
    for c in range(ncirc):                                                            # [TRANSCRIPT-23] shuffle/prover.rs:209-231
        for S in shuffles[c]:
            tr.write_scalar(R.eval_polynomial(S["product_poly"], x))
            tr.write_scalar(R.eval_polynomial(S["product_poly"], x_next))

    # 13. queries, in the reference's order (prover.rs:840-889).  A query = (point, polynomial): the polynomial OBJECT identifies
    # the commitment (PolynomialPointer compares by address, poly/query.rs:51-55), so each is wrapped once.
    class Poly:
        def __init__(self, coeffs):
            self.coeffs = coeffs

    wrap = {}

    def q(point, coeffs):
        if id(coeffs) not in wrap:
            wrap[id(coeffs)] = Poly(coeffs)
        return (point % P, wrap[id(coeffs)])

    queries = []
    for c in range(ncirc):
        for col, rot in cs.queries["advice"]:
            queries.append(q(domain.rotate_omega(x, rot), advice_polys[c][col]))
        sets = perm_sets[c]                                                           # permutation/prover.rs:287-332
        for s in sets:
            queries.append(q(x, s["poly"]))
            queries.append(q(x_next, s["poly"]))
        for s in list(reversed(sets))[1:]:
            queries.append(q(x_last, s["poly"]))
        for L in permuted[c]:                                                         # lookup/prover.rs:359-399
            queries += [q(x, L["product_poly"]), q(x, L["permuted_input_poly"]), q(x, L["permuted_table_poly"]),
                        q(x_prev, L["permuted_input_poly"]), q(x_next, L["product_poly"])]
        for S in shuffles[c]:                                                         # shuffle/prover.rs:233-254
            queries += [q(x, S["product_poly"]), q(x_next, S["product_poly"])]
    for col, rot in cs.queries["fixed"]:
        queries.append(q(domain.rotate_omega(x, rot), K.fixed_polys[col]))
    for poly in K.permutation_polys:                                                  # permutation/prover.rs:215-224
        queries.append(q(x, poly))
    queries += [q(x, h_poly), q(x, random_poly)]                                      # vanishing/prover.rs:188-205

    if multiopen == "shplonk":
        shplonk_open(params, tr, queries)
    elif multiopen == "gwc":
        gwc_open(params, tr, queries)
    else:
        raise ValueError(multiopen)
    return tr.finalize()


def permute_expression_pair(input_expression, table_expression, n, bf, rng):
    """lookup/prover.rs:410-494."""
    usable = n - (bf + 1)
    permuted_input = sorted(input_expression[:usable])      # `Fr: Ord` = order of the canonical integers
    leftover = {}
    for v in table_expression[:usable]:
        leftover[v] = leftover.get(v, 0) + 1
    permuted_table = [0] * usable
    repeated = []
    for row, v in enumerate(permuted_input):
        if row == 0 or v != permuted_input[row - 1]:
            permuted_table[row] = v
            if v in leftover:
                assert leftover[v] > 0
                leftover[v] -= 1
            else:
                raise ValueError("ConstraintSystemFailure")
        else:
            repeated.append(row)
    for v in sorted(leftover):                              # BTreeMap iteration: ascending
        for _ in range(leftover[v]):
            permuted_table[repeated.pop()] = v
    assert not repeated
    permuted_input += [rng.random_fr() for _ in range(bf + 1)]
    permuted_table += [rng.random_fr() for _ in range(bf + 1)]
    return permuted_input, permuted_table


def evaluate_h(K, advice_polys, instance_polys, challenges, y, beta, gamma, theta, lookups, shuffles, perm_sets):
    """plonk/evaluation.rs:317-623: the quotient numerator on the extended coset, constraint after constraint folded with y."""
    cs, domain = K.cs, K.domain
    size = domain.extended_len()
    rot_scale = 1 << (domain.extended_k - domain.k)
    values = [0] * size
    fixed = K.fixed_cosets
    l0, l_last, l_active = K.l0, K.l_last, K.l_active_row
    bf = cs.blinding_factors()
    for c in range(len(advice_polys)):
        advice = [domain.coeff_to_extended(p) for p in advice_polys[c]]
        instance = [domain.coeff_to_extended(p) for p in instance_polys[c]]
        ev = lambda e, idx: _eval_at(e, idx, size, rot_scale, fixed, advice, instance, challenges)
        # custom gates: Horner(PreviousValue, gate polys, y)  (evaluation.rs:225-236, 373-396)
        for idx in range(size):
            v = values[idx]
            for _, poly in cs.gates:
                v = (v * y + ev(poly, idx)) % P
            values[idx] = v
        # permutations (evaluation.rs:397-493)
        sets = perm_sets[c]
        if sets:
            chunk_len = K.cs_degree - 2
            delta_start = beta * R.FR_ZETA % P
            src = {"advice": advice, "fixed": fixed, "instance": instance}
            first, last = sets[0]["coset"], sets[-1]["coset"]
            beta_term = 1
            for idx in range(size):
                r_next = (idx + rot_scale) % size
                r_last = (idx - (bf + 1) * rot_scale) % size
                v = values[idx]
                v = (v * y + (1 - first[idx]) * l0[idx]) % P
                v = (v * y + (last[idx] * last[idx] - last[idx]) * l_last[idx]) % P
                for si in range(1, len(sets)):
                    v = (v * y + (sets[si]["coset"][idx] - sets[si - 1]["coset"][r_last]) * l0[idx]) % P
                current_delta = delta_start * beta_term % P
                for si, s in enumerate(sets):
                    columns = cs.permutation_columns[si * chunk_len:(si + 1) * chunk_len]
                    cosets = K.permutation_cosets[si * chunk_len:(si + 1) * chunk_len]
                    left = s["coset"][r_next]
                    for (kind, index), perm in zip(columns, cosets):
                        left = left * (src[kind][index][idx] + beta * perm[idx] + gamma) % P
                    right = s["coset"][idx]
                    for kind, index in columns:
                        right = right * (src[kind][index][idx] + current_delta + gamma) % P
                        current_delta = current_delta * DELTA % P
                    v = (v * y + (left - right) * l_active[idx]) % P
                values[idx] = v
                beta_term = beta_term * domain.extended_omega % P
        # lookups (evaluation.rs:496-565)
        for (_, ins, tabs), L in zip(cs.lookups, lookups[c]):
            product = domain.coeff_to_extended(L["product_poly"])
            p_in = domain.coeff_to_extended(L["permuted_input_poly"])
            p_tab = domain.coeff_to_extended(L["permuted_table_poly"])
            for idx in range(size):
                a = 0
                for e in ins:
                    a = (a * theta + ev(e, idx)) % P
                s = 0
                for e in tabs:
                    s = (s * theta + ev(e, idx)) % P
                table_value = (a + beta) * (s + gamma) % P
                r_next, r_prev = (idx + rot_scale) % size, (idx - rot_scale) % size
                a_minus_s = (p_in[idx] - p_tab[idx]) % P
                v = values[idx]
                v = (v * y + (1 - product[idx]) * l0[idx]) % P
                v = (v * y + (product[idx] * product[idx] - product[idx]) * l_last[idx]) % P
                v = (v * y + (product[r_next] * (p_in[idx] + beta) % P * (p_tab[idx] + gamma) - product[idx] * table_value) * l_active[idx]) % P
                v = (v * y + a_minus_s * l0[idx]) % P
                v = (v * y + a_minus_s * (p_in[idx] - p_in[r_prev]) % P * l_active[idx]) % P
                values[idx] = v
        # shuffles (evaluation.rs:568-620)
        for (_, ins, shs), S in zip(cs.shuffles, shuffles[c]):
            product = domain.coeff_to_extended(S["product_poly"])
            for idx in range(size):
                a = 0
                for e in ins:
                    a = (a * theta + ev(e, idx)) % P
                s = 0
                for e in shs:
                    s = (s * theta + ev(e, idx)) % P
                input_value, shuffle_value = (a + gamma) % P, (s + gamma) % P
                r_next = (idx + rot_scale) % size
                v = values[idx]
                v = (v * y + (1 - product[idx]) * l0[idx]) % P
                v = (v * y + (product[idx] * product[idx] - product[idx]) * l_last[idx]) % P
                v = (v * y + l_active[idx] * (product[r_next] * shuffle_value - product[idx] * input_value)) % P
                values[idx] = v
    return values


# =====================================================================================================================
# multi-open arguments
# =====================================================================================================================
def shplonk_open(params: Params, tr: Blake2bWrite, queries):
    """shplonk.rs:46-137 (construct_intermediate_sets) + shplonk/prover.rs:121-305."""
    n = params.n
    y = tr.squeeze_challenge()
    # commitment -> set of points, in first-seen order of the commitments; points ordered as a BTreeSet<Fr>
    commitment_points = []                                  # [(poly object, set of points)]
    super_points = set()
    for point, poly in queries:
        super_points.add(point)
        for cp in commitment_points:
            if cp[0] is poly:
                cp[1].add(point)
                break
        else:
            commitment_points.append((poly, {point}))
    rotation_sets = []                                      # [(sorted points, [poly objects])] in first-seen order of the point sets
    for poly, pts in commitment_points:
        key = tuple(sorted(pts))
        for rs in rotation_sets:
            if rs[0] == key:
                rs[1].append(poly)
                break
        else:
            rotation_sets.append((key, [poly]))
    # extend every commitment with its low-degree equivalent r(X): the interpolant of its evaluations on the set's points
    ext_sets = []
    for points, polys in rotation_sets:
        entries = []
        for poly in polys:
            evals = [R.eval_polynomial(poly.coeffs, pt) for pt in points]
            entries.append((poly, lagrange_interpolate(list(points), evals)))
        ext_sets.append((list(points), entries))
    v = tr.squeeze_challenge()

    def quotient_contribution(points, entries):             # prover.rs:143-176
        acc, power = None, 1
        for poly, low in entries:
            num = list(poly.coeffs)
            for i, c in enumerate(low):
                num[i] = (num[i] - c) % P
            num = [c * power % P for c in num]
            acc = num if acc is None else [(a + b) % P for a, b in zip(acc, num)]
            power = power * y % P
        for pt in points:                                   # div_by_vanishing: one kate_division per root
            acc = R.kate_division(acc, pt)
        return acc + [0] * (n - len(acc))

    h_x, power = None, 1
    for points, entries in ext_sets:
        qpoly = [c * power % P for c in quotient_contribution(points, entries)]
        h_x = qpoly if h_x is None else [(a + b) % P for a, b in zip(h_x, qpoly)]
        power = power * v % P
    tr.write_point(params.commit(h_x))
    u = tr.squeeze_challenge()

    l_x, power_v, z_diffs = None, 1, []
    for points, entries in ext_sets:                        # linearisation_contribution, prover.rs:219-257
        diffs = [pt for pt in sorted(super_points) if pt not in points]
        z_i = vanishing_eval(diffs, u)
        z_diffs.append(z_i)
        inner, power_y = None, 1
        for poly, low in entries:
            r_eval = R.eval_polynomial(low, u)
            contrib = list(poly.coeffs)
            contrib[0] = (contrib[0] - r_eval) % P          # `Polynomial - F` subtracts from the constant term (poly.rs:268-276)
            contrib = [c * power_y % P for c in contrib]
            inner = contrib if inner is None else [(a + b) % P for a, b in zip(inner, contrib)]
            power_y = power_y * y % P
        inner = [c * z_i % P * power_v % P for c in inner]
        l_x = inner if l_x is None else [(a + b) % P for a, b in zip(l_x, inner)]
        power_v = power_v * v % P
    zt_eval = vanishing_eval(sorted(super_points), u)
    l_x = [(a - b * zt_eval) % P for a, b in zip(l_x, h_x)]
    assert R.eval_polynomial(l_x, u) == 0                   # the debug_assertions sanity check, prover.rs:281-286
    h2 = R.kate_division(l_x, u)
    z0_inv = inv(z_diffs[0])
    h2 = [c * z0_inv % P for c in h2]
    tr.write_point(params.commit(h2))


def gwc_open(params: Params, tr: Blake2bWrite, queries):
    """gwc.rs:26-50 + gwc/prover.rs:45-90: one witness polynomial per distinct point, in first-seen order."""
    v = tr.squeeze_challenge()
    by_point = []
    for point, poly in queries:
        for entry in by_point:
            if entry[0] == point:
                entry[1].append(poly)
                break
        else:
            by_point.append((point, [poly]))
    for z, polys in by_point:
        poly_batch, eval_batch, power = None, 0, 1
        for poly in polys:
            scaled = [c * power % P for c in poly.coeffs]
            poly_batch = scaled if poly_batch is None else [(a + b) % P for a, b in zip(poly_batch, scaled)]
            eval_batch = (eval_batch + R.eval_polynomial(poly.coeffs, z) * power) % P
            power = power * v % P
        poly_batch[0] = (poly_batch[0] - eval_batch) % P
        tr.write_point(params.commit(R.kate_division(poly_batch, z)))
