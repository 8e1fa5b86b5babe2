"""CPU tests of the prover mirror (yet-another-halo2-fork_b200/plonk) over the python-int oracle ops: the host logic of
halo2_backend/src/plonk/{prover,keygen}.rs, the four argument provers, the Blake2b transcript and SHPLONK, end to end:
create_proof -> bytes -> verify_proof restatement (oracle/plonk_verifier.py), as halo2_proofs/tests/frontend_backend_split.rs:595-643
does with the real crates."""
import hashlib

import numpy as np
import pytest

import h2b200
from h2b200 import fields as F
from h2b200 import plonk
from h2b200.plonk import arguments as A
from h2b200.plonk import example, rng as RNG, shplonk, transcript as T
from oracle import plonk_ops, plonk_verifier
from oracle import pyref as R

K = 6
SECRET = 0x1234567


@pytest.fixture(scope="module")
def setup():
    circ = example.MyCircuit(K, 42)
    cc = circ.compile()
    cs = plonk.circuit.ConstraintSystemBack(cc.cs)
    ops = plonk_ops.RefOps(K, SECRET, R.EvaluationDomain(cs.degree(), K))
    vk = plonk.keygen_vk(ops, cc)
    pk = plonk.keygen_pk(ops, vk, cc)
    return circ, ops, vk, pk


def _prove(circ, ops, pk, rng, witness=None, num_threads=4):
    tr = plonk.Blake2bWrite()
    plonk.create_proof(ops, pk, circ.instances(), witness or circ.witness, rng, tr, num_threads=num_threads)
    return tr.finalize()


def test_constraint_system_shape(setup):
    """degree 6 from the lookup (plonk/circuit.rs:354-373), 5 blinding factors, 4 permutation columns in
    one set (chunk_len = degree - 2), queries collected gates -> lookups -> shuffles -> permutation columns (keygen.rs:305-341)."""
    _, _, vk, _ = setup
    cs = vk.cs
    assert cs.degree() == 6 and cs.blinding_factors() == 5 and cs.minimum_rows() == 8
    assert cs.advice_queries == [(0, 0), (1, 0), (2, 0), (0, 1), (3, 0)]
    assert cs.instance_queries == [(0, 0)]
    assert cs.fixed_queries[0] == (5, 0) and (0, 0) in cs.fixed_queries and len(cs.fixed_queries) == 7   # s_instance is never queried
    assert cs.phases() == [0, 1]
    assert vk.domain.extended_k == K + 3 and vk.domain.quotient_poly_degree == 5


def test_proof_verifies_and_is_deterministic(setup):
    circ, ops, vk, pk = setup
    proof = _prove(circ, ops, pk, RNG.OneNg())
    # 4 advice + 2 + 1 + 1 + 1 + 1 + 5 points, 5 + 7 + 1 + 4 + 2 + 5 + 2 scalars, 2 opening points
    assert len(proof) == 32 * (4 + 2 + 1 + 1 + 1 + 1 + 5 + 5 + 7 + 1 + 4 + 2 + 5 + 2 + 2)
    assert plonk_verifier.verify_proof(vk, circ.instances(), proof, SECRET, 1 << K)
    assert proof == _prove(circ, ops, pk, RNG.OneNg())
    assert proof != _prove(circ, ops, pk, RNG.OneNg(), num_threads=3), "the thread count seeds the random polynomial"
    other = _prove(circ, ops, pk, RNG.ChaCha20Rng(bytes(range(32))))
    assert other != proof and plonk_verifier.verify_proof(vk, circ.instances(), other, SECRET, 1 << K)


def test_verifier_rejects(setup):
    circ, ops, vk, pk = setup
    proof = _prove(circ, ops, pk, RNG.OneNg())
    for pos in (40, 32 * 12 + 3, len(proof) - 5):
        bad = bytearray(proof)
        bad[pos] ^= 1
        try:
            assert not plonk_verifier.verify_proof(vk, circ.instances(), bytes(bad), SECRET, 1 << K)
        except ValueError:
            pass                                              # not a curve point / not a canonical scalar
    inst = circ.instances()
    inst[0][2] += 1
    assert not plonk_verifier.verify_proof(vk, inst, proof, SECRET, 1 << K)

    def broken(column, row, value):
        def w(phase, ch):
            cols = circ.witness(phase, ch)
            if phase == 0:
                cols[column] = cols[column].copy()
                cols[column][row] = F.fr_mont(value)
            return cols
        return w
    for column, row, value in ((2, 0, 5), (1, 36, 7), (0, 8, 3)):    # gate, shuffle, copy constraint
        bad = _prove(circ, ops, pk, RNG.OneNg(), witness=broken(column, row, value))
        assert not plonk_verifier.verify_proof(vk, circ.instances(), bad, SECRET, 1 << K)
    with pytest.raises(A.ConstraintSystemFailure):                    # lookup/prover.rs:440 -> Error::ConstraintSystemFailure
        _prove(circ, ops, pk, RNG.OneNg(), witness=broken(0, 20, 12345))


def test_commit_phase_errors(setup):
    """prover.rs:343-390."""
    circ, ops, _, pk = setup
    p = plonk.ProverSingle(ops, pk, circ.instances(), RNG.OneNg(), plonk.Blake2bWrite())
    with pytest.raises(plonk.prover.ProverError, match="Committing invalid phase"):
        p.commit_phase(1, circ.witness(1, {0: 1}))
    with pytest.raises(plonk.prover.ProverError, match="expected no advice column"):
        w = circ.witness(0, {})
        w[3] = w[0]
        p.commit_phase(0, w)
    with pytest.raises(plonk.prover.ProverError, match="to have length"):
        w = circ.witness(0, {})
        w[0] = w[0][:-1]
        p.commit_phase(0, w)
    with pytest.raises(plonk.prover.ProverError, match="InvalidInstances"):
        plonk.ProverSingle(ops, pk, [], RNG.OneNg(), plonk.Blake2bWrite())


def test_transcript_protocol():
    """transcript.rs:308-322, 358-418: the byte protocol, restated with hashlib directly."""
    tw = T.Blake2bWrite()
    g = (1, 2)
    tw.common_scalar(5)
    tw.write_point(g)
    tw.write_scalar(7)
    c1 = tw.squeeze_challenge_scalar()
    c2 = tw.squeeze_challenge_scalar()
    h = hashlib.blake2b(digest_size=64, person=b"Halo2-Transcript")
    h.update(b"\x02" + (5).to_bytes(32, "little"))
    h.update(b"\x01" + (1).to_bytes(32, "little") + (2).to_bytes(32, "little"))
    h.update(b"\x02" + (7).to_bytes(32, "little"))
    h.update(b"\x00")
    assert c1 == int.from_bytes(h.copy().digest(), "little") % F.R_MOD
    h.update(b"\x00")
    assert c2 == int.from_bytes(h.copy().digest(), "little") % F.R_MOD and c1 != c2
    proof = tw.finalize()
    assert proof == (1).to_bytes(32, "little") + (7).to_bytes(32, "little")      # y = 2 is even: no sign bit
    tr = T.Blake2bRead(proof)
    tr.common_scalar(5)
    assert tr.read_point() == g and tr.read_scalar() == 7
    assert tr.squeeze_challenge_scalar() == c1
    with pytest.raises(IOError):
        tw.write_point(None)                                                      # transcript.rs:401-406


def test_point_encoding_round_trip():
    pts = [R.g1_mul(R.G1_GEN, s) for s in (1, 2, 3, 0xDEADBEEF, F.R_MOD - 1)]
    for p in pts + [None]:
        b = T.g1_to_bytes(p)
        assert len(b) == 32 and T.g1_from_bytes(b) == p
    neg = (pts[1][0], F.Q_MOD - pts[1][1])
    assert T.g1_to_bytes(neg)[:31] == T.g1_to_bytes(pts[1])[:31] and T.g1_to_bytes(neg) != T.g1_to_bytes(pts[1])
    with pytest.raises(ValueError):
        T.g1_from_bytes((4).to_bytes(32, "little"))                               # x = 4: x^3 + 3 = 67 is not a square mod q
    with pytest.raises(ValueError):
        T.fr_from_repr(F.R_MOD.to_bytes(32, "little"))


def test_chacha20_stream():
    """ChaCha20 with the all-zero key, block counter 0, zero nonce: the published keystream (RFC 7539 test vector #1 family)."""
    ks = RNG.chacha20_blocks(bytes(32), 0, 2)
    assert ks[0].astype("<u4").tobytes().hex().startswith("76b8e0ada0f13d90405d6ae55386bd28bdd219b8a08ded1aa836efcc8b770dc7")
    a = RNG.ChaCha20Rng(bytes(32))
    words = [a.next_u64() for _ in range(72)]                                     # crosses two 64-word refills
    b = RNG.ChaCha20Rng(bytes(32))
    assert [int(x) for x in b.random_fr_limbs(9).reshape(-1)] == words
    c = RNG.ChaCha20Rng(bytes(32))
    assert c.fill_bytes(13) + c.fill_bytes(3) != b"" and c.index == 4 + 1         # 13 bytes consume 4 words, 3 bytes 1 word
    d = RNG.ChaCha20Rng(bytes(32))
    assert d.next_u32() == words[0] & 0xFFFFFFFF and d.next_u64() == (words[0] >> 32) | ((words[1] & 0xFFFFFFFF) << 32)
    one = RNG.OneNg()
    assert one.random_fr() == sum(((1 << 32) | 1) << (64 * i) for i in range(8)) % F.R_MOD and one.fill_bytes(5) == b"\x01\0\0\0\x01"


def test_permute_expression_pair():
    """lookup/prover.rs:398-494 against a literal python restatement (sorted input, BTreeMap of leftovers, pop from the back)."""
    rng = np.random.default_rng(5)
    for _ in range(20):
        u = int(rng.integers(1, 40))
        table = [int(x) for x in rng.integers(0, 12, size=u)]
        inp = [table[int(i)] for i in rng.integers(0, u, size=u)]
        srt = sorted(inp)
        leftover = {}
        for t in table:
            leftover[t] = leftover.get(t, 0) + 1
        out, repeated = [0] * u, []
        for row, v in enumerate(srt):
            if row == 0 or v != srt[row - 1]:
                out[row] = v
                leftover[v] -= 1
            else:
                repeated.append(row)
        for coeff in sorted(leftover):
            for _ in range(leftover[coeff]):
                out[repeated.pop()] = coeff
        limbs = lambda xs: np.array([[x, 0, 0, x >> 1] for x in xs], dtype=np.uint64)   # two significant limbs
        pi, pt = A.permute_expression_pair(limbs(inp), limbs(table))
        assert [int(r[0]) for r in pi] == srt and [int(r[0]) for r in pt] == out
    with pytest.raises(A.ConstraintSystemFailure):
        A.permute_expression_pair(np.array([[1, 0, 0, 0]], dtype=np.uint64), np.array([[2, 0, 0, 0]], dtype=np.uint64))


def test_parallelize_chunks_and_interpolation():
    assert A.parallelize_chunks(10, 4) == [(0, 3), (3, 3), (6, 2), (8, 2)]        # arithmetic.rs:148-170
    assert A.parallelize_chunks(3, 8) == [(0, 1), (1, 1), (2, 1)]
    assert A.parallelize_chunks(16, 4) == [(0, 4), (4, 4), (8, 4), (12, 4)]
    pts, evs = [3, 7, 11, 20], [5, 1, 9, 2]
    poly = shplonk.lagrange_interpolate(pts, evs)                                 # arithmetic.rs:177-232 and its test :263-280
    assert len(poly) == 4 and [shplonk.eval_small(poly, p) for p in pts] == evs
    assert shplonk.lagrange_interpolate([4], [9]) == [9]
    assert shplonk.evaluate_vanishing_polynomial([1, 2], 5) == 12


def test_intermediate_sets_grouping():
    """shplonk.rs:46-137: commitments with the same point set share a rotation set, in first-seen order; points ascend."""
    q = [(5, "a", 1), (9, "a", 2), (5, "b", 3), (9, "b", 4), (5, "c", 5), (9, "a", 2)]
    sets, points = shplonk.construct_intermediate_sets(q, lambda x: x[1], lambda x: x[2])
    assert points == [5, 9] and len(sets) == 2
    assert sets[0].points == [5, 9] and [c[0][1] for c in sets[0].commitments] == ["a", "b"]
    assert [c[1] for c in sets[0].commitments] == [[1, 2], [3, 4]]
    assert sets[1].points == [5] and sets[1].commitments[0][1] == [5]


@pytest.mark.parametrize("minimum_degree,n_sets", [(5, 1), (None, 3)])
def test_standard_plonk_circuit(minimum_degree, n_sets):
    """halo2_proofs/benches/plonk.rs circuit (BASELINE configs[0] shape): no lookups / shuffles / instance; with the degree left
    at 3 the three permuted columns need three chained grand products (permutation/prover.rs:96-190, verifier.rs:126-143)."""
    k = 5
    circ = example.StandardPlonk(k, 7, minimum_degree)
    cc = circ.compile()
    cs = plonk.circuit.ConstraintSystemBack(cc.cs)
    assert cs.degree() == (minimum_degree or 3) and cs.blinding_factors() == 5
    ops = plonk_ops.RefOps(k, SECRET, R.EvaluationDomain(cs.degree(), k))
    vk = plonk.keygen_vk(ops, cc)
    pk = plonk.keygen_pk(ops, vk, cc)
    proof = _prove(circ, ops, pk, RNG.ChaCha20Rng(bytes(32)), num_threads=2)
    n_points = 3 + n_sets + 1 + (cs.degree() - 1) + 2
    n_scalars = 3 + 4 + 1 + 3 + (3 * n_sets - 1)
    assert len(proof) == 32 * (n_points + n_scalars)
    assert plonk_verifier.verify_proof(vk, [], proof, SECRET, 1 << k)

    def broken(phase, ch):
        cols = circ.witness(phase, ch)
        cols[1][3] = F.fr_mont(11)            # b at an "add" row: breaks the gate there and the copy b1 == c0
        return cols
    assert not plonk_verifier.verify_proof(vk, [], _prove(circ, ops, pk, RNG.OneNg(), witness=broken), SECRET, 1 << k)


def test_gwc_multiopen(setup):
    """ProverGWC / VerifierGWC (multiopen/gwc/prover.rs:45-90, verifier.rs:40-123): one witness commitment per distinct point."""
    circ, ops, vk, pk = setup
    tr = plonk.Blake2bWrite()
    plonk.create_proof(ops, pk, circ.instances(), circ.witness, RNG.OneNg(), tr, num_threads=4, multiopen="gwc")
    proof = tr.finalize()
    shplonk_proof = _prove(circ, ops, pk, RNG.OneNg())
    assert proof[:-3 * 32] == shplonk_proof[:-2 * 32]                  # same transcript up to the opening; 3 points: x, x*w, x/w
    assert plonk_verifier.verify_proof(vk, circ.instances(), proof, SECRET, 1 << K, multiopen="gwc")
    bad = bytearray(proof)
    bad[200] ^= 4
    try:
        assert not plonk_verifier.verify_proof(vk, circ.instances(), bytes(bad), SECRET, 1 << K, multiopen="gwc")
    except ValueError:
        pass


def _golden():
    import importlib.util
    import json
    import os
    here = os.path.dirname(__file__)
    spec = importlib.util.spec_from_file_location("make_golden_proofs", os.path.join(here, "golden", "make_golden_proofs.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    return gen, json.load(open(os.path.join(here, "golden", "proofs.json")))


def test_golden_proofs_regenerate():
    """tests/golden/proofs.json is what the independent prover oracle (oracle/plonk_prover.py) writes today, the package's prover
    over the python-int ops reproduces every stored proof byte for byte (two implementations of prover.rs that share no code), and
    the restated verifier accepts the stored bytes."""
    gen, gold = _golden()
    assert [c["name"] for c in gold["proofs"]] == [c["name"] for c in gen.CASES]
    for case in gold["proofs"]:
        if case["k"] <= 6:                        # the oracle at k = 8 takes several seconds per proof: checked on the GPU leg
            proof, keys, _ = gen.oracle_prove(case)
            assert proof.hex() == case["proof"] and hex(keys.transcript_repr) == case["transcript_repr"], case["name"]
        proof, vk, circ = gen.prove(case, lambda k, cs: plonk_ops.RefOps(k, SECRET, R.EvaluationDomain(cs.degree(), k)))
        assert proof.hex() == case["proof"] and hashlib.sha256(proof).hexdigest() == case["sha256"], case["name"]
        assert hex(vk.transcript_repr) == case["transcript_repr"]
        assert plonk_verifier.verify_proof(vk, circ.instances(), bytes.fromhex(case["proof"]), SECRET, 1 << case["k"],
                                           multiopen=case["multiopen"])


def test_pinned_vk_string_matches_the_oracle(setup):
    """`format!("{:?}", vk.pinned())` as the package writes it (plonk/pinned.py) == the oracle's Debug model, whose structure is
    pinned by the reference's own golden text (tests/test_oracle_prover.py); and keygen_vk takes a value from a Rust run verbatim."""
    import _plain_circuit as PC
    from h2b200.plonk import pinned
    from oracle import plonk_prover as PP
    circ, ops, vk, _ = setup
    keys = PC.oracle_keys(circ, K, SECRET)
    text = PP.render(PP.pinned_vk_node(keys.cs, keys.domain, keys.fixed_commitments, keys.permutation_commitments))
    assert pinned.verification_key(vk) == text
    assert text.startswith('PinnedVerificationKey { base_modulus: "0x30644e72e131a029b85045b68181585d97816a916871ca8d3c208c16d87cfd47", '
                           'scalar_modulus: "0x30644e72e131a029b85045b68181585d2833e84879b9709143e1f593f0000001", '
                           'domain: PinnedEvaluationDomain { k: 6, extended_k: 9, omega: 0x')
    assert "num_challenges: 1, advice_column_phase: [0, 0, 0, 1], challenge_phase: [0], gates: [Product(Var(Query(QueryBack {" in text
    assert "Var(Challenge(ChallengeMid { index: 0, phase: 0 }))" in text and 'shuffles: [Argument { name: "shuffle.0", ' in text
    assert text.endswith("] } }") and "minimum_degree: None }, fixed_commitments: [(0x" in text
    assert vk.transcript_repr == keys.transcript_repr
    assert plonk.keygen_vk(ops, circ.compile(), transcript_repr=12345).transcript_repr == 12345


@pytest.mark.parametrize("fmt", ["Processed", "RawBytes", "RawBytesUnchecked"])
def test_key_serialization_round_trip(setup, fmt):
    """halo2_proofs/tests/serialization.rs:131-195: write the proving key, read it back, prove with it, verify; plus the byte
    lengths of plonk.rs:131-139, 259-272 and the verifying-key round trip (plonk.rs:57-118)."""
    from h2b200.plonk import serialization as S
    circ, ops, vk, pk = setup
    vk_bytes = S.vk_write(vk, fmt)
    assert len(vk_bytes) == S.vk_bytes_length(vk, fmt) and vk_bytes[0] == 0x04 and vk_bytes[1] == K
    vk2 = S.vk_read(vk_bytes, fmt, vk.cs, lambda degree, k: R.EvaluationDomain(degree, k))
    assert vk2.fixed_commitments == vk.fixed_commitments and vk2.permutation_commitments == vk.permutation_commitments
    assert vk2.transcript_repr == vk.transcript_repr
    pk_bytes = S.pk_write(ops, pk, fmt)
    assert len(pk_bytes) == S.pk_bytes_length(ops, pk, fmt)
    pk2 = S.pk_read(ops, pk_bytes, fmt, vk.cs)
    assert pk2.l0 == pk.l0 and pk2.l_active_row == pk.l_active_row and pk2.fixed_cosets == pk.fixed_cosets
    assert pk2.permutation.polys == pk.permutation.polys and S.pk_write(ops, pk2, fmt) == pk_bytes
    proof = _prove(circ, ops, pk2, RNG.OneNg())
    assert proof == _prove(circ, ops, pk, RNG.OneNg())
    assert plonk_verifier.verify_proof(vk2, circ.instances(), proof, SECRET, 1 << K)


def test_coset_only_key_serialises_like_the_full_key(setup):
    """keygen_pk(extended=False) keeps no extended-domain vectors; writing it rebuilds them, byte for byte what the full key
    writes, and a key that was read back carries the coefficient forms of l_0 / l_last / l_active the coset prover needs."""
    from h2b200.plonk import serialization as S
    circ, ops, vk, pk = setup
    lean = plonk.keygen_pk(ops, vk, circ.compile(), extended=False)
    assert lean.l0 is None and lean.fixed_cosets == []
    blob = S.pk_write(ops, pk, "RawBytes")
    assert S.pk_write(ops, lean, "RawBytes") == blob and S.pk_bytes_length(ops, lean, "RawBytes") == len(blob)
    pk2 = S.pk_read(ops, blob, "RawBytes", vk.cs)
    assert pk2.l0_poly == pk.l0_poly and pk2.l_last_poly == pk.l_last_poly and pk2.l_active_poly == pk.l_active_poly
    assert pk2.l0_poly is not None


def test_key_serialization_rejects():
    from h2b200.plonk import serialization as S
    with pytest.raises(IOError, match="version"):
        S.vk_read(bytes([3, 6, 0, 0, 0, 0]), "RawBytes", None, None)
    r = S._Reader(b"\x00\x00\x00\x01" + F.R_MOD.to_bytes(32, "little"))
    with pytest.raises(IOError, match="prime field"):
        S.read_polynomial(None, r, "Processed")
    g = (1, 2)
    for fmt in S.FORMATS:
        assert S.read_point(S._Reader(S.write_point(g, fmt)), fmt) == g and len(S.write_point(g, fmt)) == S.point_byte_length(fmt)
        assert S.read_point(S._Reader(S.write_point(None, fmt)), fmt) is None
    bad = F.fq_mont(1).tobytes() + F.fq_mont(3).tobytes()                  # (1, 3) is not on the curve
    with pytest.raises(IOError, match="curve"):
        S.read_point(S._Reader(bad), "RawBytes")
    assert S.read_point(S._Reader(bad), "RawBytesUnchecked") == (1, 3)


def test_batch_proving_two_circuits(setup):
    """`Prover` over two instances of the circuit in one transcript (prover.rs:130-175, create_proof with `circuits: &[C]`,
    halo2_proofs/src/plonk/prover.rs:60-123; verifier.rs `num_proofs` loops)."""
    _, ops, vk, pk = setup
    circuits = [example.MyCircuit(K, 42), example.MyCircuit(K, 1000)]
    instances = [c.instances() for c in circuits]
    tr = plonk.Blake2bWrite()
    plonk.create_proof_batch(ops, pk, instances, [c.witness for c in circuits], RNG.ChaCha20Rng(bytes(32)), tr, num_threads=2)
    proof = tr.finalize()
    single = len(_prove(circuits[0], ops, pk, RNG.OneNg()))
    # per circuit: 4 advice + 2 + 1 + 1 + 1 points and 5 + 2 + 5 + 2 scalars more; the rest is shared
    assert len(proof) == single + 32 * (9 + 14)
    assert plonk_verifier.verify_proof_batch(vk, instances, proof, SECRET, 1 << K)
    assert not plonk_verifier.verify_proof_batch(vk, instances[::-1], proof, SECRET, 1 << K)
    with pytest.raises(plonk.prover.ProverError, match="witness.len"):
        p = plonk.Prover(ops, pk, instances, RNG.OneNg(), plonk.Blake2bWrite())
        p.commit_phase(0, [circuits[0].witness(0, {})])
