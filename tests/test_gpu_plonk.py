"""GPU parity of the whole proving pipeline: the package's prover over DeviceOps (libh2b200.so: MSM, NTT, quotient, scans) must emit
the same proof bytes as the INDEPENDENT prover oracle (oracle/plonk_prover.py: python ints, own keygen / transcript / RNG /
multi-open, no code shared with the package) given the same RNG -- north_star's "byte-identical proof transcripts given a fixed
blinding RNG", with the CPU restatement standing where the Rust binary cannot run -- and the proof must pass the verify_proof
restatement."""
import numpy as np
import pytest

import h2b200
from h2b200 import plonk
from h2b200.plonk import example, rng as RNG
from h2b200.plonk.ops import DeviceOps
from oracle import cref as C
from oracle import plonk_ops, plonk_verifier
from oracle import pyref as R
import _plain_circuit as PC

pytestmark = pytest.mark.gpu
SECRET = 0x1234567


def _device_setup(ctx, k, circ):
    cc = circ.compile()
    cs = plonk.circuit.ConstraintSystemBack(cc.cs)
    ref_dom = R.EvaluationDomain(cs.degree(), k)
    g, gl = C.kzg_setup(k, C.mont_from_int(SECRET, R.R_MOD), C.mont_from_int(ref_dom.omega, R.R_MOD))
    engine = h2b200.B200Engine(ctx)
    params = h2b200.ParamsKZG(k, g, gl, engine)
    ops = DeviceOps(params, engine, h2b200.EvaluationDomain(cs.degree(), k, ctx))
    vk = plonk.keygen_vk(ops, cc)
    pk = plonk.keygen_pk(ops, vk, cc)
    return cc, cs, ref_dom, ops, vk, pk


def _prove(circ, ops, pk, rng, num_threads=4):
    tr = plonk.Blake2bWrite()
    plonk.create_proof(ops, pk, circ.instances(), circ.witness, rng, tr, num_threads=num_threads)
    return tr.finalize()


def _prove_with(circuits, ops, pk, rng_spec, threads, multiopen="shplonk"):
    tr = plonk.Blake2bWrite()
    if isinstance(circuits, (list, tuple)):
        plonk.create_proof_batch(ops, pk, [c.instances() for c in circuits], [c.witness for c in circuits], PC.package_rng(rng_spec),
                                 tr, num_threads=threads, multiopen=multiopen)
    else:
        plonk.create_proof(ops, pk, circuits.instances(), circuits.witness, PC.package_rng(rng_spec), tr, num_threads=threads,
                           multiopen=multiopen)
    return tr.finalize()


@pytest.mark.parametrize("name,k", [("StandardPlonk", 5), ("MyCircuit", 6), ("MyCircuit", 8)])
def test_proof_bytes_match_the_cpu_restatement(ctx, name, k):
    """k = 5, 6, 8; SHPLONK and GWC; OneNg and ChaCha20; two thread counts each (the vanishing argument's per-thread seeds,
    vanishing/prover.rs:48-77): the device proof equals the independent oracle's, key material included."""
    circ = example.MyCircuit(k, 42) if name == "MyCircuit" else example.StandardPlonk(k, 7, 5)
    cc, cs, ref_dom, ops, vk, pk = _device_setup(ctx, k, circ)
    keys = PC.oracle_keys(circ, k, SECRET)
    assert keys.fixed_commitments == vk.fixed_commitments and keys.permutation_commitments == vk.permutation_commitments
    assert keys.transcript_repr == vk.transcript_repr
    before = ctx.launch_count()
    for rng_spec, threads, multiopen in (("OneNg", 4, "shplonk"), (PC.SEED_0_31, 3, "shplonk"), (PC.SEED_0_31, 4, "gwc"),
                                         ("OneNg", 3, "gwc")):
        proof = _prove_with(circ, ops, pk, rng_spec, threads, multiopen)
        assert proof == PC.oracle_proof(keys, circ, rng_spec, threads, multiopen), \
            f"GPU proof differs from the independent CPU restatement ({rng_spec[:8]}, {threads} threads, {multiopen})"
        assert plonk_verifier.verify_proof(vk, circ.instances(), proof, SECRET, 1 << k, multiopen=multiopen)
    assert ctx.launch_count() > before


def test_proof_bytes_match_the_package_prover_over_cpu_ops(ctx):
    """The same host code over the python-int ops (oracle/plonk_ops.py): isolates a device-op difference from a host-logic one."""
    k = 6
    circ = example.MyCircuit(k, 42)
    cc, cs, ref_dom, ops, vk, pk = _device_setup(ctx, k, circ)
    ref_ops = plonk_ops.RefOps(k, SECRET, ref_dom)
    ref_pk = plonk.keygen_pk(ref_ops, plonk.keygen_vk(ref_ops, cc), cc)
    assert _prove(circ, ops, pk, RNG.OneNg(), 4) == _prove(circ, ref_ops, ref_pk, RNG.OneNg(), 4)


def test_proof_verifies_k12(ctx):
    k = 12
    circ = example.MyCircuit(k, 7)
    _, _, _, ops, vk, pk = _device_setup(ctx, k, circ)
    proof = _prove(circ, ops, pk, RNG.ChaCha20Rng(bytes(32)), 8)
    assert plonk_verifier.verify_proof(vk, circ.instances(), proof, SECRET, 1 << k)
    bad = bytearray(proof)
    bad[-40] ^= 2
    try:
        assert not plonk_verifier.verify_proof(vk, circ.instances(), bytes(bad), SECRET, 1 << k)
    except ValueError:
        pass


def test_params_setup_matches_the_oracle(ctx):
    """ParamsKZG::setup (kzg/commitment.rs:64-131) on the device == the C restatement, point for point."""
    k = 7
    dom = R.EvaluationDomain(3, k)
    g, gl = C.kzg_setup(k, C.mont_from_int(SECRET, R.R_MOD), C.mont_from_int(dom.omega, R.R_MOD))
    params = h2b200.ParamsKZG.setup(k, SECRET, h2b200.B200Engine(ctx))
    assert np.array_equal(params.g.cpu().numpy().view(np.uint64), g)
    assert np.array_equal(params.g_lagrange.cpu().numpy().view(np.uint64), gl)


def test_fr_from_u512(ctx):
    """Fr::from_u512 / from_uniform_bytes: 512-bit little-endian integer mod r, edge values included."""
    engine = h2b200.B200Engine(ctx)
    ops = DeviceOps(None, engine, h2b200.EvaluationDomain(3, 4, ctx))
    rng = np.random.default_rng(3)
    limbs = rng.integers(0, 1 << 63, size=(67, 8), dtype=np.uint64) * np.uint64(2) + rng.integers(0, 2, size=(67, 8), dtype=np.uint64)
    full = np.uint64(0xFFFFFFFFFFFFFFFF)
    limbs[0] = 0
    limbs[1] = full
    limbs[2] = [int(x) for x in h2b200.fields.limbs(R.R_MOD)] + [0, 0, 0, 0]
    limbs[3] = [0, 0, 0, 0] + [int(x) for x in h2b200.fields.limbs(R.R_MOD - 1)]
    limbs[4, :4], limbs[4, 4:] = full, 0
    exp = [sum(int(row[i]) << (64 * i) for i in range(8)) % R.R_MOD for row in limbs]
    assert ops.to_ints(ops.from_u512(limbs)) == exp


def test_proof_k16_device_srs(ctx):
    """A 2^16-row proof with the SRS built on the device; checked by the verify_proof restatement."""
    k = 16
    circ = example.MyCircuit(k, 5)
    cc = circ.compile()
    cs = plonk.circuit.ConstraintSystemBack(cc.cs)
    engine = h2b200.B200Engine(ctx)
    ops = DeviceOps(h2b200.ParamsKZG.setup(k, SECRET, engine), engine, h2b200.EvaluationDomain(cs.degree(), k, ctx))
    vk = plonk.keygen_vk(ops, cc)
    pk = plonk.keygen_pk(ops, vk, cc)
    proof = _prove(circ, ops, pk, RNG.OneNg(), 16)
    assert plonk_verifier.verify_proof(vk, circ.instances(), proof, SECRET, 1 << k)


def test_proof_k20_verifies(ctx):
    """BASELINE config 4 size: a 2^20-row proof (extended domain 2^23), SRS built on the device, accepted by the verify_proof
    restatement; a second run with the same RNG gives the same bytes."""
    k = 20
    circ = example.MyCircuit(k, 42)
    cc = circ.compile()
    cs = plonk.circuit.ConstraintSystemBack(cc.cs)
    engine = h2b200.B200Engine(ctx)
    ops = DeviceOps(h2b200.ParamsKZG.setup(k, SECRET, engine), engine, h2b200.EvaluationDomain(cs.degree(), k, ctx))
    vk = plonk.keygen_vk(ops, cc)
    pk = plonk.keygen_pk(ops, vk, cc)
    proof = _prove(circ, ops, pk, RNG.OneNg(), 16)
    assert plonk_verifier.verify_proof(vk, circ.instances(), proof, SECRET, 1 << k)
    assert proof == _prove(circ, ops, pk, RNG.OneNg(), 16)


@pytest.mark.parametrize("minimum_degree", [5, None])
def test_standard_plonk_bytes_match(ctx, minimum_degree):
    """benches/plonk.rs circuit: one permutation set (minimum_degree 5, the benchmark's setting) and three chained sets (degree 3)."""
    k = 6
    circ = example.StandardPlonk(k, 9, minimum_degree)
    cc, cs, ref_dom, ops, vk, pk = _device_setup(ctx, k, circ)
    keys = PC.oracle_keys(circ, k, SECRET)
    assert keys.transcript_repr == vk.transcript_repr
    proof = _prove_with(circ, ops, pk, PC.SEED_0_31, 5)
    assert proof == PC.oracle_proof(keys, circ, PC.SEED_0_31, 5)
    assert plonk_verifier.verify_proof(vk, [], proof, SECRET, 1 << k)


def test_standard_plonk_k12(ctx):
    """BASELINE.json configs[0]: the plonk benchmark circuit at k = 12 (KZG-BN254, SHPLONK)."""
    k = 12
    circ = example.StandardPlonk(k, 3, 5)
    _, _, _, ops, vk, pk = _device_setup(ctx, k, circ)
    proof = _prove(circ, ops, pk, RNG.OneNg(), 16)
    assert plonk_verifier.verify_proof(vk, [], proof, SECRET, 1 << k)


def test_gwc_bytes_match(ctx):
    """The GWC opening (multiopen/gwc/prover.rs) on the device == the independent oracle; accepted by the VerifierGWC restatement."""
    k = 6
    circ = example.MyCircuit(k, 42)
    cc, cs, ref_dom, ops, vk, pk = _device_setup(ctx, k, circ)
    keys = PC.oracle_keys(circ, k, SECRET)
    proof = _prove_with(circ, ops, pk, "OneNg", 4, "gwc")
    assert proof == PC.oracle_proof(keys, circ, "OneNg", 4, "gwc")
    assert plonk_verifier.verify_proof(vk, circ.instances(), proof, SECRET, 1 << k, multiopen="gwc")


def test_new_entry_points_reject_bad_arguments(ctx):
    """The set-up / randomness / batch entry points return H2B_E_INVALID (never crash) on null or inconsistent arguments."""
    import ctypes
    import torch
    lib, h = ctx.lib, ctx.handle
    buf = torch.zeros((16, 8), dtype=torch.int64, device="cuda")
    p = ctypes.c_void_p(buf.data_ptr())
    assert lib.h2b_g1_fixed_base_mul_device(h, None, 4, p) != 0
    assert lib.h2b_g1_fixed_base_mul_device(h, p, 0, None) == 0
    assert lib.h2b_fr_from_u512_device(h, p, 4, None) != 0
    seeds = (ctypes.c_uint8 * 64)()
    offs = (ctypes.c_uint64 * 3)(0, 5, 3)
    assert lib.h2b_fr_random_chacha20_device(h, seeds, offs, 2, p) != 0            # decreasing offsets
    offs = (ctypes.c_uint64 * 3)(1, 5, 8)
    assert lib.h2b_fr_random_chacha20_device(h, seeds, offs, 2, p) != 0            # offsets[0] != 0
    assert lib.h2b_fr_random_chacha20_device(h, None, None, 0, None) == 0
    one = F_mont1 = h2b200.fields.fr_mont(1)
    vp = one.ctypes.data_as(ctypes.c_void_p)
    cols = (ctypes.c_void_p * 1)(buf.data_ptr())
    assert lib.h2b_coeff_to_extended_batch_device(h, cols, None, 1, 2, 3, vp, vp, vp) != 0
    assert lib.h2b_coeff_to_extended_batch_device(h, cols, cols, 1, 5, 3, vp, vp, vp) != 0     # k > extended_k
    assert lib.h2b_coeff_to_extended_batch_device(h, None, None, 0, 2, 3, vp, vp, vp) == 0


def test_random_poly_chacha20_matches_numpy(ctx):
    """h2b_fr_random_chacha20_device == rand_chacha semantics restated in numpy (plonk/rng.py), ragged chunks included."""
    engine = h2b200.B200Engine(ctx)
    ops = DeviceOps(None, engine, h2b200.EvaluationDomain(3, 4, ctx))
    seeds = [bytes([i + 1] * 32) for i in range(3)]
    chunks = [(0, 70), (70, 1), (71, 259)]
    got = ops.to_ints(ops.random_poly_chacha20(seeds, chunks))
    exp = []
    for seed, (_, length) in zip(seeds, chunks):
        exp += RNG.limbs512_to_ints(RNG.ChaCha20Rng(seed).random_fr_limbs(length))
    assert got == exp


def test_coeff_to_extended_many_matches_single(ctx):
    dom = h2b200.EvaluationDomain(6, 9, ctx)
    import torch
    cols = [torch.from_numpy(C.random_fr(1 << 9, 40 + i).view(np.int64)).cuda() for i in range(5)]
    many = dom.coeff_to_extended_many(cols)
    for c, m in zip(cols, many):
        assert torch.equal(dom.coeff_to_extended(c), m)


@pytest.mark.parametrize("shape", ["random", "small", "mixed", "constant", "two_limbs", "high_only"])
def test_lookup_permute_matches_host_restatement(ctx, shape):
    """h2b_lookup_permute_device == the numpy restatement of permute_expression_pair (lookup/prover.rs:398-494), on columns with
    many repeats and a table larger than the set of inputs; missing value -> ConstraintSystemFailure.  The shapes steer the sort
    plan: values spanning all four limbs (top limb alone decides), small integers (one pass over a few bits), a mix of both (ties on
    the top limb: the full limb sequence runs), one repeated value (nothing to sort), values below 2^100 and values that differ
    only in the top limb."""
    from h2b200.plonk import arguments as A
    engine = h2b200.B200Engine(ctx)
    for k, usable, seed in ((4, 10, 1), (10, 1000, 2), (13, (1 << 13) - 6, 3)):
        ops = DeviceOps(None, engine, h2b200.EvaluationDomain(3, k, ctx))
        n = 1 << k
        rng = np.random.default_rng(seed)
        pool = [int(x) for x in rng.integers(0, 1 << 62, size=max(3, usable // 7))]
        wide = [(p * 0x9E3779B97F4A7C15F39CC0605CEDC835 * (i + 1)) % R.R_MOD for i, p in enumerate(pool)]
        if shape == "random":
            pool = wide
        elif shape == "small":
            pool = [p % 5000 for p in pool]
        elif shape == "mixed":
            pool = [w if i % 3 == 0 else p % 5000 for i, (p, w) in enumerate(zip(pool, wide))]
        elif shape == "constant":
            pool = [wide[0]]
        elif shape == "two_limbs":
            pool = [w % (1 << 100) for w in wide]
        elif shape == "high_only":
            pool = [((w >> 192) << 192) | 77 for w in wide]
        table = [pool[int(i)] for i in rng.integers(0, len(pool), size=usable)]
        present = sorted(set(table))
        inp = [present[int(i)] for i in rng.integers(0, len(present), size=usable)]
        pad = lambda xs: xs + [12345] * (n - usable)
        d_in, d_tab = ops.from_ints(pad(inp)), ops.from_ints(pad(table))
        got_in, got_tab = ops.permute_expression_pair(d_in, d_tab, usable)
        exp_in, exp_tab = A.permute_expression_pair(ops.to_canonical(d_in, usable), ops.to_canonical(d_tab, usable))
        assert np.array_equal(ops.to_canonical(got_in, usable), exp_in)
        assert np.array_equal(ops.to_canonical(got_tab, usable), exp_tab)
        assert ops.read_ints(got_in, usable, n - usable) == [0] * (n - usable) == ops.read_ints(got_tab, usable, n - usable)
        bad = list(inp)
        bad[usable // 2] = (max(present) + 1) % R.R_MOD
        if bad[usable // 2] not in present:
            with pytest.raises(A.ConstraintSystemFailure):
                ops.permute_expression_pair(ops.from_ints(pad(bad)), d_tab, usable)


@pytest.mark.parametrize("fmt", ["Processed", "RawBytes"])
def test_key_serialization_matches_cpu_and_round_trips(ctx, fmt):
    """The serialised proving key of the device prover is byte-identical to the CPU restatement's (plonk.rs:276-287), reads back
    into device polynomials, and proves the same bytes (halo2_proofs/tests/serialization.rs:131-195)."""
    from h2b200.plonk import serialization as S
    k = 5
    circ = example.StandardPlonk(k, 9, 5)
    cc, cs, ref_dom, ops, vk, pk = _device_setup(ctx, k, circ)
    ref_ops = plonk_ops.RefOps(k, SECRET, ref_dom)
    ref_pk = plonk.keygen_pk(ref_ops, plonk.keygen_vk(ref_ops, cc), cc)
    blob = S.pk_write(ops, pk, fmt)
    assert blob == S.pk_write(ref_ops, ref_pk, fmt) and len(blob) == S.pk_bytes_length(ops, pk, fmt)
    pk2 = S.pk_read(ops, blob, fmt, cs)
    assert S.pk_write(ops, pk2, fmt) == blob
    proof = _prove(circ, ops, pk2, RNG.OneNg(), 2)
    assert proof == _prove(circ, ops, pk, RNG.OneNg(), 2)
    assert plonk_verifier.verify_proof(pk2.vk, [], proof, SECRET, 1 << k)


def test_multi_device_proof_matches_single_device():
    """One context over several device slots (MSMs sharded by point range, evaluate_h's coset NTTs dealt round-robin, SURVEY.md
    8e): the proof bytes equal the single-device proof.  On a one-GPU box the slots are three contexts' worth of streams and
    buffers on the same GPU (device list [0, 0, 0]): the sharding, per-shard host threads, peer copies and the fold run all the
    same, only the wires differ; with more GPUs the slots are distinct devices."""
    import torch
    ndev = torch.cuda.device_count()
    multi = [i % ndev for i in range(max(3, min(ndev, 4)))]
    print(f"multi-device proof over device slots {multi} ({ndev} GPU(s) visible)")
    k = 10
    circ = example.MyCircuit(k, 42)
    proofs = []
    for devices in ([0], multi):
        c = h2b200.Context(devices)
        try:
            cc = circ.compile()
            cs = plonk.circuit.ConstraintSystemBack(cc.cs)
            engine = h2b200.B200Engine(c)
            ops = DeviceOps(h2b200.ParamsKZG.setup(k, SECRET, engine), engine, h2b200.EvaluationDomain(cs.degree(), k, c))
            vk = plonk.keygen_vk(ops, cc)
            proofs.append(_prove(circ, ops, plonk.keygen_pk(ops, vk, cc), RNG.OneNg(), 8))
        finally:
            c.close()
    assert proofs[0] == proofs[1]
    assert plonk_verifier.verify_proof(vk, circ.instances(), proofs[1], SECRET, 1 << k)


def test_batch_proving_bytes_match(ctx):
    """Two circuit instances in one transcript (`Prover`, prover.rs:130-175): evaluate_h chains the second circuit's custom-gate
    Horner through PreviousValue on the device (evaluation.rs:366-395); bytes equal the independent oracle's, verifier accepts."""
    k = 6
    circuits = [example.MyCircuit(k, 42), example.MyCircuit(k, 77)]
    cc, cs, ref_dom, ops, vk, pk = _device_setup(ctx, k, circuits[0])
    keys = PC.oracle_keys(circuits[0], k, SECRET)
    instances = [c.instances() for c in circuits]
    for rng_spec, threads in (("OneNg", 4), (PC.SEED_0_31, 6)):
        proof = _prove_with(circuits, ops, pk, rng_spec, threads)
        assert proof == PC.oracle_proof(keys, circuits, rng_spec, threads)
        assert plonk_verifier.verify_proof_batch(vk, instances, proof, SECRET, 1 << k)


@pytest.mark.parametrize("name,k,workers", [("StandardPlonk", 5, 1), ("StandardPlonk3", 6, 2), ("MyCircuit", 6, 1), ("MyCircuit", 8, 3),
                                            ("MyCircuit", 10, 8)])
def test_coset_prover_bytes_match(ctx, name, k, workers):
    """evaluate_h coset by coset (plonk/cosets.py: e size-n cosets instead of one extended domain of n * e rows, dealt over
    `workers` single-device contexts -- distinct GPUs when the box has them, the same GPU otherwise): the quotient's pieces are the
    same polynomials, so the proof bytes equal the extended path's and the independent oracle's.  The key is built without any
    extended-length vector (keygen_pk(extended=False))."""
    import torch
    from h2b200.plonk.cosets import CosetProver
    circ = {"MyCircuit": lambda: example.MyCircuit(k, 42), "StandardPlonk": lambda: example.StandardPlonk(k, 7, 5),
            "StandardPlonk3": lambda: example.StandardPlonk(k, 7, None)}[name]()
    cc, cs, ref_dom, ops, vk, pk = _device_setup(ctx, k, circ)
    lean_pk = plonk.keygen_pk(ops, vk, cc, extended=False)
    assert lean_pk.l0 is None and lean_pk.fixed_cosets == [] and lean_pk.permutation.cosets == []
    ndev = torch.cuda.device_count()
    devices = [i % ndev for i in range(workers)]
    expected = _prove_with(circ, ops, pk, PC.SEED_0_31, 3)
    ops.cosets = CosetProver(ops, devices)
    try:
        e = 1 << (ops.domain.extended_k - k)
        print(f"coset prover: k = {k}, e = {e} cosets over devices {devices} ({ndev} GPU(s) visible)")
        for _ in range(2):                                       # the second proof reuses the per-coset key tables
            assert _prove_with(circ, ops, lean_pk, PC.SEED_0_31, 3) == expected
        assert _prove_with(circ, ops, lean_pk, "OneNg", 4, "gwc") == PC.oracle_proof(PC.oracle_keys(circ, k, SECRET), circ, "OneNg", 4, "gwc")
    finally:
        ops.cosets.close()
        ops.cosets = None


def test_coset_key_serialization_round_trip(ctx):
    """a key read back from bytes must drive the coset prover too (ProvingKey::read yields a fully usable key,
    plonk.rs:289-316), and a coset-only key (extended=False) must serialise to the very bytes of the full key."""
    from h2b200.plonk import serialization as S
    from h2b200.plonk.cosets import CosetProver
    k = 6
    circ = example.MyCircuit(k, 42)
    cc, cs, ref_dom, ops, vk, pk = _device_setup(ctx, k, circ)
    lean_pk = plonk.keygen_pk(ops, vk, cc, extended=False)
    blob = S.pk_write(ops, pk, "RawBytes")
    assert S.pk_write(ops, lean_pk, "RawBytes") == blob and S.pk_bytes_length(ops, lean_pk, "RawBytes") == len(blob)
    pk2 = S.pk_read(ops, blob, "RawBytes", cs)
    expected = _prove_with(circ, ops, pk, "OneNg", 4)
    ops.cosets = CosetProver(ops, [0, 0])
    try:
        assert _prove_with(circ, ops, pk2, "OneNg", 4) == expected
    finally:
        ops.cosets.close()
        ops.cosets = None


def test_coset_prover_batch_of_two_circuits(ctx):
    from h2b200.plonk.cosets import CosetProver
    k = 6
    circuits = [example.MyCircuit(k, 42), example.MyCircuit(k, 77)]
    cc, cs, ref_dom, ops, vk, pk = _device_setup(ctx, k, circuits[0])
    expected = _prove_with(circuits, ops, pk, "OneNg", 4)
    ops.cosets = CosetProver(ops, [0, 0])
    try:
        assert _prove_with(circuits, ops, plonk.keygen_pk(ops, vk, cc, extended=False), "OneNg", 4) == expected
    finally:
        ops.cosets.close()
        ops.cosets = None
