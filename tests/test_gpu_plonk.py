"""GPU parity of the whole proving pipeline: the prover mirror over DeviceOps (libh2b200.so: MSM, NTT, quotient, scans) must emit
the same proof bytes as the same code over the python-int oracle ops, given the same RNG -- north_star's "byte-identical proof
transcripts given a fixed blinding RNG", with the CPU restatement standing where the Rust binary cannot run -- and the proof must
pass the verify_proof restatement."""
import numpy as np
import pytest

import h2b200
from h2b200 import plonk
from h2b200.plonk import example, rng as RNG
from h2b200.plonk.ops import DeviceOps
from oracle import cref as C
from oracle import plonk_ops, plonk_verifier
from oracle import pyref as R

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


@pytest.mark.parametrize("k", [6, 8])
def test_proof_bytes_match_the_cpu_restatement(ctx, k):
    circ = example.MyCircuit(k, 42)
    cc, cs, ref_dom, ops, vk, pk = _device_setup(ctx, k, circ)
    ref_ops = plonk_ops.RefOps(k, SECRET, ref_dom)
    ref_vk = plonk.keygen_vk(ref_ops, cc)
    assert ref_vk.fixed_commitments == vk.fixed_commitments and ref_vk.permutation_commitments == vk.permutation_commitments
    assert ref_vk.transcript_repr == vk.transcript_repr
    ref_pk = plonk.keygen_pk(ref_ops, ref_vk, cc)
    before = ctx.launch_count()
    for make_rng, threads in ((RNG.OneNg, 4), (lambda: RNG.ChaCha20Rng(bytes(range(32))), 3)):
        proof = _prove(circ, ops, pk, make_rng(), threads)
        assert proof == _prove(circ, ref_ops, ref_pk, make_rng(), threads), "GPU proof differs from the CPU restatement"
        assert plonk_verifier.verify_proof(vk, circ.instances(), proof, SECRET, 1 << k)
    assert ctx.launch_count() > before


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
