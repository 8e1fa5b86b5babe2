"""Test helper: turns one of the package's example circuits (plonk/example.py: CompiledCircuit + witness arrays) into the plain
dict / list-of-int description the independent prover oracle (oracle/plonk_prover.py) takes.  Only tests and smoke() use it: it is
the one place where the package's data and the oracle's meet."""
import numpy as np

from oracle import pyref as R


def _ints(col):
    if isinstance(col, np.ndarray):
        return R.fr_vec_from_bytes(np.ascontiguousarray(col).tobytes())       # (n, 4) uint64 Montgomery limbs
    return [int(v) for v in col]


def plain_circuit(k, compiled):
    cs = compiled.cs
    return {
        "k": k,
        "cs": {
            "num_fixed_columns": cs.num_fixed_columns, "num_advice_columns": cs.num_advice_columns,
            "num_instance_columns": cs.num_instance_columns, "num_challenges": cs.num_challenges,
            "unblinded_advice_columns": list(cs.unblinded_advice_columns), "advice_column_phase": list(cs.advice_column_phase),
            "challenge_phase": list(cs.challenge_phase), "gates": [(name, poly) for name, poly in cs.gates],
            "permutation_columns": [tuple(c) for c in cs.permutation_columns],
            "lookups": [(l.name, list(l.input_expressions), list(l.table_expressions)) for l in cs.lookups],
            "shuffles": [(s.name, list(s.input_expressions), list(s.shuffle_expressions)) for s in cs.shuffles],
            "minimum_degree": cs.minimum_degree,
        },
        "fixed": [_ints(col) for col in compiled.preprocessing.fixed],
        "copies": [(tuple(lc), lr, tuple(rc), rr) for (lc, lr), (rc, rr) in compiled.preprocessing.copies],
    }


def plain_witness(circuit):
    """witness_fn(phase, challenges) of the oracle from the example circuit's `witness`."""
    def fn(phase, challenges):
        return [None if col is None else _ints(col) for col in circuit.witness(phase, challenges)]
    return fn


def oracle_rng(spec):
    """"OneNg" (frontend_backend_split.rs:477-489) or "ChaCha20:<64 hex digits of the seed>" -> a fresh oracle RNG."""
    from oracle import plonk_prover as PP
    if spec == "OneNg":
        return PP.OneNg()
    assert spec.startswith("ChaCha20:")
    return PP.ChaCha20Rng(bytes.fromhex(spec.split(":", 1)[1]))


def package_rng(spec):
    from h2b200.plonk import rng as RNG
    if spec == "OneNg":
        return RNG.OneNg()
    return RNG.ChaCha20Rng(bytes.fromhex(spec.split(":", 1)[1]))


SEED_0_31 = "ChaCha20:" + bytes(range(32)).hex()


def oracle_keys(circuit, k, secret, transcript_repr=None):
    """keygen by the independent oracle for one of the package's example circuits."""
    from oracle import plonk_prover as PP
    return PP.keygen(plain_circuit(k, circuit.compile()), secret, transcript_repr=transcript_repr)


def oracle_proof(keys, circuits, rng_spec, threads, multiopen="shplonk"):
    """Proof bytes by the independent oracle; `circuits`: one example circuit or a list of them (one transcript, prover.rs:130-175)."""
    from oracle import plonk_prover as PP
    if not isinstance(circuits, (list, tuple)):
        circuits = [circuits]
    return PP.create_proof_batch(keys, [c.instances() for c in circuits], [plain_witness(c) for c in circuits],
                                 oracle_rng(rng_spec), threads, multiopen)
