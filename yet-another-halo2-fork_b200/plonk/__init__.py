"""Host-side mirror of the reference's proving pipeline (halo2_backend/src/plonk + poly/kzg/multiopen/shplonk), written against
the device-polynomial interface of ops.py so that every size-n vector stays on the GPU between the MSM / NTT calls
(SURVEY.md 8a row 15, 8f rows N2-N4).  Importing it does not touch the GPU."""
from . import arguments, circuit, example, gwc, keygen, polyexpr, prover, rng, serialization, shplonk, transcript
from .keygen import keygen_pk, keygen_vk
from .prover import Prover, ProverSingle, create_proof, create_proof_batch
from .transcript import Blake2bRead, Blake2bWrite

__all__ = ["arguments", "circuit", "example", "gwc", "keygen", "polyexpr", "prover", "rng", "serialization", "shplonk", "transcript", "keygen_pk",
           "keygen_vk", "Prover", "ProverSingle", "create_proof", "create_proof_batch", "Blake2bRead", "Blake2bWrite"]
