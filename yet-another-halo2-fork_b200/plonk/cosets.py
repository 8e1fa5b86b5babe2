"""evaluate_h by coset: the quotient numerator computed on the e = 2^(extended_k - k) cosets of the size-n domain one at a time,
so that no vector of the extended length n * e is ever built, and dealt over the GPUs of the box coset by coset.

Why this is the same polynomial.  The reference evaluates every column on the extended coset {zeta * w_e^i : i < n * e}
(coeff_to_extended, domain.rs:230-244), folds the constraints row by row (evaluation.rs:317-623), divides by X^n - 1
(domain.rs:297-316) and interpolates back (extended_to_coeff, domain.rs:271-293).  Write i = e * r + j: because w_e^e = w, row i is
the point s_j * w^r with s_j = zeta * w_e^j -- the extended coset is the union of e cosets s_j * H of the size-n domain H, row r of
coset j being row e * r + j.  A rotation moves e rows on the extended domain (evaluation.rs:21-24: idx + rot * rot_scale), i.e. it
stays inside coset j and moves `rot` rows there.  So the pointwise kernels run unchanged on coset j with size n and rot_scale 1
(k == extended_k in the C ABI), given
    * the columns on coset j:        NTT_n of (a_t * s_j^t)_t            (a size-n transform of the scaled coefficients);
    * X on row r of coset j:         s_j * w^r                           (the permutation argument's delta^c * beta * X term: the C ABI
                                                                          takes the start beta * s_j and the step w as arguments);
    * 1 / (X^n - 1) on coset j:      the constant t_evaluations[j]        (X^n = zeta^n * (w_e^n)^j on the whole coset).
Back to coefficients: h has degree < n * e; h = sum_q X^(q n) h_q(X) with deg h_q < n -- the h_q ARE the pieces the vanishing
argument commits (vanishing/prover.rs:117-124).  On coset j, X^n = s_j^n, so the degree-<n interpolant of h there is
c_j = sum_q (s_j^n)^q h_q, coefficient by coefficient: c_j = iNTT_n(values_j) unscaled by s_j^-t.  With s_j^n = zeta^n * eps^j
(eps = w_e^n, a primitive e-th root of unity) this is an e-point DFT across the pieces, inverted per coefficient:
    h_q = zeta^(-n q) / e * sum_j eps^(-j q) c_j.
Every step is exact field arithmetic on the same polynomial, so the pieces -- and the proof bytes -- equal those of the extended
path (tests/test_gpu_plonk.py::test_coset_prover_*).

Memory: a device holds the coefficient form of every column (n each) and the working set of ONE coset (n per column) instead of
n * e per column; the proving key's coset tables (fixed, sigma, l_0, l_last, l_active) are kept per coset on the device that owns
the coset.  With G GPUs, coset j is evaluated on GPU j mod G; what crosses NVLink is the columns' coefficients (once per proof and
device) and one n-vector c_j per coset on the way back.
"""
from __future__ import annotations

import threading
from concurrent.futures import ThreadPoolExecutor
from typing import List, Optional, Sequence

from .. import evaluation as EV
from .. import fields as F
from ..domain import EvaluationDomain
from ..engine import B200Engine, Context
from . import circuit as C
from . import polyexpr as PX
from .keygen import FR_DELTA

R = F.R_MOD


class _Ready:
    """A column on the worker's device, possibly still in flight: get() waits for its copy once."""

    def __init__(self, tensor, event=None):
        self.tensor, self.event = tensor, event

    def get(self):
        if self.event is not None:
            self.event.synchronize()
            self.event = None
        return self.tensor


class _Worker:
    """One GPU's share: a single-device library context, the plain size-n domain on it, and the per-coset tables it owns."""

    def __init__(self, device: int, ctx: Context, k: int, own_ctx: bool):
        from .ops import DeviceOps
        self.device, self.ctx, self.own_ctx = device, ctx, own_ctx
        self.domain = EvaluationDomain(1, k, ctx)                     # quotient degree 0: extended_k == k, plain transforms
        self.ops = DeviceOps(None, B200Engine(ctx), self.domain)
        self.powers = {}                                              # j -> (s_j^t)_t, (s_j^-t)_t
        self.key = {}                                                 # j -> dict of the proving key's columns on coset j
        self.polys = None                                             # the proving key's coefficient forms on this device


class CosetProver:
    """Attach to a DeviceOps (`ops.cosets = CosetProver(ops)`) and create_proof evaluates the quotient by coset.
    devices: CUDA device index per worker (default: the devices of the ops' context; repeats are allowed and give several workers
    on one GPU, which is how a one-GPU box exercises the dealing)."""

    def __init__(self, ops, devices: Optional[Sequence[int]] = None):
        import torch
        self.torch = torch
        self.main = ops
        d = ops.domain
        self.k, self.n = d.k, d.n
        self.e = 1 << (d.extended_k - d.k)
        self.omega, self.omega_e = d.omega, d.extended_omega
        self.shifts = [F.FR_ZETA * pow(d.extended_omega, j, R) % R for j in range(self.e)]
        self.t_inv = list(d.t_evaluations)
        assert len(self.t_inv) == self.e
        self.pieces = d.quotient_poly_degree
        home = torch.cuda.current_device()
        if devices is None:
            devices = getattr(ops.ctx, "device_ids", None) or [home]
        self.workers: List[_Worker] = []
        for i, dev in enumerate(devices):
            with torch.cuda.device(dev):
                if i == 0 and dev == home:
                    self.workers.append(_Worker(dev, ops.ctx, self.k, False))
                else:
                    self.workers.append(_Worker(dev, Context([dev]), self.k, True))
        self.pool = ThreadPoolExecutor(max_workers=len(self.workers)) if len(self.workers) > 1 else None
        self._key_id = None

    def close(self):
        if self.pool:
            self.pool.shutdown()
        for w in self.workers:
            w.key.clear(); w.powers.clear(); w.polys = None
            if w.own_ctx:
                w.ctx.close()
        self.workers = []

    # ---- per-worker helpers (run with the worker's device current) ---------------------------------------------------------
    def _powers(self, w: _Worker, j: int):
        if j not in w.powers:
            ops = w.ops
            fwd = ops.grand_product(ops.full(self.n, self.shifts[j]), 1, self.n)
            inv = ops.grand_product(ops.full(self.n, pow(self.shifts[j], -1, R)), 1, self.n)
            w.powers[j] = (fwd, inv)
        return w.powers[j]

    def _to_coset(self, w: _Worker, poly, j: int):
        """coefficients (n) -> evaluations on coset j, a new tensor."""
        scaled = w.ops.poly.mul(poly, self._powers(w, j)[0])                         # a_t * s_j^t  (h2b_fr_mul_device)
        return w.domain.coeff_to_lagrange(scaled)

    def _move(self, w: _Worker, t):
        """A device tensor of the main device on the worker's device (NVLink peer copy when they differ)."""
        if t.device.index == w.device:
            return t
        out = t.to(f"cuda:{w.device}")
        self.torch.cuda.current_stream().synchronize()
        return out

    def _key_polys(self, w: _Worker, pk):
        """The proving key in coefficient form on the worker's device: fixed, sigma, l_0, l_last, l_active."""
        if w.polys is None:
            mv = lambda p: self._move(w, p)
            w.polys = {"fixed": [mv(p) for p in pk.fixed_polys], "sigma": [mv(p) for p in pk.permutation.polys],
                       "l0": mv(pk.l0_poly), "l_last": mv(pk.l_last_poly), "l_active": mv(pk.l_active_poly)}
        return w.polys

    def _key_coset(self, w: _Worker, pk, j: int):
        if j not in w.key:
            kp = self._key_polys(w, pk)
            to = lambda p: self._to_coset(w, p, j)
            w.key[j] = {"fixed": [to(p) for p in kp["fixed"]], "sigma": [to(p) for p in kp["sigma"]], "l0": to(kp["l0"]),
                        "l_last": to(kp["l_last"]), "l_active": to(kp["l_active"])}
        return w.key[j]

    # ---- one coset ---------------------------------------------------------------------------------------------------------
    def _coset(self, w: _Worker, j: int, pk, circuits, challenges, y, beta, gamma, theta):
        """evaluation.rs:317-623 on coset j -> c_j (see the module docstring), on the worker's device."""
        cs, ev, k, ctx = pk.vk.cs, pk.ev, self.k, w.ctx
        key = self._key_coset(w, pk, j)
        fixed, l0, l_last, l_active = key["fixed"], key["l0"], key["l_last"], key["l_active"]
        to = lambda p: self._to_coset(w, p.get(), j)                                  # p: a _Ready (fetched column)
        graph = lambda g, adv, inst, prev=None: g.evaluate(ctx, fixed, adv, inst, list(challenges), beta, gamma, theta, y, k, k,
                                                           previous=prev)
        values = None
        for circ in circuits:
            advice = [to(p) for p in circ["advice"]]
            instance = [to(p) for p in circ["instance"]]
            values = graph(ev.custom_gates, advice, instance, values)
            if circ["permutation"]:
                source = {"advice": advice, "fixed": fixed, "instance": instance}
                col_values = [source[kind][index] for kind, index in cs.permutation_columns]
                sets = [to(p) for p in circ["permutation"]]
                values = EV.permutation_fold(ctx, values, sets, col_values, key["sigma"], cs.degree() - 2, l0, l_last, l_active,
                                             beta, gamma, y, self.shifts[j], FR_DELTA, self.omega, cs.blinding_factors(), k, k)
                del sets
            for n_l, (product, permuted_input, permuted_table) in enumerate(circ["lookups"]):
                table_values = graph(ev.lookups[n_l], advice, instance)
                values = EV.lookup_fold(ctx, values, table_values, to(product), to(permuted_input), to(permuted_table), l0, l_last,
                                        l_active, beta, gamma, y, k, k)
                del table_values
            for n_s, product in enumerate(circ["shuffles"]):
                input_values = graph(ev.shuffles[2 * n_s], advice, instance)
                shuffle_values = graph(ev.shuffles[2 * n_s + 1], advice, instance)
                values = EV.shuffle_fold(ctx, values, input_values, shuffle_values, to(product), l0, l_last, l_active, y, k, k)
                del input_values, shuffle_values
            del advice, instance
        if values is None:
            values = w.ops.zeros(self.n)
        values = w.domain.lagrange_to_coeff(values)                                   # iNTT_n, 1/n included
        # s_j^-t, and 1 / (X^n - 1) which is the constant t_evaluations[j] on this coset
        return w.ops.poly.mul(values, self._powers(w, j)[1], scale=self.t_inv[j], out=values)

    def _run_worker(self, wi: int, pk, circuits, challenges, y, beta, gamma, theta):
        w = self.workers[wi]
        torch = self.torch
        with torch.cuda.device(w.device):
            # Columns living on another device are pulled over NVLink before the first coset (blocking peer copies: issuing them
            # asynchronously on a side stream and waiting per column at first use measured SLOWER, 56 vs 46 ms per worker at
            # k = 22 on 4 GPUs -- torch queues cross-device copies on the SOURCE device's stream, where the workers serialise).
            def fetch(t):
                return _Ready(self._move(w, t))

            local = []
            for c in circuits:                               # fetch order = use order inside _coset
                lc = {"advice": [fetch(p) for p in c["advice"]], "instance": [fetch(p) for p in c["instance"]]}
                lc["permutation"] = [fetch(p) for p in c["permutation"]]
                lc["lookups"] = [tuple(fetch(p) for p in l) for l in c["lookups"]]
                lc["shuffles"] = [fetch(p) for p in c["shuffles"]]
                local.append(lc)
            out = {}
            for j in range(wi, self.e, len(self.workers)):
                out[j] = self._coset(w, j, pk, local, challenges, y, beta, gamma, theta)
            torch.cuda.current_stream().synchronize()
            return out

    # ---- the whole quotient ----------------------------------------------------------------------------------------------------
    def evaluate_h_pieces(self, pk, advice_polys, instance_polys, challenges, y, beta, gamma, theta, lookups, shuffles, permutations):
        """Same arguments as prover.evaluate_h; returns the quotient's pieces h_0 .. h_(d-2) in coefficient form (n each) on the
        main device -- what vanishing_construct cuts out of extended_to_coeff's output."""
        torch = self.torch
        if self._key_id is not pk:                                                    # tables belong to one proving key
            for w in self.workers:
                w.key.clear(); w.polys = None
            self._key_id = pk
        circuits = []
        for adv, inst, ls, ss, perm in zip(advice_polys, instance_polys, lookups, shuffles, permutations):
            circuits.append({"advice": list(adv), "instance": list(inst),
                             "permutation": [s.permutation_product_poly for s in perm],
                             "lookups": [(l.product_poly, l.permuted_input_poly, l.permuted_table_poly) for l in ls],
                             "shuffles": [s.product_poly for s in ss]})
        args = (pk, circuits, list(challenges), y, beta, gamma, theta)
        if self.pool:
            parts = list(self.pool.map(lambda wi: self._run_worker(wi, *args), range(len(self.workers))))
        else:
            parts = [self._run_worker(0, *args)]
        home = self.workers[0].device
        c = [None] * self.e
        with torch.cuda.device(home):
            for part in parts:
                for j, t in part.items():
                    c[j] = t if t.device.index == home else t.to(f"cuda:{home}")
            torch.cuda.current_stream().synchronize()
            eps_inv = pow(pow(self.omega_e, self.n, R), -1, R)
            zeta_n_inv = pow(pow(F.FR_ZETA, self.n, R), -1, R)
            e_inv = pow(self.e, -1, R)
            pieces = []
            for q in range(self.pieces):
                scale = pow(zeta_n_inv, q, R) * e_inv % R
                pieces.append(PX.lincomb(self.main, [(scale * pow(eps_inv, j * q, R) % R, c[j]) for j in range(self.e)]))
        return pieces
