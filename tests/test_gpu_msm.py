"""GPU parity: the CUDA MSM (through the C ABI / MsmAccel mirror) against the CPU oracle.
Bit-exact: the normalised result must equal the oracle's affine point limb for limb."""
import ctypes

import numpy as np
import pytest

from oracle import cref as C
from oracle import pyref as R

pytestmark = pytest.mark.gpu

ONE_Q = C.limbs_from_int((1 << 256) % R.Q_MOD)


def check(out12, exp_affine):
    if not exp_affine.any():
        assert not out12[:4].any() and not out12[8:].any() and np.array_equal(out12[4:8], ONE_Q)
    else:
        assert np.array_equal(out12[8:], ONE_Q), "result must be normalised (z = 1)"
        assert np.array_equal(out12[:8], exp_affine)


def oracle_affine(s, b):
    return C.to_affine(C.best_multiexp(s, b))


@pytest.mark.parametrize("n", [1, 2, 3, 4, 31, 32, 33, 100, 1 << 10, (1 << 12) + 5, 1 << 14, 1 << 16])
def test_msm_matches_oracle(engine, n):
    """zal.rs:262-349 test_msm_zal restated: engine.msm == best_multiexp == msm_with_cached_base."""
    s = C.random_fr(n, 10 + n)
    b = C.gen_bases(n, 3)
    exp = oracle_affine(s, b)
    check(engine.msm(s, b), exp)
    desc = engine.get_base_descriptor(b)
    check(engine.msm_with_cached_base(s, desc), exp)
    cdesc = engine.get_coeffs_descriptor(s)
    check(engine.msm_with_cached_inputs(cdesc, desc), exp)
    check(engine.msm_with_cached_scalars(cdesc, b), exp)
    desc.release()


@pytest.mark.parametrize("c", [4, 5, 7, 8, 11, 13, 16, 18, 20])
def test_window_sizes(engine, c, knob):
    knob("msm_c", c)
    n = 3000
    s = C.random_fr(n, c)
    b = C.gen_bases(n, 9)
    check(engine.msm(s, b), oracle_affine(s, b))


@pytest.mark.parametrize("pre_c", [4, 6, 9, 13, 16, 20])
def test_precomputed_window_multiples(engine, pre_c, knob):
    """Registered bases carry 2^(c*j) * P_i for every window (one shared bucket set); any c must give the same point,
    also for a prefix of the registered bases and for identity bases / edge scalars."""
    knob("msm_pre_c", pre_c)
    n = 2500
    s = C.random_fr(n, 100 + pre_c)
    b = C.gen_bases(n, 13)
    s[0] = 0; s[1] = C.mont_from_int(R.R_MOD - 1, R.R_MOD); s[2] = C.mont_from_int(1, R.R_MOD); b[3] = 0; b[5] = b[4]
    desc = engine.get_base_descriptor(b)
    check(engine.msm_with_cached_base(s, desc), oracle_affine(s, b))
    check(engine.msm_with_cached_base(s[:777], desc), oracle_affine(s[:777], b[:777]))
    desc.release()


def test_precompute_can_be_disabled(engine, knob):
    knob("msm_precompute", 0)
    n = 4000
    s, b = C.random_fr(n, 5), C.gen_bases(n, 2)
    desc = engine.get_base_descriptor(b)
    assert desc.info() == (0, n * 64)
    check(engine.msm_with_cached_base(s, desc), oracle_affine(s, b))
    desc.release()


@pytest.mark.parametrize("chunks,precompute", [(3, "1"), (5, "0"), (2, "1")])
def test_streamed_scalar_chunks(engine, chunks, precompute, knob):
    """Host scalars are streamed in chunks that add into the same buckets (copy / compute overlap): same point."""
    knob("msm_chunks", chunks)
    knob("msm_precompute", precompute)
    knob("msm_pre_c", 11)
    n = 5001
    s, b = C.random_fr(n, 60 + chunks), C.gen_bases(n, 4)
    s[10:400] = C.mont_from_int(1, R.R_MOD)          # an oversized bucket that spans chunks
    b[1700] = 0
    desc = engine.get_base_descriptor(b)
    exp = oracle_affine(s, b)
    check(engine.msm_with_cached_base(s, desc), exp)
    check(engine.msm_with_cached_base(s, desc), exp)      # scratch reuse across calls
    desc.release()


@pytest.mark.parametrize("msm_sort", [0, 1])
@pytest.mark.parametrize("shape", ["classic_c13", "pre_c16", "pre_c20_512_bins", "pre_c22_lb12", "skewed", "chunked"])
def test_binning_variants(engine, knob, msm_sort, shape):
    """The two binning pipelines -- global-atomic count + scatter (H2B_MSM_SORT=0) and the tiled two-level one over shared-memory
    atomics (msm_bin.cuh, the default) -- give the oracle's point for every key geometry: several windows' bucket sets side by
    side (classic), one shared set with 2^15 ... 2^21 keys (1 ... 512 coarse bins of 2^10 ... 2^12 keys), skewed digits (one
    coarse bin takes almost everything), and a streamed MSM whose chunks add into the same buckets."""
    knob("msm_sort", msm_sort)
    n, seed = 20000, 77
    registered = True
    if shape == "classic_c13":
        knob("msm_c", 13); registered = False
    elif shape == "pre_c16":
        knob("msm_pre_c", 16); n = 1 << 16
    elif shape == "pre_c20_512_bins":
        knob("msm_pre_c", 20); n = (1 << 17) + 11
    elif shape == "pre_c22_lb12":
        knob("msm_pre_c", 22); n = 5000
    elif shape == "chunked":
        knob("msm_pre_c", 12); knob("msm_chunks", 3); n = 50001
    s = C.random_fr(n, seed)
    if shape == "skewed":
        knob("msm_pre_c", 18)
        rng = np.random.default_rng(5)
        sel = rng.random(n)
        s[sel < 0.5] = C.mont_from_int(1, R.R_MOD)
        s[(sel >= 0.5) & (sel < 0.8)] = 0
        s[(sel >= 0.8) & (sel < 0.9)] = C.mont_from_int(R.R_MOD - 1, R.R_MOD)
    b = C.gen_bases(n, 4)
    exp = oracle_affine(s, b)
    if registered:
        desc = engine.get_base_descriptor(b)
        check(engine.msm_with_cached_base(s, desc), exp)
        check(engine.msm_with_cached_base(s[: n // 3], desc), oracle_affine(s[: n // 3], b[: n // 3]))
        desc.release()
    else:
        check(engine.msm(s, b), exp)


@pytest.mark.parametrize("regs", [0, 1])
@pytest.mark.parametrize("rounds", [1, 2])
@pytest.mark.parametrize("shape", ["classic_c8", "pre_c9", "pre_c11_chunked", "pre_c16", "collisions", "skewed", "sparse"])
def test_batched_affine_accumulation(engine, knob, rounds, regs, shape):
    """msm_ba.cuh: the bucket accumulation with 1 or 2 rounds of batched affine additions (pairs of a bucket's entries, ONE
    inversion per block and round) gives the oracle's point: dense buckets (dozens of pairs per thread, more than MSM_BA_KCAP
    of them, odd leftovers), chunks adding into the same buckets, pairs the affine formula cannot take (repeated bases with
    equal or opposite scalars, identity bases), oversized buckets next to it, and buckets with no pair at all."""
    knob("msm_ba", rounds)
    knob("msm_ba_regs", regs)
    knob("msm_ba_min_mean", 1)
    n, registered = 30000, True
    if shape == "classic_c8":
        knob("msm_c", 8); registered = False                       # 128 buckets per window: ~230 entries per bucket (> 2 * KCAP)
    elif shape == "pre_c9":
        knob("msm_pre_c", 9); n = 20011                            # one set of 256 buckets: thousands of entries each (the long path)
    elif shape == "pre_c11_chunked":
        knob("msm_pre_c", 11); knob("msm_chunks", 3); n = 50001
    elif shape == "pre_c16":
        knob("msm_pre_c", 16); n = (1 << 17) + 3
    elif shape == "sparse":
        knob("msm_pre_c", 20); n = 4097                            # ~0.1 entries per bucket
    elif shape in ("collisions", "skewed"):
        knob("msm_pre_c", 12)
    s = C.random_fr(n, 300 + rounds)
    b = C.gen_bases(n, 6)
    if shape == "collisions":
        # a few distinct (base, scalar) pairs repeated many times: equal points and opposite points meet inside the buckets
        m = 40
        rng = np.random.default_rng(9)
        pick = rng.integers(0, m, n)
        neg = rng.random(n) < 0.3
        b = b[pick]
        s = s[pick]
        s[neg] = C.fr_sub(np.zeros_like(s[neg]), s[neg])
        b[rng.random(n) < 0.05] = 0
    if shape == "skewed":
        rng = np.random.default_rng(5)
        sel = rng.random(n)
        s[sel < 0.5] = C.mont_from_int(1, R.R_MOD)
        s[(sel >= 0.5) & (sel < 0.8)] = 0
        s[(sel >= 0.8) & (sel < 0.9)] = C.mont_from_int(R.R_MOD - 1, R.R_MOD)
    exp = oracle_affine(s, b)
    if registered:
        desc = engine.get_base_descriptor(b)
        check(engine.msm_with_cached_base(s, desc), exp)
        check(engine.msm_with_cached_base(s[: n // 3], desc), oracle_affine(s[: n // 3], b[: n // 3]))
        desc.release()
    else:
        check(engine.msm(s, b), exp)


def test_edge_scalars_and_bases(engine):
    n = 600
    s = C.random_fr(n, 5)
    b = C.gen_bases(n, 1)
    s[0] = 0
    s[1] = C.mont_from_int(1, R.R_MOD)
    s[2] = C.mont_from_int(R.R_MOD - 1, R.R_MOD)
    s[3] = C.mont_from_int((1 << 253) + 12345, R.R_MOD)
    b[4] = 0                                  # identity base (0,0)
    b[6] = b[5]; s[6] = s[5]                  # repeated (P, k), (P, k): doubling inside a bucket
    b[8] = b[7]; s[8] = C.mont_from_int((-C.int_from_mont(s[7], R.R_MOD)) % R.R_MOD, R.R_MOD)   # P, -k: cancellation
    check(engine.msm(s, b), oracle_affine(s, b))
    # empty, all-zero scalars, all-identity bases -> identity
    check(engine.msm(np.zeros((0, 4), dtype=np.uint64), np.zeros((0, 8), dtype=np.uint64)), np.zeros(8, dtype=np.uint64))
    check(engine.msm(np.zeros((50, 4), dtype=np.uint64), C.gen_bases(50, 2)), np.zeros(8, dtype=np.uint64))
    check(engine.msm(C.random_fr(50, 1), np.zeros((50, 8), dtype=np.uint64)), np.zeros(8, dtype=np.uint64))
    # single generator times k
    k = C.random_fr(1, 4)
    g = C.gen_bases(1, 1)
    check(engine.msm(k, g), C.affine_mul(g[0], k[0]))


def test_skewed_scalars_long_buckets(engine):
    """Columns of a real circuit are mostly 0/1/small values (SURVEY.md §7 hard parts): one bucket holds
    most of the points, which exercises the block-cooperative oversized-bucket path."""
    n = 1 << 15
    s = C.random_fr(n, 21)
    one = C.mont_from_int(1, R.R_MOD)
    rng = np.random.default_rng(3)
    sel = rng.random(n)
    s[sel < 0.45] = 0
    s[(sel >= 0.45) & (sel < 0.9)] = one
    small = np.where(sel >= 0.97)[0]
    for i in small:
        s[i] = C.mont_from_int(int(rng.integers(0, 1 << 16)), R.R_MOD)
    b = C.gen_bases(n, 5)
    check(engine.msm(s, b), oracle_affine(s, b))
    # every scalar identical -> every window has exactly one (huge) bucket
    s2 = np.tile(C.random_fr(1, 8), (n, 1))
    check(engine.msm(s2, b), oracle_affine(s2, b))


def test_length_mismatch_and_short_commit(engine):
    s = C.random_fr(10, 1)
    b = C.gen_bases(8, 1)
    with pytest.raises(ValueError):
        engine.msm(s, b)
    desc = engine.get_base_descriptor(b)
    with pytest.raises(ValueError):
        engine.msm_with_cached_base(s, desc)
    # fewer scalars than cached bases uses the prefix (commitment.rs:313-316)
    check(engine.msm_with_cached_base(s[:5], desc), oracle_affine(s[:5], b[:5]))


def test_commit_lagrange_identity(engine, ctx):
    """kzg/commitment.rs:381-408 test_commit_lagrange restated with a known secret:
    commit(lagrange_to_coeff(a)) == commit_lagrange(a), and both equal [p(s)]G."""
    import h2b200
    k = 6
    s_secret = 0x1234567890ABCDEF1234567890ABCDEF
    ref_dom = R.EvaluationDomain(1, k)
    g, gl = C.kzg_setup(k, C.mont_from_int(s_secret, R.R_MOD), C.mont_from_int(ref_dom.omega, R.R_MOD))
    params = h2b200.ParamsKZG(k, g, gl, engine)
    dom = h2b200.EvaluationDomain(1, k, ctx)
    a = C.fr_array_from_ints(list(range(1 << k)), R.R_MOD)           # a[i] = i, as in the reference test
    coeffs = dom.lagrange_to_coeff(a.copy())
    c1 = params.commit(engine, coeffs)
    c2 = params.commit_lagrange(engine, a)
    assert np.array_equal(c1, c2)
    p_at_s = R.eval_polynomial(C.ints_from_fr_array(coeffs, R.R_MOD), s_secret)
    exp = R.g1_mul(R.G1_GEN, p_at_s)
    assert h2b200.fields.jacobian_to_affine_ints(c1) == exp


def test_sharded_windows_combine(engine):
    """Point-range sharding (multi-GPU path): per-shard window sums + h2b_msm_combine == whole MSM."""
    n = 5000
    s = C.random_fr(n, 31)
    b = C.gen_bases(n, 7)
    parts = []
    cs = set()
    for lo, hi in [(0, 1200), (1200, 1201), (1201, 5000)]:
        desc = engine.get_base_descriptor(b[lo:hi])
        w, c = engine.msm_windows(np.ascontiguousarray(s[lo:hi]), desc, n)
        parts.append(w)
        cs.add(c)
        desc.release()
    assert len(cs) == 1
    out = engine.msm_combine(np.stack(parts), cs.pop())
    check(out, oracle_affine(s, b))


def test_linearity_full_size(engine):
    """BASELINE size 2^24 via size-independent properties: MSM(s, B) + MSM(t, B) == MSM(s + t, B) and a
    concatenation split, all on device-resident inputs."""
    import torch
    n = 1 << 24
    b = C.gen_bases(n, 1)
    s = C.random_fr(n, 1)
    desc = engine.get_base_descriptor(b)
    ds = torch.from_numpy(s.view(np.int64)).cuda()
    whole = engine.msm_with_cached_base(ds, desc)
    half = n // 2
    lo = engine.msm_with_cached_base(ds[:half], desc)
    d_hi = engine.get_base_descriptor(b[half:])
    hi = engine.msm_with_cached_base(ds[half:].contiguous(), d_hi)
    assert np.array_equal(C.affine_add(lo[:8], hi[:8]), whole[:8])
    assert C.is_on_curve(whole[:8])
    # scalar all-ones: MSM == sum of bases == [sum_{i} (1+i)]G
    ones = torch.from_numpy(np.tile(C.mont_from_int(1, R.R_MOD), (n, 1)).view(np.int64)).cuda()
    tot = engine.msm_with_cached_base(ones, desc)
    k = (n * (n + 1) // 2) % R.R_MOD
    assert np.array_equal(tot[:8], C.affine_mul(C.gen_bases(1, 1)[0], C.mont_from_int(k, R.R_MOD)))


def test_single_process_multi_device_context():
    """h2b_init with several device slots: bases sharded by point range, shards run concurrently, host folds the sums;
    batched NTT columns are dealt round-robin.  On a one-GPU box the slots share the GPU (device list [0, 0, 0]): every
    host-side path (sharding, per-shard threads, peer copies, fold) runs, only NVLink is not involved."""
    import h2b200
    ndev = h2b200.load().h2b_device_count()
    slots = [i % ndev for i in range(max(3, min(ndev, 4)))]
    print(f"multi-device context over device slots {slots} ({ndev} GPU(s) visible)")
    ctx2 = h2b200.Context(slots)
    eng2 = h2b200.B200Engine(ctx2)
    for n in (1, 3, 1000, (1 << 17) + 11):
        s, b = C.random_fr(n, 900 + n), C.gen_bases(n, 6)
        exp = oracle_affine(s, b)
        check(eng2.msm(s, b), exp)
        desc = eng2.get_base_descriptor(b)
        check(eng2.msm_with_cached_base(s, desc), exp)
        check(eng2.msm_with_cached_base(s[: n // 2 + 1], desc), oracle_affine(s[: n // 2 + 1], b[: n // 2 + 1]))
        desc.release()
    dom = h2b200.EvaluationDomain(5, 10, ctx2)
    cols = [C.random_fr(1 << 10, 70 + i) for i in range(7)]
    ref = R.EvaluationDomain(5, 10)
    m = lambda x: C.mont_from_int(x, R.R_MOD)
    exp_cols = [C.coeff_to_extended(c, 10, ref.extended_k, m(ref.extended_omega), m(ref.g_coset), m(ref.g_coset_inv)) for c in cols]
    for e, o in zip(exp_cols, dom.coeff_to_extended_batch(cols)):
        assert np.array_equal(o, e)
    import torch
    dev_cols = [torch.from_numpy(c.view(np.int64)).cuda() for c in cols]      # device-resident columns, dealt over the slots
    for e, o in zip(exp_cols, dom.coeff_to_extended_many(dev_cols)):
        assert np.array_equal(o.cpu().numpy().view(np.uint64), e)
    ctx2.close()


@pytest.mark.parametrize("n", [5, 3000, (1 << 16) + 7, 1 << 18])
def test_msm_batch_equals_column_by_column(engine, n):
    """h2b_msm_batch (the per-column commits of one prover phase, plonk/prover.rs:436-440): one pipeline, same points.
    n >= 2^16 takes the batched path (precomputed windows, one bucket set per column); small n the per-column fallback."""
    import torch
    b = C.gen_bases(n, 9)
    desc = engine.get_base_descriptor(b)
    cols = [C.random_fr(n, 500 + i) for i in range(5)]
    cols[1][::2] = 0                                   # selector-like column: half zeros ...
    cols[1][1::4] = C.mont_from_int(1, R.R_MOD)        # ... a quarter ones (one oversized bucket)
    cols[2][:] = cols[0]                               # a repeated column
    cols[3][:] = 0                                     # all zero -> identity
    exp = [engine.msm_with_cached_base(c, desc) for c in cols]
    for c, e in zip(cols[:3], exp[:3]):
        check(e, oracle_affine(c, b))
    got = engine.msm_batch(cols, desc)
    for g, e in zip(got, exp):
        assert np.array_equal(g, e)
    dev = [torch.from_numpy(c.view(np.int64)).cuda() for c in cols]
    got = engine.msm_batch(dev, desc)
    for g, e in zip(got, exp):
        assert np.array_equal(g, e)
    assert engine.msm_batch([], desc).shape == (0, 12)
    desc.release()


def test_concurrent_callers_share_one_context(engine, ctx):
    """The engine is borrowed immutably everywhere (`&self`, zal.rs:57-58) and EvaluationDomain methods are called from inside
    `parallelize` in keygen (permutation/keygen.rs:188-205): entry points must be safe to call from several host threads."""
    import threading
    import h2b200
    n = 1 << 13
    b = C.gen_bases(n, 4)
    desc = engine.get_base_descriptor(b)
    scal = [C.random_fr(n, 700 + i) for i in range(4)]
    exp_msm = [oracle_affine(s, b) for s in scal]
    dom = h2b200.EvaluationDomain(5, 10, ctx)
    ref = R.EvaluationDomain(5, 10)
    mm = lambda x: C.mont_from_int(x, R.R_MOD)
    polys = [C.random_fr(1 << 10, 800 + i) for i in range(4)]
    exp_ntt = [C.lagrange_to_coeff(p, 10, mm(ref.omega_inv), mm(ref.ifft_divisor)) for p in polys]
    errors = []

    def worker(i):
        try:
            for _ in range(5):
                check(engine.msm_with_cached_base(scal[i], desc), exp_msm[i])
                assert np.array_equal(dom.lagrange_to_coeff(polys[i].copy()), exp_ntt[i])
        except Exception as e:          # noqa: BLE001
            errors.append(repr(e))

    threads = [threading.Thread(target=worker, args=(i,)) for i in range(4)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    desc.release()


def _run_ranks(fn, world):
    import threading
    out, err = [None] * world, []

    def work(r):
        try:
            out[r] = fn(r)
        except Exception as e:          # noqa: BLE001
            err.append(repr(e))

    th = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not err, err
    return out


@pytest.mark.parametrize("world,n", [(1, 1000), (2, 5000), (3, 3 * 1400), (4, 1 << 18), (2, 1 << 17), (8, 8 * 700)])
def test_msm_sharded_peer_exchange(world, n):
    """h2b_msm_sharded: one single-GPU context per rank (here: `world` contexts in one process, dealt over the visible
    devices, mailboxes handed over as pointers; bench.py does the same across processes through CUDA IPC).  The last kernel of
    each rank's reduction stores its terms into every peer's mailbox, waits and folds: every rank returns the whole result."""
    import h2b200
    ndev = h2b200.load().h2b_device_count()
    ctxs = [h2b200.Context([r % ndev]) for r in range(world)]
    engs = [h2b200.B200Engine(c) for c in ctxs]
    for e in engs:
        e.peer_export()
    for r, e in enumerate(engs):
        e.peer_attach(r, world, None, {q: engs[q] for q in range(world) if q != r})
    s, b = C.random_fr(n, 40 + world), C.gen_bases(n, 11)
    s[7] = 0; b[9] = 0
    exp = oracle_affine(s, b)
    per = n // world
    descs = [engs[r].get_base_descriptor(np.ascontiguousarray(b[r * per:(r + 1) * per])) for r in range(world)]
    for rep in range(3):                     # three epochs: both mailbox parities and their reuse
        outs = _run_ranks(lambda r: engs[r].msm_sharded(np.ascontiguousarray(s[r * per:(r + 1) * per]), descs[r], n), world)
        for o in outs:
            check(o, exp)
    for d in descs:
        d.release()
    for c in ctxs:
        c.close()


@pytest.mark.parametrize("n", [6000, 1 << 18])
def test_msm_sharded_across_processes_over_ipc(n):
    """The deployment shape of bench.py --gpus N: one PROCESS per rank, mailboxes opened through CUDA IPC handles (the in-process
    tests above hand raw pointers over).  Two gloo ranks (tests/_dist_gpu_worker.py); equal shard sizes so that both ranks reduce
    with the same window geometry."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29541")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29541", os.path.join(root, "tests", "_dist_gpu_worker.py"), str(n)]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]


def test_msm_sharded_reports_a_missing_peer():
    import h2b200
    ctxs = [h2b200.Context([0]) for _ in range(2)]
    engs = [h2b200.B200Engine(c) for c in ctxs]
    for e in engs:
        e.peer_export()
    for r, e in enumerate(engs):
        e.peer_attach(r, 2, None, {1 - r: engs[1 - r]})
    ctxs[0].set_option("peer_timeout_ms", 300)
    s, b = C.random_fr(64, 1), C.gen_bases(64, 1)
    d = engs[0].get_base_descriptor(b)
    with pytest.raises(h2b200.H2bError) as e:
        engs[0].msm_sharded(s, d, 128)           # rank 1 never calls
    assert "rank 1" in str(e.value)
    d.release()
    for c in ctxs:
        c.close()


def _peer_group(world, options=()):
    """`world` single-GPU contexts in this process, dealt over the visible devices, exported and attached to each other."""
    import h2b200
    ndev = h2b200.load().h2b_device_count()
    ctxs = [h2b200.Context([r % ndev]) for r in range(world)]
    for c in ctxs:
        for name, value in options:
            c.set_option(name, value)
    engs = [h2b200.B200Engine(c) for c in ctxs]
    for e in engs:
        e.peer_export()
    _attach_all(engs)
    return ctxs, engs


def _attach_all(engs):
    world = len(engs)
    for r, e in enumerate(engs):
        e.peer_attach(r, world, None, {q: engs[q] for q in range(world) if q != r})


def test_msm_sharded_after_reattach_ignores_the_old_groups_mailbox():
    """h2b_peer_attach restarts the epoch at 0.  The flags, tags and slots three earlier calls left in the mailboxes
    must not satisfy the new group's epochs 1..3 (they would be folded as if they were this call's terms): attach clears them,
    and the wait accepts exactly its own epoch."""
    world, n = 2, 6000
    ctxs, engs = _peer_group(world)
    per = n // world
    b = C.gen_bases(n, 21)
    descs = [engs[r].get_base_descriptor(np.ascontiguousarray(b[r * per:(r + 1) * per])) for r in range(world)]
    for session in range(3):
        if session:
            _attach_all(engs)                # the same mailboxes, a new group
        for rep in range(3):
            s = C.random_fr(n, 500 + 10 * session + rep)          # different scalars every call: stale terms give a wrong point
            exp = oracle_affine(s, b)
            outs = _run_ranks(lambda r: engs[r].msm_sharded(np.ascontiguousarray(s[r * per:(r + 1) * per]), descs[r], n), world)
            for o in outs:
                check(o, exp)
    for d in descs:
        d.release()
    for c in ctxs:
        c.close()


def test_msm_sharded_recovers_after_a_one_sided_failure():
    """The recovery contract of include/h2b200.h: every h2b_msm_sharded call counts, failed ones too.  Rank 0's call is refused
    (unknown handle) while rank 1 publishes and times out; the next call of both ranks is in step again and correct.  A rank that
    SKIPS a call gets an error on its next one (the peer's terms for the skipped call are marked "gave up"), never a silently
    wrong point; attaching again repairs the group."""
    import h2b200
    world, n = 2, 4000
    ctxs, engs = _peer_group(world, [("peer_timeout_ms", 400)])
    per = n // world
    b = C.gen_bases(n, 22)
    descs = [engs[r].get_base_descriptor(np.ascontiguousarray(b[r * per:(r + 1) * per])) for r in range(world)]
    shard = lambda s, r: np.ascontiguousarray(s[r * per:(r + 1) * per])

    def run(seed, broken_rank=None):
        s = C.random_fr(n, seed)
        errs = [None] * world

        def call(r):
            try:
                if r == broken_rank:
                    bogus = h2b200.BaseDescriptor.__new__(h2b200.BaseDescriptor)
                    bogus.ctx, bogus.n, bogus.handle = ctxs[r], per, 0xDEAD
                    return engs[r].msm_sharded(shard(s, r), bogus, n)
                return engs[r].msm_sharded(shard(s, r), descs[r], n)
            except h2b200.H2bError as e:
                errs[r] = str(e)
                return None
        outs = _run_ranks(call, world)
        return s, outs, errs

    s, outs, errs = run(700)
    assert not any(errs)
    s, outs, errs = run(701, broken_rank=0)
    assert "unknown bases handle" in errs[0] and "timed out" in errs[1] and "rank 0" in errs[1]
    for seed in (702, 703):                  # both parities after the failure
        s, outs, errs = run(seed)
        assert not any(errs), errs
        for o in outs:
            check(o, oracle_affine(s, b))
    # rank 1 skips a call altogether: rank 0 times out; on the next call rank 0 is one epoch ahead and both ranks fail loudly
    s = C.random_fr(n, 704)
    with pytest.raises(h2b200.H2bError) as e:
        engs[0].msm_sharded(shard(s, 0), descs[0], n)
    assert "timed out" in str(e.value)
    s, outs, errs = run(705)
    assert errs[1] and "rank 0 gave up" in errs[1], errs      # the terms rank 0 published for the call rank 1 skipped are marked
    assert errs[0] and "timed out" in errs[0], errs             # rank 1 never reaches rank 0's epoch
    _attach_all(engs)
    s, outs, errs = run(706)
    assert not any(errs), errs
    for o in outs:
        check(o, oracle_affine(s, b))
    for d in descs:
        d.release()
    for c in ctxs:
        c.close()


def test_msm_sharded_without_window_multiples_at_2p20():
    """bases registered without the precomputed multiples (msm_precompute=0, or skipped for lack of memory) reduce one
    bucket set per window; at n_total = 2^20 (c = 16, 2^15 buckets) the low-latency reduction would leave 16 x 17 terms, more
    than a mailbox slot holds.  The sharded path then takes the classic reduction (one term per window) instead of failing."""
    import h2b200
    world, n = 2, 1 << 20
    C.use_all_host_cores()
    ctxs, engs = _peer_group(world, [("msm_precompute", 0)])
    per = n // world
    t0 = C.bench_t0()
    d_b = h2b200.synth.device_bases(ctxs[0], n, t0 + 5)
    d_s = h2b200.synth.chacha_scalars_device(ctxs[0], n)
    s, b = d_s.cpu().numpy().view(np.uint64), d_b.cpu().numpy().view(np.uint64)
    descs = [engs[r].get_base_descriptor(np.ascontiguousarray(b[r * per:(r + 1) * per])) for r in range(world)]
    assert all(d.info()[0] == 0 for d in descs)
    exp = oracle_affine(s, b)
    for rep in range(2):
        outs = _run_ranks(lambda r: engs[r].msm_sharded(np.ascontiguousarray(s[r * per:(r + 1) * per]), descs[r], n), world)
        for o in outs:
            check(o, exp)
    for d in descs:
        d.release()
    for c in ctxs:
        c.close()


def test_seeded_inputs_match_the_oracle(ctx, engine):
    """the bench inputs come from ChaCha20 seeds so that CPU oracle and GPU agree: device generator
    (h2b_fr_random_chacha20_device / h2b_synth_bases) == oracle/cref.py's own ChaCha20 and base generator."""
    import h2b200
    n = 5000
    d_s = h2b200.synth.chacha_scalars_device(ctx, n)
    assert np.array_equal(d_s.cpu().numpy().view(np.uint64), C.random_fr_chacha20(n, C.SEED_SCALARS))
    assert h2b200.synth.bench_t0() == C.bench_t0()
    t0 = C.bench_t0()
    d_b = h2b200.synth.device_bases(ctx, 300, t0 + 17)
    assert np.array_equal(d_b.cpu().numpy().view(np.uint64), C.gen_bases(300, t0 + 17))


def test_registration_is_reported(ctx, engine, knob):
    n = 1 << 16
    b = C.gen_bases(n, 2)
    desc = engine.get_base_descriptor(b)
    c, nbytes = desc.info()
    assert c == 17 and nbytes == n * 64 + n * 15 * 64          # 15 windows of 17 bits, 64 B per multiple
    desc.release()
    with pytest.raises(h2b200_error()):
        ctx.set_option("no_such_knob", 1)


def h2b200_error():
    import h2b200
    return h2b200.H2bError


def test_full_size_msm_matches_the_oracle_directly(ctx, engine):
    """BASELINE configs[1] at its headline size, compared with the C port itself (not through identities): 2^24 points,
    the seeded bench inputs.  ~15 s of CPU."""
    import h2b200
    n = 1 << 24
    C.use_all_host_cores()
    t0 = C.bench_t0()
    d_b = h2b200.synth.device_bases(ctx, n, t0)
    d_s = h2b200.synth.chacha_scalars_device(ctx, n)
    desc = engine.get_base_descriptor(d_b)
    got = engine.msm_with_cached_base(d_s, desc)
    desc.release()
    s, b = d_s.cpu().numpy().view(np.uint64), d_b.cpu().numpy().view(np.uint64)
    assert np.array_equal(s[:1000], C.random_fr_chacha20(1000, C.SEED_SCALARS))
    check(got, oracle_affine(s, b))


@pytest.mark.parametrize("k", [20, 22, 26])
def test_sweep_sizes_match_the_oracle_directly(ctx, engine, k):
    """BASELINE configs[1] names a sweep 2^16 .. 2^26: the sizes between the small cases above and the 2^24 headline, and the top of
    the sweep (2^26: 4 GiB of bases, ~56 GB of window multiples, ~40 s of CPU), each compared with the C port itself."""
    import torch
    import h2b200
    n = 1 << k
    C.use_all_host_cores()
    d_b = h2b200.synth.device_bases(ctx, n, C.bench_t0() + 1000 * k)
    d_s = h2b200.synth.chacha_scalars_device(ctx, n, bytes([k]) * 32)
    desc = engine.get_base_descriptor(d_b)
    got = engine.msm_with_cached_base(d_s, desc)
    desc.release()
    s, b = d_s.cpu().numpy().view(np.uint64), d_b.cpu().numpy().view(np.uint64)
    del d_s, d_b
    torch.cuda.empty_cache()
    check(got, oracle_affine(s, b))


@pytest.mark.parametrize("kind", ["mostly_zero_or_one", "below_2p16"])
def test_non_uniform_scalars_at_2p22(ctx, engine, kind):
    """SURVEY.md 8(d)'s two non-uniform scalar sets at scale (the shapes real witnesses have: selector-like columns that are 90 % zeros
    and ones; small range-checked values): a handful of buckets take nearly every point (the long-bucket path), most windows are empty."""
    import torch
    import h2b200
    n = 1 << 22
    C.use_all_host_cores()
    rng = np.random.default_rng(5)
    if kind == "mostly_zero_or_one":
        s = C.random_fr_chacha20(n, bytes([9]) * 32)
        u = rng.random(n)
        s[u < 0.45] = 0
        s[(u >= 0.45) & (u < 0.90)] = C.mont_from_int(1, R.R_MOD)
    else:
        table = C.fr_array_from_ints(list(range(1 << 16)), R.R_MOD)
        s = np.ascontiguousarray(table[rng.integers(0, 1 << 16, size=n)])
    d_b = h2b200.synth.device_bases(ctx, n, C.bench_t0() + 77)
    b = d_b.cpu().numpy().view(np.uint64)
    desc = engine.get_base_descriptor(d_b)
    got_host = engine.msm_with_cached_base(s, desc)                                               # pageable host scalars, chunked
    got_dev = engine.msm_with_cached_base(torch.from_numpy(s.view(np.int64)).cuda(), desc)       # device-resident scalars
    desc.release()
    exp = oracle_affine(s, b)
    check(got_host, exp)
    check(got_dev, exp)


def test_release_after_destroy_is_harmless():
    """Garbage-collected hosts finalise in any order: releasing a descriptor of a context that is already gone, or destroying
    a context twice, must be refused by the library, not crash it (a registry of live contexts inside libh2b200.so)."""
    import h2b200
    lib = h2b200.load()
    c = h2b200.Context([0])
    desc = h2b200.B200Engine(c).get_base_descriptor(C.gen_bases(100, 1))
    raw = ctypes.c_void_p(c.handle.value)
    handle = desc.handle
    c.close()
    assert lib.h2b_bases_release(raw, handle) == -1          # H2B_E_INVALID, no use-after-free
    lib.h2b_destroy(raw)                                      # second destroy: ignored
    desc.handle = 0
