"""GPU: the fused assignment head (log double soft-max + matchability + filter_matches,
lightglue/lightglue.py:196-254) against the reference's own functions (fixtures) and the oracle.

Stated tolerance: the kernel computes sim with bf16 tensor-core MMAs (fp32 accumulate), exponentials with
ex2.approx and the soft-max normalisers by an online log-sum-exp, so
  * inputs already bf16-representable: >= 99.5 % of the rows get the reference's partner (differences are
    near-ties of the score and rows sitting on the threshold) and matching scores agree to 2e-4 absolute
    (scores are probabilities in [0, 1]);
  * fp32 inputs through AssignmentHead (projection in fp32, operands rounded to bf16): >= 97 % agreement,
    scores within 0.05.
"""
import os

import numpy as np
import pytest
import torch

from oracle import oracle_float

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, params=[0, 1], ids=["single_sweep", "sweep_per_direction"])
def float_sweep(pkg, request):
    """Both float kernels: the single sweep with the fused column maximum (default) and one sweep per direction."""
    pkg.set_option("float_sweep", request.param)
    yield request.param
    pkg.set_option("float_sweep", 0)


def bf16_bits_to_f32(x):
    return (x.astype(np.uint16).astype(np.uint32) << 16).view(np.float32)


def _run_pair(pkg, d0, d1, z0, z1, th):
    m, n = d0.shape[0], d1.shape[0]
    bank = pkg.FloatBank.from_arrays([d0, d1], device="cuda")
    z = torch.zeros((2, bank.k_stride), dtype=torch.float32, device="cuda")
    z[0, :m] = torch.as_tensor(np.asarray(z0), dtype=torch.float32)
    z[1, :n] = torch.as_tensor(np.asarray(z1), dtype=torch.float32)
    res = pkg.assign_pairs(bank, z, [[0, 1]], filter_threshold=th)
    res.check_overflow()
    q, t, s = res.to_lists()[0]
    return np.asarray(q), np.asarray(t), np.asarray(s)


def _compare(q, t, s, m0_exp, ms0_exp, min_agree=0.995, atol=2e-4):
    assert np.all(np.diff(q) > 0)                                    # ascending query index (lightglue.py:419-425)
    m0 = np.full(m0_exp.shape[0], -1, np.int64)
    m0[q] = t
    agree = float(np.mean(m0 == m0_exp))
    assert agree >= min_agree, agree
    same = (m0 == m0_exp) & (m0 > -1)
    sc = np.zeros(m0_exp.shape[0], np.float32)
    sc[q] = s
    assert np.allclose(sc[same], ms0_exp[same], atol=atol), float(np.abs(sc[same] - ms0_exp[same]).max())
    return agree


@pytest.mark.parametrize("dim", [128, 256])
def test_assign_fixture_from_reference(pkg, golden_dir, dim):
    g = np.load(os.path.join(golden_dir, f"assign_{dim}.npz"))
    d0, d1 = bf16_bits_to_f32(g["d0_bf16"]), bf16_bits_to_f32(g["d1_bf16"])
    for name, th in (("none", None), ("thr", float(g["threshold"]))):
        q, t, s = _run_pair(pkg, d0, d1, g["z0"], g["z1"], th)
        _compare(q, t, s, g[f"m0_{name}"], g[f"ms0_{name}"])
        assert len(q) > 100


@pytest.mark.parametrize("m,n,dim", [(1, 1, 128), (3, 700, 256), (700, 3, 128), (129, 257, 256), (2048, 2048, 256),
                                     (2048, 1999, 128), (300, 5000, 128)])
def test_assign_shapes_vs_oracle(pkg, m, n, dim):
    gen = torch.Generator().manual_seed(m * 17 + n)
    w = int(1.3 * max(m, n)) + 1
    world = torch.nn.functional.normalize(torch.randn(w, dim, generator=gen), dim=1)
    d0 = 4.5 * torch.nn.functional.normalize(world[torch.randperm(w, generator=gen)[:m]] + 0.04 * torch.randn(m, dim, generator=gen), dim=1)
    d1 = 4.5 * torch.nn.functional.normalize(world[torch.randperm(w, generator=gen)[:n]] + 0.04 * torch.randn(n, dim, generator=gen), dim=1)
    d0, d1 = d0.to(torch.bfloat16).float().numpy(), d1.to(torch.bfloat16).float().numpy()
    z0 = (1.0 + 1.5 * torch.randn(m, generator=gen)).numpy()
    z1 = (1.0 + 1.5 * torch.randn(n, generator=gen)).numpy()
    for th in (None, 0.1):
        q, t, s = _run_pair(pkg, d0, d1, z0, z1, th)
        _, _, _, (m0, _, ms0, _), _ = oracle_float.assign(d0, d1, z0, z1, th)
        _compare(q, t, s, m0, ms0)


def test_assign_batched_pairs_share_nothing(pkg):
    """Several pairs in one call (different sizes per image) equal the same pairs run one by one."""
    dim, sizes = 256, [500, 777, 1024, 300]
    gen = torch.Generator().manual_seed(5)
    world = torch.nn.functional.normalize(torch.randn(1500, dim, generator=gen), dim=1)
    descs = [(4.0 * world[torch.randperm(1500, generator=gen)[:k]]).to(torch.bfloat16).float() for k in sizes]
    zs = [torch.randn(k, generator=gen) for k in sizes]
    bank = pkg.FloatBank.from_arrays(descs, device="cuda")
    z = torch.zeros((len(sizes), bank.k_stride), dtype=torch.float32, device="cuda")
    for i, zi in enumerate(zs):
        z[i, :sizes[i]] = zi
    pairs = [[0, 1], [2, 3], [3, 0], [1, 2], [2, 2]]
    res = pkg.assign_pairs(bank, z, pairs, filter_threshold=0.05)
    res.check_overflow()
    for (a, b), (q, t, s) in zip(pairs, res.to_lists()):
        q1, t1, s1 = _run_pair(pkg, descs[a].numpy(), descs[b].numpy(), zs[a].numpy(), zs[b].numpy(), 0.05)
        assert np.array_equal(q, q1) and np.array_equal(t, t1) and np.allclose(s, s1, atol=1e-6)
        assert len(q) > 20


def test_assignment_head_module_vs_fp32_restatement(pkg):
    """AssignmentHead.forward (projection + fused head) against the reference formulas in fp32 torch on the
    same device: MatchAssignment.forward + filter_matches restated inline (lightglue.py:218-254)."""
    dim, b, m, n = 256, 2, 900, 1100
    gen = torch.Generator().manual_seed(11)
    head = pkg.AssignmentHead(dim, device="cuda", filter_threshold=0.1)
    # weights that give a peaked assignment: a scaled rotation-like projection
    wq, _ = torch.linalg.qr(torch.randn(dim, dim, generator=gen))
    head.load_state_dict({"final_proj.weight": 18.0 * wq, "final_proj.bias": 0.01 * torch.randn(dim, generator=gen),
                          "matchability.weight": torch.randn(1, dim, generator=gen), "matchability.bias": torch.ones(1)})
    world = torch.nn.functional.normalize(torch.randn(1600, dim, generator=gen), dim=1)
    desc0 = torch.stack([torch.nn.functional.normalize(world[torch.randperm(1600, generator=gen)[:m]] + 0.03 * torch.randn(m, dim, generator=gen), dim=1) for _ in range(b)])
    desc1 = torch.stack([torch.nn.functional.normalize(world[torch.randperm(1600, generator=gen)[:n]] + 0.03 * torch.randn(n, dim, generator=gen), dim=1) for _ in range(b)])
    out = head(desc0, desc1)
    md0, z0 = head.project(desc0)
    md1, z1 = head.project(desc1)
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        sim = torch.einsum("bmd,bnd->bmn", md0, md1)
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    F = torch.nn.functional
    scores = F.log_softmax(sim, 2) + F.log_softmax(sim, 1) + F.logsigmoid(z0)[:, :, None] + F.logsigmoid(z1)[:, None, :]
    max0, max1 = scores.max(2), scores.max(1)
    for i in range(b):
        a0, a1 = max0.indices[i], max1.indices[i]
        mutual0 = torch.arange(m, device="cuda") == a1[a0]
        ms0 = torch.where(mutual0, max0.values[i].exp(), torch.zeros((), device="cuda"))
        m0 = torch.where(mutual0 & (ms0 > 0.1), a0, torch.full_like(a0, -1))
        got = out["matches0"][i]
        agree = float((got == m0).float().mean())
        assert agree >= 0.97, agree
        same = (got == m0) & (m0 > -1)
        assert float((out["matching_scores0"][i][same] - ms0[same]).abs().max()) <= 0.05
        assert int((m0 > -1).sum()) > 200
        q, t = out["matches"][i][:, 0], out["matches"][i][:, 1]
        assert torch.equal(got[q], t) and torch.equal(out["matches1"][i][t], q)
        assert torch.all(out["scores"][i] > 0.1)


@pytest.mark.parametrize("below", [56.0, 62.0, 68.0])
def test_assign_row_maxima_near_the_edge_of_the_common_reference_range(pkg, below):
    """Every row but one has its largest score `below` nats under the pair's reference point (one long row and the
    longest column set it), with 2,048 columns: around 60 nats the single-sweep sums come close to where flushed
    exponentials (ex2.approx.ftz, < 2^-126 each, up to n of them) would stop being negligible. The accepted range
    scales with the number of terms (kLseTinyPerTerm); on either side of that edge -- single sweep kept, or batch redone
    with per-row maxima -- partners and scores must be the oracle's."""
    dim, m, n = 128, 1200, 2048
    gen = torch.Generator().manual_seed(123)
    world = torch.nn.functional.normalize(torch.randn(2600, dim, generator=gen), dim=1)
    d0 = 4.5 * torch.nn.functional.normalize(world[torch.randperm(2600, generator=gen)[:m]] + 0.04 * torch.randn(m, dim, generator=gen), dim=1)
    d1 = 4.5 * torch.nn.functional.normalize(world[torch.randperm(2600, generator=gen)[:n]] + 0.04 * torch.randn(n, dim, generator=gen), dim=1)
    d0[17] *= (20.0 + below) / 20.25                       # reference point = |d0[17]| * 4.5 = 20 + below; other rows peak near 20
    d0, d1 = d0.to(torch.bfloat16).float().numpy(), d1.to(torch.bfloat16).float().numpy()
    z0 = (1.0 + 1.5 * torch.randn(m, generator=gen)).numpy()
    z1 = (1.0 + 1.5 * torch.randn(n, generator=gen)).numpy()
    for th in (None, 0.1):
        q, t, s = _run_pair(pkg, d0, d1, z0, z1, th)
        _, _, _, (m0, _, ms0, _), _ = oracle_float.assign(d0, d1, z0, z1, th)
        _compare(q, t, s, m0, ms0, min_agree=0.99)
        assert len(q) > 100


@pytest.mark.parametrize("case", ["all_rows_long", "one_long_row"])
def test_assign_scores_outside_the_common_reference_range(pkg, case):
    """The single-sweep log-sum-exp measures every exponential of a pair from ONE reference point (the product of the
    longest row norms). Rows whose largest score lies far below it (> ~61-69 nats) would lose their sums to the fp32
    range: the kernel flags that and the batch is redone with per-row maxima. Long descriptors (|d| = 30: scores up
    to 900) and a single outlier row that inflates the reference must give the oracle's result all the same."""
    dim, m, n = 128, 600, 520
    gen = torch.Generator().manual_seed(77)
    world = torch.nn.functional.normalize(torch.randn(900, dim, generator=gen), dim=1)
    d0 = torch.nn.functional.normalize(world[torch.randperm(900, generator=gen)[:m]] + 0.04 * torch.randn(m, dim, generator=gen), dim=1)
    d1 = torch.nn.functional.normalize(world[torch.randperm(900, generator=gen)[:n]] + 0.04 * torch.randn(n, dim, generator=gen), dim=1)
    if case == "all_rows_long":
        d0, d1 = 30.0 * d0, 30.0 * d1
    else:
        d0, d1 = 4.5 * d0, 4.5 * d1
        d0[17] *= 40.0                                     # |d| = 180: the reference point moves to 810
    d0, d1 = d0.to(torch.bfloat16).float().numpy(), d1.to(torch.bfloat16).float().numpy()
    z0 = (1.0 + 1.5 * torch.randn(m, generator=gen)).numpy()
    z1 = (1.0 + 1.5 * torch.randn(n, generator=gen)).numpy()
    for th in (None, 0.1):
        q, t, s = _run_pair(pkg, d0, d1, z0, z1, th)
        _, _, _, (m0, _, ms0, _), _ = oracle_float.assign(d0, d1, z0, z1, th)
        _compare(q, t, s, m0, ms0, min_agree=0.99)
        assert len(q) > 50


def test_assign_is_repeatable_bit_for_bit(pkg):
    """Two calls on the same inputs give the same lists and the same float32 score bits: the column sums of the
    single-sweep log-sum-exp are combined in a fixed order (lane quadrants in shared memory, row blocks in the
    finalize kernel), never with floating-point atomics."""
    dim, sizes = 128, [1500, 1300, 2048, 700, 1024, 1800]
    gen = torch.Generator().manual_seed(9)
    world = torch.nn.functional.normalize(torch.randn(2600, dim, generator=gen), dim=1)
    descs = [(4.5 * torch.nn.functional.normalize(world[torch.randperm(2600, generator=gen)[:k]] + 0.04 * torch.randn(k, dim, generator=gen), dim=1))
             .to(torch.bfloat16).float() for k in sizes]
    bank = pkg.FloatBank.from_arrays(descs, device="cuda")
    z = torch.zeros((len(sizes), bank.k_stride), dtype=torch.float32, device="cuda")
    for i, k in enumerate(sizes):
        z[i, :k] = 1.0 + torch.randn(k, generator=gen)
    pairs = [[a, b] for a in range(len(sizes)) for b in range(len(sizes)) if a != b]
    runs = []
    for _ in range(3):
        res = pkg.assign_pairs(bank, z, pairs, filter_threshold=0.05)
        res.check_overflow()
        runs.append(res.to_lists())
    total = 0
    for other in runs[1:]:
        for (q0, t0, s0), (q1, t1, s1) in zip(runs[0], other):
            assert np.array_equal(q0, q1) and np.array_equal(t0, t1)
            assert np.array_equal(np.asarray(s0, np.float32).view(np.uint32), np.asarray(s1, np.float32).view(np.uint32))
            total += len(q0)
    assert total > 5000
