"""GPU: the host-buffer entry points (the e2e leg of bench.py) give exactly the lists of the resident-bank call.
  * matcher.match_pairs_host: chunked upload on a copy stream, every image block prepared as it lands, pair
    segments released as soon as their images are resident;
  * the range forms of bank prepare / sort (b200m_bank_prepare_range_u8 / b200m_bank_sort_range_u8) fill the
    tables of the whole bank, whatever the order of the ranges."""
import importlib

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _bank(n_img=37, k=600, seed=11):
    syn = importlib.import_module("3d_reconstruction_b200.synthetic")
    rng = np.random.default_rng(seed)
    ks = rng.integers(200, k + 1, size=n_img).tolist()
    ks[5] = 0
    ks[9] = 1
    return syn.make_bank_tensor(n_img, k, seed=seed, device="cpu", k_per_image=ks), ks


def test_range_prepare_equals_whole_bank(pkg):
    (desc, kpi), ks = _bank()
    whole = pkg.DescriptorBank.from_packed(desc.cuda(), kpi)
    part = pkg.DescriptorBank.allocate(desc.shape[0], desc.shape[1], kpi, "cuda")
    part.desc.copy_(desc)
    for lo, cnt in ((30, 7), (0, 5), (12, 18), (5, 7)):          # out of order, covering every image once
        part.prepare_range(lo, cnt)
    torch.cuda.synchronize()
    assert torch.equal(part.desc, whole.desc) and torch.equal(part.sqnorm, whole.sqnorm)
    assert torch.equal(part.sorted_desc, whole.sorted_desc)
    for a, b in zip(part.sort_views(), whole.sort_views()):
        assert torch.equal(a, b)
    for (a, _), (b, _) in zip(part.table_sections(), whole.table_sections()):
        assert torch.equal(a, b)


@pytest.mark.parametrize("order", ["blocked", "row_major", "reversed"])
def test_match_pairs_host_equals_resident_bank(pkg, order):
    (desc, kpi), ks = _bank()
    n_img = desc.shape[0]
    pairs = pkg.exhaustive_pairs(n_img)
    if order == "blocked":
        pairs = pkg.blocked_order(pairs, 8)
    elif order == "reversed":
        pairs = pairs[::-1].copy()
    bank = pkg.DescriptorBank.from_packed(desc.cuda(), kpi)
    ref = pkg.match_pairs(bank, pairs, cap=700).flatten().cpu()
    host = desc.pin_memory()
    for staging in (False, True):
        got = pkg.match_pairs_host(host, kpi.numpy(), pairs, cap=700, block=8, staging=staging)
        assert torch.equal(got.offsets, ref.offsets)
        assert torch.equal(got.query_idx, ref.query_idx) and torch.equal(got.train_idx, ref.train_idx)
        assert torch.equal(got.dist.view(torch.int32), ref.dist.view(torch.int32))
        assert got.digest()[1] == ref.digest()[1]
    assert int(ref.offsets[-1]) > 1000
    # pageable host memory and a block size that does not divide the bank work too
    got = pkg.match_pairs_host(desc.clone(), kpi.numpy(), pairs, cap=700, block=5)
    assert torch.equal(got.query_idx, ref.query_idx) and torch.equal(got.offsets, ref.offsets)
    with pytest.raises(ValueError):
        pkg.match_pairs_host(host, kpi.numpy(), np.array([[0, n_img]]), cap=700)
    empty = pkg.match_pairs_host(host, kpi.numpy(), np.zeros((0, 2), np.int32), cap=700)
    assert empty.offsets.tolist() == [0]
