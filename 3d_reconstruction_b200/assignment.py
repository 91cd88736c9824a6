"""The reference matcher's assignment head (lightglue/lightglue.py:196-254) without its M x N matrices.

MatchAssignment.forward (:218-227) projects both descriptor sets, forms sim = mdesc0 . mdesc1 (:223),
turns it into a log assignment matrix with a row soft-max, a column soft-max and the matchability
logits (sigmoid_log_double_softmax, :196-208), and filter_matches (:235-254) keeps mutual arg-max pairs
whose exp(score) exceeds filter_threshold. The reference materialises sim, two log-soft-max copies and
the (M+1) x (N+1) score matrix; here TWO tensor-core sweeps per pair keep everything on chip (option "float_sweep" = 0,
the default; "float_sweep" = 1 runs one sweep per direction and pass, four per pair):
  pass 1  log-sum-exp of every row AND every column of sim from one sweep (= row and column normalisers)
  pass 2  row arg-max and column maxima of 2 sim - lse_row - lse_col + logsigmoid(z0) + logsigmoid(z1)
then the mutual check / threshold / match list (b200m_assign_pairs_bf16, include/b200m.h).

  * assign_pairs      batched: projected descriptors (FloatBank) + matchability logits + pair list
  * AssignmentHead    MatchAssignment-shaped module: holds final_proj / matchability weights (state-dict
                      compatible with `log_assignment.<i>.*`), forward(desc0, desc1) -> the reference's
                      filter_matches outputs (m0, m1, mscores0, mscores1) + the match list of :417-426.
The two small projections are plain library GEMMs (torch / cuBLAS); operands are rounded to bf16.
"""
from __future__ import annotations

import torch

from . import _lib
from .bank import TILE_T, FloatBank, round_up
from .matcher import PairMatches, _check_out, _pairs_tensor


def assign_pairs(bank: FloatBank, matchability: torch.Tensor, pairs, filter_threshold: float | None = 0.1,
                 cap: int | None = None, out: PairMatches | None = None,
                 workspace: torch.Tensor | None = None) -> PairMatches:
    """bank: projected descriptors mdesc (bf16 [n_img, k_stride, D]); matchability: float32 [n_img, k_stride]
    logits z. -> per pair the mutual arg-max matches of the log assignment matrix, ascending in query index;
    `dist` of the result holds exp(score) (matching_scores0 of the reference). filter_threshold=None keeps
    every mutual pair (lightglue.py:247-250)."""
    dev = bank.device
    z = matchability
    if not z.is_cuda or z.dtype != torch.float32 or tuple(z.shape) != (bank.n_img, bank.k_stride) or not z.is_contiguous():
        raise ValueError("matchability must be a contiguous float32 CUDA tensor [n_img, k_stride]")
    pairs_t = _pairs_tensor(pairs, dev, bank.n_img)
    P = int(pairs_t.shape[0])
    if cap is None:
        cap = bank.k_stride
    if out is None:
        out = PairMatches(torch.empty((P, cap), dtype=torch.int32, device=dev),
                          torch.empty((P, cap), dtype=torch.int32, device=dev),
                          torch.empty((P, cap), dtype=torch.float32, device=dev),
                          torch.zeros((P,), dtype=torch.int32, device=dev), int(cap))
    else:
        _check_out(out, P)
    if P == 0:
        return out
    L = _lib.lib()
    if workspace is None:
        workspace = torch.empty(int(L.b200m_workspace_bytes_assign(bank.k_stride, P)), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        st = torch.cuda.current_stream().cuda_stream
        _lib.check(L.b200m_assign_pairs_bf16(
            bank.desc.data_ptr(), z.data_ptr(), bank.k_per_image.data_ptr(), bank.n_img, bank.k_stride, bank.dim,
            pairs_t.data_ptr(), P, float(filter_threshold if filter_threshold is not None else 0.0),
            int(filter_threshold is not None),
            out.query_idx.data_ptr(), out.train_idx.data_ptr(), out.dist.data_ptr(), out.count.data_ptr(),
            int(out.cap), workspace.data_ptr(), workspace.numel(), st))
    return out


class AssignmentHead:
    """MatchAssignment (lightglue.py:211-232) + filter_matches (:235-254) as one call.

    Weights: final_proj (Linear(dim, dim)) and matchability (Linear(dim, 1)); load_state_dict accepts the
    reference's keys ('final_proj.weight', 'final_proj.bias', 'matchability.weight', 'matchability.bias')."""

    def __init__(self, dim: int, device="cuda", filter_threshold: float | None = 0.1):
        if dim not in (128, 256):
            raise ValueError("descriptor length must be 128 or 256")
        self.dim = dim
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("the assignment head runs on a CUDA device (there is no CPU path)")
        self.filter_threshold = filter_threshold
        g = torch.Generator().manual_seed(0)
        bound = dim ** -0.5                           # nn.Linear's default init range
        self.final_proj_weight = ((torch.rand(dim, dim, generator=g) * 2 - 1) * bound).to(self.device)
        self.final_proj_bias = ((torch.rand(dim, generator=g) * 2 - 1) * bound).to(self.device)
        self.matchability_weight = ((torch.rand(1, dim, generator=g) * 2 - 1) * bound).to(self.device)
        self.matchability_bias = ((torch.rand(1, generator=g) * 2 - 1) * bound).to(self.device)

    def load_state_dict(self, sd: dict, prefix: str = ""):
        for name, attr in (("final_proj.weight", "final_proj_weight"), ("final_proj.bias", "final_proj_bias"),
                           ("matchability.weight", "matchability_weight"), ("matchability.bias", "matchability_bias")):
            t = sd[prefix + name].detach().to(device=self.device, dtype=torch.float32)
            if t.shape != getattr(self, attr).shape:
                raise ValueError(f"{name}: shape {tuple(t.shape)} does not fit dim {self.dim}")
            setattr(self, attr, t.contiguous())
        return self

    def project(self, desc: torch.Tensor):
        """desc [..., K, dim] float -> (mdesc = final_proj(desc) / dim^0.25 (:220-222), z = matchability(desc) (:224))."""
        d = desc.to(device=self.device, dtype=torch.float32)
        mdesc = torch.nn.functional.linear(d, self.final_proj_weight, self.final_proj_bias) / self.dim ** 0.25
        z = torch.nn.functional.linear(d, self.matchability_weight, self.matchability_bias).squeeze(-1)
        return mdesc, z

    def scores(self, desc0: torch.Tensor, desc1: torch.Tensor):
        """MatchAssignment.scores (:229-232): sigmoid of the matchability logits."""
        return torch.sigmoid(self.project(desc0)[1]), torch.sigmoid(self.project(desc1)[1])

    def forward(self, desc0: torch.Tensor, desc1: torch.Tensor) -> dict:
        """desc0 [B, M, dim], desc1 [B, N, dim] -> {'matches0' [B,M], 'matches1' [B,N], 'matching_scores0',
        'matching_scores1', 'matches': [LongTensor [S,2]] * B, 'scores': [FloatTensor [S]] * B} -- what
        LightGlue._forward builds from filter_matches at :412-426 and returns at :442-450."""
        assert desc0.dim() == 3 and desc1.dim() == 3 and desc0.shape[0] == desc1.shape[0]
        assert desc0.shape[-1] == self.dim and desc1.shape[-1] == self.dim
        b, m, n = desc0.shape[0], desc0.shape[1], desc1.shape[1]
        dev = self.device
        ks = round_up(max(m, n, 1), TILE_T)
        md0, z0 = self.project(desc0)
        md1, z1 = self.project(desc1)
        bank = torch.zeros((2 * b, ks, self.dim), dtype=torch.bfloat16, device=dev)   # image 2i = desc0[i], 2i+1 = desc1[i]
        zz = torch.zeros((2 * b, ks), dtype=torch.float32, device=dev)
        bank[0::2, :m] = md0.to(torch.bfloat16)
        bank[1::2, :n] = md1.to(torch.bfloat16)
        zz[0::2, :m] = z0
        zz[1::2, :n] = z1
        kpi = torch.tensor([m, n] * b, dtype=torch.int32)
        pairs = torch.arange(2 * b, dtype=torch.int32).reshape(b, 2)
        # every mutual pair, the threshold is applied below: mscores0 / mscores1 are not thresholded (:245-246)
        res = assign_pairs(FloatBank(bank, kpi), zz, pairs, filter_threshold=None)
        res.check_overflow()
        m0 = torch.full((b, m), -1, dtype=torch.long, device=dev)
        m1 = torch.full((b, n), -1, dtype=torch.long, device=dev)
        ms0 = torch.zeros((b, m), device=dev)
        ms1 = torch.zeros((b, n), device=dev)
        matches, mscores = [], []
        counts = res.count.tolist()
        for i in range(b):
            s = counts[i]
            q = res.query_idx[i, :s].long()
            t = res.train_idx[i, :s].long()
            sc = res.dist[i, :s]
            ms0[i, q] = sc
            ms1[i, t] = sc
            keep = sc > self.filter_threshold if self.filter_threshold is not None else torch.ones_like(sc, dtype=torch.bool)
            q, t, sc = q[keep], t[keep], sc[keep]
            m0[i, q] = t
            m1[i, t] = q
            matches.append(torch.stack([q, t], -1))
            mscores.append(sc)
        return {"matches0": m0, "matches1": m1, "matching_scores0": ms0, "matching_scores1": ms1,
                "matches": matches, "scores": mscores}

    __call__ = forward
