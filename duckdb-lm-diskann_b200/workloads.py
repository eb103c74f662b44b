"""Synthetic workloads of BASELINE.json configs C2-C5 generated ON the GPU (bench.py's `other_configs`, scripts/
shard_bench.py). Measurement plumbing, not the query path: data sets come from a torch CUDA generator (the same
clustered-latent recipe as fixtures.make_dataset), graphs from lmd_build_graph (row f1), ground
truth from a tiled fp32 matmul - so nothing large crosses PCIe and a 100 M-row configuration is ready in seconds.
"""
import numpy as np
import torch


def gen(n, dims, seed, device, normalize=False, kind="clustered", latent=16, noise=0.05, chunk=1 << 20):
    """[n, dims] float32 on `device`. clustered: x = z P + noise * eps with the fixed projection of fixtures.make_dataset
    (numpy seed 7); iid: x ~ N(0, I) (SURVEY.md section 8d's hard case)."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    out = torch.empty((n, dims), dtype=torch.float32, device=device)
    if kind == "clustered":
        a = np.random.default_rng(7).standard_normal((dims, latent))
        q, _ = np.linalg.qr(a)
        proj = torch.from_numpy(q.T.astype(np.float32)).to(device)
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        if kind == "clustered":
            z = torch.randn((e - s, latent), generator=g, device=device)
            out[s:e] = z @ proj + noise * torch.randn((e - s, dims), generator=g, device=device)
        elif kind == "iid":
            out[s:e] = torch.randn((e - s, dims), generator=g, device=device)
        else:
            raise ValueError(kind)
    if normalize:
        out /= out.norm(dim=1, keepdim=True).clamp_min(1e-12)
    return out


def topk_scores(x, q, k, metric, base_id=0, q_tile=2500, b_tile=1 << 19):
    """(scores [nq,k], ids [nq,k]) of the k best rows of x per query; score is monotone in the metric's distance and
    comparable across shards (L2: |x|^2 - 2 q.x; IP / COSINE with unit rows: -q.x)."""
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    nq = q.shape[0]
    S = torch.empty((nq, k), device=q.device)
    I = torch.empty((nq, k), dtype=torch.int64, device=q.device)
    try:
        for q0 in range(0, nq, q_tile):
            a = q[q0:q0 + q_tile]
            bs = bi = None
            for b0 in range(0, x.shape[0], b_tile):
                b = x[b0:b0 + b_tile]
                s = a @ b.T
                s = (b * b).sum(1)[None, :] - 2.0 * s if metric == "L2" else -s
                ts, ti = torch.topk(s, min(k, b.shape[0]), dim=1, largest=False)
                ti = ti + (b0 + base_id)
                if bs is None:
                    bs, bi = ts, ti
                else:
                    cs, ci = torch.cat([bs, ts], 1), torch.cat([bi, ti], 1)
                    bs, sel = torch.topk(cs, k, dim=1, largest=False)
                    bi = torch.gather(ci, 1, sel)
            S[q0:q0 + q_tile], I[q0:q0 + q_tile] = bs, bi
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old
    return S, I


def recall_hits(found, truth):
    """number of (query, rank) pairs of `truth` [nq,k] that appear in `found` [nq,k] (device tensors)"""
    f = found.unsqueeze(2) == truth.unsqueeze(1)
    return int(f.any(dim=1).sum())


def recall(found, truth):
    return recall_hits(found, truth) / float(truth.numel())


def build_store(lmd, x, r, metric, device_index, row_ids=None, l_insert=64, alpha=1.2, l_search=100):
    """Node store over the rows of x (a device tensor) with a graph built by lmd_build_graph -> (Store, BuildStats)."""
    store = lmd.Store(x.shape[1], r, metric=lmd.METRIC_BY_NAME[metric], capacity=x.shape[0], device=device_index,
                      l_search=l_search)
    store.upload_vectors(x, row_ids=row_ids)
    bst = store.build_graph(l_insert=l_insert, alpha=alpha)
    return store, bst
