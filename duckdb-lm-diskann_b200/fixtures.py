"""Synthetic LM_DISKANN indexes for tests, smoke() and bench.py: data sets, a batch Vamana-style graph
builder and the node-block writer. NOT on the query path ("graph provenance is
free" - parity is always "same graph, same queries" between the CUDA path and the oracle).

Everything is torch so that the same code builds the 2 000-node graphs of the CPU tests here and the
1 M-node graph of the bench on the GPU box (device="cuda").

The block writer follows the reference's on-disk node layout bit for bit
(core/index_config.cpp:104-148; empty id slot = INT64_MAX, core/GraphManager.cpp:154-156,451-453;
ternary planes: bit d of word d/64, set in `pos` when x_d > 0 and in `neg` when x_d < 0,
core/ternary_quantization.hpp:146-199; planes are two SoA arrays [R][words] u64 as index_config.cpp:136-143
sizes them). The graph builder applies the reference's pruning rule (core/GraphManager.cpp:300-370: sort by
distance, keep p unless alpha * d(p, r) < d(node, p) for an already selected r).
"""
import numpy as np
import torch

INT64_MAX = np.iinfo(np.int64).max


# ----------------------------------------------------------------------------------------------
# data sets
# ----------------------------------------------------------------------------------------------
def make_dataset(n, dims, kind="clustered", seed=42, latent=16, noise=0.05, normalize=False, chunk=1 << 20):
    """clustered: x = z P + noise * eps, z ~ N(0, I_latent), P a fixed random orthonormal [latent, dims]
    (seed 7); iid: x ~ N(0, I_dims). Base and query sets use different seeds and the same P."""
    out = np.empty((n, dims), dtype=np.float32)
    rng = np.random.default_rng(seed)
    if kind == "clustered":
        prng = np.random.default_rng(7)
        a = prng.standard_normal((dims, latent))
        q, _ = np.linalg.qr(a)  # [dims, latent], orthonormal columns
        proj = q.T.astype(np.float32)
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        if kind == "clustered":
            z = rng.standard_normal((e - s, latent), dtype=np.float32)
            out[s:e] = z @ proj + noise * rng.standard_normal((e - s, dims), dtype=np.float32)
        elif kind == "iid":
            out[s:e] = rng.standard_normal((e - s, dims), dtype=np.float32)
        else:
            raise ValueError(kind)
    if normalize:
        out /= np.maximum(np.linalg.norm(out, axis=1, keepdims=True), 1e-12)
    return out


# ----------------------------------------------------------------------------------------------
# brute force (ground truth and kNN candidates)
# ----------------------------------------------------------------------------------------------
def _pair_scores(a, b, bn, metric):
    """smaller = closer; monotone in the reference's distance for each metric (used for ordering only)."""
    dot = a @ b.T
    if metric == "L2":
        return bn[None, :] - 2.0 * dot
    if metric == "IP":
        return -dot
    if metric == "COSINE":
        return -dot / torch.clamp(bn[None, :].sqrt(), min=1e-30)
    raise ValueError(metric)


def brute_force_topk(base, queries, k, metric="L2", device=None, exclude_self=False, q_tile=None, b_tile=1 << 18):
    """ids [nq, k] of the k closest base rows per query (exact fp32 matmul, tiled). exclude_self: queries ARE
    base[0:nq+offset] rows and a row must not return itself (kNN-graph use) - pass the offset as an int."""
    device = device or ("cuda" if torch.cuda.is_available() else "cpu")
    base_t = torch.as_tensor(base)
    n = base_t.shape[0]
    qs = torch.as_tensor(queries)
    nq = qs.shape[0]
    if q_tile is None:
        q_tile = 4096 if device != "cpu" else 1024
    old_tf32 = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    out = torch.empty((nq, k), dtype=torch.int64)
    try:
        base_dev = base_t.to(device) if n * base_t.shape[1] * 4 < (24 << 30) else None
        for qs0 in range(0, nq, q_tile):
            qs1 = min(nq, qs0 + q_tile)
            a = qs[qs0:qs1].to(device)
            best_s = None
            best_i = None
            for b0 in range(0, n, b_tile):
                b1 = min(n, b0 + b_tile)
                b = base_dev[b0:b1] if base_dev is not None else base_t[b0:b1].to(device)
                bn = (b * b).sum(1)
                s = _pair_scores(a, b, bn, metric)
                if exclude_self is not False:
                    off = int(exclude_self) if exclude_self is not True else 0
                    rows = torch.arange(qs0, qs1, device=device) + off
                    inside = (rows >= b0) & (rows < b1)
                    if bool(inside.any()):
                        r_idx = torch.nonzero(inside).squeeze(1)
                        s[r_idx, rows[r_idx] - b0] = float("inf")
                kk = min(k, b1 - b0)
                ts, ti = torch.topk(s, kk, dim=1, largest=False, sorted=True)
                ti = ti + b0
                if best_s is None:
                    best_s, best_i = ts, ti
                else:
                    cs = torch.cat([best_s, ts], 1)
                    ci = torch.cat([best_i, ti], 1)
                    kk2 = min(k, cs.shape[1])
                    best_s, sel = torch.topk(cs, kk2, dim=1, largest=False, sorted=True)
                    best_i = torch.gather(ci, 1, sel)
            if best_i.shape[1] < k:
                pad = torch.full((best_i.shape[0], k - best_i.shape[1]), -1, dtype=torch.int64, device=best_i.device)
                best_i = torch.cat([best_i, pad], 1)
            out[qs0:qs1] = best_i.cpu()
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old_tf32
    return out.numpy()


def recall_at_k(found_ids, truth_ids):
    """mean over queries of |found ∩ truth| / k"""
    k = truth_ids.shape[1]
    hits = 0
    for f, t in zip(found_ids, truth_ids):
        hits += len(set(int(x) for x in f[:k] if x >= 0) & set(int(x) for x in t))
    return hits / float(k * len(truth_ids))


# ----------------------------------------------------------------------------------------------
# graph builder: kNN candidates -> alpha prune -> reverse edges -> a few random long edges
# ----------------------------------------------------------------------------------------------
def _alpha_prune(x_dev, cand, alpha, keep, batch):
    """cand [n, C] candidate slots sorted by distance from the node, ascending. Returns [n, keep] slots, -1
    padded. Rule of core/GraphManager.cpp:335-370: candidate p survives unless alpha * d(p, r) < d(node, p)
    for an r selected before it; stop at `keep` neighbours. True L2 distances (sqrt) as the reference uses."""
    n, C = cand.shape
    out = torch.full((n, keep), -1, dtype=torch.int64, device=cand.device)
    for s in range(0, n, batch):
        e = min(n, s + batch)
        c = cand[s:e]
        cv = x_dev[c]  # [B, C, D]
        node = x_dev[s:e].unsqueeze(1)
        d_np = (cv - node).square().sum(-1).sqrt()  # [B, C]
        g = torch.bmm(cv, cv.transpose(1, 2))
        sq = cv.square().sum(-1)
        d_cc = (sq.unsqueeze(2) + sq.unsqueeze(1) - 2.0 * g).clamp_(min=0).sqrt_()  # [B, C, C]
        del g, cv
        alive = c >= 0
        nsel = torch.zeros(e - s, dtype=torch.int64, device=cand.device)
        sel = torch.zeros_like(alive)
        for i in range(C):
            take = alive[:, i] & (nsel < keep)
            sel[:, i] = take
            nsel += take
            # r = candidate i (if taken) prunes every later p with alpha * d(p, r) < d(node, p)
            kill = (alpha * d_cc[:, i, :] < d_np) & take.unsqueeze(1)
            kill[:, : i + 1] = False
            alive &= ~kill
        # compact the selected candidates to the front, order preserved
        rank = torch.cumsum(sel, 1) - 1
        rows = torch.arange(e - s, device=cand.device).unsqueeze(1).expand_as(c)
        o = out[s:e]
        o[rows[sel], rank[sel]] = c[sel]
    return out


def build_graph(x, r, alpha=1.2, n_cand=None, n_random=None, seed=1234, device=None, prune_batch=None, log=None):
    """Returns (nbrs [n, r] int64 slots, -1 = empty; counts [n]). Every node gets: up to r - n_random - reserve
    alpha-pruned near neighbours, reverse edges of those while room remains, then n_random random long edges.
    Neighbour lists are duplicate-free and never contain the node itself."""
    device = device or ("cuda" if torch.cuda.is_available() else "cpu")
    x_t = torch.as_tensor(x)
    n, dims = x_t.shape
    if n_random is None:
        n_random = max(2, r // 8)
    if n_cand is None:
        n_cand = min(max(2 * r, 48), 96)
    n_cand = min(n_cand, n - 1)
    keep = max(1, r - n_random - r // 4)  # leave r/4 slots for reverse edges
    if prune_batch is None:
        prune_batch = 8192 if device != "cpu" else 1024
    gen = torch.Generator(device="cpu")
    gen.manual_seed(seed)
    if log:
        log("graph: kNN candidates (n=%d, C=%d)" % (n, n_cand))
    knn = torch.as_tensor(brute_force_topk(x_t, x_t, n_cand, "L2", device=device, exclude_self=True)).to(device)
    x_dev = x_t.to(device)
    if log:
        log("graph: alpha prune (alpha=%.2f, keep=%d)" % (alpha, keep))
    fwd = _alpha_prune(x_dev, knn, alpha, keep, prune_batch)  # [n, keep]
    del knn
    nbrs = torch.full((n, r), -1, dtype=torch.int64, device=device)
    nbrs[:, :keep] = fwd
    counts = (fwd >= 0).sum(1)
    # reverse edges: (u -> v) forward gives candidate (v -> u); accept in random order while v has room
    if log:
        log("graph: reverse edges")
    u = torch.arange(n, device=device).unsqueeze(1).expand_as(fwd)[fwd >= 0]
    v = fwd[fwd >= 0]
    fwd_keys = u * n + v
    rev_keys = v * n + u
    fresh = ~torch.isin(rev_keys, fwd_keys)  # v does not already point at u
    rv, ru = v[fresh], u[fresh]
    perm = torch.randperm(rv.numel(), generator=gen).to(device)
    rv, ru = rv[perm], ru[perm]
    order = torch.sort(rv, stable=True).indices
    rv, ru = rv[order], ru[order]
    first = torch.searchsorted(rv, torch.arange(n, device=device))
    rank = torch.arange(rv.numel(), device=device) - first[rv]
    room = (r - n_random) - counts
    ok = rank < room[rv]
    rv, ru, rank = rv[ok], ru[ok], rank[ok]
    nbrs[rv, counts[rv] + rank] = ru
    counts = counts + torch.bincount(rv, minlength=n)
    # random long edges (skip a draw that hits the node itself or an existing neighbour)
    if log:
        log("graph: random edges")
    rnd = torch.randint(0, n, (n, n_random), generator=gen).to(device)
    self_idx = torch.arange(n, device=device).unsqueeze(1)
    for j in range(n_random):
        cand = rnd[:, j]
        dup = (nbrs == cand.unsqueeze(1)).any(1) | (cand == self_idx.squeeze(1))
        okj = ~dup & (counts < r)
        rows = torch.nonzero(okj).squeeze(1)
        nbrs[rows, counts[rows]] = cand[rows]
        counts = counts + okj
    return nbrs.cpu(), counts.cpu()


# ----------------------------------------------------------------------------------------------
# node-block writer (the on-disk contract, core/index_config.cpp:104-148)
# ----------------------------------------------------------------------------------------------
def layout(dims, r, vec_type=1):
    """Pure-python CalculateLayoutInternal; tests check it against the C ABI, the C oracle and the reference."""
    elem = 4 if vec_type == 1 else 1
    words = (dims + 63) // 64
    a8 = lambda v: (v + 7) // 8 * 8
    lay = {"count_off": 0}
    off = a8(2)
    lay["vector_off"] = off
    off = a8(off + elem * dims)
    lay["ids_off"] = off
    off = a8(off + 8 * r)
    lay["pos_off"] = off
    off = a8(off + r * words * 8)
    lay["neg_off"] = off
    off += r * words * 8
    lay["total_node_size"] = off
    lay["ternary_edge_bytes"] = 2 * words * 8
    lay["words"] = words
    lay["block_bytes"] = (off + 4095) // 4096 * 4096
    return lay


def ternary_planes(x_t):
    """[n, dims] float -> (pos, neg) [n, words] int64 bit planes (EncodeTernary)."""
    n, dims = x_t.shape
    words = (dims + 63) // 64
    pad = words * 64 - dims
    sh = torch.arange(64, dtype=torch.int64, device=x_t.device)

    def pack(mask):
        if pad:
            mask = torch.cat([mask, torch.zeros((n, pad), dtype=torch.bool, device=mask.device)], 1)
        return (mask.view(n, words, 64).to(torch.int64) << sh).sum(-1)  # bit 63 wraps into the sign, as intended

    return pack(x_t > 0), pack(x_t < 0)


def pack_blocks(x, nbrs, counts, r, row_ids=None, vec_type=1, device="cpu", chunk=1 << 16, block_bytes=None):
    """Writes n node blocks in the reference layout. nbrs holds SLOTS (-1 = empty); ids stored in the block are
    row_ids[slot] (row id = slot when row_ids is None). Returns a uint8 tensor [n, block_bytes] on `device`."""
    x_t = torch.as_tensor(x).to(device)
    nb = torch.as_tensor(nbrs).to(device)
    cnt = torch.as_tensor(counts).to(device)
    n, dims = x_t.shape
    lay = layout(dims, r, vec_type)
    bb = block_bytes or lay["block_bytes"]
    words = lay["words"]
    rid = None if row_ids is None else torch.as_tensor(np.asarray(row_ids, dtype=np.int64)).to(device)
    blocks = torch.zeros((n, bb), dtype=torch.uint8, device=device)
    pos, neg = ternary_planes(x_t.float())
    if vec_type == 1:
        vec_bytes = x_t.to(torch.float32).contiguous().view(torch.uint8).view(n, dims * 4)
    else:
        vec_bytes = x_t.to(torch.int8).contiguous().view(torch.uint8).view(n, dims)
    blocks[:, lay["vector_off"] : lay["vector_off"] + vec_bytes.shape[1]] = vec_bytes
    c16 = cnt.to(torch.int16).contiguous().view(torch.uint8).view(n, 2)
    blocks[:, 0:2] = c16
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        nbc = nb[s:e]
        valid = nbc >= 0
        safe = nbc.clamp(min=0)
        ids = safe if rid is None else rid[safe]
        ids = torch.where(valid, ids, torch.full_like(ids, INT64_MAX))
        blocks[s:e, lay["ids_off"] : lay["ids_off"] + 8 * r] = ids.contiguous().view(torch.uint8).view(e - s, 8 * r)
        pp = torch.where(valid.unsqueeze(2), pos[safe], torch.zeros((), dtype=torch.int64, device=device))
        nn_ = torch.where(valid.unsqueeze(2), neg[safe], torch.zeros((), dtype=torch.int64, device=device))
        nbytes = r * words * 8
        blocks[s:e, lay["pos_off"] : lay["pos_off"] + nbytes] = pp.contiguous().view(torch.uint8).view(e - s, nbytes)
        blocks[s:e, lay["neg_off"] : lay["neg_off"] + nbytes] = nn_.contiguous().view(torch.uint8).view(e - s, nbytes)
    return blocks


def build_index(n, dims, r, metric="L2", kind="clustered", seed=42, device=None, alpha=1.2, row_ids=None, log=None,
                blocks_device=None, n_random=None, n_cand=None):
    """Data set + graph + blocks in one call. Returns dict(x, nbrs, counts, blocks (torch uint8), layout)."""
    x = make_dataset(n, dims, kind=kind, seed=seed, normalize=(metric == "COSINE"))
    nbrs, counts = build_graph(x, r, alpha=alpha, device=device, log=log, n_random=n_random, n_cand=n_cand)
    blocks = pack_blocks(x, nbrs, counts, r, row_ids=row_ids, device=blocks_device or "cpu")
    return {"x": x, "nbrs": nbrs, "counts": counts, "blocks": blocks, "layout": layout(dims, r)}
