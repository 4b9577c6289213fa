"""Synthetic molecular graphs in the layout of ``pinot.graph.from_rdkit_mol`` (pinot/graph.py:192-244).

RDKit and the CSV datasets cannot be used on the benchmark box, so workloads are generated
(SURVEY.md section 8d): per molecule ``n_atoms ~ clip(round(N(25, 6^2)), 4, 60)``, a random spanning
tree with maximum degree 4, ``Poisson(3)`` ring-closure requests (most are met) between atoms at least 4 apart in index,
directed edges = all bonds u->v followed by all bonds v->u; node features (n, 117) = one-hot atomic
number (columns 0-99) + the 17 fingerprint columns of ``fp_rdkit`` (pinot/graph.py:74-130);
targets ``y ~ N(0, 1)``.  Everything is vectorised over molecules so that million-molecule pools are
generated in seconds, and seeded (default 2666, the reference's house seed, pinot/data/utils.py:21).
"""
import numpy as np
import torch

from .graph import BatchedGraph

MAX_ATOMS = 60
MIN_ATOMS = 4
ELEMENTS = np.array([6, 7, 8, 9, 16, 17])
ELEMENT_P = np.array([0.72, 0.10, 0.12, 0.02, 0.02, 0.02])
ELEMENT_MASS = np.array([12.011, 14.007, 15.999, 18.998, 32.06, 35.45], dtype=np.float32)
ELEMENT_VALENCE = np.array([4, 3, 2, 1, 2, 1], dtype=np.float32)
FEATURE_UNITS = 117


def _chunk(rng, m, mean_atoms, std_atoms, ring_rate):
    """Generate ``m`` molecules.  Returns (n_atoms (m,), bonds_u, bonds_v, bond_mol) with bonds in
    creation order per molecule (tree bonds by child index, then ring closures)."""
    n_atoms = np.clip(np.rint(rng.normal(mean_atoms, std_atoms, size=m)), MIN_ATOMS, MAX_ATOMS).astype(np.int64)
    amax = int(n_atoms.max())
    deg = np.zeros((m, amax), dtype=np.int8)
    idx = np.arange(amax)[None, :]
    parents = np.zeros((m, amax), dtype=np.int64)
    for i in range(1, amax):
        active = n_atoms > i
        cand = (idx < i) & (deg < 4)
        # uniformly random candidate: arg-max of iid uniforms restricted to the candidate set
        score = rng.random((m, amax), dtype=np.float32) * cand
        par = score.argmax(axis=1)
        parents[:, i] = par
        rows = np.nonzero(active)[0]
        deg[rows, par[rows]] += 1
        deg[rows, i] += 1
    # tree bonds: (parent -> child) for child = 1..n-1
    child = np.broadcast_to(idx, (m, amax))
    tmask = (child >= 1) & (child < n_atoms[:, None])
    t_mol = np.nonzero(tmask)[0]
    t_u = parents[tmask]
    t_v = child[tmask]
    # ring closures
    n_ring = rng.poisson(ring_rate, size=m)
    rmax = int(n_ring.max()) if m else 0
    r_mol, r_u, r_v, r_k = [], [], [], []
    adj = None
    if rmax:
        # adjacency bitmaps (amax <= 60 bits) to reject duplicate bonds
        adj = np.zeros((m, amax), dtype=np.uint64)
        np.bitwise_or.at(adj, (t_mol, t_u), np.uint64(1) << t_v.astype(np.uint64))
        np.bitwise_or.at(adj, (t_mol, t_v), np.uint64(1) << t_u.astype(np.uint64))
    accepted = np.zeros(m, dtype=np.int64)
    rows = np.arange(m)
    for k in range(3 * rmax):          # a few attempts per wanted ring: most random pairs are rejected
        a = (rng.random(m) * n_atoms).astype(np.int64)
        b = (rng.random(m) * n_atoms).astype(np.int64)
        lo, hi = np.minimum(a, b), np.maximum(a, b)
        ok = (accepted < n_ring) & (hi - lo >= 4) & (deg[rows, lo] < 4) & (deg[rows, hi] < 4)
        ok &= ((adj[rows, lo] >> hi.astype(np.uint64)) & np.uint64(1)) == 0
        sel = np.nonzero(ok)[0]
        accepted[sel] += 1
        deg[sel, lo[sel]] += 1
        deg[sel, hi[sel]] += 1
        adj[sel, lo[sel]] |= np.uint64(1) << hi[sel].astype(np.uint64)
        adj[sel, hi[sel]] |= np.uint64(1) << lo[sel].astype(np.uint64)
        r_mol.append(sel)
        r_u.append(lo[sel])
        r_v.append(hi[sel])
        r_k.append(np.full(sel.shape, k, dtype=np.int64))
    if r_mol:
        r_mol, r_u, r_v, r_k = (np.concatenate(x) for x in (r_mol, r_u, r_v, r_k))
    else:
        r_mol = r_u = r_v = r_k = np.zeros(0, dtype=np.int64)
    # order bonds per molecule: tree bonds (by child) first, ring bonds (by attempt) after
    b_mol = np.concatenate([t_mol, r_mol])
    b_u = np.concatenate([t_u, r_u])
    b_v = np.concatenate([t_v, r_v])
    key2 = np.concatenate([t_v, amax + r_k])  # tree bonds by child index, then ring bonds by attempt
    order = np.lexsort((key2, b_mol))
    return n_atoms, deg, b_u[order], b_v[order], b_mol[order]


def _features(rng, n_atoms, deg):
    m, amax = deg.shape
    mask = np.arange(amax)[None, :] < n_atoms[:, None]
    total = int(mask.sum())
    el = rng.choice(len(ELEMENTS), size=total, p=ELEMENT_P)
    h = np.zeros((total, FEATURE_UNITS), dtype=np.float32)
    r = np.arange(total)
    h[r, ELEMENTS[el]] = 1.0
    d = deg[mask].astype(np.float32)
    val = np.maximum(ELEMENT_VALENCE[el], d)
    h[:, 100] = val                      # total degree (heavy atoms + implicit H)
    h[:, 101] = val                      # total valence
    h[:, 102] = d                        # explicit valence
    h[:, 103] = 0.0                      # formal charge
    aromatic = (rng.random(total) < 0.3) & (d >= 2)
    h[:, 104] = aromatic
    h[:, 105] = ELEMENT_MASS[el]
    ring6 = aromatic | ((rng.random(total) < 0.1) & (d >= 2))
    ring5 = (~ring6) & (rng.random(total) < 0.08) & (d >= 2)
    h[:, 108] = ring5                    # in ring of size 5
    h[:, 109] = ring6                    # in ring of size 6
    hyb = np.where(aromatic, 1, np.where(rng.random(total) < 0.15, 1, 2))  # SP2 / SP3
    hyb = np.where((rng.random(total) < 0.02) & ~aromatic, 0, hyb)         # a few SP
    h[r, 112 + hyb] = 1.0
    return h, ELEMENTS[el].astype(np.float32)[:, None]


def make_batch(n_molecules, seed=2666, mean_atoms=25.0, std_atoms=6.0, ring_rate=3.0, device="cpu",
               chunk=32768, with_type=False):
    """Generate a batched graph of ``n_molecules`` synthetic molecules and targets ``y`` (B, 1).

    The result is identical to ``dgl.batch`` of the individual molecules (``tests`` check this
    against :func:`make_molecules` + :func:`pinot_b200.graph.batch`)."""
    rng = np.random.default_rng(seed)
    srcs, dsts, feats, types, bnn, bne = [], [], [], [], [], []
    node_off = 0
    done = 0
    while done < n_molecules:
        m = min(chunk, n_molecules - done)
        n_atoms, deg, b_u, b_v, b_mol = _chunk(rng, m, mean_atoms, std_atoms, ring_rate)
        h, z = _features(rng, n_atoms, deg)
        starts = np.concatenate([[0], np.cumsum(n_atoms)[:-1]]) + node_off
        nb = np.bincount(b_mol, minlength=m)
        # per molecule: [u->v for all bonds] then [v->u for all bonds]
        gu = b_u + starts[b_mol]
        gv = b_v + starts[b_mol]
        bond_start = np.concatenate([[0], np.cumsum(nb)[:-1]])
        pos_in_mol = np.arange(b_mol.shape[0]) - bond_start[b_mol]
        e_start = 2 * bond_start
        fwd_pos = e_start[b_mol] + pos_in_mol
        rev_pos = fwd_pos + nb[b_mol]
        src = np.empty(2 * b_mol.shape[0], dtype=np.int64)
        dst = np.empty_like(src)
        src[fwd_pos], dst[fwd_pos] = gu, gv
        src[rev_pos], dst[rev_pos] = gv, gu
        srcs.append(src)
        dsts.append(dst)
        feats.append(h)
        types.append(z)
        bnn.append(n_atoms)
        bne.append(2 * nb)
        node_off += int(n_atoms.sum())
        done += m
    src = torch.from_numpy(np.concatenate(srcs))
    dst = torch.from_numpy(np.concatenate(dsts))
    bnn = np.concatenate(bnn)
    bne = np.concatenate(bne)
    g = BatchedGraph(src, dst, node_off, bnn.tolist(), bne.tolist())
    g.ndata["h"] = torch.from_numpy(np.concatenate(feats))
    if with_type:
        g.ndata["type"] = torch.from_numpy(np.concatenate(types))
    y = torch.from_numpy(rng.standard_normal((n_molecules, 1)).astype(np.float32))
    if str(device) != "cpu":
        g = g.to(device)
        y = y.to(device)
    return g, y


def make_molecules(n_molecules, seed=2666, **kwargs):
    """The same molecules as :func:`make_batch`, as a list of single-molecule graphs + targets."""
    from .graph import unbatch
    g, y = make_batch(n_molecules, seed=seed, **kwargs)
    return unbatch(g), y
