"""Design study (CPU, numpy): per-ray and per-warp traversal work of candidate acceleration structures over the
traced primitives of random_scene, on a ray population drawn with the oracle (primary + diffuse bounces).
    python tools/trav_sim.py [W H SPP]
Counts inner visits / box tests / exact tests per ray for BVH2 (SAH), its BVH4 / BVH8 collapses and a uniform grid,
and replays 32-lane warps under the kernel's inner/leaf majority vote to get warp-level iteration counts."""
import os, sys, math, random
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from oracle import binding as O
from raytracingoneweekend_b200 import _abi as A

W, H, SPP = (int(x) for x in (sys.argv[1:4] + [96, 54, 4][len(sys.argv) - 1:]))
rng = np.random.default_rng(7)
osc = O.OracleScene().random_scene(1)
sph = osc.get("spheres")
n_s = len(sph)
M = np.array([np.array(s.m_local_to_world[:], np.float64).reshape(4, 4) for s in sph])
Mi = np.array([np.array(s.m_world_to_local[:], np.float64).reshape(4, 4) for s in sph])
ctr = M[:, :3, 3]
ext = np.sqrt((M[:, :3, :3] ** 2).sum(axis=2))
lo = ctr - ext
hi = ctr + ext
big = ext.max(axis=1) > 50
prims = [i for i in range(n_s) if not big[i]]
print("spheres", n_s, "always-list", int(big.sum()), "bvh prims", len(prims))
P_lo = lo[prims] - 1e-4
P_hi = hi[prims] + 1e-4
P_Mi = Mi[prims]
NP = len(prims)

# ---- ray population ----------------------------------------------------------------------------
cam = O.camera_new((13, 2, 3), (0, 0, 0), (0, 1, 0), 20.0, W / H, 0.1, 10.0)
par = A.RenderParams(W, H, SPP, 50, 0.001, 100.0, 1, 0, 0, 0, 0, 1, 0)
gens = []  # list of (origins, dirs, pixel)
for s in range(SPP):
    o, d = O.primary_rays(cam, par, 1, s)
    px = np.arange(W * H)
    for g in range(4):
        gens.append((o.copy(), d.copy(), px.copy(), g))
        ids, t, pts, nrm = osc.trace_rays(o, d)
        hit = ids != 0
        if not hit.any():
            break
        u = rng.normal(size=(hit.sum(), 3))
        u /= np.linalg.norm(u, axis=1, keepdims=True)
        nd = nrm[hit] + u
        nd /= np.linalg.norm(nd, axis=1, keepdims=True)
        o, d, px = pts[hit].astype(np.float32), nd.astype(np.float32), px[hit]
RO = np.concatenate([g[0] for g in gens]).astype(np.float64)
RD = np.concatenate([g[1] for g in gens]).astype(np.float64)
RP = np.concatenate([g[2] for g in gens])
RG = np.concatenate([np.full(len(g[2]), g[3]) for g in gens])
NR = len(RO)
print("rays", NR, "by generation", [int((RG == g).sum()) for g in range(4)])

# initial t bound = hit on the always-list (ground sphere) else tmax
def ground_t(o, d):
    best = 100.0
    for i in np.nonzero(big)[0]:
        ol = Mi[i, :3, :3] @ o + Mi[i, :3, 3]
        dl = Mi[i, :3, :3] @ d
        a = dl @ dl; hb = ol @ dl; c = ol @ ol - 1
        disc = hb * hb - a * c
        if disc < 0: continue
        sq = math.sqrt(disc)
        for r in ((-hb - sq) / a, (-hb + sq) / a):
            if 0.001 <= r <= best:
                best = r
                break
    return best

def exact_t(k, o, d, tmax):
    ol = P_Mi[k, :3, :3] @ o + P_Mi[k, :3, 3]
    dl = P_Mi[k, :3, :3] @ d
    a = dl @ dl; hb = ol @ dl; c = ol @ ol - 1
    disc = hb * hb - a * c
    if disc < 0: return None
    sq = math.sqrt(disc)
    for r in ((-hb - sq) / a, (-hb + sq) / a):
        if 0.001 <= r <= tmax:
            return r
    return None

def slab(blo, bhi, o, inv, tmax):
    t0 = (blo - o) * inv; t1 = (bhi - o) * inv
    tn = max(np.minimum(t0, t1).max(), 0.001); tf = min(np.maximum(t0, t1).min(), tmax)
    return tn if tn <= tf else None

# ---- BVH2 (binned SAH, one primitive per leaf) -----------------------------------------------------
class Node:
    __slots__ = ("lo", "hi", "kids", "prim", "axis")
def area(l, h):
    d = h - l
    return d[0] * d[1] + d[1] * d[2] + d[2] * d[0]
def build(idx):
    n = Node(); n.lo = P_lo[idx].min(axis=0); n.hi = P_hi[idx].max(axis=0); n.prim = -1; n.kids = []
    if len(idx) == 1:
        n.prim = idx[0]; return n
    c = 0.5 * (P_lo[idx] + P_hi[idx])
    best = (1e30, None)
    for a in range(3):
        order = np.argsort(c[:, a], kind="stable")
        srt = [idx[j] for j in order]
        L_lo = np.minimum.accumulate(P_lo[srt], axis=0); L_hi = np.maximum.accumulate(P_hi[srt], axis=0)
        R_lo = np.minimum.accumulate(P_lo[srt][::-1], axis=0)[::-1]; R_hi = np.maximum.accumulate(P_hi[srt][::-1], axis=0)[::-1]
        for k in range(1, len(srt)):
            cost = area(L_lo[k - 1], L_hi[k - 1]) * k + area(R_lo[k], R_hi[k]) * (len(srt) - k)
            if cost < best[0]: best = (cost, (srt, k))
    srt, k = best[1]
    n.kids = [build(srt[:k]), build(srt[k:])]
    return n
root2 = build(list(range(NP)))

def collapse(n, width):
    """wide node: children = up to `width` subtrees, opening the largest-area inner child first"""
    if n.prim >= 0: return n
    kids = list(n.kids)
    while len(kids) < width:
        cand = [(area(k.lo, k.hi), j) for j, k in enumerate(kids) if k.prim < 0]
        if not cand: break
        _, j = max(cand)
        k = kids.pop(j)
        kids.extend(k.kids)
    w = Node(); w.lo = n.lo; w.hi = n.hi; w.prim = -1
    w.kids = [collapse(k, width) for k in kids]
    return w
def count_nodes(n):
    return 0 if n.prim >= 0 else 1 + sum(count_nodes(k) for k in n.kids)
def depth(n):
    return 0 if n.prim >= 0 else 1 + max(depth(k) for k in n.kids)
roots = {2: root2, 4: collapse(root2, 4), 8: collapse(root2, 8)}
for w, r in roots.items():
    print("BVH%d nodes %d depth %d" % (w, count_nodes(r), depth(r)))

def trav_bvh(root, o, d, t0):
    """returns op sequence 'I' (inner visit) / 'L' (exact test), boxes tested, final t"""
    inv = 1.0 / np.where(np.abs(d) < 1e-30, 1e-30, d)
    ops = []; boxes = 0; best = t0
    stack = [(0.0, root)]
    while stack:
        te, n = stack.pop()
        if te > best: continue
        if n.prim >= 0:
            ops.append("L")
            r = exact_t(n.prim, o, d, best)
            if r is not None: best = r
            continue
        ops.append("I")
        hits = []
        for k in n.kids:
            boxes += 1
            t = slab(k.lo, k.hi, o, inv, best)
            if t is not None: hits.append((t, id(k), k))
        hits.sort(reverse=True)
        for t, _, k in hits: stack.append((t, k))
    return ops, boxes, best

# ---- uniform grid ---------------------------------------------------------------------------------
def make_grid(res):
    glo = P_lo.min(axis=0) - 1e-3; ghi = P_hi.max(axis=0) + 1e-3
    dims = np.maximum(1, np.array(res))
    h = (ghi - glo) / dims
    cells = {}
    for k in range(NP):
        a = np.clip(np.floor((P_lo[k] - 1e-3 * h - glo) / h).astype(int), 0, dims - 1)
        b = np.clip(np.floor((P_hi[k] + 1e-3 * h - glo) / h).astype(int), 0, dims - 1)
        for x in range(a[0], b[0] + 1):
            for y in range(a[1], b[1] + 1):
                for z in range(a[2], b[2] + 1):
                    cells.setdefault((x, y, z), []).append(k)
    ent = sum(len(v) for v in cells.values())
    print("grid", tuple(dims), "entries", ent, "non-empty cells", len(cells), "of", int(dims.prod()))
    return glo, ghi, dims, h, cells
def trav_grid(G, o, d, t0):
    glo, ghi, dims, h, cells = G
    inv = 1.0 / np.where(np.abs(d) < 1e-30, 1e-30, d)
    ops = []; best = t0
    te = slab(glo, ghi, o, inv, best)
    if te is None: return ["S"], 0, best
    ops.append("S")
    p = o + te * d
    c = np.clip(np.floor((p - glo) / h).astype(int), 0, dims - 1)
    step = np.where(d >= 0, 1, -1)
    nb = glo + (c + (step > 0)) * h
    tn = (nb - o) * inv
    dt = h * np.abs(inv)
    tested = set(); cand = 0
    while True:
        for k in cells.get((c[0], c[1], c[2]), ()):
            ops.append("C"); cand += 1   # candidate box test (no mailbox: duplicates are re-tested)
            if slab(P_lo[k], P_hi[k], o, inv, best) is not None:
                ops.append("L")
                r = exact_t(k, o, d, best)
                if r is not None: best = r
        a = int(np.argmin(tn))
        if tn[a] > best: break
        c[a] += step[a]
        if c[a] < 0 or c[a] >= dims[a]: break
        tn[a] += dt[a]
        ops.append("S")
    return ops, cand, best

# ---- run ----------------------------------------------------------------------------------------------
sel = rng.permutation(NR)[: int(os.environ.get("NSIM", "6000"))]
T0 = {i: ground_t(RO[i], RD[i]) for i in sel}
structs = {"bvh2": lambda o, d, t: trav_bvh(roots[2], o, d, t), "bvh4": lambda o, d, t: trav_bvh(roots[4], o, d, t),
           "bvh8": lambda o, d, t: trav_bvh(roots[8], o, d, t)}
for res in ((23, 1, 23), (23, 2, 23), (46, 2, 46), (32, 3, 32)):
    G = make_grid(res)
    structs["grid%dx%dx%d" % res] = (lambda G: lambda o, d, t: trav_grid(G, o, d, t))(G)
COST = {"I2": 65, "I4": 105, "I8": 190, "L": 135, "S": 22, "C": 22, "vote": 17}
seqs = {}
for name, f in structs.items():
    tot = {"I": 0, "L": 0, "S": 0, "C": 0}; boxes = 0; ref = None
    seqs[name] = {}
    for i in sel:
        ops, b, best = f(RO[i], RD[i], T0[i])
        seqs[name][i] = ops
        for c in ops: tot[c] += 1
        boxes += b
    n = len(sel)
    ik = "I" + name[3] if name.startswith("bvh") else None
    cost = sum(tot[c] * (COST[ik] if c == "I" else COST[c]) for c in tot if tot[c])
    print("%-14s per ray: inner %.2f leaf %.2f step %.2f cand %.2f boxes %.1f  -> thread-instr %.0f" % (name, tot["I"] / n, tot["L"] / n, tot["S"] / n, tot["C"] / n, boxes / n, cost / n))

# warp replay: 32 random rays of the population per call, majority vote between cheap ops (I / S / C) and L
def replay(name, calls=300, lanes=23):
    ik = "I" + name[3] if name.startswith("bvh") else None
    tot = 0.0; it_cheap = 0; it_leaf = 0
    keys = list(seqs[name].keys())
    for _ in range(calls):
        rays = [list(seqs[name][random.choice(keys)]) for _ in range(lanes)]
        pos = [0] * lanes
        while True:
            cheap = sum(1 for r, p in zip(rays, pos) if p < len(r) and r[p] != "L")
            leaf = sum(1 for r, p in zip(rays, pos) if p < len(r) and r[p] == "L")
            if cheap + leaf == 0: break
            if cheap >= leaf:
                it_cheap += 1
                kinds = set(r[p] for r, p in zip(rays, pos) if p < len(r) and r[p] != "L")
                tot += COST["vote"] + sum(COST[ik] if k == "I" else COST[k] for k in kinds)  # S and C bodies diverge
                pos = [p + 1 if p < len(r) and r[p] != "L" else p for r, p in zip(rays, pos)]
            else:
                it_leaf += 1
                tot += COST["vote"] + COST["L"]
                pos = [p + 1 if p < len(r) and r[p] == "L" else p for r, p in zip(rays, pos)]
    return tot / calls, it_cheap / calls, it_leaf / calls
random.seed(3)
for name in structs:
    c, a, b = replay(name)
    print("%-14s warp call (23 lanes): %.0f warp-instr, %.1f cheap iterations, %.1f leaf iterations" % (name, c, a, b))
