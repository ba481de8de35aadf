"""A host-side MODEL of the cooperative closest-hit query of k_render_bvh (csrc/rt_bvh_render.cuh: bvh_coop_query) -- not the
CUDA code (tests/test_gpu_parity.py::test_lbvh_straggler_mode_does_not_change_a_bit runs that), but the three claims its
correctness rests on, checked on CPU where every intermediate state can be inspected:

  1. the closest hit is the lexicographic (t, index) MINIMUM over the spheres the candidate test accepts, so the order in which
     a traversal meets them cannot change it -- 32 nodes per round from one shared work list give what the near-first
     depth-first traversal gives, ties (coincident spheres) included;
  2. a box may be culled against ANY bound that is no smaller than the final closest t (`tnear <= bound`, not `<`): the bound
     of the warp, a stale one, or the best hit of a traversal that was abandoned half-way -- which is why a query may RESTART
     from the root with the best hit so far (the straggler mode drops a lane's private stack) ...
  3. ... and why a work list that overflows may hand the query to the plain traversal, again with the best so far.

The geometry is double precision here; the claims are about order and bounds, not about rounding.
"""
import numpy as np
import pytest

KMIN = 1e-3


def sphere_hit(o, d, c, r):
    """nearest root in (KMIN, inf), else None (the shape of src/raytracer.hxx:52-92)"""
    oc = o - c
    a, b, cc = d @ d, oc @ d, oc @ oc - r * r
    disc = b * b - a * cc
    if disc <= 0:
        return None
    sq = np.sqrt(disc)
    for t in ((-b - sq) / a, (-b + sq) / a):
        if KMIN < t < np.inf:
            return t
    return None


def better(t, i, bt, bi):
    return t < bt or (t == bt and i < bi)


class Bvh:
    """median-split binary tree over sphere boxes; node = (lo, hi, left, right), leaf child = ~sphere index"""

    def __init__(self, c, r):
        self.c, self.r, self.nodes = c, r, []
        self.root = self.build(np.arange(len(c)))

    def box(self, idx):
        return (self.c[idx] - self.r[idx, None]).min(0), (self.c[idx] + self.r[idx, None]).max(0)

    def build(self, idx):
        if len(idx) == 1:
            return ~int(idx[0])
        lo, hi = self.box(idx)
        ax = int(np.argmax(hi - lo))
        order = idx[np.argsort(self.c[idx, ax], kind="stable")]
        me = len(self.nodes)
        self.nodes.append(None)
        half = len(order) // 2
        self.nodes[me] = (lo, hi, self.build(order[:half]), self.build(order[half:]))
        return me

    def child_box(self, ch):
        if ch < 0:
            return self.c[~ch] - self.r[~ch], self.c[~ch] + self.r[~ch]
        return self.nodes[ch][0], self.nodes[ch][1]


def box_tnear(o, inv, lo, hi, pad):
    """conservative slab test: the padded box's entry distance, or None when the ray misses it"""
    with np.errstate(invalid="ignore", over="ignore"):     # a ray parallel to an axis: 0 * inf, dropped by fmin / fmax
        t0, t1 = (lo - pad - o) * inv, (hi + pad - o) * inv
    tn, tf = np.fmax.reduce(np.fmin(t0, t1)), np.fmin.reduce(np.fmax(t0, t1))
    return tn if (tn <= tf and tf >= KMIN) else None


def visit(bvh, node, o, d, inv, pad, bound, bt, bi):
    """one node: leaf children tested against (bt, bi), inner children that survive `tnear <= bound` returned (near first)"""
    _, _, left, right = bvh.nodes[node]
    out = []
    for ch in (left, right):
        tn = box_tnear(o, inv, *bvh.child_box(ch), pad)
        if tn is None or not tn <= bound:
            continue
        if ch < 0:
            t = sphere_hit(o, d, bvh.c[~ch], bvh.r[~ch])
            if t is not None and better(t, ~ch, bt, bi):
                bt, bi = t, ~ch
        else:
            out.append((tn, ch))
    out.sort()
    return [ch for _, ch in out], bt, bi


def sequential(bvh, o, d, pad, bt=np.inf, bi=-1, max_visits=None):
    """near-first depth-first traversal with a private stack; max_visits: abandon it half-way and return the best so far"""
    inv = 1.0 / d
    stack, n = [bvh.root], 0
    while stack:
        if max_visits is not None and n >= max_visits:
            break
        node = stack.pop()
        n += 1
        kids, bt, bi = visit(bvh, node, o, d, inv, pad, bt, bt, bi)
        stack.extend(reversed(kids))                         # near child on top
    return bt, bi, n


def cooperative(bvh, o, d, pad, bt=np.inf, bi=-1, width=32, cap=None, rng=None):
    """`width` nodes per round from one shared work list; every lane of a round culls against the bound the round STARTED with
    (the warp merges its best once per round).  rng: shuffle which lane gets which node (no order may matter).
    Returns (t, index, fits): fits = False when the list would have held more than cap entries (the query is handed back)."""
    inv = 1.0 / d
    work = [bvh.root]
    while work:
        take = work[-width:]
        del work[-width:]
        if rng is not None:
            rng.shuffle(take)
        bound, found, pushed = bt, [], []
        for node in take:                                    # the lanes of one round: all see `bound`, none sees another's find
            kids, t, i = visit(bvh, node, o, d, inv, pad, bound, bt, bi)
            found.append((t, i))
            pushed.extend(reversed(kids))
        for t, i in found:                                   # the (t, index) minimum over the warp
            if better(t, i, bt, bi):
                bt, bi = t, i
        if cap is not None and len(work) + len(pushed) > cap:
            return bt, bi, False
        work.extend(pushed)
    return bt, bi, True


def scene(rng, n=260):
    c = rng.uniform([-12, 0, -12], [12, 0.8, 12], size=(n, 3))
    r = rng.uniform(0.1, 0.45, n)
    c[:, 1] = r                                             # spheres standing on a plane, like the 1M-sphere scene
    c[n - 20:n - 10] = c[n - 10:]                           # ten coincident pairs: exact ties in t, the lower index must win
    r[n - 20:n - 10] = r[n - 10:]
    c = np.vstack([[0, -1000, 0], c]); r = np.concatenate([[1000.0], r])      # the ground
    return c, r


def rays(rng, c, r, n):
    pick = rng.integers(0, len(c), n)
    dirs = rng.normal(size=(n, 3)); dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    o = c[pick] + dirs * np.minimum(r[pick], 2.0)[:, None] * rng.choice([0.0, 1.0, 1.0, 1.5, 6.0], (n, 1))
    d = rng.normal(size=(n, 3)) * rng.choice([1e-2, 1.0, 1.0, 30.0], (n, 1))
    d[: n // 4, 1] *= 0.02                                   # a quarter of them skim the plane: the long traversals
    return o, d


@pytest.fixture(scope="module")
def world():
    rng = np.random.default_rng(20261105)
    c, r = scene(rng)
    return Bvh(c, r), c, r, rng


def brute(c, r, o, d):
    bt, bi = np.inf, -1
    for i in range(len(c)):
        t = sphere_hit(o, d, c[i], r[i])
        if t is not None and better(t, i, bt, bi):
            bt, bi = t, i
    return bt, bi


def test_any_order_gives_the_brute_force_minimum(world):
    bvh, c, r, rng = world
    O, D = rays(rng, c, r, 160)
    hits = ties = 0
    for o, d in zip(O, D):
        want = brute(c, r, o, d)
        assert sequential(bvh, o, d, 1e-9)[:2] == want
        for width, shuffle in ((32, None), (32, rng), (5, rng), (1, None)):
            t, i, fits = cooperative(bvh, o, d, 1e-9, width=width, rng=shuffle)
            assert fits and (t, i) == want, (width, want, (t, i))
        hits += want[1] >= 0
        ties += want[1] >= len(c) - 20
    assert hits > 60 and ties > 0                            # the case means something: hits, misses and tie-broken hits


def test_padding_only_adds_work(world):
    """A conservative (padded) box test visits more nodes and returns the same hit."""
    bvh, c, r, rng = world
    O, D = rays(rng, c, r, 40)
    for o, d in zip(O, D):
        t0, i0, n0 = sequential(bvh, o, d, 0.0)
        t1, i1, n1 = sequential(bvh, o, d, 0.7)
        assert (t0, i0) == (t1, i1) and n1 >= n0
        assert cooperative(bvh, o, d, 0.7)[:2] == (t0, i0)


def test_restart_from_the_root_with_the_best_so_far(world):
    """The straggler mode drops a lane's private stack: abandon the sequential traversal after k visits, keep only its best
    hit, start again from the root cooperatively."""
    bvh, c, r, rng = world
    O, D = rays(rng, c, r, 60)
    for o, d in zip(O, D):
        want = brute(c, r, o, d)
        total = sequential(bvh, o, d, 1e-9)[2]
        for k in sorted({0, 1, total // 3, total // 2, max(total - 1, 0)}):
            bt, bi, _ = sequential(bvh, o, d, 1e-9, max_visits=k)
            assert (bt, bi) == (np.inf, -1) or bt >= want[0]                     # a real candidate: never better than the answer
            t, i, fits = cooperative(bvh, o, d, 1e-9, bt=bt, bi=bi)
            assert fits and (t, i) == want


def test_overflow_hands_the_query_back(world):
    """A work list of 4 entries (and boxes padded as for a far ray) overflows on many queries; the plain traversal, started with
    the best so far, finishes them."""
    bvh, c, r, rng = world
    O, D = rays(rng, c, r, 80)
    overflowed = 0
    for o, d in zip(O, D):
        want = brute(c, r, o, d)
        t, i, fits = cooperative(bvh, o, d, 0.7, cap=4)
        if not fits:
            overflowed += 1
            t, i, _ = sequential(bvh, o, d, 0.7, bt=t, bi=i)
        assert (t, i) == want
    assert overflowed > 10
