"""Seeded synthetic stereo frames: smooth background + rectangles/discs + pixel noise.

Used by tests (small counts), bench.py (batches) -- pure numpy so it runs on the GPU box without /root/reference.
"""
import numpy as np


def _value_noise(rng, h, w, octaves=3, amp=60.0):
    out = np.zeros((h, w), np.float32)
    for o in range(octaves):
        n = 4 << o
        g = rng.random((n + 2, n + 2), dtype=np.float32)
        ys = np.linspace(0, n, h, endpoint=False, dtype=np.float32)
        xs = np.linspace(0, n, w, endpoint=False, dtype=np.float32)
        y0 = ys.astype(np.int32); x0 = xs.astype(np.int32)
        fy = (ys - y0)[:, None]; fx = (xs - x0)[None, :]
        a = g[y0][:, x0]; b = g[y0][:, x0 + 1]; c = g[y0 + 1][:, x0]; d = g[y0 + 1][:, x0 + 1]
        out += (amp / (1 << o)) * ((a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy)
    return out


def synth_left(seed, w=752, h=480, nshapes=400):
    rng = np.random.default_rng(1234 + seed)
    img = 40.0 + _value_noise(rng, h, w)
    for _ in range(nshapes):
        cx, cy = int(rng.integers(0, w)), int(rng.integers(0, h))
        s = int(rng.integers(4, 40))
        g = float(rng.integers(0, 256))
        kind, t, frac = rng.random(), rng.random() * np.pi, rng.random()
        x0, x1, y0, y1 = max(cx - s, 0), min(cx + s + 1, w), max(cy - s, 0), min(cy + s + 1, h)
        yy, xx = np.mgrid[y0:y1, x0:x1]
        if kind < 0.3:
            m = (xx - cx) ** 2 + (yy - cy) ** 2 <= (s // 2) ** 2
        elif kind < 0.65:
            u = (xx - cx) * np.cos(t) + (yy - cy) * np.sin(t)
            v = -(xx - cx) * np.sin(t) + (yy - cy) * np.cos(t)
            m = (np.abs(u) <= s // 2) & (np.abs(v) <= max(2, s // 3))
        else:
            m = (np.abs(xx - cx) <= s // 2) & (np.abs(yy - cy) <= max(2, int(s * frac)))
        img[y0:y1, x0:x1][m] = g
    return img, rng


def synth_stereo(seed, w=752, h=480, nshapes=400, dmin=2.0, dmax=60.0):
    """Rectified pair: right = left warped horizontally by a smooth disparity field, + independent noise."""
    base, rng = synth_left(seed, w, h, nshapes)
    disp = dmin + (dmax - dmin) * (_value_noise(rng, h, w, octaves=2, amp=1.0) / 1.5)
    xs = np.arange(w, dtype=np.float32)[None, :] + disp       # right(x) = left(x + d)
    x0 = np.clip(np.floor(xs).astype(np.int32), 0, w - 1)
    x1 = np.clip(x0 + 1, 0, w - 1)
    f = xs - np.floor(xs)
    rows = np.arange(h)[:, None]
    right = base[rows, x0] * (1 - f) + base[rows, x1] * f
    left = base + rng.normal(0, 2.0, base.shape)
    right = right + rng.normal(0, 2.0, base.shape)
    return (np.clip(np.rint(left), 0, 255).astype(np.uint8), np.clip(np.rint(right), 0, 255).astype(np.uint8))


def rectify_maps(w, h, k1=-0.28, k2=0.07, tilt=0.012, shift=(3.25, -2.5), scale=1.04):
    """Deterministic float32 rectification maps (M1 = x, M2 = y) of a radial-distortion + small-rotation camera, using
    only correctly-rounded float32 operations (+ - * /), so every machine computes identical bits.  Parts of the
    destination map outside the source on purpose (constant-0 border of cv::remap)."""
    f = np.float32
    yy, xx = np.meshgrid(np.arange(h, dtype=np.float32), np.arange(w, dtype=np.float32), indexing="ij")
    cx, cy, fl = f(w / 2 - 0.5), f(h / 2 - 0.5), f(0.6 * w)
    x = (xx - cx) / fl * f(scale); y = (yy - cy) / fl * f(scale)
    xr = x - f(tilt) * y; yr = y + f(tilt) * x
    r2 = xr * xr + yr * yr
    d = f(1) + f(k1) * r2 + f(k2) * r2 * r2
    return (xr * d * fl + cx + f(shift[0])).astype(np.float32), (yr * d * fl + cy + f(shift[1])).astype(np.float32)


KEYPOINT = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
PROJ_QUERY = np.dtype([("x", "<f4"), ("y", "<f4"), ("r", "<f4"), ("xr", "<f4"), ("angle", "<f4"),
                       ("min_level", "<i4"), ("max_level", "<i4"), ("flags", "<i4")])


def projection_case(seed, n, nq, w=752, h=480, form="map", th=1.0, flip=0.07, p_dup=0.2, p_free=0.15, p_occupied=0.1,
                    stereo=True):
    """One SearchByProjection call's inputs: a frame (keys, descriptors, mvuRight, occupied keys) and nq already
    projected queries that mostly re-observe the frame's keys (moved a few pixels, noisy descriptor, predicted level
    off by one now and then), with deliberate competition: p_dup of the queries copy another query's target, p_free of
    them do not block (Observations() == 0).  form: "map" (:155-241), "last" (:1440-1593), "keyframe" (:1595-1721) or "sim3" (:406-518)
    chooses how radius, level range and flags are derived, as the reference's three callers derive them."""
    rng = np.random.default_rng(9000 + seed)
    F = np.float32
    scale = (F(1.2) ** np.arange(8)).astype(F)
    keys = np.zeros(n, KEYPOINT)
    keys["x"] = (rng.random(n) * (w + 40) - 20).astype(F); keys["y"] = (rng.random(n) * (h + 40) - 20).astype(F)
    keys["angle"] = (rng.random(n) * 360).astype(F)
    keys["octave"] = np.minimum(rng.geometric(0.35, n) - 1, 7)
    keys["size"], keys["class_id"] = 31, -1
    desc = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    u_right = np.where(rng.random(n) < (0.7 if stereo else 0.0), keys["x"] - (rng.random(n) * 50 + 1).astype(F), F(-1)).astype(F)
    occupied = (rng.random(n) < p_occupied).astype(np.uint8)
    q = np.zeros(nq, PROJ_QUERY)
    qdesc = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    if n and nq:
        tgt = rng.integers(0, n, nq)
        dup = np.nonzero(rng.random(nq) < p_dup)[0]
        tgt[dup] = tgt[rng.integers(0, nq, len(dup))]
        sig = 1.5 if form == "map" else 4.0
        q["x"] = keys["x"][tgt] + rng.normal(0, sig, nq).astype(F); q["y"] = keys["y"][tgt] + rng.normal(0, sig, nq).astype(F)
        lvl = np.clip(keys["octave"][tgt] + rng.choice([0, 0, 0, 1, -1], nq), 0, 7)
        q["angle"] = np.mod(keys["angle"][tgt] + rng.normal(12, 4, nq), 360).astype(F)
        wild = rng.random(nq) < 0.12
        q["angle"][wild] = (rng.random(int(wild.sum())) * 360).astype(F)
        qdesc[:] = desc[tgt] ^ np.packbits(rng.random((nq, 256)) < flip, axis=1)
        far = rng.random(nq) < 0.1                                     # descriptors that should fail the threshold
        qdesc[far] = rng.integers(0, 256, (int(far.sum()), 32), dtype=np.uint8)
        ur = u_right[tgt]
        q["xr"] = np.where(rng.random(nq) < 0.85, ur + rng.normal(0, 1.0, nq), ur + 80).astype(F)
        if form == "map":
            r = np.where(rng.random(nq) < 0.5, F(2.5), F(4.0)).astype(F)          # RadiusByViewingCos, :243-249
            if th != 1.0:
                r = r * F(th)
            q["r"] = r * scale[lvl]
            q["min_level"], q["max_level"] = lvl - 1, lvl
            q["flags"] = 1 | 4 | np.where(rng.random(nq) < p_free, 0, 2)
        elif form == "last":
            q["r"] = F(th) * scale[lvl]
            mode = rng.integers(0, 3)                                             # bForward / bBackward / neither
            if mode == 0:
                q["min_level"], q["max_level"] = lvl, -1
            elif mode == 1:
                q["min_level"], q["max_level"] = 0, lvl
            else:
                q["min_level"], q["max_level"] = lvl - 1, lvl + 1
            q["flags"] = 1 | 4 | np.where(rng.random(nq) < p_free, 0, 2)
        elif form == "fuse":                                                      # :937-1087: error gate, nothing blocks
            q["r"] = F(th) * scale[lvl]
            q["min_level"], q["max_level"] = lvl - 1, lvl
            q["flags"] = 1 | 8
        elif form == "sim3":                                                      # :406-518: levels lvl-1 .. lvl
            q["r"] = F(th) * scale[lvl]
            q["min_level"], q["max_level"] = lvl - 1, lvl
            q["flags"] = 1 | 2
        else:
            q["r"] = F(th) * scale[lvl]
            q["min_level"], q["max_level"] = lvl - 1, lvl + 1
            q["flags"] = 1 | 2
        q["flags"] &= np.where(rng.random(nq) < 0.06, ~1, ~0)                     # not in view / bad / outlier
    return keys, desc, u_right, occupied, q, qdesc


def bow_case(seed, n1, n2, nodes=400, flip=0.06, p_valid1=0.8, p_valid2=0.85, dup=0.25, with_partner=False):
    """Inputs of one ORBmatcher::SearchByBoW call: two frames whose first min(n1, n2) features are re-observations
    (noisy descriptor, rotated a little), every feature quantised into one of `nodes` vocabulary nodes (a re-observed
    feature lands in its partner's node 9 times out of 10), near-duplicate descriptors inside nodes so that several
    queries compete for one candidate, and node ids that exist in only one frame so the merge has to skip.
    Returns (k1, d1, valid1, fv1, k2, d2, valid2, fv2); fvX = {node id: [feature indices in insertion order]}."""
    rng = np.random.default_rng(7000 + seed)
    F = np.float32

    def keys(n):
        k = np.zeros(n, KEYPOINT)
        k["x"] = (rng.random(n) * 752).astype(F); k["y"] = (rng.random(n) * 480).astype(F)
        k["angle"] = (rng.random(n) * 360).astype(F); k["octave"] = rng.integers(0, 8, n); k["size"], k["class_id"] = 31, -1
        return k
    k1, k2 = keys(n1), keys(n2)
    d1 = rng.integers(0, 256, (n1, 32), dtype=np.uint8); d2 = rng.integers(0, 256, (n2, 32), dtype=np.uint8)
    node1 = rng.integers(0, max(nodes, 1), n1) * 7 + 3
    node2 = rng.integers(0, max(nodes, 1), n2) * 7 + 3 + (rng.random(n2) < 0.2)          # some ids frame 1 never has
    m = min(n1, n2)
    if m:
        nd = int(m * dup)
        src = rng.integers(0, m, nd); dst = rng.integers(0, m, nd)
        d1[dst] = d1[src] ^ np.packbits(rng.random((nd, 256)) < 0.02, axis=1); node1[dst] = node1[src]
        d2[:m] = d1[:m] ^ np.packbits(rng.random((m, 256)) < flip, axis=1)
        same = rng.random(m) < 0.9
        node2[:m] = np.where(same, node1[:m], node2[:m])
        k2["angle"][:m] = np.mod(k1["angle"][:m] + rng.normal(10, 4, m), 360).astype(F)
        wild = rng.random(m) < 0.12
        k2["angle"][:m][wild] = (rng.random(int(wild.sum())) * 360).astype(F)
    p2 = rng.permutation(n2)
    k2, d2, node2 = k2[p2], d2[p2], node2[p2]
    partner = np.where(p2 < m, p2, -1)                # frame-1 feature a frame-2 feature re-observes

    def featvec(node, n):
        fv = {}
        for i in rng.permutation(n):                  # DBoW2 pushes features in descriptor order; any order must work
            fv.setdefault(int(node[i]), []).append(int(i))
        return fv
    valid1 = (rng.random(n1) < p_valid1).astype(np.uint8); valid2 = (rng.random(n2) < p_valid2).astype(np.uint8)
    out = (k1, d1, valid1, featvec(node1, n1), k2, d2, valid2, featvec(node2, n2))
    return out + (partner,) if with_partner else out


def triangulation_case(seed, n1, n2, nodes=400):
    """Inputs of one ORBmatcher::SearchForTriangulation call: bow_case frames where a re-observed feature sits on its
    partner's epipolar line (sideways motion: same row, one pixel of noise, a slightly perturbed F12 so every term of
    the line equation is exercised), stereo flags on both sides, an epipole inside the image.
    Returns (k1, d1, valid1, stereo1, fv1, k2, d2, valid2, stereo2, fv2, F12, ex, ey, scale_factors, level_sigma2)."""
    k1, d1, v1, fv1, k2, d2, v2, fv2, partner = bow_case(seed, n1, n2, nodes, with_partner=True)
    rng = np.random.default_rng(8000 + seed)
    F = np.float32
    has = partner >= 0
    k2["x"][has] = (k1["x"][partner[has]] - rng.random(int(has.sum())) * 40).astype(F)
    k2["y"][has] = (k1["y"][partner[has]] + rng.normal(0, 1.0, int(has.sum()))).astype(F)
    k2["octave"][has] = k1["octave"][partner[has]]
    F12 = np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], F) + rng.normal(0, 2e-5, (3, 3)).astype(F)
    stereo1 = (rng.random(n1) < 0.5).astype(np.uint8); stereo2 = (rng.random(n2) < 0.5).astype(np.uint8)
    sf = (F(1.2) ** np.arange(8)).astype(F)
    return k1, d1, v1, stereo1, fv1, k2, d2, v2, stereo2, fv2, F12, F(400.0), F(240.0), sf, (sf * sf).astype(F)


def sim3_case(seed, n1, n2, w=752, h=480, th=7.5, flip=0.07, p_mp=0.8, p_already=0.1):
    """Inputs of one ORBmatcher::SearchBySim3 call, already projected: two key frames (keys, descriptors) whose first
    min(n1, n2) * 0.7 features observe the same scene points; every feature may hold a map point (p_mp) whose descriptor
    is a noisy copy of the feature's; queries12[i1] = map point i1 of KF1 projected into KF2 (lands a few pixels from its
    partner key when it has one, anywhere otherwise), queries21 the other way round; a share of the map points is
    "already matched" / bad / out of view (flags = 0).  Near-duplicate descriptors and shared windows make the two
    directions disagree now and then.  Returns (k1, d1, k2, d2, q12, qd12, q21, qd21)."""
    rng = np.random.default_rng(12000 + seed)
    F = np.float32
    scale = (F(1.2) ** np.arange(8)).astype(F)

    def frame(n):
        k = np.zeros(n, KEYPOINT)
        k["x"] = (rng.random(n) * (w + 30) - 15).astype(F); k["y"] = (rng.random(n) * (h + 30) - 15).astype(F)
        k["angle"] = (rng.random(n) * 360).astype(F)
        k["octave"] = np.minimum(rng.geometric(0.35, n) - 1, 7)
        k["size"], k["class_id"] = 31, -1
        return k, rng.integers(0, 256, (n, 32), dtype=np.uint8)
    k1, d1 = frame(n1)
    k2, d2 = frame(n2)
    m = int(0.7 * min(n1, n2))
    partner12 = np.full(n1, -1); partner21 = np.full(n2, -1)
    if m:
        p2 = rng.permutation(n2)[:m]
        partner12[:m] = p2; partner21[p2] = np.arange(m)
        d2[p2] = d1[:m] ^ np.packbits(rng.random((m, 256)) < flip, axis=1)
        k2["octave"][p2] = np.clip(k1["octave"][:m] + rng.choice([0, 0, 0, 1, -1], m), 0, 7)
        nd = m // 6                                                    # near-duplicate descriptors inside KF2: rivals
        if nd:
            src = rng.integers(0, m, nd); dst = rng.permutation(n2)[:nd]
            d2[dst] = d2[p2[src]] ^ np.packbits(rng.random((nd, 256)) < 0.02, axis=1)
            k2["x"][dst] = k2["x"][p2[src]] + rng.normal(0, 3, nd).astype(F); k2["y"][dst] = k2["y"][p2[src]] + rng.normal(0, 3, nd).astype(F)
            k2["octave"][dst] = k2["octave"][p2[src]]

    def queries(n, k_other, partner, d_own):
        q = np.zeros(n, PROJ_QUERY)
        qd = d_own ^ np.packbits(rng.random((n, 256)) < 0.03, axis=1)    # pMP->GetDescriptor(): close to the observing feature's
        has = partner >= 0
        q["x"] = (rng.random(n) * w).astype(F); q["y"] = (rng.random(n) * h).astype(F)
        lvl = rng.integers(0, 8, n)
        if has.any():
            q["x"][has] = k_other["x"][partner[has]] + rng.normal(0, 2.5, int(has.sum())).astype(F)
            q["y"][has] = k_other["y"][partner[has]] + rng.normal(0, 2.5, int(has.sum())).astype(F)
            lvl[has] = np.clip(k_other["octave"][partner[has]] + rng.choice([0, 0, 1, 1, 2, -1], int(has.sum())), 0, 7)
        q["r"] = F(th) * scale[lvl]
        q["min_level"], q["max_level"] = lvl - 1, lvl
        q["flags"] = (rng.random(n) < p_mp) & ~(rng.random(n) < p_already)
        return q, qd
    q12, qd12 = queries(n1, k2, partner12, d1)
    q21, qd21 = queries(n2, k1, partner21, d2)
    return k1, d1, k2, d2, q12, qd12, q21, qd21
