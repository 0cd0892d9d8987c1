"""Scenarios for the LocalMapping / LoopClosing matchers, shared by the CPU tests (oracle against the
reference's compiled ORBmatcher.cc) and the GPU tests (liborbgpu against both, and the C++ shim against the reference).

The poses are chosen so that the reference's fp32 projection algebra is exact and can be restated here without
depending on how cv::Mat orders its sums: identity rotations, unit intrinsics, every point at depth 1, so a point
(X, Y, 1) projects to (u, v) = (X, Y) after at most one rounding per coordinate."""
import numpy as np

from ar_orbslam2_b200.synth import SplitMix64

W, H = 752, 480
TH_LOW, TH_HIGH = 50, 100


def bow_groups(ka, kb, dx, dy):
    """Stand-in for two DBoW2 FeatureVectors: vocabulary node = 40 px cell (of the un-shifted position), as CSR over the
    nodes both key frames hold, ascending node id."""
    def node_of(k, sx, sy):
        node = ((k["y"] - sy) // 40).astype(np.int64) * 19 + ((k["x"] - sx) // 40).astype(np.int64)
        return node, np.argsort(node, kind="stable")
    na, oa = node_of(ka, 0, 0)
    nb, ob = node_of(kb, dx, dy)
    off1, off2, idx1, idx2 = [0], [0], [], []
    for c in np.intersect1d(na, nb):
        idx1 += oa[na[oa] == c].tolist(); idx2 += ob[nb[ob] == c].tolist()
        off1.append(len(idx1)); off2.append(len(idx2))
    return off1, idx1, off2, idx2


def triangulation_case(scene, seed=3):
    """KF1 = frame A, KF2 = frame A shifted by (7, 4): the epipolar geometry of a pure image-plane translation,
    F12 = [t]x with t = (7, 4, 0), so x1' F12 x2 = 0 for x2 = x1 + t."""
    ka, kb = scene["ka"], scene["kb"]
    off1, idx1, off2, idx2 = bow_groups(ka, kb, 7, 4)
    rng = SplitMix64(seed)
    usable1 = np.array([rng.next() % 8 != 0 for _ in range(len(ka))], np.uint8)
    usable2 = np.array([rng.next() % 8 != 0 for _ in range(len(kb))], np.uint8)
    ur1 = np.where(np.arange(len(ka)) % 3 == 0, ka["x"] - 5.0, -1.0).astype(np.float32)
    ur2 = np.where(np.arange(len(kb)) % 4 == 0, kb["x"] - 5.0, -1.0).astype(np.float32)
    t = np.array([7.0, 4.0, 0.0])
    F12 = np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]], np.float32) / np.float32(8.0)
    sigma2 = (scene["scale"] * scene["scale"]).astype(np.float32)
    # the epipole sits next to a monocular feature of KF2 whose counterpart in KF1 is monocular too, so that the "too close
    # to the epipole" test (:747-753) rejects a pair that would otherwise match
    j = -1
    for cand in np.flatnonzero((ur2 < 0) & (usable2 != 0)):
        near = np.flatnonzero((np.abs(ka["x"] + 7 - kb["x"][cand]) < 1.5) & (np.abs(ka["y"] + 4 - kb["y"][cand]) < 1.5) & (ur1 < 0) &
                              (usable1 != 0) & (ka["octave"] == kb["octave"][cand]))
        if len(near) and min(int(np.unpackbits(scene["da"][i] ^ scene["db"][cand]).sum()) for i in near) < 25:
            j = int(cand)
            break
    assert j >= 0
    return dict(off1=off1, idx1=idx1, off2=off2, idx2=idx2, usable1=usable1, usable2=usable2, ur1=ur1, ur2=ur2, F12=F12,
                ex=np.float32(kb["x"][j] + 1.25), ey=np.float32(kb["y"][j] - 0.5), sigma2=sigma2)


def _levels_and_ranges(octave, dist, seed):
    """mfMaxDistance such that PredictScale (ceil(log(max/dist)/log 1.2), MapPoint.cc:357-366) lands on the feature's octave or
    the one above (so both ends of the [level-1, level] filter are exercised), robustly against the last bit of dist."""
    rng = SplitMix64(seed)
    level = np.clip(octave + np.array([rng.next() % 2 for _ in range(len(octave))]), 0, 7).astype(np.int32)
    maxd = (dist * np.power(1.2, level - 0.5)).astype(np.float32)
    return level, maxd, np.zeros(len(octave), np.float32)


def fuse_case(scene, seed=11):
    """Key frame = frame B at the identity pose with unit intrinsics; candidate map points = frame A's features moved by
    the (7, 4) shift plus a little noise, at depth 1, carrying frame A's descriptors."""
    ka, kb = scene["ka"], scene["kb"]
    n, N = len(ka), len(kb)
    rng = SplitMix64(seed)
    jit = np.array([[(rng.next() % 9 - 4) / 4.0, (rng.next() % 9 - 4) / 4.0] for _ in range(n)], np.float32)
    X = (ka["x"] + np.float32(7.0) + jit[:, 0]).astype(np.float32)
    Y = (ka["y"] + np.float32(4.0) + jit[:, 1]).astype(np.float32)
    world = np.stack([X, Y, np.ones(n, np.float32)], 1).astype(np.float32)
    dist = np.sqrt(X.astype(np.float64) ** 2 + Y.astype(np.float64) ** 2 + 1.0)
    level, maxd, mind = _levels_and_ranges(ka["octave"], dist, seed + 1)
    normal = (world / np.linalg.norm(world, axis=1, keepdims=True)).astype(np.float32)   # PO.Pn = dist > 0.5 dist
    normal[::7] *= -1                                                                      # these fail the 60-degree test
    present = np.array([(0 if r % 23 == 0 else 2 if r % 19 == 0 else 1) for r in (rng.next() for _ in range(n))], np.uint8)
    nobs = np.array([rng.next() % 6 for _ in range(n)], np.int32)
    # what the key frame already holds: other map points at some features, and a few of the candidates themselves
    kf_mp_obs = np.array([(int(r % 6) if r % 3 == 0 else -1) for r in (rng.next() for _ in range(N))], np.int32)
    in_kf = np.full(n, -1, np.int32)
    free = np.flatnonzero(kf_mp_obs < 0)
    for j, i in enumerate(range(5, n, 29)):
        if present[i] and j < len(free):
            in_kf[i] = free[-1 - j]
    ur_kf = np.where(np.arange(N) % 3 == 0, kb["x"] - np.float32(5.0), -1.0).astype(np.float32)
    mbf = np.float32(5.0)
    return dict(world=world, normal=normal, present=present, maxd=maxd, mind=mind, nobs=nobs, in_kf=in_kf, kf_mp_obs=kf_mp_obs,
                level=level, u=X, v=Y, ur=(X - mbf).astype(np.float32), ur_kf=ur_kf, mbf=mbf, desc=scene["da"],
                geom_ok=(np.arange(n) % 7 != 0) & (X >= 0) & (X < W) & (Y >= 0) & (Y < H))


def fuse_flat(matcher, kf, scene, c, variant, th):
    """ORBmatcher::Fuse restated around the flat C ABI: queries -> SearchWindowBest -> the reference's bookkeeping
    (ORBmatcher.cc:954-973 / :1075-1088) on plain arrays.  Returns the same dict as oracle.refm_fuse."""
    n, N = len(c["u"]), kf.N
    inv_sigma2 = (np.float32(1.0) / (scene["scale"] * scene["scale"]).astype(np.float32)).astype(np.float32)
    radius = (np.float32(th) * scene["scale"][c["level"]]).astype(np.float32)
    valid = (c["present"] != 0) & c["geom_ok"]
    if variant == 1:
        valid &= (c["present"] == 1) & (c["in_kf"] < 0)      # isBad / spAlreadyFound are fixed for the whole call (:1009)
    gate = dict(gate_ur=c["ur"], inv_level_sigma2=inv_sigma2) if variant == 0 else {}
    best_idx, best_dist = matcher.SearchWindowBest(kf, c["u"], c["v"], radius, c["level"] - 1, c["level"], c["desc"],
                                                   valid=valid.astype(np.uint8), **gate)
    kf_state = np.where(c["kf_mp_obs"] >= 0, 1000000 + np.arange(N), -1).astype(np.int32)
    kf_obs = {1000000 + i: int(c["kf_mp_obs"][i]) for i in range(N) if c["kf_mp_obs"][i] >= 0}
    pt_obs, pt_repl, kfmp_repl = c["in_kf"].copy(), np.full(n, -1, np.int32), np.full(N, -1, np.int32)
    for i in np.flatnonzero(c["in_kf"] >= 0):
        kf_state[c["in_kf"][i]] = i
    bad = {i: c["present"][i] == 2 for i in range(n)}
    bad.update({k: False for k in kf_obs})
    nobs = {i: int(c["nobs"][i]) for i in range(n)}
    nobs.update(kf_obs)
    nfused = 0
    for i in range(n):
        if not c["present"][i]:
            continue
        if variant == 0 and (bad[i] or pt_obs[i] >= 0):       # isBad() || IsInKeyFrame(pKF), evaluated in loop order (:851)
            continue
        if not valid[i] or best_dist[i] > TH_LOW:
            continue
        j = int(best_idx[i])
        holder = int(kf_state[j])
        if holder >= 0:
            if not bad[holder]:
                if variant == 1:
                    pt_repl[i] = holder
                elif nobs[holder] > nobs[i]:                   # pMP->Replace(pMPinKF)
                    pt_repl[i] = holder; bad[i] = True
                else:                                          # pMPinKF->Replace(pMP)
                    bad[holder] = True
                    if holder >= 1000000:
                        kfmp_repl[holder - 1000000] = i
                    else:
                        pt_repl[holder] = i
        else:
            if pt_obs[i] < 0:                                  # AddObservation is a no-op for an already observing point
                pt_obs[i] = j; nobs[i] += 1
            kf_state[j] = i
        nfused += 1
    return dict(kf_state=kf_state, pt_obs=pt_obs, pt_repl=pt_repl, kfmp_repl=kfmp_repl, nfused=nfused)


def sim3_case(scene, seed=21):
    """KF1 = frame A, KF2 = frame B; identity rotations, s12 = 1, t12 = (-7, -4, 0): a map point (x, y, 1) of KF1 lands at
    (x + 7, y + 4) in KF2 and vice versa."""
    ka, kb = scene["ka"], scene["kb"]
    rng = SplitMix64(seed)

    def side(k, s):
        n = len(k)
        has = np.array([(0 if r % 5 == 0 else 2 if r % 31 == 0 else 1) for r in (rng.next() for _ in range(n))], np.uint8)
        world = np.stack([k["x"], k["y"], np.ones(n, np.float32)], 1).astype(np.float32)
        return has, world
    has1, world1 = side(ka, 0)
    has2, world2 = side(kb, 1)
    t12 = np.array([-7.0, -4.0, 0.0], np.float32)
    u12, v12 = (ka["x"] + np.float32(7.0)).astype(np.float32), (ka["y"] + np.float32(4.0)).astype(np.float32)
    u21, v21 = (kb["x"] + t12[0]).astype(np.float32), (kb["y"] + t12[1]).astype(np.float32)
    d1 = np.sqrt(u12.astype(np.float64) ** 2 + v12.astype(np.float64) ** 2 + 1.0)
    d2 = np.sqrt(u21.astype(np.float64) ** 2 + v21.astype(np.float64) ** 2 + 1.0)
    lvl1, maxd1, mind1 = _levels_and_ranges(ka["octave"], d1, seed + 1)
    lvl2, maxd2, mind2 = _levels_and_ranges(kb["octave"], d2, seed + 2)
    pre12 = np.full(len(ka), -1, np.int32)
    good2 = np.flatnonzero(has2 == 1)
    for j, i in enumerate(range(3, len(ka), 41)):
        pre12[i] = good2[(j * 7) % len(good2)]
    return dict(has1=has1, world1=world1, maxd1=maxd1, mind1=mind1, lvl1=lvl1, u12=u12, v12=v12, has2=has2, world2=world2, maxd2=maxd2,
                mind2=mind2, lvl2=lvl2, u21=u21, v21=v21, pre12=pre12, t12=t12, desc1=scene["da"], desc2=scene["db"])


def sim3_flat(matcher, kf1, kf2, scene, c, th):
    """SearchBySim3 around the flat C ABI: two SearchWindowBest calls and the mutual-agreement pass (:1308-1326)."""
    N1, N2 = kf1.N, kf2.N
    already1 = c["pre12"] >= 0
    already2 = np.zeros(N2, bool)
    already2[c["pre12"][already1]] = True
    in1 = (c["u21"] >= 0) & (c["u21"] < W) & (c["v21"] >= 0) & (c["v21"] < H)      # KF2's points inside KF1's image
    in2 = (c["u12"] >= 0) & (c["u12"] < W) & (c["v12"] >= 0) & (c["v12"] < H)
    v1 = (c["has1"] == 1) & ~already1 & in2
    v2 = (c["has2"] == 1) & ~already2 & in1
    r1 = (np.float32(th) * scene["scale"][c["lvl1"]]).astype(np.float32)
    r2 = (np.float32(th) * scene["scale"][c["lvl2"]]).astype(np.float32)
    b1, d1 = matcher.SearchWindowBest(kf2, c["u12"], c["v12"], r1, c["lvl1"] - 1, c["lvl1"], c["desc1"], valid=v1.astype(np.uint8))
    b2, d2 = matcher.SearchWindowBest(kf1, c["u21"], c["v21"], r2, c["lvl2"] - 1, c["lvl2"], c["desc2"], valid=v2.astype(np.uint8))
    m1 = np.where(v1 & (d1 <= TH_HIGH), b1, -1)
    m2 = np.where(v2 & (d2 <= TH_HIGH), b2, -1)
    m12 = c["pre12"].copy()
    nfound = 0
    for i1 in range(N1):
        if m1[i1] >= 0 and m2[m1[i1]] == i1:
            m12[i1] = m1[i1]; nfound += 1
    return nfound, m12
