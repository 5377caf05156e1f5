"""CPU: the order-free form of the best / runner-up update of ORBmatcher::SearchByProjection (ORBmatcher.cc:80-120)
that k_search.cu relies on, checked exhaustively-at-random against the literal sequential loop; and the C++ oracle's
orc_search_by_projection against a literal Python restatement of the same reference lines on small frames."""
import math
import random

import numpy as np

from orbslam3lib_b200 import scenes


def _sequential(el):
    bd, bl, bd2, bl2, bi = 256, -1, 256, -1, -1
    for i, (d, l) in enumerate(el):
        if d < bd:
            bd2, bd, bl2, bl, bi = bd, d, bl, l, i
        elif d < bd2:
            bl2, bd2 = l, d
    return bd, bl, bi, bd2, bl2


def _order_free(el):
    keys = [(d, i) for i, (d, _) in enumerate(el)]
    if not keys or min(keys)[0] >= 256:
        return 256, -1, -1, 256, -1
    m1, p1 = min(keys)
    before = [k for k in keys if k[1] < p1]
    after = [k for k in keys if k[1] > p1]
    mp, ms = (min(before) if before else None), (min(after) if after else None)
    sec = mp if mp is not None and (ms is None or mp[0] <= ms[0]) else ms
    if sec is None or sec[0] >= 256:
        return m1, el[p1][1], p1, 256, -1
    return m1, el[p1][1], p1, sec[0], el[sec[1]][1]


def test_runner_up_rule_equals_sequential_loop():
    rng = random.Random(7)
    for _ in range(60000):
        n = rng.randint(0, 9)
        el = [(rng.choice([rng.randint(0, 5), rng.randint(0, 256)]), rng.randint(0, 3)) for _ in range(n)]
        assert _sequential(el) == _order_free(el), el


def _py_search(keys, desc, u_right, bounds, scale, pts, th, nn_ratio):
    """literal restatement of Frame.cc:405-436, 672-751 and ORBmatcher.cc:44-143 (Nleft == -1), numpy float32."""
    f32 = np.float32
    min_x, min_y, max_x, max_y = [f32(v) for v in bounds]
    gw, gh = f32(f32(64) / f32(max_x - min_x)), f32(f32(48) / f32(max_y - min_y))

    def c_round(v):
        v = float(v)
        return int(math.floor(v + 0.5) if v >= 0 else math.ceil(v - 0.5))
    grid = {}
    for i in range(len(keys)):
        px = c_round(f32(f32(keys["x"][i] - min_x) * gw))
        py = c_round(f32(f32(keys["y"][i] - min_y) * gh))
        if 0 <= px < 64 and 0 <= py < 48:
            grid.setdefault((px, py), []).append(i)
    bits = np.unpackbits(desc, axis=1)
    blocked = np.zeros(len(keys), bool)
    match = np.full(len(pts), -1, np.int32)
    for ip in range(len(pts)):
        mp = pts[ip]
        if not (mp["flags"] & 1) or (mp["flags"] & 2):
            continue
        lvl = int(mp["level"])
        r = f32(2.5) if float(mp["view_cos"]) > 0.998 else f32(4.0)
        if th != 1.0:
            r = f32(r * f32(th))
        rr = f32(r * f32(scale[lvl]))
        x, y = f32(mp["proj_x"]), f32(mp["proj_y"])
        cx0 = max(0, int(math.floor(f32(f32(f32(x - min_x) - rr) * gw))))
        cx1 = min(63, int(math.ceil(f32(f32(f32(x - min_x) + rr) * gw))))
        cy0 = max(0, int(math.floor(f32(f32(f32(y - min_y) - rr) * gh))))
        cy1 = min(47, int(math.ceil(f32(f32(f32(y - min_y) + rr) * gh))))
        if cx0 >= 64 or cx1 < 0 or cy0 >= 48 or cy1 < 0:
            continue
        cand = []
        for ix in range(cx0, cx1 + 1):
            for iy in range(cy0, cy1 + 1):
                for j in grid.get((ix, iy), []):
                    o = int(keys["octave"][j])
                    if o < lvl - 1 or o > lvl:
                        continue
                    if abs(f32(keys["x"][j] - x)) < rr and abs(f32(keys["y"][j] - y)) < rr:
                        cand.append(j)
        mbits = np.unpackbits(mp["desc"])
        bd, bl, bd2, bl2, bi = 256, -1, 256, -1, -1
        for j in cand:
            if blocked[j]:
                continue
            if u_right is not None and u_right[j] > 0 and abs(f32(mp["proj_xr"] - u_right[j])) > rr:
                continue
            d = int(np.count_nonzero(mbits != bits[j]))
            if d < bd:
                bd2, bd, bl2, bl, bi = bd, d, bl, int(keys["octave"][j]), j
            elif d < bd2:
                bl2, bd2 = int(keys["octave"][j]), d
        if bd <= 100:
            lim = f32(f32(nn_ratio) * f32(bd2))
            if bl == bl2 and f32(bd) > lim:
                continue
            if bl != bl2 or f32(bd) <= lim:
                match[ip] = bi
                blocked[bi] = bool(mp["flags"] & 4)
    return match


def test_oracle_search_by_projection_vs_python(oracle):
    for seed, th in ((11, 1.0), (12, 3.0), (13, 8.0)):
        img = scenes.make("texture", 320, 240, 6000 + seed)
        ex = oracle.OracleExtractor(500, 1.2, 8, 20, 7)
        _, kps, desc = ex(img)
        u_right = np.where(scenes.rand_u32(len(kps), seed) % 3 == 0, -1.0, kps["x"] - 7.5).astype(np.float32)
        pts = scenes.track_points(kps, desc, u_right, 8, 400, seed)
        m, owner, nm = oracle.search_by_projection(kps, desc, u_right, None, (0, 0, 320, 240), ex.scale, pts, th=th,
                                                   nn_ratio=0.8)
        ref = _py_search(kps, desc, u_right, (0, 0, 320, 240), ex.scale, pts, th, 0.8)
        np.testing.assert_array_equal(m, ref)
        assert nm == int((m >= 0).sum()) and nm > 100
        for j in np.unique(m[m >= 0]):
            assert owner[j] == np.nonzero(m == j)[0].max()


def _py_search_stereo(keys, desc, nleft, ltr, rtl, bounds, scale, pts, th, nn_ratio):
    """literal restatement of ORBmatcher.cc:44-214 for Nleft != -1 (both branches), numpy float32."""
    f32 = np.float32
    min_x, min_y, max_x, max_y = [f32(v) for v in bounds]
    gw, gh = f32(f32(64) / f32(max_x - min_x)), f32(f32(48) / f32(max_y - min_y))

    def c_round(v):
        v = float(v)
        return int(math.floor(v + 0.5) if v >= 0 else math.ceil(v - 0.5))
    grid, grid_r = {}, {}
    for i in range(len(keys)):
        px = c_round(f32(f32(keys["x"][i] - min_x) * gw))
        py = c_round(f32(f32(keys["y"][i] - min_y) * gh))
        if 0 <= px < 64 and 0 <= py < 48:
            if i < nleft:
                grid.setdefault((px, py), []).append(i)
            else:
                grid_r.setdefault((px, py), []).append(i - nleft)
    bits = np.unpackbits(desc, axis=1)
    blocked = np.zeros(len(keys), bool)
    ml = np.full(len(pts), -1, np.int32)
    mr = np.full(len(pts), -1, np.int32)
    nmatches = 0

    def area(x, y, rr, lvl, right):
        cx0 = max(0, int(math.floor(f32(f32(f32(x - min_x) - rr) * gw))))
        cx1 = min(63, int(math.ceil(f32(f32(f32(x - min_x) + rr) * gw))))
        cy0 = max(0, int(math.floor(f32(f32(f32(y - min_y) - rr) * gh))))
        cy1 = min(47, int(math.ceil(f32(f32(f32(y - min_y) + rr) * gh))))
        out = []
        if cx0 >= 64 or cx1 < 0 or cy0 >= 48 or cy1 < 0:
            return out
        g, off = (grid_r, nleft) if right else (grid, 0)
        for ix in range(cx0, cx1 + 1):
            for iy in range(cy0, cy1 + 1):
                for j in g.get((ix, iy), []):
                    o = int(keys["octave"][off + j])
                    if o < lvl - 1 or o > lvl:
                        continue
                    if abs(f32(keys["x"][off + j] - x)) < rr and abs(f32(keys["y"][off + j] - y)) < rr:
                        out.append(j)
        return out

    def best_of(cand, off, mbits):
        bd, bl, bd2, bl2, bi = 256, -1, 256, -1, -1
        for j in cand:
            if blocked[off + j]:
                continue
            d = int(np.count_nonzero(mbits != bits[off + j]))
            if d < bd:
                bd2, bd, bl2, bl, bi = bd, d, bl, int(keys["octave"][off + j]), j
            elif d < bd2:
                bl2, bd2 = int(keys["octave"][off + j]), d
        return bd, bl, bd2, bl2, bi
    for ip in range(len(pts)):
        mp = pts[ip]
        fl = int(mp["flags"])
        if not (fl & 1) and not (fl & 8):
            continue
        if fl & 2:
            continue
        obs = bool(fl & 4)
        mbits = np.unpackbits(mp["desc"])
        skip_right = False
        if fl & 1:
            lvl = int(mp["level"])
            r = f32(2.5) if float(mp["view_cos"]) > 0.998 else f32(4.0)
            if th != 1.0:
                r = f32(r * f32(th))
            cand = area(f32(mp["proj_x"]), f32(mp["proj_y"]), f32(r * f32(scale[lvl])), lvl, False)
            if cand:
                bd, bl, bd2, bl2, bi = best_of(cand, 0, mbits)
                if bd <= 100:
                    lim = f32(f32(nn_ratio) * f32(bd2))
                    if bl == bl2 and f32(bd) > lim:
                        skip_right = True
                    elif bl != bl2 or f32(bd) <= lim:
                        ml[ip] = bi
                        blocked[bi] = obs
                        if ltr[bi] != -1:
                            blocked[ltr[bi] + nleft] = obs
                            nmatches += 1
                        nmatches += 1
        if skip_right:
            continue
        if fl & 8:
            lvl = int(mp["level_r"])
            if lvl != -1:
                r = f32(2.5) if float(mp["view_cos_r"]) > 0.998 else f32(4.0)
                cand = area(f32(mp["proj_xr"]), f32(mp["proj_yr"]), f32(r * f32(scale[lvl])), lvl, True)
                if not cand:
                    continue
                bd, bl, bd2, bl2, bi = best_of(cand, nleft, mbits)
                if bd <= 100:
                    if bl == bl2 and f32(bd) > f32(f32(nn_ratio) * f32(bd2)):
                        continue
                    if rtl[bi] != -1:
                        blocked[rtl[bi]] = obs
                        nmatches += 1
                    mr[ip] = bi
                    blocked[bi + nleft] = obs
                    nmatches += 1
    return ml, mr, nmatches


def test_oracle_search_by_projection_stereo_vs_python(oracle):
    for seed, th in ((21, 1.0), (22, 4.0)):
        left = scenes.make("texture", 320, 240, 6100 + seed)
        right = scenes.stereo_right_rectified(left, 6100 + seed)
        ex = oracle.OracleExtractor(400, 1.2, 8, 20, 7)
        _, kl, dl = ex(left)
        _, kr, dr = ex(right)
        keys, desc, nleft, ltr, rtl = scenes.stereo_camera_frame(kl, dl, kr, dr, seed)
        pts = scenes.track_points_stereo(keys, desc, nleft, ltr, 8, 500, seed)
        ml, mr, owner, nm = oracle.search_by_projection_stereo(keys, desc, None, nleft, ltr, rtl, (0, 0, 320, 240), ex.scale,
                                                               pts, th=th, nn_ratio=0.8)
        rl, rr, rn = _py_search_stereo(keys, desc, nleft, ltr, rtl, (0, 0, 320, 240), ex.scale, pts, th, 0.8)
        np.testing.assert_array_equal(ml, rl)
        np.testing.assert_array_equal(mr, rr)
        assert nm == rn and (ml >= 0).sum() > 100 and (mr >= 0).sum() > 20


def _py_search_init(k1, d1, k2, d2, bounds, prev, window, nn_ratio, check_ori):
    """literal restatement of ORBmatcher.cc:649-764 + ComputeThreeMaxima (:2061-2102), numpy float32."""
    f32 = np.float32
    min_x, min_y, max_x, max_y = [f32(v) for v in bounds]
    gw, gh = f32(f32(64) / f32(max_x - min_x)), f32(f32(48) / f32(max_y - min_y))

    def c_round(v):
        v = float(v)
        return int(math.floor(v + 0.5) if v >= 0 else math.ceil(v - 0.5))
    grid = {}
    for i in range(len(k2)):
        px = c_round(f32(f32(k2["x"][i] - min_x) * gw))
        py = c_round(f32(f32(k2["y"][i] - min_y) * gh))
        if 0 <= px < 64 and 0 <= py < 48:
            grid.setdefault((px, py), []).append(i)
    b1, b2 = np.unpackbits(d1, axis=1), np.unpackbits(d2, axis=1)
    prev = np.array(prev, np.float32).copy()
    m12 = np.full(len(k1), -1, np.int32)
    m21 = np.full(len(k2), -1, np.int32)
    mdist = np.full(len(k2), 2**31 - 1, np.int64)
    hist = [[] for _ in range(30)]
    factor = f32(1.0) / f32(30)
    r = f32(window)
    nmatches = 0
    for i1 in range(len(k1)):
        if k1["octave"][i1] > 0:
            continue
        x, y = prev[i1]
        cx0 = max(0, int(math.floor(f32(f32(f32(x - min_x) - r) * gw))))
        cx1 = min(63, int(math.ceil(f32(f32(f32(x - min_x) + r) * gw))))
        cy0 = max(0, int(math.floor(f32(f32(f32(y - min_y) - r) * gh))))
        cy1 = min(47, int(math.ceil(f32(f32(f32(y - min_y) + r) * gh))))
        if cx0 >= 64 or cx1 < 0 or cy0 >= 48 or cy1 < 0:
            continue
        cand = []
        for ix in range(cx0, cx1 + 1):
            for iy in range(cy0, cy1 + 1):
                for j in grid.get((ix, iy), []):
                    if k2["octave"][j] != 0:
                        continue
                    if abs(f32(k2["x"][j] - x)) < r and abs(f32(k2["y"][j] - y)) < r:
                        cand.append(j)
        if not cand:
            continue
        bd, bd2, bi = 2**31 - 1, 2**31 - 1, -1
        for j in cand:
            d = int(np.count_nonzero(b1[i1] != b2[j]))
            if mdist[j] <= d:
                continue
            if d < bd:
                bd2, bd, bi = bd, d, j
            elif d < bd2:
                bd2 = d
        if bd <= 50 and f32(bd) < f32(f32(bd2) * f32(nn_ratio)):
            if m21[bi] >= 0:
                m12[m21[bi]] = -1
                nmatches -= 1
            m12[i1], m21[bi], mdist[bi] = bi, i1, bd
            nmatches += 1
            if check_ori:
                rot = f32(k1["angle"][i1] - k2["angle"][bi])
                if rot < 0:
                    rot = f32(rot + f32(360))
                b = c_round(f32(rot * factor))
                hist[0 if b == 30 else b].append(i1)
    if check_ori:
        mx = [0, 0, 0]
        ind = [-1, -1, -1]
        for i in range(30):
            s = len(hist[i])
            if s > mx[0]:
                mx, ind = [s, mx[0], mx[1]], [i, ind[0], ind[1]]
            elif s > mx[1]:
                mx, ind = [mx[0], s, mx[1]], [ind[0], i, ind[1]]
            elif s > mx[2]:
                mx[2], ind[2] = s, i
        if mx[1] < f32(0.1) * f32(mx[0]):
            ind[1] = ind[2] = -1
        elif mx[2] < f32(0.1) * f32(mx[0]):
            ind[2] = -1
        for i in range(30):
            if i in ind:
                continue
            for i1 in hist[i]:
                if m12[i1] >= 0:
                    m12[i1] = -1
                    nmatches -= 1
    for i1 in range(len(k1)):
        if m12[i1] >= 0:
            prev[i1] = (k2["x"][m12[i1]], k2["y"][m12[i1]])
    return m12, prev, nmatches


def test_oracle_search_for_initialization_vs_python(oracle):
    cases = [("texture", 320, 240, 7100, 100, 0.9, True), ("texture", 320, 240, 7101, 30, 0.9, False),
             ("checker", 320, 240, 7102, 100, 0.95, True), ("texture", 320, 240, 7103, 10, 0.9, True)]
    total = 0
    for kind, w, h, seed, window, ratio, ori in cases:
        cur = scenes.make(kind, w, h, seed)
        prv = scenes.previous_frame(cur, seed)
        ex = oracle.OracleExtractor(1000, 1.2, 8, 20, 7)
        _, k1, d1 = ex(prv)
        _, k2, d2 = ex(cur)
        pm = np.stack([k1["x"], k1["y"]], 1)
        m12, pm2, nm = oracle.search_for_initialization(k1, d1, k2, d2, (0, 0, w, h), pm, window, ratio, ori)
        rm12, rpm, rnm = _py_search_init(k1, d1, k2, d2, (0, 0, w, h), pm, window, ratio, ori)
        np.testing.assert_array_equal(m12, rm12)
        np.testing.assert_array_equal(pm2, rpm)
        assert nm == rnm == int((m12 >= 0).sum())
        total += nm
    assert total > 150


def _py_search_frame(keys, desc, u_right, occupied, nleft, bounds, scale, mbf, pts, th, direction, check_ori, keyframe=False,
                     orb_dist=100):
    """literal restatement of ORBmatcher.cc:1683-1921 (LastFrame form) / :1923-2059 (KeyFrame form) with Frame.cc:405-436,
    672-751 and ComputeThreeMaxima :2061-2102, numpy float32.  Returns (mvpMapPoints as point indices, nmatches)."""
    f32 = np.float32
    n = len(keys)
    min_x, min_y, max_x, max_y = [f32(v) for v in bounds]
    gw, gh = f32(f32(64) / f32(max_x - min_x)), f32(f32(48) / f32(max_y - min_y))

    def c_round(v):
        v = float(v)
        return int(math.floor(v + 0.5) if v >= 0 else math.ceil(v - 0.5))
    grids = ({}, {})
    for i in range(n):
        px = c_round(f32(f32(keys["x"][i] - min_x) * gw))
        py = c_round(f32(f32(keys["y"][i] - min_y) * gh))
        if 0 <= px < 64 and 0 <= py < 48:
            if i < nleft:
                grids[0].setdefault((px, py), []).append(i)
            else:
                grids[1].setdefault((px, py), []).append(i - nleft)

    def area(x, y, r, min_level, max_level, right):
        out = []
        cx0 = max(0, int(math.floor(f32(f32(f32(x - min_x) - r) * gw))))
        cx1 = min(63, int(math.ceil(f32(f32(f32(x - min_x) + r) * gw))))
        cy0 = max(0, int(math.floor(f32(f32(f32(y - min_y) - r) * gh))))
        cy1 = min(47, int(math.ceil(f32(f32(f32(y - min_y) + r) * gh))))
        if cx0 >= 64 or cx1 < 0 or cy0 >= 48 or cy1 < 0:
            return out
        check = min_level > 0 or max_level >= 0
        for ix in range(cx0, cx1 + 1):
            for iy in range(cy0, cy1 + 1):
                for j in grids[1 if right else 0].get((ix, iy), []):
                    k = keys[nleft + j] if right else keys[j]
                    if check:
                        if k["octave"] < min_level:
                            continue
                        if max_level >= 0 and k["octave"] > max_level:
                            continue
                    if abs(f32(k["x"] - x)) < r and abs(f32(k["y"] - y)) < r:
                        out.append(j)
        return out
    bits = np.unpackbits(desc, axis=1)
    blocked = np.zeros(n, bool) if occupied is None else np.asarray(occupied).astype(bool).copy()
    mvp = np.full(n, -1, np.int64)
    rot_hist = [[] for _ in range(30)]
    factor = f32(1.0) / f32(30)
    nmatches = 0
    stereo_cam = nleft != n

    def levels(o):
        if keyframe:
            return o - 1, o + 1
        return (o, -1) if direction == 1 else ((0, o) if direction == 2 else (o - 1, o + 1))

    def push(angle_src, slot):
        rot = f32(f32(angle_src) - f32(keys["angle"][slot]))
        if keyframe and (rot < 0.0 or rot > 360.0):
            rot = f32(math.fmod(float(rot), 360.0))
        if rot < 0.0:
            rot = f32(rot + f32(360.0))
        b = c_round(f32(rot * factor))
        if b == 30:
            b = 0
        rot_hist[b].append(slot)
    for ip in range(len(pts)):
        p = pts[ip]
        if not (p["flags"] & 1):
            continue
        u, v = f32(p["u"]), f32(p["v"])
        if u < min_x or u > max_x or v < min_y or v > max_y:
            continue
        o = int(p["level"])
        radius = f32(f32(th) * f32(scale[o]))
        lo, hi = levels(o)
        cand = area(u, v, radius, lo, hi, False)
        if not cand:
            continue
        mbits = np.unpackbits(p["desc"])
        blocks = True if keyframe else bool(p["flags"] & 4)
        bd, bi = 256, -1
        for j in cand:
            if blocked[j]:
                continue
            if not keyframe and not stereo_cam and u_right is not None and u_right[j] > 0:
                ur = f32(u - f32(f32(mbf) * f32(p["invz"])))
                if abs(f32(ur - u_right[j])) > radius:
                    continue
            d = int(np.count_nonzero(mbits != bits[j]))
            if d < bd:
                bd, bi = d, j
        if bd <= (orb_dist if keyframe else 100):
            mvp[bi] = ip
            blocked[bi] = blocks
            nmatches += 1
            if check_ori:
                push(p["angle"], bi)
        if stereo_cam and not keyframe:
            cand = area(f32(p["ur"]), f32(p["vr"]), radius, lo, hi, True)
            bd, bi = 256, -1
            for j in cand:
                if blocked[j + nleft]:
                    continue
                d = int(np.count_nonzero(mbits != bits[j + nleft]))
                if d < bd:
                    bd, bi = d, j
            if bd <= 100:
                mvp[bi + nleft] = ip
                blocked[bi + nleft] = blocks
                nmatches += 1
                if check_ori:
                    push(p["angle"], bi + nleft)
    if check_ori:
        m1 = m2 = m3 = 0
        i1 = i2 = i3 = -1
        for i in range(30):
            s = len(rot_hist[i])
            if s > m1:
                m3, m2, m1, i3, i2, i1 = m2, m1, s, i2, i1, i
            elif s > m2:
                m3, m2, i3, i2 = m2, s, i2, i
            elif s > m3:
                m3, i3 = s, i
        if m2 < f32(0.1) * f32(m1):
            i2 = i3 = -1
        elif m3 < f32(0.1) * f32(m1):
            i3 = -1
        for i in range(30):
            if i not in (i1, i2, i3):
                for slot in rot_hist[i]:
                    mvp[slot] = -1
                    nmatches -= 1
    return mvp, nmatches


def test_oracle_search_by_projection_frame_vs_python(oracle):
    """mono / rectified frames (mvuRight gate, all three level modes) and stereo-camera frames (right search)."""
    for seed, th, direction in ((21, 7.0, 0), (22, 15.0, 1), (23, 15.0, 2), (24, 4.0, 0)):
        img = scenes.make("texture", 320, 240, 6100 + seed)
        ex = oracle.OracleExtractor(500, 1.2, 8, 20, 7)
        _, kps, desc = ex(img)
        n = len(kps)
        u_right = np.where(scenes.rand_u32(n, seed) % 3 == 0, -1.0, kps["x"] - 7.5).astype(np.float32)
        occupied = (scenes.rand_u32(n, seed + 5) % 12 == 0).astype(np.uint8)
        pts = scenes.frame_points(kps, desc, n, None, u_right, 8, 350, seed)
        ml, mr, owner, nm = oracle.search_by_projection_frame(kps, desc, u_right, occupied, n, (0, 0, 320, 240), ex.scale, 47.9,
                                                              pts, th=th, direction=direction)
        ref_owner, ref_nm = _py_search_frame(kps, desc, u_right, occupied, n, (0, 0, 320, 240), ex.scale, 47.9, pts, th,
                                             direction, True)
        np.testing.assert_array_equal(owner, ref_owner)
        assert nm == ref_nm and nm > 60 and (mr == -1).all()
    # stereo-camera frame: left features then right features
    imgl = scenes.make("texture", 320, 240, 6130)
    imgr = scenes.stereo_right(imgl, 6130)
    ex = oracle.OracleExtractor(400, 1.2, 8, 20, 7)
    _, kl, dl = ex(imgl)
    _, kr, dr = ex(imgr)
    keys, desc, nleft, ltr, rtl = scenes.stereo_camera_frame(kl, dl, kr, dr, 6130)
    for seed, th, direction, ori in ((31, 7.0, 0, True), (32, 12.0, 1, True), (33, 7.0, 0, False)):
        pts = scenes.frame_points(keys, desc, nleft, ltr, None, 8, 400, seed)
        ml, mr, owner, nm = oracle.search_by_projection_frame(keys, desc, None, None, nleft, (0, 0, 320, 240), ex.scale, 0.0, pts,
                                                              th=th, direction=direction, check_orientation=ori)
        ref_owner, ref_nm = _py_search_frame(keys, desc, None, None, nleft, (0, 0, 320, 240), ex.scale, 0.0, pts, th, direction,
                                             ori)
        np.testing.assert_array_equal(owner, ref_owner)
        assert nm == ref_nm and (mr >= 0).sum() > 20


def test_oracle_search_by_projection_keyframe_vs_python(oracle):
    for seed, th, dist in ((41, 10.0, 100), (42, 3.0, 64)):
        img = scenes.make("texture", 320, 240, 6200 + seed)
        ex = oracle.OracleExtractor(500, 1.2, 8, 20, 7)
        _, kps, desc = ex(img)
        n = len(kps)
        occupied = (scenes.rand_u32(n, seed + 5) % 6 == 0).astype(np.uint8)
        pts = scenes.frame_points(kps, desc, n, None, None, 8, 350, seed)
        m, owner, nm = oracle.search_by_projection_keyframe(kps, desc, occupied, (0, 0, 320, 240), ex.scale, pts, th=th,
                                                            orb_dist=dist)
        ref_owner, ref_nm = _py_search_frame(kps, desc, None, occupied, n, (0, 0, 320, 240), ex.scale, 0.0, pts, th, 0, True,
                                             keyframe=True, orb_dist=dist)
        np.testing.assert_array_equal(owner, ref_owner)
        assert nm == ref_nm and nm > 40
        assert (owner[occupied.astype(bool)] == -1).all()    # taken slots are never reassigned


def _py_search_bow(dk, ak, vk, fvk, df, af, nleft, fvf, ratio, check_ori):
    """literal restatement of ORBmatcher.cc:224-425 with std::map feature vectors as dicts."""
    f32 = np.float32
    kf = {int(i): [int(x) for x in fvk[2][fvk[1][j]:fvk[1][j + 1]]] for j, i in enumerate(fvk[0])}
    fr = {int(i): [int(x) for x in fvf[2][fvf[1][j]:fvf[1][j + 1]]] for j, i in enumerate(fvf[0])}
    bk, bf = np.unpackbits(dk, axis=1), np.unpackbits(df, axis=1)
    vp = np.full(len(df), -1, np.int64)
    hist = [[] for _ in range(30)]
    nm = 0

    def c_round(v):
        v = float(v)
        return int(math.floor(v + 0.5) if v >= 0 else math.ceil(v - 0.5))

    def push(ikf, slot):
        rot = f32(f32(ak[ikf]) - f32(af[slot]))
        if rot < 0.0:
            rot = f32(rot + f32(360.0))
        b = c_round(f32(rot * (f32(1.0) / f32(30))))
        hist[0 if b == 30 else b].append(slot)
    for node in sorted(set(kf) & set(fr)):
        for ikf in kf[node]:
            if not vk[ikf]:
                continue
            b1, i1, b2, b1r, i1r, b2r = 256, -1, 256, 256, -1, 256
            for jf in fr[node]:
                if vp[jf] >= 0:
                    continue
                d = int(np.count_nonzero(bk[ikf] != bf[jf]))
                if nleft == -1 or jf < nleft:
                    if d < b1:
                        b2, b1, i1 = b1, d, jf
                    elif d < b2:
                        b2 = d
                else:
                    if d < b1r:
                        b2r, b1r, i1r = b1r, d, jf
                    elif d < b2r:
                        b2r = d
            if b1 <= 50:
                if f32(b1) < f32(f32(ratio) * f32(b2)):
                    vp[i1] = ikf
                    if check_ori:
                        push(ikf, i1)
                    nm += 1
                if b1r <= 50:
                    vp[i1r] = ikf
                    if check_ori:
                        push(ikf, i1r)
                    nm += 1
    if check_ori:
        m1 = m2 = m3 = 0
        i1 = i2 = i3 = -1
        for i in range(30):
            s = len(hist[i])
            if s > m1:
                m3, m2, m1, i3, i2, i1 = m2, m1, s, i2, i1, i
            elif s > m2:
                m3, m2, i3, i2 = m2, s, i2, i
            elif s > m3:
                m3, i3 = s, i
        if m2 < f32(0.1) * f32(m1):
            i2 = i3 = -1
        elif m3 < f32(0.1) * f32(m1):
            i3 = -1
        for i in range(30):
            if i not in (i1, i2, i3):
                for slot in hist[i]:
                    vp[slot] = -1
                    nm -= 1
    return vp, nm


def bow_case(oracle, seed, w=320, h=240, nf=500, stereo=False):
    """key frame = the previous frame, frame = the current one (two extractions); for stereo the frame's slots are the left
    features followed by the right ones."""
    cur = scenes.make("texture", w, h, seed)
    ex = oracle.OracleExtractor(nf, 1.2, 8, 20, 7)
    _, kk, dk = ex(scenes.previous_frame(cur, seed))
    _, kf, df = ex(cur)
    nleft = -1
    if stereo:
        _, kr, dr = ex(scenes.previous_frame(scenes.previous_frame(cur, seed), seed + 11))    # a view close to the key frame's
        nleft = len(kf)
        kf, df = np.concatenate([kf, kr]), np.concatenate([df, dr])
    vk = (scenes.rand_u32(len(kk), seed + 3) % 8 != 0).astype(np.uint8)
    ak = np.where(scenes.rand_u32(len(kk), seed + 4) % 9 == 0, (kk["angle"] + 97.0) % 360.0, kk["angle"]).astype(np.float32)
    return dk, ak, vk, scenes.feature_vector(dk), df, kf["angle"].astype(np.float32), nleft, scenes.feature_vector(df)


def test_oracle_search_by_bow_vs_python(oracle):
    for seed, stereo, ratio, ori in ((6301, False, 0.7, True), (6302, False, 0.9, True), (6303, True, 0.7, True),
                                     (6304, True, 0.9, False)):
        dk, ak, vk, fvk, df, af, nleft, fvf = bow_case(oracle, seed, stereo=stereo)
        m, nm = oracle.search_by_bow(dk, ak, vk, fvk, df, af, nleft, fvf, ratio, ori)
        ref, rnm = _py_search_bow(dk, ak, vk, fvk, df, af, nleft, fvf, ratio, ori)
        np.testing.assert_array_equal(m, ref)
        assert nm == rnm and nm > 40
        if stereo:
            assert (m[nleft:] >= 0).sum() > 10    # right-camera slots matched through the `|| true` branch


# ---- MapPoints projected into a key frame: ORBmatcher.cc:428-647 (Sim3 projection), :1155-1462 (Fuse), :1464-1681 (SearchBySim3)
class _PyKeyFrame:
    """KeyFrame grid (the Frame's, Frame.cc:405-436), KeyFrame::GetFeaturesInArea (KeyFrame.cc:706-750) and IsInImage
    (:752-755) in numpy float32."""

    def __init__(self, keys, bounds):
        f32 = np.float32
        self.keys = keys
        self.min_x, self.min_y, self.max_x, self.max_y = [f32(v) for v in bounds]
        self.gw = f32(f32(64) / f32(self.max_x - self.min_x))
        self.gh = f32(f32(48) / f32(self.max_y - self.min_y))
        self.grid = {}
        for i in range(len(keys)):
            px = self._round(f32(f32(keys["x"][i] - self.min_x) * self.gw))
            py = self._round(f32(f32(keys["y"][i] - self.min_y) * self.gh))
            if 0 <= px < 64 and 0 <= py < 48:
                self.grid.setdefault((px, py), []).append(i)

    @staticmethod
    def _round(v):
        v = float(v)
        return int(math.floor(v + 0.5) if v >= 0 else math.ceil(v - 0.5))

    def is_in_image(self, x, y):
        return x >= self.min_x and x < self.max_x and y >= self.min_y and y < self.max_y

    def features_in_area(self, x, y, r):
        f32 = np.float32
        out = []
        cx0 = max(0, int(math.floor(f32(f32(f32(x - self.min_x) - r) * self.gw))))
        if cx0 >= 64:
            return out
        cx1 = min(63, int(math.ceil(f32(f32(f32(x - self.min_x) + r) * self.gw))))
        if cx1 < 0:
            return out
        cy0 = max(0, int(math.floor(f32(f32(f32(y - self.min_y) - r) * self.gh))))
        if cy0 >= 48:
            return out
        cy1 = min(47, int(math.ceil(f32(f32(f32(y - self.min_y) + r) * self.gh))))
        if cy1 < 0:
            return out
        for ix in range(cx0, cx1 + 1):
            for iy in range(cy0, cy1 + 1):
                for j in self.grid.get((ix, iy), []):
                    k = self.keys[j]
                    if abs(f32(k["x"] - x)) < r and abs(f32(k["y"] - y)) < r:
                        out.append(j)
        return out


def _py_kf_best(kf, bits, p, th, scale, skip=None, gate=None):
    """the candidate loop shared by :488-509, :1259-1312, :1423-1440, :1552-1575: (bestDist, bestIdx) or None when the point
    does not reach it."""
    f32 = np.float32
    if not (p["flags"] & 1):
        return None
    u, v = f32(p["u"]), f32(p["v"])
    if not kf.is_in_image(u, v):
        return None
    lvl = int(p["level"])
    radius = f32(f32(th) * f32(scale[lvl]))
    cand = kf.features_in_area(u, v, radius)
    if not cand:
        return None
    pb = np.unpackbits(p["desc"])
    best_dist, best_idx = 256, -1
    for idx in cand:
        if skip is not None and skip[idx]:
            continue
        o = int(kf.keys["octave"][idx])
        if o < lvl - 1 or o > lvl:
            continue
        if gate is not None and not gate(idx, o):
            continue
        d = int((pb != bits[idx]).sum())
        if d < best_dist:
            best_dist, best_idx = d, idx
    return best_dist, best_idx


def _py_search_sim3(keys, desc, occupied, bounds, scale, pts, th, ratio):
    f32 = np.float32
    kf = _PyKeyFrame(keys, bounds)
    bits = np.unpackbits(desc, axis=1)
    matched = np.zeros(len(keys), bool) if occupied is None else np.asarray(occupied).astype(bool).copy()
    match = np.full(len(pts), -1, np.int64)
    nm = 0
    for i in range(len(pts)):
        r = _py_kf_best(kf, bits, pts[i], th, scale, skip=matched)
        if r is None:
            continue
        if f32(r[0]) <= f32(f32(50) * f32(ratio)):
            matched[r[1]] = True
            match[i] = r[1]
            nm += 1
    return match, nm


def _py_fuse(keys, desc, u_right, bounds, scale, inv_sigma2, bf, pts, th, check):
    f32 = np.float32
    kf = _PyKeyFrame(keys, bounds)
    bits = np.unpackbits(desc, axis=1)
    best = np.full(len(pts), -1, np.int64)
    for i in range(len(pts)):
        p = pts[i]
        u, v = f32(p["u"]), f32(p["v"])
        ur = f32(u - f32(f32(bf) * f32(p["invz"])))

        def gate(idx, o):
            kpx, kpy = f32(keys["x"][idx]), f32(keys["y"][idx])
            kpr = f32(-1.0) if u_right is None else f32(u_right[idx])
            ex, ey = f32(u - kpx), f32(v - kpy)
            if kpr >= 0:
                er = f32(ur - kpr)
                e2 = f32(f32(f32(ex * ex) + f32(ey * ey)) + f32(er * er))
                return not (float(f32(e2 * f32(inv_sigma2[o]))) > 7.8)
            e2 = f32(f32(ex * ex) + f32(ey * ey))
            return not (float(f32(e2 * f32(inv_sigma2[o]))) > 5.99)
        r = _py_kf_best(kf, bits, p, th, scale, gate=gate if check else None)
        if r is not None and r[0] <= 50:
            best[i] = r[1]
    return best


def _py_search_by_sim3(k1, d1, k2, d2, b1, b2, scale, p1, p2, th):
    kf1, kf2 = _PyKeyFrame(k1, b1), _PyKeyFrame(k2, b2)
    bits1, bits2 = np.unpackbits(d1, axis=1), np.unpackbits(d2, axis=1)
    vn1, vn2 = np.full(len(k1), -1, np.int64), np.full(len(k2), -1, np.int64)
    for i1 in range(len(k1)):
        r = _py_kf_best(kf2, bits2, p1[i1], th, scale)
        if r is not None and r[1] >= 0 and r[0] <= 100:
            vn1[i1] = r[1]
    for i2 in range(len(k2)):
        r = _py_kf_best(kf1, bits1, p2[i2], th, scale)
        if r is not None and r[1] >= 0 and r[0] <= 100:
            vn2[i2] = r[1]
    m12 = np.full(len(k1), -1, np.int64)
    for i1 in range(len(k1)):
        if vn1[i1] >= 0 and vn2[vn1[i1]] == i1:
            m12[i1] = vn1[i1]
    return m12, vn1, vn2


def kf_case(oracle, seed, npts=350, w=320, h=240, nf=500):
    """one key frame (features of an extraction) and MapPoints projected into it (scenes.frame_points: jittered
    re-observations, level -1 / 0 / +1, bit flips, duplicates, some not reaching the search), with a few projections placed
    exactly on the bounds (IsInImage is >= min and < max)."""
    img = scenes.make("texture", w, h, seed)
    ex = oracle.OracleExtractor(nf, 1.2, 8, 20, 7)
    _, kps, desc = ex(img)
    pts = scenes.frame_points(kps, desc, len(kps), None, None, 8, npts, seed)
    if npts >= 8:
        pts["u"][3], pts["v"][5], pts["u"][6], pts["v"][7] = w, h, 0.0, 0.0
        pts["flags"][[3, 5, 6, 7]] |= 1
    return kps, desc, ex, pts


def test_oracle_search_by_projection_sim3_vs_python(oracle):
    for seed, th, ratio in ((51, 10.0, 1.0), (52, 4.0, 1.6), (53, 10.0, 0.5)):
        kps, desc, ex, pts = kf_case(oracle, 6400 + seed)
        occupied = (scenes.rand_u32(len(kps), seed + 5) % 7 == 0).astype(np.uint8)
        m, owner, nm = oracle.search_by_projection_sim3(kps, desc, occupied, (0, 0, 320, 240), ex.scale, pts, th, ratio)
        ref, rnm = _py_search_sim3(kps, desc, occupied, (0, 0, 320, 240), ex.scale, pts, th, ratio)
        np.testing.assert_array_equal(m, ref)
        assert nm == rnm and nm > (60 if ratio >= 1.0 else 20)
        assert m[3] == -1 and m[5] == -1                      # u == mnMaxX / v == mnMaxY are outside
        got = m[m >= 0]
        assert len(np.unique(got)) == len(got)                 # a taken slot is never taken again
        assert not occupied[got].any()
        np.testing.assert_array_equal(np.nonzero(owner >= 0)[0], np.sort(got))


def test_oracle_fuse_vs_python(oracle):
    for seed, th, check in ((61, 3.0, True), (62, 6.0, True), (63, 4.0, False)):
        kps, desc, ex, pts = kf_case(oracle, 6500 + seed)
        n = len(kps)
        u_right = np.where(scenes.rand_u32(n, seed) % 3 == 0, -1.0, kps["x"] - 8.0).astype(np.float32)
        inv_sigma2 = (1.0 / (ex.scale.astype(np.float32) ** 2)).astype(np.float32)
        bi, bd, nf = oracle.fuse(kps, desc, u_right, (0, 0, 320, 240), ex.scale, inv_sigma2, 47.9, pts, th, check)
        ref = _py_fuse(kps, desc, u_right, (0, 0, 320, 240), ex.scale, inv_sigma2, 47.9, pts, th, check)
        np.testing.assert_array_equal(bi, ref)
        assert nf == (ref >= 0).sum() and nf > 40
        assert ((bd >= 0) == (bi >= 0)).all() and bd.max() <= 50
    # the gate matters: without it more points find a feature
    kps, desc, ex, pts = kf_case(oracle, 6561)
    inv_sigma2 = (1.0 / (ex.scale.astype(np.float32) ** 2)).astype(np.float32)
    a = oracle.fuse(kps, desc, None, (0, 0, 320, 240), ex.scale, inv_sigma2, 0.0, pts, 6.0, True)[0]
    b = oracle.fuse(kps, desc, None, (0, 0, 320, 240), ex.scale, inv_sigma2, 0.0, pts, 6.0, False)[0]
    assert (a >= 0).sum() < (b >= 0).sum()


def sim3_case(oracle, seed, w=320, h=240, nf=400):
    """two key frames seeing the same scene (current / previous frame); KF1's MapPoint i1 is the re-observation of a KF2
    feature and vice versa, built with scenes.frame_points on the OTHER frame's features."""
    cur = scenes.make("texture", w, h, seed)
    ex = oracle.OracleExtractor(nf, 1.2, 8, 20, 7)
    _, k1, d1 = ex(cur)
    _, k2, d2 = ex(scenes.previous_frame(cur, seed))
    p1 = scenes.frame_points(k2, d2, len(k2), None, None, 8, len(k1), seed + 1, dup_frac=0.05)
    p2 = scenes.frame_points(k1, d1, len(k1), None, None, 8, len(k2), seed + 2, dup_frac=0.05)
    # make a good share of the pairs mutual: point i1 looks at feature j of KF2 and point j of KF2 looks back at i1
    r = scenes.rand_u32(len(k1), seed + 3)
    for i1 in range(0, len(k1), 2):
        j = int(r[i1] % len(k2))
        p1[i1]["u"], p1[i1]["v"], p1[i1]["level"], p1[i1]["desc"], p1[i1]["flags"] = k2["x"][j], k2["y"][j], k2["octave"][j], d2[j], 1
        p2[j]["u"], p2[j]["v"], p2[j]["level"], p2[j]["desc"], p2[j]["flags"] = k1["x"][i1], k1["y"][i1], k1["octave"][i1], d1[i1], 1
    return k1, d1, k2, d2, ex, p1, p2


def test_oracle_search_by_sim3_vs_python(oracle):
    for seed, th in ((6601, 7.5), (6602, 3.0)):
        k1, d1, k2, d2, ex, p1, p2 = sim3_case(oracle, seed)
        b = (0, 0, 320, 240)
        m12, v1, v2, nf = oracle.search_by_sim3(k1, d1, k2, d2, b, b, ex.scale, p1, p2, th)
        r12, r1, r2 = _py_search_by_sim3(k1, d1, k2, d2, b, b, ex.scale, p1, p2, th)
        np.testing.assert_array_equal(v1, r1)
        np.testing.assert_array_equal(v2, r2)
        np.testing.assert_array_equal(m12, r12)
        assert nf == (r12 >= 0).sum() and nf > 30
        assert (v1 >= 0).sum() > nf    # some one-directional matches fail the agreement


# ---- key frame <-> key frame through the vocabulary: ORBmatcher.cc:766-906 (SearchByBoW) and :908-1153 (SearchForTriangulation)
def _merge_nodes(fv1, fv2):
    """the std::map merge of :789-877 / :959-1121: yields (features of node in KF1, features of node in KF2) per common node."""
    ids1, st1, ix1 = fv1
    ids2, st2, ix2 = fv2
    a, b = 0, 0
    while a != len(ids1) and b != len(ids2):
        if ids1[a] == ids2[b]:
            yield ix1[st1[a]:st1[a + 1]], ix2[st2[b]:st2[b + 1]]
            a += 1
            b += 1
        elif ids1[a] < ids2[b]:
            a = int(np.searchsorted(ids1, ids2[b], side="left"))
        else:
            b = int(np.searchsorted(ids2, ids1[a], side="left"))


def _py_three_maxima(rot_hist):
    max1 = max2 = max3 = 0
    ind1 = ind2 = ind3 = -1
    for i in range(30):
        s = len(rot_hist[i])
        if s > max1:
            max3, max2, max1 = max2, max1, s
            ind3, ind2, ind1 = ind2, ind1, i
        elif s > max2:
            max3, max2 = max2, s
            ind3, ind2 = ind2, i
        elif s > max3:
            max3, ind3 = s, i
    if np.float32(max2) < np.float32(0.1) * np.float32(max1):
        ind2 = ind3 = -1
    elif np.float32(max3) < np.float32(0.1) * np.float32(max1):
        ind3 = -1
    return ind1, ind2, ind3


def _py_bin(a1, a2):
    f32 = np.float32
    rot = f32(f32(a1) - f32(a2))
    if rot < 0.0:
        rot = f32(rot + f32(360.0))
    v = float(f32(rot * (f32(1.0) / f32(30))))
    b = int(math.floor(v + 0.5))
    return 0 if b == 30 else b


def _py_search_bow_kf(d1, a1, v1, fv1, d2, a2, v2, fv2, ratio, check_ori):
    f32 = np.float32
    bits1, bits2 = np.unpackbits(d1, axis=1), np.unpackbits(d2, axis=1)
    m12 = np.full(len(d1), -1, np.int64)
    matched2 = np.zeros(len(d2), bool)
    rot_hist = [[] for _ in range(30)]
    nm = 0
    for n1, n2 in _merge_nodes(fv1, fv2):
        for idx1 in n1:
            if not v1[idx1]:
                continue
            best1, best_idx2, best2 = 256, -1, 256
            for idx2 in n2:
                if matched2[idx2] or not v2[idx2]:
                    continue
                d = int((bits1[idx1] != bits2[idx2]).sum())
                if d < best1:
                    best2, best1, best_idx2 = best1, d, idx2
                elif d < best2:
                    best2 = d
            if best1 < 50 and f32(best1) < f32(f32(ratio) * f32(best2)):
                m12[idx1] = best_idx2
                matched2[best_idx2] = True
                if check_ori:
                    rot_hist[_py_bin(a1[idx1], a2[best_idx2])].append(idx1)
                nm += 1
    if check_ori:
        keep = _py_three_maxima(rot_hist)
        for i in range(30):
            if i not in keep:
                for idx1 in rot_hist[i]:
                    m12[idx1] = -1
                    nm -= 1
    return m12, nm


def _py_search_tri(k1, d1, free1, st1, fv1, k2, d2, free2, st2, fv2, F, ep, scale, sigma2, coarse, check_ori):
    f32 = np.float32
    F = np.asarray(F, f32).reshape(3, 3)
    bits1, bits2 = np.unpackbits(d1, axis=1), np.unpackbits(d2, axis=1)
    m12 = np.full(len(k1), -1, np.int64)
    rot_hist = [[] for _ in range(30)]
    nm = 0

    def constrain(kp1, kp2, unc):
        x1, y1, x2, y2 = f32(kp1["x"]), f32(kp1["y"]), f32(kp2["x"]), f32(kp2["y"])
        a = f32(f32(f32(x1 * F[0, 0]) + f32(y1 * F[1, 0])) + F[2, 0])
        b = f32(f32(f32(x1 * F[0, 1]) + f32(y1 * F[1, 1])) + F[2, 1])
        c = f32(f32(f32(x1 * F[0, 2]) + f32(y1 * F[1, 2])) + F[2, 2])
        num = f32(f32(f32(a * x2) + f32(b * y2)) + c)
        den = f32(f32(a * a) + f32(b * b))
        if den == 0:
            return False
        dsqr = f32(f32(num * num) / den)
        return float(dsqr) < 3.84 * float(f32(unc))
    for n1, n2 in _merge_nodes(fv1, fv2):
        for idx1 in n1:
            if not free1[idx1]:
                continue
            s1 = st1 is not None and bool(st1[idx1])
            best_dist, best_idx2 = 50, -1
            for idx2 in n2:
                if not free2[idx2]:
                    continue
                s2 = st2 is not None and bool(st2[idx2])
                d = int((bits1[idx1] != bits2[idx2]).sum())
                if d > 50 or d > best_dist:
                    continue
                kp2 = k2[idx2]
                if not s1 and not s2:
                    ex, ey = f32(f32(ep[0]) - f32(kp2["x"])), f32(f32(ep[1]) - f32(kp2["y"]))
                    if f32(f32(ex * ex) + f32(ey * ey)) < f32(f32(100) * f32(scale[kp2["octave"]])):
                        continue
                if coarse or constrain(k1[idx1], kp2, sigma2[kp2["octave"]]):
                    best_idx2, best_dist = idx2, d
            if best_idx2 >= 0:
                m12[idx1] = best_idx2
                nm += 1
                if check_ori:
                    rot_hist[_py_bin(k1["angle"][idx1], k2["angle"][best_idx2])].append(idx1)
    if check_ori:
        keep = _py_three_maxima(rot_hist)
        for i in range(30):
            if i not in keep:
                for idx1 in rot_hist[i]:
                    m12[idx1] = -1
                    nm -= 1
    return m12, nm


def kf_pair_case(oracle, seed, w=320, h=240, nf=500):
    """KF1 = the current frame, KF2 = the previous one (image shifted by (+3, +2)): arrays for both BoW matchers."""
    cur = scenes.make("texture", w, h, seed)
    ex = oracle.OracleExtractor(nf, 1.2, 8, 20, 7)
    _, k1, d1 = ex(cur)
    _, k2, d2 = ex(scenes.previous_frame(cur, seed))
    k1, k2 = k1.copy(), k2.copy()
    r1, r2 = scenes.rand_u32(len(k1), seed + 3), scenes.rand_u32(len(k2), seed + 4)
    k1["angle"] = np.where(r1 % 9 == 0, (k1["angle"] + 97.0) % 360.0, k1["angle"])    # rotation-histogram losers
    has1, has2 = (r1 % 5 != 0), (r2 % 5 != 0)           # feature has a (good) MapPoint
    st1, st2 = (r1 % 3 == 0).astype(np.uint8), (r2 % 3 == 1).astype(np.uint8)    # mvuRight >= 0
    scale = ex.scale.astype(np.float32)
    # pure image translation t = (3, 2, 0): F12 = [t]x with K = I; the epipole is put inside the image to exercise its gate
    F = np.array([[0, 0, 2], [0, 0, -3], [-2, 3, 0]], np.float32)
    return dict(k1=k1, d1=d1, k2=k2, d2=d2, has1=has1.astype(np.uint8), has2=has2.astype(np.uint8), st1=st1, st2=st2,
                fv1=scenes.feature_vector(d1), fv2=scenes.feature_vector(d2), F=F, ep=np.array([w * 0.4, h * 0.45], np.float32),
                scale=scale, sigma2=(scale * scale).astype(np.float32))


def test_oracle_search_by_bow_kf_vs_python(oracle):
    for seed, ratio, ori in ((6701, 0.75, True), (6702, 0.9, True), (6703, 0.6, False)):
        c = kf_pair_case(oracle, seed)
        a1, a2 = c["k1"]["angle"].astype(np.float32), c["k2"]["angle"].astype(np.float32)
        m, nm = oracle.search_by_bow_kf(c["d1"], a1, c["has1"], c["fv1"], c["d2"], a2, c["has2"], c["fv2"], ratio, ori)
        ref, rnm = _py_search_bow_kf(c["d1"], a1, c["has1"], c["fv1"], c["d2"], a2, c["has2"], c["fv2"], ratio, ori)
        np.testing.assert_array_equal(m, ref)
        assert nm == rnm and nm > 40
        got = m[m >= 0]
        assert len(np.unique(got)) == len(got) and c["has2"][got].all() and c["has1"][m >= 0].all()


def test_oracle_search_for_triangulation_vs_python(oracle):
    for seed, coarse, ori, stereo in ((6801, False, True, False), (6802, True, True, False), (6803, False, False, True),
                                      (6804, False, True, True)):
        c = kf_pair_case(oracle, seed)
        free1, free2 = 1 - c["has1"], 1 - c["has2"]
        free1[::2], free2[::3] = 1, 1    # most features are still without a MapPoint in local mapping
        s1, s2 = (c["st1"], c["st2"]) if stereo else (None, None)
        args = (c["k1"], c["d1"], free1, s1, c["fv1"], c["k2"], c["d2"], free2, s2, c["fv2"], c["F"], c["ep"], c["scale"],
                c["sigma2"], coarse, ori)
        m, nm = oracle.search_for_triangulation(*args)
        ref, rnm = _py_search_tri(*args)
        np.testing.assert_array_equal(m, ref)
        assert nm == rnm and nm > 30
        assert free1[m >= 0].all() and free2[m[m >= 0]].all()
    # the constraint matters: coarse keeps more pairs
    c = kf_pair_case(oracle, 6811)
    free = np.ones(max(len(c["k1"]), len(c["k2"])), np.uint8)
    base = (c["k1"], c["d1"], free, None, c["fv1"], c["k2"], c["d2"], free, None, c["fv2"], c["F"], c["ep"], c["scale"], c["sigma2"])
    strict = oracle.search_for_triangulation(*base, False, False)[1]
    loose = oracle.search_for_triangulation(*base, True, False)[1]
    assert 0 < strict < loose


# ---- MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:331-401)
def observation_blocks(seed, npoints=60, max_n=70):
    """descriptor blocks of MapPoints: N observations each (1 .. max_n, a few empty), every observation = a base descriptor
    with a different number of flipped bits, so medians differ and ties occur (several rows with the same median)."""
    r = scenes.rand_u32(npoints * 3, seed).reshape(3, npoints)
    counts = (r[0] % np.uint32(max_n)).astype(np.int64) + 1
    counts[r[1] % 11 == 0] = 0
    counts[:4] = (1, 2, 3, max_n)
    off = np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)
    total = int(off[-1])
    base = scenes.rand_u32(npoints * 8, seed + 1).reshape(npoints, 8).view(np.uint8).reshape(npoints, 32)
    desc = np.repeat(base, counts, axis=0).copy()
    nflip = (scenes.rand_u32(total, seed + 2) % np.uint32(60)).astype(np.int64)
    pos = (scenes.rand_u32(total * 60, seed + 3) % np.uint32(256)).reshape(total, 60).astype(np.int64)
    for k in range(60):
        rows = np.nonzero(nflip > k)[0]
        q = pos[rows, k]
        desc[rows, q >> 3] ^= (1 << (q & 7)).astype(np.uint8)
    return desc, off


def _py_distinctive(desc, off):
    best = np.full(len(off) - 1, -1, np.int64)
    med = np.full(len(off) - 1, -1, np.int64)
    for p in range(len(off) - 1):
        blk = np.unpackbits(desc[off[p]:off[p + 1]], axis=1).astype(np.int32)
        n = len(blk)
        if n == 0:
            continue
        dist = (blk[:, None, :] != blk[None, :, :]).sum(axis=2)
        best_median, best_idx = 2 ** 31 - 1, 0
        for i in range(n):
            v = sorted(int(x) for x in dist[i])
            m = v[int(0.5 * (n - 1))]
            if m < best_median:
                best_median, best_idx = m, i
        best[p], med[p] = best_idx, best_median
    return best, med


def test_oracle_distinctive_descriptors_vs_python(oracle):
    desc, off = observation_blocks(6900)
    bi, bm = oracle.distinctive_descriptors(desc, off)
    ref_i, ref_m = _py_distinctive(desc, off)
    np.testing.assert_array_equal(bi, ref_i)
    np.testing.assert_array_equal(bm, ref_m)
    n = np.diff(off)
    assert (bi[n == 0] == -1).all() and (bi[n == 1] == 0).all() and (bm[n == 1] == 0).all()
    assert (bi[n == 2] == 0).all() and (bm[n == 2] == 0).all()    # median of (0, d) is element int(0.5) = 0: the first row wins
    assert len(np.unique(bi[n > 10])) > 5


# ---- SearchForTriangulation on two-camera (KannalaBrandt8) key frames: ORBmatcher.cc:908-1153 with :1014-1043
def _py_search_tri_kb8(oracle, c, fv1, fv2, free1, free2, rig, coarse, check_ori):
    """literal restatement of the loop with the camera / pose selection of :1015-1042; the constraint itself
    (KannalaBrandt8::epipolarConstrain = TriangulateMatches > 0.0001) is the oracle's Eigen-pinned function."""
    k1, d1, k2, d2 = c["keys1"], c["desc1"], c["keys2"], c["desc2"]
    bits1, bits2 = np.unpackbits(d1, axis=1), np.unpackbits(d2, axis=1)
    m12 = np.full(len(k1), -1, np.int64)
    rot_hist = [[] for _ in range(30)]
    nm = 0
    for n1, n2 in _merge_nodes(fv1, fv2):
        for idx1 in n1:
            if not free1[idx1]:
                continue
            right1 = idx1 >= c["nleft1"]
            best_dist, best_idx2 = 50, -1
            for idx2 in n2:
                if not free2[idx2]:
                    continue
                d = int((bits1[idx1] != bits2[idx2]).sum())
                if d > 50 or d > best_dist:
                    continue
                right2 = idx2 >= c["nleft2"]
                ok = coarse
                if not ok:
                    if right1 and right2:
                        T, cam_a, cam_b = c["t12"][3], rig["cam2"], rig["cam2"]    # Trr, mpCamera2, mpCamera2
                    elif right1 and not right2:
                        T, cam_a, cam_b = c["t12"][2], rig["cam2"], rig["cam1"]    # Trl
                    elif not right1 and right2:
                        T, cam_a, cam_b = c["t12"][1], rig["cam1"], rig["cam2"]    # Tlr
                    else:
                        T, cam_a, cam_b = c["t12"][0], rig["cam1"], rig["cam1"]    # Tll
                    g = rig.copy()
                    g["cam1"], g["cam2"], g["Rlr"], g["tlr"] = cam_a, cam_b, T[:9], T[9:]
                    z = oracle.kb_triangulate_matches(g, k1["x"][idx1], k1["y"][idx1], k2["x"][idx2], k2["y"][idx2],
                                                      c["sigma2"][k1["octave"][idx1]], c["sigma2"][k2["octave"][idx2]])
                    ok = z > np.float32(0.0001)
                if ok:
                    best_idx2, best_dist = idx2, d
            if best_idx2 >= 0:
                m12[idx1] = best_idx2
                nm += 1
                if check_ori:
                    rot_hist[_py_bin(k1["angle"][idx1], k2["angle"][best_idx2])].append(idx1)
    if check_ori:
        keep = _py_three_maxima(rot_hist)
        for i in range(30):
            if i not in keep:
                for idx1 in rot_hist[i]:
                    m12[idx1] = -1
                    nm -= 1
    return m12, nm


def kb8_tri_case(seed, w=640, h=400, npts=900):
    rig = scenes.fisheye_rig(w, h, 0.11, seed)
    c = scenes.kb8_keyframe_pair(rig, w, h, seed, npts)
    fv1, fv2 = scenes.feature_vector(c["desc1"]), scenes.feature_vector(c["desc2"])
    free1 = (scenes.rand_u32(len(c["keys1"]), seed + 20) % 6 != 0).astype(np.uint8)
    free2 = (scenes.rand_u32(len(c["keys2"]), seed + 21) % 6 != 0).astype(np.uint8)
    return rig, c, fv1, fv2, free1, free2


def test_oracle_search_for_triangulation_kb8_vs_python(oracle):
    for seed, coarse, ori in ((7001, False, True), (7002, False, False), (7003, True, True)):
        rig, c, fv1, fv2, free1, free2 = kb8_tri_case(seed, npts=500)
        m, nm, slack = oracle.search_for_triangulation_kb8(c["keys1"], c["desc1"], free1, c["nleft1"], fv1, c["keys2"], c["desc2"],
                                                           free2, c["nleft2"], fv2, rig, c["t12"], c["sigma2"], coarse, ori)
        ref, rnm = _py_search_tri_kb8(oracle, c, fv1, fv2, free1, free2, rig, coarse, ori)
        np.testing.assert_array_equal(m, ref)
        assert nm == rnm and nm > 150
        got = m >= 0
        assert free1[got].all() and free2[m[got]].all()
        # all four camera combinations occur among the matches
        r1, r2 = np.nonzero(got)[0] >= c["nleft1"], m[got] >= c["nleft2"]
        assert all(((r1 == a) & (r2 == b)).sum() > 5 for a in (False, True) for b in (False, True))
    # the constraint removes the displaced observations: fewer matches than the coarse form
    rig, c, fv1, fv2, free1, free2 = kb8_tri_case(7004, npts=500)
    a = (c["keys1"], c["desc1"], free1, c["nleft1"], fv1, c["keys2"], c["desc2"], free2, c["nleft2"], fv2, rig, c["t12"], c["sigma2"])
    assert oracle.search_for_triangulation_kb8(*a, False, False)[1] < oracle.search_for_triangulation_kb8(*a, True, False)[1]
