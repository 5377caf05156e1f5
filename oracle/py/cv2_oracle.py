"""Second oracle: GENUINE OpenCV primitives (cv2) + the ORB-SLAM glue of ORBextractor_old.cc restated in Python.

Test infrastructure only.  Its job is to pin oracle/cpp (the dependency-free restatement) against the real
OpenCV implementations of resize / copyMakeBorder / FAST / GaussianBlur / fastAtan2 / BFMatcher that the reference
links (OpenCV 4.2.0 there, cv2 4.13.0 in this container), and to generate tests/golden/.
cv2 exists only in the build container's image; nothing on the GPU run path imports this module.
"""
import math

import cv2
import numpy as np

EDGE = 19
PATCH = 31
HALF_PATCH = 15


def cv_round(v):
    """cvRound: round half to even."""
    return int(np.rint(v))


def _f32(v):
    return np.float32(v)


# ---- libstdc++ std::sort (introsort + final insertion sort), restated so that ties come out as in oracle/cpp ----
def _std_sort(a, less):
    n = len(a)
    if n == 0:
        return

    def move_median_to_first(res, x, y, z):
        if less(a[x], a[y]):
            if less(a[y], a[z]):
                a[res], a[y] = a[y], a[res]
            elif less(a[x], a[z]):
                a[res], a[z] = a[z], a[res]
            else:
                a[res], a[x] = a[x], a[res]
        elif less(a[x], a[z]):
            a[res], a[x] = a[x], a[res]
        elif less(a[y], a[z]):
            a[res], a[z] = a[z], a[res]
        else:
            a[res], a[y] = a[y], a[res]

    def unguarded_partition(first, last, pivot):
        while True:
            while less(a[first], a[pivot]):
                first += 1
            last -= 1
            while less(a[pivot], a[last]):
                last -= 1
            if not first < last:
                return first
            a[first], a[last] = a[last], a[first]
            first += 1

    def adjust_heap(first, hole, length, value):
        top = hole
        child = hole
        while child < (length - 1) // 2:
            child = 2 * (child + 1)
            if less(a[first + child], a[first + child - 1]):
                child -= 1
            a[first + hole] = a[first + child]
            hole = child
        if (length & 1) == 0 and child == (length - 2) // 2:
            child = 2 * (child + 1)
            a[first + hole] = a[first + child - 1]
            hole = child - 1
        parent = (hole - 1) // 2
        while hole > top and less(a[first + parent], value):
            a[first + hole] = a[first + parent]
            hole = parent
            parent = (hole - 1) // 2
        a[first + hole] = value

    def heap_sort(first, last):
        length = last - first
        if length >= 2:
            parent = (length - 2) // 2
            while True:
                adjust_heap(first, parent, length, a[first + parent])
                if parent == 0:
                    break
                parent -= 1
        while last - first > 1:
            last -= 1
            value = a[last]
            a[last] = a[first]
            adjust_heap(first, 0, last - first, value)

    def introsort_loop(first, last, depth):
        while last - first > 16:
            if depth == 0:
                heap_sort(first, last)
                return
            depth -= 1
            mid = first + (last - first) // 2
            move_median_to_first(first, first + 1, mid, last - 1)
            cut = unguarded_partition(first + 1, last, first)
            introsort_loop(cut, last, depth)
            last = cut

    def unguarded_linear_insert(last):
        val = a[last]
        nxt = last - 1
        while less(val, a[nxt]):
            a[last] = a[nxt]
            last = nxt
            nxt -= 1
        a[last] = val

    def insertion_sort(first, last):
        for i in range(first + 1, last):
            if less(a[i], a[first]):
                val = a[i]
                a[first + 1:i + 1] = a[first:i]
                a[first] = val
            else:
                unguarded_linear_insert(i)

    introsort_loop(0, n, 2 * (n.bit_length() - 1))
    if n > 16:
        insertion_sort(0, 16)
        for i in range(16, n):
            unguarded_linear_insert(i)
    else:
        insertion_sort(0, n)


class _Node:
    __slots__ = ("keys", "ULx", "ULy", "URx", "BRy", "no_more", "alive")

    def __init__(self):
        self.keys = []
        self.no_more = False
        self.alive = True


def _divide(n):
    """ExtractorNode::DivideNode (ORBextractor_old.cc:482-538); keys are (x, y, response, order)."""
    half_x = int(math.ceil(float(_f32(n.URx - n.ULx)) / 2))
    half_y = int(math.ceil(float(_f32(n.BRy - n.ULy)) / 2))
    c = [_Node() for _ in range(4)]
    c[0].ULx, c[0].ULy, c[0].URx, c[0].BRy = n.ULx, n.ULy, n.ULx + half_x, n.ULy + half_y
    c[1].ULx, c[1].ULy, c[1].URx, c[1].BRy = n.ULx + half_x, n.ULy, n.URx, n.ULy + half_y
    c[2].ULx, c[2].ULy, c[2].URx, c[2].BRy = n.ULx, n.ULy + half_y, n.ULx + half_x, n.BRy
    c[3].ULx, c[3].ULy, c[3].URx, c[3].BRy = n.ULx + half_x, n.ULy + half_y, n.URx, n.BRy
    mx, my = c[0].URx, c[0].BRy
    for k in n.keys:
        if k[0] < mx:
            (c[0] if k[1] < my else c[2]).keys.append(k)
        elif k[1] < my:
            c[1].keys.append(k)
        else:
            c[3].keys.append(k)
    for ch in c:
        ch.no_more = len(ch.keys) == 1
    return c


def distribute_oct_tree(keys, min_x, max_x, min_y, max_y, n_target):
    """ORBextractor::DistributeOctTree (ORBextractor_old.cc:557-781). keys: list of (x, y, response)."""
    n_ini = max(1, int(np.round(_f32(max_x - min_x) / _f32(max_y - min_y))))
    h_x = _f32(max_x - min_x) / _f32(n_ini)
    nodes = []    # front of the std::list = index 0
    ini = []
    for i in range(n_ini):
        nd = _Node()
        nd.ULx, nd.ULy = int(h_x * _f32(i)), 0
        nd.URx, nd.BRy = int(h_x * _f32(i + 1)), max_y - min_y
        nodes.append(nd)
        ini.append(nd)
    for k in keys:
        ini[int(_f32(k[0]) / h_x)].keys.append(k)
    kept = []
    for nd in nodes:
        if len(nd.keys) == 1:
            nd.no_more = True
            kept.append(nd)
        elif nd.keys:
            kept.append(nd)
    nodes = kept

    def less(e1, e2):    # compareNodes :540-555
        if e1[0] != e2[0]:
            return e1[0] < e2[0]
        return e1[1].ULx < e2[1].ULx

    finish = False
    while not finish:
        prev_size = len(nodes)
        n_to_expand = 0
        size_and_node = []
        front = []
        rest = []
        for nd in nodes:
            if nd.no_more:
                rest.append(nd)
                continue
            for ch in _divide(nd):
                if ch.keys:
                    front.insert(0, ch)
                    if len(ch.keys) > 1:
                        n_to_expand += 1
                        size_and_node.append((len(ch.keys), ch))
        nodes = front + rest
        if len(nodes) >= n_target or len(nodes) == prev_size:
            finish = True
        elif len(nodes) + n_to_expand * 3 > n_target:
            while not finish:
                prev_size = len(nodes)
                prev = list(size_and_node)
                size_and_node = []
                _std_sort(prev, less)
                for j in range(len(prev) - 1, -1, -1):
                    nd = prev[j][1]
                    for ch in _divide(nd):
                        if ch.keys:
                            nodes.insert(0, ch)
                            if len(ch.keys) > 1:
                                size_and_node.append((len(ch.keys), ch))
                    nodes.remove(nd)
                    if len(nodes) >= n_target:
                        break
                if len(nodes) >= n_target or len(nodes) == prev_size:
                    finish = True
    out = []
    for nd in nodes:
        best = nd.keys[0]
        for k in nd.keys[1:]:
            if k[2] > best[2]:
                best = k
        out.append(best)
    return out


class Cv2Extractor:
    """ORBextractor_old.cc on top of real OpenCV calls."""

    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7):
        self.nfeatures, self.nlevels, self.ini_th, self.min_th = nfeatures, nlevels, ini_th, min_th
        sf = float(_f32(scale_factor))    # double member holding the float argument
        self.scale = [_f32(1.0)]
        for _ in range(1, nlevels):
            self.scale.append(_f32(float(self.scale[-1]) * sf))
        self.inv_scale = [_f32(1.0) / s for s in self.scale]
        factor = _f32(1.0 / sf)
        n_des = _f32(nfeatures) * (_f32(1) - factor) / (_f32(1) - _f32(math.pow(float(factor), float(nlevels))))
        self.features_per_level = []
        total = 0
        for _ in range(nlevels - 1):
            self.features_per_level.append(cv_round(n_des))
            total += self.features_per_level[-1]
            n_des = _f32(n_des * factor)
        self.features_per_level.append(max(nfeatures - total, 0))
        umax = [0] * (HALF_PATCH + 1)
        vmax = int(math.floor(float(_f32(HALF_PATCH) * np.sqrt(_f32(2.0)) / _f32(2) + _f32(1))))
        vmin = int(math.ceil(float(_f32(HALF_PATCH) * np.sqrt(_f32(2.0)) / _f32(2))))
        for v in range(vmax + 1):
            umax[v] = cv_round(math.sqrt(HALF_PATCH * HALF_PATCH - v * v))
        v0 = 0
        for v in range(HALF_PATCH, vmin - 1, -1):
            while umax[v0] == umax[v0 + 1]:
                v0 += 1
            umax[v] = v0
            v0 += 1
        self.umax = umax
        self.pattern = None

    def set_pattern(self, pattern_i8):
        self.pattern = np.asarray(pattern_i8, np.int32).reshape(512, 2)

    def compute_pyramid(self, image):
        self.pyr, self.pyr_padded = [], []
        h, w = image.shape
        for level in range(self.nlevels):
            s = self.inv_scale[level]
            sz = (cv_round(_f32(w) * s), cv_round(_f32(h) * s))
            if level == 0:
                roi = image
            else:
                roi = cv2.resize(self.pyr[level - 1], sz, 0, 0, cv2.INTER_LINEAR)
            padded = cv2.copyMakeBorder(roi, EDGE, EDGE, EDGE, EDGE, cv2.BORDER_REFLECT_101)
            self.pyr.append(np.ascontiguousarray(padded[EDGE:-EDGE, EDGE:-EDGE]))
            self.pyr_padded.append(padded)

    def compute_keypoints(self):
        self.candidates, self.level_kps = [], []
        W = _f32(35)
        for level in range(self.nlevels):
            img = self.pyr[level]
            rows, cols = img.shape
            min_bx = min_by = EDGE - 3
            max_bx, max_by = cols - EDGE + 3, rows - EDGE + 3
            width, height = _f32(max_bx - min_bx), _f32(max_by - min_by)
            n_cols, n_rows = int(width / W), int(height / W)
            cand = []
            if n_cols > 0 and n_rows > 0:
                w_cell = int(math.ceil(float(width / _f32(n_cols))))
                h_cell = int(math.ceil(float(height / _f32(n_rows))))
                det_ini = cv2.FastFeatureDetector_create(self.ini_th, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
                det_min = cv2.FastFeatureDetector_create(self.min_th, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
                for i in range(n_rows):
                    ini_y = min_by + i * h_cell
                    max_y = ini_y + h_cell + 6
                    if ini_y >= max_by - 3:
                        continue
                    max_y = min(max_y, max_by)
                    for j in range(n_cols):
                        ini_x = min_bx + j * w_cell
                        max_x = ini_x + w_cell + 6
                        if ini_x >= max_bx - 6:
                            continue
                        max_x = min(max_x, max_bx)
                        cell = np.ascontiguousarray(img[ini_y:max_y, ini_x:max_x])
                        kps = det_ini.detect(cell)
                        if not kps:
                            kps = det_min.detect(cell)
                        for kp in kps:
                            cand.append((int(kp.pt[0]) + j * w_cell, int(kp.pt[1]) + i * h_cell, int(kp.response)))
            self.candidates.append(cand)
            if cand or (n_cols > 0 and n_rows > 0):
                kept = distribute_oct_tree(cand, min_bx, max_bx, min_by, max_by, self.features_per_level[level])
            else:
                kept = []
            size = float(int(_f32(PATCH) * self.scale[level]))
            out = []
            for (x, y, r) in kept:
                px, py = x + min_bx, y + min_by
                out.append([float(px), float(py), size, self.ic_angle(img, px, py), float(r), level])
            self.level_kps.append(out)

    def ic_angle(self, img, x, y):
        m01 = m10 = 0
        im = img.astype(np.int64)
        for u in range(-HALF_PATCH, HALF_PATCH + 1):
            m10 += u * int(im[y, x + u])
        for v in range(1, HALF_PATCH + 1):
            d = self.umax[v]
            vs = 0
            for u in range(-d, d + 1):
                p, m = int(im[y + v, x + u]), int(im[y - v, x + u])
                vs += p - m
                m10 += u * (p + m)
            m01 += v * vs
        return float(cv2.fastAtan2(float(m01), float(m10)))

    def describe(self, blurred, x, y, angle_deg):
        factor_pi = _f32(math.pi / float(_f32(180.0)))
        ang = _f32(_f32(angle_deg) * factor_pi)
        a, b = _f32(math.cos(float(ang))), _f32(math.sin(float(ang)))
        px = self.pattern[:, 0].astype(np.float32)
        py = self.pattern[:, 1].astype(np.float32)
        ry = np.rint((px * b).astype(np.float32) + (py * a).astype(np.float32)).astype(np.int64)
        rx = np.rint((px * a).astype(np.float32) - (py * b).astype(np.float32)).astype(np.int64)
        vals = blurred[y + ry, x + rx].astype(np.int32)
        bits = (vals[0::2] < vals[1::2]).astype(np.uint8)
        return np.packbits(bits.reshape(32, 8), axis=1, bitorder="little").reshape(32)

    def __call__(self, image, lapping=(0, 0)):
        """-> (mono_index, keypoints list [x,y,size,angle,response,octave], descriptors (n,32))"""
        self.compute_pyramid(image)
        self.compute_keypoints()
        n = sum(len(k) for k in self.level_kps)
        kps = [None] * n
        desc = np.zeros((n, 32), np.uint8)
        mono, stereo = 0, n - 1
        self.blurred = []
        for level in range(self.nlevels):
            lk = self.level_kps[level]
            if not lk:
                self.blurred.append(None)
                continue
            blurred = cv2.GaussianBlur(self.pyr[level].copy(), (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
            self.blurred.append(blurred)
            s = self.scale[level]
            for k in lk:
                d = self.describe(blurred, int(k[0]), int(k[1]), k[3])
                x, y = _f32(k[0]), _f32(k[1])
                if level != 0:
                    x, y = _f32(x * s), _f32(y * s)
                rec = [float(x), float(y), k[2], k[3], k[4], k[5]]
                if lapping[0] <= x <= lapping[1]:
                    kps[stereo], desc[stereo] = rec, d
                    stereo -= 1
                else:
                    kps[mono], desc[mono] = rec, d
                    mono += 1
        return mono, kps, desc


def knn2(query, train):
    """cv2.BFMatcher(NORM_HAMMING).knnMatch(query, train, k=2) -> (idx (nq,2) int32, dist (nq,2) int32)."""
    m = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(np.ascontiguousarray(query), np.ascontiguousarray(train), k=2)
    idx = np.full((len(m), 2), -1, np.int32)
    dist = np.full((len(m), 2), -1, np.int32)
    for i, row in enumerate(m):
        for k, dm in enumerate(row):
            idx[i, k], dist[i, k] = dm.trainIdx, int(dm.distance)
    return idx, dist


def stereo_matches_cv2(pyr_l, pyr_r, scale, inv_scale, kps_l, desc_l, kps_r, desc_r, mb, mbf):
    """Frame::ComputeStereoMatches (Frame.cc:827-997) with the genuine cv2.norm(NORM_L1) for the window SAD and
    numpy float32 for the reference's float arithmetic; an independent second statement that pins the C++ oracle.
    pyr_l / pyr_r: lists of level ROIs (mvImagePyramid); kps_*: structured arrays with x, y, octave; desc_*: (n,32) u8.
    -> (uRight f32 (n,), depth f32 (n,))"""
    f32 = np.float32
    th_high, th_low = 100, 50
    n = len(kps_l)
    u_right = np.full(n, -1.0, f32)
    depth = np.full(n, -1.0, f32)
    th_orb = (th_high + th_low) // 2
    n_rows = pyr_l[0].shape[0]
    rows = [[] for _ in range(n_rows)]
    for ir in range(len(kps_r)):
        y = f32(kps_r["y"][ir])
        r = f32(2.0) * f32(scale[int(kps_r["octave"][ir])])
        maxr = int(np.ceil(f32(y + r)))
        minr = int(np.floor(f32(y - r)))
        for yi in range(minr, maxr + 1):
            rows[yi].append(ir)
    mb, mbf = f32(mb), f32(mbf)
    min_d = f32(0)
    max_d = f32(mbf / mb)
    bits_l = np.unpackbits(np.ascontiguousarray(desc_l, np.uint8).reshape(-1, 32), axis=1)
    bits_r = np.unpackbits(np.ascontiguousarray(desc_r, np.uint8).reshape(-1, 32), axis=1)
    dist_idx = []
    for il in range(n):
        level = int(kps_l["octave"][il])
        v_l, u_l = f32(kps_l["y"][il]), f32(kps_l["x"][il])
        cand = rows[int(v_l)]
        if not cand:
            continue
        min_u, max_u = f32(u_l - max_d), f32(u_l - min_d)
        if max_u < 0:
            continue
        best, best_r = th_high, 0
        for ir in cand:
            o = int(kps_r["octave"][ir])
            if o < level - 1 or o > level + 1:
                continue
            u_r = f32(kps_r["x"][ir])
            if min_u <= u_r <= max_u:
                d = int(np.count_nonzero(bits_l[il] != bits_r[ir]))
                if d < best:
                    best, best_r = d, ir
        if best >= th_orb:
            continue
        u_r0 = f32(kps_r["x"][best_r])
        sf = f32(inv_scale[level])

        def c_round(v):    # round(): half away from zero
            v = float(v)
            return f32(np.floor(v + 0.5) if v >= 0 else np.ceil(v - 0.5))
        su_l, sv_l, su_r0 = c_round(f32(u_l * sf)), c_round(f32(v_l * sf)), c_round(f32(u_r0 * sf))
        w, L = 5, 5
        il_img = pyr_l[level][int(sv_l) - w:int(sv_l) + w + 1, int(su_l) - w:int(su_l) + w + 1]
        assert il_img.shape == (11, 11)
        ini_u, end_u = f32(su_r0 + L - w), f32(su_r0 + L + w + 1)
        if ini_u < 0 or end_u >= pyr_r[level].shape[1]:
            continue
        best_sad, best_inc = 2 ** 31 - 1, 0
        dists = np.zeros(2 * L + 1, f32)
        for inc in range(-L, L + 1):
            c0 = int(su_r0) + inc - w
            ir_img = pyr_r[level][int(sv_l) - w:int(sv_l) + w + 1, c0:c0 + 2 * w + 1]
            assert c0 >= 0 and ir_img.shape == (11, 11)
            d = f32(cv2.norm(np.ascontiguousarray(il_img), np.ascontiguousarray(ir_img), cv2.NORM_L1))
            if d < best_sad:
                best_sad, best_inc = int(d), inc
            dists[L + inc] = d
        if best_inc == -L or best_inc == L:
            continue
        d1, d2, d3 = dists[L + best_inc - 1], dists[L + best_inc], dists[L + best_inc + 1]
        with np.errstate(divide="ignore", invalid="ignore"):
            delta = f32(f32(d1 - d3) / f32(f32(2.0) * f32(f32(d1 + d3) - f32(f32(2.0) * d2))))
        if delta < -1 or delta > 1:
            continue
        best_u = f32(f32(scale[level]) * f32(f32(su_r0 + f32(best_inc)) + delta))
        disp = f32(u_l - best_u)
        if disp >= min_d and disp < max_d:
            if disp <= 0:
                disp = f32(0.01)
                best_u = f32(float(u_l) - 0.01)
            depth[il] = f32(mbf / disp)
            u_right[il] = best_u
            dist_idx.append((best_sad, il))
    if dist_idx:
        dist_idx.sort()
        median = f32(dist_idx[len(dist_idx) // 2][0])
        th = f32(f32(f32(1.5) * f32(1.4)) * median)
        for d, il in reversed(dist_idx):
            if f32(d) < th:
                break
            u_right[il] = -1
            depth[il] = -1
    return u_right, depth
