"""Second, independent restatement of ORBextractor::operator() built on the REAL OpenCV binary.

TEST INFRASTRUCTURE ONLY (pins oracle/orb_oracle.cpp; never imported by the product).

Every OpenCV primitive the reference calls is executed by cv2 itself here
(`cv2.resize`, `cv2.copyMakeBorder`, `cv2.FastFeatureDetector` on ROI views,
`cv2.GaussianBlur`, `cv2.fastAtan2`), libm's `sincosf` is called through ctypes, and only
the in-tree logic of src/ORBextractor.cc (cell loop :781-872, DistributeOctTree :555-779,
IC_Angle :76-103, computeOrbDescriptor :107-146, ordering :1120-1166) is re-expressed in
Python.  tests/test_oracle_vs_cv2.py compares this pipeline with the C++ oracle end to end;
tests/golden/make_golden.py records that they agreed when the goldens were made.
"""
import ctypes
import math
import re
from pathlib import Path

import cv2
import numpy as np

_libm = ctypes.CDLL("libm.so.6")
_libm.sincosf.argtypes = [ctypes.c_float, ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_float)]

EDGE = 19
HALF_PATCH = 15
F32 = np.float32


def _pattern():
    txt = (Path(__file__).resolve().parent / "orb_pattern_oracle.inc").read_text()
    txt = "\n".join(l for l in txt.splitlines() if not l.startswith("//"))
    return np.array([int(v) for v in re.findall(r"-?\d+", txt)], np.int32).reshape(512, 2)


def sincosf(angle_f32):
    s, c = ctypes.c_float(), ctypes.c_float()
    _libm.sincosf(ctypes.c_float(float(angle_f32)), ctypes.byref(s), ctypes.byref(c))
    return F32(c.value), F32(s.value)


def cv_round(v):
    """cvRound: round half to even (np.rint does that)."""
    return int(np.rint(v))


class Node:
    __slots__ = ("keys", "UL", "UR", "BL", "BR", "no_more", "alive")

    def __init__(self):
        self.keys = []
        self.no_more = False
        self.alive = True


def _divide(n):
    half_x = int(math.ceil(float(F32(n.UR[0] - n.UL[0]) / F32(2))))
    half_y = int(math.ceil(float(F32(n.BR[1] - n.UL[1]) / F32(2))))
    n1, n2, n3, n4 = Node(), Node(), Node(), Node()
    n1.UL = n.UL
    n1.UR = (n.UL[0] + half_x, n.UL[1])
    n1.BL = (n.UL[0], n.UL[1] + half_y)
    n1.BR = (n.UL[0] + half_x, n.UL[1] + half_y)
    n2.UL, n2.UR, n2.BL, n2.BR = n1.UR, n.UR, n1.BR, (n.UR[0], n.UL[1] + half_y)
    n3.UL, n3.UR, n3.BL, n3.BR = n1.BL, n1.BR, n.BL, (n1.BR[0], n.BL[1])
    n4.UL, n4.UR, n4.BL, n4.BR = n3.UR, n2.BR, n3.BR, n.BR
    for kp in n.keys:
        if kp[0] < n1.UR[0]:
            (n1 if kp[1] < n1.BR[1] else n3).keys.append(kp)
        elif kp[1] < n1.BR[1]:
            n2.keys.append(kp)
        else:
            n4.keys.append(kp)
    for c in (n1, n2, n3, n4):
        if len(c.keys) == 1:
            c.no_more = True
    return n1, n2, n3, n4


def _std_sort(v, less):
    """libstdc++ std::sort (introsort: median-of-3 quicksort to depth 2*log2(n), heapsort
    fallback, then insertion sort with threshold 16) — needed because compareNodes leaves ties."""
    n = len(v)
    if n < 2:
        return

    def insertion(lo, hi, guarded):
        for i in range(lo, hi):
            val = v[i]
            if guarded and less(val, v[lo0]):
                v[lo0 + 1:i + 1] = v[lo0:i]
                v[lo0] = val
            else:
                j = i
                while less(val, v[j - 1]):
                    v[j] = v[j - 1]
                    j -= 1
                v[j] = val

    def heapsort(lo, hi):
        import functools
        seg = v[lo:hi]
        # libstdc++ __partial_sort(first,last,last) = make_heap + sort_heap; any correct
        # heap-sort gives a different tie order, so reproduce push/pop exactly.
        m = len(seg)

        def adjust(hole, length, val):
            top = hole
            child = hole
            while child < (length - 1) // 2:
                child = 2 * (child + 1)
                if less(seg[child], seg[child - 1]):
                    child -= 1
                seg[hole] = seg[child]
                hole = child
            if (length & 1) == 0 and child == (length - 2) // 2:
                child = 2 * (child + 1)
                seg[hole] = seg[child - 1]
                hole = child - 1
            parent = (hole - 1) // 2
            while hole > top and less(seg[parent], val):
                seg[hole] = seg[parent]
                hole = parent
                parent = (hole - 1) // 2
            seg[hole] = val

        if m >= 2:
            parent = (m - 2) // 2
            while True:
                adjust(parent, m, seg[parent])
                if parent == 0:
                    break
                parent -= 1
        last = m
        while last > 1:
            last -= 1
            val = seg[last]
            seg[last] = seg[0]
            adjust(0, last, val)
        v[lo:hi] = seg

    def introsort(lo, hi, depth):
        while hi - lo > 16:
            if depth == 0:
                heapsort(lo, hi)
                return
            depth -= 1
            mid = lo + (hi - lo) // 2
            a, b, c = lo + 1, mid, hi - 1
            # __move_median_to_first(lo, a, b, c)
            if less(v[a], v[b]):
                if less(v[b], v[c]):
                    m = b
                elif less(v[a], v[c]):
                    m = c
                else:
                    m = a
            elif less(v[a], v[c]):
                m = a
            elif less(v[b], v[c]):
                m = c
            else:
                m = b
            v[lo], v[m] = v[m], v[lo]
            # __unguarded_partition(lo+1, hi, pivot=lo)
            first, last = lo + 1, hi
            while True:
                while less(v[first], v[lo]):
                    first += 1
                last -= 1
                while less(v[lo], v[last]):
                    last -= 1
                if not first < last:
                    break
                v[first], v[last] = v[last], v[first]
                first += 1
            introsort(first, hi, depth)
            hi = first

    introsort(0, n, 2 * (n.bit_length() - 1))
    lo0 = 0
    if n > 16:
        insertion(1, 16, True)
        for i in range(16, n):  # unguarded
            val = v[i]
            j = i
            while less(val, v[j - 1]):
                v[j] = v[j - 1]
                j -= 1
            v[j] = val
    else:
        insertion(1, n, True)


def distribute_octree(keys, min_x, max_x, min_y, max_y, n_target):
    """keys: list of (x, y, response) floats relative to (min_x, min_y)."""
    # C round(): half away from zero (value is positive)
    n_ini = int(math.floor(float(F32(max_x - min_x) / F32(max_y - min_y)) + 0.5))
    h_x = F32(max_x - min_x) / F32(n_ini)
    nodes = []  # python list used as the std::list (front = index 0)
    ini = []
    for i in range(n_ini):
        n = Node()
        n.UL = (int(h_x * F32(i)), 0)
        n.UR = (int(h_x * F32(i + 1)), 0)
        n.BL = (n.UL[0], max_y - min_y)
        n.BR = (n.UR[0], max_y - min_y)
        nodes.append(n)
        ini.append(n)
    for kp in keys:
        ini[int(F32(kp[0]) / h_x)].keys.append(kp)
    kept = []
    for n in nodes:
        if len(n.keys) == 1:
            n.no_more = True
            kept.append(n)
        elif len(n.keys) > 1:
            kept.append(n)
    nodes = kept

    finish = False
    size_and_node = []
    while not finish:
        prev_size = len(nodes)
        n_to_expand = 0
        size_and_node = []
        front = []
        rest = []
        for n in nodes:
            if n.no_more:
                rest.append(n)
                continue
            for c in _divide(n):
                if c.keys:
                    front.insert(0, c)
                    if len(c.keys) > 1:
                        n_to_expand += 1
                        size_and_node.append((len(c.keys), c))
        nodes = front + rest
        if len(nodes) >= n_target or len(nodes) == prev_size:
            finish = True
        elif len(nodes) + n_to_expand * 3 > n_target:
            while not finish:
                prev_size = len(nodes)
                prev = size_and_node
                size_and_node = []
                _std_sort(prev, lambda a, b: a[0] < b[0] or (a[0] == b[0] and a[1].UL[0] < b[1].UL[0]))
                for j in range(len(prev) - 1, -1, -1):
                    node = prev[j][1]
                    for c in _divide(node):
                        if c.keys:
                            nodes.insert(0, c)
                            if len(c.keys) > 1:
                                size_and_node.append((len(c.keys), c))
                    nodes.remove(node)  # identity-based (Node has no __eq__)
                    if len(nodes) >= n_target:
                        break
                if len(nodes) >= n_target or len(nodes) == prev_size:
                    finish = True
    out = []
    for n in nodes:
        best = n.keys[0]
        for kp in n.keys[1:]:
            if kp[2] > best[2]:
                best = kp
        out.append(best)
    return out


class Cv2Extractor:
    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, fma=True):
        self.nfeatures, self.nlevels, self.ini_th, self.min_th, self.fma = nfeatures, nlevels, ini_th, min_th, fma
        sf64 = float(F32(scale_factor))
        self.sf = [F32(1.0)]
        for i in range(1, nlevels):
            self.sf.append(F32(float(self.sf[-1]) * sf64))
        self.inv = [F32(1.0) / s for s in self.sf]
        factor = F32(1.0 / sf64)
        nd = F32(nfeatures) * (F32(1) - factor) / (F32(1) - F32(float(factor) ** nlevels))
        self.per_level, tot = [], 0
        for _ in range(nlevels - 1):
            self.per_level.append(cv_round(nd))
            tot += self.per_level[-1]
            nd = nd * factor
        self.per_level.append(max(nfeatures - tot, 0))
        um = [0] * (HALF_PATCH + 1)
        vmax = int(math.floor(float(F32(HALF_PATCH) * F32(math.sqrt(2.0)) / F32(2) + F32(1))))
        vmin = int(math.ceil(float(F32(HALF_PATCH) * F32(math.sqrt(2.0)) / F32(2))))
        for v in range(vmax + 1):
            um[v] = cv_round(math.sqrt(HALF_PATCH * HALF_PATCH - v * v))
        v0 = 0
        for v in range(HALF_PATCH, vmin - 1, -1):
            while um[v0] == um[v0 + 1]:
                v0 += 1
            um[v] = v0
            v0 += 1
        self.umax = um
        self.pattern = _pattern()
        self.det_ini = cv2.FastFeatureDetector_create(ini_th, True)
        self.det_min = cv2.FastFeatureDetector_create(min_th, True)

    def pyramid(self, image):
        self.pyr, self.pyr_bordered = [], []
        for level in range(self.nlevels):
            w = cv_round(F32(image.shape[1]) * self.inv[level])
            h = cv_round(F32(image.shape[0]) * self.inv[level])
            if level != 0:
                lvl = cv2.resize(self.pyr[level - 1], (w, h), fx=0, fy=0, interpolation=cv2.INTER_LINEAR)
                flag = cv2.BORDER_REFLECT_101 | cv2.BORDER_ISOLATED
            else:
                lvl = np.ascontiguousarray(image)
                flag = cv2.BORDER_REFLECT_101
            self.pyr.append(lvl)
            self.pyr_bordered.append(cv2.copyMakeBorder(lvl, EDGE, EDGE, EDGE, EDGE, flag))
        return self.pyr

    def ic_angle(self, im, px, py):
        m01 = m10 = 0
        row = im[py].astype(np.int64)
        for u in range(-HALF_PATCH, HALF_PATCH + 1):
            m10 += u * int(row[px + u])
        for v in range(1, HALF_PATCH + 1):
            d = self.umax[v]
            us = np.arange(-d, d + 1)
            plus = im[py + v, px - d:px + d + 1].astype(np.int64)
            minus = im[py - v, px - d:px + d + 1].astype(np.int64)
            m01 += v * int((plus - minus).sum())
            m10 += int((us * (plus + minus)).sum())
        return F32(cv2.fastAtan2(float(F32(m01)), float(F32(m10))))

    def keypoints(self):
        self.candidates, self.levels = [], []
        W = F32(35)
        for level in range(self.nlevels):
            im = self.pyr[level]
            min_bx = min_by = EDGE - 3
            max_bx, max_by = im.shape[1] - EDGE + 3, im.shape[0] - EDGE + 3
            width, height = F32(max_bx - min_bx), F32(max_by - min_by)
            n_cols, n_rows = int(width / W), int(height / W)
            w_cell = int(math.ceil(float(width / F32(n_cols))))
            h_cell = int(math.ceil(float(height / F32(n_rows))))
            keys = []
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
                    roi = im[ini_y:max_y, ini_x:max_x]
                    cell = self.det_ini.detect(roi)
                    if not cell:
                        cell = self.det_min.detect(roi)
                    for kp in cell:
                        keys.append((F32(kp.pt[0]) + F32(j * w_cell), F32(kp.pt[1]) + F32(i * h_cell), F32(kp.response)))
            self.candidates.append(keys)
            sel = distribute_octree(keys, min_bx, max_bx, min_by, max_by, self.per_level[level])
            size = F32(int(F32(31) * self.sf[level]))
            out = []
            for (x, y, r) in sel:
                out.append([F32(x + F32(min_bx)), F32(y + F32(min_by)), size, F32(-1), r, level])
            self.levels.append(out)
        for level in range(self.nlevels):
            for kp in self.levels[level]:
                kp[3] = self.ic_angle(self.pyr[level], cv_round(kp[0]), cv_round(kp[1]))

    def descriptor(self, kp, img):
        factor_pi = F32(math.pi / float(F32(180.0)))
        angle = F32(kp[3]) * factor_pi
        a, b = sincosf(angle)
        cx, cy = cv_round(kp[0]), cv_round(kp[1])
        px = self.pattern[:, 0].astype(np.float32)
        py = self.pattern[:, 1].astype(np.float32)
        if self.fma:
            ya = (py * a).astype(np.float32).astype(np.float64)
            yb = (py * b).astype(np.float32).astype(np.float64)
            fy = (px.astype(np.float64) * float(b) + ya).astype(np.float32)
            fx = (px.astype(np.float64) * float(a) - yb).astype(np.float32)
        else:
            fy = (px * b).astype(np.float32) + (py * a).astype(np.float32)
            fx = (px * a).astype(np.float32) - (py * b).astype(np.float32)
        iy = np.rint(fy).astype(np.int64)
        ix = np.rint(fx).astype(np.int64)
        vals = img[cy + iy, cx + ix].astype(np.int32)
        bits = (vals[0::2] < vals[1::2]).astype(np.uint8).reshape(32, 8)
        return (bits << np.arange(8, dtype=np.uint8)).sum(axis=1).astype(np.uint8)

    def extract(self, image, lap=(0, 0)):
        from . import KP_DTYPE
        self.pyramid(image)
        self.keypoints()
        n = sum(len(l) for l in self.levels)
        kps = np.zeros(n, KP_DTYPE)
        desc = np.zeros((n, 32), np.uint8)
        mono, stereo = 0, n - 1
        for level in range(self.nlevels):
            if not self.levels[level]:
                continue
            working = self.pyr[level].copy()
            working = cv2.GaussianBlur(working, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
            scale = self.sf[level]
            for kp in self.levels[level]:
                d = self.descriptor(kp, working)
                x, y = kp[0], kp[1]
                if level != 0:
                    x, y = F32(x * scale), F32(y * scale)
                if lap[0] <= x <= lap[1]:
                    dst = stereo
                    stereo -= 1
                else:
                    dst = mono
                    mono += 1
                kps[dst] = (x, y, kp[2], kp[3], kp[4], level, -1)
                desc[dst] = d
        return kps, desc, mono
