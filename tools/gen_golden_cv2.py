#!/usr/bin/env python3
"""Golden vectors from a second, independent composition of the reference algorithm over REAL OpenCV.

The reference (/root/reference/src/ORBextractor.cc, src/Frame.cc:889-1078) cannot be compiled in the build
container (needs OpenCV C++), and ships no golden vectors (test/test_Stereo.cpp:230-245 has no assertions).
This script follows the reference's control flow in Python and calls the genuine OpenCV primitives through
cv2 (4.13.0) wherever the reference calls OpenCV: cv2.resize, cv2.copyMakeBorder, cv2.FastFeatureDetector,
cv2.GaussianBlur, cv2.fastAtan2, cv2.norm.  Its outputs are committed under tests/golden/ and pin the C++
oracle (tests/test_oracle_golden.py) -- and through it the CUDA path.

Frozen choices (documented in DESIGN.md): quadtree ties on node size are broken by creation order, newest
first (the reference breaks them by heap address, src/ORBextractor.cc:591,684); `mb` is passed explicitly
(the reference reads it uninitialised, src/Frame.cc:920 vs :322).

usage: python tools/gen_golden_cv2.py   (writes tests/golden/*.npz; needs tests/data/*.pgm)
"""
import math
import sys
from pathlib import Path

import cv2
import numpy as np

ROOT = Path(__file__).resolve().parent.parent
cv2.setNumThreads(1)
f32 = np.float32
EDGE, HALF, PATCH = 19, 15, 31


def read_pgm(p):
    d = Path(p).read_bytes()
    parts = d.split(b"\n", 3)
    w, h = map(int, parts[1].split())
    return np.frombuffer(parts[3], np.uint8).reshape(h, w).copy()


def cv_round(v):  # round half to even on a float32/float64 value
    return int(np.rint(v))


def load_pattern():
    txt = (ROOT / "oracle/orb_pattern_31.inc").read_text()
    vals = [int(t) for line in txt.splitlines() if not line.startswith("//") for t in line.split(",") if t.strip()]
    return np.array(vals, np.int32).reshape(512, 2)


class RefExtractor:
    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini=20, mn=7):
        self.nf, self.nl, self.ini, self.mn = nfeatures, nlevels, ini, mn
        sf = float(f32(scale_factor))  # ctor arg is float, member is double
        self.scale = [f32(1.0)]
        for i in range(1, nlevels):
            self.scale.append(f32(float(self.scale[-1]) * sf))
        self.inv = [f32(1.0) / s for s in self.scale]
        factor = f32(1.0 / sf)
        des = f32(nfeatures) * (f32(1) - factor) / (f32(1) - f32(math.pow(float(factor), float(nlevels))))
        self.quota, tot = [], 0
        for _ in range(nlevels - 1):
            q = cv_round(des)
            self.quota.append(q)
            tot += q
            des = f32(des * factor)
        self.quota.append(max(nfeatures - tot, 0))
        self.pattern = load_pattern()
        umax = [0] * (HALF + 1)
        vmax = int(math.floor(f32(HALF) * f32(math.sqrt(2.0)) / f32(2) + f32(1)))
        vmin = int(math.ceil(f32(HALF) * f32(math.sqrt(2.0)) / f32(2)))
        for v in range(vmax + 1):
            umax[v] = cv_round(math.sqrt(HALF * HALF - v * v))
        v0 = 0
        for v in range(HALF, vmin - 1, -1):
            while umax[v0] == umax[v0 + 1]:
                v0 += 1
            umax[v] = v0
            v0 += 1
        self.umax = umax
        self.fast_ini = cv2.FastFeatureDetector_create(threshold=ini, nonmaxSuppression=True,
                                                       type=cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
        self.fast_min = cv2.FastFeatureDetector_create(threshold=mn, nonmaxSuppression=True,
                                                       type=cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)

    def compute_pyramid(self, image):
        self.pyr_padded, self.pyr = [], []
        for l in range(self.nl):
            w = cv_round(f32(image.shape[1]) * self.inv[l])
            h = cv_round(f32(image.shape[0]) * self.inv[l])
            if l == 0:
                padded = cv2.copyMakeBorder(image, EDGE, EDGE, EDGE, EDGE, cv2.BORDER_REFLECT_101)
            else:
                lvl = cv2.resize(self.pyr[l - 1], (w, h), interpolation=cv2.INTER_LINEAR)
                padded = cv2.copyMakeBorder(lvl, EDGE, EDGE, EDGE, EDGE, cv2.BORDER_REFLECT_101)
            self.pyr_padded.append(padded)
            self.pyr.append(padded[EDGE:EDGE + h, EDGE:EDGE + w])

    def detect_level(self, l):
        im = self.pyr[l]
        minB, maxBX, maxBY = EDGE - 3, im.shape[1] - EDGE + 3, im.shape[0] - EDGE + 3
        width, height = f32(maxBX - minB), f32(maxBY - minB)
        nC, nR = int(width / f32(30)), int(height / f32(30))
        if nC <= 0 or nR <= 0:
            return []
        wC, hC = int(math.ceil(width / f32(nC))), int(math.ceil(height / f32(nR)))
        out = []
        for i in range(nR):
            iniY = minB + i * hC
            maxY = iniY + hC + 6
            if iniY >= maxBY - 3:
                continue
            maxY = min(maxY, maxBY)
            for j in range(nC):
                iniX = minB + j * wC
                maxX = iniX + wC + 6
                if iniX >= maxBX - 6:
                    continue
                maxX = min(maxX, maxBX)
                cell = np.ascontiguousarray(im[iniY:maxY, iniX:maxX])
                kps = self.fast_ini.detect(cell, None)
                if len(kps) == 0:
                    kps = self.fast_min.detect(cell, None)
                for k in kps:
                    out.append((int(k.pt[0]) + j * wC, int(k.pt[1]) + i * hC, int(k.response)))
        return out

    # --- quadtree: literal list semantics, nodes are dicts, list = python list (front = index 0)
    @staticmethod
    def _divide(n, c):
        halfX = int(math.ceil(f32(n["UR"][0] - n["UL"][0]) / f32(2)))
        halfY = int(math.ceil(f32(n["BR"][1] - n["UL"][1]) / f32(2)))
        UL, UR, BL, BR = n["UL"], n["UR"], n["BL"], n["BR"]
        n1 = dict(UL=UL, UR=(UL[0] + halfX, UL[1]), BL=(UL[0], UL[1] + halfY), BR=(UL[0] + halfX, UL[1] + halfY), pts=[])
        n2 = dict(UL=n1["UR"], UR=UR, BL=n1["BR"], BR=(UR[0], UL[1] + halfY), pts=[])
        n3 = dict(UL=n1["BL"], UR=n1["BR"], BL=BL, BR=(n1["BR"][0], BL[1]), pts=[])
        n4 = dict(UL=n3["UR"], UR=n2["BR"], BL=n3["BR"], BR=BR, pts=[])
        for idx in n["pts"]:
            x, y = c[idx][0], c[idx][1]
            if x < n1["UR"][0]:
                (n1 if y < n1["BR"][1] else n3)["pts"].append(idx)
            elif y < n1["BR"][1]:
                n2["pts"].append(idx)
            else:
                n4["pts"].append(idx)
        kids = [n1, n2, n3, n4]
        for k in kids:
            k["nomore"] = len(k["pts"]) == 1
        return kids

    def distribute(self, c, minX, maxX, minY, maxY, N):
        ratio = f32(maxX - minX) / f32(maxY - minY)
        nIni = int(math.floor(ratio + f32(0.5)))  # round() half away from zero, positive argument
        if nIni <= 0 or not c:
            return []
        hX = f32(maxX - minX) / f32(nIni)
        seq = [0]
        nodes = []
        for i in range(nIni):
            nodes.append(dict(UL=(int(hX * f32(i)), 0), UR=(int(hX * f32(i + 1)), 0),
                              BL=(int(hX * f32(i)), maxY - minY), BR=(int(hX * f32(i + 1)), maxY - minY),
                              pts=[], nomore=False, seq=seq[0]))
            seq[0] += 1
        ini = list(nodes)
        for i, p in enumerate(c):
            ini[int(f32(p[0]) / hX)]["pts"].append(i)
        kept = []
        for n in nodes:
            if len(n["pts"]) == 1:
                n["nomore"] = True
                kept.append(n)
            elif len(n["pts"]) > 1:
                kept.append(n)
        nodes = kept

        def push_children(kids, expandable):
            cnt = 0
            for k in kids:
                if not k["pts"]:
                    continue
                k["seq"] = seq[0]
                seq[0] += 1
                nodes.insert(0, k)
                if len(k["pts"]) > 1:
                    cnt += 1
                    expandable.append(k)
            return cnt

        finish = False
        while not finish:
            prev = len(nodes)
            n_to_expand = 0
            expandable = []
            for n in list(nodes):  # nodes existing at the start of the pass, in list order
                if n["nomore"]:
                    continue
                n_to_expand += push_children(self._divide(n, c), expandable)
                nodes.remove(n)
            if len(nodes) >= N or len(nodes) == prev:
                finish = True
            elif len(nodes) + n_to_expand * 3 > N:
                while not finish:
                    prev = len(nodes)
                    prev_exp = sorted(expandable, key=lambda k: (len(k["pts"]), k["seq"]))
                    expandable = []
                    for n in reversed(prev_exp):
                        push_children(self._divide(n, c), expandable)
                        nodes.remove(n)
                        if len(nodes) >= N:
                            break
                    if len(nodes) >= N or len(nodes) == prev:
                        finish = True
        res = []
        for n in nodes:
            best = n["pts"][0]
            for k in n["pts"][1:]:
                if c[k][2] > c[best][2]:
                    best = k
            res.append(best)
        return res

    def ic_angle(self, l, x, y):
        im = self.pyr_padded[l].astype(np.int64)
        cy, cx = y + EDGE, x + EDGE
        m01 = m10 = 0
        for u in range(-HALF, HALF + 1):
            m10 += u * int(im[cy, cx + u])
        for v in range(1, HALF + 1):
            d = self.umax[v]
            vs = 0
            for u in range(-d, d + 1):
                p, m = int(im[cy + v, cx + u]), int(im[cy - v, cx + u])
                vs += p - m
                m10 += u * (p + m)
            m01 += v * vs
        return f32(cv2.fastAtan2(float(f32(m01)), float(f32(m10))))

    def descriptor(self, blurred, x, y, angle):
        factorPI = f32(math.pi / 180.0)
        ang = f32(angle * factorPI)
        a, b = f32(np.cos(ang, dtype=f32)), f32(np.sin(ang, dtype=f32))
        px, py = self.pattern[:, 0].astype(f32), self.pattern[:, 1].astype(f32)
        rr = np.rint(px * b + py * a).astype(np.int64)
        cc = np.rint(px * a - py * b).astype(np.int64)
        vals = blurred[y + rr, x + cc].astype(np.int32)
        bits = (vals[0::2] < vals[1::2]).astype(np.uint8).reshape(32, 8)
        return (bits << np.arange(8, dtype=np.uint8)).sum(axis=1).astype(np.uint8)

    def extract(self, image):
        self.compute_pyramid(image)
        allk, cands = [], []
        for l in range(self.nl):
            im = self.pyr[l]
            c = self.detect_level(l)
            cands.append(np.array(c, np.int32).reshape(-1, 3))
            keep = self.distribute(c, 16, im.shape[1] - 16, 16, im.shape[0] - 16, self.quota[l])
            size = f32(int(f32(PATCH) * self.scale[l]))
            ks = []
            for idx in keep:
                x, y, r = c[idx]
                ks.append([f32(x + 16), f32(y + 16), size, f32(-1), f32(r), l])
            allk.append(ks)
        for l in range(self.nl):
            for k in allk[l]:
                k[3] = self.ic_angle(l, int(k[0]), int(k[1]))
        kps, desc = [], []
        for l in range(self.nl):
            if not allk[l]:
                continue
            work = self.pyr[l].copy()
            blurred = cv2.GaussianBlur(work, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
            for k in allk[l]:
                desc.append(self.descriptor(blurred, int(k[0]), int(k[1]), k[3]))
                if l != 0:
                    k[0] = f32(k[0] * self.scale[l])
                    k[1] = f32(k[1] * self.scale[l])
                kps.append(k)
        K = np.zeros(len(kps), dtype=[("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                                      ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
        for i, k in enumerate(kps):
            K[i] = (k[0], k[1], k[2], k[3], k[4], k[5], -1)
        D = np.array(desc, np.uint8).reshape(-1, 32)
        return K, D, cands


def hamming(a, b):
    return int(np.unpackbits(np.bitwise_xor(a, b)).sum())


def stereo_classic(exL, exR, KL, DL, KR, DR, bf, b):
    """src/Frame.cc:889-1078 with cv2.norm for the SAD."""
    N = len(KL)
    uR_out = np.full(N, -1, f32)
    depth = np.full(N, -1, f32)
    best_d = np.full(N, 100, np.int32)
    best_i = np.full(N, -1, np.int32)
    nRows = exL.pyr[0].shape[0]
    rows = [[] for _ in range(nRows)]
    for iR in range(len(KR)):
        y = f32(KR["y"][iR])
        r = f32(2.0) * exL.scale[KR["octave"][iR]]
        for yi in range(int(math.floor(f32(y - r))), int(math.ceil(f32(y + r))) + 1):
            rows[yi].append(iR)
    minD, maxD = f32(0), f32(bf) / f32(b)
    dist_idx = []
    for iL in range(N):
        lvl = int(KL["octave"][iL]); vL = f32(KL["y"][iL]); uL = f32(KL["x"][iL])
        cand = rows[int(vL)]
        if not cand:
            continue
        minU, maxU = f32(uL - maxD), f32(uL - minD)
        if maxU < 0:
            continue
        bd, bi = 100, 0
        for iR in cand:
            o = int(KR["octave"][iR])
            if o < lvl - 1 or o > lvl + 1:
                continue
            u = f32(KR["x"][iR])
            if minU <= u <= maxU:
                d = hamming(DL[iL], DR[iR])
                if d < bd:
                    bd, bi = d, iR
        if bd < 100:
            best_d[iL], best_i[iL] = bd, bi
        if bd >= 75:
            continue
        sf = exL.inv[lvl]
        # C round(): half away from zero; arguments are positive float32 products
        su = int(math.floor(float(f32(uL * sf)) + 0.5))
        sv = int(math.floor(float(f32(vL * sf)) + 0.5))
        sr = int(math.floor(float(f32(f32(KR["x"][bi]) * sf)) + 0.5))
        w = L = 5
        imL, imR = exL.pyr[lvl], exR.pyr[lvl]
        IL = imL[sv - w:sv + w + 1, su - w:su + w + 1].astype(f32)
        IL = IL - IL[w, w]
        if sr + L - w < 0 or sr + L + w + 1 >= imR.shape[1]:
            continue
        best, binc, dists = 2 ** 31 - 1, 0, [f32(0)] * (2 * L + 1)
        for inc in range(-L, L + 1):
            IR = imR[sv - w:sv + w + 1, sr + inc - w:sr + inc + w + 1].astype(f32)
            IR = IR - IR[w, w]
            dist = f32(cv2.norm(IL, IR, cv2.NORM_L1))
            if dist < best:
                best, binc = int(dist), inc
            dists[L + inc] = dist
        if binc in (-L, L):
            continue
        d1, d2, d3 = dists[L + binc - 1], dists[L + binc], dists[L + binc + 1]
        with np.errstate(all="ignore"):
            delta = f32(f32(d1 - d3) / f32(f32(2.0) * f32(f32(d1 + d3) - f32(f32(2.0) * d2))))
        if delta < -1 or delta > 1:
            continue
        bu = f32(exL.scale[lvl] * f32(f32(f32(sr) + f32(binc)) + delta))
        disp = f32(uL - bu)
        if minD <= disp < maxD:
            if disp <= 0:
                disp = f32(0.01)
                bu = f32(float(uL) - 0.01)
            depth[iL] = f32(f32(bf) / disp)
            uR_out[iL] = bu
            dist_idx.append((best, iL))
    if dist_idx:
        dist_idx.sort()
        median = f32(dist_idx[len(dist_idx) // 2][0])
        th = f32(f32(f32(1.5) * f32(1.4)) * median)
        for d, i in reversed(dist_idx):
            if d < th:
                break
            uR_out[i] = -1
            depth[i] = -1
    return uR_out, depth, best_d, best_i


def main():
    out = ROOT / "tests/golden"
    out.mkdir(parents=True, exist_ok=True)
    bf = 47.90639384423901
    b = bf / 435.2046959714599
    res = {}
    for name in ("EuRoC_l", "EuRoC_r", "TUM_l", "TUM_r"):
        im = read_pgm(ROOT / "tests/data" / f"{name}.pgm")
        ex = RefExtractor()
        K, D, cands = ex.extract(im)
        res[name] = (ex, K, D)
        lvl_px = {f"pyr{l}": ex.pyr[l] for l in (1, 4, 7)}  # a few levels; full pyramid is large
        np.savez_compressed(out / f"extract_{name}.npz", kps=K, desc=D,
                            cand_counts=np.array([len(c) for c in cands], np.int32),
                            cand0=cands[0], cand3=cands[3], cand7=cands[7], **lvl_px)
        print(name, len(K), [len(c) for c in cands])
    for pair in ("EuRoC", "TUM"):
        exL, KL, DL = res[f"{pair}_l"]
        exR, KR, DR = res[f"{pair}_r"]
        u, z, bd, bi = stereo_classic(exL, exR, KL, DL, KR, DR, bf, b)
        np.savez_compressed(out / f"stereo_{pair}.npz", uRight=u, depth=z, best_dist=bd, best_idx=bi,
                            bf=f32(bf), b=f32(b))
        print(pair, "stereo matches", int((z > 0).sum()), "hamming<100", int((bi >= 0).sum()))


if __name__ == "__main__":
    sys.exit(main())
