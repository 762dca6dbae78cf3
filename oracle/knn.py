"""Oracle, matching half: kNN(k=2) under L2, Lowe ratio test, mutual-NN.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

Reference call sites restated here (paths relative to /root/reference):
* ``utils/find_matches.py:144``  ``cv2.BFMatcher(norm, crossCheck=False)``
* ``utils/find_matches.py:147``  ``bf.knnMatch(desc1, desc2, k=2)``
* ``utils/find_matches.py:150-153`` the ratio loop ``m.distance < 0.75 * n.distance``

The arithmetic behind ``knnMatch`` lives in opencv-python (pinned ==4.11.0 in
``requirements.txt:2``, not vendored).  Its published semantics (SURVEY.md 3.2,
probed against cv2 4.13): per query row the two nearest train rows, ascending
distance, LOWEST train index wins ties; ``DMatch.distance`` is
``float32(sqrt(float32(d2)))`` where d2 is the exact integer squared distance
(exact in float32 because d2 <= 128*255^2 < 2^24).
"""
from __future__ import annotations

import numpy as np

RATIO = 0.75  # utils/find_matches.py:152


# --------------------------------------------------------------------------
# exact integer restatement
# --------------------------------------------------------------------------
def sqdist_matrix(d1: np.ndarray, d2: np.ndarray) -> np.ndarray:
    """Exact integer ||a-b||^2 for integer-valued descriptors (any dtype).

    float32 GEMM is exact here: every partial sum is an integer < 2^24.
    """
    a = np.ascontiguousarray(d1, dtype=np.float32)
    b = np.ascontiguousarray(d2, dtype=np.float32)
    dot = (a @ b.T).astype(np.int64)
    na = (a.astype(np.int64) ** 2).sum(1)
    nb = (b.astype(np.int64) ** 2).sum(1)
    return na[:, None] + nb[None, :] - 2 * dot


def knn2_l2_numpy(d1: np.ndarray, d2: np.ndarray):
    """(idx[Nq,2] int64, sqd[Nq,2] int64): two nearest train rows per query row.

    Tie rule: lowest train index first, for both the best and the second best
    (== ``np.argsort(d2, kind='stable')[:, :2]``).  Requires Nt >= 2.
    """
    D = sqdist_matrix(d1, d2)
    nq, nt = D.shape
    if nt < 2:
        raise ValueError("knn2 needs at least two train descriptors")
    rows = np.arange(nq)
    i1 = D.argmin(1)                      # first occurrence == lowest index
    v1 = D[rows, i1]
    D[rows, i1] = np.iinfo(np.int64).max
    i2 = D.argmin(1)
    v2 = D[rows, i2]
    return np.stack([i1, i2], 1).astype(np.int64), np.stack([v1, v2], 1)


def distance_f32(sqd: np.ndarray) -> np.ndarray:
    """``DMatch.distance`` as cv2 reports it: correctly rounded f32 sqrt of the
    exact integer squared distance."""
    return np.sqrt(np.asarray(sqd).astype(np.float32)).astype(np.float32)


def ratio_test_arrays(idx: np.ndarray, sqd: np.ndarray, ratio: float = RATIO):
    """The literal ratio loop of utils/find_matches.py:150-153 on arrays.

    Python evaluates ``m.distance < 0.75 * n.distance`` in doubles on two
    float32 values; that is reproduced bit for bit here.
    Returns (queryIdx int64[M], trainIdx int64[M], distance float32[M]).
    """
    dist = distance_f32(sqd)
    keep = dist[:, 0].astype(np.float64) < ratio * dist[:, 1].astype(np.float64)
    q = np.nonzero(keep)[0].astype(np.int64)
    return q, idx[q, 0].astype(np.int64), dist[q, 0]


def mutual_nn_numpy(d1: np.ndarray, d2: np.ndarray) -> np.ndarray:
    """nn21[Nt]: for every train row the nearest query row (lowest index ties).

    Mutual-NN keeps (q, t) iff nn21[t] == q.  Not in the reference code or data
    (README.md:108 only); equivalent to ``BFMatcher(NORM_L2, crossCheck=True)``.
    """
    return sqdist_matrix(d1, d2).argmin(0).astype(np.int64)


def match_features(d1, d2, ratio: float = RATIO, mutual: bool = False):
    """NumPy restatement of ``ImageMatcher.match_features`` (+ optional mutual).

    Returns (queryIdx, trainIdx, distance_f32) in ascending queryIdx order.
    """
    idx, sqd = knn2_l2_numpy(d1, d2)
    q, t, d = ratio_test_arrays(idx, sqd, ratio)
    if mutual:
        nn21 = mutual_nn_numpy(d1, d2)
        ok = nn21[t] == q
        q, t, d = q[ok], t[ok], d[ok]
    return q, t, d


# --------------------------------------------------------------------------
# the reference's own OpenCV calls
# --------------------------------------------------------------------------
def knn2_l2_cv2(d1: np.ndarray, d2: np.ndarray):
    """utils/find_matches.py:144-147 with NORM_L2.  Returns (idx, dist_f32)."""
    import cv2
    bf = cv2.BFMatcher(cv2.NORM_L2, crossCheck=False)
    a = np.ascontiguousarray(d1, dtype=np.float32)
    b = np.ascontiguousarray(d2, dtype=np.float32)
    matches = bf.knnMatch(a, b, k=2)
    idx = np.array([[m.trainIdx, n.trainIdx] for m, n in matches], dtype=np.int64)
    dist = np.array([[m.distance, n.distance] for m, n in matches], dtype=np.float32)
    return idx.reshape(-1, 2), dist.reshape(-1, 2)


def ratio_test(matches, ratio: float = RATIO):
    """The literal loop, on cv2 DMatch pairs (utils/find_matches.py:150-153)."""
    good = []
    for m, n in matches:
        if m.distance < ratio * n.distance:
            good.append(m)
    return good


def match_features_cv2(d1, d2, ratio: float = RATIO):
    """``ImageMatcher.match_features`` with NORM_L2, returned as arrays."""
    import cv2
    bf = cv2.BFMatcher(cv2.NORM_L2, crossCheck=False)
    a = np.ascontiguousarray(d1, dtype=np.float32)
    b = np.ascontiguousarray(d2, dtype=np.float32)
    good = ratio_test(bf.knnMatch(a, b, k=2), ratio)
    q = np.array([m.queryIdx for m in good], dtype=np.int64)
    t = np.array([m.trainIdx for m in good], dtype=np.int64)
    d = np.array([m.distance for m in good], dtype=np.float32)
    return q, t, d


def mutual_nn_cv2(d1, d2):
    """(q, t) pairs kept by ``BFMatcher(NORM_L2, crossCheck=True).match``."""
    import cv2
    bf = cv2.BFMatcher(cv2.NORM_L2, crossCheck=True)
    a = np.ascontiguousarray(d1, dtype=np.float32)
    b = np.ascontiguousarray(d2, dtype=np.float32)
    ms = bf.match(a, b)
    q = np.array([m.queryIdx for m in ms], dtype=np.int64)
    t = np.array([m.trainIdx for m in ms], dtype=np.int64)
    o = np.argsort(q, kind="stable")
    return q[o], t[o]
