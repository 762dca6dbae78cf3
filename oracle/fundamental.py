"""Oracle, geometry half: RANSAC fundamental matrix + geometric verification.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

Reference call sites restated here (paths relative to /root/reference):
* ``utils/find_matches.py:282``  ``cv2.findFundamentalMat(pts1, pts2, cv2.FM_RANSAC, 3.0)``
* ``utils/find_matches.py:157-201`` ``ImageMatcher.geometric_verification``
* ``utils/find_matches.py:203-214`` ``ImageMatcher.verify_match_quality``

``findFundamentalMat`` / ``computeCorrespondEpilines`` live in opencv-python
(pinned ==4.11.0, ``requirements.txt:2``; not vendored).  The restatement of
``computeCorrespondEpilines`` follows its published behaviour: the line
``l = F x`` (or ``F^T x``) is formed in double precision, scaled so that
a^2+b^2 = 1, then stored as float32 when the points are float32.
"""
from __future__ import annotations

import numpy as np

RANSAC_THRESH = 3.0    # utils/find_matches.py:282
VERIFY_THRESH = 3.0    # utils/find_matches.py:157
MIN_SPREAD = 20        # utils/find_matches.py:185
MIN_INLIERS = 15       # utils/find_matches.py:203
MIN_RATIO = 0.3
MAX_ERROR = 2.0
MIN_MATCHES = 5        # utils/find_matches.py:274


def find_fundamental_cv2(pts1, pts2, thresh: float = RANSAC_THRESH):
    """The reference's call, verbatim (utils/find_matches.py:282)."""
    import cv2
    F, mask = cv2.findFundamentalMat(pts1, pts2, cv2.FM_RANSAC, thresh)
    return F, mask


def _epilines(pts: np.ndarray, which: int, F: np.ndarray) -> np.ndarray:
    """Restatement of cv2.computeCorrespondEpilines for float32 points.

    which == 1: points are in image 1, lines (in image 2) are F x.
    which == 2: points are in image 2, lines (in image 1) are F^T x.
    """
    Fd = np.asarray(F, dtype=np.float64)
    if which == 2:
        Fd = Fd.T
    x = pts[:, 0].astype(np.float64)
    y = pts[:, 1].astype(np.float64)
    a = Fd[0, 0] * x + Fd[0, 1] * y + Fd[0, 2]
    b = Fd[1, 0] * x + Fd[1, 1] * y + Fd[1, 2]
    c = Fd[2, 0] * x + Fd[2, 1] * y + Fd[2, 2]
    nu = a * a + b * b
    with np.errstate(divide="ignore", invalid="ignore"):
        nu = np.where(nu != 0, 1.0 / np.sqrt(nu), 1.0)
    return np.stack([a * nu, b * nu, c * nu], 1).astype(np.float32)


def _verification_from_lines(pts1, pts2, lines1, lines2, threshold):
    """utils/find_matches.py:165-201, literally (NumPy, float32 inputs)."""
    errors1 = np.abs(np.sum(np.multiply(pts1, lines1[:, :2]), axis=1) + lines1[:, 2]) / \
        np.sqrt(np.sum(np.square(lines1[:, :2]), axis=1))
    errors2 = np.abs(np.sum(np.multiply(pts2, lines2[:, :2]), axis=1) + lines2[:, 2]) / \
        np.sqrt(np.sum(np.square(lines2[:, :2]), axis=1))
    symmetric_errors = (errors1 + errors2) / 2
    inlier_mask = symmetric_errors < threshold
    reproj_error = np.mean(symmetric_errors[inlier_mask]) if np.any(inlier_mask) else float("inf")
    if np.any(inlier_mask):
        pts1_std = np.std(pts1[inlier_mask], axis=0)
        pts2_std = np.std(pts2[inlier_mask], axis=0)
        well_distributed = bool(np.all(pts1_std > MIN_SPREAD) and np.all(pts2_std > MIN_SPREAD))
    else:
        well_distributed = False
    return {
        "metrics": {
            "total_matches": len(pts1),
            "inliers": np.sum(inlier_mask),
            "inlier_ratio": np.mean(inlier_mask),
            "reprojection_error": reproj_error,
            "symmetric_error": np.mean(symmetric_errors),
            "well_distributed": well_distributed,
        },
        "inlier_mask": inlier_mask,
        "symmetric_errors": symmetric_errors,
    }


def geometric_verification(pts1, pts2, F, threshold: float = VERIFY_THRESH):
    """NumPy restatement of ``ImageMatcher.geometric_verification``."""
    pts1 = np.asarray(pts1, dtype=np.float32).reshape(-1, 2)
    pts2 = np.asarray(pts2, dtype=np.float32).reshape(-1, 2)
    lines1 = _epilines(pts2, 2, F)   # :160
    lines2 = _epilines(pts1, 1, F)   # :162
    return _verification_from_lines(pts1, pts2, lines1, lines2, threshold)


def geometric_verification_cv2(pts1, pts2, F, threshold: float = VERIFY_THRESH):
    """The reference's body with its own cv2 calls (utils/find_matches.py:160-163)."""
    import cv2
    pts1 = np.asarray(pts1, dtype=np.float32).reshape(-1, 2)
    pts2 = np.asarray(pts2, dtype=np.float32).reshape(-1, 2)
    lines1 = cv2.computeCorrespondEpilines(pts2.reshape(-1, 1, 2), 2, F).reshape(-1, 3)
    lines2 = cv2.computeCorrespondEpilines(pts1.reshape(-1, 1, 2), 1, F).reshape(-1, 3)
    return _verification_from_lines(pts1, pts2, lines1, lines2, threshold)


def verify_match_quality(res, min_inliers=MIN_INLIERS, min_ratio=MIN_RATIO, max_error=MAX_ERROR):
    """utils/find_matches.py:203-214."""
    m = res["metrics"]
    return bool(all([
        m["inliers"] >= min_inliers,
        m["inlier_ratio"] >= min_ratio,
        m["reprojection_error"] <= max_error,
        m["well_distributed"],
    ]))


# --------------------------------------------------------------------------
# error measures used by the RANSAC parity checks (float64)
# --------------------------------------------------------------------------
def epipolar_sq(pts1, pts2, F):
    """(d1^2, d2^2): squared point-to-epiline distances in image 2 and image 1.

    cv2's own RANSAC mask is ``max(d1^2, d2^2) < thresh^2``.
    """
    F = np.asarray(F, dtype=np.float64)
    x1 = np.concatenate([np.asarray(pts1, np.float64), np.ones((len(pts1), 1))], 1)
    x2 = np.concatenate([np.asarray(pts2, np.float64), np.ones((len(pts2), 1))], 1)
    l2 = x1 @ F.T          # lines in image 2
    l1 = x2 @ F            # lines in image 1
    r = np.sum(x2 * l2, 1)
    d2 = r * r / (l2[:, 0] ** 2 + l2[:, 1] ** 2)
    d1 = r * r / (l1[:, 0] ** 2 + l1[:, 1] ** 2)
    return d2, d1


def sampson_sq(pts1, pts2, F):
    """Sampson (first-order geometric) squared error."""
    F = np.asarray(F, dtype=np.float64)
    x1 = np.concatenate([np.asarray(pts1, np.float64), np.ones((len(pts1), 1))], 1)
    x2 = np.concatenate([np.asarray(pts2, np.float64), np.ones((len(pts2), 1))], 1)
    l2 = x1 @ F.T
    l1 = x2 @ F
    r = np.sum(x2 * l2, 1)
    return r * r / (l2[:, 0] ** 2 + l2[:, 1] ** 2 + l1[:, 0] ** 2 + l1[:, 1] ** 2)
