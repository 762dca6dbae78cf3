"""Oracle, control flow: ``ImageMatcher.process_image_pair`` from descriptors on.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

Follows utils/find_matches.py:246-310 from the point where keypoints and
descriptors exist (the north-star path starts at descriptors; feature
extraction is hoisted to once per image).  Status codes name
the reference's four skip gates.
"""
from __future__ import annotations

import numpy as np

from . import fundamental as fm
from . import knn

OK, TOO_FEW_FEATURES, TOO_FEW_MATCHES, NO_MODEL, QUALITY_FAIL = 0, 1, 2, 3, 4


def process_pair(desc1, kp1, desc2, kp2, mutual: bool = False, use_cv2: bool = True):
    """One pair through kNN+ratio(+mutual) -> RANSAC F -> verification.

    desc*: [n,128] integer-valued; kp*: float32 [n,2].  Returns a dict with
    ``status`` and, past the match gate, the arrays the reference saves
    (utils/find_matches.py:312-327).
    """
    out = {"status": OK}
    n1, n2 = len(desc1), len(desc2)
    if n1 < 1 or n2 < 1:                      # :268
        out["status"] = TOO_FEW_FEATURES
        return out
    if n2 < 2:
        # knnMatch rows have length 1 -> `for m, n in matches` raises (:151),
        # caught per pair at :348 -> the pair is skipped.
        out["status"] = TOO_FEW_FEATURES
        return out
    if use_cv2 and not mutual:
        q, t, d = knn.match_features_cv2(desc1, desc2)
    else:
        q, t, d = knn.match_features(desc1, desc2, mutual=mutual)
    out.update(queryIdx=q, trainIdx=t, distance=d)
    if len(q) < fm.MIN_MATCHES:               # :274
        out["status"] = TOO_FEW_MATCHES
        return out
    pts1 = np.float32(np.asarray(kp1)[q]).reshape(-1, 2)   # :278
    pts2 = np.float32(np.asarray(kp2)[t]).reshape(-1, 2)   # :279
    out.update(pts1=pts1, pts2=pts2)
    F, _ = fm.find_fundamental_cv2(pts1, pts2)             # :282
    if F is None or F.shape != (3, 3):                     # :284
        out["status"] = NO_MODEL
        return out
    res = fm.geometric_verification_cv2(pts1, pts2, F)     # :288
    out.update(F=F, inlier_mask=res["inlier_mask"], metrics=res["metrics"],
               symmetric_errors=res["symmetric_errors"])
    if not fm.verify_match_quality(res):                   # :290
        out["status"] = QUALITY_FAIL
    return out


def process_range(descs, kps, pairs=None, mutual: bool = False):
    """All pairs i<j (utils/find_matches.py:329-350) over in-memory features."""
    n = len(descs)
    if pairs is None:
        pairs = [(i, j) for i in range(n) for j in range(i + 1, n)]
    return {(i, j): process_pair(descs[i], kps[i], descs[j], kps[j], mutual=mutual)
            for i, j in pairs}
