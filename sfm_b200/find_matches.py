"""Drop-in ``ImageMatcher`` for the preprocess stage, on the B200 matching path.

Mirrors the reference class (utils/find_matches.py:10-364; exported from
utils/__init__.py:3) -- same method names, argument meaning, return types, skip
gates, output files and CSV columns -- so ``main.py preprocess``
(main.py:189-213) and the reconstruct / image-graph / export stages run
unchanged on what it writes.  Not mirrored: ``visualize_features_with_mask`` (:26,
a debugging plot), and ``visualizations/<pair>_matches.png`` (:293-295) is written
only when ``write_visualizations`` is set (``--visualize``): drawing a PNG per
passing pair is host work that no later stage reads.  What differs is where the
arithmetic happens:

  * features are extracted ONCE per image (the reference re-detects both images
    for every pair, :265-266) and uploaded once;
  * ``process_image_range`` hands the whole pair list to the GPU pair engine
    (kNN k=2 -> ratio -> RANSAC F -> verification, sfm_b200/engine.py) instead
    of looping over pairs in Python (:344-350);
  * the per-stage methods (``match_features``, ``geometric_verification``, ...)
    call the same kernels for a single pair.

There is no CPU fallback: without the CUDA library or a GPU every matching call
raises (``lib.SfmError``).  Feature extraction stays on the host with OpenCV,
as in the reference.
"""
from __future__ import annotations

import logging
from pathlib import Path
from types import SimpleNamespace

import numpy as np

from . import dist as D
from . import features as feat
from . import lib as L

CSV_COLUMNS = ["pair_name", "img1", "img2", "num_matches", "num_inliers", "inlier_ratio",
               "reprojection_error", "well_distributed"]          # utils/find_matches.py:301-310


def _dmatch(q: int, t: int, d: float):
    """cv2.DMatch when OpenCV is importable (what the reference returns, :141),
    else an object with the same three attributes."""
    try:
        import cv2
        return cv2.DMatch(int(q), int(t), float(d))
    except Exception:                                              # pragma: no cover
        return SimpleNamespace(queryIdx=int(q), trainIdx=int(t), distance=float(d), imgIdx=0)


class ImageMatcher:
    """utils/find_matches.py:10.  One instance per run, single-threaded caller."""

    def __init__(self, data_dir, device: int = 0, mutual: bool = False, write_visualizations: bool = False,
                 writer_threads: int = 8):
        self.data_dir = Path(data_dir)
        self.image_dir = self.data_dir / "images"
        self.silhouette_dir = self.data_dir / "silhouettes"
        self.matches_dir = self.data_dir / "matches"
        self.fund_dir = self.data_dir / "fundamental"
        self.corr_dir = self.data_dir / "correspondences"
        self.viz_dir = self.data_dir / "visualizations"
        for d in (self.matches_dir, self.fund_dir, self.corr_dir, self.viz_dir):   # :21-22
            d.mkdir(parents=True, exist_ok=True)
        self.results = []                                                          # :24
        self.device = device
        self.mutual = mutual            # README.md:108 mutual-NN check; off = the reference's code path
        self.write_visualizations = write_visualizations
        self.writer_threads = int(writer_threads)   # 0: write the pair files synchronously
        self._engine = None
        self.last_batch = None          # PairBatchResult of the last process_image_range

    # ------------------------------------------------------------------ plumbing
    def _get_engine(self):
        if self._engine is None:
            from .engine import PairEngine
            self._engine = PairEngine(self.device, mutual=self.mutual)
        return self._engine

    def _register(self, descs, kps):
        """Pack per-image (desc u8 [n,128], kps f32 [n,2]) lists and register them on the GPU."""
        import torch
        counts = [len(d) for d in descs]
        off = np.zeros(len(descs) + 1, np.int64)
        off[1:] = np.cumsum(counts)
        total = int(off[-1])
        desc = np.zeros((total, 128), np.uint8)
        kp = np.zeros((total, 2), np.float32)
        for i, (d, k) in enumerate(zip(descs, kps)):
            d = np.asarray(d)
            if len(d):
                d8 = d.astype(np.uint8)
                if not np.array_equal(d8.astype(d.dtype), d):
                    raise ValueError("descriptors must be integer-valued in 0..255 (SIFT)")
                desc[off[i]:off[i + 1]] = d8.reshape(-1, 128)
                if k is not None:
                    kp[off[i]:off[i + 1]] = np.asarray(k, np.float32).reshape(-1, 2)
        eng = self._get_engine()
        eng.upload_images(torch.from_numpy(desc).pin_memory(), torch.from_numpy(kp).pin_memory(), off)
        return eng

    # ------------------------------------------------------------------ features
    def load_mask(self, idx):
        """utils/find_matches.py:47-72."""
        mask = feat.load_mask(self.silhouette_dir, idx)
        if mask is None:
            logging.warning(f"Mask not found: {self.silhouette_dir / f'{idx:05d}.pgm'}")
        return mask

    def detect_features(self, img, mask=None):
        """utils/find_matches.py:74-139 -> (list[cv2.KeyPoint], descriptors), SIFT recipe of
        SURVEY.md 0.2 (the shipped artefacts were produced with SIFT, not the FAST+ORB in the file)."""
        import cv2
        pts, desc = feat.detect_sift(img, mask)
        kps = [cv2.KeyPoint(float(x), float(y), 1.0) for x, y in pts]
        return kps, desc

    # ------------------------------------------------------------------ per-stage calls
    def match_features(self, desc1, desc2):
        """utils/find_matches.py:141-155: kNN k=2 (L2) + Lowe ratio 0.75 -> list of DMatch,
        ascending queryIdx.  Raises ValueError when desc2 has fewer than 2 rows, as the
        reference's ``for m, n in matches`` does (:151)."""
        desc1, desc2 = np.asarray(desc1), np.asarray(desc2)
        if len(desc2) < 2 and len(desc1) > 0:
            raise ValueError("not enough values to unpack (expected 2, got %d)" % len(desc2))
        eng = self._register([desc1, desc2], [None, None])
        ctx = eng.ctx
        pairs = np.array([(0, 1)], np.int32)
        knn, ro = ctx.knn2(pairs)
        rev = ro_rev = None
        if self.mutual:
            rev, ro_rev = ctx.knn2(np.array([(1, 0)], np.int32))
        mq, mt, md2, cnt = ctx.filter_matches(pairs, ro, knn, eng.ratio, ro_rev, rev)
        n = int(cnt.cpu()[0])
        q, t = mq[:n].cpu().numpy(), mt[:n].cpu().numpy()
        d = np.sqrt(md2[:n].cpu().numpy().astype(np.float32))
        return [_dmatch(a, b, c) for a, b, c in zip(q, t, d)]

    def geometric_verification(self, pts1, pts2, F, threshold=3.0):
        """utils/find_matches.py:157-201 on the GPU (sfm_verify_given_F)."""
        import torch
        pts1 = np.ascontiguousarray(np.asarray(pts1, np.float32).reshape(-1, 2))
        pts2 = np.ascontiguousarray(np.asarray(pts2, np.float32).reshape(-1, 2))
        ctx = self._get_engine().ctx
        prm = L.default_params(verify_thresh=float(threshold))
        off = np.array([0, len(pts1)], np.int64)
        o = ctx.verify_given_F(off, torch.from_numpy(pts1).to(ctx.device), torch.from_numpy(pts2).to(ctx.device),
                               torch.from_numpy(np.asarray(F, np.float64).reshape(1, 9)).to(ctx.device), prm)
        m = L.metrics_to_numpy(o["metrics"])[0]
        mask = o["mask"].cpu().numpy().astype(bool)
        return _verification_dict(m, mask, o["sym"].cpu().numpy())

    def verify_match_quality(self, geometric_results, min_inliers=15, min_ratio=0.3, max_error=2.0):
        """utils/find_matches.py:203-214 (scalar logic; nothing to accelerate)."""
        m = geometric_results["metrics"]
        return bool(m["inliers"] >= min_inliers and m["inlier_ratio"] >= min_ratio
                    and m["reprojection_error"] <= max_error and m["well_distributed"])

    def visualize_matches(self, img1, img2, kp1, kp2, matches, geometric_results, save_path):
        """utils/find_matches.py:216-244: the two images side by side, one line per match coloured by
        the inlier flag, the inlier count as a caption.  kp1 / kp2: cv2.KeyPoint lists or [n,2] arrays."""
        import cv2
        mask = geometric_results["inlier_mask"]
        met = geometric_results["metrics"]
        h1, w1 = img1.shape[:2]
        h2, w2 = img2.shape[:2]
        canvas = np.zeros((max(h1, h2), w1 + w2, 3), np.uint8)
        canvas[:h1, :w1] = img1 if img1.ndim == 3 else img1[..., None]
        canvas[:h2, w1:] = img2 if img2.ndim == 3 else img2[..., None]

        def xy(k, i):
            pt = k[i].pt if hasattr(k[i], "pt") else k[i]
            return float(pt[0]), float(pt[1])

        for n, m in enumerate(matches):
            (x1, y1), (x2, y2) = xy(kp1, m.queryIdx), xy(kp2, m.trainIdx)
            a, b = (int(x1), int(y1)), (int(x2 + w1), int(y2))
            colour = (0, 255, 0) if mask[n] else (0, 0, 255)
            cv2.line(canvas, a, b, colour, 1)
            cv2.circle(canvas, a, 3, colour, -1)
            cv2.circle(canvas, b, 3, colour, -1)
        cv2.putText(canvas, f"Inliers: {met['inliers']}/{len(matches)} ({met['inlier_ratio']:.1%})", (10, 30),
                    cv2.FONT_HERSHEY_SIMPLEX, 1, (255, 255, 255), 2)
        cv2.imwrite(str(save_path), canvas)

    def _visualize_pair(self, img1_path, img2_path, pair_name, pts1, pts2, mask):
        """visualize_matches from the packed results of a passing pair (row m matches row m)."""
        import cv2
        img1, img2 = cv2.imread(str(img1_path)), cv2.imread(str(img2_path))
        if img1 is None or img2 is None:
            return
        ms = [SimpleNamespace(queryIdx=i, trainIdx=i) for i in range(len(pts1))]
        res = {"inlier_mask": mask, "metrics": {"inliers": int(mask.sum()),
                                                "inlier_ratio": float(mask.sum()) / max(len(mask), 1)}}
        self.visualize_matches(img1, img2, pts1, pts2, ms, res, self.viz_dir / f"{pair_name}_matches.png")

    # ------------------------------------------------------------------ pairs
    def process_image_pair(self, img1_path, img2_path, pair_name, mask=False):
        """utils/find_matches.py:246-310 for one pair: dict (CSV row) or None on a skip gate."""
        import cv2
        img1_path, img2_path = Path(img1_path), Path(img2_path)
        img1, img2 = cv2.imread(str(img1_path)), cv2.imread(str(img2_path))
        if img1 is None or img2 is None:
            raise ValueError(f"Could not read images: {img1_path}, {img2_path}")          # :253
        m1 = self.load_mask(int(img1_path.stem)) if mask else None
        m2 = self.load_mask(int(img2_path.stem)) if mask else None
        k1, d1 = feat.detect_sift(img1, m1)
        k2, d2 = feat.detect_sift(img2, m2)
        if len(k1) < 1 or len(k2) < 1:                                                     # :268
            return None
        eng = self._register([d1, d2], [k1, k2])
        res = eng.match_pairs(np.array([(0, 1)], np.int32), only_ok=True)
        rows = self._emit(res, [(img1_path, img2_path, pair_name)])
        return rows[0] if rows else None

    def save_pair_data(self, pair_name, pts1, pts2, F, inlier_mask, matches):
        """utils/find_matches.py:312-327: same files, keys and dtypes."""
        inlier_mask = np.asarray(inlier_mask, bool)
        np.save(self.corr_dir / f"{pair_name}_pts1.npy", pts1[inlier_mask])
        np.save(self.corr_dir / f"{pair_name}_pts2.npy", pts2[inlier_mask])
        np.savez(self.fund_dir / f"{pair_name}_F.npz", F=F, mask=inlier_mask, pts1=pts1, pts2=pts2)
        np.savez(self.matches_dir / f"{pair_name}_matches.npz",
                 queryIdx=np.array([m.queryIdx for m in matches]),
                 trainIdx=np.array([m.trainIdx for m in matches]),
                 distance=np.array([m.distance for m in matches]),
                 inlier_mask=inlier_mask)

    def _save_pair_arrays(self, pair_name, pts1, pts2, F, inlier_mask, q, t, dist):
        """save_pair_data without building DMatch objects (identical files)."""
        np.save(self.corr_dir / f"{pair_name}_pts1.npy", pts1[inlier_mask])
        np.save(self.corr_dir / f"{pair_name}_pts2.npy", pts2[inlier_mask])
        np.savez(self.fund_dir / f"{pair_name}_F.npz", F=F, mask=inlier_mask, pts1=pts1, pts2=pts2)
        np.savez(self.matches_dir / f"{pair_name}_matches.npz", queryIdx=q.astype(np.int64),
                 trainIdx=t.astype(np.int64), distance=dist.astype(np.float64), inlier_mask=inlier_mask)

    def _write_pair(self, names_p, pair_args, viz):
        self._save_pair_arrays(*pair_args)
        if viz:
            self._visualize_pair(names_p[0], names_p[1], pair_args[0], pair_args[1], pair_args[2], pair_args[4])

    def _emit(self, res, names):
        """Write the files of every passing pair of a PairBatchResult and build the CSV rows.
        The four files per pair (:312-327) are written by a small thread pool (row N1 of SURVEY 8f:
        at scale the writer, not the GPU, sets the pace; file I/O and zlib release the GIL); all
        writes have finished when this returns, as in the reference.  A pair whose files cannot be
        written is logged and loses its row -- that pair only, as the reference's per-pair
        ``except`` does (:348-350)."""
        rows, jobs = [], []
        pool = None
        if self.writer_threads > 0:
            from concurrent.futures import ThreadPoolExecutor
            pool = ThreadPoolExecutor(max_workers=self.writer_threads)
        try:
            for p, (img1_path, img2_path, pair_name) in enumerate(names):
                if res.status[p] != L.PAIR_OK:
                    continue
                sl = res.pair_slice(p)
                mask = res.inlier_mask[sl]
                m = res.metrics[p]
                args = (pair_name, res.pts1[sl], res.pts2[sl], res.F[p], mask, res.query_idx[sl],
                        res.train_idx[sl], res.distance(p))
                n = int(m["total_matches"])
                row = {
                    "pair_name": pair_name, "img1": Path(img1_path).name, "img2": Path(img2_path).name,
                    "num_matches": n, "num_inliers": np.int64(mask.sum()),
                    "inlier_ratio": np.float64(mask.sum()) / np.float64(n),               # np.mean(bool), :193
                    "reprojection_error": np.float32(m["reprojection_error"]),
                    "well_distributed": bool(m["well_distributed"]),
                }
                if pool is None:
                    try:
                        self._write_pair((img1_path, img2_path), args, self.write_visualizations)
                        rows.append(row)
                    except Exception as e:
                        logging.error(f"Error processing pair {pair_name}: {e}")
                else:
                    jobs.append((pool.submit(self._write_pair, (img1_path, img2_path), args,
                                             self.write_visualizations), row))
            for job, row in jobs:
                try:
                    job.result()
                    rows.append(row)
                except Exception as e:
                    logging.error(f"Error processing pair {row['pair_name']}: {e}")
        finally:
            if pool is not None:
                pool.shutdown(wait=True)
        return rows

    def collect_features(self, start_idx, end_idx, mask=False, ids=None):
        """Features of every existing image id in the inclusive range (or of `ids`), extracted ONCE
        per image (hoisted out of the reference's per-pair loop, :265-266):
        {id: (path, kps f32 [n,2], desc u8 [n,128])}."""
        import cv2
        feats = {}
        for i in (range(start_idx, end_idx + 1) if ids is None else ids):
            p = feat.find_image(self.image_dir, i)
            if p is None:
                continue
            try:
                img = cv2.imread(str(p))
                if img is None:
                    raise ValueError(f"Could not read images: {p}")                      # :253
                m = self.load_mask(i) if mask else None
                k, d = feat.detect_sift(img, m)
                feats[i] = (p, k, d)
            except Exception as e:                                                      # :348-350
                logging.error(f"Error processing image {i}: {e}")
        return feats

    @staticmethod
    def _valid_features(feats):
        """Drop images whose descriptors the GPU path cannot take (not [n,128] integer values 0..255):
        logged per image, only their pairs are lost (the reference loses one pair at a time, :348-350)."""
        ok = {}
        for i, (path, kps, desc) in feats.items():
            d = np.asarray(desc)
            try:
                if d.size and (d.ndim != 2 or d.shape[1] != 128):
                    raise ValueError(f"descriptors must be [n,128], got {d.shape}")
                if d.size and not np.array_equal(d.astype(np.uint8).astype(d.dtype), d):
                    raise ValueError("descriptors must be integer-valued in 0..255 (SIFT)")
                if len(np.asarray(kps).reshape(-1, 2)) != len(d):
                    raise ValueError("keypoints and descriptors differ in length")
                ok[i] = (path, kps, desc)
            except Exception as e:
                logging.error(f"Error processing image {i}: {e}")
        return ok

    def process_features(self, feats, start_idx, end_idx):
        """The pair loop of process_image_range (:331-350) over in-memory features: every pair
        i<j of the inclusive range whose two images exist, as ONE batch on the GPU."""
        feats = self._valid_features(feats)
        ids = sorted(feats)
        pos = {i: n for n, i in enumerate(ids)}
        names, pairs = [], []
        for i in range(start_idx, end_idx):
            for j in range(i + 1, end_idx + 1):
                if i in feats and j in feats:
                    names.append((feats[i][0], feats[j][0], f"pair_{i}_{j}"))
                    pairs.append((pos[i], pos[j]))
        if not pairs:
            return
        try:
            eng = self._register([feats[i][2] for i in ids], [feats[i][1] for i in ids])
            res = eng.match_pairs(np.array(pairs, np.int32), only_ok=True)
            self.last_batch = res
            self.last_names = names
            self.results.extend(self._emit(res, names))
        except L.SfmError:
            raise                                   # no GPU / no library: fail loudly, never fall back
        except Exception as e:
            logging.error(f"Error processing pairs: {e}")

    def process_image_range(self, start_idx, end_idx, mask=False):
        """utils/find_matches.py:329-350: all pairs i<j of the existing images in the inclusive
        range; never raises for a bad pair.  `mask` is used for truthiness only (main.py:199).
        Under torch.distributed (one process per GPU) the images and the pairs are split over the
        ranks."""
        rank, world = D.rank_world()
        if world > 1:
            return self._process_image_range_ranks(start_idx, end_idx, mask, rank, world)
        self.process_features(self.collect_features(start_idx, end_idx, mask), start_idx, end_idx)

    def _process_image_range_ranks(self, start_idx, end_idx, mask, rank, world):
        """One process per GPU: rank r extracts the features of the r-th contiguous block of images,
        one all-gather replicates descriptors and keypoints on every GPU, the pair list is cut into
        contiguous ranges of equal n_i * n_j, every rank matches its range and writes the files of
        its own pairs (names are per pair: the directories equal a 1-GPU run's), and the CSV rows
        meet on rank 0 in pair order.  No collective runs between upload and row gather."""
        import torch
        if torch.cuda.is_available():
            torch.cuda.set_device(self.device)
        ids = [i for i in range(start_idx, end_idx + 1) if feat.find_image(self.image_dir, i) is not None]
        b0, b1 = D.image_blocks(len(ids), world)[rank]
        mine = self._valid_features(self.collect_features(start_idx, end_idx, mask, ids=ids[b0:b1]))
        ok = D.all_gather_objects(sorted(mine))                    # an image that failed is absent everywhere
        ids = [i for part in ok for i in part]
        my_ids = ok[rank]
        eng = self._get_engine()
        dev = eng.ctx.device
        ld = np.concatenate([np.asarray(mine[i][2]).astype(np.uint8).reshape(-1, 128) for i in my_ids]
                            or [np.zeros((0, 128), np.uint8)])
        lk = np.concatenate([np.asarray(mine[i][1], np.float32).reshape(-1, 2) for i in my_ids]
                            or [np.zeros((0, 2), np.float32)])
        desc, kps, off = D.replicate_images(torch.from_numpy(ld).to(dev), torch.from_numpy(lk).to(dev),
                                            [len(mine[i][2]) for i in my_ids])
        pos = {i: n for n, i in enumerate(ids)}
        paths = {i: feat.find_image(self.image_dir, i) for i in ids}
        names, pairs = [], []
        for i in range(start_idx, end_idx):
            for j in range(i + 1, end_idx + 1):
                if i in pos and j in pos:
                    names.append((paths[i], paths[j], f"pair_{i}_{j}"))
                    pairs.append((pos[i], pos[j]))
        rows, err = [], None
        if pairs:
            pairs = np.array(pairs, np.int32)
            lo, hi = D.partition_pairs(pairs, np.diff(off), world)[rank]
            try:
                eng.set_images(desc, kps, off)
                if hi > lo:
                    res = eng.match_pairs(pairs[lo:hi], only_ok=True)
                    self.last_batch = res
                    self.last_names = names[lo:hi]
                    rows = self._emit(res, names[lo:hi])
            except Exception as e:                                  # raised on every rank AFTER the collective
                err = e
                logging.error(f"Error processing pairs on rank {rank}: {e}")
        # the rows and a failure flag travel together: no rank is left waiting in a collective
        parts = D.all_gather_objects((rows, None if err is None else (type(err).__name__, str(err))))
        fatal = [(r, e) for r, (_, e) in enumerate(parts) if e is not None and e[0] == "SfmError"]
        if isinstance(err, L.SfmError):
            raise err
        if fatal:
            raise L.SfmError(f"rank {fatal[0][0]} failed: {fatal[0][1][1]}")
        if rank == 0:
            self.results.extend(r for part, _ in parts for r in part)

    def save_results(self, output_csv):
        """utils/find_matches.py:352-364 (raises KeyError with zero rows, as the reference does)."""
        import pandas as pd
        if D.rank_world()[0] != 0:
            return                                  # the rows were gathered on rank 0
        df = pd.DataFrame(self.results)
        df.to_csv(output_csv, index=False)
        logging.info(f"Results saved to {output_csv}")
        print("\nMatching Summary:")
        print(f"Total pairs processed: {len(self.results)}")
        print(f"Average matches per pair: {df['num_matches'].mean():.1f}")
        print(f"Average inliers per pair: {df['num_inliers'].mean():.1f}")
        print(f"Average inlier ratio: {df['inlier_ratio'].mean():.3f}")
        print(f"Average reprojection error: {df['reprojection_error'].mean():.3f}")
        print(f"Pairs with well-distributed points: {df['well_distributed'].sum()}")


def _verification_dict(m, mask, sym):
    """The dict of geometric_verification (utils/find_matches.py:190-201) from kernel outputs."""
    n = len(mask)
    return {
        "metrics": {
            "total_matches": n,
            "inliers": np.int64(mask.sum()),
            "inlier_ratio": (np.float64(mask.sum()) / np.float64(n)) if n else np.float64("nan"),
            "reprojection_error": np.float32(m["reprojection_error"]),
            "symmetric_error": np.float32(m["symmetric_error"]),
            "well_distributed": bool(m["well_distributed"]),
        },
        "inlier_mask": mask,
        "symmetric_errors": sym,
    }


def run_preprocessing(data_dir, start_idx=0, end_idx=35, mask=None, device=None, mutual=False,
                      visualize=False) -> bool:
    """``main.py preprocess`` (main.py:189-213): returns the stage's bool.  Under ``torchrun`` the
    process group is joined (NCCL) and the GPU is the local rank's."""
    import os
    if device is None:
        device = int(os.environ.get("LOCAL_RANK", 0))
    if int(os.environ.get("WORLD_SIZE", 1)) > 1:
        import torch
        import torch.distributed as dist
        if not dist.is_initialized():
            torch.cuda.set_device(device)
            dist.init_process_group("nccl", device_id=torch.device("cuda", device))
    try:
        matcher = ImageMatcher(data_dir, device=device, mutual=mutual, write_visualizations=visualize)
        matcher.process_image_range(start_idx, end_idx, mask)
        matcher.save_results(Path(data_dir) / "matching_results.csv")
        return True
    except L.SfmError:
        raise
    except Exception as e:
        logging.error(f"Preprocessing failed: {e}")
        return False
