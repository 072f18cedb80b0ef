"""Batched replay of the reference's `--only_optimization` mode: re-run ONLY the feature-metric refinement
from the matches / PnP results the pipeline cached on disk.

Reference (s2dhm/pose_prediction/sparse_to_dense_predictor.py):
  * `read_npz` (:29-39) loads `<output>/cached_matches/*.npz`; every entry is a dict with the keys written at
    :193-204 (`reference_filename, success, query_2D, reference_2D, points_3D, num_matches, num_inliers,
    inlier_mask, quaternion, matrix`);
  * :206-224 turns an entry into a `solve_pnp.Prediction` (inliers only);
  * :243-257 refines one query at a time and appends a row to `result_frame`, written with `sep=";"` (:285);
  * `save` (:289-304) writes the pose files, one `name qw qx qy qz tx ty tz` line per query.

Here the per-query loop becomes ONE `fmpnp_refine_batch` launch over all cached queries: the hypercolumns
are turned into device-resident problems one after the other (Sobel + pack + descriptor gather kernels),
then the whole slice is refined together.  File formats are unchanged.
"""
import os
from collections import namedtuple

import numpy as np
import torch

import fmpnp_b200 as _fm
from pose_prediction import matrix_utils
from pose_prediction.optimize_feature_pnp import build_problem

# same fields, same order as s2dhm/pose_prediction/solve_pnp.py:7-8
Prediction = namedtuple("Prediction", "success num_matches num_inliers reference_inliers query_inliers points_3d "
                                      "quaternion matrix reference_filename reference_keypoints inlier_mask")
SUMMARY_COLUMNS = ["reference_image_origin", "query_image_origin", "num_initial_matches", "num_final_matches",
                   "initial_cost", "final_cost", "track_pickle_path"]


def read_npz(dirpath):
    """{key: record} from every .npz under `<dirpath>cached_matches/` (reference :29-39; like the reference,
    `dirpath` is used as a prefix, so it should end with a path separator)."""
    folder = dirpath + "cached_matches/"
    files = {}
    for _, _, names in os.walk(folder):
        for fn in sorted(names):
            with np.load(folder + fn, allow_pickle=True) as npz:
                files.update({key: npz[key].item() for key in npz})
        break
    return files


def write_npz(path, records):
    """Writes {key: record} the way the pipeline caches it (`np.savez(cache_filename, **cache_dict)`, :228-231)."""
    os.makedirs(os.path.dirname(path), exist_ok=True)
    np.savez(path, **records)


def prediction_from_record(f_data):
    """A cached record -> Prediction restricted to the RANSAC inliers (reference :210-221)."""
    m = f_data["inlier_mask"]
    return Prediction(success=f_data["success"], num_matches=f_data["num_matches"], num_inliers=f_data["num_inliers"],
                      reference_inliers=f_data["reference_2D"][m], query_inliers=f_data["query_2D"][m],
                      points_3d=f_data["points_3D"][m], quaternion=f_data["quaternion"], matrix=f_data["matrix"],
                      reference_filename=f_data["reference_filename"], reference_keypoints=None, inlier_mask=m)


def query_of_key(key):
    """Cache keys are `query` or `query:reference` (:205 writes the latter, :209 reads the former)."""
    return key.split(":")[0]


def refine_cached(records, hypercolumn_fn, intrinsics_fn, image_shape, n_iters, loss_fn="squared_loss", lambda_=0.01,
                  ratio_threshold=None, feature_pyramid=None, output_converter=None):
    """Refines every successful cached prediction in one batched launch.

    records          {key: record} from read_npz
    hypercolumn_fn   image filename -> (1, C, H, W) hypercolumn (what `net.compute_hypercolumn([f])[0]` returns)
    intrinsics_fn    query filename -> 3x3 K
    returns (rows, poses): rows = list of summary rows (SUMMARY_COLUMNS order), poses = the reference's
    `output` list: for every query the initial export followed by the refined one (:236-240, :253, :283).
    """
    keys = [k for k in records]
    preds = {k: prediction_from_record(records[k]) for k in keys}
    todo = [k for k in keys if preds[k].success]
    problems = []
    for k in todo:
        q = query_of_key(k)
        problems.append(build_problem(hypercolumn_fn(q), hypercolumn_fn(preds[k].reference_filename), preds[k],
                                      torch.as_tensor(np.asarray(intrinsics_fn(q))), image_shape))
    levels = None
    if feature_pyramid is not None:
        if any(ts is not None or ks is not None for _, _, ts, ks in feature_pyramid):
            raise NotImplementedError("replay batches channel-slice pyramids only")
        levels = [(int(s), int(e)) for s, e, _, _ in feature_pyramid]
    res = {}
    if problems:
        out = _fm.refine_batch(problems, n_iters, loss=loss_fn, lambda_=lambda_, ratio_threshold=ratio_threshold,
                               levels=levels)
        res = dict(zip(todo, out["results"]))
    name = output_converter or (lambda f: f)
    rows, poses = [], []
    for k in keys:
        p, q = preds[k], query_of_key(k)
        export = [name(q), *list(p.quaternion), *list(np.asarray(p.matrix)[:3, 3])]
        poses.append(list(export))
        if k in res:
            r = res[k]
            T = np.eye(4)
            T[:3, :3] = np.array(r["R"]).reshape(3, 3)
            export[1:5], export[5:] = list(matrix_utils.matrix_quaternion(T)), list(np.array(r["t"]))
            rows.append([p.reference_filename, q, p.num_matches, int(r["best_num_inliers"]), float(r["initial_cost"]),
                         float(r["best_cost"]), None])
        else:
            rows.append([p.reference_filename, q, None, None, None, None, None])
        poses.append(export)
    return rows, poses


def write_summary(path, rows):
    """`result_frame.to_csv(path, sep=";")` (:285): header with an unnamed index column, empty cells for None."""
    with open(path, "w") as f:
        f.write(";" + ";".join(SUMMARY_COLUMNS) + "\n")
        for i, row in enumerate(rows):
            f.write(";".join([str(i)] + ["" if v is None else (repr(float(v)) if isinstance(v, float) else str(v))
                                         for v in row]) + "\n")


def write_poses(path, poses):
    """`save` (:289-304): odd entries (refined) to `path`, even entries (initial) to `*_initial.txt`, space separated."""
    def dump(p, lines):
        with open(p, "w") as f:
            for row in lines:
                f.write(" ".join(str(v) for v in row) + "\n")
    dump(path, poses[1::2])
    dump(path.replace(".txt", "_initial.txt"), poses[::2])
