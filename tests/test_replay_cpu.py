"""CPU: the on-disk formats of the `--only_optimization` replay (reference sparse_to_dense_predictor.py:29-39,
193-224, 257, 285, 289-304): cache round trip, Prediction reconstruction, summary CSV identical to what pandas writes."""
import io
import os

import numpy as np
import pytest


def _record(seed, n=12, success=True):
    g = np.random.default_rng(seed)
    mask = np.sort(g.permutation(n)[: n // 2])
    M = np.eye(4)
    M[:3, 3] = g.normal(size=3)
    return {"reference_filename": "ref_%d.jpg" % seed, "success": success, "query_2D": g.uniform(0, 256, (n, 2)),
            "reference_2D": g.uniform(0, 256, (n, 2)), "points_3D": g.normal(size=(n, 3)), "num_matches": n,
            "num_inliers": len(mask), "inlier_mask": mask, "quaternion": np.array([1.0, 0.0, 0.0, 0.0]), "matrix": M}


def test_cache_round_trip_and_prediction(tmp_path):
    from pose_prediction import replay
    recs = {"q0.jpg:ref_0.jpg": _record(0), "q1.jpg": _record(1, success=False)}
    out = str(tmp_path) + os.sep
    replay.write_npz(out + "cached_matches/part0.npz", {"q0.jpg:ref_0.jpg": recs["q0.jpg:ref_0.jpg"]})
    replay.write_npz(out + "cached_matches/part1.npz", {"q1.jpg": recs["q1.jpg"]})
    back = replay.read_npz(out)
    assert set(back) == set(recs)
    for k in recs:
        for f in recs[k]:
            assert np.array_equal(np.asarray(back[k][f]), np.asarray(recs[k][f])), (k, f)
    p = replay.prediction_from_record(back["q0.jpg:ref_0.jpg"])
    r = recs["q0.jpg:ref_0.jpg"]
    assert p._fields == ("success", "num_matches", "num_inliers", "reference_inliers", "query_inliers", "points_3d",
                         "quaternion", "matrix", "reference_filename", "reference_keypoints", "inlier_mask")
    assert np.array_equal(p.points_3d, r["points_3D"][r["inlier_mask"]])
    assert np.array_equal(p.reference_inliers, r["reference_2D"][r["inlier_mask"]])
    assert replay.query_of_key("q0.jpg:ref_0.jpg") == "q0.jpg" and replay.query_of_key("q1.jpg") == "q1.jpg"


def test_summary_csv_is_what_pandas_writes(tmp_path):
    pd = pytest.importorskip("pandas")
    from pose_prediction import replay
    rows = [["ref_0.jpg", "q0.jpg", 120, 97, 0.29881234567890123, 0.2919, None], ["ref_1.jpg", "q1.jpg", None, None, None, None, None],
            ["ref_2.jpg", "q2.jpg", 64, 64, 1e-05, 3.5, None]]
    frame = pd.DataFrame(columns=replay.SUMMARY_COLUMNS)
    for i, r in enumerate(rows):
        frame.loc[i] = r
    buf = io.StringIO()
    frame.to_csv(buf, sep=";")
    path = str(tmp_path / "featurePnP_summary.csv")
    replay.write_summary(path, rows)
    assert open(path).read() == buf.getvalue()


def test_pose_files(tmp_path):
    from pose_prediction import replay
    poses = [["a.jpg", 1.0, 0.0, 0.0, 0.0, 0.1, 0.2, 0.3], ["a.jpg", 0.9, 0.1, 0.0, 0.0, 0.11, 0.21, 0.31]]
    path = str(tmp_path / "poses.txt")
    replay.write_poses(path, poses)
    assert open(path).read() == "a.jpg 0.9 0.1 0.0 0.0 0.11 0.21 0.31\n"
    assert open(path.replace(".txt", "_initial.txt")).read() == "a.jpg 1.0 0.0 0.0 0.0 0.1 0.2 0.3\n"
