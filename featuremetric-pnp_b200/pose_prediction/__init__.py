"""Device-resident drop-in for the caller of the hot path, `s2dhm/pose_prediction/optimize_feature_pnp.py`."""
