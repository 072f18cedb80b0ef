"""Rotation-matrix -> quaternion conversion used by the adapter (reference
s2dhm/pose_prediction/matrix_utils.py:29-74, `matrix_quaternion`, itself Gohlke's transformations.py).
Host-side numpy on one 4x4 matrix per query; not on the data-parallel path."""
import math

import numpy as np


def matrix_quaternion(matrix, isprecise=True):
    """Quaternion (w, x, y, z), w >= 0, of the rotation block of a homogeneous 4x4 matrix.

    isprecise=True assumes an exact rotation matrix (the branch the adapter uses): Shepperd's
    selection of the largest of {trace, M00, M11, M22} with the reference's scaling by M[3, 3].
    isprecise=False takes the dominant eigenvector of the symmetric 4x4 matrix K."""
    M = np.asarray(matrix, dtype=np.float64)[:4, :4]
    if isprecise:
        q = np.empty(4)
        tr = M[0, 0] + M[1, 1] + M[2, 2] + M[3, 3]
        if tr > M[3, 3]:
            s = tr
            q[:] = (s, M[2, 1] - M[1, 2], M[0, 2] - M[2, 0], M[1, 0] - M[0, 1])
        else:
            i = 0  # the reference's comparisons are strict: ties keep the lower index
            if M[1, 1] > M[0, 0]:
                i = 1
            if M[2, 2] > M[i, i]:
                i = 2
            j, k = (i + 1) % 3, (i + 2) % 3
            s = M[i, i] - (M[j, j] + M[k, k]) + M[3, 3]
            v = np.empty(3)
            v[i] = s
            v[j] = M[i, j] + M[j, i]
            v[k] = M[k, i] + M[i, k]
            q[0] = M[k, j] - M[j, k]
            q[1:] = v
        q *= 0.5 / math.sqrt(s * M[3, 3])
    else:
        m = M
        K = np.array([[m[0, 0] - m[1, 1] - m[2, 2], 0.0, 0.0, 0.0],
                      [m[0, 1] + m[1, 0], m[1, 1] - m[0, 0] - m[2, 2], 0.0, 0.0],
                      [m[0, 2] + m[2, 0], m[1, 2] + m[2, 1], m[2, 2] - m[0, 0] - m[1, 1], 0.0],
                      [m[2, 1] - m[1, 2], m[0, 2] - m[2, 0], m[1, 0] - m[0, 1], m[0, 0] + m[1, 1] + m[2, 2]]]) / 3.0
        w, V = np.linalg.eigh(K)
        q = V[[3, 0, 1, 2], int(np.argmax(w))]
    if q[0] < 0.0:
        q = -q
    return q
