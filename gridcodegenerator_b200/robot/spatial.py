"""Numeric 6-D spatial algebra used to build robot models (host side, fp64).

Conventions follow Featherstone and the reference's tables: 6-vectors are
[angular(3); linear(3)], X is the *motion* transform parent -> child with
TL = BR = E and TR = 0 (reference: helpers/_topology_helpers.py:224, :323-328).
"""
import numpy as np

SNAP_EPS = 1e-12


def snap(a):
    """Clean values that are zero / +-1 up to rounding (cos(pi/2) = 6e-17 etc.)."""
    a = np.array(a, dtype=np.float64)
    a[np.abs(a) < SNAP_EPS] = 0.0
    for t in (1.0, -1.0):
        a[np.abs(a - t) < SNAP_EPS] = t
    return a


def skew(r):
    x, y, z = r
    return np.array([[0.0, -z, y], [z, 0.0, -x], [-y, x, 0.0]])


def rot_axis(axis, th):
    """Coordinate rotation E for a frame rotated by th about principal axis 0/1/2."""
    c, s = np.cos(th), np.sin(th)
    if axis == 0:
        return np.array([[1, 0, 0], [0, c, s], [0, -s, c]], dtype=np.float64)
    if axis == 1:
        return np.array([[c, 0, -s], [0, 1, 0], [s, 0, c]], dtype=np.float64)
    return np.array([[c, s, 0], [-s, c, 0], [0, 0, 1]], dtype=np.float64)


def rpy_to_E(rpy):
    """URDF rpy (fixed-axis roll, pitch, yaw) -> coordinate transform E = R^T,
    R = Rz(yaw) Ry(pitch) Rx(roll)."""
    r, p, y = rpy
    return snap(rot_axis(0, r) @ rot_axis(1, p) @ rot_axis(2, y))


def X_from_E_r(E, r):
    """Pluecker motion transform for a frame translated by r then rotated by E."""
    X = np.zeros((6, 6))
    X[:3, :3] = E
    X[3:, 3:] = E
    X[3:, :3] = -E @ skew(r)
    return X


def spatial_inertia(mass, com, Ic):
    """6x6 spatial inertia about the frame origin; Ic is the 3x3 rotational
    inertia about the centre of mass in frame axes."""
    C = skew(com)
    I = np.zeros((6, 6))
    I[:3, :3] = Ic + mass * (C @ C.T)
    I[:3, 3:] = mass * C
    I[3:, :3] = mass * C.T
    I[3:, 3:] = mass * np.eye(3)
    return I


def transform_inertia_to_parent(X, I):
    """I expressed in the parent frame of X (X: parent -> child)."""
    return X.T @ I @ X
