"""Coordinates of the six-joint-chain floating base (robot/urdf.py: FLOATING_CHAIN) and their relation to the pose /
body-twist coordinates that floating-base code usually carries (the reference emitter's convention is a quaternion
base with the body twist as velocity, helpers/_topology_helpers.py:231-255).  Small NumPy helpers for the HOST side
of an application (building batches, reading results); nothing here is on the GPU path.

Chain: translations x, y, z along the parent's axes, then rotations about x (roll), y (pitch), z (yaw), each about the
axis carried by the previous joint:   p_world = (x, y, z),   R_world<-base = Rx(roll) Ry(pitch) Rz(yaw).
Body twist of the base link (6-vector [angular; linear], expressed in the base frame) = J(q6) qd6.
Singular at pitch = +-90 degrees (J loses rank), like every Euler-angle base.
"""
from __future__ import annotations

import numpy as np


def _rx(a):
    c, s = np.cos(a), np.sin(a)
    return np.array([[1, 0, 0], [0, c, -s], [0, s, c]])


def _ry(a):
    c, s = np.cos(a), np.sin(a)
    return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]])


def _rz(a):
    c, s = np.cos(a), np.sin(a)
    return np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]])


def pose_from_chain(q6):
    """(position in the world frame, rotation matrix world <- base) of the chain coordinates [x y z roll pitch yaw]."""
    q6 = np.asarray(q6, dtype=np.float64)
    return q6[:3].copy(), _rx(q6[3]) @ _ry(q6[4]) @ _rz(q6[5])


def rotation_from_quaternion(quat_xyzw):
    """Unit quaternion (x, y, z, w) -> rotation matrix (the reference's base state orders it this way:
    s_q = [x y z qx qy qz qw], helpers/_topology_helpers.py:240-255)."""
    x, y, z, w = np.asarray(quat_xyzw, dtype=np.float64) / np.linalg.norm(quat_xyzw)
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def chain_from_pose(position, rotation):
    """Chain coordinates [x y z roll pitch yaw] of a base pose; `rotation` is a 3x3 matrix world <- base or a unit
    quaternion (x, y, z, w).  pitch in [-pi/2, pi/2]."""
    R = np.asarray(rotation, dtype=np.float64)
    if R.shape == (4,):
        R = rotation_from_quaternion(R)
    pitch = np.arcsin(np.clip(R[0, 2], -1.0, 1.0))
    roll = np.arctan2(-R[1, 2], R[2, 2])
    yaw = np.arctan2(-R[0, 1], R[0, 0])
    return np.concatenate([np.asarray(position, dtype=np.float64), [roll, pitch, yaw]])


def twist_jacobian(q6):
    """J(q6), 6 x 6: body twist of the base link [omega; v] (base frame) = J qd6 for the chain rates
    qd6 = d/dt [x y z roll pitch yaw]."""
    q6 = np.asarray(q6, dtype=np.float64)
    _, R = pose_from_chain(q6)
    J = np.zeros((6, 6))
    J[3:, :3] = R.T                                    # world-frame translation rates seen in the base frame
    Rz, Ry = _rz(q6[5]), _ry(q6[4])
    J[:3, 3] = (Ry @ Rz).T @ np.array([1.0, 0, 0])     # roll axis, carried through pitch and yaw
    J[:3, 4] = Rz.T @ np.array([0, 1.0, 0])            # pitch axis, carried through yaw
    J[:3, 5] = np.array([0, 0, 1.0])                   # yaw axis is the base z axis
    return J


def chain_rates_from_twist(q6, twist):
    """qd6 of a body twist [omega; v] given in the base frame."""
    return np.linalg.solve(twist_jacobian(q6), np.asarray(twist, dtype=np.float64))


def generalized_force_to_wrench(q6, c6):
    """The base wrench [moment; force] (base frame) that the chain's first six generalized forces represent:
    c6 = J^T wrench."""
    return np.linalg.solve(twist_jacobian(q6).T, np.asarray(c6, dtype=np.float64))
