"""URDF ingestion: the stdlib-XML reader + robot-model layer that stand in for the un-vendored
URDFParser package (reference README.md:10-21).  Every shipped robot goes through this path, so its corner cases are
tested here: fixed-joint merging, links without <inertial>, the default axis, prismatic joints, branched trees with
DFS pre-order ids, negative and general joint axes (joint frame re-oriented so the motion subspace stays one-hot,
reference helpers/_spatial_algebra_helpers.py:13)."""
import os

import numpy as np
import pytest

from gridcodegenerator_b200.robot.urdf import URDFParser, parse_urdf_string
from oracle import Oracle

G = 9.81


def _urdf(body: str) -> str:
    return f'<robot name="t">{body}</robot>'


def _link(name, mass=None, com="0 0 0", inertia=(0, 0, 0, 0, 0, 0)):
    if mass is None:
        return f'<link name="{name}"/>'
    ixx, ixy, ixz, iyy, iyz, izz = inertia
    return (f'<link name="{name}"><inertial><origin xyz="{com}" rpy="0 0 0"/><mass value="{mass}"/>'
            f'<inertia ixx="{ixx}" ixy="{ixy}" ixz="{ixz}" iyy="{iyy}" iyz="{iyz}" izz="{izz}"/></inertial></link>')


def _joint(name, jtype, parent, child, xyz="0 0 0", rpy="0 0 0", axis=None, damping=None):
    ax = f'<axis xyz="{axis}"/>' if axis else ""
    dy = f'<dynamics damping="{damping}"/>' if damping is not None else ""
    return (f'<joint name="{name}" type="{jtype}"><parent link="{parent}"/><child link="{child}"/>'
            f'<origin xyz="{xyz}" rpy="{rpy}"/>{ax}{dy}</joint>')


def test_fixed_joint_children_are_merged_into_the_moving_body():
    a = parse_urdf_string(_urdf(
        _link("base") + _link("l1", 2.0, "0 0 0.1", (0.01, 0, 0, 0.02, 0, 0.03)) + _link("tool", 0.5, "0.05 0 0", (0.001, 0, 0, 0.001, 0, 0.001))
        + _joint("j1", "revolute", "base", "l1", axis="0 0 1")
        + _joint("f", "fixed", "l1", "tool", xyz="0 0 0.3", rpy="0 0 0")))
    assert a.get_num_pos() == 1 and a.get_num_fixed_joints() == 1
    I = a.get_Imat_by_id(0)
    assert abs(I[3, 3] - 2.5) < 1e-12                              # masses add
    com = np.array([I[2, 4], I[0, 5], I[1, 3]]) / 2.5              # m * skew(c) block
    assert np.allclose(com, (2.0 * np.array([0, 0, 0.1]) + 0.5 * np.array([0.05, 0, 0.3])) / 2.5, atol=1e-12)
    assert a.get_fixed_joint_by_name("f").get_parent() == "l1"


def test_link_without_inertial_and_default_axis_and_damping():
    r = parse_urdf_string(_urdf(
        _link("base") + _link("massless") + _link("l2", 1.0, "0 0.2 0", (0.01, 0, 0, 0.01, 0, 0.01))
        + _joint("j1", "revolute", "base", "massless")               # no <axis>: URDF default (1 0 0)
        + _joint("j2", "continuous", "massless", "l2", xyz="0 0 0.1", axis="0 1 0", damping=0.7)))
    assert r.get_num_pos() == 2
    assert np.array_equal(r.get_S_by_id(0), [1, 0, 0, 0, 0, 0])
    assert np.array_equal(r.get_S_by_id(1), [0, 1, 0, 0, 0, 0])
    assert np.abs(r.get_Imat_by_id(0)).max() == 0.0
    assert r.get_damping_by_id(1) == 0.7 and r.get_joint_by_id(1).get_type() == "revolute"
    c = Oracle(r).rnea(np.zeros(2), np.array([0.0, 2.0]), np.zeros(2), gravity=0.0)[0]
    assert abs(c[1] - 1.4) < 1e-12                                 # only the damping term d * qd


def test_prismatic_joint_and_branched_tree_ids():
    r = parse_urdf_string(_urdf(
        _link("base") + "".join(_link(n, 1.0, "0 0 0.1", (0.01, 0, 0, 0.01, 0, 0.01)) for n in ("a", "a1", "a2", "b", "b1"))
        + _joint("ja", "revolute", "base", "a", axis="0 0 1") + _joint("jb", "prismatic", "base", "b", axis="0 0 1")
        + _joint("ja1", "revolute", "a", "a1", xyz="0.1 0 0", axis="0 1 0") + _joint("ja2", "revolute", "a", "a2", xyz="-0.1 0 0", axis="1 0 0")
        + _joint("jb1", "prismatic", "b", "b1", xyz="0 0.1 0", axis="1 0 0")))
    n = r.get_num_pos()
    assert n == 5
    par = r.get_parent_id_array()
    assert all(par[j] < j for j in range(n))                        # parent id < child id
    for j in range(n):
        st = r.get_subtree_by_id(j)
        assert st == list(range(j, j + len(st)))                    # every subtree is a contiguous id range
    names = [r.get_joint_by_id(j).get_name() for j in range(n)]
    assert names.index("ja1") > names.index("ja") and names.index("jb1") == names.index("jb") + 1
    jb = names.index("jb")
    assert np.array_equal(r.get_S_by_id(jb), [0, 0, 0, 0, 0, 1]) and r.get_joint_by_id(jb).get_type() == "prismatic"
    assert sorted(r.get_leaf_nodes()) == sorted(names.index(x) for x in ("ja1", "ja2", "jb1"))
    # a prismatic joint along z carries the whole weight of its subtree
    c = Oracle(r).rnea(np.zeros(n), np.zeros(n), np.zeros(n))[0]
    assert abs(c[jb] - 2.0 * G) < 1e-10


def _two_link(axis1, axis2):
    return parse_urdf_string(_urdf(
        _link("base") + _link("l1", 1.3, "0.02 -0.03 0.11", (0.02, 0.001, -0.002, 0.03, 0.0005, 0.025))
        + _link("l2", 0.9, "-0.04 0.05 0.07", (0.011, -0.0007, 0.0004, 0.012, 0.0009, 0.008))
        + _joint("j1", "revolute", "base", "l1", xyz="0 0 0.1", rpy="0.3 -0.2 0.5", axis=axis1, damping=0.1)
        + _joint("j2", "revolute", "l1", "l2", xyz="0.05 0.02 0.25", rpy="-0.4 0.1 0.2", axis=axis2)))


@pytest.mark.parametrize("flip", [(True, False), (False, True), (True, True)])
def test_negative_axis_equals_mirrored_model(flip):
    """axis = -e_k is the same mechanism as axis = +e_k with q -> -q: tau_neg(q, qd, qdd) = -tau_pos(-q, -qd, -qdd) on
    the flipped joints, for RNEA, its gradient, Minv and FD gradient (1e-12); also through the reference's own
    test_rnea when the reference tree is present (build container only)."""
    pos = _two_link("0 0 1", "0 1 0")
    neg = _two_link("0 0 -1" if flip[0] else "0 0 1", "0 -1 0" if flip[1] else "0 1 0")
    assert np.array_equal(neg.get_S_by_id(0), [0, 0, 1, 0, 0, 0]) and np.array_equal(neg.get_S_by_id(1), [0, 1, 0, 0, 0, 0])
    sg = np.array([-1.0 if f else 1.0 for f in flip])
    op, on = Oracle(pos), Oracle(neg)
    rng = np.random.default_rng(5)
    for _ in range(4):
        q, qd, qdd, u = (rng.uniform(-1, 1, 2) for _ in range(4))
        cn = on.rnea(q, qd, qdd)[0]
        cp = op.rnea(sg * q, sg * qd, sg * qdd)[0]
        assert np.abs(cn - sg * cp).max() < 1e-12
        S = np.diag(sg)
        assert np.abs(on.minv(q) - S @ op.minv(sg * q) @ S).max() < 1e-11
        gp = op.fd_grad(sg * q, sg * qd, sg * u)
        gn = on.fd_grad(q, qd, u)
        S2 = np.diag(np.concatenate([sg, sg]))
        assert np.abs(gn - S @ gp @ S2).max() < 1e-9
        if os.path.isdir("/root/reference"):
            import contextlib
            import io
            import sys
            sys.path.insert(0, "/root")
            from reference import GRiDCodeGenerator as RefGen
            with contextlib.redirect_stdout(io.StringIO()):
                rn = np.asarray(RefGen(neg).test_rnea(q, qd, qdd)[0])
                rp = np.asarray(RefGen(pos).test_rnea(sg * q, sg * qd, sg * qdd)[0])
            assert np.abs(rn - sg * rp).max() < 1e-12 and np.abs(rn - cn).max() < 1e-12


def test_general_axis_pendulum_physics():
    """A joint about an arbitrary unit axis a: gravity torque a . (p x m g z) and inertia a^T (Ic + m(|p|^2 - p p^T)) a."""
    a = np.array([1.0, 2.0, -0.5])
    a /= np.linalg.norm(a)
    p = np.array([0.1, -0.2, 0.3])
    m, Ic = 1.7, np.diag([0.02, 0.03, 0.04])
    r = parse_urdf_string(_urdf(
        _link("base") + _link("l", m, " ".join(map(str, p)), (0.02, 0, 0, 0.03, 0, 0.04))
        + _joint("j", "revolute", "base", "l", axis=" ".join(map(str, a)))))
    orc = Oracle(r)
    tau_g = orc.rnea(np.zeros(1), np.zeros(1), np.zeros(1))[0][0]
    assert abs(tau_g - a @ np.cross(p, [0, 0, m * G])) < 1e-10
    M = 1.0 / orc.minv(np.zeros(1))[0, 0]
    assert abs(M - a @ (Ic + m * (p @ p * np.eye(3) - np.outer(p, p))) @ a) < 1e-10


def test_urdfparser_facade_and_errors(tmp_path):
    f = tmp_path / "p.urdf"
    f.write_text(_urdf(_link("base") + _link("l", 1.0) + _joint("j", "revolute", "base", "l", axis="0 0 1")))
    r = URDFParser().parse(str(f))
    assert r.name == "p" and r.get_num_pos() == 1 and r.floating_base is False
    with pytest.raises(ValueError):
        parse_urdf_string("<notarobot/>")
    with pytest.raises(ValueError):
        parse_urdf_string(_urdf(_link("a") + _link("b")))                                  # two roots
    with pytest.raises(ValueError):
        parse_urdf_string(_urdf(_link("base") + _link("l") + _joint("j", "planar", "base", "l")))


# ---------------------------------------------------------------- floating base
def _floating_pair():
    """(fixed-base HyQ, HyQ hanging from a free joint expanded into the six-joint chain)"""
    from gridcodegenerator_b200.robot import LIBRARY_DIR, load_robot
    fixed = load_robot("hyq")
    floating = URDFParser().parse(os.path.join(LIBRARY_DIR, "hyq.urdf"), floating_base=True)
    return fixed, floating


def test_floating_base_flag_expands_the_root_into_a_six_joint_chain():
    from gridcodegenerator_b200.robot import load_robot
    fixed, fl = _floating_pair()
    assert fl.get_num_pos() == fixed.get_num_pos() + 6 and fl.floating_base_chain and not fl.floating_base
    types = [fl.get_joint_by_id(j).get_type() for j in range(6)]
    assert types == ["prismatic"] * 3 + ["revolute"] * 3
    S = [int(np.argmax(fl.get_S_by_id(j))) for j in range(6)]
    assert S == [3, 4, 5, 0, 1, 2]                                   # translations x y z, then rotations x y z
    assert fl.get_parent_id_array()[:7] == [-1, 0, 1, 2, 3, 4, 5]
    # the five intermediate links are massless; the sixth body is the trunk
    for j in range(5):
        assert np.abs(np.asarray(fl.get_Imat_by_id(j), dtype=float)).max() == 0.0
    assert np.allclose(np.asarray(fl.get_Imat_by_id(5), dtype=float), np.asarray(fixed.base_Imat, dtype=float), atol=1e-12)
    # the shipped hyq_floating.urdf (<joint type="floating">) is the same model
    lib = load_robot("hyq_floating")
    assert lib.get_num_pos() == fl.get_num_pos() and lib.get_parent_id_array() == fl.get_parent_id_array()
    q = np.random.default_rng(3).uniform(-1, 1, fl.get_num_pos())
    for j in range(fl.get_num_pos()):
        assert np.abs(np.asarray(lib.get_Xmat_Func_by_id(j)(q[j]), dtype=float) -
                      np.asarray(fl.get_Xmat_Func_by_id(j)(q[j]), dtype=float)).max() < 1e-14


def test_floating_base_dynamics_reduce_to_the_fixed_base_robot():
    """Base pose at the origin, base at rest: the leg torques of the floating model are the fixed-base robot's; the
    leg block of M is the fixed-base M; the base z force carries the robot's weight; and a base that is FREE
    (zero base wrench) in free fall makes every joint torque of a motionless robot vanish."""
    fixed, fl = _floating_pair()
    of, ol = Oracle(fixed), Oracle(fl)
    n = fixed.get_num_pos()
    rng = np.random.default_rng(11)
    mass = 53.433 + 4 * (2.93 + 2.638 + 0.881)
    for _ in range(3):
        q, qd, qdd = (rng.uniform(-1, 1, n) for _ in range(3))
        z6 = np.zeros(6)
        cl = ol.rnea(np.concatenate([z6, q]), np.concatenate([z6, qd]), np.concatenate([z6, qdd]))[0]
        cf = of.rnea(q, qd, qdd)[0]
        assert np.abs(cl[6:] - cf).max() < 1e-10
        Ml, Mf = ol.crba(np.concatenate([z6, q])), of.crba(q)
        assert np.abs(Ml[6:, 6:] - Mf).max() < 1e-10
        assert np.abs(Ml[:3, :3] - mass * np.eye(3)).max() < 1e-9    # translations of the whole robot: total mass
        c0 = ol.rnea(np.concatenate([z6, q]), np.zeros(n + 6), np.zeros(n + 6))[0]
        assert abs(c0[2] - mass * G) < 1e-9 and abs(c0[0]) < 1e-9 and abs(c0[1]) < 1e-9
        # forward dynamics of the free robot under zero generalized force: the base falls with g, joints do not move
        base = np.concatenate([rng.uniform(-1, 1, 3), rng.uniform(-0.5, 0.5, 3)])          # any base pose
        qdd_free = ol.fd(np.concatenate([base, q]), np.zeros(n + 6), np.zeros(n + 6))
        # damping is zero at rest, so the only force is gravity: uniform acceleration -g along world z
        assert np.abs(qdd_free[6:]).max() < 1e-8 and np.abs(qdd_free[3:6]).max() < 1e-8
        assert np.abs(qdd_free[:3] - np.array([0.0, 0.0, -G])).max() < 1e-8


def test_floating_joint_in_the_middle_of_a_tree():
    """type="floating" is not only for the root: a free-flying payload on top of a one-joint arm."""
    r = parse_urdf_string(_urdf(
        _link("base") + _link("arm", 1.0, "0 0 0.2", (0.01, 0, 0, 0.01, 0, 0.002)) + _link("payload", 2.0, "0 0 0", (0.02, 0, 0, 0.02, 0, 0.02))
        + _joint("j1", "revolute", "base", "arm", axis="0 1 0")
        + _joint("free", "floating", "arm", "payload", xyz="0 0 0.4")))
    assert r.get_num_pos() == 7 and r.floating_base_chain
    names = [r.get_joint_by_id(j).get_name() for j in range(7)]
    assert names == ["j1", "free_px", "free_py", "free_pz", "free_rx", "free_ry", "free_rz"]
    c = Oracle(r).rnea(np.zeros(7), np.zeros(7), np.zeros(7))[0]
    assert abs(c[3] - 2.0 * G) < 1e-10                               # the z translation carries the payload's weight


def test_floating_chain_coordinate_helpers():
    """robot/floating.py against the model itself: the pose of the chain coordinates is the product of the robot's own
    homogeneous joint transforms, the twist Jacobian reproduces the base-link velocity the oracle's RNEA computes, the
    conversions round-trip, and the base wrench of a robot at rest is its weight along world z."""
    import sympy as sp
    from gridcodegenerator_b200.robot import floating as fl
    from gridcodegenerator_b200.robot.model import THETA
    _, robot = _floating_pair()
    orc = Oracle(robot)
    n = robot.get_num_pos()
    rng = np.random.default_rng(17)
    hom = [sp.lambdify(THETA, M, "numpy") for M in robot.get_Xmats_hom_ordered_by_id()[:6]]
    for _ in range(4):
        q6 = np.concatenate([rng.uniform(-1, 1, 3), rng.uniform(-1.2, 1.2, 3)])
        T = np.eye(4)
        for j in range(6):
            T = T @ np.array(hom[j](q6[j]), dtype=np.float64)
        p, R = fl.pose_from_chain(q6)
        assert np.abs(T[:3, 3] - p).max() < 1e-12 and np.abs(T[:3, :3] - R).max() < 1e-12
        assert np.abs(fl.chain_from_pose(p, R) - q6).max() < 1e-12
        # quaternion (x, y, z, w) of R by the standard formula -> same coordinates
        w = 0.5 * np.sqrt(max(1e-16, 1 + np.trace(R)))
        quat = np.array([(R[2, 1] - R[1, 2]) / (4 * w), (R[0, 2] - R[2, 0]) / (4 * w), (R[1, 0] - R[0, 1]) / (4 * w), w])
        assert np.abs(fl.chain_from_pose(p, quat) - q6).max() < 1e-10
        # twist of the base link: column 5 of the oracle's v is the base link's spatial velocity
        qd6 = rng.uniform(-1, 1, 6)
        q = np.concatenate([q6, rng.uniform(-1, 1, n - 6)])
        qd = np.concatenate([qd6, np.zeros(n - 6)])
        v = orc.rnea(q, qd, np.zeros(n))[1]
        J = fl.twist_jacobian(q6)
        assert np.abs(J @ qd6 - v[:, 5]).max() < 1e-10
        assert np.abs(fl.chain_rates_from_twist(q6, J @ qd6) - qd6).max() < 1e-10
        # robot at rest: generalized base forces = J^T [moment; force]; the force is the weight along WORLD z
        c = orc.rnea(q, np.zeros(n), np.zeros(n))[0]
        wrench = fl.generalized_force_to_wrench(q6, c[:6])
        mass = 53.433 + 4 * (2.93 + 2.638 + 0.881)
        assert np.abs(R @ wrench[3:] - np.array([0, 0, mass * G])).max() < 1e-8
