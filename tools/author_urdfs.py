#!/usr/bin/env python
"""Writes the robot descriptions shipped in gridcodegenerator_b200/robot/library.

No URDF exists in the reference tree or on the box, so the
descriptions are authored here: kinematic structure, joint-axis mix and mass
distribution follow the public KUKA LBR iiwa 14, IIT HyQ and Boston Dynamics
Atlas (v5) models; digits are approximate.  Parity is defined against the
reference algorithms run on the *same* robot object, so it does not depend on
matching an upstream URDF digit for digit.
"""
import os

PI = 3.141592653589793
HP = PI / 2

OUT = os.path.join(os.path.dirname(__file__), "..", "gridcodegenerator_b200", "robot", "library")


def link(name, mass, com, inertia, com_rpy=(0, 0, 0)):
    ixx, iyy, izz = inertia[:3]
    ixy, ixz, iyz = (inertia[3:] + (0, 0, 0))[:3] if len(inertia) > 3 else (0, 0, 0)
    return (
        f'  <link name="{name}">\n'
        f"    <inertial>\n"
        f'      <origin xyz="{com[0]} {com[1]} {com[2]}" rpy="{com_rpy[0]} {com_rpy[1]} {com_rpy[2]}"/>\n'
        f'      <mass value="{mass}"/>\n'
        f'      <inertia ixx="{ixx}" ixy="{ixy}" ixz="{ixz}" iyy="{iyy}" iyz="{iyz}" izz="{izz}"/>\n'
        f"    </inertial>\n"
        f"  </link>\n"
    )


def joint(name, jtype, parent, child, xyz, rpy=(0, 0, 0), axis=(0, 0, 1), damping=0.0, limits=(-3.0, 3.0, 300.0, 10.0)):
    s = f'  <joint name="{name}" type="{jtype}">\n'
    s += f'    <parent link="{parent}"/>\n    <child link="{child}"/>\n'
    s += f'    <origin xyz="{xyz[0]} {xyz[1]} {xyz[2]}" rpy="{rpy[0]} {rpy[1]} {rpy[2]}"/>\n'
    if jtype != "fixed":
        s += f'    <axis xyz="{axis[0]} {axis[1]} {axis[2]}"/>\n'
        s += f'    <limit lower="{limits[0]}" upper="{limits[1]}" effort="{limits[2]}" velocity="{limits[3]}"/>\n'
        s += f'    <dynamics damping="{damping}"/>\n'
    s += "  </joint>\n"
    return s


def write(name, body):
    os.makedirs(OUT, exist_ok=True)
    with open(os.path.join(OUT, name + ".urdf"), "w") as f:
        f.write(f'<?xml version="1.0"?>\n<robot name="{name}">\n{body}</robot>\n')


X, Y, Z = (1, 0, 0), (0, 1, 0), (0, 0, 1)


def iiwa14():
    b = link("iiwa_link_0", 5.0, (-0.1, 0, 0.07), (0.05, 0.06, 0.03))
    L = [
        ("iiwa_link_1", 4.0, (0, -0.03, 0.12), (0.1, 0.09, 0.02)),
        ("iiwa_link_2", 4.0, (0.0003, 0.059, 0.042), (0.05, 0.018, 0.044)),
        ("iiwa_link_3", 3.0, (0, 0.03, 0.13), (0.08, 0.075, 0.01)),
        ("iiwa_link_4", 2.7, (0, 0.067, 0.034), (0.03, 0.01, 0.029)),
        ("iiwa_link_5", 1.7, (0.0001, 0.021, 0.076), (0.02, 0.018, 0.005)),
        ("iiwa_link_6", 1.8, (0, 0.0006, 0.0004), (0.005, 0.0036, 0.0047)),
        ("iiwa_link_7", 0.3, (0, 0, 0.02), (0.001, 0.001, 0.001)),
    ]
    J = [
        ((0, 0, 0.1575), (0, 0, 0), (-2.9668, 2.9668, 320, 1.4835)),
        ((0, 0, 0.2025), (HP, 0, PI), (-2.0942, 2.0942, 320, 1.4835)),
        ((0, 0.2045, 0), (HP, 0, PI), (-2.9668, 2.9668, 176, 1.7452)),
        ((0, 0, 0.2155), (HP, 0, 0), (-2.0942, 2.0942, 176, 1.3089)),
        ((0, 0.1845, 0), (-HP, PI, 0), (-2.9668, 2.9668, 110, 2.2688)),
        ((0, 0, 0.2155), (HP, 0, 0), (-2.0942, 2.0942, 40, 2.3561)),
        ((0, 0.081, 0), (-HP, PI, 0), (-3.0541, 3.0541, 40, 2.3561)),
    ]
    prev = "iiwa_link_0"
    for i, ((lname, m, com, ine), (xyz, rpy, lim)) in enumerate(zip(L, J)):
        b += link(lname, m, com, ine)
        b += joint(f"iiwa_joint_{i+1}", "revolute", prev, lname, xyz, rpy, Z, damping=0.5, limits=lim)
        prev = lname
    # flange as a fixed kinematic target
    b += link("iiwa_link_ee", 0.0, (0, 0, 0), (0, 0, 0))
    b += joint("iiwa_joint_ee", "fixed", prev, "iiwa_link_ee", (0, 0, 0.045))
    write("iiwa14", b)


def hyq():
    b = _hyq_body()
    write("hyq", b)
    # the same robot with a free joint between a massless world link and the trunk: the URDF reader expands
    # type="floating" into a six-joint chain (robot/urdf.py: FLOATING_CHAIN) -> 18 degrees of freedom
    fb = ('  <link name="world"/>\n  <joint name="floating_base" type="floating">\n    <parent link="world"/>\n'
          '    <child link="trunk"/>\n    <origin xyz="0 0 0" rpy="0 0 0"/>\n  </joint>\n')
    write("hyq_floating", fb + b)


def _hyq_body():
    b = link("trunk", 53.433, (0.056, 0.0215, 0.00358), (1.5725, 8.5015, 9.1954, 0.03, -0.05, 0.0))
    legs = {
        "lf": (0.3735, 0.207, 1, 1),
        "rf": (0.3735, -0.207, 1, -1),
        "lh": (-0.3735, 0.207, -1, 1),
        "rh": (-0.3735, -0.207, -1, -1),
    }
    for leg, (x, y, fx, fy) in legs.items():
        hip, upper, lower, foot = f"{leg}_hipassembly", f"{leg}_upperleg", f"{leg}_lowerleg", f"{leg}_foot"
        b += link(hip, 2.93, (0.04263 * fx, 0.0 * fy, -0.16931 * 0.1), (0.0050, 0.0111, 0.0112, -0.0003 * fx * fy, 0.0002 * fx, 0.0))
        b += joint(f"{leg}_haa_joint", "revolute", "trunk", hip, (x, y, 0.0), (0, 0, 0), X, damping=0.1, limits=(-1.22, 0.44, 120, 12))
        b += link(upper, 2.638, (0.0 + 0.02 * fx, -0.0003 * fy, -0.15074), (0.0368, 0.0423, 0.0058, 0.0, 0.0018 * fx, 0.0))
        b += joint(f"{leg}_hfe_joint", "revolute", hip, upper, (0.08 * fx, 0.0, 0.0), (0, 0, 0), Y, damping=0.1, limits=(-0.87, 1.22, 150, 12))
        b += link(lower, 0.881, (0.0 - 0.01 * fx, 0.0001 * fy, -0.1254), (0.0108, 0.0110, 0.0004, 0.0, -0.0003 * fx, 0.0))
        b += joint(f"{leg}_kfe_joint", "revolute", upper, lower, (0.0, 0.0, -0.35), (0, 0, 0), Y, damping=0.1, limits=(-2.44, -0.35, 150, 12))
        b += link(foot, 0.0, (0, 0, 0), (0, 0, 0))
        b += joint(f"{leg}_foot_joint", "fixed", lower, foot, (0.0, 0.0, -0.33))
    return b


def atlas():
    b = link("pelvis", 9.509, (0.0111, 0.0, 0.0271), (0.1244, 0.0949, 0.1347, 0.0008, -0.0007, -0.0005))
    d = 0.1

    def chain(parent, specs):
        nonlocal b
        for jname, axis, child, xyz, mass, com, ine in specs:
            b += link(child, mass, com, ine)
            b += joint(jname, "revolute", parent, child, xyz, (0, 0, 0), axis, damping=d)
            parent = child
        return parent

    # torso
    chain("pelvis", [
        ("back_bkz", Z, "ltorso", (-0.0125, 0, 0), 2.27, (-0.0113, 0.0, 0.0747), (0.0039, 0.0034, 0.0017, 0.0, -0.0005, 0.0)),
        ("back_bky", Y, "mtorso", (0, 0, 0.162), 0.799, (-0.0082, -0.0132, 0.0307), (0.00045, 0.00065, 0.00038, 0.0, 0.0, 0.0)),
        ("back_bkx", X, "utorso", (0, 0, 0.05), 84.409, (-0.0623, 0.0023, 0.3157), (1.577, 1.602, 0.565, 0.0103, 0.0348, 0.0083)),
    ])
    chain("utorso", [
        ("neck_ry", Y, "head", (0.2546, 0, 0.6215), 1.42, (-0.0754, 0.0, 0.0358), (0.0039, 0.0041, 0.0035, 0.0, -0.0008, 0.0)),
    ])
    for side, s in (("l", 1.0), ("r", -1.0)):
        chain("utorso", [
            (f"{side}_arm_shz", Z, f"{side}_clav", (0.1406, 0.2256 * s, 0.4776), 4.466, (0.0, 0.048 * s, -0.084), (0.011, 0.009, 0.004, 0.0, 0.0, 0.004 * s)),
            (f"{side}_arm_shx", X, f"{side}_scap", (0, 0.11 * s, -0.245), 3.899, (0.0, 0.0 * s, 0.0), (0.00319, 0.00583, 0.00583, 0.0, 0.0, 0.0)),
            (f"{side}_arm_ely", Y, f"{side}_uarm", (0, 0.187 * s, -0.016), 4.386, (0.0, 0.065 * s, 0.0), (0.00656, 0.00358, 0.00656, 0.0, 0.0, 0.0)),
            (f"{side}_arm_elx", X, f"{side}_larm", (0, 0.119 * s, 0.0092), 3.248, (0.0, 0.0 * s, 0.0), (0.00265, 0.00446, 0.00446, 0.0, 0.0, 0.0)),
            (f"{side}_arm_wry", Y, f"{side}_ufarm", (0, 0.29955 * s, -0.00921), 2.4798, (0.00015, 0.08296 * s, 0.00037), (0.01267, 0.00264, 0.01300, 0.0, 0.0, 0.0)),
            (f"{side}_arm_wrx", X, f"{side}_lfarm", (0, 0, 0), 0.648, (0.00017, -0.02515 * s, 0.00163), (0.00076, 0.00083, 0.00077, 0.0, 0.0, 0.0)),
            (f"{side}_arm_wry2", Y, f"{side}_hand", (0, 0.08 * s, 0), 0.5839, (-0.00016, 0.05 * s, 0.00266), (0.00039, 0.00029, 0.00041, 0.0, 0.0, 0.0)),
        ])
    for side, s in (("l", 1.0), ("r", -1.0)):
        chain("pelvis", [
            (f"{side}_leg_hpz", Z, f"{side}_uglut", (0, 0.089 * s, 0), 1.959, (0.00529, -0.00344 * s, 0.00313), (0.00074, 0.00069, 0.00135, 0.0, 0.0, 0.0)),
            (f"{side}_leg_hpx", X, f"{side}_lglut", (0, 0, -0.02), 0.898, (0.0133, 0.0170 * s, -0.0312), (0.00069, 0.00413, 0.00387, 0.0, 0.0001, 0.0)),
            (f"{side}_leg_hpy", Y, f"{side}_uleg", (0.05, 0.0225 * s, -0.066), 8.204, (0.0, 0.0, -0.21), (0.09, 0.09, 0.02, 0.0, 0.0, 0.0)),
            (f"{side}_leg_kny", Y, f"{side}_lleg", (-0.05, 0, -0.374), 4.515, (0.001, 0.0, -0.187), (0.077, 0.076, 0.01, 0.0, -0.0007, 0.0)),
            (f"{side}_leg_aky", Y, f"{side}_talus", (0, 0, -0.422), 0.125, (0.0, 0.0, 0.0), (1.017e-05, 8.284e-06, 1.302e-05, 0.0, 0.0, 0.0)),
            (f"{side}_leg_akx", X, f"{side}_foot", (0, 0, -0.01), 2.41, (0.027, 0.0, -0.067), (0.002, 0.007, 0.008, 0.0, 0.0, 0.0)),
        ])
    write("atlas", b)


def toys():
    # 2-link planar pendulum about y, 3-link chain with mixed axes + a prismatic joint, branched toy
    b = link("base", 1.0, (0, 0, 0), (0.01, 0.01, 0.01))
    b += link("l1", 1.0, (0, 0, -0.25), (0.02, 0.02, 0.001)) + joint("j1", "revolute", "base", "l1", (0, 0, 0), (0, 0, 0), Y, 0.05)
    b += link("l2", 0.7, (0, 0, -0.2), (0.01, 0.01, 0.001)) + joint("j2", "revolute", "l1", "l2", (0, 0, -0.5), (0, 0, 0), Y, 0.05)
    write("pendulum2", b)

    b = link("base", 1.0, (0, 0, 0), (0.01, 0.01, 0.01))
    b += link("l1", 2.0, (0.01, 0.02, 0.1), (0.03, 0.025, 0.01, 0.001, -0.002, 0.0015)) + joint("j1", "revolute", "base", "l1", (0, 0, 0.1), (0, 0, 0.3), Z, 0.1)
    b += link("l2", 1.5, (0.1, 0.0, 0.02), (0.004, 0.02, 0.021, 0.0, 0.001, 0.0)) + joint("j2", "revolute", "l1", "l2", (0.05, 0, 0.2), (0.2, -0.4, 0.1), Y, 0.2)
    b += link("l3", 0.8, (0.0, 0.05, 0.05), (0.006, 0.004, 0.005, 0.0005, 0.0, -0.0004)) + joint("j3", "revolute", "l2", "l3", (0.2, 0.01, 0), (HP, 0.1, 0), X, 0.0)
    write("chain3", b)

    # prismatic joint: valid for RNEA / Minv / FD / ABA / CRBA; the reference's gradient
    # algorithm assumes revolute joints (its Mxf term, _test.py:311, :437)
    b = link("base", 1.0, (0, 0, 0), (0.01, 0.01, 0.01))
    b += link("l1", 2.0, (0.01, 0.02, 0.1), (0.03, 0.025, 0.01, 0.001, -0.002, 0.0015)) + joint("j1", "revolute", "base", "l1", (0, 0, 0.1), (0, 0, 0.3), Z, 0.1)
    b += link("l2", 1.5, (0.1, 0.0, 0.02), (0.004, 0.02, 0.021, 0.0, 0.001, 0.0)) + joint("j2", "prismatic", "l1", "l2", (0.05, 0, 0.2), (0.2, -0.4, 0.1), X, 0.2)
    b += link("l3", 0.8, (0.0, 0.05, 0.05), (0.006, 0.004, 0.005, 0.0005, 0.0, -0.0004)) + joint("j3", "revolute", "l2", "l3", (0.2, 0.01, 0), (HP, 0.1, 0), X, 0.0)
    write("slider3", b)

    b = link("base", 1.0, (0, 0, 0), (0.01, 0.01, 0.01))
    b += link("t1", 3.0, (0, 0, 0.1), (0.05, 0.05, 0.02)) + joint("jt1", "revolute", "base", "t1", (0, 0, 0.05), (0, 0, 0), Z, 0.1)
    b += link("a1", 1.0, (0.1, 0, 0), (0.002, 0.01, 0.01)) + joint("ja1", "revolute", "t1", "a1", (0, 0.1, 0.2), (0, 0, 0), Y, 0.0)
    b += link("a2", 0.5, (0.08, 0, 0.01), (0.001, 0.004, 0.004)) + joint("ja2", "revolute", "a1", "a2", (0.2, 0, 0), (0.3, 0, 0), X, 0.02)
    b += link("b1", 1.2, (0.1, 0, 0), (0.002, 0.012, 0.012)) + joint("jb1", "revolute", "t1", "b1", (0, -0.1, 0.2), (0, 0.2, 0), Y, 0.0)
    b += link("b2", 0.4, (0.0, 0.02, 0.07), (0.003, 0.003, 0.0008)) + joint("jb2", "revolute", "b1", "b2", (0.2, 0, 0), (0, 0, 0), Z, 0.3)
    b += link("c1", 0.9, (0, 0.05, 0), (0.004, 0.001, 0.004)) + joint("jc1", "revolute", "base", "c1", (0.2, 0, 0), (0, 0, 1.0), X, 0.0)
    write("branched6", b)


if __name__ == "__main__":
    iiwa14()
    hyq()
    atlas()
    toys()
    print("wrote", sorted(os.listdir(OUT)))
