"""Minimal URDF reader (stdlib xml.etree) producing the Robot model.

Stands in for URDFParser(...).parse(urdf) which the reference's README names
as the producer of its `robot` argument (reference README.md:10-21) but does
not vendor.  Only what rigid-body dynamics needs is read: link inertials,
joint type / origin / axis / dynamics damping / limits.
"""
from __future__ import annotations

import os
import xml.etree.ElementTree as ET
from typing import List

from .model import JointDesc, LinkDesc, Robot


def _floats(text, n, default):
    if text is None:
        return tuple(default)
    vals = [float(x) for x in text.split()]
    if len(vals) != n:
        raise ValueError(f"expected {n} numbers, got {text!r}")
    return tuple(vals)


# A 6-DoF (free) joint as a chain of six one-DoF joints over five massless links: translations along x, y, z of the
# parent frame, then rotations about x, y, z (roll, pitch, yaw applied in that order along the chain).  Every algorithm
# of the hot path then treats a floating-base robot as the fixed-base tree it already handles -- including the
# reference's own CPU oracle (_test.py has no floating-base branch), which is what pins these models.  Coordinates are
# q = [x y z roll pitch yaw | joints], qd = their rates (NOT the reference emitter's quaternion + body-twist
# convention, helpers/_topology_helpers.py:231-255, whose robot-side half lives in the un-vendored URDFParser);
# singular at pitch = +-90 degrees like every Euler-angle base.
FLOATING_CHAIN = (("px", "prismatic", (1.0, 0.0, 0.0)), ("py", "prismatic", (0.0, 1.0, 0.0)),
                  ("pz", "prismatic", (0.0, 0.0, 1.0)), ("rx", "revolute", (1.0, 0.0, 0.0)),
                  ("ry", "revolute", (0.0, 1.0, 0.0)), ("rz", "revolute", (0.0, 0.0, 1.0)))


def _expand_floating(links: List[LinkDesc], joints: List[JointDesc]) -> List[JointDesc]:
    """Replace every <joint type="floating"> by the six-joint chain above (five massless links are added)."""
    out: List[JointDesc] = []
    for jd in joints:
        if jd.jtype != "floating":
            out.append(jd)
            continue
        parent = jd.parent
        for i, (tag, jtype, axis) in enumerate(FLOATING_CHAIN):
            last = i == len(FLOATING_CHAIN) - 1
            child = jd.child if last else f"{jd.child}__{tag}"
            if not last:
                links.append(LinkDesc(name=child))                 # massless
            j = JointDesc(name=f"{jd.name}_{tag}", jtype=jtype, parent=parent, child=child)
            j.axis = axis
            if i == 0:
                j.xyz, j.rpy = jd.xyz, jd.rpy
            out.append(j)
            parent = child
    return out


def parse_urdf_string(xml_text: str, name: str | None = None, floating_base: bool = False) -> Robot:
    root = ET.fromstring(xml_text)
    if root.tag != "robot":
        raise ValueError("not a URDF: root element is <%s>" % root.tag)
    rname = name or root.get("name", "robot")
    links: List[LinkDesc] = []
    joints: List[JointDesc] = []
    for le in root.findall("link"):
        ld = LinkDesc(name=le.get("name"))
        ine = le.find("inertial")
        if ine is not None:
            o = ine.find("origin")
            if o is not None:
                ld.com = _floats(o.get("xyz"), 3, (0, 0, 0))
                ld.com_rpy = _floats(o.get("rpy"), 3, (0, 0, 0))
            m = ine.find("mass")
            if m is not None:
                ld.mass = float(m.get("value"))
            it = ine.find("inertia")
            if it is not None:
                ld.inertia = tuple(float(it.get(k, "0")) for k in ("ixx", "ixy", "ixz", "iyy", "iyz", "izz"))
        links.append(ld)
    for je in root.findall("joint"):
        jd = JointDesc(
            name=je.get("name"),
            jtype=je.get("type"),
            parent=je.find("parent").get("link"),
            child=je.find("child").get("link"),
        )
        o = je.find("origin")
        if o is not None:
            jd.xyz = _floats(o.get("xyz"), 3, (0, 0, 0))
            jd.rpy = _floats(o.get("rpy"), 3, (0, 0, 0))
        a = je.find("axis")
        if a is not None:
            jd.axis = _floats(a.get("xyz"), 3, (0, 0, 1))
        elif jd.jtype != "fixed":
            jd.axis = (1.0, 0.0, 0.0)  # URDF default axis
        d = je.find("dynamics")
        if d is not None:
            jd.damping = float(d.get("damping", "0"))
        lim = je.find("limit")
        if lim is not None:
            jd.limits = tuple(float(lim.get(k, "0")) for k in ("lower", "upper", "effort", "velocity"))
        joints.append(jd)
    if floating_base:
        # the root link (never a child) hangs from a new massless world link by a free joint
        children = {j.child for j in joints}
        roots = [l.name for l in links if l.name not in children]
        if len(roots) != 1:
            raise ValueError(f"floating_base=True needs exactly one root link, found {roots}")
        links.insert(0, LinkDesc(name="floating_base_world"))
        joints.insert(0, JointDesc(name="floating_base", jtype="floating", parent="floating_base_world", child=roots[0]))
    had_floating = any(j.jtype == "floating" for j in joints)
    joints = _expand_floating(links, joints)
    robot = Robot(rname, links, joints)
    # True when a free joint was expanded into the six-joint chain.  `floating_base` itself stays False: that flag
    # switches consumers to the quaternion convention (reference helpers/_topology_helpers.py:231), which this is not.
    robot.floating_base_chain = had_floating
    return robot


def parse_urdf(path: str, name: str | None = None, floating_base: bool = False) -> Robot:
    with open(path, "r") as f:
        txt = f.read()
    return parse_urdf_string(txt, name or os.path.splitext(os.path.basename(path))[0], floating_base)


class URDFParser:
    """Drop-in for `URDFParser().parse(path)` as used with the reference."""

    def parse(self, filename, alpha_tie_breaker=False, floating_base=False):
        """floating_base=True: the root link gets a free joint to the world, modelled as a six-joint chain
        (FLOATING_CHAIN above; the returned robot is a fixed-base tree with six more joints and
        `floating_base_chain = True`)."""
        return parse_urdf(filename, floating_base=floating_base)
