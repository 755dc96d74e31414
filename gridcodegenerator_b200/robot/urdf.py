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


def parse_urdf_string(xml_text: str, name: str | None = None) -> Robot:
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
    return Robot(rname, links, joints)


def parse_urdf(path: str, name: str | None = None) -> Robot:
    with open(path, "r") as f:
        txt = f.read()
    return parse_urdf_string(txt, name or os.path.splitext(os.path.basename(path))[0])


class URDFParser:
    """Drop-in for `URDFParser().parse(path)` as used with the reference."""

    def parse(self, filename, alpha_tie_breaker=False, floating_base=False):
        if floating_base:
            raise NotImplementedError("floating-base models are out of scope (SURVEY.md section 2 row 18)")
        return parse_urdf(filename)
