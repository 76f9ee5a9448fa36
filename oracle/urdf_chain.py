"""URDF -> KDL chain descriptor, CPU restatement.  TEST INFRASTRUCTURE ONLY (see oracle/README.md).

Follows the reference's vendored parser and robot constructor:
  - tup::TinyURDFParser::init / buildKDLTree / parseXMLJoint / rawArrayToEigenVector
    (ilqr_planner/ilqr_planner/lib/tinyurdfparser/TinyURDFParser/TinyURDFParser.cpp:68-160,163-218,270-304)
  - KDL::Tree::getChain(base, tip)                 (orocos_kdl, serial chain base -> tip)
  - KDL::Segment(joint, f_tip): tip = joint.pose(0)^-1 * f_tip; KDL::Joint normalises its axis
  - sim::KDLRobot ctor: extra fixed segment "robot_custom_tip" with Rotation::EulerZYX(rpy0, rpy1, rpy2)
    (ilqr_planner/ilqr_planner/src/sim/KDLRobot.cpp:60-66)

Returns a list of dicts {joint_type, axis[3], origin[3], tip_R[9] (row-major), tip_p[3]} — the fields of
`ilqrb_segment` in include/ilqrb.h.
"""
import math
import xml.etree.ElementTree as ET

import numpy as np

JOINT_FIXED, JOINT_ROT, JOINT_TRANS = 0, 1, 2
_TYPE = {"prismatic": JOINT_TRANS, "revolute": JOINT_ROT, "continuous": JOINT_ROT, "fixed": JOINT_FIXED}


def rot_rpy(roll, pitch, yaw):
    """KDL::Rotation::RPY (frames.cpp)."""
    ca1, sa1 = math.cos(yaw), math.sin(yaw)
    cb1, sb1 = math.cos(pitch), math.sin(pitch)
    cc1, sc1 = math.cos(roll), math.sin(roll)
    return np.array(
        [
            [ca1 * cb1, ca1 * sb1 * sc1 - sa1 * cc1, ca1 * sb1 * cc1 + sa1 * sc1],
            [sa1 * cb1, sa1 * sb1 * sc1 + ca1 * cc1, sa1 * sb1 * cc1 - ca1 * sc1],
            [-sb1, cb1 * sc1, cb1 * cc1],
        ]
    )


def rot_euler_zyx(alfa, beta, gamma):
    """KDL::Rotation::EulerZYX(Alfa,Beta,Gamma) = RPY(Gamma,Beta,Alfa)."""
    return rot_rpy(gamma, beta, alfa)


def _vec3(raw):
    """rawArrayToEigenVector: space separated, exactly 3 numbers (TinyURDFParser.cpp:270-304)."""
    toks = [t for t in raw.rstrip().split(" ") if t != ""]
    if len(toks) > 3:
        raise RuntimeError("[urdf_parsing] Expected a 3d array got more, check arrays format.")
    if len(toks) < 3:
        raise RuntimeError("[urdf_parsing] Expected a 3d array got less, check arrays format.")
    return np.array([float(t) for t in toks])


def parse_urdf(urdf_text):
    root = ET.fromstring(urdf_text)
    if root.tag != "robot":
        raise RuntimeError("[urdf_parsing] Root element should be named <robot> not <%s>" % root.tag)
    joints, links = {}, []
    for j in root.findall("joint"):  # only top-level children (TinyURDFParser.cpp:75)
        name, typ = j.get("name"), j.get("type")
        if name is None:
            raise RuntimeError('[urdf_parsing] Mandatory attribute "name" is missing for one of the joint')
        if typ is None:
            raise RuntimeError('[urdf_parsing] Mandatory attribute "type" is missing for joint ' + name)
        if typ not in _TYPE:
            raise RuntimeError('[urdf_parsing] Unknown joint type "%s"' % typ)
        rec = dict(name=name, type=typ, parent=j.find("parent").get("link"), child=j.find("child").get("link"),
                   xyz=np.zeros(3), rpy=np.zeros(3), axis=np.array([1.0, 0.0, 0.0]))
        org = j.find("origin")
        if org is not None:
            if org.get("rpy") is not None:
                rec["rpy"] = _vec3(org.get("rpy"))
            if org.get("xyz") is not None:
                rec["xyz"] = _vec3(org.get("xyz"))
        ax = j.find("axis")
        if typ in ("revolute", "continuous", "prismatic", "planar") and ax is not None and ax.get("xyz") is not None:
            rec["axis"] = _vec3(ax.get("xyz"))
        joints[name] = rec
    for l in root.findall("link"):
        links.append(l.get("name"))
    if not joints or not links:
        raise RuntimeError("[urdf_parsing] URDF file does not have <joint> or <link> tags")
    return joints, links


def kdl_chain_from_urdf(urdf_text, base, tip, transform_rpy=(0, 0, 0), transform_xyz=(0, 0, 0)):
    joints, links = parse_urdf(urdf_text)
    # link -> "parent" joint pointer as TUP sets it (Q17): the LAST joint in std::map (name) order whose parent is the link
    link_joint = {}
    for name in sorted(joints):
        link_joint[joints[name]["parent"]] = joints[name]
    child_to_joint = {j["child"]: j for j in joints.values()}
    if tip not in child_to_joint and tip != base:
        raise RuntimeError("[KDLRobot] Unable to build kinematic chain from %s to %s" % (base, tip))
    # walk tip -> base through child links
    path, cur = [], tip
    while cur != base:
        if cur not in child_to_joint:
            raise RuntimeError("[KDLRobot] Unable to build kinematic chain from %s to %s" % (base, tip))
        j = child_to_joint[cur]
        path.append(j)
        cur = j["parent"]
    path.reverse()
    root_links = set(l for l in links if l not in child_to_joint)
    segs = []
    for cur in path:
        R = rot_rpy(*cur["rpy"])
        xyz = cur["xyz"].copy()
        # joints hanging directly off a root link are reached with hook == "root"?  No: the root link itself is the
        # segment added with hook "root" (start_ is a JOINT element, TinyURDFParser.cpp:49,124-133); its children see
        # hook = root link name.  So every URDF joint gets `previous_joint` = link_joint[parent link].
        prev = link_joint[cur["parent"]]
        if prev["type"] == "fixed":
            jt, axis, origin = JOINT_FIXED, np.zeros(3), np.zeros(3)
        else:
            jt = _TYPE[prev["type"]]
            axis = R @ prev["axis"]
            axis = axis / np.linalg.norm(axis)  # KDL::Joint ctor normalises
            origin = xyz.copy()
        # KDL::Segment: tip = joint.pose(0)^-1 * f_tip ; joint.pose(0) = (I, origin) for Rot/Trans axis joints
        tip_R = R
        tip_p = xyz - origin
        segs.append(dict(joint_type=jt, axis=axis, origin=origin, tip_R=tip_R.reshape(-1), tip_p=tip_p))
    # KDLRobot.cpp:60-66
    Rv = rot_euler_zyx(transform_rpy[0], transform_rpy[1], transform_rpy[2])
    segs.append(dict(joint_type=JOINT_FIXED, axis=np.zeros(3), origin=np.zeros(3), tip_R=Rv.reshape(-1), tip_p=np.array(transform_xyz, dtype=float)))
    _ = root_links
    return segs


def chain_dof(segs):
    return sum(1 for s in segs if s["joint_type"] != JOINT_FIXED)
