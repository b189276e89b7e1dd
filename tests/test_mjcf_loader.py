"""MJCF-subset loader (sspp_model.cpp parse_mjcf): the input format in front of the scoring path.

The reference hands its scenes to MuJoCo's compiler (mj_loadXML, /root/reference/include/sspp.h:47);
here the same files are read without MuJoCo, so the constructs the shipped scenes use
(default classes + childclass /root/reference/mjcf/robocrane/robocrane.xml:5-45, <exclude> :199-206,
fromto, euler in degrees) and their neighbours in the MJCF schema (frames, includes, several joints
in one body) are pinned here against hand-computed poses and MuJoCo's documented rules.
"""
import math

import numpy as np
import pytest

from sspp_b200 import capi
from oracle import pyoracle

pyoracle.build()


def parse(text):
    return capi.Model(text=text)


def frames(model, q, dof):
    o1 = pyoracle.O1(model, 0, dof)
    return o1.geom_frames(np.asarray(q, float))


def wrap(world, extra="", compiler='<compiler angle="degree"/>'):
    return f"<mujoco>{compiler}{extra}<worldbody>{world}</worldbody></mujoco>"


def rotz(t):
    c, s = math.cos(t), math.sin(t)
    return np.array([[c, -s, 0], [s, c, 0], [0, 0, 1.0]])


def roty(t):
    c, s = math.cos(t), math.sin(t)
    return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]])


def rotx(t):
    c, s = math.cos(t), math.sin(t)
    return np.array([[1, 0, 0], [0, c, -s], [0, s, c]])


# ------------------------------------------------------------------------------------------------
# default classes
# ------------------------------------------------------------------------------------------------
DEFAULTS = """
<default>
  <geom contype="3" margin="0.01"/>
  <joint axis="0 1 0"/>
  <default class="outer">
    <geom conaffinity="5" size="0.1"/>
    <default class="inner"><geom contype="9"/></default>
  </default>
</default>"""


def test_nested_default_classes_inherit_and_childclass_applies():
    m = parse(wrap("""
      <geom name="g_main" type="sphere" size="0.2"/>
      <body name="b" childclass="outer" pos="0 0 1">
        <joint name="j"/>
        <geom name="g_child"/>
        <geom name="g_inner" class="inner"/>
        <geom name="g_own" contype="7" size="0.3"/>
        <body name="c"><geom name="g_deep"/></body>
        <body name="d" childclass="inner"><geom name="g_override"/></body>
      </body>""", DEFAULTS))
    a = m.arrays()
    by = {m.geom_name(g): g for g in range(a["ngeom"])}
    ct, ca, sz, mg = a["geom_contype"], a["geom_conaffinity"], a["geom_size"], a["geom_margin"]
    assert (ct[by["g_main"]], ca[by["g_main"]], mg[by["g_main"]]) == (3, 1, 0.01)
    assert (ct[by["g_child"]], ca[by["g_child"]], sz[by["g_child"], 0]) == (3, 5, 0.1)
    assert (ct[by["g_inner"]], ca[by["g_inner"]], sz[by["g_inner"], 0]) == (9, 5, 0.1)
    assert (ct[by["g_own"]], ca[by["g_own"]], sz[by["g_own"], 0]) == (7, 5, 0.3)
    assert (ct[by["g_deep"]], ca[by["g_deep"]]) == (3, 5)            # childclass reaches grandchildren
    assert (ct[by["g_override"]], ca[by["g_override"]]) == (9, 5)    # a nearer childclass wins
    assert np.allclose(a["jnt_axis"][m.body_id("b")], [0, 1, 0])     # joint defaults of the main class


def test_unknown_class_and_bad_numbers_are_errors():
    with pytest.raises(capi.SsppError, match="unknown default class"):
        parse(wrap('<geom class="nope" type="sphere" size="1"/>'))
    with pytest.raises(capi.SsppError, match="size is incomplete"):
        parse(wrap('<geom type="box" size="1 1"/>'))
    with pytest.raises(capi.SsppError, match="pos needs 3"):
        parse(wrap('<geom type="sphere" size="1" pos="1 2"/>'))
    with pytest.raises(capi.SsppError, match="repeated body name"):
        parse(wrap('<body name="a"/><body name="a"/>'))
    with pytest.raises(capi.SsppError, match="ref other than 0"):
        parse(wrap('<body><joint ref="0.2"/><geom size="1"/></body>'))
    with pytest.raises(capi.SsppError, match="explicit geom pairs"):
        parse(wrap('<geom name="a" size="1"/>', '<contact><pair geom1="a" geom2="a"/></contact>'))
    with pytest.raises(capi.SsppError, match="collidable mesh"):
        parse(wrap('<geom type="mesh" mesh="m"/>'))
    parse(wrap('<geom type="mesh" mesh="m" contype="0" conaffinity="0"/><geom size="1"/>'))   # visual mesh: skipped


# ------------------------------------------------------------------------------------------------
# orientation attributes and geom sizes
# ------------------------------------------------------------------------------------------------
def test_orientation_forms_agree():
    t = math.radians(30)
    q = f"{math.cos(t / 2)} 0 0 {math.sin(t / 2)}"
    forms = [f'quat="{q}"', 'euler="0 0 30"', 'axisangle="0 0 2 30"',
             f'xyaxes="{math.cos(t)} {math.sin(t)} 0 {-math.sin(t)} {math.cos(t)} 0"']
    world = "".join(f'<geom type="box" size="1 2 3" pos="{i} 0 0" {f}/>' for i, f in enumerate(forms))
    m = parse(wrap(world))
    _, xmat = frames(m, [], 0)
    for i in range(len(forms)):
        assert np.allclose(xmat[i], rotz(t), atol=1e-12), forms[i]


def test_euler_sequence_and_radians():
    # default eulerseq "xyz" is intrinsic: R = Rx Ry Rz; upper case letters are extrinsic
    m = parse(wrap('<geom type="box" size="1 1 1" euler="0.3 -0.4 0.5"/>', compiler='<compiler angle="radian"/>'))
    _, xmat = frames(m, [], 0)
    assert np.allclose(xmat[0], rotx(0.3) @ roty(-0.4) @ rotz(0.5), atol=1e-12)
    m = parse(wrap('<geom type="box" size="1 1 1" euler="0.3 -0.4 0.5"/>', compiler='<compiler angle="radian" eulerseq="XYZ"/>'))
    _, xmat = frames(m, [], 0)
    assert np.allclose(xmat[0], rotz(0.5) @ roty(-0.4) @ rotx(0.3), atol=1e-12)
    m = parse(wrap('<geom type="box" size="1 1 1" euler="0.3 -0.4 0.5"/>', compiler='<compiler angle="radian" eulerseq="zyx"/>'))
    _, xmat = frames(m, [], 0)
    assert np.allclose(xmat[0], rotz(0.3) @ roty(-0.4) @ rotx(0.5), atol=1e-12)


def test_fromto_sets_centre_axis_and_half_length():
    m = parse(wrap("""
      <geom name="cap" type="capsule" size="0.05" fromto="1 0 0 1 0 2"/>
      <geom name="cyl" type="cylinder" size="0.1" fromto="0 0 0 3 4 0"/>
      <geom name="box" type="box" size="0.1 0.2" fromto="0 1 1 0 1 3"/>"""))
    a = m.arrays()
    xpos, xmat = frames(m, [], 0)
    assert np.allclose(xpos[0], [1, 0, 1]) and np.allclose(a["geom_size"][0, :2], [0.05, 1.0])
    assert np.allclose(np.abs(xmat[0][:, 2]), [0, 0, 1])
    assert np.allclose(xpos[1], [1.5, 2, 0]) and np.allclose(a["geom_size"][1, :2], [0.1, 2.5])
    assert np.allclose(np.abs(xmat[1][:, 2]), [0.6, 0.8, 0], atol=1e-12)
    assert np.allclose(xpos[2], [0, 1, 2]) and np.allclose(a["geom_size"][2], [0.1, 0.2, 1.0])
    for k in range(3):      # proper rotations
        assert np.allclose(xmat[k] @ xmat[k].T, np.eye(3), atol=1e-12) and np.linalg.det(xmat[k]) > 0.999


def test_default_geom_type_is_sphere_and_plane_needs_no_size():
    m = parse(wrap('<geom size="0.4"/><geom type="plane" size="0 0 0.01"/><geom type="plane"/>'))
    a = m.arrays()
    assert list(a["geom_type"]) == [2, 0, 0]      # mjGEOM_SPHERE, mjGEOM_PLANE


# ------------------------------------------------------------------------------------------------
# frames, includes, entities
# ------------------------------------------------------------------------------------------------
def test_frames_compose_onto_geoms_and_bodies():
    m = parse(wrap("""
      <frame pos="1 0 0" euler="0 0 90">
        <geom name="a" type="box" size="1 2 3" pos="1 0 0"/>
        <frame pos="0 0 2">
          <body name="b" pos="0 1 0"><geom name="c" type="sphere" size="0.1" pos="1 0 0"/></body>
        </frame>
      </frame>"""))
    xpos, xmat = frames(m, [], 0)
    assert np.allclose(xpos[0], [1, 1, 0], atol=1e-12) and np.allclose(xmat[0], rotz(math.pi / 2), atol=1e-12)
    # body b: frame (1,0,0)+Rz90*(0,0,2)+Rz90*(0,1,0) = (0,0,2); geom c: + Rz90*(1,0,0) = (0,1,2)
    a = m.arrays()
    assert np.allclose(a["body_pos"][m.body_id("b")], [0, 0, 2], atol=1e-12)
    assert np.allclose(xpos[1], [0, 1, 2], atol=1e-12)


def test_include_is_resolved_against_the_including_file(tmp_path):
    (tmp_path / "parts").mkdir()
    (tmp_path / "parts" / "defs.xml").write_text('<mujocoinclude><default><geom contype="6"/></default></mujocoinclude>')
    (tmp_path / "parts" / "arm.xml").write_text(
        '<mujocoinclude><body name="arm" pos="0 0 1"><joint axis="0 0 1"/><geom name="link" type="capsule" size="0.05 0.2"/>'
        '<include file="tip.xml"/></body></mujocoinclude>')
    (tmp_path / "parts" / "tip.xml").write_text('<mujocoinclude><geom name="tip" size="0.02" pos="0 0 0.3"/></mujocoinclude>')
    (tmp_path / "scene.xml").write_text(
        '<mujoco><include file="parts/defs.xml"/><worldbody><geom name="floor" type="plane"/>'
        '<include file="parts/arm.xml"/></worldbody></mujoco>')
    m = capi.Model(path=str(tmp_path / "scene.xml"))
    a = m.arrays()
    assert [m.geom_name(g) for g in range(a["ngeom"])] == ["floor", "link", "tip"]
    assert list(a["geom_contype"]) == [6, 6, 6]
    with pytest.raises(capi.SsppError, match="no directory to resolve"):
        parse('<mujoco><include file="x.xml"/><worldbody/></mujoco>')
    (tmp_path / "loop.xml").write_text('<mujoco><include file="loop.xml"/><worldbody/></mujoco>')
    with pytest.raises(capi.SsppError, match="nesting deeper"):
        capi.Model(path=str(tmp_path / "loop.xml"))


def test_xml_details_comments_entities_cdata_and_quotes():
    m = parse("""<?xml version="1.0"?>
      <!-- header -->
      <mujoco model='a &amp; b'>
        <worldbody><!-- inside -->
          <geom name='g&lt;1&gt;' size='0.5'/> text is ignored <![CDATA[ <geom size="9"/> ]]>
        </worldbody>
      </mujoco>""")
    assert m.arrays()["ngeom"] == 1 and m.geom_name(0) == "g<1>"
    with pytest.raises(capi.SsppError, match="mismatched"):
        parse("<mujoco><worldbody></mujoco>")
    with pytest.raises(capi.SsppError, match="expected <mujoco>"):
        parse("<robot/>")


# ------------------------------------------------------------------------------------------------
# kinematic tree: several joints per body, anchors, slide joints
# ------------------------------------------------------------------------------------------------
def test_several_joints_in_one_body_act_in_order():
    # MuJoCo applies a body's joints in declaration order, each in the frame left by the earlier ones
    m = parse(wrap("""
      <geom name="floor" type="plane"/>
      <body name="wrist" pos="0 0 1">
        <joint name="yaw" axis="0 0 1"/>
        <joint name="pitch" axis="0 1 0" pos="0 0 0.5"/>
        <joint name="reach" type="slide" axis="1 0 0"/>
        <geom name="hand" type="box" size="0.1 0.1 0.1" pos="0.2 0 0"/>
        <body name="finger" pos="0.4 0 0"><geom name="pad" size="0.02"/></body>
      </body>""", compiler='<compiler angle="radian"/>'))
    a = m.arrays()
    assert a["nq"] == 3 and a["nbody"] == 5           # world + 3 links standing for "wrist" + finger
    wrist = m.body_id("wrist")
    assert wrist == 3 and m.body_name(1) == "" and m.body_name(2) == ""
    assert list(a["body_weld"][[1, 2, 3]]) == [1, 1, 1]            # one body as far as the collision filters go
    assert a["body_weld"][m.body_id("finger")] == 1
    yaw, pitch, reach = 0.7, -0.4, 0.25
    xpos, xmat = frames(m, [yaw, pitch, reach], 3)
    R = rotz(yaw) @ roty(pitch)
    anchor = np.array([0, 0, 0.5])
    origin = np.array([0, 0, 1.0]) + rotz(yaw) @ (anchor - roty(pitch) @ anchor) + R @ np.array([reach, 0, 0])
    hand = origin + R @ np.array([0.2, 0, 0])
    pad = origin + R @ np.array([0.4, 0, 0])
    assert np.allclose(xmat[1], R, atol=1e-12)
    assert np.allclose(xpos[1], hand, atol=1e-12) and np.allclose(xpos[2], pad, atol=1e-12)
    # wrist (3 links) and finger (welded to it) never collide with each other; both see the floor
    o1 = pyoracle.O1(m, 0, 3)
    pairs = {tuple(sorted(p)) for p in o1.pairs().tolist()}
    assert pairs == {(0, 1), (0, 2)}


def test_parent_child_filter_sees_through_joint_chains():
    m = parse(wrap("""
      <body name="base"><joint axis="0 0 1"/><geom name="g0" size="0.1"/>
        <body name="two" pos="0 0 0.3"><joint axis="1 0 0"/><joint axis="0 1 0"/><geom name="g1" size="0.1"/>
          <body name="three" pos="0 0 0.3"><joint axis="1 0 0"/><geom name="g2" size="0.1"/></body>
        </body>
      </body>"""))
    o1 = pyoracle.O1(m, 0, 4)
    pairs = {tuple(sorted(p)) for p in o1.pairs().tolist()}
    assert pairs == {(0, 2)}          # g0-g1 and g1-g2 are parent-child, g0-g2 is not


def test_exclude_and_masks_filter_pairs():
    xml = wrap("""
      <geom name="floor" type="plane" contype="1" conaffinity="1"/>
      <body name="a" pos="0 0 1"><joint/><geom name="ga" size="0.1" contype="2" conaffinity="1"/></body>
      <body name="b" pos="1 0 1"><joint/><geom name="gb" size="0.1" contype="1" conaffinity="2"/></body>
      <body name="c" pos="2 0 1"><joint/><geom name="gc" size="0.1" contype="4" conaffinity="4"/></body>""",
               '<contact><exclude body1="a" body2="b"/></contact>')
    m = parse(xml)
    o1 = pyoracle.O1(m, 0, 3)
    pairs = {tuple(sorted(p)) for p in o1.pairs().tolist()}
    # floor-ga: (1&1)|(2&1) -> yes; floor-gb: (1&2)|(1&1) -> yes; floor-gc: no; ga-gb excluded; gc sees nobody
    assert pairs == {(0, 1), (0, 2)}
    m2 = parse(xml.replace('<contact><exclude body1="a" body2="b"/></contact>', ""))
    pairs2 = {tuple(sorted(p)) for p in pyoracle.O1(m2, 0, 3).pairs().tolist()}
    assert pairs2 == {(0, 1), (0, 2), (1, 2)}
    with pytest.raises(capi.SsppError, match="unknown body"):
        parse(xml.replace('body2="b"', 'body2="zzz"'))


def test_shipped_scenes_still_parse_to_the_same_tables():
    from tests.helpers import ROOT
    for name, nb, ng in (("planner.xml", None, None), ("stacking.xml", None, None), ("robocrane.xml", None, None), ("robocrane7.xml", None, None)):
        m = capi.Model(path=str(ROOT / "fixtures" / name))
        a = m.arrays()
        assert a["ngeom"] > 0 and a["nbody"] > 1
        assert all(m.body_name(b) != "" for b in range(a["nbody"]))      # no body of these scenes needs a link chain


# ------------------------------------------------------------------------------------------------
# the constructs above, through model compilation and both oracles
# ------------------------------------------------------------------------------------------------
def test_armlab_scene_compiles_and_oracles_agree():
    """fixtures/armlab.xml (frames, a two-joint body, fromto capsules, spheres, a cylinder, nested
    classes, an exclude): the FP64 oracle (own FK over the flat model) and the FP32 replica (the
    compiled kernel tables) must see the same contacts."""
    from tests.helpers import armlab_case, o1_for, o2_for
    case = armlab_case()
    m = case.model
    a = m.arrays()
    assert a["nq"] == 6 and a["nbody"] == 7 and m.body_name(4) == "" and m.body_id("wrist") == 5
    o1, o2 = o1_for(case), o2_for(case)
    pairs = {tuple(sorted(p)) for p in o1.pairs().tolist()}
    by = {m.geom_name(g): g for g in range(a["ngeom"])}
    assert (by["jar"], by["pad"]) in pairs and (by["jar"], by["tip"]) in pairs and (by["jar"], by["palm"]) in pairs
    assert not any(by["jar"] in p and (by["fore_link"] in p or by["wrist_ball"] in p) for p in pairs)   # masks
    assert (by["column"], by["fore_link"]) not in pairs                  # <exclude body1="base" body2="fore">
    assert (by["column"], by["wrist_ball"]) in pairs                     # ... but not its descendants
    assert (by["wrist_ball"], by["pad"]) not in pairs                    # finger is the child of the (two-link) wrist
    rng = np.random.default_rng(5)
    Q = rng.uniform(-2.0, 2.0, (3000, 6)); Q[:, 5] = rng.uniform(0, 0.04, 3000)
    bad = amb = hit = 0
    for q in Q:
        c1, ncon, md = o1.point_cost(q)
        c2 = o2.point_cost(q)
        hit += c1 != 0.0
        if (c1 != 0.0) != (c2 != 0.0):
            # first contact within 1e-4 of a margin (0, 0.001, 0.002 in this scene): FP32 may fall on either side
            if min(abs(md - t) for t in (0.0, 1e-3, 2e-3)) <= 1e-4: amb += 1
            else: bad += 1
    assert bad == 0 and amb <= 15, (bad, amb)
    assert 0.05 < hit / len(Q) < 0.95
