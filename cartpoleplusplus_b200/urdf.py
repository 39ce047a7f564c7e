"""URDF -> per-body constants for the three boxes of the cartpole++ scene.

The reference hands `models/ground.urdf`, `models/cart.urdf` and `models/pole.urdf`
to pybullet (`p.loadURDF`, bullet_cartpole.py:154-156).  Here the host parses them once
into mass / inertia / extent / contact constants; nothing else of URDF is needed for a
single-link box body.

How pybullet turns the file into a body ([BULLET-RECALL], examples/Importers/ImportURDFDemo
`URDF2Bullet.cpp`, `BulletUrdfImporter.cpp`; parameters, not literals):
  * `loadURDF` is called without flags, so the local inertia diagonal is *recomputed from
    the collision shape* (box formula m/12 (b^2+c^2)) and then multiplied by the link's
    `<inertia_scaling>` (3.0 for cart and pole).  `inertia_from_file=True` keeps the file's
    `<inertia>` instead (URDF_USE_INERTIA_FROM_FILE), still scaled.
  * `<lateral_friction>` sets the collision object's friction; absent => Bullet default 0.5.
  * `<restitution>` absent => 0.  `<contact_cfm>/<contact_erp>` are parsed but not applied
    to contact rows; `<rolling_friction value="0.0">` adds no rows.
"""
from __future__ import annotations

import dataclasses
import os
import xml.etree.ElementTree as ET
from typing import Optional, Tuple

MODEL_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "models")

BULLET_DEFAULT_FRICTION = 0.5  # btCollisionObject::m_friction
BULLET_DEFAULT_RESTITUTION = 0.0


@dataclasses.dataclass
class BoxBody:
    name: str
    mass: float
    inertia_file: Tuple[float, float, float]  # <inertia ixx iyy izz> as written
    box_size: Tuple[float, float, float]  # full extents of the collision box
    lateral_friction: float
    restitution: float
    inertia_scaling: float
    rolling_friction: float = 0.0
    contact_cfm: Optional[float] = None
    contact_erp: Optional[float] = None
    rgba: Tuple[float, float, float, float] = (1.0, 1.0, 1.0, 1.0)  # <visual><material><color rgba>: the pixel observation's colours
    # compound bodies (parse_compound_cart_urdf): the friction of the half that touches the ground, and the composite
    # inertia diagonal with <inertia_scaling> already applied link by link
    ground_friction: Optional[float] = None
    composite_inertia: Optional[Tuple[float, float, float]] = None

    @property
    def half_extents(self) -> Tuple[float, float, float]:
        return tuple(0.5 * s for s in self.box_size)

    def box_inertia(self) -> Tuple[float, float, float]:
        """btBoxShape::calculateLocalInertia: m/12 * (ly^2+lz^2, lx^2+lz^2, lx^2+ly^2)."""
        lx, ly, lz = (2.0 * h for h in self.half_extents)
        m = self.mass
        return (m / 12.0 * (ly * ly + lz * lz), m / 12.0 * (lx * lx + lz * lz), m / 12.0 * (lx * lx + ly * ly))

    def effective_inertia(self, inertia_from_file: bool = False) -> Tuple[float, float, float]:
        if self.mass == 0.0:
            return (0.0, 0.0, 0.0)
        if self.composite_inertia is not None:
            return self.composite_inertia
        base = self.inertia_file if inertia_from_file else self.box_inertia()
        return tuple(self.inertia_scaling * v for v in base)


def _value(elem, tag, default=None):
    node = elem.find(tag) if elem is not None else None
    if node is None:
        return default
    return float(node.attrib["value"])


def parse_box_urdf(path: str) -> BoxBody:
    """Parse a single-link URDF whose collision geometry is one origin-centred box."""
    root = ET.parse(path).getroot()
    if root.tag != "robot":
        raise ValueError(f"{path}: not a URDF (root <{root.tag}>)")
    links = root.findall("link")
    if len(links) != 1 or root.findall("joint"):
        raise ValueError(f"{path}: expected exactly one link and no joints")
    link = links[0]
    inertial = link.find("inertial")
    if inertial is None:
        raise ValueError(f"{path}: <inertial> missing")
    mass = float(inertial.find("mass").attrib["value"])
    inertia = inertial.find("inertia").attrib
    for off in ("ixy", "ixz", "iyz"):
        if float(inertia.get(off, 0.0)) != 0.0:
            raise ValueError(f"{path}: off-diagonal inertia not supported")
    collision = link.find("collision")
    if collision is None:
        raise ValueError(f"{path}: <collision> missing")
    origin = collision.find("origin")
    if origin is not None:
        xyz = [float(v) for v in origin.attrib.get("xyz", "0 0 0").split()]
        rpy = [float(v) for v in origin.attrib.get("rpy", "0 0 0").split()]
        if any(xyz) or any(rpy):
            raise ValueError(f"{path}: collision origin offset not supported")
    box = collision.find("geometry/box")
    if box is None:
        raise ValueError(f"{path}: collision geometry must be a <box>")
    size = tuple(float(v) for v in box.attrib["size"].split())
    if len(size) != 3:
        raise ValueError(f"{path}: box size needs 3 values")
    contact = link.find("contact")
    color = link.find("visual/material/color")
    rgba = tuple(float(v) for v in color.attrib["rgba"].split()) if color is not None and "rgba" in color.attrib else (1.0, 1.0, 1.0, 1.0)
    return BoxBody(
        name=root.attrib.get("name", os.path.basename(path)),
        mass=mass,
        inertia_file=(float(inertia["ixx"]), float(inertia["iyy"]), float(inertia["izz"])),
        box_size=size,
        lateral_friction=_value(contact, "lateral_friction", BULLET_DEFAULT_FRICTION),
        restitution=_value(contact, "restitution", BULLET_DEFAULT_RESTITUTION),
        inertia_scaling=_value(contact, "inertia_scaling", 1.0),
        rolling_friction=_value(contact, "rolling_friction", 0.0),
        contact_cfm=_value(contact, "contact_cfm"),
        contact_erp=_value(contact, "contact_erp"),
        rgba=rgba,
    )


def parse_compound_cart_urdf(path: str, inertia_from_file: bool = False) -> BoxBody:
    """The fork's two-link cart (reference models/double_cart_1.urdf:15-80): two box links glued by a fixed joint, the
    upper one with friction against the pole, the lower one frictionless against the ground.

    pybullet builds a btMultiBody whose base is the parent link and whose one link hangs on a fixed (0-DOF) joint: it
    moves as ONE rigid body - the composite of the two halves - with two colliders that keep their own friction
    ([BULLET-RECALL]).  This parser returns that composite as a single box body:
      * mass = sum; inertia about the composite centre = sum over the links of (link inertia x its <inertia_scaling>,
        about the link's own centre: box formula unless inertia_from_file) + the parallel-axis term m d^2, which the
        scaling does not touch;
      * collision box = the union of the two flush halves (checked: same footprint, stacked without gap or overlap);
      * lateral_friction = the UPPER half's (what the pole touches), ground_friction = the LOWER half's (what the ground
        touches).
    Exactness of the single-box reduction: the pole touches the upper half's top face and the ground the lower half's
    bottom face - where the two-collider world and this body produce the same contact points with the same friction.  A
    pole leaning against the cart's side, or a cart lying on its side, would touch BOTH halves in the two-collider world;
    here every cart-pole contact then takes the upper friction and every ground-cart contact the lower one (stated in
    DESIGN.md section 11; the fork's game ends an episode long before)."""
    root = ET.parse(path).getroot()
    links = {l.attrib["name"]: l for l in root.findall("link")}
    joints = root.findall("joint")
    if len(links) != 2 or len(joints) != 1 or joints[0].attrib.get("type") != "fixed":
        raise ValueError(f"{path}: expected two links glued by one fixed joint")
    jo = joints[0].find("origin")
    if jo is not None and (any(float(v) for v in jo.attrib.get("xyz", "0 0 0").split()) or any(float(v) for v in jo.attrib.get("rpy", "0 0 0").split())):
        raise ValueError(f"{path}: joint origin offset not supported")
    parts = []
    for name in (joints[0].find("parent").attrib["link"], joints[0].find("child").attrib["link"]):
        link = links[name]
        inertial, collision, contact = link.find("inertial"), link.find("collision"), link.find("contact")
        com = [float(v) for v in inertial.find("origin").attrib.get("xyz", "0 0 0").split()]
        cxyz = [float(v) for v in collision.find("origin").attrib.get("xyz", "0 0 0").split()]
        if com != cxyz or com[0] or com[1]:
            raise ValueError(f"{path}: link {name}: collision box must be centred on the link's centre of mass, on the z axis")
        size = tuple(float(v) for v in collision.find("geometry/box").attrib["size"].split())
        ia = inertial.find("inertia").attrib
        color = link.find("visual/material/color")
        parts.append(dict(mass=float(inertial.find("mass").attrib["value"]), z=com[2], size=size,
                          inertia_file=(float(ia["ixx"]), float(ia["iyy"]), float(ia["izz"])),
                          friction=_value(contact, "lateral_friction", BULLET_DEFAULT_FRICTION),
                          restitution=_value(contact, "restitution", BULLET_DEFAULT_RESTITUTION),
                          scaling=_value(contact, "inertia_scaling", 1.0),
                          rgba=tuple(float(v) for v in color.attrib["rgba"].split()) if color is not None else (1.0, 1.0, 1.0, 1.0)))
    upper, lower = sorted(parts, key=lambda q: -q["z"])
    if upper["size"][:2] != lower["size"][:2] or abs((upper["z"] - upper["size"][2] / 2) - (lower["z"] + lower["size"][2] / 2)) > 1e-12:
        raise ValueError(f"{path}: the two boxes must share a footprint and be flush")
    mass = upper["mass"] + lower["mass"]
    zc = (upper["mass"] * upper["z"] + lower["mass"] * lower["z"]) / mass
    top, bottom = upper["z"] + upper["size"][2] / 2, lower["z"] - lower["size"][2] / 2
    if abs(zc - (top + bottom) / 2) > 1e-12:
        raise ValueError(f"{path}: composite centre of mass off the box centre: not supported")
    inertia = [0.0, 0.0, 0.0]
    for q in (upper, lower):
        lx, ly, lz = q["size"]
        own = q["inertia_file"] if inertia_from_file else (q["mass"] / 12 * (ly * ly + lz * lz), q["mass"] / 12 * (lx * lx + lz * lz), q["mass"] / 12 * (lx * lx + ly * ly))
        d2 = (q["z"] - zc) ** 2
        for i, par in enumerate((d2, d2, 0.0)):
            inertia[i] += q["scaling"] * own[i] + q["mass"] * par
    return BoxBody(name=root.attrib.get("name", os.path.basename(path)), mass=mass, inertia_file=tuple(inertia),
                   box_size=(upper["size"][0], upper["size"][1], top - bottom), lateral_friction=upper["friction"],
                   restitution=upper["restitution"], inertia_scaling=1.0, rgba=upper["rgba"],
                   ground_friction=lower["friction"], composite_inertia=tuple(inertia))


def load_scene(model_dir: Optional[str] = None, cart_model: str = "cart"):
    """(ground, cart, pole); cart_model 'double_cart' loads the fork's compound cart (models/double_cart.urdf)."""
    if cart_model == "double_cart":
        d = model_dir or MODEL_DIR
        return (parse_box_urdf(os.path.join(d, "ground.urdf")), parse_compound_cart_urdf(os.path.join(d, "double_cart.urdf")),
                parse_box_urdf(os.path.join(d, "pole.urdf")))
    return _load_scene(model_dir)


def _load_scene(model_dir: Optional[str] = None):
    """(ground, cart, pole) parsed from `model_dir` (default: the packaged models)."""
    d = model_dir or MODEL_DIR
    return tuple(parse_box_urdf(os.path.join(d, f"{n}.urdf")) for n in ("ground", "cart", "pole"))
