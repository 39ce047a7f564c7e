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


def load_scene(model_dir: Optional[str] = None):
    """(ground, cart, pole) parsed from `model_dir` (default: the packaged models)."""
    d = model_dir or MODEL_DIR
    return tuple(parse_box_urdf(os.path.join(d, f"{n}.urdf")) for n in ("ground", "cart", "pole"))
