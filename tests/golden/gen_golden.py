"""Regenerates tests/golden/*.json from the read-only reference tree (run in the build
container only; /root/reference does not exist on the GPU box).

  model_constants.json : the literal numbers of models/{ground,cart,pole}.urdf
  add_opts_defaults.json: every flag of bullet_cartpole.add_opts with its default,
                          extracted with `ast` (the module itself cannot be imported:
                          it needs gym and pybullet)
"""
import ast
import json
import os
import xml.etree.ElementTree as ET

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def urdf_constants(name):
    root = ET.parse(os.path.join(REF, "models", name + ".urdf")).getroot()
    link = root.find("link")
    out = {"mass": float(link.find("inertial/mass").attrib["value"])}
    inertia = link.find("inertial/inertia").attrib
    out["inertia"] = [float(inertia[k]) for k in ("ixx", "iyy", "izz")]
    out["offdiag"] = [float(inertia[k]) for k in ("ixy", "ixz", "iyz")]
    out["box"] = [float(v) for v in link.find("collision/geometry/box").attrib["size"].split()]
    contact = link.find("contact")
    out["contact"] = {} if contact is None else {c.tag: float(c.attrib["value"]) for c in contact}
    return out


def add_opts_defaults():
    tree = ast.parse(open(os.path.join(REF, "bullet_cartpole.py")).read())
    fn = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "add_opts"][0]
    flags = {}
    for node in ast.walk(fn):
        if isinstance(node, ast.Call) and getattr(node.func, "attr", "") == "add_argument":
            name = node.args[0].value
            kw = {k.arg: k.value for k in node.keywords}
            if "action" in kw:
                flags[name] = False
            else:
                flags[name] = ast.literal_eval(kw["default"]) if "default" in kw else None
    return flags


def compound_constants(name):
    """the two links of a double_cart URDF (models/double_cart_1.urdf:15-80) and its joint"""
    root = ET.parse(os.path.join(REF, "models", name + ".urdf")).getroot()
    out = {"links": {}, "joint": {"type": root.find("joint").attrib["type"], "parent": root.find("joint/parent").attrib["link"],
                                   "child": root.find("joint/child").attrib["link"]}}
    for link in root.findall("link"):
        inertia = link.find("inertial/inertia").attrib
        out["links"][link.attrib["name"]] = {
            "mass": float(link.find("inertial/mass").attrib["value"]),
            "com": [float(v) for v in link.find("inertial/origin").attrib["xyz"].split()],
            "inertia": [float(inertia[k]) for k in ("ixx", "iyy", "izz")],
            "box": [float(v) for v in link.find("collision/geometry/box").attrib["size"].split()],
            "box_origin": [float(v) for v in link.find("collision/origin").attrib["xyz"].split()],
            "contact": {c.tag: float(c.attrib["value"]) for c in link.find("contact")}}
    return out


if __name__ == "__main__":
    json.dump({n: compound_constants(n) for n in ("double_cart_1", "double_cart_2")},
              open(os.path.join(HERE, "double_cart_constants.json"), "w"), indent=1, sort_keys=True)
    json.dump({n: urdf_constants(n) for n in ("ground", "cart", "pole")},
              open(os.path.join(HERE, "model_constants.json"), "w"), indent=1, sort_keys=True)
    json.dump(add_opts_defaults(), open(os.path.join(HERE, "add_opts_defaults.json"), "w"), indent=1, sort_keys=True)
    print("written")
